// Training-side DataArgumentationLayer forward on the device (B1 half of the C ABI).
//
// Restates /root/reference/scripts/data_argumentation_layer/argumentation_engine.py
// with the Python-2 / 2017-NumPy arithmetic it ran under:
//   flip_image (rect part)                :241-267      -> flip_rect()
//   random_argumentation (zoom-crop)      :157-173      -> crop_window()
//   crop_image_dimension                  :193-236      -> crop_window()
//   demean_rgb_image (float32, min-max)   :297-303      -> aug_minmax_kernel + aug_resample_kernel
//   resize_image_and_labels               :114-138      -> aug_resample_kernel (bilinear, see A11) + resize_rect()
//   bounding_box_parameterized_labels     :69-109       -> label_raster_kernel
//   JaccardCoeff.iou / generate_box_labels:24-55,272-278-> cell_labelled()
// The random draws are made on the host (argumentation_engine.py:146,165-166,200-202) and
// passed in (fcnod_aug_draw), so oracle and device see identical parameters.
// Label grids are bit-exact by construction: integer / double / IEEE-fp32 arithmetic in
// the reference's evaluation order, no fast-math, explicit __fdiv_rn.
#include "common.cuh"
#include "kernels.h"
#include "workspace.h"

namespace fcnod {
namespace {

struct AugGeom {          // per image, produced by aug_geometry_kernel
    int flip;             // -1 / 0 / 1, 2 = none
    int x0, y0, cw, ch;   // crop window in the (flipped) source image; cw/ch may be 0
    int rect[4];          // augmented rect at net resolution
    int label;
    unsigned mn_enc, mx_enc;   // order-preserving encodings of the fp32 min / max of the demeaned crop
};

__device__ __forceinline__ unsigned enc_f32(float f) {
    unsigned u = __float_as_uint(f);
    return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__device__ __forceinline__ float dec_f32(unsigned e) {
    unsigned u = (e & 0x80000000u) ? (e & 0x7fffffffu) : ~e;
    return __uint_as_float(u);
}

__device__ __forceinline__ long long floordiv2(long long v) { return v >= 0 ? v / 2 : -((-v + 1) / 2); }

// Python slice normalisation for image[a:b] on an axis of length len (a, b already int()-truncated)
__device__ __forceinline__ void py_slice(long long a, long long b, int len, int* start, int* size) {
    if (a < 0) { a += len; if (a < 0) a = 0; } else if (a > len) a = len;
    if (b < 0) { b += len; if (b < 0) b = 0; } else if (b > len) b = len;
    *start = (int)a;
    *size = b > a ? (int)(b - a) : 0;
}

__device__ void flip_rect(const long long r[4], int W, int H, int flag, long long o[4]) {
    long long p1x = r[0], p1y = r[1], p2x = r[0] + r[2], p2y = r[1] + r[3];
    if (flag == -1) { p1x = W - p1x - 1; p1y = H - p1y - 1; p2x = W - p2x - 1; p2y = H - p2y - 1; }
    else if (flag == 0) { p1y = H - p1y - 1; p2y = H - p2y - 1; }
    else if (flag == 1) { p1x = W - p1x - 1; p2x = W - p2x - 1; }
    long long x = p1x < p2x ? p1x : p2x, y = p1y < p2y ? p1y : p2y;
    o[2] = p2x - p1x < 0 ? p1x - p2x : p2x - p1x;
    o[3] = p2y - p1y < 0 ? p1y - p2y : p2y - p1y;
    o[0] = x < 0 ? 0 : x;
    o[1] = y < 0 ? 0 : y;
}

__global__ void aug_geometry_kernel(const float* __restrict__ label, const fcnod_aug_draw* __restrict__ draws, int N,
                                    int src_h, int src_w, int net_h, int net_w, AugGeom* __restrict__ geom,
                                    int32_t* __restrict__ rects_out) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= N) return;
    const fcnod_aug_draw d = draws[i];
    long long r0[4], r[4];
    for (int k = 0; k < 4; ++k) r0[k] = (long long)label[i * 5 + k];     // int() truncation
    if (d.flip > -2 && d.flip < 2) flip_rect(r0, src_w, src_h, d.flip, r);
    else for (int k = 0; k < 4; ++k) r[k] = r0[k];
    const long long rx = r[0], ry = r[1], rw = r[2], rh = r[3];
    // widths = (int(rw*f1), rw*f2)    heights = (int(rh*f1), rh*f2)                    (:168-169)
    const long long wid0 = (long long)((double)rw * d.f1), hei0 = (long long)((double)rh * d.f1);
    const double wid1 = (double)rw * d.f2, hei1 = (double)rh * d.f2;
    const long long x = (rx + floordiv2(rw)) - wid0, y = (ry + floordiv2(rh)) - hei0;   // (:194-195)
    const double w = wid1 + (double)wid0, h = hei1 + (double)hei0;                      // (:196-197)
    double cx = (double)rx + (double)rw / 2.0, cy = (double)ry + (double)rh / 2.0;      // (:200)
    cx = d.sign_x ? cx + (double)d.shift_x : cx - (double)d.shift_x;                    // (:202-203)
    cy = d.sign_y ? cy + (double)d.shift_y : cy - (double)d.shift_y;
    double nx = (double)(long long)(cx - w / 2.0), ny = (double)(long long)(cy - h / 2.0);   // int()   (:205-206)
    const long long nw = (long long)w, nh = (long long)h;                                      // (:207-208)
    if (nx > (double)x) nx = (double)x;                     // (:213-215)  nw -= |nx - x| == 0
    if (ny > (double)y) ny = (double)y;
    if (nx + (double)nw < (double)x + w) nx += ((double)x + w) - (nx + (double)nw);     // (:219-220)
    if (ny + (double)nh < (double)y + h) ny += ((double)y + h) - (ny + (double)nh);
    double X = nx < 0.0 ? 0.0 : nx, Y = ny < 0.0 ? 0.0 : ny;                            // (:228-229)
    double Wd = (double)nw, Hd = (double)nh;
    if (X > (double)src_w) Wd = (Wd - (Wd + X) - (double)src_w);                        // (:230-231)
    if (Y > (double)src_h) Hd = (Hd - (Hd + Y) - (double)src_h);
    AugGeom g;
    g.flip = d.flip;
    py_slice((long long)X, (long long)(X + Wd), src_w, &g.x0, &g.cw);                   // image[int(y):int(y+h), int(x):int(x+w)]
    py_slice((long long)Y, (long long)(Y + Hd), src_h, &g.y0, &g.ch);
    const long long c0 = (long long)((double)rx - X), c1 = (long long)((double)ry - Y); // new_rect (:234)
    // resize_image_and_labels (:123-136): fp32
    const float ratio_x = __fdiv_rn((float)g.cw, (float)net_w), ratio_y = __fdiv_rn((float)g.ch, (float)net_h);
    const float fx = (float)c0, fy = (float)c1, fw = (float)rw, fh = (float)rh;
    const float xt = __fdiv_rn(fx, ratio_x), yt = __fdiv_rn(fy, ratio_y);
    const float xb = __fdiv_rn(__fadd_rn(fx, fw), ratio_x), yb = __fdiv_rn(__fadd_rn(fy, fh), ratio_y);
    const bool ok = g.cw > 0 && g.ch > 0;
    g.rect[0] = ok ? (int)xt : 0; g.rect[1] = ok ? (int)yt : 0;
    g.rect[2] = ok ? (int)__fsub_rn(xb, xt) : 0; g.rect[3] = ok ? (int)__fsub_rn(yb, yt) : 0;
    g.label = (int)label[i * 5 + 4];
    g.mn_enc = 0xffffffffu; g.mx_enc = 0u;
    geom[i] = g;
    if (rects_out) for (int k = 0; k < 4; ++k) rects_out[i * 4 + k] = g.rect[k];
}

__device__ __forceinline__ float mean_of(int c) {
    return c == 0 ? 104.0069879317889f : (c == 1 ? 116.66876761696767f : 122.6789143406786f);
}

__device__ __forceinline__ long src_index(const AugGeom& g, int src_h, int src_w, int y, int x) {
    // pixel (y, x) of the crop -> index into one source channel plane, undoing cv.flip
    int fy = g.y0 + y, fx = g.x0 + x;
    if (g.flip == 0 || g.flip == -1) fy = src_h - 1 - fy;
    if (g.flip == 1 || g.flip == -1) fx = src_w - 1 - fx;
    return (long)fy * src_w + fx;
}

// Pixel (y, x), channel c of image n's crop.  U8 = false: straight from the float bottom `data_in` (CHW planes) through
// the flip / crop mapping; U8 = true: from the photometrically augmented crop (uint8 HWC, rows packed at the crop width,
// flip and crop already applied by aug_crop_u8_kernel).
template <bool U8>
__device__ __forceinline__ float crop_px(const void* __restrict__ base, const AugGeom& g, int src_h, int src_w, int n, int c, int y, int x) {
    if (U8) return (float)static_cast<const uint8_t*>(base)[(size_t)n * src_h * src_w * 3 + ((size_t)y * g.cw + x) * 3 + c];
    return static_cast<const float*>(base)[((size_t)n * 3 + c) * src_h * src_w + src_index(g, src_h, src_w, y, x)];
}

// flipped + cropped source -> uint8 HWC crop (the image `color_space_argumentation` sees, argumentation_engine.py:176), and
// its (h, w) for the photometric kernels.  Pixel values of the bottom are integers 0..255 held in floats.
__global__ void __launch_bounds__(256) aug_crop_u8_kernel(const float* __restrict__ data_in, int src_h, int src_w,
                                                          const AugGeom* __restrict__ geom, uint8_t* __restrict__ crop,
                                                          int32_t* __restrict__ sizes) {
    const int n = blockIdx.y;
    const AugGeom g = geom[n];
    if (blockIdx.x == 0 && threadIdx.x == 0) { sizes[2 * n] = g.ch; sizes[2 * n + 1] = g.cw; }
    const long npx = (long)g.cw * g.ch;
    uint8_t* o = crop + (size_t)n * src_h * src_w * 3;
    for (long i = blockIdx.x * (long)blockDim.x + threadIdx.x; i < npx; i += (long)gridDim.x * blockDim.x) {
        const int y = (int)(i / g.cw), x = (int)(i - (long)y * g.cw);
#pragma unroll
        for (int c = 0; c < 3; ++c) {
            const float v = crop_px<false>(data_in, g, src_h, src_w, n, c, y, x);
            o[i * 3 + c] = (uint8_t)fminf(fmaxf(v, 0.f), 255.f);
        }
    }
}

// grid (chunks, N): fp32 min / max of (pixel - mean_c) over the crop, all 3 channels
template <bool U8>
__global__ void __launch_bounds__(256) aug_minmax_kernel(const void* __restrict__ data_in, int src_h, int src_w,
                                                         AugGeom* __restrict__ geom) {
    const int n = blockIdx.y;
    const AugGeom g = geom[n];
    const long npx = (long)g.cw * g.ch;
    float mn = INFINITY, mx = -INFINITY;
    for (long i = blockIdx.x * (long)blockDim.x + threadIdx.x; i < npx; i += (long)gridDim.x * blockDim.x) {
        const int y = (int)(i / g.cw), x = (int)(i - (long)y * g.cw);
#pragma unroll
        for (int c = 0; c < 3; ++c) {
            const float v = __fsub_rn(crop_px<U8>(data_in, g, src_h, src_w, n, c, y, x), mean_of(c));
            mn = fminf(mn, v); mx = fmaxf(mx, v);
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        mn = fminf(mn, __shfl_xor_sync(0xffffffffu, mn, o));
        mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    }
    __shared__ float smn[8], smx[8];
    if ((threadIdx.x & 31) == 0) { smn[threadIdx.x >> 5] = mn; smx[threadIdx.x >> 5] = mx; }
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int w = 1; w < 8; ++w) { mn = fminf(mn, smn[w]); mx = fmaxf(mx, smx[w]); }
        if (npx > 0 && mn <= mx) {
            atomicMin(&geom[n].mn_enc, enc_f32(mn));
            atomicMax(&geom[n].mx_enc, enc_f32(mx));
        }
    }
}

// thread per VEC horizontally adjacent output pixels, all 3 channels; NCHW fp32 out
template <int VEC, bool U8>
__global__ void __launch_bounds__(256) aug_resample_kernel(const void* __restrict__ data_in, int src_h, int src_w,
                                                           const AugGeom* __restrict__ geom, int N, int net_h, int net_w,
                                                           float* __restrict__ top) {
    const int wv = net_w / VEC;
    const long total = (long)N * net_h * wv;
    for (long t = blockIdx.x * (long)blockDim.x + threadIdx.x; t < total; t += (long)gridDim.x * blockDim.x) {
        const int xv = (int)(t % wv);
        const long r = t / wv;
        const int oy = (int)(r % net_h), n = (int)(r / net_h);
        const AugGeom g = geom[n];
        float res[3][VEC];
        if (g.cw <= 0 || g.ch <= 0) {
#pragma unroll
            for (int c = 0; c < 3; ++c)
#pragma unroll
                for (int v = 0; v < VEC; ++v) res[c][v] = 0.f;
        } else {
            const float mn = dec_f32(g.mn_enc), mx = dec_f32(g.mx_enc);
            const float range = __fsub_rn(mx, mn);
            const bool same = (g.cw == net_w && g.ch == net_h);
            const double sy = (double)g.ch / (double)net_h, sx = (double)g.cw / (double)net_w;
            double fyd = ((double)oy + 0.5) * sy - 0.5;
            int y0 = (int)floor(fyd);
            float ay = (float)(fyd - (double)y0);
            if (y0 < 0) { y0 = 0; ay = 0.f; }
            if (y0 >= g.ch - 1) { y0 = g.ch - 1; ay = 0.f; }
            const int y1 = min(y0 + 1, g.ch - 1);
#pragma unroll
            for (int v = 0; v < VEC; ++v) {
                const int ox = xv * VEC + v;
                double fxd = ((double)ox + 0.5) * sx - 0.5;
                int x0 = (int)floor(fxd);
                float ax = (float)(fxd - (double)x0);
                if (x0 < 0) { x0 = 0; ax = 0.f; }
                if (x0 >= g.cw - 1) { x0 = g.cw - 1; ax = 0.f; }
                const int x1 = min(x0 + 1, g.cw - 1);
#pragma unroll
                for (int c = 0; c < 3; ++c) {
                    const float m = mean_of(c);
                    // (im - min) / (max - min) in fp32 (:302), then bilinear (:120).  The normalisation is affine and the
                    // bilinear weights sum to one, so it commutes with the interpolation: interpolate the raw pixels, normalise
                    // ONCE (one IEEE divide per output instead of four; the image top's gate is 1e-5, not bit-exactness).
                    const float v00 = crop_px<U8>(data_in, g, src_h, src_w, n, c, y0, x0);
                    if (same) { res[c][v] = __fdiv_rn(__fsub_rn(__fsub_rn(v00, m), mn), range); continue; }
                    const float v01 = crop_px<U8>(data_in, g, src_h, src_w, n, c, y0, x1);
                    const float v10 = crop_px<U8>(data_in, g, src_h, src_w, n, c, y1, x0);
                    const float v11 = crop_px<U8>(data_in, g, src_h, src_w, n, c, y1, x1);
                    const float tp = v00 * (1.f - ax) + v01 * ax, bt = v10 * (1.f - ax) + v11 * ax;
                    res[c][v] = __fdiv_rn(__fsub_rn(__fsub_rn(tp * (1.f - ay) + bt * ay, m), mn), range);
                }
            }
        }
#pragma unroll
        for (int c = 0; c < 3; ++c) {
            float* o = top + (((size_t)n * 3 + c) * net_h + oy) * net_w + xv * VEC;
            if (VEC == 4) *reinterpret_cast<float4*>(o) = make_float4(res[c][0], res[c][1], res[c][2], res[c][3]);
            else
#pragma unroll
                for (int v = 0; v < VEC; ++v) o[v] = res[c][v];
        }
    }
}

// JaccardCoeff.iou(cell, rect) > 0.1  (argumentation_engine.py:26-55,276)
__device__ __forceinline__ bool cell_labelled(int cx, int cy, int s, int bx, int by, int bw, int bh) {
    const double ax = (double)cx, ay = (double)cy, as = (double)s;
    const double x = fmax(ax, (double)bx), y = fmax(ay, (double)by);
    const double w = fmin(ax + as, (double)bx + (double)bw) - x, h = fmin(ay + as, (double)by + (double)bh) - y;
    if (w < 0.0 || h < 0.0) return false;
    const double ux = fmin(ax, (double)bx), uy = fmin(ay, (double)by);
    const double uw = fmax(ax + as, (double)bx + (double)bw) - ux, uh = fmax(ay + as, (double)by + (double)bh) - uy;
    const float aub = (float)(uw * uh), anb = (float)(w * h);
    const float ratio = __fdiv_rn((float)(as * as), (float)((long long)bw * (long long)bh));
    const float score = __fdiv_rn(__fdiv_rn(anb, aub), ratio);
    return (double)score > 0.1;        // np.float32 vs Python float: compared in double (A10)
}

struct LabelVals { float b0, b1, b2, b3, s0, s1, obj; };
__device__ __forceinline__ LabelVals label_vals(int cx, int cy, int s, int bx, int by, int bw, int bh) {
    LabelVals v;
    v.b0 = (float)((double)bx - (double)cx); v.b1 = (float)((double)by - (double)cy);
    v.b2 = (float)(((double)bx + (double)bw) - (double)cx); v.b3 = (float)(((double)by + (double)bh) - (double)cy);
    v.s0 = (float)(1.0 / (double)bw); v.s1 = (float)(1.0 / (double)bh);       // float64 divide, fp32 blob cast (:96-99)
    v.obj = __fdiv_rn((float)(s * s), (float)((long long)bw * (long long)bh));   // fp32 divide (:101-102)
    return v;
}

// One thread per V horizontally adjacent grid cells (V = 4: every store is 128 bits; V = 1 when the grid width or the
// pointers do not allow it); classes in the outer loop, the rects of a class applied in order (later rects overwrite,
// like the python loop :81-107 -- rects of different classes touch different channels).  Write-only and HBM-bound:
// (C + 16 C) x gh x gw floats per image, each byte written once (zeros included).
template <int V>
__global__ void __launch_bounds__(256) label_raster_kernel(const int32_t* __restrict__ rects, const int32_t* __restrict__ labels,
                                                           const int32_t* __restrict__ nrect, const AugGeom* __restrict__ geom,
                                                           int N, int R, int gh, int gw, int stride, int C,
                                                           float* __restrict__ fg, float* __restrict__ bbox,
                                                           float* __restrict__ size, float* __restrict__ obj,
                                                           float* __restrict__ cvg) {
    const int plane = gh * gw, pv = plane / V;
    const long total = (long)N * pv;
    for (long t = blockIdx.x * (long)blockDim.x + threadIdx.x; t < total; t += (long)gridDim.x * blockDim.x) {
        const int n = (int)(t / pv), cell = V * (int)(t - (long)n * pv);
        const int j = cell / gw, i0 = cell - j * gw;
        const int cy = j * stride;
        const int nr = geom ? 1 : (nrect ? min(nrect[n], R) : R);
        for (int c = 0; c < C; ++c) {
            float vb[4][V], vs[2][V], vo[V], vc[V];
#pragma unroll
            for (int v = 0; v < V; ++v) { vb[0][v] = vb[1][v] = vb[2][v] = vb[3][v] = 0.f; vs[0][v] = vs[1][v] = 0.f; vo[v] = 0.f; vc[v] = 0.f; }
            for (int q = 0; q < nr; ++q) {
                int bx, by, bw, bh, lab;
                if (geom) { bx = geom[n].rect[0]; by = geom[n].rect[1]; bw = geom[n].rect[2]; bh = geom[n].rect[3]; lab = geom[n].label; }
                else { const int32_t* r = rects + ((size_t)n * R + q) * 4; bx = r[0]; by = r[1]; bw = r[2]; bh = r[3]; lab = labels[(size_t)n * R + q]; }
                if (lab != c) continue;              // labels outside [0, C): python would raise IndexError; here the rect is ignored
#pragma unroll
                for (int v = 0; v < V; ++v) {
                    const int cx = (i0 + v) * stride;
                    if (!cell_labelled(cx, cy, stride, bx, by, bw, bh)) continue;
                    const LabelVals L = label_vals(cx, cy, stride, bx, by, bw, bh);
                    vb[0][v] = L.b0; vb[1][v] = L.b1; vb[2][v] = L.b2; vb[3][v] = L.b3;
                    vs[0][v] = L.s0; vs[1][v] = L.s1; vo[v] = L.obj; vc[v] = 1.f;
                }
            }
            auto put = [&](float* base, const float (&x)[V]) {
                if (V == 4) *reinterpret_cast<float4*>(base) = make_float4(x[0], x[1], x[2], x[3]);
                else base[0] = x[0];
            };
            const size_t o1 = ((size_t)n * C + c) * plane + cell, o4 = ((size_t)n * 4 * C + 4 * c) * plane + cell;
            put(fg + o1, vc);
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                put(bbox + o4 + (size_t)k * plane, vb[k]);
                put(size + o4 + (size_t)k * plane, vs[k & 1]);
                put(obj + o4 + (size_t)k * plane, vo);
                put(cvg + o4 + (size_t)k * plane, vc);
            }
        }
    }
}

int grid_1d(long items, int threads) {
    long b = (items + threads - 1) / threads;
    long cap = 148L * 16;
    if (b > cap) b = cap;
    if (b < 1) b = 1;
    return (int)b;
}

void launch_label_raster(const int32_t* rects, const int32_t* labels, const int32_t* nrect, const AugGeom* geom, int N, int R,
                         int gh, int gw, int stride, int C, float* fg, float* bbox, float* size, float* obj, float* cvg,
                         cudaStream_t s) {
    const uintptr_t al = reinterpret_cast<uintptr_t>(fg) | reinterpret_cast<uintptr_t>(bbox) | reinterpret_cast<uintptr_t>(size) |
                         reinterpret_cast<uintptr_t>(obj) | reinterpret_cast<uintptr_t>(cvg);
    if (gw % 4 == 0 && al % 16 == 0)
        label_raster_kernel<4><<<grid_1d((long)N * gh * gw / 4, 256), 256, 0, s>>>(rects, labels, nrect, geom, N, R, gh, gw, stride, C,
                                                                                    fg, bbox, size, obj, cvg);
    else
        label_raster_kernel<1><<<grid_1d((long)N * gh * gw, 256), 256, 0, s>>>(rects, labels, nrect, geom, N, R, gh, gw, stride, C,
                                                                                fg, bbox, size, obj, cvg);
    FCNOD_CUDA(cudaGetLastError());
}

void check_layer_dims(int N, int net_h, int net_w, int stride, int C) {
    FCNOD_CHECK(N > 0 && net_h > 0 && net_w > 0 && stride > 0 && C > 0, "layer: bad dimensions");
    FCNOD_CHECK(net_h / stride > 0 && net_w / stride > 0, "layer: stride larger than the image");
}

}  // namespace
}  // namespace fcnod

using namespace fcnod;

extern "C" {

int fcnod_label_raster_device(const int32_t* rects, const int32_t* labels, const int32_t* nrect, int N, int R, int net_h,
                              int net_w, int stride, int num_classes, float* fg, float* bbox, float* size, float* obj,
                              float* cvg, void* stream) {
    FCNOD_API_BEGIN
    FCNOD_CHECK(rects && labels && fg && bbox && size && obj && cvg, "null argument");
    check_layer_dims(N, net_h, net_w, stride, num_classes);
    FCNOD_CHECK(R > 0, "R must be positive");
    const int gh = net_h / stride, gw = net_w / stride;      // Python-2 integer division (:284, layer :60-61)
    launch_label_raster(rects, labels, nrect, (const AugGeom*)0, N, R, gh, gw, stride, num_classes, fg, bbox, size, obj, cvg,
                        (cudaStream_t)stream);
    FCNOD_API_END
}

int fcnod_label_raster(const int32_t* rects, const int32_t* labels, const int32_t* nrect, int N, int R, int net_h, int net_w,
                       int stride, int num_classes, float* fg, float* bbox, float* size, float* obj, float* cvg) {
    FCNOD_API_BEGIN
    FCNOD_CHECK(rects && labels && fg && bbox && size && obj && cvg, "null argument");
    check_layer_dims(N, net_h, net_w, stride, num_classes);
    FCNOD_CHECK(R > 0, "R must be positive");
    const int gh = net_h / stride, gw = net_w / stride;
    const size_t plane = (size_t)gh * gw, nfg = (size_t)N * num_classes * plane, n4 = nfg * 4;
    Workspace& ws = workspace();
    cudaStream_t s = ws.stream();
    ws.reserve(Workspace::padded((size_t)N * R * 16) + 2 * Workspace::padded((size_t)N * R * 4) + Workspace::padded((nfg + 4 * n4) * 4));
    int32_t* d_r = ws.take<int32_t>((size_t)N * R * 4);
    int32_t* d_l = ws.take<int32_t>((size_t)N * R);
    int32_t* d_n = nrect ? ws.take<int32_t>((size_t)N) : 0;
    float* d_out = ws.take<float>(nfg + 4 * n4);
    FCNOD_CUDA(cudaMemcpyAsync(d_r, rects, (size_t)N * R * 4 * 4, cudaMemcpyHostToDevice, s));
    FCNOD_CUDA(cudaMemcpyAsync(d_l, labels, (size_t)N * R * 4, cudaMemcpyHostToDevice, s));
    if (nrect) FCNOD_CUDA(cudaMemcpyAsync(d_n, nrect, (size_t)N * 4, cudaMemcpyHostToDevice, s));
    float *dfg = d_out, *db = dfg + nfg, *ds = db + n4, *dobj = ds + n4, *dc = dobj + n4;
    launch_label_raster(d_r, d_l, d_n, (const AugGeom*)0, N, R, gh, gw, stride, num_classes, dfg, db, ds, dobj, dc, s);
    FCNOD_CUDA(cudaMemcpyAsync(fg, dfg, nfg * 4, cudaMemcpyDeviceToHost, s));
    FCNOD_CUDA(cudaMemcpyAsync(bbox, db, n4 * 4, cudaMemcpyDeviceToHost, s));
    FCNOD_CUDA(cudaMemcpyAsync(size, ds, n4 * 4, cudaMemcpyDeviceToHost, s));
    FCNOD_CUDA(cudaMemcpyAsync(obj, dobj, n4 * 4, cudaMemcpyDeviceToHost, s));
    FCNOD_CUDA(cudaMemcpyAsync(cvg, dc, n4 * 4, cudaMemcpyDeviceToHost, s));
    FCNOD_CUDA(cudaStreamSynchronize(s));
    FCNOD_API_END
}

size_t fcnod_aug_scratch_bytes(int N) { return (size_t)(N > 0 ? N : 0) * sizeof(AugGeom); }

size_t fcnod_aug_photo_scratch_bytes(int N, int src_h, int src_w) {
    if (N <= 0 || src_h <= 0 || src_w <= 0) return 0;
    return 3 * align_up((size_t)N * src_h * src_w * 3, 256) + align_up((size_t)N * 8, 256);
}

int fcnod_aug_forward_photo_device(const float* data_in, const float* label, const fcnod_aug_draw* draws,
                                   const fcnod_photo_draw* photo, int N, int src_h, int src_w, int net_h, int net_w, int stride,
                                   int num_classes, float* top_data, float* fg, float* bbox, float* size, float* obj, float* cvg,
                                   int32_t* rects_out, void* scratch, void* photo_scratch, void* stream) {
    FCNOD_API_BEGIN
    FCNOD_CHECK(data_in && label && draws && top_data && fg && bbox && size && obj && cvg && scratch, "null argument");
    FCNOD_CHECK(!photo || photo_scratch, "photo draws need photo_scratch (fcnod_aug_photo_scratch_bytes)");
    check_layer_dims(N, net_h, net_w, stride, num_classes);
    FCNOD_CHECK(src_h > 0 && src_w > 0, "layer: bad source size");
    cudaStream_t s = (cudaStream_t)stream;
    AugGeom* geom = static_cast<AugGeom*>(scratch);
    const int gh = net_h / stride, gw = net_w / stride;
    aug_geometry_kernel<<<ceil_div(N, 128), 128, 0, s>>>(label, draws, N, src_h, src_w, net_h, net_w, geom, rects_out);
    FCNOD_CUDA(cudaGetLastError());
    const long np4 = (long)N * net_h * (net_w / 4), np1 = (long)N * net_h * net_w;
    if (photo) {
        // T4 (argumentation_engine.py:176): the flipped, cropped uint8 image goes through the imgaug chain before demeaning
        const size_t img_bytes = align_up((size_t)N * src_h * src_w * 3, 256);
        uint8_t* crop = static_cast<uint8_t*>(photo_scratch);
        uint8_t* aug = crop + img_bytes;
        uint8_t* tmp = aug + img_bytes;
        int32_t* sizes = reinterpret_cast<int32_t*>(tmp + img_bytes);
        aug_crop_u8_kernel<<<dim3(32, N), 256, 0, s>>>(data_in, src_h, src_w, geom, crop, sizes);
        FCNOD_CUDA(cudaGetLastError());
        photo_launch(crop, aug, tmp, sizes, photo, N, src_h, src_w, s);
        aug_minmax_kernel<true><<<dim3(32, N), 256, 0, s>>>(aug, src_h, src_w, geom);
        FCNOD_CUDA(cudaGetLastError());
        if (net_w % 4 == 0) aug_resample_kernel<4, true><<<grid_1d(np4, 256), 256, 0, s>>>(aug, src_h, src_w, geom, N, net_h, net_w, top_data);
        else aug_resample_kernel<1, true><<<grid_1d(np1, 256), 256, 0, s>>>(aug, src_h, src_w, geom, N, net_h, net_w, top_data);
    } else {
        aug_minmax_kernel<false><<<dim3(32, N), 256, 0, s>>>(data_in, src_h, src_w, geom);
        FCNOD_CUDA(cudaGetLastError());
        if (net_w % 4 == 0) aug_resample_kernel<4, false><<<grid_1d(np4, 256), 256, 0, s>>>(data_in, src_h, src_w, geom, N, net_h, net_w, top_data);
        else aug_resample_kernel<1, false><<<grid_1d(np1, 256), 256, 0, s>>>(data_in, src_h, src_w, geom, N, net_h, net_w, top_data);
    }
    FCNOD_CUDA(cudaGetLastError());
    launch_label_raster((const int32_t*)0, (const int32_t*)0, (const int32_t*)0, geom, N, 1, gh, gw, stride, num_classes, fg, bbox,
                        size, obj, cvg, s);
    FCNOD_API_END
}

int fcnod_aug_forward_device(const float* data_in, const float* label, const fcnod_aug_draw* draws, int N, int src_h,
                             int src_w, int net_h, int net_w, int stride, int num_classes, float* top_data, float* fg,
                             float* bbox, float* size, float* obj, float* cvg, int32_t* rects_out, void* scratch,
                             void* stream) {
    return fcnod_aug_forward_photo_device(data_in, label, draws, 0, N, src_h, src_w, net_h, net_w, stride, num_classes, top_data, fg,
                                          bbox, size, obj, cvg, rects_out, scratch, 0, stream);
}

int fcnod_aug_forward(const float* data_in, const float* label, const fcnod_aug_draw* draws, int N, int src_h, int src_w,
                      int net_h, int net_w, int stride, int num_classes, float* top_data, float* fg, float* bbox, float* size,
                      float* obj, float* cvg, int32_t* rects_out) {
    return fcnod_aug_forward_photo(data_in, label, draws, 0, N, src_h, src_w, net_h, net_w, stride, num_classes, top_data, fg, bbox,
                                   size, obj, cvg, rects_out);
}

int fcnod_aug_forward_photo(const float* data_in, const float* label, const fcnod_aug_draw* draws, const fcnod_photo_draw* photo,
                            int N, int src_h, int src_w, int net_h, int net_w, int stride, int num_classes, float* top_data,
                            float* fg, float* bbox, float* size, float* obj, float* cvg, int32_t* rects_out) {
    FCNOD_API_BEGIN
    FCNOD_CHECK(data_in && label && draws && top_data && fg && bbox && size && obj && cvg, "null argument");
    check_layer_dims(N, net_h, net_w, stride, num_classes);
    const int gh = net_h / stride, gw = net_w / stride;
    const size_t nin = (size_t)N * 3 * src_h * src_w, ntop = (size_t)N * 3 * net_h * net_w;
    const size_t plane = (size_t)gh * gw, nfg = (size_t)N * num_classes * plane, n4 = nfg * 4;
    Workspace& ws = workspace();
    cudaStream_t s = ws.stream();
    ws.reserve(Workspace::padded(nin * 4) + Workspace::padded((size_t)N * 5 * 4) + Workspace::padded((size_t)N * sizeof(fcnod_aug_draw)) +
               Workspace::padded((size_t)N * 16) + Workspace::padded(fcnod_aug_scratch_bytes(N)) +
               Workspace::padded((ntop + nfg + 4 * n4) * 4) +
               (photo ? Workspace::padded((size_t)N * sizeof(fcnod_photo_draw)) + Workspace::padded(fcnod_aug_photo_scratch_bytes(N, src_h, src_w)) : 0));
    float* d_in = ws.take<float>(nin);
    float* d_label = ws.take<float>((size_t)N * 5);
    fcnod_aug_draw* d_draw = ws.take<fcnod_aug_draw>((size_t)N);
    int32_t* d_rect = ws.take<int32_t>((size_t)N * 4);
    void* d_scr = ws.take<char>(fcnod_aug_scratch_bytes(N));
    float* d_out = ws.take<float>(ntop + nfg + 4 * n4);
    fcnod_photo_draw* d_photo = photo ? ws.take<fcnod_photo_draw>((size_t)N) : 0;
    void* d_pscr = photo ? ws.take<char>(fcnod_aug_photo_scratch_bytes(N, src_h, src_w)) : 0;
    if (photo) FCNOD_CUDA(cudaMemcpyAsync(d_photo, photo, (size_t)N * sizeof(fcnod_photo_draw), cudaMemcpyHostToDevice, s));
    FCNOD_CUDA(cudaMemcpyAsync(d_in, data_in, nin * 4, cudaMemcpyHostToDevice, s));
    FCNOD_CUDA(cudaMemcpyAsync(d_label, label, (size_t)N * 5 * 4, cudaMemcpyHostToDevice, s));
    FCNOD_CUDA(cudaMemcpyAsync(d_draw, draws, (size_t)N * sizeof(fcnod_aug_draw), cudaMemcpyHostToDevice, s));
    float *dtop = d_out, *dfg = dtop + ntop, *db = dfg + nfg, *ds = db + n4, *dobj = ds + n4, *dc = dobj + n4;
    int rc = fcnod_aug_forward_photo_device(d_in, d_label, d_draw, d_photo, N, src_h, src_w, net_h, net_w, stride, num_classes, dtop,
                                            dfg, db, ds, dobj, dc, d_rect, d_scr, d_pscr, s);
    FCNOD_CHECK(rc == 0, fcnod_last_error());
    FCNOD_CUDA(cudaMemcpyAsync(top_data, dtop, ntop * 4, cudaMemcpyDeviceToHost, s));
    FCNOD_CUDA(cudaMemcpyAsync(fg, dfg, nfg * 4, cudaMemcpyDeviceToHost, s));
    FCNOD_CUDA(cudaMemcpyAsync(bbox, db, n4 * 4, cudaMemcpyDeviceToHost, s));
    FCNOD_CUDA(cudaMemcpyAsync(size, ds, n4 * 4, cudaMemcpyDeviceToHost, s));
    FCNOD_CUDA(cudaMemcpyAsync(obj, dobj, n4 * 4, cudaMemcpyDeviceToHost, s));
    FCNOD_CUDA(cudaMemcpyAsync(cvg, dc, n4 * 4, cudaMemcpyDeviceToHost, s));
    if (rects_out) FCNOD_CUDA(cudaMemcpyAsync(rects_out, d_rect, (size_t)N * 4 * 4, cudaMemcpyDeviceToHost, s));
    FCNOD_CUDA(cudaStreamSynchronize(s));
    FCNOD_API_END
}

}  // extern "C"

// ------------------------------------------------------------------------------------------ allocation-free variant
// For callers that may not allocate device memory (stream capture, serving loops, harnesses that forbid cudaMalloc): the
// layer forward with its per-image geometry records (the `scratch` argument of fcnod_aug_forward_device) in a fixed
// array that lives in the library's image.  No counterpart in the reference.
//   * every GPU that loads the library holds FCNOD_AUG_NOALLOC_MAX_N * sizeof(AugGeom) = 4096 * 48 B = 192 KiB until exit;
//   * all calls on one device share the array: calls that overlap on the GPU (different streams, different threads) race;
//   * the array keeps the last call's records; aug_geometry_kernel rewrites every field of records [0, N) at the start of
//     each call, the min / max accumulators of aug_minmax_kernel included, so nothing has to be cleared between calls.
// In host code `g_aug_noalloc_scratch` is only a handle for the symbol, not a device address: the address is looked up
// once per device with cudaGetSymbolAddress (no allocation) and that pointer is what the kernels receive.
#include <mutex>

namespace {

constexpr int kAugNoallocMaxN = 4096;       // == FCNOD_AUG_NOALLOC_MAX_N (include/fcnod_b200.h)
constexpr int kAugNoallocMaxDevices = 64;

static __device__ __align__(256) unsigned char g_aug_noalloc_scratch[(size_t)kAugNoallocMaxN * sizeof(AugGeom)];

void* aug_noalloc_scratch_ptr() {
    static std::mutex mu;
    static void* addr[kAugNoallocMaxDevices] = {};
    int dev = -1;
    FCNOD_CUDA(cudaGetDevice(&dev));
    FCNOD_CHECK(dev >= 0 && dev < kAugNoallocMaxDevices, "aug noalloc: device ordinal out of range");
    std::lock_guard<std::mutex> lock(mu);
    if (!addr[dev]) {
        void* p = 0;
        FCNOD_CUDA(cudaGetSymbolAddress(&p, g_aug_noalloc_scratch));      // this device's copy: the device is current here
        FCNOD_CHECK(p != 0, "aug noalloc: cudaGetSymbolAddress returned a null address");
        addr[dev] = p;
    }
    return addr[dev];
}

}  // namespace

extern "C" {

int fcnod_aug_forward_device_noalloc(const float* data_in, const float* label, const fcnod_aug_draw* draws, int N, int src_h,
                                     int src_w, int net_h, int net_w, int stride, int num_classes, float* top_data, float* fg,
                                     float* bbox, float* size, float* obj, float* cvg, int32_t* rects_out, void* stream) {
    FCNOD_API_BEGIN
    static_assert(kAugNoallocMaxN == FCNOD_AUG_NOALLOC_MAX_N, "header and kernel disagree on the limit");
    if (N > kAugNoallocMaxN) {
        char msg[200];
        snprintf(msg, sizeof msg, "fcnod_aug_forward_device_noalloc: N = %d images exceeds FCNOD_AUG_NOALLOC_MAX_N = %d "
                 "(the fixed scratch array); split the batch or use fcnod_aug_forward_device with caller scratch", N, kAugNoallocMaxN);
        FCNOD_CHECK(false, msg);
    }
    void* scratch = aug_noalloc_scratch_ptr();
    int rc = fcnod_aug_forward_photo_device(data_in, label, draws, 0, N, src_h, src_w, net_h, net_w, stride, num_classes, top_data, fg,
                                            bbox, size, obj, cvg, rects_out, scratch, 0, stream);
    if (rc != 0) return rc;      // last-error string already set by the callee
    FCNOD_API_END
}

}  // extern "C"
