// Post-net stage on the device: coverage threshold -> grid-box decode -> cv::groupRectangles
// clustering -> height filter / confidence -> rescale to source pixels.
//
// Restates, for B frames at once, the per-frame Python of
// /root/reference/scripts/fcn_object_detector.py:
//   per-class loop            :89-90,104-118      (skip coverage channel 0, bbox channels 4*index..)
//   gridbox_to_boxes          :357-394            (row-major candidate order, float32 + int -> float64)
//   vote_boxes                :337-351            (cv.groupRectangles on [x1,y1,x2,y2] "as if" x,y,w,h)
//   resize_detection          :396-405            (double scale, truncation toward zero)
// and OpenCV's groupRectangles (objdetect/cascadedetect.cpp: SimilarRects + partition +
// fp32 mean with round-half-even + containment filter).  Integer / double arithmetic only:
// results are bit-exact against the oracle (tests/test_postnet_gpu.py).
//
// Kernel 1 (decode_group_kernel): one CTA per (frame, foreground class).  Coverage is read
//   with coalesced loads, candidates are compacted IN ROW-MAJOR ORDER with warp ballots +
//   a cross-warp scan, kept in shared memory, clustered there: every unordered SimilarRects pair
//   is tested once and similar pairs are merged at once in a lock-free union-find (larger root
//   hooked under the smaller) = connected components, class order = smallest member index,
//   exactly what cv::partition yields; reduced with shared-memory atomics and filtered.
// Kernel 2 (assemble_kernel): one warp per frame concatenates the per-class lists in
//   class order and rescales.
#include <cmath>
#include <cstring>
#include <map>
#include <mutex>
#include <vector>

#include "common.cuh"
#include "kernels.h"
#include "workspace.h"

namespace fcnod {
namespace {

constexpr int kMaxCand = 1024;     // grid cells per (frame, class) handled in shared memory
constexpr int kThreads = 1024;

struct PostParams {
    float prob_thresh;
    int group_threshold;
    double eps;
    int cell_w, cell_h;
    int rows_used;
    int skip_bg;
    int min_height;
    int net_w, net_h;
};

// exclusive, order-preserving block scan of a 0/1 flag; returns this thread's offset and the block total
__device__ __forceinline__ int block_scan_flag(bool flag, int* warp_tot /*[8]*/, int* total) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    unsigned m = __ballot_sync(0xffffffffu, flag);
    int pre = __popc(m & ((1u << lane) - 1u));
    if (lane == 0) warp_tot[warp] = __popc(m);
    __syncthreads();
    int base = 0, tot = 0;
#pragma unroll
    for (int w = 0; w < kThreads / 32; ++w) {
        int t = warp_tot[w];
        if (w < warp) base += t;
        tot += t;
    }
    __syncthreads();
    *total = tot;
    return base + pre;
}

// Lock-free union-find over shared memory.  Invariant: lab[x] <= x, and only roots (lab[x] == x) are ever hooked, so
// the structure is a forest whose roots are the smallest members of their trees.
__device__ __forceinline__ int uf_find(volatile int* lab, int x) {
    const int x0 = x;
    int p = lab[x];
    if (p == x) return x;
    do { x = p; p = lab[x]; } while (p != x);
    // compress the starting node with a plain store: x is an ancestor of x0 whatever else happened meanwhile, so a lost
    // race only leaves a longer (still valid) path
    lab[x0] = x;
    return x;
}
constexpr int kDeltaTable = 8192;   // T(m) for 0 <= m < 8192 (sum of two box sides); larger m take the FP64 path

// T(m) = floor(eps * m * 0.5), evaluated in double exactly as cv::SimilarRects does
__global__ void delta_table_kernel(double eps, int* __restrict__ tab) {
    const int m = blockIdx.x * blockDim.x + threadIdx.x;
    if (m < kDeltaTable) tab[m] = (int)floor(eps * (double)m * 0.5);
}

__global__ void __launch_bounds__(kThreads) decode_group_kernel(const float* __restrict__ coverage,
                                                               const float* __restrict__ bboxes, int c_cvg, int c_bbox,
                                                               int gh, int gw, PostParams P,
                                                               const int* __restrict__ delta_tab,
                                                               int32_t* __restrict__ cls_det /*[B][ncls][MAX_DET][6]*/,
                                                               int32_t* __restrict__ cls_cnt /*[B][ncls]*/) {
    __shared__ int rx[kMaxCand], ry[kMaxCand], rw[kMaxCand], rh[kMaxCand];
    __shared__ int lab[kMaxCand];
    __shared__ int warp_tot[kThreads / 32];
    __shared__ int s_any;
    // dynamic shared memory: per-class sums / counts / containment margins / surviving-class list, 8 arrays of gh*gw ints
    extern __shared__ unsigned adj[];
    const int ncell_all = gh * gw;
    int* const sx = reinterpret_cast<int*>(adj);
    int* const sy = sx + ncell_all;
    int* const sw = sy + ncell_all;
    int* const sh = sw + ncell_all;
    int* const cnt = sh + ncell_all;
    int* const dxr = cnt + ncell_all;
    int* const dyr = dxr + ncell_all;
    int* const strong = dyr + ncell_all;

    const int cls = blockIdx.x, b = blockIdx.y, ncls = gridDim.x;
    const int cch = cls + (P.skip_bg ? 1 : 0);
    const float* cv = coverage + ((size_t)b * c_cvg + cch) * gh * gw;
    const float* bb = bboxes + ((size_t)b * c_bbox + 4 * cls) * gh * gw;
    const int plane = gh * gw;
    const int ncell = P.rows_used * gw;
    int32_t* out = cls_det + ((size_t)b * ncls + cls) * FCNOD_MAX_DET * 6;

    if (threadIdx.x == 0) s_any = 0;
    __syncthreads();

    // ---- 1. threshold + ordered compaction + decode (float64) + round-half-even to int
    int n = 0;
    for (int base = 0; base < ncell; base += kThreads) {
        const int i = base + threadIdx.x;
        bool f = false;
        if (i < ncell) f = cv[i] >= P.prob_thresh;
        int tot;
        const int off = block_scan_flag(f, warp_tot, &tot);
        if (f) {
            const int y = i / gw, x = i - y * gw;
            const double mx = (double)(x * P.cell_w), my = (double)(y * P.cell_h);
            const double x1 = (double)bb[i] + mx, y1 = (double)bb[plane + i] + my;
            const double x2 = (double)bb[2 * plane + i] + mx, y2 = (double)bb[3 * plane + i] + my;
            if (x1 != 0.0 || y1 != 0.0 || x2 != 0.0 || y2 != 0.0) s_any = 1;   // `propose_boxes.any()` (:339)
            const int k = n + off;
            rx[k] = __double2int_rn(x1); ry[k] = __double2int_rn(y1);
            rw[k] = __double2int_rn(x2); rh[k] = __double2int_rn(y2);
        }
        n += tot;
    }
    __syncthreads();
    if (n == 0 || !s_any) {
        if (threadIdx.x == 0) cls_cnt[b * ncls + cls] = 0;
        return;
    }

    // ---- 2. cv::partition with SimilarRects(eps): connected components, label = smallest member index
    for (int i = threadIdx.x; i < n; i += kThreads) lab[i] = i;
    if (threadIdx.x == 0) s_any = 1;     // re-used: 1 = every coordinate is small enough for 32-bit arithmetic
    __syncthreads();
    if (P.group_threshold > 0) {
        // Every unordered pair (i, j < i) is tested ONCE and, when similar, merged straight
        // away in a lock-free union-find over `lab` (hook the larger root under the smaller one with atomicCAS): no
        // similarity matrix, no propagation sweeps.  Links always point to a smaller index, so the root of a component
        // is its smallest member -- the class order cv::partition produces -- whatever order the merges happen in.
        // OpenCV: delta = eps * (min(w) + min(h)) * 0.5 in double; similar iff all four |d| <= delta.  The |d| are
        // integers, so the test is |d| <= floor(delta) =: T(m), m = min(w) + min(h); T comes from a table built with that
        // very double expression (delta_table_kernel), FP64 only outside the table (FP64 issue is scarce on this part).
        // A WARP takes rows r and n-1-r (r + (n-1-r) partners: balanced) and walks each row 32 partners at a time: lane = j,
        // control flow is warp-uniform, the similar lanes are found with one ballot and merged together.
        for (int i = threadIdx.x; i < n; i += kThreads) {
            const int lim = 1 << 28;
            if (abs(rx[i]) >= lim || abs(ry[i]) >= lim || abs(rw[i]) >= lim || abs(rh[i]) >= lim) s_any = 0;
        }
        __syncthreads();
        const bool small = s_any != 0;
        const bool pre = P.eps > 0.0;
        volatile int* vlab = lab;
        const int lane = threadIdx.x & 31;
        for (int r = threadIdx.x >> 5; 2 * r < n; r += kThreads / 32) {      // warp-uniform
            for (int pass = 0; pass < 2; ++pass) {
                const int i = pass ? n - 1 - r : r;
                if (pass && i == r) break;
                const int xi = rx[i], yi = ry[i], wi = rw[i], hi = rh[i];
                // T is monotone in m and m <= wi + hi: |d| > T(wi + hi) rejects most pairs before the table look-up
                const int mi = wi + hi;
                int Tmax = 0;
                if (small) {
                    Tmax = mi < 0 ? -1 : (mi < kDeltaTable ? __ldg(delta_tab + mi) : (int)floor(P.eps * (double)mi * 0.5));
                    if (pre && Tmax < 0) continue;
                }
                for (int j0 = 0; j0 < i; j0 += 32) {
                    const int j = j0 + lane;
                    const bool valid = j < i;
                    bool sim = false;
                    if (small) {
                        int xj = 0, yj = 0, d0 = 0, d1 = 0;
                        bool cand = false;
                        if (valid) {
                            xj = rx[j]; yj = ry[j];
                            d0 = abs(xi - xj); d1 = abs(yi - yj);
                            cand = !pre || (d0 <= Tmax && d1 <= Tmax);
                        }
                        if (!__any_sync(0xffffffffu, cand)) continue;
                        if (cand) {
                            const int wj = rw[j], hj = rh[j];
                            const int m = min(wi, wj) + min(hi, hj);
                            int T;
                            if (m < 0) T = -1;
                            else if (m < kDeltaTable) T = __ldg(delta_tab + m);
                            else T = (int)floor(P.eps * (double)m * 0.5);
                            const int d2 = abs(xi + wi - xj - wj), d3 = abs(yi + hi - yj - hj);
                            sim = d0 <= T && d1 <= T && d2 <= T && d3 <= T;
                        }
                    } else if (valid) {
                        const int xj = rx[j], yj = ry[j], wj = rw[j], hj = rh[j];
                        const long long m = (long long)min(wi, wj) + (long long)min(hi, hj);
                        long long T;
                        if (m < 0) T = -1;
                        else if (m < kDeltaTable) T = __ldg(delta_tab + (int)m);
                        else T = (long long)floor(P.eps * (double)m * 0.5);
                        const long long d0 = llabs((long long)xi - xj), d1 = llabs((long long)yi - yj);
                        const long long d2 = llabs((long long)xi + wi - xj - wj), d3 = llabs((long long)yi + hi - yj - hj);
                        sim = d0 <= T && d1 <= T && d2 <= T && d3 <= T;
                    }
                    if (!__any_sync(0xffffffffu, sim)) continue;
                    // merge i with all its similar j of this chunk: every tree involved is hooked under the smallest root
                    // seen; a failed hook (the root was taken by another warp meanwhile) is found again in the next round
                    for (;;) {
                        const int ri = __shfl_sync(0xffffffffu, uf_find(vlab, i), 0);
                        const int rj = sim ? uf_find(vlab, j) : ri;
                        const int tt = min(__reduce_min_sync(0xffffffffu, rj), ri);
                        bool again = false;
                        if (rj != tt) { atomicCAS(lab + rj, rj, tt); again = true; }
                        if (ri != tt) { if (lane == 0) atomicCAS(lab + ri, ri, tt); again = true; }
                        if (!__any_sync(0xffffffffu, again)) break;
                    }
                }
            }
        }
        __syncthreads();
        for (int i = threadIdx.x; i < n; i += kThreads) {   // flatten: every member points at its root
            const int root = uf_find(vlab, i);
            vlab[i] = root;
        }
    }
    // ---- 3. per-class integer sums
    __syncthreads();
    for (int i = threadIdx.x; i < n; i += kThreads) sx[i] = sy[i] = sw[i] = sh[i] = cnt[i] = 0;
    __syncthreads();
    for (int i = threadIdx.x; i < n; i += kThreads) {
        const int r = lab[i];
        atomicAdd(&sx[r], rx[i]); atomicAdd(&sy[r], ry[i]);
        atomicAdd(&sw[r], rw[i]); atomicAdd(&sh[r], rh[i]);
        atomicAdd(&cnt[r], 1);
    }
    __syncthreads();
    // ---- 4. mean rect: saturate_cast<int>(sum * (1.f / n))  (fp32, round half even)
    for (int i = threadIdx.x; i < n; i += kThreads) {
        if (lab[i] != i) continue;
        if (P.group_threshold > 0) {
            const float s = __fdiv_rn(1.f, (float)cnt[i]);
            sx[i] = __float2int_rn(__fmul_rn((float)sx[i], s)); sy[i] = __float2int_rn(__fmul_rn((float)sy[i], s));
            sw[i] = __float2int_rn(__fmul_rn((float)sw[i], s)); sh[i] = __float2int_rn(__fmul_rn((float)sh[i], s));
        }
        // containment margins of this class, once per class instead of once per pair (FP64 is scarce)
        dxr[i] = __double2int_rn((double)sw[i] * P.eps);
        dyr[i] = __double2int_rn((double)sh[i] * P.eps);
    }
    __syncthreads();
    // ---- 5. threshold on member count: the surviving classes, compacted in index (= cv::partition class) order ...
    int ns = 0;
    for (int base = 0; base < n; base += kThreads) {
        const int i = base + threadIdx.x;
        const bool f = i < n && lab[i] == i && (P.group_threshold <= 0 || cnt[i] > P.group_threshold);
        int tot;
        const int off = block_scan_flag(f, warp_tot, &tot);
        if (f) strong[ns + off] = i;
        ns += tot;
    }
    __syncthreads();
    // ... then the containment filter among them and vote_boxes' height filter; ordered output
    int nout = 0;
    for (int base = 0; base < ns; base += kThreads) {
        const int k = base + threadIdx.x;
        bool keep = false;
        int i = 0, n1 = 0;
        if (k < ns) {
            i = strong[k];
            n1 = cnt[i];
            keep = true;
            if (P.group_threshold > 0) {
                const int x1 = sx[i], y1 = sy[i], w1 = sw[i], h1 = sh[i];
                for (int q = 0; q < ns; ++q) {
                    const int j = strong[q];
                    if (j == i) continue;
                    const int n2 = cnt[j];
                    const int x2 = sx[j], y2 = sy[j], w2 = sw[j], h2 = sh[j];
                    const int dx = dxr[j], dy = dyr[j];
                    if (x1 >= x2 - dx && y1 >= y2 - dy && x1 + w1 <= x2 + w2 + dx && y1 + h1 <= y2 + h2 + dy &&
                        (n2 > max(3, n1) || n1 < 3)) { keep = false; break; }
                }
            }
            if (keep && (sh[i] - sy[i]) < P.min_height) keep = false;    // rect[3]-rect[1] >= 20 (:346)
        }
        int tot;
        const int off = block_scan_flag(keep, warp_tot, &tot);
        if (keep) {
            const int o_k = nout + off;
            if (o_k < FCNOD_MAX_DET) {
                int32_t* o = out + o_k * 6;
                o[0] = sx[i]; o[1] = sy[i]; o[2] = sw[i]; o[3] = sh[i];
                o[4] = (int)log((double)n1);                               // int(math.log(weight)) (:347,:123)
                o[5] = n1;
            }
        }
        nout += tot;
    }
    // the untruncated number of kept clusters: more than FCNOD_MAX_DET means the list was cut (the reference has no cap)
    if (threadIdx.x == 0) cls_cnt[b * ncls + cls] = nout;
}

// ---- standalone decode = gridbox_to_boxes (scripts/fcn_object_detector.py:357-394) for every (frame, class) ------
// threshold -> ordered compaction -> float64 boxes, no clustering.  HBM-bound: every coverage / bbox value of the used
// rows is read exactly once with 128-bit loads (a thread owns 4 consecutive cells), candidates leave as 32-byte box
// records + (cell, score).  One CTA of 256 threads per (frame, class): 784 cells = 196 threads x 4 cells.
constexpr int kDecThreads = 256;
__global__ void __launch_bounds__(kDecThreads) decode_kernel(const float* __restrict__ coverage, const float* __restrict__ bboxes,
                                                            int c_cvg, int c_bbox, int gh, int gw, PostParams P, int cap,
                                                            double* __restrict__ out_boxes /*[B][ncls][cap][4]*/,
                                                            int32_t* __restrict__ out_cell /*[B][ncls][cap]*/,
                                                            float* __restrict__ out_score /*[B][ncls][cap]*/,
                                                            int32_t* __restrict__ out_count /*[B][ncls]*/) {
    __shared__ int warp_tot[kDecThreads / 32];
    const int cls = blockIdx.x, b = blockIdx.y, ncls = gridDim.x;
    const int cch = cls + (P.skip_bg ? 1 : 0);
    const int plane = gh * gw, ncell = P.rows_used * gw;
    const float* cv = coverage + ((size_t)b * c_cvg + cch) * plane;
    const float* bb = bboxes + ((size_t)b * c_bbox + 4 * cls) * plane;
    const size_t slot = (size_t)b * ncls + cls;
    double* ob = out_boxes + slot * cap * 4;
    int32_t* oc = out_cell + slot * cap;
    float* os = out_score + slot * cap;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const bool vec = (plane % 4 == 0) && ((reinterpret_cast<uintptr_t>(coverage) | reinterpret_cast<uintptr_t>(bboxes)) % 16 == 0);
    int n = 0;
    for (int base = 0; base < ncell; base += 4 * kDecThreads) {
        const int i0 = base + 4 * (int)threadIdx.x;
        float c[4], v[4][4];
#pragma unroll
        for (int k = 0; k < 4; ++k) { c[k] = -INFINITY; v[0][k] = v[1][k] = v[2][k] = v[3][k] = 0.f; }
        if (vec && i0 + 3 < ncell) {
            const float4 c4 = __ldg(reinterpret_cast<const float4*>(cv + i0));
            c[0] = c4.x; c[1] = c4.y; c[2] = c4.z; c[3] = c4.w;
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                const float4 t = __ldg(reinterpret_cast<const float4*>(bb + (size_t)q * plane + i0));
                v[q][0] = t.x; v[q][1] = t.y; v[q][2] = t.z; v[q][3] = t.w;
            }
        } else {
#pragma unroll
            for (int k = 0; k < 4; ++k)
                if (i0 + k < ncell) {
                    c[k] = __ldg(cv + i0 + k);
#pragma unroll
                    for (int q = 0; q < 4; ++q) v[q][k] = __ldg(bb + (size_t)q * plane + i0 + k);
                }
        }
        int cnt = 0;
        bool f[4];
#pragma unroll
        for (int k = 0; k < 4; ++k) { f[k] = i0 + k < ncell && c[k] >= P.prob_thresh; cnt += f[k] ? 1 : 0; }
        // exclusive scan of the per-thread counts: warp shuffle scan + cross-warp totals (order = row-major cell order)
        int incl = cnt;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { const int t = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += t; }
        if (lane == 31) warp_tot[warp] = incl;
        __syncthreads();
        int wbase = 0, tot = 0;
#pragma unroll
        for (int w = 0; w < kDecThreads / 32; ++w) { const int t = warp_tot[w]; if (w < warp) wbase += t; tot += t; }
        __syncthreads();
        int k_out = n + wbase + incl - cnt;
#pragma unroll
        for (int k = 0; k < 4; ++k)
            if (f[k]) {
                if (k_out < cap) {
                    const int i = i0 + k, y = i / gw, x = i - y * gw;
                    const double mx = (double)(x * P.cell_w), my = (double)(y * P.cell_h);
                    double2* o2 = reinterpret_cast<double2*>(ob + (size_t)k_out * 4);
                    o2[0] = make_double2((double)v[0][k] + mx, (double)v[1][k] + my);
                    o2[1] = make_double2((double)v[2][k] + mx, (double)v[3][k] + my);
                    oc[k_out] = i;
                    os[k_out] = c[k];
                }
                ++k_out;
            }
        n += tot;
    }
    if (threadIdx.x == 0) out_count[slot] = n;          // untruncated; rows beyond `cap` are not stored
}

__global__ void assemble_kernel(const int32_t* __restrict__ cls_det, const int32_t* __restrict__ cls_cnt, int B, int ncls,
                                const int32_t* __restrict__ src_hw, int net_w, int net_h, int32_t* __restrict__ out_det,
                                int32_t* __restrict__ out_label, int32_t* __restrict__ out_weight,
                                int32_t* __restrict__ out_count) {
    const int b = blockIdx.x * (blockDim.x / 32) + (threadIdx.x >> 5);
    const int lane = threadIdx.x & 31;
    if (b >= B) return;
    double diffx = 1.0, diffy = 1.0;
    if (src_hw) {
        diffx = (double)src_hw[2 * b + 1] / (double)net_w;
        diffy = (double)src_hw[2 * b] / (double)net_h;
    }
    int total = 0, total_all = 0;
    for (int c = 0; c < ncls; ++c) {
        const int k_all = cls_cnt[b * ncls + c], k = min(k_all, FCNOD_MAX_DET);
        const int32_t* src = cls_det + ((size_t)b * ncls + c) * FCNOD_MAX_DET * 6;
        for (int i = lane; i < k; i += 32) {
            const int d = total + i;
            if (d >= FCNOD_MAX_DET) break;
            int32_t* o = out_det + ((size_t)b * FCNOD_MAX_DET + d) * 5;
            o[0] = (int)((double)src[i * 6 + 0] * diffx);
            o[1] = (int)((double)src[i * 6 + 1] * diffy);
            o[2] = (int)((double)src[i * 6 + 2] * diffx);
            o[3] = (int)((double)src[i * 6 + 3] * diffy);
            o[4] = src[i * 6 + 4];
            out_label[b * FCNOD_MAX_DET + d] = c;
            out_weight[b * FCNOD_MAX_DET + d] = src[i * 6 + 5];
        }
        total += k;
        total_all += k_all;
    }
    // out_count = every cluster the reference would have returned; only the first FCNOD_MAX_DET of them are stored
    if (lane == 0) out_count[b] = total_all;
}

}  // namespace

size_t postnet_scratch_bytes(int B, int c_cvg) {
    return (size_t)B * c_cvg * (FCNOD_MAX_DET * 6 + 1) * sizeof(int32_t) + kDeltaTable * sizeof(int32_t);
}
int postnet_num_launches() { return 2; }

namespace {
// Per device: the kernel's dynamic shared-memory limit (set once) and the SimilarRects threshold tables T(m), one per
// eps value seen, built on the host with the very double expression of delta_table_kernel (IEEE multiply + floor: same
// bits) and kept for the life of the process -- a replayed CUDA graph does not rebuild it every step.
struct PostDevState {
    bool attr_set = false;
    std::map<unsigned long long, int*> tables;     // bit pattern of eps -> device table
};
std::mutex g_post_mutex;
PostDevState g_post_state[64];

// The table for `eps` on the current device, or nullptr when it is not cached and cannot be built right now (the
// calling thread is capturing a graph: the caller then builds it in-stream into its scratch area).
const int* postnet_prepare_locked(double eps, cudaStream_t stream) {
    int dev = 0;
    FCNOD_CUDA(cudaGetDevice(&dev));
    FCNOD_CHECK(dev >= 0 && dev < 64, "postnet: device ordinal out of range");
    std::lock_guard<std::mutex> lock(g_post_mutex);
    PostDevState& st = g_post_state[dev];
    cudaStreamCaptureStatus cap = cudaStreamCaptureStatusNone;
    if (stream) FCNOD_CUDA(cudaStreamIsCapturing(stream, &cap));
    if (!st.attr_set) {
        FCNOD_CHECK(cap == cudaStreamCaptureStatusNone, "postnet: first use on this device inside a graph capture; call "
                    "fcnod_postnet_prepare (or any post-net entry point) once before capturing");
        FCNOD_CUDA(cudaFuncSetAttribute(decode_group_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024));
        st.attr_set = true;
    }
    unsigned long long key;
    memcpy(&key, &eps, 8);
    std::map<unsigned long long, int*>::iterator it = st.tables.find(key);
    if (it != st.tables.end()) return it->second;
    if (cap != cudaStreamCaptureStatusNone) return nullptr;
    std::vector<int> host(kDeltaTable);
    for (int m = 0; m < kDeltaTable; ++m) host[m] = (int)floor(eps * (double)m * 0.5);
    int* d = 0;
    FCNOD_CUDA(cudaMalloc(&d, kDeltaTable * sizeof(int)));
    cudaError_t e = cudaMemcpy(d, host.data(), kDeltaTable * sizeof(int), cudaMemcpyHostToDevice);
    if (e != cudaSuccess) { cudaFree(d); FCNOD_CUDA(e); }
    st.tables[key] = d;
    return d;
}
}  // namespace

void postnet_prepare(double eps) { postnet_prepare_locked(eps, 0); }

void postnet_launch(const float* coverage, const float* bboxes, int B, int c_cvg, int c_bbox, int gh, int gw,
                    const int32_t* src_hw, const fcnod_postnet_params& p, void* scratch, int32_t* out_det,
                    int32_t* out_label, int32_t* out_weight, int32_t* out_count, cudaStream_t stream) {
    FCNOD_CHECK(B > 0 && c_cvg > 0 && gh > 0 && gw > 0, "postnet: bad dimensions");
    FCNOD_CHECK(p.stride > 0 && p.net_w >= p.stride && p.net_h >= p.stride, "postnet: bad stride / net size");
    const int ncls = c_cvg - (p.skip_background ? 1 : 0);
    FCNOD_CHECK(ncls > 0, "postnet: no foreground coverage channel");
    FCNOD_CHECK(c_bbox >= 4 * ncls, "postnet: bboxes has fewer than 4 channels per foreground class");
    FCNOD_CHECK(gh * gw <= kMaxCand, "postnet: grid larger than 1024 cells is not supported");
    PostParams P;
    P.prob_thresh = p.prob_thresh;
    P.group_threshold = p.min_boxes;
    P.eps = p.eps;
    const int grid_sz_x = p.net_w / p.stride, grid_sz_y = p.net_h / p.stride;   // int(im_sz / stride)   (:362-363)
    P.cell_w = p.net_w / grid_sz_x;                                              // Py2 int division     (:368-369)
    P.cell_h = p.net_h / grid_sz_y;
    P.rows_used = std::min(gh, std::min(grid_sz_y, grid_sz_x));                  // net_cvg[0:gy][0:gx]  (:371)
    P.skip_bg = p.skip_background ? 1 : 0;
    P.min_height = p.min_height;
    P.net_w = p.net_w; P.net_h = p.net_h;
    int32_t* cls_det = static_cast<int32_t*>(scratch);
    int32_t* cls_cnt = cls_det + (size_t)B * ncls * FCNOD_MAX_DET * 6;
    const int* delta_tab = postnet_prepare_locked(p.eps, stream);
    if (!delta_tab) {   // new eps while capturing: the table is built in-stream (and becomes part of the graph)
        int* tab = static_cast<int*>(scratch) + (size_t)B * c_cvg * (FCNOD_MAX_DET * 6 + 1);
        delta_table_kernel<<<kDeltaTable / 256, 256, 0, stream>>>(p.eps, tab);
        FCNOD_CUDA(cudaGetLastError());
        delta_tab = tab;
    }
    const size_t adj_bytes = (size_t)8 * gh * gw * sizeof(unsigned);
    decode_group_kernel<<<dim3(ncls, B), kThreads, adj_bytes, stream>>>(coverage, bboxes, c_cvg, c_bbox, gh, gw, P, delta_tab, cls_det, cls_cnt);
    FCNOD_CUDA(cudaGetLastError());
    assemble_kernel<<<ceil_div(B, 4), 128, 0, stream>>>(cls_det, cls_cnt, B, ncls, src_hw, p.net_w, p.net_h, out_det,
                                                        out_label, out_weight, out_count);
    FCNOD_CUDA(cudaGetLastError());
}

}  // namespace fcnod

using namespace fcnod;

extern "C" {

void fcnod_postnet_default_params(fcnod_postnet_params* p) {
    if (!p) return;
    p->prob_thresh = 0.5f; p->min_boxes = 3; p->eps = 0.2;
    p->net_w = 448; p->net_h = 448; p->stride = 16; p->skip_background = 1; p->min_height = 20;
}

size_t fcnod_postnet_scratch_bytes(int B, int c_cvg) { return postnet_scratch_bytes(B, c_cvg); }

int fcnod_postnet_prepare(const fcnod_postnet_params* p) {
    FCNOD_API_BEGIN
    FCNOD_CHECK(p, "null argument");
    postnet_prepare(p->eps);
    FCNOD_API_END
}

int fcnod_postnet_device(const float* coverage, const float* bboxes, int B, int c_cvg, int c_bbox, int gh, int gw,
                         const int32_t* src_hw, const fcnod_postnet_params* p, void* scratch, int32_t* out_det,
                         int32_t* out_label, int32_t* out_weight, int32_t* out_count, void* stream) {
    FCNOD_API_BEGIN
    FCNOD_CHECK(coverage && bboxes && p && scratch && out_det && out_label && out_weight && out_count, "null argument");
    postnet_launch(coverage, bboxes, B, c_cvg, c_bbox, gh, gw, src_hw, *p, scratch, out_det, out_label, out_weight,
                   out_count, (cudaStream_t)stream);
    FCNOD_API_END
}

int fcnod_decode_device(const float* coverage, const float* bboxes, int B, int c_cvg, int c_bbox, int gh, int gw,
                        const fcnod_postnet_params* p, int cap, double* out_boxes, int32_t* out_cell, float* out_score,
                        int32_t* out_count, void* stream) {
    FCNOD_API_BEGIN
    FCNOD_CHECK(coverage && bboxes && p && out_boxes && out_cell && out_score && out_count, "null argument");
    FCNOD_CHECK(B > 0 && c_cvg > 0 && gh > 0 && gw > 0 && cap > 0, "decode: bad dimensions");
    FCNOD_CHECK(p->stride > 0 && p->net_w >= p->stride && p->net_h >= p->stride, "decode: bad stride / net size");
    const int ncls = c_cvg - (p->skip_background ? 1 : 0);
    FCNOD_CHECK(ncls > 0, "decode: no foreground coverage channel");
    FCNOD_CHECK(c_bbox >= 4 * ncls, "decode: bboxes has fewer than 4 channels per foreground class");
    FCNOD_CHECK(B <= 65535, "decode: at most 65535 frames per launch");
    PostParams P;
    memset(&P, 0, sizeof(P));
    P.prob_thresh = p->prob_thresh;
    const int grid_sz_x = p->net_w / p->stride, grid_sz_y = p->net_h / p->stride;
    P.cell_w = p->net_w / grid_sz_x; P.cell_h = p->net_h / grid_sz_y;
    P.rows_used = std::min(gh, std::min(grid_sz_y, grid_sz_x));
    P.skip_bg = p->skip_background ? 1 : 0;
    decode_kernel<<<dim3(ncls, B), kDecThreads, 0, (cudaStream_t)stream>>>(coverage, bboxes, c_cvg, c_bbox, gh, gw, P, cap, out_boxes,
                                                                          out_cell, out_score, out_count);
    FCNOD_CUDA(cudaGetLastError());
    FCNOD_API_END
}

int fcnod_postnet(const float* coverage, const float* bboxes, int B, int c_cvg, int c_bbox, int gh, int gw,
                  const int32_t* src_hw, const fcnod_postnet_params* p, int32_t* out_det, int32_t* out_label,
                  int32_t* out_weight, int32_t* out_count) {
    FCNOD_API_BEGIN
    FCNOD_CHECK(coverage && bboxes && p && out_det && out_label && out_weight && out_count, "null argument");
    FCNOD_CHECK(B > 0 && c_cvg > 0 && c_bbox > 0 && gh > 0 && gw > 0, "postnet: bad dimensions");
    const size_t ncv = (size_t)B * c_cvg * gh * gw, nbb = (size_t)B * c_bbox * gh * gw;
    const size_t M = FCNOD_MAX_DET, scr = postnet_scratch_bytes(B, c_cvg);
    Workspace& ws = workspace();
    cudaStream_t s = ws.stream();
    ws.reserve(Workspace::padded(ncv * 4) + Workspace::padded(nbb * 4) + Workspace::padded((size_t)B * 2 * 4) +
               Workspace::padded((size_t)B * (M * 7 + 1) * 4) + Workspace::padded(scr));
    float* d_cv = ws.take<float>(ncv);
    float* d_bb = ws.take<float>(nbb);
    int32_t* d_hw = src_hw ? ws.take<int32_t>((size_t)B * 2) : 0;
    int32_t* d_det = ws.take<int32_t>((size_t)B * (M * 7 + 1));      // det | label | weight | count, contiguous
    int32_t* d_lab = d_det + (size_t)B * M * 5;
    int32_t* d_w = d_lab + (size_t)B * M;
    int32_t* d_cnt = d_w + (size_t)B * M;
    void* d_scr = ws.take<char>(scr);
    FCNOD_CUDA(cudaMemcpyAsync(d_cv, coverage, ncv * 4, cudaMemcpyHostToDevice, s));
    FCNOD_CUDA(cudaMemcpyAsync(d_bb, bboxes, nbb * 4, cudaMemcpyHostToDevice, s));
    if (src_hw) FCNOD_CUDA(cudaMemcpyAsync(d_hw, src_hw, (size_t)B * 2 * 4, cudaMemcpyHostToDevice, s));
    FCNOD_CUDA(cudaMemsetAsync(d_det, 0, (size_t)B * (M * 7 + 1) * 4, s));
    postnet_launch(d_cv, d_bb, B, c_cvg, c_bbox, gh, gw, d_hw, *p, d_scr, d_det, d_lab, d_w, d_cnt, s);
    FCNOD_CUDA(cudaMemcpyAsync(out_det, d_det, (size_t)B * M * 5 * 4, cudaMemcpyDeviceToHost, s));
    FCNOD_CUDA(cudaMemcpyAsync(out_label, d_lab, (size_t)B * M * 4, cudaMemcpyDeviceToHost, s));
    FCNOD_CUDA(cudaMemcpyAsync(out_weight, d_w, (size_t)B * M * 4, cudaMemcpyDeviceToHost, s));
    FCNOD_CUDA(cudaMemcpyAsync(out_count, d_cnt, (size_t)B * 4, cudaMemcpyDeviceToHost, s));
    FCNOD_CUDA(cudaStreamSynchronize(s));
    FCNOD_API_END
}

int fcnod_decode(const float* coverage, const float* bboxes, int B, int c_cvg, int c_bbox, int gh, int gw,
                 const fcnod_postnet_params* p, int cap, double* out_boxes, int32_t* out_cell, float* out_score,
                 int32_t* out_count) {
    FCNOD_API_BEGIN
    FCNOD_CHECK(coverage && bboxes && p && out_boxes && out_cell && out_score && out_count, "null argument");
    FCNOD_CHECK(B > 0 && c_cvg > 0 && c_bbox > 0 && gh > 0 && gw > 0 && cap > 0, "decode: bad dimensions");
    const int ncls = c_cvg - (p->skip_background ? 1 : 0);
    FCNOD_CHECK(ncls > 0, "decode: no foreground coverage channel");
    const size_t ncv = (size_t)B * c_cvg * gh * gw, nbb = (size_t)B * c_bbox * gh * gw, slots = (size_t)B * ncls;
    Workspace& ws = workspace();
    cudaStream_t s = ws.stream();
    ws.reserve(Workspace::padded(ncv * 4) + Workspace::padded(nbb * 4) + Workspace::padded(slots * cap * 32) +
               2 * Workspace::padded(slots * cap * 4) + Workspace::padded(slots * 4));
    float* d_cv = ws.take<float>(ncv);
    float* d_bb = ws.take<float>(nbb);
    double* d_box = ws.take<double>(slots * cap * 4);
    int32_t* d_cell = ws.take<int32_t>(slots * cap);
    float* d_score = ws.take<float>(slots * cap);
    int32_t* d_cnt = ws.take<int32_t>(slots);
    FCNOD_CUDA(cudaMemcpyAsync(d_cv, coverage, ncv * 4, cudaMemcpyHostToDevice, s));
    FCNOD_CUDA(cudaMemcpyAsync(d_bb, bboxes, nbb * 4, cudaMemcpyHostToDevice, s));
    int rc = fcnod_decode_device(d_cv, d_bb, B, c_cvg, c_bbox, gh, gw, p, cap, d_box, d_cell, d_score, d_cnt, s);
    FCNOD_CHECK(rc == 0, fcnod_last_error());
    FCNOD_CUDA(cudaMemcpyAsync(out_boxes, d_box, slots * cap * 32, cudaMemcpyDeviceToHost, s));
    FCNOD_CUDA(cudaMemcpyAsync(out_cell, d_cell, slots * cap * 4, cudaMemcpyDeviceToHost, s));
    FCNOD_CUDA(cudaMemcpyAsync(out_score, d_score, slots * cap * 4, cudaMemcpyDeviceToHost, s));
    FCNOD_CUDA(cudaMemcpyAsync(out_count, d_cnt, slots * 4, cudaMemcpyDeviceToHost, s));
    FCNOD_CUDA(cudaStreamSynchronize(s));
    FCNOD_API_END
}

}  // extern "C"
