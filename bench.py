#!/usr/bin/env python
"""Benchmark of the detection hot path (BASELINE.json metric: frames/sec @448x448, FCN fwd + decode + NMS).

    python bench.py --gpus N --steps K --warmup W            # this repo's B200 path
    python bench.py --impl reference --gpus N --steps K ...  # the CPU restatement of the reference, host cores
    python bench.py --global-batch 256 ...                   # strong scaling: 256 frames split over the ranks (config 4)

One "step" = one pass of the whole hot path (network forward + coverage decode + groupRectangles
clustering) over one batch of 64 synthetic 448x448 frames per GPU (BASELINE config 3:
`deploy.prototxt inference batch 64 @448x448 bf16 on 1 B200`), replayed as one CUDA graph.
Prints ONE JSON line (rank 0).  See DESIGN.md section "Measurement" for every field.
"""
import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

BATCH = 64                       # frames per GPU per step (BASELINE config 3)
FLOP_PER_FRAME = 2 * 7803973632  # 59 conv layers, 7.804 GMAC
METRIC = 'frames/sec @448x448 FCN fwd+decode+NMS'
UNIT = 'frames/s'
DECODE_BYTES_PER_FRAME = 47040   # 3 fg coverage maps + 12 bbox maps, fp32, 28x28
RASTER_BYTES_PER_IMAGE = 53312   # (1 + 4*4) x 28 x 28 fp32 written, C = 1


def peaks():
    p = os.path.join(ROOT, 'MEASURED_PEAKS.json')
    if os.path.exists(p):
        d = json.load(open(p))
        return dict(bf16_burst=d['bf16_tflops'], bf16_sustained=d.get('bf16_tflops_sustained', d['bf16_tflops']),
                    hbm=d['hbm_gbs'], which='measured (MEASURED_PEAKS.json)')
    return dict(bf16_burst=1590.0, bf16_sustained=1400.0, hbm=6650.0, which='fallback (B200_PROFILING.md)')


def ncu_conv_summary():
    """From the newest committed `ncu --set full` summary of the conv launches under profiles/ (tools/gpu_round.sh):
    dram bytes per launch, and the time-weighted sm__pipe_tensor_cycles_active (% of peak, elapsed) over all launches."""
    import csv
    import glob
    files = sorted(glob.glob(os.path.join(ROOT, 'profiles', '*_conv_tc2_ncu_full.csv')))
    if not files:
        return None, None, None
    rows = {r[0]: r for r in csv.reader(open(files[-1])) if r}
    try:
        scale = {'byte': 1.0, 'Kbyte': 1e3, 'Mbyte': 1e6, 'Gbyte': 1e9}
        tot = 0.0
        for k in ('dram__bytes_read.sum', 'dram__bytes_write.sum'):
            tot += sum(float(v) for v in rows[k][2:]) * scale[rows[k][1]]
        n = len(rows['dram__bytes_read.sum']) - 2
        t = [float(v) for v in rows['gpu__time_duration.sum'][2:]]
        a = [float(v) for v in rows['sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed'][2:]]
        pipe = sum(x * y for x, y in zip(t, a)) / sum(t)
        return tot / n, pipe, os.path.basename(files[-1])
    except (KeyError, ValueError, ZeroDivisionError):
        return None, None, None


class ClockSampler(object):
    """nvidia-smi clocks + throttle reasons sampled DURING the timed region."""
    Q = ('clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,'
         'clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap')

    def __init__(self, index):
        self.rows = []
        self.proc = None
        try:
            self.proc = subprocess.Popen(['nvidia-smi', '-i', str(index), '--query-gpu=' + self.Q,
                                          '--format=csv,noheader,nounits', '-lms', '100'],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append((time.time(), line.strip()))

    def stop(self, t0, t1):
        if self.proc is None:
            return {'sm_mhz': None, 'sm_max_mhz': None, 'reasons': ['nvidia-smi unavailable']}
        time.sleep(0.15)
        self.proc.terminate()
        sm, mx, reasons = [], None, set()
        names = ['hw_slowdown', 'hw_thermal_slowdown', 'sw_thermal_slowdown', 'sw_power_cap']
        for ts, line in self.rows:
            f = [x.strip() for x in line.split(',')]
            if len(f) < 7:
                continue
            try:
                mx = float(f[1])
                if t0 - 0.05 <= ts <= t1 + 0.15:
                    sm.append(float(f[0]))
                    for n, v in zip(names, f[3:7]):
                        if v.lower().startswith('active'):
                            reasons.add(n)
            except ValueError:
                continue
        return {'sm_mhz': float(np.median(sm)) if sm else None, 'sm_max_mhz': mx, 'reasons': sorted(reasons),
                'samples': len(sm)}


def model_files():
    """deploy.prototxt + synthetic snapshot: xavier fillers of the prototxt (seed 0) with heads biased so that frames
    yield detections (caffemodel.firing_heads) -- the post-net's clustering, containment filter, assembly and rescale
    then run on real candidates inside the timed region, as they do with a trained snapshot."""
    from fcn_object_detector_b200.caffemodel import write_synthetic_caffemodel
    from fcn_object_detector_b200.netspec import write_deploy_prototxt
    d = '/tmp/fcnod_bench_%d' % os.getpid()
    os.makedirs(d, exist_ok=True)
    proto = write_deploy_prototxt(os.path.join(d, 'deploy.prototxt'))
    cm = write_synthetic_caffemodel(os.path.join(d, 'snapshot_iter_1000.caffemodel'), seed=0, firing=True)
    return proto, cm


# ----------------------------------------------------------------------------------------------- reference arm
def cpu_reference_fps(frames, reps, proto, cm, threads=None):
    """The oracle (CPU restatement of the reference: torch-CPU fp32 net + numpy/python post-net) timed on the
    host cores.  Checker code used as the *baseline being timed*, never as the product.
    Returns (per-rep total seconds, per-rep net seconds, per-rep post-net seconds, threads)."""
    import torch
    from oracle import net_ref, postnet_ref, prototxt
    from oracle.caffemodel_io import read_caffemodel
    if threads:
        torch.set_num_threads(threads)
    Ls = prototxt.layers(prototxt.load(proto))
    W = read_caffemodel(cm)
    x = np.random.RandomState(2).uniform(0, 1, (frames, 3, 448, 448)).astype(np.float32)
    tot, tnet, tpost = [], [], []
    for _ in range(reps):
        t0 = time.perf_counter()
        o = net_ref.forward(Ls, W, x, want=['coverage', 'bboxes'])
        t1 = time.perf_counter()
        for b in range(frames):
            postnet_ref.postnet(o['coverage'][b], o['bboxes'][b], (448, 448, 3))
        t2 = time.perf_counter()
        tot.append(t2 - t0); tnet.append(t1 - t0); tpost.append(t2 - t1)
    return tot, tnet, tpost, torch.get_num_threads()


def run_reference(args):
    rank = int(os.environ.get('RANK', '0'))
    if rank != 0:
        return
    proto, cm = model_files()
    frames = 8
    # torchrun exports OMP_NUM_THREADS=1; the reference arm takes every host core this process may run on
    ncpu = len(os.sched_getaffinity(0))
    for _ in range(max(args.warmup, 1)):
        cpu_reference_fps(frames, 1, proto, cm, threads=ncpu)
    times, tnet, tpost, threads = cpu_reference_fps(frames, args.steps, proto, cm, threads=ncpu)
    tot = sum(times)
    fps = frames * len(times) / tot
    line = {'impl': 'reference', 'metric': METRIC, 'value': fps, 'unit': UNIT, 'n_gpus': args.gpus, 'steps': args.steps,
            'warmup': args.warmup, 'ms_per_step': 1e3 * tot / len(times), 'higher_is_better': True, 'scaling': 'weak',
            'vs_baseline': None, 'dtype': 'f32', 'data': 'synthetic',
            'config': {'workload': 'deploy.prototxt 448x448 fwd+decode+NMS, CPU restatement of the reference '
                                   '(Caffe/Python-2 reference not runnable here), %d frames per step' % frames},
            'cpu_baseline': {'value': fps, 'unit': UNIT, 'cores': threads, 'kind': 'port',
                             'sample': '%d steps x %d frames, torch-CPU fp32 net + python post-net' % (args.steps, frames),
                             'net_ms_per_frame': 1e3 * sum(tnet) / (frames * len(times)),
                             'postnet_ms_per_frame': 1e3 * sum(tpost) / (frames * len(times))},
            'e2e': {'value': fps, 'unit': UNIT, 'h2d_bytes_per_step': 0, 'd2h_bytes_per_step': 0},
            'gpu_launches': 0}
    print(json.dumps(line))


# ----------------------------------------------------------------------------------------------- extra sections (N = 1)
def _events_ms(torch, fn, reps):
    fn()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / reps


def bandwidth_kernels(torch, L, _lib, stream, pk, reps=20):
    """decode (gridbox_to_boxes) and label_raster are launch-latency bound at the configured batches (3 MB / 1.7 MB):
    their HBM roofline is measured at 4096 frames / images per launch, as SURVEY.md 8(d) asks, plus the config batch."""
    out = {}
    p = _lib.default_postnet_params()
    for B in (64, 4096):
        g = torch.Generator(device='cuda').manual_seed(5)
        cov = torch.rand((B, 4, 28, 28), device='cuda', generator=g) * 0.55       # ~9 % of the cells >= 0.5
        bb = torch.randn((B, 16, 28, 28), device='cuda', generator=g) * 20
        cap = 784
        boxes = torch.empty((B, 3, cap, 4), dtype=torch.float64, device='cuda')
        cell = torch.empty((B, 3, cap), dtype=torch.int32, device='cuda')
        score = torch.empty((B, 3, cap), dtype=torch.float32, device='cuda')
        cnt = torch.empty((B, 3), dtype=torch.int32, device='cuda')

        def run():
            _lib.check(L.fcnod_decode_device(cov.data_ptr(), bb.data_ptr(), B, 4, 16, 28, 28, C.byref(p), cap, boxes.data_ptr(),
                                             cell.data_ptr(), score.data_ptr(), cnt.data_ptr(), stream))
        ms = _events_ms(torch, run, reps)
        ncand = int(cnt.sum().item())
        algo = B * DECODE_BYTES_PER_FRAME + ncand * 40 + B * 3 * 4      # maps read + 40 B per candidate + counts written
        out['decode_b%d' % B] = {'frames': B, 'ms': ms, 'candidates': ncand, 'algorithmic_bytes': algo,
                                 'achieved_gbs': algo / ms / 1e6, 'frac_of_hbm_peak': algo / ms / 1e6 / pk['hbm']}
        del cov, bb, boxes, cell, score, cnt
    rs = np.random.RandomState(1)
    for N in (32, 4096):
        rects = np.zeros((N, 1, 4), np.int32)
        for i in range(N):
            w, h = int(rs.randint(32, 257)), int(rs.randint(32, 257))
            rects[i, 0] = [int(rs.randint(0, 448 - w)), int(rs.randint(0, 448 - h)), w, h]
        d_r = torch.from_numpy(rects).cuda(); d_l = torch.zeros((N, 1), dtype=torch.int32, device='cuda')
        fg = torch.empty((N, 1, 28, 28), device='cuda')
        o4 = [torch.empty((N, 4, 28, 28), device='cuda') for _ in range(4)]

        def run():
            _lib.check(L.fcnod_label_raster_device(d_r.data_ptr(), d_l.data_ptr(), None, N, 1, 448, 448, 16, 1, fg.data_ptr(),
                                                   o4[0].data_ptr(), o4[1].data_ptr(), o4[2].data_ptr(), o4[3].data_ptr(), stream))
        ms = _events_ms(torch, run, reps)
        algo = N * (RASTER_BYTES_PER_IMAGE + 20)
        out['label_raster_n%d' % N] = {'images': N, 'ms': ms, 'algorithmic_bytes': algo, 'achieved_gbs': algo / ms / 1e6,
                                       'frac_of_hbm_peak': algo / ms / 1e6 / pk['hbm']}
    out['note'] = ('HBM-bound kernels; peak = %s GB/s (%s). The 4096-frame launches are the roofline measurement '
                   '(the configured batches move 3 MB / 1.7 MB and are launch-latency bound).' % (pk['hbm'], pk['which']))
    return out


def layer_section(torch, L, _lib, pk):
    """BASELINE config 2: DataArgumentationLayer forward, param_str '448, 448,16, 1', batch 32, 480x640 sources,
    one bbox per image -- images/s through the Python layer (host blobs, page-locked like Caffe's), the device-only
    time, and the CPU oracle beside it."""
    import random
    import fcn_object_detector_b200.caffe_shim as caffe
    from fcn_object_detector_b200.data_argumentation_layer import DataArgumentationLayer
    from oracle import layer_ref
    rs = np.random.RandomState(1)
    N, Hs, Ws = 32, 480, 640
    bottom = [caffe.Blob((N, 3, Hs, Ws), pinned=True), caffe.Blob((N, 1, 1, 5), pinned=True)]
    bottom[0].data[...] = rs.randint(0, 256, (N, 3, Hs, Ws)).astype(np.float32)
    rng = random.Random(1234)
    draws_d, draws = [], []
    for i in range(N):
        w, h = int(rs.randint(32, 257)), int(rs.randint(32, 257))
        rect = [int(rs.randint(0, Ws - w)), int(rs.randint(0, Hs - h)), w, h]
        bottom[1].data[i, 0, 0] = rect + [0]
        d = layer_ref.draw_params(rng, Ws, Hs, rect)
        draws_d.append(d)
        a = _lib.AugDraw()
        for k, v in d.items():
            setattr(a, k, v)
        draws.append(a)
    top = [caffe.Blob((1,), pinned=True) for _ in range(6)]
    layer = DataArgumentationLayer()
    layer.param_str = '448, 448,16, 1'
    layer.setup(bottom, top)
    layer.reshape(bottom, top)
    for _ in range(3):
        layer.forward(bottom, top, draws=draws)
    reps = 10
    t0 = time.perf_counter()
    for _ in range(reps):
        layer.reshape(bottom, top)
        layer.forward(bottom, top, draws=draws)
    e2e_ms = (time.perf_counter() - t0) * 1e3 / reps
    # device only: inputs resident
    d_in = torch.from_numpy(np.asarray(bottom[0].data)).cuda(); d_lab = torch.from_numpy(np.asarray(bottom[1].data).reshape(N, 5)).cuda()
    darr = (_lib.AugDraw * N)(*draws)
    d_draw = torch.from_numpy(np.frombuffer(bytes(darr), dtype=np.uint8).copy()).cuda()
    tops = [torch.empty(t.data.shape, device='cuda') for t in top]
    rects = torch.empty((N, 4), dtype=torch.int32, device='cuda')
    scr = torch.empty(int(L.fcnod_aug_scratch_bytes(N)), dtype=torch.uint8, device='cuda')
    stream = torch.cuda.current_stream().cuda_stream

    def run():
        _lib.check(L.fcnod_aug_forward_device(d_in.data_ptr(), d_lab.data_ptr(), d_draw.data_ptr(), N, Hs, Ws, 448, 448, 16, 1,
                                              tops[0].data_ptr(), tops[1].data_ptr(), tops[2].data_ptr(), tops[3].data_ptr(),
                                              tops[4].data_ptr(), tops[5].data_ptr(), rects.data_ptr(), scr.data_ptr(), stream))
    dev_ms = _events_ms(torch, run, 20)
    h2d = int(bottom[0].data.nbytes + bottom[1].data.nbytes + N * C.sizeof(_lib.AugDraw))
    d2h = int(sum(t.data.nbytes for t in top) + N * 16)
    # CPU oracle: the reference's numpy / python layer restated, 1 core (as the reference runs it), bounded sample
    ncpu_imgs = 8
    data_np = np.asarray(bottom[0].data)[:ncpu_imgs]; lab_np = np.asarray(bottom[1].data)[:ncpu_imgs]
    layer_ref.layer_forward(data_np[:1], lab_np[:1], draws_d[:1])
    t0 = time.perf_counter()
    layer_ref.layer_forward(data_np, lab_np, draws_d[:ncpu_imgs])
    cpu_s = time.perf_counter() - t0
    return {'workload': "DataArgumentationLayer.forward, param_str '448, 448,16, 1', batch 32, 480x640 fp32 sources, one bbox per "
                        "image (BASELINE config 2); colour augmentation off",
            'value': N / (e2e_ms * 1e-3), 'unit': 'images/s', 'ms_per_batch': e2e_ms,
            'api': 'DataArgumentationLayer.reshape + forward on page-locked host blobs -> fcnod_aug_forward',
            'h2d_bytes_per_batch': h2d, 'd2h_bytes_per_batch': d2h,
            'pcie_floor_ms_at_55GBs': (h2d + d2h) / 55e9 * 1e3,
            'device_only': {'ms_per_batch': dev_ms, 'images_per_s': N / (dev_ms * 1e-3), 'kernels': 4},
            'cpu_baseline': {'value': ncpu_imgs / cpu_s, 'unit': 'images/s', 'cores': 1, 'kind': 'port',
                             'sample': '%d images of the same batch, numpy/python restatement of argumentation_engine.py' % ncpu_imgs}}


def latency_section(torch, det, proto, cm, calls=200):
    """BASELINE config 1 as the node sees it: run_detector on ONE 448x448 BGR frame per call (host uint8 in, boxes out),
    bf16 plan; p50 / p99 over `calls` calls."""
    img = np.random.RandomState(0).randint(0, 256, (448, 448, 3)).astype(np.uint8)
    for _ in range(5):
        det.run_detector(img)
    ts = []
    for _ in range(calls):
        t0 = time.perf_counter()
        det.run_detector(img)
        ts.append((time.perf_counter() - t0) * 1e3)
    ts = np.sort(np.asarray(ts))
    return {'workload': 'run_detector(one 448x448 BGR uint8 frame), batch 1, bf16 plan (BASELINE config 1 input)', 'calls': calls,
            'p50_ms': float(ts[len(ts) // 2]), 'p99_ms': float(ts[int(len(ts) * 0.99) - 1]), 'mean_ms': float(ts.mean()),
            'min_ms': float(ts[0])}


# ----------------------------------------------------------------------------------------------- B200 arm
def run_ours(args):
    import torch
    import torch.distributed as dist
    from fcn_object_detector_b200 import _lib, build
    from fcn_object_detector_b200.detector import FCNObjectDetector
    rank = int(os.environ.get('RANK', '0'))
    world = int(os.environ.get('WORLD_SIZE', '1'))
    local = int(os.environ.get('LOCAL_RANK', '0'))
    if not torch.cuda.is_available():
        raise SystemExit('bench.py: no CUDA device; this path has no CPU fallback')
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group('nccl', device_id=torch.device('cuda', local))
    if rank == 0:
        build.build()
    if world > 1:
        dist.barrier()
    L = _lib.lib()
    proto, cm = model_files()
    det = FCNObjectDetector(pretrained_weights=cm, deployment_prototxt=proto, precision='bf16', device_id=local)
    net = det.net
    h = net.handle
    strong = args.global_batch > 0
    if strong:
        if args.global_batch % world:
            raise SystemExit('--global-batch must be a multiple of the number of ranks')
        B = args.global_batch // world      # config 4: 256 frames split evenly over the GPUs
    else:
        B = BATCH                           # config 3: 64 frames per GPU (weak scaling)
    # every rank gets its own frames; no data-path collective
    g = torch.Generator(device='cpu').manual_seed(2 + rank)
    data_host = torch.rand((B, 3, 448, 448), generator=g, dtype=torch.float32).pin_memory()
    data = data_host.cuda()
    # a non-default stream (the legacy default stream cannot be captured into a graph); made torch's current
    # stream so that torch.cuda.Event timings and the kernels share it
    tstream = torch.cuda.Stream()
    torch.cuda.set_stream(tstream)
    stream = tstream.cuda_stream
    p = _lib.default_postnet_params()
    _lib.check(L.fcnod_detect_graph_build(h, data.data_ptr(), B, C.byref(p), stream))
    launches_per_step = L.fcnod_net_launches_per_forward(h) + 2          # + decode_group_kernel + assemble_kernel

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    for _ in range(args.warmup):
        _lib.check(L.fcnod_detect_graph_launch(h, stream))
    barrier()
    sampler = ClockSampler(local) if rank == 0 else None
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.time()
    e0.record()
    for _ in range(args.steps):
        _lib.check(L.fcnod_detect_graph_launch(h, stream))
    e1.record()
    barrier()
    t1 = time.time()
    ms = e0.elapsed_time(e1)
    clocks = sampler.stop(t0, t1) if sampler else None
    if world > 1:
        t = torch.tensor([ms], device='cuda')
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    fps = world * B * args.steps / (ms * 1e-3)

    # ---- e2e: the public predict-on-image calls with HOST buffers (uint8 BGR frames in pinned memory):
    # H2D of the frames, prelude, net, post-net, D2H of the detections, every step.  Headline: the streaming call
    # (predict_stream: two batches in flight, the upload of step i+1 overlaps the compute of step i; two distinct
    # pinned frame buffers alternate); the one-call-per-batch form (predict_batch) is reported beside it.
    imgs2 = [torch.randint(0, 256, (B, 448, 448, 3), dtype=torch.uint8, generator=g).pin_memory().numpy() for _ in range(2)]
    imgs = imgs2[0]
    e_steps = max(3, min(args.steps, 20))

    def timed(fn):
        barrier()
        t0 = time.perf_counter()
        r = fn()
        torch.cuda.synchronize()
        t_ms = (time.perf_counter() - t0) * 1e3
        if world > 1:
            t = torch.tensor([t_ms], device='cuda')
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            t_ms = float(t.item())
        return r, t_ms

    det.predict_batch(imgs)
    det.predict_batch(imgs)
    res, sync_ms = timed(lambda: [det.predict_batch(imgs2[i & 1]) for i in range(e_steps)][-1])
    sync_fps = world * B * e_steps / (sync_ms * 1e-3)
    list(det.predict_stream(imgs2[i & 1] for i in range(3)))
    res, e_ms = timed(lambda: list(det.predict_stream(imgs2[i & 1] for i in range(e_steps)))[-1])
    e_fps = world * B * e_steps / (e_ms * 1e-3)
    n_det = int(sum(len(r[0]) for r in res))

    line = None
    if rank == 0:
        pk = peaks()
        # ---- per-launch timing of the plan (CUDA events on the launching stream), dominant kernel = conv_tc
        _lib.check(L.fcnod_detect_graph_build(h, data.data_ptr(), B, C.byref(p), stream))   # restore this batch's state
        plan = net.describe().splitlines()
        reps = 5
        per = []
        for i, desc in enumerate(plan):
            L.fcnod_net_run_step(h, i, 1, stream)
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            _lib.check(L.fcnod_net_run_step(h, i, reps, stream))
            b.record()
            torch.cuda.synchronize()
            per.append(a.elapsed_time(b) / reps)
        # post-net (2 launches) timed alone on the net's own output maps
        cov_p, bb_p = C.c_void_p(), C.c_void_p()
        _lib.check(L.fcnod_net_output_device(h, b'coverage', C.byref(cov_p)))
        _lib.check(L.fcnod_net_output_device(h, b'bboxes', C.byref(bb_p)))
        scr = torch.empty(int(L.fcnod_postnet_scratch_bytes(B, 4)), dtype=torch.uint8, device='cuda')
        o_det = torch.empty((B, _lib.FCNOD_MAX_DET, 5), dtype=torch.int32, device='cuda')
        o_lab = torch.empty((B, _lib.FCNOD_MAX_DET), dtype=torch.int32, device='cuda')
        o_w = torch.empty((B, _lib.FCNOD_MAX_DET), dtype=torch.int32, device='cuda')
        o_cnt = torch.empty((B,), dtype=torch.int32, device='cuda')

        def run_post():
            _lib.check(L.fcnod_postnet_device(cov_p, bb_p, B, 4, 16, 28, 28, None, C.byref(p), scr.data_ptr(), o_det.data_ptr(),
                                              o_lab.data_ptr(), o_w.data_ptr(), o_cnt.data_ptr(), stream))
        post_ms = _events_ms(torch, run_post, reps)
        from fcn_object_detector_b200.netspec import conv_shapes
        import re
        shapes = {n: (co, ci, k) for n, co, ci, k, _b in conv_shapes()}

        def step_macs(d):
            """algorithmic MACs per frame of one conv_tc launch (a launch may hold several fused convolutions)"""
            names = re.match(r'\d+: conv_tc (.+?) \d+x\d+/\d+ ', d).group(1).split(' + ')
            dims = d.split(' -> ')[1].split('(')[1].split(')')[0].split(',')
            out_hw = int(dims[1]) * int(dims[2])
            return sum(shapes[n][0] * shapes[n][1] * shapes[n][2] ** 2 * out_hw for n in names)
        macs = {d: step_macs(d) for d in plan if 'conv_tc' in d}
        tc_ms = sum(t for t, d in zip(per, plan) if 'conv_tc' in d)
        tc_flop = sum(2.0 * macs[d] * B for d in plan if 'conv_tc' in d)
        n_tc = sum('conv_tc' in d for d in plan)
        n_pool = sum('+maxpool3x3' in d for d in plan)
        tot_ms = sum(per) + post_ms
        ach = tc_flop / (tc_ms * 1e-3) / 1e12
        # the same without the launches that also pool or normalise (their time includes the max-pool / LRN work that used to
        # be launches of its own: ten pools, and the LRN half of the LRN+pool launch)
        def pure(d):
            return 'conv_tc' in d and '+maxpool3x3' not in d and '+lrn(' not in d
        pure_ms = sum(t for t, d in zip(per, plan) if pure(d))
        pure_flop = sum(2.0 * macs[d] * B for d in plan if pure(d))
        n_lrn = sum('+lrn(' in d for d in plan)
        ach_pure = pure_flop / (pure_ms * 1e-3) / 1e12
        whole = world * B * args.steps * FLOP_PER_FRAME / (ms * 1e-3) / 1e12 / world    # per GPU
        traffic, ncu_pipe, ncu_src = ncu_conv_summary()
        # Each conv launch is timed ALONE for a few tens of microseconds at full clocks: the denominator is the BURST
        # cuBLAS bf16 peak of MEASURED_PEAKS.json (the sustained figure is reported beside it).
        roof = {'bound': 'tensor',
                'kernel': 'conv_tc2_kernel (%d launches per step; %d of them also run the inception max-pool on their A-operand path, '
                          '%d the LRN in their epilogue)' % (n_tc, n_pool, n_lrn),
                'achieved': ach, 'peak': pk['bf16_burst'], 'peak_source': pk['which'] + ', burst cuBLAS bf16 (launches timed alone)',
                'unit': 'TFLOP/s', 'frac': ach / pk['bf16_burst'],
                'peak_sustained': pk['bf16_sustained'], 'frac_vs_sustained': ach / pk['bf16_sustained'],
                'excluding_pool_and_lrn_fused_launches': {'launches': n_tc - n_pool - n_lrn, 'achieved': ach_pure,
                                                          'frac': ach_pure / pk['bf16_burst'], 'share_of_conv_flop': pure_flop / tc_flop},
                'whole_step_achieved': whole, 'whole_step_frac': whole / pk['bf16_burst'],
                'ncu_tensor_pipe_pct_time_weighted': ncu_pipe,
                'traffic': traffic, 'traffic_unit': 'bytes per launch (dram read + write, ncu)', 'ncu_source': ncu_src,
                'share_of_step': tc_ms / tot_ms, 'avg_launch_ms': tc_ms / n_tc, 'flop_per_launch': tc_flop / n_tc,
                'sum_of_launches_ms': tot_ms, 'graph_step_ms': ms / args.steps,
                'note': 'algorithmic FLOP = 2 x MACs of the convolutions of each launch (SURVEY.md 8a) x batch; the launches of a '
                        'step overlap in the graph (independent inception branches run on separate graph branches), so the step '
                        'is shorter than the sum of its launches'}
        os.makedirs(os.path.join(ROOT, 'gpurun_out'), exist_ok=True)
        with open(os.path.join(ROOT, 'gpurun_out', 'per_launch_ms.json'), 'w') as f:
            json.dump([{'step': i, 'ms': t, 'desc': d} for i, (t, d) in enumerate(zip(per, plan))] +
                      [{'step': len(plan), 'ms': post_ms, 'desc': 'postnet decode_group_kernel + assemble_kernel'}], f, indent=1)
        # ---- CPU baseline: the oracle on the host cores, bounded sample; extra sections (N = 1 only)
        cpu = layer = latency = bwk = train = None
        if world == 1 and not args.no_cpu_baseline:
            frames = 4
            ncpu = len(os.sched_getaffinity(0))
            cpu_reference_fps(frames, 1, proto, cm, threads=ncpu)
            times, tnet, tpost, threads = cpu_reference_fps(frames, 3, proto, cm, threads=ncpu)
            nf = frames * len(times)
            cpu = {'value': nf / sum(times), 'unit': UNIT, 'cores': threads, 'kind': 'port',
                   'sample': '3 x %d frames (+1 warm-up), torch-CPU fp32 net + python post-net restatement' % frames,
                   'net_ms_per_frame': 1e3 * sum(tnet) / nf, 'postnet_ms_per_frame': 1e3 * sum(tpost) / nf}
        if world == 1 and not args.no_extras:
            bwk = bandwidth_kernels(torch, L, _lib, stream, pk)
            layer = layer_section(torch, L, _lib, pk)
            latency = latency_section(torch, det, proto, cm)
            # BASELINE config 5 (training step, batch 32) in short form; the full line: --workload train
            try:
                import bench_train
                train = bench_train.run_ours(args, section=True)
            except Exception as e:       # the headline line must survive a failure of an extra section
                train = {'error': '%s: %s' % (type(e).__name__, e)}
        wl = ('deploy.prototxt fwd + coverage decode + groupRectangles NMS, batch %d/GPU @448x448, bf16 tcgen05 convs, synthetic '
              'xavier weights (seed 0) with firing heads, one CUDA graph per step' % B)
        line = {'metric': METRIC, 'value': fps, 'unit': UNIT, 'n_gpus': world, 'steps': args.steps, 'warmup': args.warmup,
                'ms_per_step': ms / args.steps, 'higher_is_better': True, 'scaling': 'strong' if strong else 'weak',
                'vs_baseline': None, 'dtype': 'bf16', 'data': 'synthetic',
                'config': {'workload': wl, 'global_batch': world * B, 'per_gpu_batch': B,
                           'parallelism': 'frame-parallel x%d, no collectives' % world,
                           'l2': 'inputs (%.0f MB/step) + activations (GBs) exceed the 126 MB L2; no flush needed' % (B * 3 * 448 * 448 * 4 / 1e6)},
                'clocks': clocks,
                'e2e': {'value': e_fps, 'unit': UNIT, 'h2d_bytes_per_step': int(imgs.nbytes),
                        'd2h_bytes_per_step': int(B * (_lib.FCNOD_MAX_DET * 7 + 1) * 4),
                        'api': 'FCNObjectDetector.predict_stream(batches of uint8 BGR frames) -> fcnod_detect_submit / _collect, '
                               'two batches in flight', 'one_call_per_batch': {'value': sync_fps, 'unit': UNIT,
                               'api': 'FCNObjectDetector.predict_batch -> fcnod_detect'}, 'steps': e_steps,
                        'detections_last_step': n_det},
                'gpu_launches': launches_per_step * args.steps,
                'roofline': roof, 'cpu_baseline': cpu,
                'tflops_whole_step': world * B * args.steps * FLOP_PER_FRAME / (ms * 1e-3) / 1e12,
                'latency_batch1': latency, 'layer': layer, 'bandwidth_kernels': bwk, 'train': train, 'plan': plan}
        print(json.dumps(line))
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--gpus', type=int, default=1)
    ap.add_argument('--steps', type=int, default=20)
    ap.add_argument('--warmup', type=int, default=3)
    ap.add_argument('--impl', default='ours', choices=['ours', 'reference'])
    ap.add_argument('--global-batch', type=int, default=0,
                    help='strong scaling (BASELINE config 4): this many frames per step split evenly over the ranks')
    ap.add_argument('--workload', default='detect', choices=['detect', 'train'],
                    help="detect = BASELINE config 3 / 4 (the headline metric); train = BASELINE config 5 (bench_train.py)")
    ap.add_argument('--train-precision', default='bf16', choices=['fp32', 'bf16'],
                    help='--workload train: fp32 = SIMT parity engine, bf16 = tcgen05 GEMM engine (fp32 master weights)')
    ap.add_argument('--train-collective', default='peer', choices=['peer', 'nccl'],
                    help='--workload train, N > 1: peer = reduce-scatter + ADAM + all-gather in one kernel over NVLink peer memory; '
                         'nccl = all-reduce of the gradient buffer + update')
    ap.add_argument('--no-cpu-baseline', action='store_true')
    ap.add_argument('--no-extras', action='store_true', help='skip the layer / latency / bandwidth-kernel sections')
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == 'ours' else args.warmup
    if args.workload == 'train':
        import bench_train
        (bench_train.run_reference if args.impl == 'reference' else bench_train.run_ours)(args)
    elif args.impl == 'reference':
        run_reference(args)
    else:
        run_ours(args)


if __name__ == '__main__':
    main()
