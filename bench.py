#!/usr/bin/env python
"""Benchmark of the detection hot path (BASELINE.json metric: frames/sec @448x448, FCN fwd + decode + NMS).

    python bench.py --gpus N --steps K --warmup W            # this repo's B200 path
    python bench.py --impl reference --gpus N --steps K ...  # the CPU restatement of the reference, host cores

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


def peaks():
    p = os.path.join(ROOT, 'MEASURED_PEAKS.json')
    if os.path.exists(p):
        d = json.load(open(p))
        return dict(bf16_burst=d['bf16_tflops'], bf16_sustained=d.get('bf16_tflops_sustained', d['bf16_tflops']),
                    hbm=d['hbm_gbs'], which='measured')
    return dict(bf16_burst=1590.0, bf16_sustained=1400.0, hbm=6650.0, which='fallback')


def ncu_conv_traffic():
    """dram__bytes_read.sum + dram__bytes_write.sum per conv_tc2 launch, from the newest committed `ncu --set full`
    summary under profiles/ (tools/gpu_round.sh); (None, None) when there is none."""
    import csv
    import glob
    files = sorted(glob.glob(os.path.join(ROOT, 'profiles', '*_conv_tc2_ncu_full.csv')))
    if not files:
        return None, None
    rows = {r[0]: r for r in csv.reader(open(files[-1])) if r}
    try:
        scale = {'byte': 1.0, 'Kbyte': 1e3, 'Mbyte': 1e6, 'Gbyte': 1e9}
        tot = 0.0
        for k in ('dram__bytes_read.sum', 'dram__bytes_write.sum'):
            tot += sum(float(v) for v in rows[k][2:]) * scale[rows[k][1]]
        n = len(rows['dram__bytes_read.sum']) - 2
        return tot / n, os.path.basename(files[-1])
    except (KeyError, ValueError):
        return None, None


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
    from fcn_object_detector_b200.caffemodel import write_synthetic_caffemodel
    from fcn_object_detector_b200.netspec import write_deploy_prototxt
    d = '/tmp/fcnod_bench_%d' % os.getpid()
    os.makedirs(d, exist_ok=True)
    proto = write_deploy_prototxt(os.path.join(d, 'deploy.prototxt'))
    cm = write_synthetic_caffemodel(os.path.join(d, 'snapshot_iter_1000.caffemodel'), seed=0)
    return proto, cm


# ----------------------------------------------------------------------------------------------- reference arm
def cpu_reference_fps(frames, reps, proto, cm, threads=None):
    """The oracle (CPU restatement of the reference: torch-CPU fp32 net + numpy/python post-net) timed on the
    host cores.  Checker code used as the *baseline being timed*, never as the product."""
    import torch
    from oracle import net_ref, postnet_ref, prototxt
    from oracle.caffemodel_io import read_caffemodel
    if threads:
        torch.set_num_threads(threads)
    Ls = prototxt.layers(prototxt.load(proto))
    W = read_caffemodel(cm)
    x = np.random.RandomState(2).uniform(0, 1, (frames, 3, 448, 448)).astype(np.float32)

    def step():
        o = net_ref.forward(Ls, W, x, want=['coverage', 'bboxes'])
        for b in range(frames):
            postnet_ref.postnet(o['coverage'][b], o['bboxes'][b], (448, 448, 3))
    times = []
    for _ in range(reps):
        t = time.perf_counter()
        step()
        times.append(time.perf_counter() - t)
    return times, torch.get_num_threads()


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
    times, threads = cpu_reference_fps(frames, args.steps, proto, cm, threads=ncpu)
    tot = sum(times)
    fps = frames * len(times) / tot
    line = {'impl': 'reference', 'metric': METRIC, 'value': fps, 'unit': UNIT, 'n_gpus': args.gpus, 'steps': args.steps,
            'warmup': args.warmup, 'ms_per_step': 1e3 * tot / len(times), 'higher_is_better': True, 'scaling': 'weak',
            'vs_baseline': None, 'dtype': 'f32', 'data': 'synthetic',
            'config': {'workload': 'deploy.prototxt 448x448 fwd+decode+NMS, CPU restatement of the reference '
                                   '(Caffe/Python-2 reference not runnable here), %d frames per step' % frames},
            'cpu_baseline': {'value': fps, 'unit': UNIT, 'cores': threads, 'kind': 'port',
                             'sample': '%d steps x %d frames, torch-CPU fp32 net + python post-net' % (args.steps, frames)},
            'e2e': {'value': fps, 'unit': UNIT, 'h2d_bytes_per_step': 0, 'd2h_bytes_per_step': 0},
            'gpu_launches': 0}
    print(json.dumps(line))


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
    B = BATCH
    # every rank gets its own frames (weak scaling: 64 frames per GPU per step); no data-path collective
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
    launches_per_step = L.fcnod_net_launches_per_forward(h) + 2

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
        _lib.check(L.fcnod_detect_graph_build(h, data.data_ptr(), B, C.byref(p), stream))   # restore batch-64 state
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
        run_post()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(reps):
            run_post()
        b.record()
        torch.cuda.synchronize()
        post_ms = a.elapsed_time(b) / reps
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
        tot_ms = sum(per) + post_ms
        ach = tc_flop / (tc_ms * 1e-3) / 1e12
        traffic, traffic_src = ncu_conv_traffic()
        roof = {'bound': 'tensor', 'kernel': 'conv_tc2_kernel (%d launches per step)' % n_tc, 'achieved': ach,
                'peak': pk['bf16_sustained'], 'peak_source': pk['which'] + ' sustained cuBLAS bf16', 'unit': 'TFLOP/s',
                'frac': ach / pk['bf16_sustained'], 'traffic': traffic, 'traffic_unit': 'bytes per launch (dram read + write, ncu)',
                'traffic_source': traffic_src, 'share_of_step': tc_ms / tot_ms,
                'avg_launch_ms': tc_ms / n_tc, 'flop_per_launch': tc_flop / n_tc}
        os.makedirs(os.path.join(ROOT, 'gpurun_out'), exist_ok=True)
        with open(os.path.join(ROOT, 'gpurun_out', 'per_launch_ms.json'), 'w') as f:
            json.dump([{'step': i, 'ms': t, 'desc': d} for i, (t, d) in enumerate(zip(per, plan))] +
                      [{'step': len(plan), 'ms': post_ms, 'desc': 'postnet decode_group_kernel + assemble_kernel'}], f, indent=1)
        # ---- CPU baseline: the oracle on the host cores, bounded sample
        cpu = None
        if world == 1 and not args.no_cpu_baseline:
            frames = 4
            ncpu = len(os.sched_getaffinity(0))
            cpu_reference_fps(frames, 1, proto, cm, threads=ncpu)
            times, threads = cpu_reference_fps(frames, 3, proto, cm, threads=ncpu)
            cpu = {'value': frames * len(times) / sum(times), 'unit': UNIT, 'cores': threads, 'kind': 'port',
                   'sample': '3 x %d frames (+1 warm-up), torch-CPU fp32 net + python post-net restatement' % frames}
        line = {'metric': METRIC, 'value': fps, 'unit': UNIT, 'n_gpus': world, 'steps': args.steps, 'warmup': args.warmup,
                'ms_per_step': ms / args.steps, 'higher_is_better': True, 'scaling': 'weak', 'vs_baseline': None,
                'dtype': 'bf16', 'data': 'synthetic',
                'config': {'workload': 'deploy.prototxt fwd + coverage decode + groupRectangles NMS, batch %d/GPU @448x448, '
                                       'bf16 tcgen05 convs, synthetic xavier weights (seed 0), one CUDA graph per step' % B,
                           'global_batch': world * B, 'parallelism': 'frame-parallel x%d, no collectives' % world,
                           'l2': 'inputs (154 MB/step) + activations (GBs) exceed the 126 MB L2; no flush needed'},
                'clocks': clocks,
                'e2e': {'value': e_fps, 'unit': UNIT, 'h2d_bytes_per_step': int(imgs.nbytes),
                        'd2h_bytes_per_step': int(B * (_lib.FCNOD_MAX_DET * 7 + 1) * 4),
                        'api': 'FCNObjectDetector.predict_stream(batches of uint8 BGR frames) -> fcnod_detect_submit / _collect, '
                               'two batches in flight', 'one_call_per_batch': {'value': sync_fps, 'unit': UNIT,
                               'api': 'FCNObjectDetector.predict_batch -> fcnod_detect'}, 'steps': e_steps,
                        'detections_last_step': n_det},
                'gpu_launches': launches_per_step * args.steps,
                'roofline': roof, 'cpu_baseline': cpu,
                'tflops_whole_step': world * B * args.steps * FLOP_PER_FRAME / (ms * 1e-3) / 1e12}
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
    ap.add_argument('--no-cpu-baseline', action='store_true')
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == 'ours' else args.warmup
    if args.impl == 'reference':
        run_reference(args)
    else:
        run_ours(args)


if __name__ == '__main__':
    main()
