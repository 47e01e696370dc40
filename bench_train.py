"""Benchmark of one training iteration (BASELINE config 5: `train_val.prototxt end-to-end training step batch 32/GPU with fused
augmentation layer, NCCL gradient allreduce`).  Run through bench.py:

    python bench.py --workload train --gpus N --steps K --warmup W
    python bench.py --workload train --impl reference ...          # the CPU restatement (torch autograd + ADAM), host cores

One "step" = DataArgumentationLayer forward on the device (32 sources of 480x640 -> `data` + five label grids) -> forward at
TRAIN phase -> fused loss head -> backward -> in-place all-reduce of the flat gradient buffer (NCCL, N > 1) -> ADAM
ApplyUpdate.  fp32 engine (csrc/train_net.cu).  value = images/s over all ranks, inputs resident; e2e = the same with the
sources uploaded from pinned host memory and the two losses read back every step.
"""
import ctypes as C
import json
import os
import random
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

BATCH = 32
SRC_H, SRC_W = 480, 640
NET = 448
METRIC = 'images/sec, train_val.prototxt training step @448x448 (augmentation layer + fwd + bwd + allreduce + ADAM)'
UNIT = 'images/s'
CONV_MACS_PER_FRAME = 7803973632      # SURVEY.md 8(a), deploy heads; the train heads (1 / 4 channels) are 12 MMAC smaller


def train_files():
    from fcn_object_detector_b200 import netspec
    from fcn_object_detector_b200.caffemodel import write_synthetic_caffemodel
    d = '/tmp/fcnod_bench_train_%d' % os.getpid()
    os.makedirs(d, exist_ok=True)
    proto = netspec.write_train_val_prototxt(os.path.join(d, 'train_val.prototxt'))
    cm = write_synthetic_caffemodel(os.path.join(d, 'init.caffemodel'), seed=0, num_cvg=1, num_bbox=4)
    return proto, cm


def layer_inputs(n, seed):
    """Sources, labels and host-made draws of the augmentation layer (BASELINE config 2's input, one bbox per image)."""
    from fcn_object_detector_b200 import _lib
    from fcn_object_detector_b200.data_argumentation_layer import draw_augmentation
    rs = np.random.RandomState(seed)
    rng = random.Random(seed)
    data = rs.randint(0, 256, (n, 3, SRC_H, SRC_W)).astype(np.float32)
    lab = np.zeros((n, 5), np.float32)
    arr = (_lib.AugDraw * n)()
    for i in range(n):
        w, h = int(rs.randint(64, 257)), int(rs.randint(64, 257))
        rect = [int(rs.randint(0, SRC_W - w)), int(rs.randint(0, SRC_H - h)), w, h]
        lab[i] = rect + [0]
        arr[i] = draw_augmentation(rng, SRC_W, SRC_H, rect)
    return data, lab, np.frombuffer(bytes(arr), dtype=np.uint8).copy()


def cpu_step_seconds(frames, reps, threads):
    """The oracle's training iteration (torch autograd over the restated graph + restated ADAM) on the host cores."""
    import torch
    from oracle import prototxt, train_net_ref
    from oracle.caffemodel_io import read_caffemodel
    torch.set_num_threads(threads)
    proto, cm = train_files()
    layers = prototxt.layers(prototxt.parse(open(proto).read()))
    W = {k: [np.asarray(a, np.float32) for a in v] for k, v in read_caffemodel(cm).items()}
    rs = np.random.RandomState(0)
    data = rs.uniform(0, 255, (frames, 3, NET, NET)).astype(np.float32)
    g = NET // 16
    grids = [rs.uniform(0, 1, (frames, c, g, g)).astype(np.float32) for c in (1, 4, 4, 4, 4)]
    mask = (rs.uniform(0, 1, (frames, 1024, g, g)) >= 0.4).astype(np.uint8)
    mults = train_net_ref.multipliers(layers)
    state, out = {}, []
    for it in range(reps + 1):
        t0 = time.perf_counter()
        r = train_net_ref.forward_backward(layers, W, data, grids, dropout_mask=mask)
        W = train_net_ref.adam_update(W, r['grads'], state, it, mults=mults)
        out.append(time.perf_counter() - t0)
    return out[1:]


def run_reference(args):
    if int(os.environ.get('RANK', '0')) != 0:
        return
    ncpu = len(os.sched_getaffinity(0))
    frames = 4
    times = cpu_step_seconds(frames, max(1, min(args.steps, 3)), ncpu)
    v = frames * len(times) / sum(times)
    print(json.dumps({
        'impl': 'reference', 'metric': METRIC, 'value': v, 'unit': UNIT, 'n_gpus': args.gpus, 'steps': len(times), 'warmup': 1,
        'ms_per_step': 1e3 * sum(times) / len(times), 'higher_is_better': True, 'scaling': 'weak', 'vs_baseline': None, 'dtype': 'f32',
        'data': 'synthetic',
        'config': {'workload': 'train_val.prototxt training step @448x448, CPU restatement of the reference (torch autograd fp32 + '
                               'restated ADAM), %d frames per step (bounded sample of the batch-32 step), label grids given' % frames},
        'cpu_baseline': {'value': v, 'unit': UNIT, 'cores': ncpu, 'kind': 'port',
                         'sample': '%d steps of %d frames (+1 warm-up)' % (len(times), frames)},
        'e2e': {'value': v, 'unit': UNIT, 'h2d_bytes_per_step': 0, 'd2h_bytes_per_step': 0}}))


def run_ours(args, section=False):
    """section=True: called from the default (detect) bench line at N = 1: returns a compact dict instead of printing."""
    import torch
    import torch.distributed as dist
    from fcn_object_detector_b200 import _lib
    from fcn_object_detector_b200.trainer import DataParallelSolver, Solver, allreduce_scale
    rank = int(os.environ.get('RANK', '0'))
    world = int(os.environ.get('WORLD_SIZE', '1'))
    local = int(os.environ.get('LOCAL_RANK', '0'))
    torch.cuda.set_device(local)
    if world > 1 and not dist.is_initialized():
        dist.init_process_group('nccl', device_id=torch.device('cuda', local))
    L = _lib.lib()
    proto, cm = train_files()
    B = BATCH
    prec = getattr(args, 'train_precision', 'bf16')
    solver = Solver(proto, cm, batch=B, height=NET, width=NET, device=local, seed=1 + rank, precision=prec)
    stream = torch.cuda.current_stream().cuda_stream

    src, lab, draws = layer_inputs(B, seed=10 + rank)
    h_src = [torch.from_numpy(src).pin_memory(), torch.from_numpy(np.roll(src, 1, axis=0).copy()).pin_memory()]
    h_lab, h_draw = torch.from_numpy(lab).pin_memory(), torch.from_numpy(draws).pin_memory()
    d_src, d_lab, d_draw = h_src[0].cuda(), h_lab.cuda(), h_draw.cuda()
    g = NET // 16
    tops = [torch.empty(s, device='cuda') for s in [(B, 3, NET, NET), (B, 1, g, g)] + [(B, 4, g, g)] * 4]
    rects = torch.empty((B, 4), dtype=torch.int32, device='cuda')
    scr = torch.empty(int(L.fcnod_aug_scratch_bytes(B)), dtype=torch.uint8, device='cuda')
    coll = getattr(args, 'train_collective', 'peer') if world > 1 else 'none'
    dp = DataParallelSolver(solver, collective=coll) if world > 1 else None
    losses_host = np.zeros(2, np.float32)

    def aug():
        _lib.check(L.fcnod_aug_forward_device(d_src.data_ptr(), d_lab.data_ptr(), d_draw.data_ptr(), B, SRC_H, SRC_W, NET, NET, 16, 1,
                                              *[t.data_ptr() for t in tops], rects.data_ptr(), scr.data_ptr(), stream))

    def fwd_bwd():
        solver.forward_backward_device([t.data_ptr() for t in tops], stream=stream)

    def update():
        solver.apply_update(1.0, stream=stream)

    def step():
        aug()
        if dp is not None:
            dp.step_device([t.data_ptr() for t in tops])       # forward + backward + collective + update
        else:
            fwd_bwd(); update()

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, n):
        barrier()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(n):
            fn()
        b.record()
        torch.cuda.synchronize()
        ms = a.elapsed_time(b)
        if world > 1:
            t = torch.tensor([ms], device='cuda')
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
        return ms

    steps, warm = (max(2, min(args.steps, 5)), 3) if section else (args.steps, max(args.warmup, 3))
    for _ in range(warm):
        step()
    sampler = None
    if rank == 0 and not section:
        from bench import ClockSampler
        sampler = ClockSampler(local)
    t0 = time.time()
    ms = timed(step, steps)
    t1 = time.time()
    clocks = sampler.stop(t0, t1) if sampler else None
    value = world * B * steps / (ms * 1e-3)
    l0 = solver.losses(stream)

    # end to end: sources from pinned host memory every step, losses back every step.  The upload of step i + 1 runs on a copy
    # stream into the other of two device buffers while step i computes (the detector's streaming form does the same).
    d_src2 = [d_src, torch.empty_like(d_src)]
    copy_stream = torch.cuda.Stream()
    up_ev = [torch.cuda.Event(), torch.cuda.Event()]
    free_ev = [torch.cuda.Event(), torch.cuda.Event()]

    def upload(i):
        b = i & 1
        with torch.cuda.stream(copy_stream):
            copy_stream.wait_event(free_ev[b])             # the step that read this buffer has finished with it
            d_src2[b].copy_(h_src[b], non_blocking=True)
            up_ev[b].record(copy_stream)

    def e2e_step(i, last):
        b = i & 1
        cur = torch.cuda.current_stream()
        cur.wait_event(up_ev[b])
        d_lab.copy_(h_lab, non_blocking=True)
        d_draw.copy_(h_draw, non_blocking=True)
        _lib.check(L.fcnod_aug_forward_device(d_src2[b].data_ptr(), d_lab.data_ptr(), d_draw.data_ptr(), B, SRC_H, SRC_W, NET, NET, 16, 1,
                                              *[t.data_ptr() for t in tops], rects.data_ptr(), scr.data_ptr(), stream))
        free_ev[b].record(cur)
        if not last:
            upload(i + 1)
        if dp is not None:
            dp.step_device([t.data_ptr() for t in tops])
        else:
            fwd_bwd(); update()
        _lib.check(L.fcnod_trainer_losses(solver.handle, _lib.ptr(losses_host), C.c_void_p(stream)))
    e_steps = max(2, min(steps, 10))
    for b in (0, 1):
        free_ev[b].record(torch.cuda.current_stream())
    upload(0); e2e_step(0, True)
    barrier()
    t0 = time.perf_counter()
    upload(0)
    for i in range(e_steps):
        e2e_step(i, i == e_steps - 1)
    torch.cuda.synchronize()
    e_ms = (time.perf_counter() - t0) * 1e3
    if world > 1:
        t = torch.tensor([e_ms], device='cuda')
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        e_ms = float(t.item())
    e_value = world * B * e_steps / (e_ms * 1e-3)
    h2d = int(h_src[0].numel() * 4 + h_lab.numel() * 4 + h_draw.numel())

    out = None
    if rank == 0:
        phases = {}
        if world == 1:        # (with several ranks a phase timed on rank 0 alone would leave the others out of the collective)
            for name, fn in (('augmentation_layer', aug), ('forward_backward', fwd_bwd), ('adam_update', update)):
                fn()
                phases[name + '_ms'] = timed(fn, 3) / 3
        # launches per step, counted from the layer list: forward 1 per layer that computes (+1 dropout mask), head 6,
        # backward 5 per convolution (mask, bias, wgrad, pack, dgrad; conv1 has no pack / dgrad, the heads no mask) + 1 per
        # other layer, 1 memset, 1 ADAM, 4 of the augmentation layer
        n_conv, n_pool, n_lrn = 59, 12, 2
        launches = (1 + n_conv + n_pool + n_lrn + 2 + 1) + 6 + (5 * n_conv - 2 - 2 + n_pool + n_lrn + 1 + 1) + 1 + 1 + 4
        if prec == 'bf16':
            # the ncu launch list of one bf16 step (profiles/r02zz_train_step_bf16_launches.csv: 228) + the layer's 4; since then the
            # ReLU / bias-gradient pass runs once per Concat blob instead of once per convolution: 9 blobs x (4 - 1) launches fewer
            launches = 228 + 4 - (0 if os.environ.get('FCNOD_TRAIN_RB_FUSE', '1') == '0' else 27)
        flop = 3 * 2 * CONV_MACS_PER_FRAME * B            # forward + data gradient + weight gradient (conv1 has no dgrad: -4%)
        from bench import peaks
        pk = peaks()
        gemm = None
        if prec == 'bf16':
            # roofline of the dominant kernel: every GEMM launch of one step timed with CUDA events on the launching stream
            # (3 steps, the fastest sum per kind), algorithmic FLOP = 2 M N K per launch
            runs = [solver.gemm_times(fwd_bwd) for _ in range(3)]
            gemm = {k: min((r[k] for r in runs), key=lambda d: d['ms']) for k in runs[0]}
            g_ms = sum(v['ms'] for v in gemm.values())
            g_flop = sum(v['flop'] for v in gemm.values())
            g_n = sum(v['launches'] for v in gemm.values())
        fb_ms = phases['forward_backward_ms'] if world == 1 else ms / steps
        out = {
            'metric': METRIC, 'value': value, 'unit': UNIT, 'n_gpus': world, 'steps': steps, 'warmup': warm, 'ms_per_step': ms / steps,
            'higher_is_better': True, 'scaling': 'weak', 'vs_baseline': None, 'dtype': 'f32' if prec == 'fp32' else 'bf16', 'data': 'synthetic',
            'config': {'workload': 'train_val.prototxt training step, batch %d/GPU @448x448: DataArgumentationLayer forward on the '
                                   'device (480x640 sources) + TRAIN forward + loss head + backward + %s + ADAM (solver.prototxt '
                                   'values), %s' % (B, ('NCCL all-reduce of %d fp32 gradients' % solver.n_params if coll == 'nccl' else 'reduce-scatter + ADAM + all-gather of %d fp32 parameters in one kernel over NVLink peer memory' % solver.n_params) if world > 1
                                                                         else 'no all-reduce (1 GPU)',
                                                                         'fp32 SIMT parity engine' if prec == 'fp32' else
                                                                         'bf16 engine: convolutions as tcgen05 GEMMs (forward, dgrad, wgrad), fp32 master weights'),
                       'l2': 'activations of a step (>6 GB) exceed L2; no flush needed'},
            'losses_after_timed_steps': l0,
            'e2e': {'value': e_value, 'unit': UNIT, 'h2d_bytes_per_step': h2d, 'd2h_bytes_per_step': 8, 'steps': e_steps,
                    'api': 'pinned host sources / labels / draws -> fcnod_aug_forward_device -> fcnod_trainer_forward_backward_device '
                           '-> %s -> fcnod_trainer_losses' % {'none': 'fcnod_trainer_apply_update', 'nccl': 'all_reduce -> fcnod_trainer_apply_update',
                                                               'peer': 'fcnod_trainer_apply_update_peer'}[coll]},
            'collective': coll,
            'gpu_launches': launches * steps,
            'phases': phases,
            'roofline': ({'bound': 'tensor', 'kernel': 'gemm_tc_kernel (tcgen05; %d launches per step: forward %d, data gradient %d, weight '
                                                       'gradient %d)' % (g_n, gemm['forward']['launches'], gemm['dgrad']['launches'], gemm['wgrad']['launches']),
                          'achieved': g_flop / (g_ms * 1e-3) / 1e12, 'peak': pk['bf16_burst'], 'unit': 'TFLOP/s',
                          'frac': g_flop / (g_ms * 1e-3) / 1e12 / pk['bf16_burst'], 'peak_source': pk['which'] + ', burst (launches timed one by one)',
                          'avg_launch_ms': g_ms / g_n, 'flop_per_launch': g_flop / g_n, 'share_of_step': g_ms / (ms / steps),
                          'by_kind': {k: dict(v, tflops=v['flop'] / (v['ms'] * 1e-3) / 1e12) for k, v in gemm.items()},
                          'whole_step_achieved': flop / (fb_ms * 1e-3) / 1e12, 'whole_step_frac': flop / (fb_ms * 1e-3) / 1e12 / pk['bf16_burst'],
                          'traffic': None,
                          'note': 'achieved = algorithmic FLOP (2 M N K) of the GEMM launches of one step / their summed CUDA-event times; '
                                  'whole_step = 3 x forward convolution FLOP x batch over the whole forward + backward time (im2col, pooling, '
                                  'LRN and the other bandwidth launches included)'} if gemm else
                         {'bound': 'tensor', 'kernel': 'conv_simt_kernel / wgrad_simt_kernel (fp32 FFMA, not on the tensor pipe)',
                          'achieved': flop / (fb_ms * 1e-3) / 1e12, 'peak': pk['bf16_burst'], 'unit': 'TFLOP/s',
                          'frac': flop / (fb_ms * 1e-3) / 1e12 / pk['bf16_burst'], 'traffic': None,
                          'note': 'fp32 parity engine: 3 x forward convolution FLOP x batch over the forward + backward time'}),
        }
        if clocks:
            out['clocks'] = clocks
        if not args.no_cpu_baseline:
            ncpu = len(os.sched_getaffinity(0))
            frames = 2
            t = cpu_step_seconds(frames, 1 if section else 2, ncpu)
            out['cpu_baseline'] = {'value': frames * len(t) / sum(t), 'unit': UNIT, 'cores': ncpu, 'kind': 'port',
                                   'sample': '%d step(s) of %d frames (+1 warm-up), torch autograd fp32 + restated ADAM' % (len(t), frames)}
    solver.close()
    if section:
        return out
    if rank == 0:
        print(json.dumps(out))
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
