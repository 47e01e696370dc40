"""Training iterations of `models/train_val.prototxt` on the B200 path (BASELINE config 5).

Host mirror of what `caffe train --solver=bounding_box/solver.prototxt` (/root/reference/train/train.sh:19) does per
iteration: Net::ForwardBackward over the network at TRAIN phase and the ADAM ApplyUpdate of
/root/reference/train/bounding_box/solver.prototxt:3-15, with Caffe's NCCL solver's gradient all-reduce when several
processes train together (one process per GPU, `torch.distributed`).  The arithmetic is in csrc/train_net.cu behind the C
ABI (`fcnod_trainer_*`); there is no CPU path.

    solver = Solver(train_val_prototxt, snapshot_caffemodel, batch=32)      # reads solver.prototxt values by default
    layer.forward(bottom, top)                                              # DataArgumentationLayer: data + 5 label grids
    losses = solver.step(top[0].data, [t.data for t in top[1:]])            # one iteration on this GPU
    solver.snapshot('snapshot_iter_%d.caffemodel' % solver.iter)

Data-parallel: `DataParallelSolver` wraps a Solver inside an initialised process group: every rank runs forward + backward
on its own minibatch, the flat gradient buffer is all-reduced in place (SUM) and ApplyUpdate scales it by 1 / world_size.
"""
import ctypes as C
import re

import numpy as np

from . import _lib


def read_solver_prototxt(path):
    """The fields of a solver.prototxt the trainer uses (train/bounding_box/solver.prototxt:3-15).  Only ADAM with
    lr_policy "step" (or "fixed") is implemented: anything else is refused rather than trained differently."""
    text = open(path).read()
    vals = {}
    for line in text.splitlines():
        line = line.split('#', 1)[0].strip()
        m = re.match(r'^(\w+)\s*:\s*(.+)$', line)
        if m:
            vals[m.group(1)] = m.group(2).strip().strip('"')
    typ = vals.get('solver_type', vals.get('type', 'SGD')).upper()
    if typ != 'ADAM':
        raise ValueError('solver type %r is not supported (only ADAM, as train/bounding_box/solver.prototxt uses)' % typ)
    policy = vals.get('lr_policy', 'fixed')
    if policy not in ('step', 'fixed'):
        raise ValueError('lr_policy %r is not supported (step / fixed)' % policy)
    if int(vals.get('iter_size', 1)) != 1:
        raise ValueError('iter_size != 1 is not supported')
    if float(vals.get('clip_gradients', -1)) >= 0:
        raise ValueError('clip_gradients is not supported')
    if vals.get('regularization_type', 'L2') != 'L2':
        raise ValueError('only L2 regularization is supported')
    return dict(base_lr=float(vals['base_lr']), gamma=float(vals.get('gamma', 1.0)) if policy == 'step' else 1.0,
                stepsize=int(vals.get('stepsize', 0)) if policy == 'step' else 0,
                momentum=float(vals.get('momentum', 0.9)), momentum2=float(vals.get('momentum2', 0.999)),
                delta=float(vals.get('delta', 1e-8)), weight_decay=float(vals.get('weight_decay', 0.0)),
                max_iter=int(vals.get('max_iter', 0)), snapshot=int(vals.get('snapshot', 0)),
                snapshot_prefix=vals.get('snapshot_prefix', 'snapshot'))


class _DeviceArray(object):
    """A device buffer owned by the trainer, exposed through __cuda_array_interface__ so that torch can wrap it without
    copying (`torch.as_tensor(obj, device='cuda')`): the all-reduce runs in place on the trainer's gradient buffer."""

    def __init__(self, ptr, n, owner):
        self._owner = owner
        self.__cuda_array_interface__ = {'shape': (int(n),), 'typestr': '<f4', 'data': (int(ptr), False), 'version': 2,
                                         'strides': None}


class Solver(object):
    def __init__(self, prototxt, caffemodel, batch, height=448, width=448, device=0, solver_prototxt=None, seed=1, precision='fp32',
                 **overrides):
        """precision: 'fp32' (parity engine) or 'bf16' (tensor cores: bf16 activations and gradients, fp32 master weights)."""
        L = _lib.lib()
        sp = _lib.SolverParams()
        _lib.check(L.fcnod_solver_defaults(C.byref(sp)))
        if solver_prototxt:
            for k, v in read_solver_prototxt(solver_prototxt).items():
                if hasattr(sp, k):
                    setattr(sp, k, v)
        for k, v in overrides.items():
            if not hasattr(sp, k):
                raise TypeError('unknown solver parameter %r' % k)
            setattr(sp, k, v)
        sp.seed = seed
        sp.precision = {'fp32': 0, 'bf16': 1}[precision]
        self.precision = precision
        self.params = sp
        self.handle = C.c_void_p()
        _lib.check(L.fcnod_trainer_create(str(prototxt).encode(), str(caffemodel).encode(), device, batch, height, width,
                                          C.byref(sp), C.byref(self.handle)))
        dims = (C.c_int * 7)()
        n = C.c_size_t()
        _lib.check(L.fcnod_trainer_info(self.handle, dims, C.byref(n)))
        self.batch, self.channels, self.height, self.width, self.num_classes, self.grid_h, self.grid_w = [int(d) for d in dims]
        self.n_params = int(n.value)
        self.device = device
        ptrs = [C.c_void_p() for _ in range(4)]
        _lib.check(L.fcnod_trainer_buffers(self.handle, *[C.byref(p) for p in ptrs]))
        self.param_ptr, self.grad_ptr, self.m_ptr, self.v_ptr = [int(p.value) for p in ptrs]
        self.layers = []
        i = 0
        while True:
            name = L.fcnod_trainer_param_layer(self.handle, i)
            if not name:
                break
            self.layers.append(name.decode())
            i += 1

    # ---- the iteration
    @property
    def iter(self):
        return int(_lib.lib().fcnod_trainer_iteration(self.handle))

    def _check_tops(self, data, grids):
        gh, gw = self.grid_h, self.grid_w
        want = [(self.batch, self.channels, self.height, self.width), (self.batch, self.num_classes, gh, gw)] + \
               [(self.batch, 4 * self.num_classes, gh, gw)] * 4
        arrs = [np.ascontiguousarray(a, dtype=np.float32) for a in [data] + list(grids)]
        for a, w in zip(arrs, want):
            if tuple(a.shape) != w:
                raise ValueError('trainer: blob of shape %s where the train_val.prototxt contract wants %s' % (a.shape, w))
        return arrs

    def step(self, data, grids):
        """One solver iteration from HOST arrays (the layer's six tops).  Returns dict(loss_bbox, loss_coverage, loss)."""
        arrs = self._check_tops(data, grids)
        losses = np.zeros(2, np.float32)
        _lib.check(_lib.lib().fcnod_trainer_step(self.handle, *[_lib.ptr(a) for a in arrs], _lib.ptr(losses)))
        return self._loss_dict(losses)

    def forward_backward_device(self, ptrs, stream=0):
        """ptrs: six device pointers (data + the five grids).  Gradients land in the flat gradient buffer."""
        _lib.check(_lib.lib().fcnod_trainer_forward_backward_device(self.handle, *[C.c_void_p(int(p)) for p in ptrs], C.c_void_p(stream)))

    def apply_update(self, grad_scale=1.0, stream=0):
        _lib.check(_lib.lib().fcnod_trainer_apply_update(self.handle, C.c_float(grad_scale), C.c_void_p(stream)))

    def losses(self, stream=0):
        losses = np.zeros(2, np.float32)
        _lib.check(_lib.lib().fcnod_trainer_losses(self.handle, _lib.ptr(losses), C.c_void_p(stream)))
        return self._loss_dict(losses)

    def _loss_dict(self, losses):
        w = float(self.params.bbox_loss_weight)
        return dict(loss_bbox=float(losses[0]), loss_coverage=float(losses[1]), loss=float(losses[1]) + w * float(losses[0]))

    def gemm_times(self, run_step):
        """Time the GEMM launches of one step (bf16 engine): run_step() must run one forward + backward.  Returns
        {'forward' | 'dgrad' | 'wgrad': dict(ms, flop, launches)} summed over the launches of each kind."""
        L = _lib.lib()
        _lib.check(L.fcnod_trainer_set_timing(self.handle, 1))
        try:
            run_step()
            ms = (C.c_float * 3)()
            fl = (C.c_double * 3)()
            n = (C.c_int * 3)()
            _lib.check(L.fcnod_trainer_gemm_times(self.handle, ms, fl, n))
        finally:
            _lib.check(L.fcnod_trainer_set_timing(self.handle, 0))
        return {k: dict(ms=float(ms[i]), flop=float(fl[i]), launches=int(n[i])) for i, k in enumerate(('forward', 'dgrad', 'wgrad'))}

    def grad_buffer(self):
        """The flat gradient buffer as an object torch can wrap in place (see DataParallelSolver)."""
        return _DeviceArray(self.grad_ptr, self.n_params, self)

    # ---- introspection
    def param(self, layer, index, which='value'):
        code = {'value': 0, 'grad': 1, 'm': 2, 'v': 3}[which]
        n = C.c_size_t(0)
        _lib.check(_lib.lib().fcnod_trainer_param_read(self.handle, layer.encode(), index, code, None, C.byref(n)))
        out = np.zeros(int(n.value), np.float32)
        _lib.check(_lib.lib().fcnod_trainer_param_read(self.handle, layer.encode(), index, code, _lib.ptr(out), C.byref(n)))
        return out

    def blob(self, name, which='data'):
        dims = (C.c_int * 3)()
        L = _lib.lib()
        _lib.check(L.fcnod_trainer_blob_read(self.handle, name.encode(), 0 if which == 'data' else 1, dims, None, 0))
        out = np.zeros((self.batch, dims[0], dims[1], dims[2]), np.float32)
        _lib.check(L.fcnod_trainer_blob_read(self.handle, name.encode(), 0 if which == 'data' else 1, dims, _lib.ptr(out), out.size))
        return out

    def dropout_mask(self):
        dims = (C.c_int * 3)()
        L = _lib.lib()
        _lib.check(L.fcnod_trainer_dropout_mask(self.handle, None, 0, dims))
        out = np.zeros((self.batch, dims[0], dims[1], dims[2]), np.uint8)
        _lib.check(L.fcnod_trainer_dropout_mask(self.handle, _lib.ptr(out), out.size, dims))
        return out

    def pool_argmax(self, layer):
        """The element the last forward selected in every window of a Pooling layer (dh * k + dw, unclipped window)."""
        dims = (C.c_int * 3)()
        L = _lib.lib()
        _lib.check(L.fcnod_trainer_pool_argmax(self.handle, layer.encode(), None, 0, dims))
        out = np.zeros((self.batch, dims[0], dims[1], dims[2]), np.uint8)
        _lib.check(L.fcnod_trainer_pool_argmax(self.handle, layer.encode(), _lib.ptr(out), out.size, dims))
        return out

    def weights(self, shapes=None):
        """{layer: [W (Cout,Cin,k,k), b]} as a .caffemodel holds them."""
        out = {}
        for name in self.layers:
            b = self.param(name, 1)
            w = self.param(name, 0)
            k2cin = w.size // b.size
            if shapes and name in shapes:
                w = w.reshape(shapes[name])
            else:
                w = w.reshape(b.size, k2cin)
            out[name] = [w, b]
        return out

    def snapshot(self, path, shapes):
        """Write the current weights as a .caffemodel (what Caffe's Solver::Snapshot leaves for the detector to load).
        shapes: {layer: (Cout, Cin, k, k)} (netspec.conv_shape_dict gives them for the shipped network)."""
        from .caffemodel import write_caffemodel
        w = self.weights(shapes)
        write_caffemodel(path, [(name, 'Convolution', w[name]) for name in self.layers])
        return path

    def close(self):
        if self.handle:
            _lib.lib().fcnod_trainer_destroy(self.handle)
            self.handle = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def allreduce_scale(world_size):
    """Caffe's NCCL solver sums the gradients of the solvers and scales by 1 / solver_count before ApplyUpdate."""
    return 1.0 / float(world_size)


class DataParallelSolver(object):
    """One process per GPU inside an initialised `torch.distributed` group (NCCL on the GPUs).  Every rank holds a full
    replica and runs forward + backward on its own minibatch; then, per iteration,

    collective='nccl': ONE in-place all-reduce (SUM) of the flat gradient buffer, ApplyUpdate with grad_scale = 1 / world_size
                       (Caffe's NCCL solver's rule);
    collective='peer': ONE kernel per rank over NVLink peer memory (the ranks' gradient and parameter buffers are mapped into
                       every process through CUDA IPC): the rank sums its shard of all gradient buffers, updates the shard
                       (ADAM moments live only on the owner) and stores the new parameters into every replica -- reduce-scatter,
                       update and all-gather without an intermediate buffer.  Two one-element all-reduces bracket it as
                       stream-ordered barriers.  One node, at most 8 ranks.

    Ranks start from the same snapshot, so replicas stay identical (bit for bit with 'peer': every replica receives the same
    stored values)."""

    def __init__(self, solver, group=None, collective='nccl'):
        import torch
        import torch.distributed as dist
        self.solver = solver
        self.dist = dist
        self.group = group
        self.world = dist.get_world_size(group)
        self.rank = dist.get_rank(group)
        self.collective = collective if self.world > 1 else 'nccl'
        self.grads = torch.as_tensor(solver.grad_buffer(), device='cuda:%d' % solver.device)
        if self.collective == 'peer':
            L = _lib.lib()
            hp, hg = C.create_string_buffer(64), C.create_string_buffer(64)
            _lib.check(L.fcnod_trainer_ipc_handles(solver.handle, hp, hg))
            mine = (bytes(hp.raw), bytes(hg.raw))
            allh = [None] * self.world
            dist.all_gather_object(allh, mine, group=group)
            ps = b''.join(h[0] for h in allh)
            gs = b''.join(h[1] for h in allh)
            _lib.check(L.fcnod_trainer_open_peers(solver.handle, self.world, self.rank, ps, gs))
            self._flag = torch.zeros(1, device='cuda:%d' % solver.device)
        elif self.collective != 'nccl':
            raise ValueError("collective must be 'nccl' or 'peer'")

    def _barrier(self):
        # stream-ordered: the next kernel on this stream starts only after every rank has reached this point on its stream
        self.dist.all_reduce(self._flag, op=self.dist.ReduceOp.SUM, group=self.group)

    def step_device(self, ptrs):
        import torch
        s = torch.cuda.current_stream().cuda_stream
        self.solver.forward_backward_device(ptrs, stream=s)
        if self.collective == 'peer':
            self._barrier()                  # every rank's gradients are complete
            _lib.check(_lib.lib().fcnod_trainer_apply_update_peer(self.solver.handle, C.c_void_p(s)))
            self._barrier()                  # every rank's shard has been stored into this replica
            return
        if self.world > 1:
            self.dist.all_reduce(self.grads, op=self.dist.ReduceOp.SUM, group=self.group)
        self.solver.apply_update(allreduce_scale(self.world), stream=s)
