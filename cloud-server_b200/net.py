"""Host-side mirror of the reference's graph interface (construct_distribute.py:208-311) over the
C ABI: ``Net(net_config)`` interprets the same JSON ``cnn()`` does; forward / backward / optimizer
steps run in the CUDA library.  PyTorch supplies device tensors, streams and (for data parallel
training) the NCCL process group -- plumbing only."""
import ctypes as C
import json

import numpy as np
import torch

from . import _lib
from .context import default_context

LOSS = {"entropy": 0, "mse": 1}
# optimizer_name values of the UI (apps/construction/util/options.py:22-31) -> library ids.
# "compat" = what the reference actually runs whatever the config says: AdagradOptimizer(1e-4)
# (construct_distribute.py:372-373).
OPT = {"AdagradOptimizer": 0, "Adagrad": 0, "adagrad": 0, "GradientDescentOptimizer": 1, "SGD": 1, "sgd": 1,
       "AdamOptimizer": 2, "Adam": 2, "adam": 2}


class _DevMem:
    """Exposes library-owned device memory to torch through __cuda_array_interface__."""

    def __init__(self, ptr, count, owner):
        self._owner = owner
        self.__cuda_array_interface__ = {"shape": (int(count),), "typestr": "<f4", "data": (int(ptr), False),
                                         "version": 2, "strides": None}


class Net:
    def __init__(self, net_config, ctx=None, in_h=28, in_w=28, in_c=1, classes=10, seed=1):
        self.ctx = ctx or default_context()
        if not isinstance(net_config, str):
            net_config = json.dumps(net_config)
        self.config = net_config
        self.classes, self.in_features = classes, in_h * in_w * in_c
        h = C.c_void_p()
        L = _lib.lib()
        _lib.check(self.ctx.handle, L.nc_net_create(self.ctx.handle, net_config.encode(), in_h, in_w, in_c, classes,
                                                    C.c_uint64(seed), C.byref(h)))
        self.handle = h
        self.param_count = int(L.nc_net_param_count(h))
        self.tensors = []
        for i in range(L.nc_net_tensor_count(h)):
            off, size, nd = C.c_int64(), C.c_int64(), C.c_int()
            dims = (C.c_int64 * 4)()
            _lib.check(self.ctx.handle, L.nc_net_tensor_info(h, i, C.byref(off), C.byref(size), C.byref(nd), C.byref(dims)))
            self.tensors.append((off.value, size.value, tuple(dims[k] for k in range(nd.value))))
        dev = self.ctx.tdev
        self.params = torch.as_tensor(_DevMem(L.nc_net_params(h), self.param_count, self), device=dev)
        self.grads = torch.as_tensor(_DevMem(L.nc_net_grads(h), self.param_count, self), device=dev)
        self._loss = torch.zeros(1, dtype=torch.float32, device=dev)
        self._correct = torch.zeros(1, dtype=torch.int32, device=dev)

    def close(self):
        if getattr(self, "handle", None):
            _lib.lib().nc_net_destroy(self.handle)
            self.handle = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---- parameters (flat, variable-creation order; conv HWIO, dense [in][out])
    def set_params(self, flat):
        flat = np.ascontiguousarray(np.asarray(flat, np.float32).reshape(-1))
        if flat.size != self.param_count:
            raise ValueError("expected %d parameters, got %d" % (self.param_count, flat.size))
        _lib.check(self.ctx.handle, _lib.lib().nc_net_set_params(self.handle, flat.ctypes.data_as(C.c_void_p)))

    def get_params(self):
        out = np.empty(self.param_count, np.float32)
        _lib.check(self.ctx.handle, _lib.lib().nc_net_get_params(self.handle, out.ctypes.data_as(C.c_void_p)))
        return out

    def profile(self, fn):
        """Run fn() (forward / backward / train_step calls on this net) with per-stage CUDA-event timing on:
        returns [(stage name, milliseconds), ...] in launch order."""
        L = _lib.lib()
        _lib.check(self.ctx.handle, L.nc_net_profile(self.handle, 1))
        try:
            fn()
            ms = (C.c_float * 512)()
            names = C.create_string_buffer(1 << 16)
            cnt = C.c_int()
            _lib.check(self.ctx.handle, L.nc_net_profile_read(self.handle, self.ctx._stream(), ms, 512, names, len(names), C.byref(cnt)))
        finally:
            L.nc_net_profile(self.handle, 0)
        nm = names.value.decode().split("\n")
        return [(nm[i], float(ms[i])) for i in range(min(cnt.value, 512))]

    def set_math(self, math):
        _lib.check(self.ctx.handle, _lib.lib().nc_net_set_math(self.handle, int(math)))

    @property
    def global_step(self):
        return int(_lib.lib().nc_net_global_step(self.handle))

    @property
    def arena_generation(self):
        """changes whenever the library moved its activation arena / workspaces (a larger batch came along)"""
        return int(_lib.lib().nc_net_arena_generation(self.handle))

    def reserve(self, max_batch):
        _lib.check(self.ctx.handle, _lib.lib().nc_net_reserve(self.handle, int(max_batch)))

    def get_state(self, opt):
        slots = _lib.lib().nc_net_state_slots(self.handle, opt)
        out = np.empty(slots * self.param_count, np.float32)
        if slots:
            _lib.check(self.ctx.handle, _lib.lib().nc_net_get_state(self.handle, opt, out.ctypes.data_as(C.c_void_p)))
        return out

    def set_state(self, opt, state, global_step):
        p = None
        if state is not None and len(state):
            state = np.ascontiguousarray(state, np.float32)
            p = state.ctypes.data_as(C.c_void_p)
        _lib.check(self.ctx.handle, _lib.lib().nc_net_set_state(self.handle, opt, p, int(global_step)))

    # ---- compute
    def _x(self, x):
        if x.dtype != torch.float32 or not x.is_cuda:
            raise ValueError("x must be a float32 CUDA tensor [N, %d]" % self.in_features)
        x = x.reshape(x.shape[0], -1).contiguous()
        if x.shape[1] != self.in_features:
            raise ValueError("x has %d features, the net takes %d" % (x.shape[1], self.in_features))
        return x

    def forward(self, x):
        x = self._x(x)
        out = torch.empty((x.shape[0], self.classes), dtype=torch.float32, device=x.device)
        _lib.check(self.ctx.handle, _lib.lib().nc_net_forward(self.handle, C.c_void_p(x.data_ptr()), x.shape[0],
                                                              C.c_void_p(out.data_ptr()), self.ctx._stream()))
        return out

    def predict(self, x):
        x = self._x(x)
        out = torch.empty((x.shape[0],), dtype=torch.int32, device=x.device)
        _lib.check(self.ctx.handle, _lib.lib().nc_net_predict(self.handle, C.c_void_p(x.data_ptr()), x.shape[0],
                                                              C.c_void_p(out.data_ptr()), self.ctx._stream()))
        return out

    def backward(self, x, y, loss_name="entropy", grad_scale=1.0):
        """forward + loss + gradients into self.grads; returns (loss[1], correct[1]) device tensors."""
        x = self._x(x)
        y = y.reshape(x.shape[0], -1).contiguous()
        if y.dtype != torch.float32 or y.shape[1] != self.classes:
            raise ValueError("y must be float32 one-hot [N, %d]" % self.classes)
        if loss_name not in LOSS:
            raise ValueError("loss_function returns None for %r" % (loss_name,))
        _lib.check(self.ctx.handle, _lib.lib().nc_net_backward(
            self.handle, C.c_void_p(x.data_ptr()), C.c_void_p(y.data_ptr()), x.shape[0], LOSS[loss_name],
            C.c_float(grad_scale), C.c_void_p(self._loss.data_ptr()), C.c_void_p(self._correct.data_ptr()), self.ctx._stream()))
        return self._loss, self._correct

    def apply(self, opt="compat", lr=1e-4):
        oid, lr = (0, 1e-4) if opt == "compat" else (OPT[opt] if isinstance(opt, str) else int(opt), lr)
        _lib.check(self.ctx.handle, _lib.lib().nc_net_apply(self.handle, oid, C.c_float(lr), self.ctx._stream()))

    def train_step(self, x, y, loss_name="entropy", opt="compat", lr=1e-4):
        """One sess.run(train_op) (construct_distribute.py:413).  On a net attached to a communicator
        (attach_nccl) x, y are this rank's slice of the global batch and every rank ends with the
        weights of the global-batch step; the returned loss / hit count stay per-rank."""
        if self.world > 1:
            # one C call: backward with dLoss / world, the tail of the gradient buffer (the dense layers) all-reduced on a
            # side stream while the convolution gradients are still being computed, the head after, then the update
            x = self._x(x)
            y = y.reshape(x.shape[0], -1).contiguous()
            if y.dtype != torch.float32 or y.shape[1] != self.classes:
                raise ValueError("y must be float32 one-hot [N, %d]" % self.classes)
            if loss_name not in LOSS:
                raise ValueError("loss_function returns None for %r" % (loss_name,))
            oid, lr = (0, 1e-4) if opt == "compat" else (OPT[opt] if isinstance(opt, str) else int(opt), lr)
            _lib.check(self.ctx.handle, _lib.lib().nc_net_train_step(
                self.handle, C.c_void_p(x.data_ptr()), C.c_void_p(y.data_ptr()), x.shape[0], LOSS[loss_name], oid, C.c_float(lr),
                C.c_void_p(self._loss.data_ptr()), C.c_void_p(self._correct.data_ptr()), self.ctx._stream()))
            return self._loss, self._correct
        r = self.backward(x, y, loss_name, 1.0)
        self.apply(opt, lr)
        return r

    # ---- data parallel inside the library (nc_net_attach_nccl)
    world = 1
    rank = 0

    def attach_nccl(self, unique_id, rank, world, libnccl_path=None):
        """Collective: joins this net to an NCCL communicator of `world` processes (one per GPU).
        `unique_id` is the 128 bytes rank 0 got from new_nccl_unique_id()."""
        if len(unique_id) != 128:
            raise ValueError("an NCCL unique id is 128 bytes")
        buf = (C.c_ubyte * 128).from_buffer_copy(bytes(unique_id))
        path = (libnccl_path or find_libnccl()).encode()
        with torch.cuda.device(self.ctx.tdev):
            _lib.check(self.ctx.handle, _lib.lib().nc_net_attach_nccl(self.handle, path, buf, int(rank), int(world)))
        self.world, self.rank = int(world), int(rank)

    def attach_p2p(self, gather):
        """Map the other ranks' gradient buffers (CUDA IPC) so that the gradient sum runs as the library's own NVLink
        peer-memory kernel instead of ncclAllReduce.  `gather(bytes) -> [bytes per rank, rank order]` is the host's channel
        (DataParallel passes torch.distributed.all_gather_object).  Returns False (and stays on NCCL) where the GPUs
        cannot map each other."""
        buf = (C.c_ubyte * 128)()
        _lib.check(self.ctx.handle, _lib.lib().nc_net_p2p_handles(self.handle, buf))
        parts = gather(bytes(buf))
        blob = b"".join(parts)
        arr = (C.c_ubyte * len(blob)).from_buffer_copy(blob)
        rc = _lib.lib().nc_net_p2p_attach(self.handle, arr, int(self.rank), int(self.world))
        if rc == _lib.NC_EUNSUPPORTED:
            return False
        _lib.check(self.ctx.handle, rc)
        return True

    def p2p_timed_out(self):
        v = C.c_int(0)
        _lib.check(self.ctx.handle, _lib.lib().nc_net_p2p_status(self.handle, C.byref(v)))
        return bool(v.value)

    def detach_nccl(self):
        _lib.check(self.ctx.handle, _lib.lib().nc_net_detach_nccl(self.handle))
        self.world = 1

    def allreduce_grads(self):
        _lib.check(self.ctx.handle, _lib.lib().nc_net_allreduce_grads(self.handle, self.ctx._stream()))


class GraphedTrainStep:
    """One sess.run(train_op) (construct_distribute.py:413) captured into a CUDA graph and replayed: the ~30 kernel
    launches of a LeNet step cost one graph launch.  Fixed batch size, loss and optimizer; Adagrad (what the reference
    runs) or SGD only -- Adam's step size is a host-side function of global_step.  On a net attached to a communicator
    (attach_nccl) the captured step contains the NCCL all-reduces of the gradient buffer (issued on the library's side
    stream, which joins the capture through events): every rank must build and replay its graph in step with the others.

        step = GraphedTrainStep(net, 256)
        loss, hits = step(x, y)          # x, y: CUDA tensors [256, features], [256, classes]; copied into the graph's inputs
    """

    def __init__(self, net, batch, loss_name="entropy", opt="compat", lr=1e-4):
        oid = 0 if opt == "compat" else (OPT[opt] if isinstance(opt, str) else int(opt))
        if oid == 2:
            raise ValueError("Adam's bias-corrected step size changes every step on the host: it cannot be replayed from a graph")
        self.net, self.batch = net, int(batch)
        dev = net.ctx.tdev
        self.x = torch.zeros((self.batch, net.in_features), dtype=torch.float32, device=dev)
        self.y = torch.zeros((self.batch, net.classes), dtype=torch.float32, device=dev)
        self.y[:, 0] = 1.0
        # one eager step on a side stream sizes every workspace (nothing may allocate during capture); its effect on
        # the weights, the optimizer state and the step counter is undone afterwards
        w0, st0, gs0 = net.params.clone(), net.get_state(oid), net.global_step
        side = torch.cuda.Stream(device=dev)
        side.wait_stream(torch.cuda.current_stream(dev))
        with torch.cuda.stream(side):
            net.train_step(self.x, self.y, loss_name, opt, lr)
        torch.cuda.current_stream(dev).wait_stream(side)
        torch.cuda.synchronize(dev)
        net.params.copy_(w0)
        net.set_state(oid, st0, gs0)
        self._oid, self._args = oid, (loss_name, opt, lr)
        self._capture()

    def _capture(self):
        """(Re)capture: the graph holds raw device addresses of the net's activation arena and workspaces, which move
        when a later forward / predict / backward brings a larger batch (nc_net_arena_generation)."""
        net, gs0 = self.net, self.net.global_step
        self.graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(self.graph):
            self.loss, self.correct = net.train_step(self.x, self.y, *self._args)
        net.set_state(self._oid, None, gs0)      # the captured call bumped the host-side counter without running anything
        self._gen = net.arena_generation

    def __call__(self, x, y):
        if self.net.arena_generation != self._gen:
            self._capture()                # the buffers only ever grow: they are large enough for this batch, just elsewhere
        self.x.copy_(x.reshape(self.batch, -1), non_blocking=True)
        self.y.copy_(y.reshape(self.batch, -1), non_blocking=True)
        self.graph.replay()
        _lib.check(self.net.ctx.handle, _lib.lib().nc_net_advance_step(self.net.handle, 1))
        return self.loss, self.correct

    # ---- batches that live in (pinned) host memory: the copy of batch i + 1 runs on a copy stream under the step on batch i
    def prefetch(self, hx, hy):
        """Start the host -> device copy of a coming batch (pinned CPU tensors [batch, features], [batch, classes]) on the
        copy stream.  At most two batches may be in flight; run_prefetched() consumes them in order."""
        dev = self.net.ctx.tdev
        if getattr(self, "_copy", None) is None:
            self._copy = torch.cuda.Stream(device=dev)
            self._stage = [(torch.empty_like(self.x), torch.empty_like(self.y)) for _ in range(2)]
            self._ready = [torch.cuda.Event(), torch.cuda.Event()]
            self._free = [None, None]
            self._fill, self._queue = 0, []
        if len(self._queue) >= 2:
            raise RuntimeError("two batches are already in flight: call run_prefetched() first")
        slot = self._fill
        self._fill ^= 1
        if self._free[slot] is not None:
            self._copy.wait_event(self._free[slot])          # the step that read this staging slot has copied it out
        with torch.cuda.stream(self._copy):
            self._stage[slot][0].copy_(hx.reshape(self.batch, -1), non_blocking=True)
            self._stage[slot][1].copy_(hy.reshape(self.batch, -1), non_blocking=True)
            self._ready[slot].record(self._copy)
        self._queue.append(slot)

    def run_prefetched(self):
        """One step on the oldest prefetched batch; returns (loss, correct) device tensors like __call__."""
        if not getattr(self, "_queue", None):
            raise RuntimeError("no batch in flight: call prefetch(hx, hy) first")
        slot = self._queue.pop(0)
        main = torch.cuda.current_stream(self.net.ctx.tdev)
        main.wait_event(self._ready[slot])
        out = self(self._stage[slot][0], self._stage[slot][1])
        if self._free[slot] is None:
            self._free[slot] = torch.cuda.Event()
        self._free[slot].record(main)
        return out


def find_libnccl():
    """Path of the NCCL this process should share with torch: the one already mapped if torch loaded it,
    else the wheel's copy, else the loader's default."""
    import os
    try:
        with open("/proc/self/maps") as f:
            for line in f:
                if "libnccl.so" in line:
                    return line.split()[-1]
    except OSError:
        pass
    try:
        import nvidia.nccl as _n
        for d in list(getattr(_n, "__path__", [])):
            cand = os.path.join(d, "lib", "libnccl.so.2")
            if os.path.exists(cand):
                return cand
    except ImportError:
        pass
    return "libnccl.so.2"


def new_nccl_unique_id(ctx=None, libnccl_path=None):
    """128-byte NCCL unique id (call on rank 0, hand the bytes to every rank)."""
    ctx = ctx or default_context()
    buf = (C.c_ubyte * 128)()
    _lib.check(ctx.handle, _lib.lib().nc_nccl_unique_id(ctx.handle, (libnccl_path or find_libnccl()).encode(), buf))
    return bytes(buf)


class DataParallel:
    """Synchronous data parallel training, one process per GPU: every rank holds the same weights
    and a contiguous slice of the global batch; gradients are summed with one NCCL all-reduce of
    the flat gradient buffer (loss is a batch mean, so each rank scales its dLoss by 1/world).
    The reference's own 'distribution' is a 1 ps + 1 worker gRPC parameter server
    (construct_distribute.py:344-357, cmd.py:63-69); the parity target is the single-device step at
    the global batch."""

    def __init__(self, net, group=None, allow_local_norm=False, in_library=None):
        """in_library: sum the gradients (and the batch-norm moments) with the library's own NCCL communicator
        instead of torch.distributed.  Default: only when the net has 'norm' layers, which need it --
        normalize() uses batch statistics (construct_distribute.py:155-165) and per-rank moments would differ
        from the single-device step at the global batch.  allow_local_norm=True keeps per-rank moments."""
        import torch.distributed as dist
        self.net, self.dist, self.group = net, dist, group
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        cfg = getattr(net, "config", None)
        has_norm = bool(cfg and '"norm"' in cfg)
        if in_library is None:
            in_library = has_norm and not allow_local_norm
        self.in_library = bool(in_library) and self.world > 1
        self.p2p = False          # True: gradient sums run as the library's NVLink peer-memory kernel (in_library only)
        src0 = dist.get_global_rank(group, 0) if (group is not None and self.world > 1) else 0
        if self.world > 1:
            dist.broadcast(net.params, src=src0, group=group)  # identical initial weights
        if self.in_library:
            if not hasattr(net, "attach_nccl"):
                raise NotImplementedError("data parallel training of a net with 'norm' layers needs the library's "
                                          "communicator (Net.attach_nccl); pass allow_local_norm=True for per-rank moments")
            rank = dist.get_rank(group)
            box = [new_nccl_unique_id(net.ctx) if rank == 0 else None]
            dist.broadcast_object_list(box, src=src0, group=group)
            net.attach_nccl(box[0], rank, self.world)
            import os
            if self.world <= 8 and not os.environ.get("NC_DP_NO_P2P"):
                def gather(b):
                    parts = [None] * self.world
                    dist.all_gather_object(parts, b, group=group)
                    return parts
                ok = net.attach_p2p(gather)
                # every rank must take the same path
                flag = torch.tensor([1 if ok else 0], device=net.ctx.tdev)
                dist.all_reduce(flag, op=dist.ReduceOp.MIN, group=group)
                self.p2p = bool(flag.item())
                if not self.p2p and ok:
                    raise RuntimeError("peer mapping succeeded on this rank but not on all: set NC_DP_NO_P2P=1")

    def train_step(self, x_local, y_local, loss_name="entropy", opt="compat", lr=1e-4):
        if self.in_library:
            return self.net.train_step(x_local, y_local, loss_name, opt, lr)
        loss, correct = self.net.backward(x_local, y_local, loss_name, 1.0 / self.world)
        if self.world > 1:
            self.dist.all_reduce(self.net.grads, group=self.group)
        self.net.apply(opt, lr)
        return loss, correct
