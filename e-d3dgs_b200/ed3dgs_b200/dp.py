"""View-batch data parallelism (new relative to the reference, which is single-GPU:
train.py:171-187 loops over the views of a batch on one device).

One process per GPU (torchrun), Gaussians + deformation network replicated, the views of a step
split across ranks, and ONE all-reduce(sum) per step over a flat fp32 gradient arena:
per Gaussian xyz 3 + f_dc 3 + f_rest 45 + opacity 1 + scaling 3 + rotation 4 + embedding 32 = 91
floats, plus the deformation network. Parameters' .grad tensors are views into the arena, so
autograd accumulates straight into it and no flatten/unflatten copy exists. The only collective is
this all-reduce: the per-view path itself needs no exchange."""
import os

import torch
import torch.distributed as dist


def init_from_env(backend=None):
    """Initialises torch.distributed from torchrun's env (RANK/LOCAL_RANK/WORLD_SIZE/MASTER_*)."""
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world > 1 and not dist.is_initialized():
        if backend is None:
            backend = "nccl" if torch.cuda.is_available() else "gloo"
        if backend == "nccl":
            torch.cuda.set_device(local)
        dist.init_process_group(backend=backend, rank=rank, world_size=world)
    return rank, local, world


class GradArena:
    """Flat gradient buffer whose slices are the parameters' .grad."""

    def __init__(self, params):
        self.params = [p for p in params if p.requires_grad]
        total = sum(p.numel() for p in self.params)
        dev = self.params[0].device
        self.flat = torch.zeros(total, dtype=torch.float32, device=dev)
        off = 0
        for p in self.params:
            n = p.numel()
            p.grad = self.flat[off:off + n].view_as(p)
            off += n

    def zero(self):
        self.flat.zero_()

    def nbytes(self):
        return self.flat.numel() * 4

    def allreduce(self, scale=None):
        """Sum over ranks (no-op for world size 1); optional scale (1/global batch)."""
        if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
            world = dist.get_world_size()
            if scale is not None and abs(scale * world - 1.0) < 1e-12 and dist.get_backend() == "nccl":
                dist.all_reduce(self.flat, op=dist.ReduceOp.AVG)   # the mean in the collective: no second pass
                return self.flat
            dist.all_reduce(self.flat, op=dist.ReduceOp.SUM)
        if scale is not None and scale != 1.0:
            self.flat.mul_(scale)
        return self.flat


class GradSync:
    """Per-step gradient exchange WITHOUT a flat arena. zero() drops the .grad tensors (the reference's own
    ``optimizer.zero_grad(set_to_none=True)``, train.py:448), so autograd hands over the tensors the backward kernels
    wrote instead of adding them into preallocated ones (no zero fill of 364 B per Gaussian, no accumulate kernels).
    allreduce() reduces every gradient in ONE NCCL group (ncclGroupStart / ncclGroupEnd: a single fused collective
    launch); gradients that are slices of one allocation (all Linear parameters of the deformation network come out
    of the backward as slices of one buffer) travel as that one allocation."""

    def __init__(self, params):
        self.params = [p for p in params if p.requires_grad]

    def zero(self):
        for p in self.params:
            p.grad = None

    def tensors(self):
        """The distinct allocations behind the current .grad tensors, in parameter order."""
        seen, out = {}, []
        for p in self.params:
            g = p.grad
            if g is None:
                continue
            st = g.untyped_storage()
            key = st.data_ptr()
            if key in seen:
                continue
            seen[key] = True
            if g.is_contiguous() and g.storage_offset() == 0 and g.numel() * g.element_size() == st.nbytes():
                out.append(g)
            else:   # a slice of a larger buffer: reduce the whole buffer once
                out.append(torch.empty(0, dtype=g.dtype, device=g.device).set_(st))
        return out

    def nbytes(self):
        return sum(t.numel() * t.element_size() for t in self.tensors())

    def allreduce(self, scale=None):
        ts = self.tensors()
        if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
            world = dist.get_world_size()
            avg = scale is not None and abs(scale * world - 1.0) < 1e-12 and dist.get_backend() == "nccl"
            op = dist.ReduceOp.AVG if avg else dist.ReduceOp.SUM
            done = False
            if dist.get_backend() == "nccl" and ts:
                try:
                    from torch.distributed.distributed_c10d import _coalescing_manager
                    with _coalescing_manager(device=ts[0].device, async_ops=False):
                        for t in ts:
                            dist.all_reduce(t, op=op)
                    done = True
                except (ImportError, TypeError, RuntimeError):
                    done = False
            if not done:
                for t in ts:
                    dist.all_reduce(t, op=op)
            if avg:
                return ts
        if scale is not None and scale != 1.0:
            for t in ts:
                t.mul_(scale)
        return ts


class OverlapSync(GradSync):
    """GradSync with the exchange split in two and the first half hidden behind the deformation network's backward.

    Bucket A = the gradients of the per-Gaussian rasterizer inputs (xyz, f_dc, f_rest, opacity, scaling, rotation:
    59 of the 91 floats per Gaussian). They are the *upstream* gradients of the deformation node, i.e. final once the
    rasterizer backward and the activations backward have run. Bucket B = embedding + deformation network.
    With the overlap on, `deformation._DeformFunction.backward` issues its kernels on a side stream with `sm_reserve`
    SMs left free; allreduce() then (1) reduces bucket A -- NCCL's stream only waits for the main stream, so the
    collective runs next to the MLP backward --, (2) makes the main stream wait for the side stream, (3) reduces bucket B.
    Host logic only: without CUDA (gloo tests) or with one rank it degrades to GradSync."""

    def __init__(self, params, early, sm_reserve=8):
        super().__init__(params)
        early_ids = {id(p) for p in early}
        self.early = [p for p in self.params if id(p) in early_ids]
        self.late = [p for p in self.params if id(p) not in early_ids]
        self.sm_reserve = sm_reserve
        self.side = None
        self.active = False

    def enable(self):
        """Switch the side stream on (needs CUDA and more than one rank); returns whether it is active."""
        from . import deformation
        multi = dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1
        if multi and torch.cuda.is_available() and self.params and self.params[0].is_cuda:
            self.side = torch.cuda.Stream(device=self.params[0].device)
            deformation.set_backward_stream(self.side, self.sm_reserve)
            self.active = True
        return self.active

    def disable(self):
        from . import deformation
        self.join()
        deformation.set_backward_stream(None)
        self.side, self.active = None, False

    def join(self):
        """Main stream waits for the side stream; releases what the side stream was using."""
        from . import deformation
        if self.side is not None:
            torch.cuda.current_stream(self.side.device).wait_stream(self.side)
        deformation._overlap["keep"].clear()

    def _reduce(self, params, scale, collective):
        keep = self.params
        self.params = params
        try:
            return GradSync.allreduce(self, scale) if collective else self.tensors()
        finally:
            self.params = keep

    def allreduce(self, scale=None, collective=True):
        """collective=False: only the stream join (bench's 'steps without the collective' breakdown)."""
        if not self.active:
            return GradSync.allreduce(self, scale) if collective else self.tensors()
        a = self._reduce(self.early, scale, collective)
        self.join()
        b = self._reduce(self.late, scale, collective)
        return a + b


def shard_views(n_views, rank, world):
    """Indices of the views rank handles in a step of n_views (contiguous, ragged tail allowed)."""
    per = (n_views + world - 1) // world
    lo = min(n_views, rank * per)
    return list(range(lo, min(n_views, lo + per)))


def reduce_densification_stats(radii, visibility, viewspace_grad):
    """The three extra reductions densification needs (train.py:189-190,346-348): max of radii,
    or of visibility, sum of the screen-space gradient."""
    if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
        dist.all_reduce(radii, op=dist.ReduceOp.MAX)
        v = visibility.to(torch.int32)
        dist.all_reduce(v, op=dist.ReduceOp.MAX)
        visibility = v.bool()
        dist.all_reduce(viewspace_grad, op=dist.ReduceOp.SUM)
    return radii, visibility, viewspace_grad
