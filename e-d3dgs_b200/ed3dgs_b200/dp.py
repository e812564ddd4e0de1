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
