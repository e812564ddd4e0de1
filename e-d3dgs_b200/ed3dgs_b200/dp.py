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
            manager = None
            if dist.get_backend() == "nccl" and ts:
                # Only the availability of the (private) coalescing manager is probed. Once a collective has been
                # enqueued an error must surface: retrying tensor by tensor would reduce some gradients twice on
                # this rank and leave the other ranks with mismatched collectives.
                try:
                    from torch.distributed.distributed_c10d import _coalescing_manager
                    manager = _coalescing_manager(device=ts[0].device, async_ops=False)
                except (ImportError, TypeError):
                    manager = None
            if manager is not None:
                with manager:
                    for t in ts:
                        dist.all_reduce(t, op=op)
            else:
                for t in ts:
                    dist.all_reduce(t, op=op)
            if avg:
                return ts
        if scale is not None and scale != 1.0:
            for t in ts:
                t.mul_(scale)
        return ts


class CopyEngineExchange:
    """Mean over ranks of a list of tensors whose TRANSFERS use no SM: a reduce-scatter and an all-gather made of
    peer-to-peer `cudaMemcpyAsync` copies (NVLink copy engines) between symmetric-memory buffers, with one small reduction
    kernel in between. It is what can run next to persistent kernels that own every SM: NCCL's
    all-reduce needs tens of CTAs to reach its bandwidth.

        pack      the tensors, back to back, into this rank's symmetric buffer `buf`                      (local copies)
        barrier   every rank's buffer is packed
        pull      chunk `rank` of every peer's buffer into `stage[peer]`                                  (P2P copies)
        reduce    buf[chunk rank] = mean over peers of stage            (nobody pulls that chunk of MY buffer: no barrier)
        barrier   every rank's reduced chunk is in place, all pulls of the first phase have completed
        pull      every peer's reduced chunk into the same chunk of `buf`                                 (P2P copies)
        barrier   nobody repacks its buffer while a peer still reads it

    Everything is issued on `self.stream`; the results are views of `buf` (valid until the next call)."""

    def __init__(self, max_floats, device, group=None):
        import torch.distributed._symmetric_memory as symm
        group = group if group is not None else dist.group.WORLD
        self.world, self.rank = dist.get_world_size(group), dist.get_rank(group)
        self.align = self.world * 64                       # chunks start on 256-byte boundaries
        cap = (int(max_floats) + self.align - 1) // self.align * self.align
        self.cap = cap
        self.buf = symm.empty(cap, dtype=torch.float32, device=device)
        self.hdl = symm.rendezvous(self.buf, group)
        self.peers = [self.hdl.get_buffer(r, (cap,), torch.float32) for r in range(self.world)]
        self.stage = torch.empty(self.world, cap // self.world, dtype=torch.float32, device=device)
        self.stream = torch.cuda.Stream(device=device)
        self.timeout_ms = 120000  # a rank that never arrives makes the barrier kernel trap instead of hanging the box

    def allreduce_mean(self, tensors, like=None):
        """like: tensors that fix the layout (one per entry of `tensors`, same order on every rank). An entry of
        `tensors` may then be None -- this rank has no contribution -- and travels as zeros, so that chunk size, offsets
        and the number of barriers never depend on rank-local state (a rank with a missing gradient would otherwise
        make its peers average unrelated floats, or leave them in a barrier until the time-out traps)."""
        like = list(like) if like is not None else list(tensors)
        if len(like) != len(tensors) or any(t is None for t in like):
            raise ValueError("CopyEngineExchange: `like` needs one tensor per entry")
        for t, ref in zip(tensors, like):
            if t is not None and t.numel() != ref.numel():
                raise ValueError(f"CopyEngineExchange: gradient of {t.numel()} floats for a slot of {ref.numel()}")
        total = sum(ref.numel() for ref in like)
        used = (total + self.align - 1) // self.align * self.align
        if used > self.cap:
            raise ValueError(f"CopyEngineExchange: {total} floats do not fit the symmetric buffer of {self.cap}")
        chunk = used // self.world
        rank, world, buf = self.rank, self.world, self.buf
        self.stream.wait_stream(torch.cuda.current_stream(buf.device))
        outs = []
        with torch.cuda.stream(self.stream):
            off = 0
            for t, ref in zip(tensors, like):
                n = ref.numel()
                if t is None:
                    buf[off:off + n].zero_()
                else:
                    buf[off:off + n].copy_(t.detach().reshape(-1))
                outs.append(buf[off:off + n].view(ref.shape))
                off += n
            if used > total:
                buf[total:used].zero_()
            self.hdl.barrier(channel=0, timeout_ms=self.timeout_ms)
            mine = slice(rank * chunk, (rank + 1) * chunk)
            for step in range(world):
                r = (rank - step) % world
                self.stage[r, :chunk].copy_(self.peers[r][mine])
            torch.mean(self.stage[:, :chunk], dim=0, out=buf[mine])
            self.hdl.barrier(channel=0, timeout_ms=self.timeout_ms)
            for step in range(1, world):
                r = (rank - step) % world
                theirs = slice(r * chunk, (r + 1) * chunk)
                buf[theirs].copy_(self.peers[r][theirs])
            self.hdl.barrier(channel=0, timeout_ms=self.timeout_ms)
        return outs


def deformation_ran():
    """True when a deformation backward of this step went to the side stream (its tensors are parked until the join)."""
    from . import deformation
    return len(deformation._overlap["keep"]) > 0


class OverlapSync(GradSync):
    """GradSync with the exchange split up and most of it hidden behind the deformation network's backward.

    Bucket A = the gradients of the per-Gaussian rasterizer inputs (xyz, f_dc, f_rest, opacity, scaling, rotation:
    59 of the 91 floats per Gaussian). They are the *upstream* gradients of the deformation node, i.e. final once the
    rasterizer backward and the activations backward have run. Bucket B = embedding + deformation network.
    With the overlap on, `deformation._DeformFunction.backward` issues its kernels on a side stream with `sm_reserve`
    SMs left free; allreduce() then (1) exchanges bucket A -- through CopyEngineExchange (transport "ce": copy engines
    over NVLink, no SMs) or through NCCL (transport "nccl": its stream only waits for the main stream) -- next to the MLP
    backward, (1b, "ce" with `embedding`) exchanges the embedding gradient from the event the library records once it is
    final, (2) makes the main stream wait for the side stream, (3) reduces what is left of bucket B through NCCL.
    Measurements and the choice of defaults: DESIGN.md section 6.
    Host logic only: without CUDA (gloo tests) or with one rank it degrades to GradSync."""

    def __init__(self, params, early, sm_reserve=8, transport="nccl", embedding=None, sm_reserve_last=8,
                 shared_leaves=False):
        """shared_leaves: set it when the loss has other terms on the leaves of the deformation node (the reference's
        embedding regulariser and temporal smoothness term, train.py:229-242). The overlap is only sound for ONE
        deformation backward per step whose leaves receive nothing else and start the step with .grad = None (the
        gradients leave the node while the side stream still writes them); a second backward before join() or a leaf
        that already holds a .grad is detected and serialised by the node itself, other loss terms cannot be, hence
        this flag (deformation.py, comment above `_overlap`).
        transport: "nccl" = bucket A through NCCL next to the backward (needs `sm_reserve` SMs and a matching
        NCCL_MAX_CTAS); "ce" = bucket A through CopyEngineExchange (no SMs for the transfers).
        embedding (transport "ce" only): the per-Gaussian embedding parameter. Its gradient is final after the chain kernel
        of the network's last stage, one weight-gradient launch before the backward ends; the library records an event
        there (ed3dgs_deform_set_emb_ready_event) and a second exchange starts from it, so that only the network's own
        parameters (a few MB) are left for the end of the step."""
        super().__init__(params)
        early_ids = {id(p) for p in early}
        self.early = [p for p in self.params if id(p) in early_ids]
        self.embedding = embedding if (embedding is not None and transport == "ce") else None
        self.late = [p for p in self.params if id(p) not in early_ids and p is not self.embedding]
        self.ce_emb = None
        self.emb_event = None
        self.sm_reserve_last = sm_reserve_last
        self.sm_reserve = sm_reserve
        self.shared_leaves = bool(shared_leaves)
        self.transport = transport
        self.side = None
        self.ce = None
        self.active = False

    def enable(self):
        """Switch the side stream on (needs CUDA and more than one rank); returns whether it is active."""
        from . import deformation
        multi = dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1
        if multi and torch.cuda.is_available() and self.params and self.params[0].is_cuda:
            dev = self.params[0].device
            if self.transport == "ce":
                self.ce = CopyEngineExchange(sum(p.numel() for p in self.early) + 64 * len(self.early), dev)
                if self.embedding is not None:
                    from . import _lib
                    import ctypes
                    self.ce_emb = CopyEngineExchange(self.embedding.numel(), dev)
                    with torch.cuda.device(dev):     # the library keeps event and reserves per (current) device
                        self.emb_event = torch.cuda.Event()
                        self.emb_event.record(torch.cuda.current_stream(dev))      # creates the cudaEvent_t
                        _lib.lib().ed3dgs_deform_set_emb_ready_event(ctypes.c_void_p(self.emb_event.cuda_event))
                        _lib.lib().ed3dgs_deform_set_sm_reserve_last(int(self.sm_reserve_last))
            self.side = torch.cuda.Stream(device=dev)
            # with the copy-engine transport nothing reduces the pass-through gradients in place: no copies needed
            deformation.set_backward_stream(self.side, self.sm_reserve, clone=self.ce is None,
                                            shared_leaves=self.shared_leaves)
            self.active = True
        return self.active

    def disable(self):
        from . import deformation
        self.join()
        deformation.set_backward_stream(None)
        self._clear_event()
        self.side, self.active = None, False

    def _clear_event(self):
        """The library holds the raw cudaEvent_t: take it back before the torch.cuda.Event can be destroyed."""
        if self.emb_event is not None:
            from . import _lib
            with torch.cuda.device(self.ce_emb.stream.device):
                _lib.lib().ed3dgs_deform_set_emb_ready_event(None)
                _lib.lib().ed3dgs_deform_set_sm_reserve_last(0)
            self.emb_event = None

    def __del__(self):
        try:   # an OverlapSync dropped without disable() must not leave a dangling event handle in the library
            self._clear_event()
        except Exception:  # noqa: BLE001  (interpreter shutdown: the library may be gone already)
            pass

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
        if self.ce is not None and collective:
            world = dist.get_world_size()
            if scale is None or abs(scale * world - 1.0) > 1e-12:
                raise ValueError("OverlapSync(transport='ce') computes the mean over ranks: pass scale = 1 / world")
            # Layout and route are the same on every rank by construction: every early parameter has its slot
            # (zeros where this rank has no gradient), and the embedding travels through ce_emb whenever that exchange
            # exists (set up collectively in enable()), whether or not the deformation ran here.
            ps = list(self.early)
            outs = self.ce.allreduce_mean([p.grad for p in ps], like=ps)       # on the exchange's own stream
            late = self.late
            if self.ce_emb is not None:
                if self.embedding.grad is not None and deformation_ran():
                    # starts when the library's event fires (after the last stage's chain kernel), not at the end of
                    # the backward
                    self.ce_emb.stream.wait_event(self.emb_event)
                elif self.side is not None:
                    # no side-stream backward on this rank this step (or one that was serialised): whatever wrote the
                    # gradient is ordered by the main and side streams, not by the library's event
                    self.ce_emb.stream.wait_stream(self.side)
                outs += self.ce_emb.allreduce_mean([self.embedding.grad], like=[self.embedding])
                ps = ps + [self.embedding]
            elif self.embedding is not None:
                late = late + [self.embedding]
            self.join()
            main = torch.cuda.current_stream(self.ce.stream.device)
            main.wait_stream(self.ce.stream)
            if self.ce_emb is not None:
                main.wait_stream(self.ce_emb.stream)
            for p, o in zip(ps, outs):
                p.grad = o
            return outs + self._reduce(late, scale, collective)
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
