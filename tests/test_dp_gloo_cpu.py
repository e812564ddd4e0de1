"""CPU, world size 2 over gloo: the data-parallel plumbing of ed3dgs_b200/dp.py (view sharding,
gradient arena whose slices are the parameters' .grad, one all-reduce per step, densification
reductions)."""
import os
import socket

import torch
import torch.distributed as dist
import torch.multiprocessing as mp


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world),
                      LOCAL_RANK=str(rank))
    from ed3dgs_b200 import dp
    r, _l, w = dp.init_from_env("gloo")
    assert (r, w) == (rank, world)
    torch.manual_seed(0)
    lin = torch.nn.Linear(4, 3)
    emb = torch.nn.Parameter(torch.ones(5, 2))
    arena = dp.GradArena(list(lin.parameters()) + [emb])
    assert lin.weight.grad.data_ptr() == arena.flat.data_ptr()
    views = dp.shard_views(5, rank, world)
    arena.zero()
    loss = sum((lin(torch.full((1, 4), float(v + 1))).sum() + (emb * (v + 1)).sum()) for v in views)
    loss.backward()
    flat = arena.allreduce(1.0 / 5)
    # every rank now holds the mean gradient over all 5 views
    expect_emb = sum(v + 1 for v in range(5)) / 5.0
    ok = torch.allclose(emb.grad, torch.full((5, 2), expect_emb)) and emb.grad.data_ptr() != 0
    expect_w = torch.full((3, 4), sum(v + 1 for v in range(5)) / 5.0)
    ok = ok and torch.allclose(lin.weight.grad, expect_w) and flat.numel() == 12 + 3 + 10
    radii, vis, g2d = dp.reduce_densification_stats(torch.tensor([rank, 3 - rank]), torch.tensor([rank == 0, False]),
                                                    torch.ones(2, 3) * (rank + 1))
    ok = ok and radii.tolist() == [1, 3] and vis.tolist() == [True, False] and float(g2d.sum()) == 18.0
    out[rank] = bool(ok) and views == ([0, 1, 2] if rank == 0 else [3, 4])
    dist.destroy_process_group()


def test_grad_arena_allreduce_world2():
    port = _free_port()
    mgr = mp.Manager()
    out = mgr.dict()
    mp.spawn(_worker, args=(2, port, out), nprocs=2, join=True)
    assert dict(out) == {0: True, 1: True}


def _worker_sync(rank, world, port, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world),
                      LOCAL_RANK=str(rank))
    from ed3dgs_b200 import dp
    dp.init_from_env("gloo")
    torch.manual_seed(0)
    lin = torch.nn.Linear(4, 3)
    emb = torch.nn.Parameter(torch.ones(5, 2))
    sync = dp.GradSync(list(lin.parameters()) + [emb])
    ok = True
    for _step in range(2):   # the second step checks that zero() really starts from nothing
        sync.zero()
        assert lin.weight.grad is None and emb.grad is None
        # the two Linear gradients arrive as slices of ONE buffer, like the deformation network's backward hands them over
        flat = torch.zeros(64 + 3)
        w = float(rank + 1)
        flat[:12] = w
        flat[64:67] = 2 * w
        lin.weight.grad = flat[:12].view(3, 4)
        lin.bias.grad = flat[64:67]
        (emb * w).sum().backward()
        ts = sync.tensors()
        ok = ok and len(ts) == 2 and ts[0].numel() == 67 and sync.nbytes() == 4 * (67 + 10)
        sync.allreduce(1.0 / world)
        ok = ok and torch.allclose(lin.weight.grad, torch.full((3, 4), 1.5)) and torch.allclose(lin.bias.grad, torch.full((3,), 3.0))
        ok = ok and torch.allclose(emb.grad, torch.full((5, 2), 1.5))
    out[rank] = bool(ok)
    dist.destroy_process_group()


def test_grad_sync_allreduce_world2():
    port = _free_port()
    mgr = mp.Manager()
    out = mgr.dict()
    mp.spawn(_worker_sync, args=(2, port, out), nprocs=2, join=True)
    assert dict(out) == {0: True, 1: True}


def test_shard_views_ragged():
    from ed3dgs_b200 import dp
    assert [dp.shard_views(8, r, 8) for r in range(8)] == [[i] for i in range(8)]
    assert dp.shard_views(3, 3, 4) == [] and dp.shard_views(3, 2, 4) == [2]


def _worker_overlap(rank, world, port, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world),
                      LOCAL_RANK=str(rank))
    from ed3dgs_b200 import dp
    dp.init_from_env("gloo")
    torch.manual_seed(0)
    xyz = torch.nn.Parameter(torch.ones(6, 3))
    emb = torch.nn.Parameter(torch.ones(6, 2))
    lin = torch.nn.Linear(2, 2)
    sync = dp.OverlapSync([xyz, emb] + list(lin.parameters()), early=[xyz], sm_reserve=8)
    ok = [p is xyz for p in sync.early] == [True] and len(sync.late) == 3
    ok = ok and sync.enable() is False          # no CUDA here: stays a plain two-bucket exchange
    for _step in range(2):
        sync.zero()
        w = float(rank + 1)
        ((xyz * w).sum() + (lin(emb) * w).sum()).backward()
        lw = lin.weight.detach().clone()
        ts = sync.allreduce(1.0 / world)
        ok = ok and len(ts) == 4
        ok = ok and torch.allclose(xyz.grad, torch.full((6, 3), 1.5))
        ok = ok and torch.allclose(emb.grad, (lw.sum(0) * 1.5).expand(6, 2))
        ok = ok and len(sync.allreduce(1.0 / world, collective=False)) == 4   # the join-only form bench.py uses
    sync.disable()
    out[rank] = bool(ok)
    dist.destroy_process_group()


def test_overlap_sync_world2():
    port = _free_port()
    mgr = mp.Manager()
    out = mgr.dict()
    mp.spawn(_worker_overlap, args=(2, port, out), nprocs=2, join=True)
    assert dict(out) == {0: True, 1: True}


def test_overlapped_backward_serialises_when_its_leaves_are_shared():
    """Host logic of deformation._must_serialise (the guard of the side-stream backward): a leaf that already holds a
    gradient, an earlier backward still parked, or the caller's shared_leaves declaration each force the join."""
    import weakref
    from ed3dgs_b200 import deformation as D
    a = torch.nn.Parameter(torch.zeros(3))
    b = torch.nn.Parameter(torch.zeros(3))
    refs = [weakref.ref(a), weakref.ref(b)]
    saved = dict(D._overlap)
    try:
        D._overlap.update(stream=None, shared_leaves=False, serialised=0)
        D._overlap["keep"].clear()
        assert not D._must_serialise(refs)
        b.grad = torch.ones(3)
        assert D._must_serialise(refs)            # AccumulateGrad would add in place, on the main stream
        b.grad = None
        D._overlap["keep"].append(("parked",))
        assert D._must_serialise(refs)            # second backward of the step
        D._overlap["keep"].clear()
        D._overlap["shared_leaves"] = True
        assert D._must_serialise(refs)
        D._overlap["shared_leaves"] = False
        del b
        assert not D._must_serialise(refs)        # a dead weak reference is not a hazard
    finally:
        D._overlap["keep"].clear()
        D._overlap.update({k: v for k, v in saved.items() if k != "keep"})
