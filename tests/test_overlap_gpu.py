"""Single-GPU checks of the two mechanisms dp.OverlapSync relies on (the NCCL side needs two devices:
tests/dp_overlap_check.py under torchrun): the SM reserve of the persistent kernels and the deformation backward issued
on a side stream must both leave every gradient unchanged (up to the arrival order of fp32 atomics)."""
import pytest
import torch

import test_deform_gpu as T

pytestmark = pytest.mark.gpu


def _grads(rec):
    net, finals, subs, leaves, _ = T.run_product(rec)
    out = {"in_" + k: v.grad.detach().clone() for k, v in leaves.items()}
    out.update({"p_" + k: p.grad.detach().clone() for k, p in net.named_parameters() if p.grad is not None})
    out.update({f"final{i}": x.detach().clone() for i, x in enumerate(finals)})
    return out


def _compare(a, b, tol=2e-5):
    assert a.keys() == b.keys()
    for k in a:
        scale = float(a[k].abs().max()) + 1e-30
        err = float((a[k] - b[k]).abs().max()) / scale
        assert err <= tol, (k, err)


@pytest.mark.parametrize("shape", [dict(N=40_000 + 5, W=128, D=1), dict(N=3000, W=64, D=2)])
def test_sm_reserve_leaves_results_unchanged(shape):
    from ed3dgs_b200 import _lib
    rec = T.oracle_record(kink_free=True, **shape)
    ref = _grads(rec)
    lib = _lib.lib()
    assert lib.ed3dgs_deform_set_sm_reserve(8) == 0
    try:
        got = _grads(rec)
    finally:
        assert lib.ed3dgs_deform_set_sm_reserve(0) == 8
    _compare(ref, got)


@pytest.mark.parametrize("shape", [dict(N=40_000 + 5, W=128, D=1), dict(N=3000, W=64, D=2)])
def test_backward_on_side_stream_leaves_results_unchanged(shape):
    from ed3dgs_b200 import deformation
    rec = T.oracle_record(kink_free=True, **shape)
    ref = _grads(rec)
    side = torch.cuda.Stream()
    deformation.set_backward_stream(side, sm_reserve=8)
    try:
        from ed3dgs_b200.deformation import deform_network  # noqa: F401
        net, finals, subs, leaves, _ = T.run_product(rec)   # ends with a device-wide synchronize
        assert len(deformation._overlap["keep"]) == 1
        torch.cuda.current_stream().wait_stream(side)
        got = {"in_" + k: v.grad.detach().clone() for k, v in leaves.items()}
        got.update({"p_" + k: p.grad.detach().clone() for k, p in net.named_parameters() if p.grad is not None})
        got.update({f"final{i}": x.detach().clone() for i, x in enumerate(finals)})
    finally:
        deformation.set_backward_stream(None)
    _compare(ref, got)


def test_copy_engine_exchange_single_rank():
    """dp.CopyEngineExchange on a one-rank NCCL group (the driver's GPU box has one device; tests/dp_overlap_check.py and
    tests/dp_sweep.sh run it on 2-8 ranks): pack, barriers, reduction kernel and the views handed back. Skipped where
    symmetric memory cannot be set up."""
    import os
    import socket
    import torch.distributed as dist
    from ed3dgs_b200 import dp
    if dist.is_initialized():
        pytest.skip("a process group already exists in this process")
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    keys = ("MASTER_ADDR", "MASTER_PORT", "RANK", "WORLD_SIZE", "LOCAL_RANK")
    saved = {k: os.environ.get(k) for k in keys}
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK="0", WORLD_SIZE="1", LOCAL_RANK="0")
    dist.init_process_group("nccl", rank=0, world_size=1)
    try:
        try:
            ex = dp.CopyEngineExchange(10_000, torch.device("cuda:0"))
        except Exception as e:  # noqa: BLE001
            pytest.skip(f"symmetric memory unavailable: {type(e).__name__}: {e}")
        g = torch.Generator(device="cuda").manual_seed(1)
        ts = [torch.randn(1000, 3, device="cuda", generator=g), torch.randn(77, device="cuda", generator=g),
              torch.randn(5, 16, 3, device="cuda", generator=g)]
        outs = ex.allreduce_mean(ts)
        torch.cuda.current_stream().wait_stream(ex.stream)
        torch.cuda.synchronize()
        for a, b in zip(ts, outs):
            assert a.shape == b.shape and torch.equal(a, b)
        with pytest.raises(ValueError):
            ex.allreduce_mean([torch.zeros(20_000, device="cuda")])
        # fixed layout: a gradient this rank does not have keeps its slot and travels as zeros
        outs = ex.allreduce_mean([ts[0], None, ts[2]], like=ts)
        torch.cuda.current_stream().wait_stream(ex.stream)
        torch.cuda.synchronize()
        assert [o.shape for o in outs] == [t.shape for t in ts]
        assert torch.equal(outs[0], ts[0]) and torch.equal(outs[2], ts[2]) and not outs[1].any()
        with pytest.raises(ValueError):
            ex.allreduce_mean([ts[0], None], like=ts)
    finally:
        dist.destroy_process_group()
        for k, v in saved.items():
            if v is None:
                os.environ.pop(k, None)
            else:
                os.environ[k] = v


def test_second_backward_before_the_join_is_serialised():
    """The overlap is only sound for one deformation backward per step (deformation.py, comment above `_overlap`):
    a second one before the join, a leaf that already holds a .grad, or a caller that declares other loss terms on the
    same leaves must make the main stream wait for the side stream before the gradients leave the node."""
    from ed3dgs_b200 import deformation
    rec = T.oracle_record(kink_free=True, N=3000, W=64, D=2)
    ref = _grads(rec)
    side = torch.cuda.Stream()
    deformation.set_backward_stream(side, sm_reserve=8)
    try:
        T.run_product(rec)
        assert deformation._overlap["serialised"] == 0 and len(deformation._overlap["keep"]) == 1
        net, finals, subs, leaves, _ = T.run_product(rec)      # no join in between
        assert deformation._overlap["serialised"] == 1 and len(deformation._overlap["keep"]) == 2
        got = {"in_" + k: v.grad.detach().clone() for k, v in leaves.items()}     # main stream only: already ordered
        got.update({"p_" + k: p.grad.detach().clone() for k, p in net.named_parameters() if p.grad is not None})
        got.update({f"final{i}": x.detach().clone() for i, x in enumerate(finals)})
        _compare(ref, got)
        deformation.set_backward_stream(side, sm_reserve=8, shared_leaves=True)
        T.run_product(rec)
        assert deformation._overlap["serialised"] == 1
    finally:
        deformation.set_backward_stream(None)
