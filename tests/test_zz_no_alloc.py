"""ed3dgs_dist_cuda2_static: distCUDA2 for callers that have no way to allocate device memory.

A host that handed its whole GPU budget to one pool at start-up (or a foreign-language binding that can only
pass the arrays it already owns) cannot give ed3dgs_dist_cuda2 the workspace it asks for. The static variant
keeps that workspace inside the library image instead. This file checks, through the C ABI, that it returns
exactly what the workspace variant returns -- distCUDA2 is an exact 3-NN mean, so the library's tests hold it
to bit equality (tests/test_raster_gpu.py::test_dist_cuda2_bit_exact) and so does this one -- on two calls in a
row: the second call finds the first call's keys, sort buffers and look-back words still in the array and has
to give the right answer for a different cloud regardless.

The file name sorts last on purpose: `pytest -x` stops at the first failure, and a failure here must not hide
the parity tests. It passes alone in a fresh process (first use of the array on the device) and after the
rest of the suite (library warm, other kernels' leftovers in memory).
"""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _with_workspace(lib, _lib, pts):
    P = pts.shape[0]
    out = torch.full((P,), -1.0, dtype=torch.float32, device=pts.device)
    nbytes = lib.ed3dgs_dist_cuda2_workspace_bytes(P)
    work = torch.empty(nbytes, dtype=torch.uint8, device=pts.device)
    _lib.check(lib.ed3dgs_dist_cuda2(P, pts.data_ptr(), out.data_ptr(), work.data_ptr(), nbytes,
                                     _lib.stream_handle()))
    return out


def _static(lib, _lib, pts, out):
    # inputs and outputs exist before the call; the call itself is handed nothing else
    _lib.check(lib.ed3dgs_dist_cuda2_static(pts.shape[0], pts.data_ptr(), out.data_ptr(), _lib.stream_handle()))
    return out


def test_static_scratch_matches_workspace_variant_twice_in_a_row():
    from ed3dgs_b200 import _lib, synthetic
    lib = _lib.lib()
    dev = torch.device("cuda:0")
    with torch.cuda.device(dev):
        # two different clouds, the larger first, so the second call runs over the first one's leftovers
        clouds = [synthetic.make_gaussians(50_000, seed=11, device=dev)["xyz"].contiguous(),
                  synthetic.make_gaussians(7_001, seed=12, device=dev)["xyz"].contiguous()]
        expected = [_with_workspace(lib, _lib, c) for c in clouds]
        outs = [torch.full((c.shape[0],), -1.0, dtype=torch.float32, device=dev) for c in clouds]
        torch.cuda.synchronize()
        # the workspace variant above ran the same kernels, so their code is resident; from here on neither
        # torch's allocator nor the driver's free-memory count may move (the smaller cloud alone needs 0.4 MB)
        held, free_before = torch.cuda.memory_reserved(dev), torch.cuda.mem_get_info(dev)[0]
        for c, o in zip(clouds, outs):      # back to back on one stream: ordered, so no race on the array
            _static(lib, _lib, c, o)
        torch.cuda.synchronize()
        assert torch.cuda.memory_reserved(dev) == held
        assert torch.cuda.mem_get_info(dev)[0] >= free_before - (256 << 10)
        for o, e in zip(outs, expected):
            assert torch.isfinite(o).all() and (o >= 0).all()
            assert torch.equal(o, e)
        # and once more with the first cloud: same answer from a dirty array
        again = _static(lib, _lib, clouds[0], torch.empty_like(outs[0]))
        torch.cuda.synchronize()
        assert torch.equal(again, expected[0])


def test_static_scratch_at_its_limit_and_beyond():
    from ed3dgs_b200 import _lib, synthetic
    lib = _lib.lib()
    dev = torch.device("cuda:0")
    limit = lib.ed3dgs_dist_cuda2_static_max_points()
    assert limit == 262144
    with torch.cuda.device(dev):
        pts = synthetic.make_gaussians(limit, seed=13, device=dev)["xyz"].contiguous()
        got = _static(lib, _lib, pts, torch.empty(limit, dtype=torch.float32, device=dev))
        assert torch.equal(got, _with_workspace(lib, _lib, pts))
        # one point more: refused before anything is launched, output untouched, message names the limit
        big = torch.zeros((limit + 1, 3), dtype=torch.float32, device=dev)
        out = torch.full((limit + 1,), -7.0, dtype=torch.float32, device=dev)
        rc = lib.ed3dgs_dist_cuda2_static(limit + 1, big.data_ptr(), out.data_ptr(), _lib.stream_handle())
        assert rc == 1   # ED3DGS_ERR_ARG
        msg = lib.ed3dgs_last_error().decode()
        assert str(limit) in msg and "ed3dgs_dist_cuda2" in msg
        torch.cuda.synchronize()
        assert (out == -7.0).all()
        # empty input is a no-op, as for the workspace variant
        assert lib.ed3dgs_dist_cuda2_static(0, None, None, _lib.stream_handle()) == 0
