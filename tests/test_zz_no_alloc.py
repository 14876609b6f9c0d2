"""vcr_dist2_knn3_noalloc: distCUDA2 for hosts that cannot give the library any device memory at call time.

Who needs it: a program whose free GPU memory already belongs to someone else's allocator (a fresh allocation would
fail), a plug-in that must not allocate once it is running, or code called from within an allocator callback.  The
entry point keeps its scratch in a fixed array inside the library instead of taking it from the caller.

The test drives the new entry point through the C ABI (ctypes) and compares it with the ordinary entry point
(vcr_dist2_knn3, caller-provided scratch) on the same points — twice in a row, the second time with a different point
count, because both calls share the built-in array and the second one starts from whatever the first one left there.
The tolerance is the one the other kNN tests use (rtol 1e-5).  It also checks that an input larger than the built-in
scratch is refused with a message instead of being run.

This file is named so that pytest collects it last: if it fails, a `-x` run has already been through every other test.
It passes on its own in a fresh process as well as after the rest of the suite.
"""
import ctypes

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _stream():
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


def _with_caller_scratch(n, pts):
    out = torch.empty(pts.shape[0], dtype=torch.float32, device=pts.device)
    scratch = torch.empty(int(n.vcr_knn_scratch_bytes(pts.shape[0])), dtype=torch.uint8, device=pts.device)
    n.check(n.vcr_dist2_knn3(pts.shape[0], pts.data_ptr(), out.data_ptr(), scratch.data_ptr(), _stream()), "vcr_dist2_knn3")
    return out


def _with_builtin_scratch(n, pts):
    out = torch.full((pts.shape[0],), -1.0, dtype=torch.float32, device=pts.device)
    n.check(n.vcr_dist2_knn3_noalloc(pts.shape[0], pts.data_ptr(), out.data_ptr(), _stream()), "vcr_dist2_knn3_noalloc")
    return out


def test_builtin_scratch_matches_caller_scratch_twice_in_a_row():
    import vcr_native as n
    dev = torch.device("cuda")
    limit = n.vcr_knn_noalloc_max_points()
    assert limit == 65536
    g = torch.Generator().manual_seed(2024)
    # first call: close to the limit; second call: fewer points, different distribution, on the dirty array
    first = (torch.randn((60000, 3), generator=g) * 2.0).to(dev).contiguous()
    second = torch.cat([torch.rand((9000, 3), generator=g), torch.randn((4321, 3), generator=g) * 0.1 + 5.0]).to(dev).contiguous()
    for pts in (first, second):
        got = _with_builtin_scratch(n, pts)
        want = _with_caller_scratch(n, pts)
        torch.cuda.synchronize()
        assert torch.isfinite(got).all() and (got >= 0).all()
        np.testing.assert_allclose(got.cpu().numpy(), want.cpu().numpy(), rtol=1e-5, atol=1e-12)
    # and once more with the first input: the result must not depend on what ran in between
    again = _with_builtin_scratch(n, first)
    np.testing.assert_allclose(again.cpu().numpy(), _with_caller_scratch(n, first).cpu().numpy(), rtol=1e-5, atol=1e-12)


def test_too_many_points_are_refused_not_run():
    import vcr_native as n
    dev = torch.device("cuda")
    P = n.vcr_knn_noalloc_max_points() + 1
    pts = torch.zeros((P, 3), device=dev)
    out = torch.full((P,), -1.0, device=dev)
    rc = n.vcr_dist2_knn3_noalloc(P, pts.data_ptr(), out.data_ptr(), _stream())
    assert rc != 0
    msg = n.vcr_last_error().decode()
    assert "exceed" in msg and "65536" in msg
    torch.cuda.synchronize()
    assert float(out.min()) == -1.0 and float(out.max()) == -1.0, "a refused call must not touch the output"
    # empty input is a no-op, as for the ordinary entry point
    assert n.vcr_dist2_knn3_noalloc(0, None, None, _stream()) == 0
