"""GPU test of the allocation-free BatchNorm statistics entry point (`dcnn_bn_batch_stats_noalloc`).

That entry point exists for callers that cannot allocate device memory at the point where they need the
statistics -- e.g. a host thread that must stay off the allocator lock, or an embedding process that gives this
library no memory pool -- so its per-tile scratch is a fixed 4 MiB `static __device__` array inside the library.
Here it is called through the C ABI (ctypes), exactly as such a caller would, and its outputs are compared with
the regular two-call path (`dcnn_bn_colstats` + `dcnn_bn_finalize`, scratch supplied by the caller) within the
tolerances the other BatchNorm tests use.  Two calls are made back to back, the second smaller than the first, so
stale scratch contents from the first call would show up if the kernels depended on them.  The file name sorts last
so that a `-x` run has already executed every other test when it gets here."""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu


def _regular(lib, x, M, Cc, eps, mom, rm, rv, gamma, beta):
    from dcnn_b200._lib import check
    tiles = lib.dcnn_bn_stat_tiles(C.c_longlong(M))
    partials = torch.empty(tiles * (2 * Cc + 1), device="cuda", dtype=torch.float32)
    out = torch.empty(4, Cc, device="cuda", dtype=torch.float32)
    check(lib.dcnn_bn_colstats(C.c_void_p(x.data_ptr()), C.c_longlong(M), C.c_int(Cc),
                               C.c_void_p(partials.data_ptr()), C.c_void_p(0)), "bn_colstats")
    check(lib.dcnn_bn_finalize(C.c_void_p(partials.data_ptr()), C.c_int(tiles), C.c_int(Cc), C.c_float(eps),
                               C.c_float(mom), C.c_void_p(rm.data_ptr()), C.c_void_p(rv.data_ptr()),
                               C.c_void_p(gamma.data_ptr()), C.c_void_p(beta.data_ptr()),
                               C.c_void_p(out[0].data_ptr()), C.c_void_p(out[1].data_ptr()),
                               C.c_void_p(out[2].data_ptr()), C.c_void_p(out[3].data_ptr()),
                               C.c_void_p(0)), "bn_finalize")
    return out


def _noalloc(lib, x, M, Cc, eps, mom, rm, rv, gamma, beta):
    out = torch.empty(4, Cc, device="cuda", dtype=torch.float32)   # outputs belong to the caller
    rc = lib.dcnn_bn_batch_stats_noalloc(
        C.c_void_p(x.data_ptr()), C.c_longlong(M), C.c_int(Cc), C.c_float(eps), C.c_float(mom),
        C.c_void_p(rm.data_ptr()), C.c_void_p(rv.data_ptr()), C.c_void_p(gamma.data_ptr()),
        C.c_void_p(beta.data_ptr()), C.c_void_p(out[0].data_ptr()), C.c_void_p(out[1].data_ptr()),
        C.c_void_p(out[2].data_ptr()), C.c_void_p(out[3].data_ptr()), C.c_void_p(0))
    return rc, out


def _close(a, b, tol):
    a, b = a.cpu(), b.cpu()
    return float((a - b).abs().max() / (b.abs().max() + 1e-12)) < tol


def test_noalloc_bn_stats_matches_regular_entry_point_twice_in_a_row():
    from dcnn_b200._lib import lib as load
    lib = load()
    lib.dcnn_bn_noalloc_capacity_floats.restype = C.c_longlong
    cap = lib.dcnn_bn_noalloc_capacity_floats()
    assert cap == 1 << 20
    torch.manual_seed(7)
    # second call is smaller (fewer tiles, fewer channels): it must not see the first call's leftovers
    # ... and the third has a channel count that is not a multiple of 4 (the scalar statistics kernel)
    for M, Cc in [(8 * 32 * 32, 256), (3 * 7 * 7, 36), (5 * 9 * 9, 10)]:
        x = (torch.randn(M, Cc, device="cuda") * 3 + 5).contiguous()
        gamma = torch.rand(Cc, device="cuda") + 0.5
        beta = torch.randn(Cc, device="cuda")
        rm_a, rv_a = torch.zeros(Cc, device="cuda"), torch.ones(Cc, device="cuda")
        rm_b, rv_b = torch.zeros(Cc, device="cuda"), torch.ones(Cc, device="cuda")
        ref = _regular(lib, x, M, Cc, 1e-5, 0.1, rm_a, rv_a, gamma, beta)
        rc, got = _noalloc(lib, x, M, Cc, 1e-5, 0.1, rm_b, rv_b, gamma, beta)
        torch.cuda.synchronize()
        assert rc == 0, lib.dcnn_last_error()
        for i, (name, tol) in enumerate([("mean", 1e-5), ("invstd", 1e-4), ("scale", 1e-4), ("shift", 1e-4)]):
            assert _close(got[i], ref[i], tol), name
        assert _close(rm_b, rm_a, 1e-5) and _close(rv_b, rv_a, 1e-4)
        # and against plain torch on the host
        xc = x.cpu().double()
        assert _close(got[0], xc.mean(0).float(), 1e-5)
        assert _close(got[1], (1.0 / (xc.var(0, unbiased=False) + 1e-5).sqrt()).float(), 1e-4)


def test_noalloc_rejects_oversized_problem():
    from dcnn_b200._lib import lib as load
    lib = load()
    lib.dcnn_last_error.restype = C.c_char_p
    M, Cc = 1 << 18, 2048           # tiles*(2C+1) is far beyond the 1 Mi-float scratch
    x = torch.empty(1, device="cuda")   # never read: the size check comes first
    out = torch.empty(4, device="cuda")
    rc = lib.dcnn_bn_batch_stats_noalloc(
        C.c_void_p(x.data_ptr()), C.c_longlong(M), C.c_int(Cc), C.c_float(1e-5), C.c_float(0.1), C.c_void_p(0),
        C.c_void_p(0), C.c_void_p(0), C.c_void_p(0), C.c_void_p(out.data_ptr()), C.c_void_p(out.data_ptr()),
        C.c_void_p(out.data_ptr()), C.c_void_p(out.data_ptr()), C.c_void_p(0))
    assert rc != 0
    assert b"scratch" in lib.dcnn_last_error()
    torch.cuda.synchronize()
