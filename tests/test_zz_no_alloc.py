"""GPU: vrgpu_potential_direct_noalloc (include/vrgpu.h) -- the allocation-free variant of the direct-sum potential, whose
scratch is a file-scope device array of the library -- against the existing entry point vrgpu_potential (pp_threshold < 0) on
the same particles, two calls in a row with different inputs (the array keeps the previous call's contents, so the second
call shows that nothing stale is read), then the first input again.  Tolerance: 1e-10 relative, as the direct sum is
asserted against the oracle in tests/test_gpu_unbind.py; groups below 4096 members take the same kernel in both entry
points and must agree bit for bit.  The entry point has no counterpart in the
reference.

Collected last on purpose.  The test needs no fixture of the other GPU tests: it loads the
library itself and creates its own context for the comparison call, so it passes alone in a fresh process and after the
others.  torch holds the caller's device buffers -- the entry point itself takes plain device pointers and allocates nothing."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

G, EPS = 43.0211, 0.01


def _group(n, seed):
    rng = np.random.default_rng(seed)
    pos = (rng.normal(size=(n, 3)) * 0.35 + 5.0).astype(np.float32).astype(np.float64)   # float32-quantised, like a snapshot
    mass = rng.uniform(0.5, 2.0, size=n)
    return pos, mass


def _noalloc(lib, torch, pos, mass, n=None, stream=None):
    n = len(pos) if n is None else n
    dpos = torch.from_numpy(np.ascontiguousarray(pos)).cuda()
    dmass = torch.from_numpy(np.ascontiguousarray(mass)).cuda()
    dphi = torch.full((max(n, 1),), np.nan, dtype=torch.float64, device="cuda")
    err = C.create_string_buffer(256)
    torch.cuda.synchronize()
    rc = lib.vrgpu_potential_direct_noalloc(0, stream, dpos.data_ptr(), dmass.data_ptr(), n, G, EPS, dphi.data_ptr(), err, len(err))
    torch.cuda.synchronize()
    return rc, err.value.decode(), dphi.cpu().numpy()[:n]


def _existing(ctx, vrgpu, pos, mass):
    n = len(pos)
    parts = vrgpu.particles_from_arrays(pos, np.zeros_like(pos), mass, np.arange(n))
    return ctx.potential(parts, np.array([0, n]), np.arange(n), G, eps=EPS, pp_threshold=-1)


def test_noalloc_potential_matches_existing_entry_point():
    import torch
    import vrgpu
    assert torch.cuda.is_available(), "no CUDA device: the product path has no CPU fallback"
    torch.cuda.set_device(0)
    lib = vrgpu.load_library()
    ctx = vrgpu.Context(0)
    try:
        a = _group(3000, 1)                      # below SYM_MIN: the same kernel in both entry points
        b = _group(20000, 2)                     # the existing entry point takes its symmetric kernel (fp64 atomics)
        want_a, want_b = _existing(ctx, vrgpu, *a), _existing(ctx, vrgpu, *b)
        # two calls in a row: the big group first, so that the smaller one finds its leftovers in the shared scratch
        rc, msg, got_b = _noalloc(lib, torch, *b)
        assert rc == 0, msg
        rc, msg, got_a = _noalloc(lib, torch, *a)
        assert rc == 0, msg
        assert np.all(np.isfinite(got_a)) and np.all(np.isfinite(got_b)) and np.all(got_a < 0)
        assert np.array_equal(got_a, want_a)                                    # same kernel, same order of operations
        assert np.max(np.abs(got_b - want_b) / np.abs(want_b)) < 1e-10
        # and again, on a stream of the caller's, with the inputs of the first call
        s = torch.cuda.Stream()
        with torch.cuda.stream(s):
            rc, msg, again_b = _noalloc(lib, torch, *b, stream=s.cuda_stream)
        assert rc == 0, msg
        assert np.array_equal(again_b, got_b)                                   # no atomics: run-to-run identical
        # unit masses through a NULL mass pointer
        dpos = torch.from_numpy(a[0]).cuda()
        dphi = torch.empty(len(a[0]), dtype=torch.float64, device="cuda")
        assert lib.vrgpu_potential_direct_noalloc(-1, None, dpos.data_ptr(), None, len(a[0]), G, EPS, dphi.data_ptr(), None, 0) == 0
        torch.cuda.synchronize()
        want_unit = _existing(ctx, vrgpu, a[0], np.ones(len(a[0])))
        assert np.array_equal(dphi.cpu().numpy(), want_unit)
    finally:
        ctx.close()


def test_noalloc_rejects_what_does_not_fit():
    import torch
    import vrgpu
    torch.cuda.set_device(0)
    lib = vrgpu.load_library()
    nmax = vrgpu.NOALLOC_MAX_PARTICLES
    pos, mass = _group(16, 3)
    rc, msg, _ = _noalloc(lib, torch, pos, mass, n=nmax + 1)     # rejected before any pointer is touched
    assert rc == vrgpu.EINVAL and str(nmax) in msg and "exceed" in msg
    rc, msg, _ = _noalloc(lib, torch, pos, mass, n=-1)
    assert rc == vrgpu.EINVAL and "negative" in msg
    err = C.create_string_buffer(128)
    assert lib.vrgpu_potential_direct_noalloc(0, None, None, None, 8, G, EPS, None, err, len(err)) == vrgpu.EINVAL
    assert b"null device pointer" in err.value
    assert lib.vrgpu_potential_direct_noalloc(0, None, None, None, 0, G, EPS, None, None, 0) == 0   # empty group: nothing to do
    # the largest size that fits
    big = _group(nmax, 4)
    rc, msg, got = _noalloc(lib, torch, *big)
    assert rc == 0, msg
    ctx = vrgpu.Context(0)
    try:
        want = _existing(ctx, vrgpu, *big)
    finally:
        ctx.close()
    assert np.max(np.abs(got - want) / np.abs(want)) < 1e-10
