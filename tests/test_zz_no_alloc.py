"""GPU: ragen_group_norm_noalloc (the allocation-free variant of ragen_group_norm, static device scratch) against the
existing entry point, through the C ABI, on two calls in a row.  Collected last on purpose (pytest -x).

The statistics are fp64 atomics in both entry points, so the two agree to the last bit except for the order of the atomic
adds; the bar is the 1e-5 of the other normalisation tests."""
import ctypes

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
RTOL = 1e-5
MAX_GROUPS = 4096  # RAGEN_NOALLOC_MAX_GROUPS in include/ragen_b200.h


def _call(fn, ctx, score, penalty, group, n_groups, method):
    from ragen_b200 import _lib

    out = torch.empty_like(score)  # caller-owned output; the entry points themselves allocate nothing user-visible
    rc = fn(ctx.handle, score.numel(), _lib.ptr(score), _lib.ptr(penalty), _lib.ptr(group), int(n_groups), int(method),
            _lib.ptr(out), ctx.stream())
    return rc, out


def test_noalloc_group_norm_equals_group_norm_on_two_calls_in_a_row():
    from ragen_b200 import _lib

    ctx = _lib.Context.get(0)
    lib = ctx.lib
    rng = np.random.default_rng(11)
    # call 1: many groups, large statistics; call 2: fewer groups with other values - stale sums of call 1 in the shared
    # static array would show up here.  Then method / size variations, and call 1's shape again.
    cases = [(4096, 256, 2, 50.0), (1024, 64, 2, 1.0), (1024, 64, 3, 1.0), (1024, 64, 1, 1.0), (1024, 1, 2, 3.0),
             (512, 512, 2, 1.0), (100, 7, 0, 1.0), (4096 * 3, MAX_GROUPS, 2, 9.0), (4096, 256, 2, 50.0)]
    for B, G, method, scale in cases:
        score = torch.from_numpy((rng.standard_normal(B) * scale).astype(np.float32)).cuda()
        penalty = torch.from_numpy(rng.choice([0.0, -0.1, -0.2], B).astype(np.float32)).cuda()
        group = torch.from_numpy(rng.integers(0, G, B).astype(np.int32)).cuda()
        rc0, want = _call(lib.ragen_group_norm, ctx, score, penalty, group, G, method)
        rc1, got = _call(lib.ragen_group_norm_noalloc, ctx, score, penalty, group, G, method)
        assert rc0 == 0 and rc1 == 0, lib.ragen_last_error()
        torch.cuda.synchronize()
        want, got = want.cpu().double(), got.cpu().double()
        s = max(1.0, float(want.abs().max()))
        assert float((got - want).abs().max()) <= RTOL * s, (B, G, method)
    # penalty may be NULL, as in ragen_group_norm
    rc0, want = _call(lib.ragen_group_norm, ctx, score, None, group, G, 2)
    rc1, got = _call(lib.ragen_group_norm_noalloc, ctx, score, None, group, G, 2)
    assert rc0 == 0 and rc1 == 0
    assert float((got.cpu().double() - want.cpu().double()).abs().max()) <= RTOL * max(1.0, float(want.abs().max()))


def test_noalloc_group_norm_rejects_sizes_beyond_the_static_scratch():
    from ragen_b200 import _lib

    ctx = _lib.Context.get(0)
    B = 2 * (MAX_GROUPS + 1)
    score = torch.zeros(B, dtype=torch.float32, device="cuda")
    group = (torch.arange(B, dtype=torch.int32, device="cuda") // 2).contiguous()
    rc, _ = _call(ctx.lib.ragen_group_norm_noalloc, ctx, score, None, group, MAX_GROUPS + 1, 2)
    assert rc == -1  # RAGEN_EINVAL
    msg = ctx.lib.ragen_last_error().decode()
    assert "RAGEN_NOALLOC_MAX_GROUPS" in msg and str(MAX_GROUPS) in msg
    assert isinstance(ctypes.c_int(rc).value, int)
