"""The offsets entry point that brings its own scratch (kas_patch_event_offsets_static).

It exists for callers that cannot obtain device memory when they need the patch-log
offsets: a service that sized all of its HBM at start-up and runs with a fixed budget,
a plug-in living in a host process whose memory pool belongs to another framework, or
a control loop that cannot take the device-wide synchronisation a cudaMalloc/cudaFree
pair brings.  The library keeps the scan's tile sums in a fixed 64 KB array of its own
instead, which every call on a device shares.

Checked here, through the C ABI: the result equals kas_patch_event_offsets (caller's
scratch) bit for bit on two calls in a row -- a large fleet first, then a smaller one,
so the second call runs over whatever the first left in the shared array -- and a fleet
beyond the built-in limit is refused.  This file sorts last on purpose: a run that stops
at the first failure has already run every other test.
"""
import ctypes as C

import numpy as np
import pytest

from tests import helpers

pytestmark = pytest.mark.gpu


def _offsets(torch, engine, d_llen, T, Q, policy, d_go, R, d_cur0, static):
    parr = (engine.Policy * 1)(engine.Policy(*policy))
    args = engine.ReplayArgs(engine.ptr(d_llen), None, 0, T, Q, R, engine.ptr(d_go),
                             engine.ptr(d_cur0), parr, 1, None, None, None)
    off = torch.full((R + 1,), -1, dtype=torch.int64, device='cuda')
    L = engine.lib()
    if static:
        st = L.kas_patch_event_offsets_static(C.byref(args), 0, engine.ptr(off), engine.stream_ptr())
    else:
        n = L.kas_patch_events_scratch_bytes(R)
        scratch = torch.empty(max(n, 8), dtype=torch.uint8, device='cuda')
        st = L.kas_patch_event_offsets(C.byref(args), 0, engine.ptr(off), engine.ptr(scratch), n,
                                       engine.stream_ptr())
    assert st == 0, engine.lib().kas_strerror(st)
    torch.cuda.synchronize()
    return off.cpu().numpy()


def test_static_scratch_offsets_equal_caller_scratch_offsets():
    import torch
    from autoscaler import engine
    rng = np.random.default_rng(2026)
    first = True
    for T, Q, grouped, policy in [(12, 200_003, False, (1, 0, 3)),     # 49 scan tiles
                                  (30, 9_000, True, (2, 1, 9))]:       # fewer tiles, stale sums behind
        llen, go = helpers.random_fleet(rng, T, Q, grouped=grouped)
        R = Q if go is None else len(go) - 1
        d_llen = torch.from_numpy(llen).cuda()
        d_go = None if go is None else torch.from_numpy(go.view(np.int64)).cuda()
        d_cur0 = torch.from_numpy(rng.integers(0, 5, size=R).astype(np.int32)).cuda()
        if first:
            # this is the first call on the device: it also looks the built-in array up
            got = _offsets(torch, engine, d_llen, T, Q, policy, d_go, R, d_cur0, static=True)
            want = _offsets(torch, engine, d_llen, T, Q, policy, d_go, R, d_cur0, static=False)
            first = False
        else:
            want = _offsets(torch, engine, d_llen, T, Q, policy, d_go, R, d_cur0, static=False)
            got = _offsets(torch, engine, d_llen, T, Q, policy, d_go, R, d_cur0, static=True)
        assert got[0] == 0 and (np.diff(got) >= 0).all()
        assert (got == want).all()
        assert got[-1] > 0


def test_static_scratch_refuses_fleets_beyond_its_limit():
    import torch
    from autoscaler import engine
    parr = (engine.Policy * 1)(engine.Policy(1, 0, 1))
    d_llen = torch.zeros(8, dtype=torch.int32, device='cuda')
    off = torch.zeros(8, dtype=torch.int64, device='cuda')
    too_many = engine.STATIC_SCRATCH_MAX_RESOURCES + 1
    args = engine.ReplayArgs(engine.ptr(d_llen), None, 0, 1, too_many, too_many, None, None, parr, 1,
                             None, None, None)
    st = engine.lib().kas_patch_event_offsets_static(C.byref(args), 0, engine.ptr(off),
                                                     engine.stream_ptr())
    assert st == engine.KAS_ERR_STATIC_LIMIT
    assert b'built-in scratch' in engine.lib().kas_strerror(st)
    torch.cuda.synchronize()
    assert int(off.sum().item()) == 0          # nothing was launched
