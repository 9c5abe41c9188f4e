"""mpcb_step_static -- the allocation-free entry point (include/mpcb.h; no counterpart in the reference)
-- against mpcb_step on identical inputs, two calls in a row, through the C ABI.

The file is named so that pytest collects it last.  It must pass alone in a fresh
process and after the other GPU tests (the static scratch is per process and keeps the last call's contents).

Tolerance: the one the step parity tests use for applied controls (1e-6); integer outputs array_equal.  The launch
order is the only thing the scratch changes, so the outputs are in fact expected to be bit-identical."""
import ctypes as C

import numpy as np
import pytest

from conftest import golden, load_routes

pytestmark = pytest.mark.gpu

T = 20


def _inputs(reps):
    g = golden(f"random_T{T}.npz")
    tile = lambda a: np.ascontiguousarray(np.concatenate([a] * reps, axis=0))
    d = dict(state=tile(g["state"]), route=tile(g["route"].astype(np.int32)), ncourse=tile(g["ncourse"].astype(np.int32)),
             L=tile(g["L"]), target=tile(g["target_in"].astype(np.int32)), warm=tile(g["has_warm"].astype(np.int32)),
             oa=tile(np.nan_to_num(g["oa_in"])), od=tile(np.nan_to_num(g["od_in"])),
             di_ai=tile(np.stack([g["di_in"], g["ai_in"]], axis=1)))
    # make the copies distinct problems: move every state a little along its own heading
    n = len(d["state"])
    shift = 0.02 * (np.arange(n) // len(g["state"]))
    d["state"][:, 0] += shift * np.cos(d["state"][:, 2])
    d["state"][:, 1] += shift * np.sin(d["state"][:, 2])
    return d


class _Arm:
    """One set of caller-owned device buffers for mpcb_step / mpcb_step_static."""

    def __init__(self, torch, d, dev):
        t = lambda a: torch.as_tensor(a).to(dev).contiguous()
        self.B = len(d["state"])
        self.state, self.route, self.ncourse, self.L = t(d["state"]), t(d["route"]), t(d["ncourse"]), t(d["L"])
        self.target, self.warm, self.oa, self.od, self.di_ai = t(d["target"]), t(d["warm"]), t(d["oa"]), t(d["od"]), t(d["di_ai"])
        z = lambda *s, dtype=torch.float64: torch.zeros(*s, dtype=dtype, device=dev)
        self.traj, self.xref, self.xdev = z(self.B, 4, T + 1), z(self.B, 4, T + 1), z(self.B)
        self.status, self.iters = z(self.B, dtype=torch.int32), z(self.B, dtype=torch.int32)

    def args(self, B=None):
        p = lambda x: C.c_void_p(x.data_ptr())
        return [B or self.B] + [p(x) for x in (self.state, self.route, self.ncourse, self.L, self.target, self.warm, self.oa,
                                               self.od, self.traj, self.xref, self.di_ai, self.xdev, self.status, self.iters)]

    def outputs(self):
        return {k: getattr(self, k).cpu().numpy() for k in
                ("target", "warm", "oa", "od", "traj", "xref", "di_ai", "xdev", "status", "iters")}


def _same(a, b, what):
    for k in ("target", "warm", "status", "iters"):
        assert np.array_equal(a[k], b[k]), f"{what}: {k} differs"
    assert np.array_equal(a["xref"], b["xref"]), f"{what}: xref differs"
    ok = a["status"] <= 1
    assert ok.sum() > 0.9 * len(ok)
    assert np.abs(a["di_ai"] - b["di_ai"]).max() < 1e-6, what
    for k in ("oa", "od", "traj"):
        assert np.abs(a[k][ok] - b[k][ok]).max() < 1e-6, f"{what}: {k}"
    assert np.array_equal(np.isnan(a["xdev"]), np.isnan(b["xdev"]))
    assert np.nanmax(np.abs(a["xdev"] - b["xdev"])) < 1e-6


def test_static_scratch_step_matches_mpcb_step_twice():
    import torch
    from mpc_for_av_at_intersection_b200 import MpcConfig, _lib
    from mpc_for_av_at_intersection_b200.batched import BatchedMPC, CourseSet
    d = _inputs(reps=4)                                   # 1600 problems: more than one wave of resident warps
    B = len(d["state"])
    eng = BatchedMPC(MpcConfig.from_json(T=T), CourseSet(load_routes(False)), _lib.STATIC_MAX_BATCH + 8)
    L, h, dev = eng._lib, eng._h, eng.dev
    ref, new = _Arm(torch, d, dev), _Arm(torch, d, dev)
    stream = C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
    free0 = None
    for call in range(2):                                # the second call starts from the first one's warm start and order
        _lib.check(L.mpcb_step(h, *ref.args(), stream), h, "mpcb_step")
        _lib.check(L.mpcb_step_static(h, *new.args(), stream), h, "mpcb_step_static")
        torch.cuda.synchronize()
        _same(ref.outputs(), new.outputs(), f"call {call}")
        free = torch.cuda.mem_get_info(dev)[0]
        if free0 is not None:
            assert free == free0, "mpcb_step_static must not allocate device memory"
        free0 = free
    assert (new.outputs()["iters"] > 0).all()
    # a different batch size next: the kept slot -> problem table is no permutation of it and must be reset
    small = {k: v[:300].copy() for k, v in d.items()}
    ref2, new2 = _Arm(torch, small, dev), _Arm(torch, small, dev)
    for call in range(2):
        _lib.check(L.mpcb_step(h, *ref2.args(), stream), h, "mpcb_step")
        _lib.check(L.mpcb_step_static(h, *new2.args(), stream), h, "mpcb_step_static")
        torch.cuda.synchronize()
        _same(ref2.outputs(), new2.outputs(), f"small call {call}")
    # beyond the stated maximum: rejected with a clear error, nothing launched
    n0 = eng.launches
    rc = L.mpcb_step_static(h, *new.args(B=_lib.STATIC_MAX_BATCH + 1), stream)
    assert rc == -1 and b"MPCB_STATIC_MAX_BATCH" in L.mpcb_last_error(h)
    assert eng.launches == n0
    eng.close()


def test_static_scratch_rejects_long_horizons():
    from mpc_for_av_at_intersection_b200 import MpcConfig
    from mpc_for_av_at_intersection_b200.batched import BatchedMPC, CourseSet
    eng = BatchedMPC(MpcConfig.from_json(T=40), CourseSet(load_routes(False)), 8)
    a = [C.c_void_p(256)] * 14                            # never dereferenced: the call is refused before any launch
    a[3] = None
    rc = eng._lib.mpcb_step_static(eng._h, 8, *a, None)
    assert rc == -1 and b"use mpcb_step" in eng._lib.mpcb_last_error(eng._h)
    eng.close()
