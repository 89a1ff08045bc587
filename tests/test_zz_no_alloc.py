"""hm_root_policy_apply_noalloc (include/hm_engine.h): the allocation-free variant of the root-policy entry point,
whose kernel scratch is a static device array.  It has no counterpart in the reference,
so the check is against the existing entry point hm_root_policy_apply: the same kernel on the same
inputs, hence bit-identical priors, picks and statuses — on two calls in a row that share the static scratch, the
second with other sizes than the first.  This file is named to be collected last."""
import ctypes as C
import os

import numpy as np
import pytest

from conftest import GOLD


def _dev(t, a):
    """numpy array -> device tensor holding the same bytes (test-side allocation; the entry point allocates nothing)."""
    a = np.ascontiguousarray(a)
    return t.from_numpy(a.view(np.uint8).reshape(-1).copy()).cuda()


def _host(x, dtype, shape=(-1,)):
    return x.cpu().numpy().view(dtype).reshape(shape)


def _noalloc(torch, L, pos, moves, offsets, priors, temperature=0.0, dirichlet_alpha=0.0, dirichlet_epsilon=0.0,
             noise_seed=0, position_hash=None, captures=None, uniforms=None):
    from hivemind_b200._lib import RootPolicy

    n, n_moves = pos.shape[0], int(offsets[-1])
    d_pos, d_moves = _dev(torch, pos), _dev(torch, np.asarray(moves, np.uint32))
    d_off, d_pri = _dev(torch, np.asarray(offsets, np.uint32)), _dev(torch, np.asarray(priors, np.float32))
    d_hash = None if position_hash is None else _dev(torch, np.asarray(position_hash, np.uint64))
    d_cap = None if captures is None else _dev(torch, np.asarray(captures, np.uint8))
    d_uni = None if uniforms is None else _dev(torch, np.asarray(uniforms, np.float64))
    d_picks = torch.zeros(2 * n * 4, dtype=torch.uint8, device="cuda")
    d_status = torch.zeros(n * 4, dtype=torch.uint8, device="cuda")
    ptr = lambda x: None if x is None else x.data_ptr()  # noqa: E731
    rp = RootPolicy(float(temperature), float(dirichlet_alpha), float(dirichlet_epsilon), int(noise_seed), ptr(d_hash),
                    ptr(d_cap), ptr(d_uni))
    torch.cuda.synchronize()
    ok = L.hm_root_policy_apply_noalloc(0, d_pos.data_ptr(), n, d_moves.data_ptr(), d_off.data_ptr(), n_moves,
                                        d_pri.data_ptr(), C.byref(rp), d_picks.data_ptr(), d_status.data_ptr(), None)
    assert ok == 1, L.hm_last_error().decode()
    torch.cuda.synchronize()
    return _host(d_pri, np.float32), _host(d_picks, np.uint32, (n, 2)), _host(d_status, np.int32)


@pytest.mark.gpu
def test_noalloc_root_policy_matches_the_allocating_entry_point(gold_positions, gold_forward):
    import torch

    from hivemind_b200 import _lib
    from hivemind_b200.engine import Engine
    from test_root_policy import syn_case, uniforms_from_draws

    L = _lib.load()
    gold_root = dict(np.load(os.path.join(GOLD, "root_policy.npz")))
    e = Engine(0, 256, 1)
    try:
        # call 1: every stage (temperature, Dirichlet noise through the static scratch, joint sampling) on real lists
        off, packed, moves = gold_positions["offsets"], gold_positions["packed"], gold_positions["moves"]
        u = uniforms_from_draws(gold_root["canonical"], off, str(gold_root["draw_order"]))
        kw = dict(temperature=1.25, dirichlet_alpha=float(gold_root["alpha"]), dirichlet_epsilon=float(gold_root["epsilon"]),
                  noise_seed=99, position_hash=gold_root["seeds"], captures=gold_root["captures"], uniforms=u)
        want = e.rootPolicyApply(packed, moves, off, gold_forward["priors_f32"], **kw)
        assert want is not None
        got = _noalloc(torch, L, packed, moves, off, gold_forward["priors_f32"], **kw)
        assert np.array_equal(got[0], want[0]) and np.array_equal(got[1], want[2]) and np.array_equal(got[2], want[3])
        assert not np.array_equal(got[0], gold_forward["priors_f32"])  # the stages really ran

        # call 2, straight after: other positions, other list sizes, another noise key — on top of call 1's scratch
        spacked, smoves, soff, spri, scap = syn_case(gold_root)
        su = uniforms_from_draws(gold_root["syn_canonical"], soff, str(gold_root["draw_order"]))
        kw = dict(temperature=0.0, dirichlet_alpha=0.3, dirichlet_epsilon=0.25, noise_seed=7,
                  position_hash=gold_root["syn_seeds"].astype(np.uint64), captures=scap, uniforms=su)
        want = e.rootPolicyApply(spacked, smoves, soff, spri, **kw)
        assert want is not None
        got = _noalloc(torch, L, spacked, smoves, soff, spri, **kw)
        assert np.array_equal(got[0], want[0]) and np.array_equal(got[1], want[2]) and np.array_equal(got[2], want[3])

        # sizes beyond the static array are refused with a clear error, before anything is launched
        rp = _lib.RootPolicy(0.0, 0.0, 0.0, 0, None, None, None)
        one = torch.zeros(64, dtype=torch.uint8, device="cuda").data_ptr()
        assert L.hm_root_policy_apply_noalloc(0, one, 1, one, one, _lib.ROOT_NOALLOC_MAX_MOVES + 1, one, C.byref(rp),
                                              None, None, None) == 0
        assert "HM_ROOT_NOALLOC_MAX_MOVES" in L.hm_last_error().decode()
    finally:
        e.close()
