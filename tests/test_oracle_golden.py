"""CPU: the plain-C oracle (oracle/hm_oracle.c) against golden vectors produced by the
reference itself (oracle/make_golden.py) and against the reference's known-answer tests
(engine/tests/test_move_gen.cc, cited per test)."""
import numpy as np
import pytest

from conftest import formula_logits
from oracle.pyoracle import POS_DTYPE, ValueConfig


def _pos(raw):
    return np.ascontiguousarray(raw).view(POS_DTYPE).reshape(-1)


def test_planes_bit_exact_fp32_fp16_u8(oracle, gold_positions):
    g = gold_positions
    assert np.array_equal(oracle.encode_planes(g["packed"], "f32").view(np.uint32), g["planes_f32"].view(np.uint32))
    assert np.array_equal(oracle.encode_planes(g["packed"], "f16"), g["planes_f16"])
    assert np.array_equal(oracle.encode_planes(g["packed"], "u8"), g["planes_u8"])


def test_bf16_is_rn_cast_of_fp32(oracle, gold_positions):
    import torch

    f32 = gold_positions["planes_f32"]
    want = torch.from_numpy(f32).to(torch.bfloat16).view(torch.int16).numpy().view(np.uint16)
    assert np.array_equal(oracle.encode_planes(gold_positions["packed"], "bf16"), want)


def test_policy_tables_match_reference(oracle):
    import os
    from conftest import GOLD

    t = np.load(os.path.join(GOLD, "policy_tables.npz"))
    normal, drop = oracle.policy_tables()
    assert np.array_equal(normal, t["normal"].astype(np.int32))
    assert np.array_equal(drop, t["drop"].astype(np.int32))
    # 1792 board-move labels + 22 knight under-promotions + 304 drops = the 2118 real labels
    assert (normal[..., 0] >= 0).sum() == 2 * 1792
    assert len(np.unique(normal[0][normal[0] >= 0])) == 1814 and (drop[0] >= 0).sum() == 304


def test_policy_index_all_legal_moves(oracle, gold_positions):
    g = gold_positions
    got = np.array([oracle.policy_index(m, s) for m, s in zip(g["all_moves"], g["all_stm"])], np.int32)
    assert np.array_equal(got, g["all_idx"])
    assert (g["all_idx"] == -1).any()  # R/B under-promotions occur in the sample


def test_priors_bit_exact_on_formula_logits(oracle, gold_positions):
    g = gold_positions
    n = g["packed"].shape[0]
    logits = formula_logits(n)
    got = oracle.batch_priors_f16(logits[:, 0].view(np.uint16), logits[:, 1].view(np.uint16), g["packed"], g["moves"], g["offsets"])
    assert np.array_equal(got.view(np.uint32), g["priors_formula_f16"].view(np.uint32))


# ---- the reference's own known-answer tests, restated ---------------------------------

def test_kat_start_history_planes_empty(oracle, gold_kat):  # test_move_gen.cc:68-83
    p = oracle.encode_planes(_pos(gold_kat["start_white.packed"]), "f32")[0].reshape(74, 64)
    assert np.array_equal(p, gold_kat["start_white.planes_f32"].reshape(74, 64))
    for off in (0, 37):
        assert not p[off + 32: off + 37].any()


def test_kat_half_matches_float_conversion(oracle, gold_kat):  # test_move_gen.cc:85-103
    pk = _pos(gold_kat["half_black_tadv.packed"])
    f = oracle.encode_planes(pk, "f32")[0]
    h = oracle.encode_planes(pk, "f16")[0]
    assert np.array_equal(h, gold_kat["half_black_tadv.planes_f16"])
    assert np.array_equal(h.view(np.float16), f.astype(np.float16))


def test_kat_oriented_last_move_and_clock(oracle, gold_kat):  # test_move_gen.cc:105-125
    G1, F3, G8, F6 = 6, 21, 62, 45
    w = oracle.encode_planes(_pos(gold_kat["g1f3_white.packed"]), "f32")[0].reshape(74, 64)
    assert w[32, G1] == 1.0 and w[33, F3] == 1.0 and w[34, 0] == np.float32(1.0) / np.float32(50.0)
    b = oracle.encode_planes(_pos(gold_kat["g1f3_black.packed"]), "f32")[0].reshape(74, 64)
    assert b[32, G8] == 1.0 and b[33, F6] == 1.0
    assert w[32].sum() == 1 and w[33].sum() == 1


def test_kat_repetition_context(oracle, gold_kat):  # test_move_gen.cc:145-157
    p = oracle.encode_planes(_pos(gold_kat["repetition_white.packed"]), "f32")[0].reshape(74, 64)
    assert (p[35] == 1.0).all() and (p[36] == 0.0).all()


def test_kat_normalize_extreme_and_nonfinite(oracle):  # test_move_gen.cc:643-657
    p = oracle.normalize_logits([1000.0, 999.0, -1000.0])
    assert np.isfinite(p).all() and abs(p.sum() - 1.0) < 1e-6 and p[0] > p[1]
    f = oracle.normalize_logits([np.nan, -np.inf])
    assert f[0] == 0.5 and f[1] == 0.5


def test_kat_pass_uses_logit_without_floor(oracle, gold_kat):  # test_move_gen.cc:659-671
    acts = gold_kat["start_white.actions0"]
    pol = np.zeros(4672, np.float32)
    pol[0] = -20.0  # POLICY_INDEX["pass"] == 0
    p = oracle.priors_f32(pol, acts, int(gold_kat["start_white.stm0"]))
    assert p.size == acts.size and p[-1] < 1e-6


def test_kat_half_policy_matches_float_policy(oracle, gold_kat):  # test_move_gen.cc:673-693
    acts = gold_kat["start_white.actions0"]
    f = ((np.arange(4672) % 17) - 8).astype(np.float32) * 0.25
    pf = oracle.priors_f32(f, acts, 0)
    ph = oracle.priors_f16(f.astype(np.float16).view(np.uint16), acts, 0)
    assert np.array_equal(pf, ph)


def test_kat_low_prior_checking_move(oracle, gold_kat):  # test_move_gen.cc:695-729
    acts = gold_kat["low_prior_check.actions0"]
    uci = list(gold_kat["low_prior_check.uci0"])
    pol = np.full(4672, -10.0, np.float32)
    qi, ci = uci.index("h1h3"), uci.index("Q@g8")
    pol[oracle.policy_index(acts[qi], 0)] = 5.0
    pol[oracle.policy_index(acts[ci], 0)] = -5.0
    p = oracle.priors_f32(pol, acts, 0)
    assert p[ci] < p[qi]


def test_kat_fast_index_matches_label_lookup(oracle, gold_kat):  # test_move_gen.cc:747-767
    # label -> index map rebuilt from the documented channel layout via the oracle tables is
    # covered by test_policy_tables_match_reference; here: black-to-move mirroring on the FEN pair
    for bn in range(2):
        acts = gold_kat[f"fast_index_fen.actions{bn}"][:-1]
        stm = int(gold_kat[f"fast_index_fen.stm{bn}"])
        uci = list(gold_kat[f"fast_index_fen.uci{bn}"])[:-1]
        for m, u in zip(acts, uci):
            idx = oracle.policy_index(m, stm)
            if u[-1] in "rb" and len(u) == 5:
                assert idx == -1
            else:
                assert 64 <= idx < 4672, (u, idx)
    assert int(gold_kat["fast_index_fen.stm0"]) == 1 and int(gold_kat["fast_index_fen.stm1"]) == 0


# ---- value shaping (searchthread.cc:569-619) --------------------------------------------

def _same_shaped(got, want):
    """Within 1 ulp of the reference build (g++ -O2 -march=x86-64-v3 contracts `(1-w)*s + w*v` into an FMA,
    the restatement rounds both products); NaN exactly where the reference has NaN; the sign of a zero is not
    part of the fixture (it is read back as 0.0f + value from the root's value sum)."""
    got, want = np.asarray(got, np.float32), np.asarray(want, np.float32)
    assert np.array_equal(np.isnan(got), np.isnan(want))
    m = ~np.isnan(want)
    assert np.abs(got[m] - want[m]).max() <= 1.2e-7


def test_value_shaping_pinned_by_the_reference_search(oracle):
    """tests/golden/value_shaping.npz = what the reference's OWN SearchThread::process_batch
    (searchthread.cc:569-621, compiled unmodified against a scripted Engine, oracle/d5_pin.cc) backed up for
    344 fp16 (value, wdl, moves_left) inputs incl. non-finite ones under 5 runtime configurations."""
    import os

    from conftest import GOLD

    g = np.load(os.path.join(GOLD, "value_shaping.npz"))
    cases = g["cases_u16"]
    assert cases.shape[0] >= 256
    for i, (en, w, disc, contempt) in enumerate(g["configs"]):
        cfg = ValueConfig(int(en), float(w), float(disc), float(contempt))
        got = [oracle.shape_value_half(int(r[0]), r[1:4], int(r[4]), cfg) for r in cases]
        _same_shaped(got, g[f"shaped_{i}"])


def test_value_shaping_fixture_is_what_the_reference_computes(ref):
    """Where the compiled reference exists: regenerate a slice of the fixture live."""
    import os

    from conftest import GOLD
    from oracle.pyoracle import REF_D5_SO

    if not os.path.exists(REF_D5_SO):
        pytest.skip("oracle/_ref/libhmref_d5.so not built")
    g = np.load(os.path.join(GOLD, "value_shaping.npz"))
    for i in (0, 3):
        en, w, disc, contempt = g["configs"][i]
        live = ref.d5_shape(g["cases_u16"], ValueConfig(int(en), float(w), float(disc), float(contempt)))
        assert np.array_equal(live.view(np.uint32), g[f"shaped_{i}"].view(np.uint32))


def test_value_shaping_cases(oracle):
    cfg = ValueConfig.default()
    v = oracle.shape_value(0.5, [0.0, 0.0, 0.0], 0.0, cfg)
    assert abs(v - 0.375) < 1e-7                                  # 0.75*0.5 + 0.25*0
    assert oracle.shape_value(float("nan"), None, None, cfg) == 0.0
    assert oracle.shape_value(3.0, None, None, cfg) == 1.0
    w = oracle.shape_value(1.0, [-10.0, 0.0, 10.0], 1.0, cfg)
    assert abs(w - (0.75 + 0.25 * (np.exp(0) - np.exp(-20)) / (np.exp(0) + np.exp(-10) + np.exp(-20))) * 0.995) < 1e-6
    # non-finite WDL logit -> scalar value only (searchthread.cc:581-583)
    assert abs(oracle.shape_value(0.5, [np.inf, 0.0, 0.0], 0.0, cfg) - 0.5) < 1e-7
    # MovesLeftDiscountingPrefersFastWinAndDistantLoss (test_move_gen.cc:784-802)
    fast = oracle.shape_value(0.8, None, 0.1, cfg)
    slow = oracle.shape_value(0.8, None, 0.9, cfg)
    assert fast > slow
    assert oracle.shape_value(-0.8, None, 0.9, cfg) > oracle.shape_value(-0.8, None, 0.1, cfg)
