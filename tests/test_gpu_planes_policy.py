"""GPU parity, stages P and D, through the C ABI (include/hm_engine.h):
bit-exact planes (fp32 / fp16 / bf16 / u8), policy indices, masked softmax, value shaping."""
import numpy as np
import pytest

from conftest import formula_logits, random_packed

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def engine():
    from hivemind_b200.engine import Engine

    e = Engine(0, 8, 1)
    yield e
    e.close()


def test_planes_bit_exact_vs_reference_golden(engine, gold_positions):
    g = gold_positions
    assert np.array_equal(engine.encodePlanes(g["packed"], "f32").view(np.uint32), g["planes_f32"].view(np.uint32))
    assert np.array_equal(engine.encodePlanes(g["packed"], "f16"), g["planes_f16"])
    assert np.array_equal(engine.encodePlanes(g["packed"], "u8"), g["planes_u8"])


def test_planes_reference_kat_positions(engine, gold_kat):
    from oracle.pyoracle import POS_DTYPE

    for name in ("start_white", "half_black_tadv", "g1f3_white", "g1f3_black", "repetition_white", "low_prior_check", "fast_index_fen"):
        pk = np.ascontiguousarray(gold_kat[name + ".packed"]).view(POS_DTYPE).reshape(-1)
        assert np.array_equal(engine.encodePlanes(pk, "f32")[0], gold_kat[name + ".planes_f32"]), name
        assert np.array_equal(engine.encodePlanes(pk, "f16")[0], gold_kat[name + ".planes_f16"]), name


@pytest.mark.parametrize("n,seed", [(1, 1), (3, 2), (4097, 3), (65536, 4)])
def test_planes_bit_exact_vs_oracle_random_records(engine, oracle, n, seed):
    """Ragged and large batches of synthetic records (every field exercised, incl. flips,
    drops as last move, rule50 > 50, pockets up to 16)."""
    pos = random_packed(n, seed)
    for dt in ("f32", "f16", "bf16", "u8"):
        got = engine.encodePlanes(pos, dt)
        want = oracle.encode_planes(pos, dt)
        assert np.array_equal(got.view(np.uint8), want.view(np.uint8)), dt


def test_planes_empty_batch(engine):
    from oracle.pyoracle import POS_DTYPE

    assert engine.encodePlanes(np.zeros(0, POS_DTYPE), "f32").shape == (0, 74, 8, 8)


def test_policy_indices_match_reference(engine, gold_positions):
    g = gold_positions
    assert np.array_equal(engine.policyIndices(g["all_moves"], g["all_stm"]), g["all_idx"])
    # exhaustive: every (from,to) normal move, queen/knight promotions, drops, both colours
    moves, stm = [], []
    for c in (0, 1):
        for f in range(64):
            for t in range(64):
                moves += [(f << 6) | t, (3 << 12) | (5 << 16) | (f << 6) | t, (3 << 12) | (2 << 16) | (f << 6) | t,
                          (3 << 12) | (4 << 16) | (f << 6) | t]
                stm += [c] * 4
        for t in range(64):
            for pt in range(0, 8):
                moves.append((4 << 12) | (pt << 16) | t)
                stm.append(c)
    moves = np.array(moves, np.uint32)
    stm = np.array(stm, np.uint8)
    from oracle.pyoracle import Oracle

    o = Oracle()
    want = np.array([o.policy_index(m, s) for m, s in zip(moves, stm)], np.int32)
    assert np.array_equal(engine.policyIndices(moves, stm), want)


def test_priors_on_formula_logits_match_reference(engine, gold_positions):
    """Identical fp16 logits in -> the reference's float priors out (<= 1 ulp; the device uses a
    double-precision exp where the reference uses glibc expf)."""
    g = gold_positions
    logits = formula_logits(g["packed"].shape[0])
    got = engine.policyPriorsHalf(logits[:, 0], logits[:, 1], g["packed"], g["moves"], g["offsets"])
    want = g["priors_formula_f16"]
    ulp = np.abs(got.view(np.int32).astype(np.int64) - want.view(np.int32).astype(np.int64))
    assert ulp.max() <= 1, ulp.max()
    assert (ulp == 0).mean() > 0.98
    sums = np.add.reduceat(got, g["offsets"][:-1].astype(np.int64))
    assert np.allclose(sums, 1.0, atol=1e-6)


def test_priors_edge_cases(engine, oracle, gold_positions):
    g = gold_positions
    n = 4
    pos = g["packed"][:n]
    offs = g["offsets"][: 2 * n + 1]
    moves = g["moves"][: offs[-1]]
    la = np.zeros((n, 4672), np.float16)
    lb = np.zeros((n, 4672), np.float16)
    la[0, :] = np.nan            # all non-finite -> uniform over all (utils.h:136-141)
    lb[0, :] = -np.inf
    la[1, :] = 60000.0           # huge equal logits
    lb[1, ::2] = np.inf          # +inf is non-finite: masked, the rest uniform
    la[2, 0] = -20.0             # pass logit without floor
    got = engine.policyPriorsHalf(la, lb, pos, moves, offs)
    want = oracle.batch_priors_f16(la.view(np.uint16), lb.view(np.uint16), pos, moves, offs)
    assert np.allclose(got, want, rtol=0, atol=1e-7, equal_nan=True)
    a0 = got[offs[0]:offs[1]]
    assert np.allclose(a0, 1.0 / a0.size)


def test_value_shaping_matches_oracle(engine, oracle):
    rng = np.random.default_rng(5)
    n = 4096
    value = rng.uniform(-1.5, 1.5, n).astype(np.float16)
    wdl = (rng.standard_normal((n, 3)) * 3).astype(np.float16)
    ml = rng.uniform(-0.2, 1.2, n).astype(np.float16)
    value[:4] = [np.nan, np.inf, -np.inf, 0.5]
    wdl[4] = [np.inf, 0, 0]
    wdl[5] = [np.nan, 0, 0]
    got = engine.shapeValuesHalf(value, wdl, ml)
    want = np.array([oracle.shape_value_half(int(v), w, int(m)) for v, w, m in
                     zip(value.view(np.uint16), wdl.view(np.uint16), ml.view(np.uint16))], np.float32)
    assert np.allclose(got, want, rtol=0, atol=2e-7)


def test_value_shaping_matches_the_reference_search(engine):
    """D5 against the reference's own compiled SearchThread::process_batch (tests/golden/value_shaping.npz,
    oracle/make_golden_d5.py): <= 1 ulp (the g++ reference build contracts one FMA), NaN where the reference has
    NaN, every runtime configuration of the fixture through hm_engine_set_value_config."""
    import os

    from conftest import GOLD

    g = np.load(os.path.join(GOLD, "value_shaping.npz"))
    c = g["cases_u16"]
    value, wdl, ml = c[:, 0].copy().view(np.float16), c[:, 1:4].copy().view(np.float16), c[:, 4].copy().view(np.float16)
    try:
        for i, (en, w, disc, contempt) in enumerate(g["configs"]):
            engine.setValueConfig(bool(en), float(w), float(disc), float(contempt))
            got = engine.shapeValuesHalf(value, wdl, ml)
            want = g[f"shaped_{i}"]
            assert np.array_equal(np.isnan(got), np.isnan(want)), i
            m = ~np.isnan(want)
            assert np.abs(got[m] - want[m]).max() <= 1.2e-7, i
    finally:
        engine.setValueConfig()
