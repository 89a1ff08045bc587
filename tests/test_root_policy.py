"""Root policy stages: temperature (tools/selfplay.cc:167-185), root Dirichlet
noise (search/node.h:288-315) and raw-policy joint sampling (tools/selfplay.cc:301-376 with
environment/joint_action.h:14-110).  CPU tests pin the C restatement (oracle/hm_oracle.c) against the
fixture generated from the reference (tests/golden/root_policy.npz, oracle/make_golden_root.py) and,
where oracle/_ref exists, against the reference code live; GPU tests compare the device kernel
(csrc/hm_root.cu, through the C ABI) with both."""
import os

import numpy as np
import pytest

from conftest import GOLD

SALT = (0x9E3779B97F4A7C15, 0xBF58476D1CE4E5B9)  # node.h:313-314


@pytest.fixture(scope="module")
def gold_root():
    return dict(np.load(os.path.join(GOLD, "root_policy.npz")))


def uniforms_from_draws(canonical, offsets, order):
    """The host recipe of INTEGRATION.md: canonical doubles in the order the reference build consumes them
    (draw_order: 'BA' for g++, which evaluates the arguments of selfplay.cc:365 right to left; a list of
    < 2 entries draws nothing) -> [n,3] {uA, uB, uJoint}."""
    n = canonical.shape[0]
    u = np.zeros((n, 3))
    for g in range(n):
        na, nb = int(offsets[2 * g + 1] - offsets[2 * g]), int(offsets[2 * g + 2] - offsets[2 * g + 1])
        k = 0
        for b in ((1, 0) if order == "BA" else (0, 1)):
            if (nb if b else na) >= 2:
                u[g, b] = canonical[g, k]
                k += 1
        u[g, 2] = canonical[g, k] if k < 3 else 0.0
    return u


def flags_of(packed, g):
    fa, fb = int(packed[g]["board"][0]["flags"]), int(packed[g]["board"][1]["flags"])
    team, tadv = (fa >> 5) & 1, (fa >> 6) & 1
    return int((fa & 1) == team), int((fb & 1) != team), tadv


def oracle_picks(oracle, packed, moves, offsets, priors, captures, u):
    n = packed.shape[0]
    picks, st = np.zeros((n, 2), np.uint32), np.zeros(n, np.int32)
    for g in range(n):
        a0, a1, b1 = int(offsets[2 * g]), int(offsets[2 * g + 1]), int(offsets[2 * g + 2])
        st[g], picks[g, 0], picks[g, 1] = oracle.sample_joint(packed[g:g + 1], moves[a0:a1], priors[a0:a1], captures[a0:a1],
                                                              moves[a1:b1], priors[a1:b1], captures[a1:b1], u[g])
    return picks, st


def syn_case(gold_root):
    from oracle.pyoracle import POS_DTYPE

    packed = np.ascontiguousarray(gold_root["syn_packed"]).view(POS_DTYPE).reshape(-1)
    return packed, gold_root["syn_moves"], gold_root["syn_offsets"], gold_root["syn_priors"], gold_root["syn_captures"]


def check_against_reference_rc(picks, st, ref_picks, ref_rc, offsets=None, ref_draws=None):
    ok = ref_rc >= 0
    assert np.array_equal(np.minimum(st[ok], 1), ref_rc[ok])  # 2 = resampled over a single legal pair
    assert np.array_equal(picks[ok], ref_picks[ok])
    assert np.all(st[~ok] == -3)  # HM_ROOT_NO_LEGAL_JOINT where the reference throws
    if ref_draws is not None:  # the status tells the host how many canonical doubles the reference consumed
        sizes = np.diff(offsets.astype(np.int64)).reshape(-1, 2)
        mine = (sizes[:, 0] >= 2).astype(np.int64) + (sizes[:, 1] >= 2) + (st == 1)
        assert np.array_equal(mine[ok], ref_draws[ok])


# ---------------------------------------------------------------- oracle vs reference fixture (CPU)
def test_mt19937_64_canonical_matches_reference(oracle, gold_root):
    for s, want in zip(gold_root["seeds"][:64], gold_root["canonical"][:64]):
        assert np.array_equal(oracle.canonical(int(s), 3), want)
    # std::mt19937_64 default-seed known answer (ISO C++ [rand.predef]: 10000th value of seed 5489)
    from oracle.pyoracle import _ptr  # noqa: F401
    big = oracle.canonical(5489, 10000)
    assert big[-1] == float(9981545732273789042) / 2.0**64


def test_temperature_oracle_bit_exact(oracle, gold_root, gold_positions, gold_forward):
    pri, off = gold_forward["priors_f32"], gold_positions["offsets"]
    for ti, T in enumerate(gold_root["temps"]):
        want = gold_root[f"priors_t{ti}"]
        for k in range(off.size - 1):
            lo, hi = int(off[k]), int(off[k + 1])
            assert np.array_equal(oracle.apply_temperature(pri[lo:hi], float(T)), want[lo:hi])
    with pytest.raises(ValueError):
        oracle.apply_temperature(pri[:4], 0.0)  # selfplay.cc:168-170 throws
    assert np.array_equal(oracle.apply_temperature(np.zeros(5, np.float32), 0.5), np.full(5, 0.2, np.float32))  # :177-181


def test_joint_sampling_oracle_matches_reference_picks(oracle, gold_root, gold_positions):
    order = str(gold_root["draw_order"])
    off, packed, moves = gold_positions["offsets"], gold_positions["packed"], gold_positions["moves"]
    u = uniforms_from_draws(gold_root["canonical"], off, order)
    for ti in range(len(gold_root["temps"])):
        picks, st = oracle_picks(oracle, packed, moves, off, gold_root[f"priors_t{ti}"], gold_root["captures"], u)
        check_against_reference_rc(picks, st, gold_root[f"picks_t{ti}"], gold_root[f"rc_t{ti}"], off, gold_root[f"draws_t{ti}"])


def test_joint_sampling_oracle_synthetic_rules(oracle, gold_root):
    packed, moves, off, pri, cap = syn_case(gold_root)
    u = uniforms_from_draws(gold_root["syn_canonical"], off, str(gold_root["draw_order"]))
    picks, st = oracle_picks(oracle, packed, moves, off, pri, cap, u)
    check_against_reference_rc(picks, st, gold_root["syn_picks"], gold_root["syn_rc"], off, gold_root["syn_draws"])
    assert (gold_root["syn_rc"] == 1).sum() > 100  # the resample path is really exercised


def test_dirichlet_mix_oracle_matches_reference(oracle, gold_root, gold_positions, gold_forward):
    pri, off = gold_forward["priors_f32"], gold_positions["offsets"]
    nz, want = gold_root["noise_ref"], gold_root["mix_ref"]
    for k in range(off.size - 1):
        lo, hi = int(off[k]), int(off[k + 1])
        got = oracle.dirichlet_mix(pri[lo:hi], nz[lo:hi], float(gold_root["epsilon"]))
        # the reference build may contract (1-eps)*p + x into an FMA (g++ -ffp-contract=fast): <= 1 ulp
        assert np.all(np.abs(got.view(np.int32).astype(np.int64) - want[lo:hi].view(np.int32)) <= 1)
        if hi - lo > 1:
            assert abs(float(got.sum()) - 1.0) < 1e-5


def test_oracle_matches_live_reference(oracle, ref, gold_root):
    packed, moves, off, pri, cap = syn_case(gold_root)
    for g in range(0, 256):
        a0, a1, b1 = int(off[2 * g]), int(off[2 * g + 1]), int(off[2 * g + 2])
        a_on, b_on, tadv = flags_of(packed, g)
        seed = int(gold_root["syn_seeds"][g])
        rc, ia, ib, draws = ref.sample_joint(moves[a0:a1], pri[a0:a1], cap[a0:a1], moves[a1:b1], pri[a1:b1], cap[a1:b1],
                                             a_on, b_on, tadv, seed)
        assert (rc, ia, ib, draws) == (gold_root["syn_rc"][g], *gold_root["syn_picks"][g], gold_root["syn_draws"][g])
    assert np.array_equal(ref.canonical(123, 5), oracle.canonical(123, 5))


# ---------------------------------------------------------------- device kernel (GPU)
@pytest.fixture(scope="module")
def engine():
    from hivemind_b200 import weights as W
    from hivemind_b200.engine import Engine

    e = Engine(0, 256, 1)
    assert e.loadNetworkFromMemory(W.random_hvmw(seed=0))
    yield e
    e.close()


@pytest.mark.gpu
def test_device_temperature_and_picks_match_reference(engine, oracle, gold_root, gold_positions, gold_forward):
    off, packed, moves = gold_positions["offsets"], gold_positions["packed"], gold_positions["moves"]
    u = uniforms_from_draws(gold_root["canonical"], off, str(gold_root["draw_order"]))
    pri = gold_forward["priors_f32"]
    for ti, T in enumerate(gold_root["temps"]):
        out = engine.rootPolicyApply(packed, moves, off, pri, temperature=float(T), captures=gold_root["captures"], uniforms=u)
        assert out is not None
        got, _, picks, st = out
        want = gold_root[f"priors_t{ti}"]
        if T == 1.0:
            assert np.array_equal(got, want)  # pow(x, 1) is exact: bit-identical to the reference
        else:  # device pow vs glibc pow: both < 1 ulp in double, the float cast may differ in the last bit
            d = np.abs(got.view(np.int32).astype(np.int64) - want.view(np.int32))
            assert d.max() <= 1 and (d > 0).mean() < 1e-3
        check_against_reference_rc(picks, st, gold_root[f"picks_t{ti}"], gold_root[f"rc_t{ti}"], off, gold_root[f"draws_t{ti}"])


@pytest.mark.gpu
def test_device_joint_sampling_synthetic_rules(engine, gold_root):
    packed, moves, off, pri, cap = syn_case(gold_root)
    u = uniforms_from_draws(gold_root["syn_canonical"], off, str(gold_root["draw_order"]))
    out = engine.rootPolicyApply(packed, moves, off, pri, captures=cap, uniforms=u)
    assert out is not None
    got, _, picks, st = out
    assert np.array_equal(got, pri)  # temperature 0 = stage skipped
    check_against_reference_rc(picks, st, gold_root["syn_picks"], gold_root["syn_rc"], off, gold_root["syn_draws"])


@pytest.mark.gpu
def test_device_dirichlet_noise(engine, oracle, gold_root, gold_positions, gold_forward):
    from scipy import stats

    off, packed, moves = gold_positions["offsets"], gold_positions["packed"], gold_positions["moves"]
    pri = gold_forward["priors_f32"]
    n = packed.shape[0]
    alpha, eps = float(gold_root["alpha"]), float(gold_root["epsilon"])
    hashes = gold_root["seeds"]
    out = engine.rootPolicyApply(packed, moves, off, pri, dirichlet_alpha=alpha, dirichlet_epsilon=eps, noise_seed=99,
                                 position_hash=hashes)
    assert out is not None
    got, noise, _, _ = out
    sizes = np.diff(off.astype(np.int64))
    multi = np.repeat(sizes > 1, sizes)
    # the mix (node.h:299-311) is exact given the variates the kernel drew
    for k in range(2 * n):
        lo, hi = int(off[k]), int(off[k + 1])
        assert np.array_equal(got[lo:hi], oracle.dirichlet_mix(pri[lo:hi], noise[lo:hi], eps))
    assert np.array_equal(got[~multi], pri[~multi]) and np.all(noise[~multi] == 0)  # node.h:292-294
    # the variates are Gamma(alpha, 1) like std::gamma_distribution<float>(alpha, 1): moments + KS, and the same
    # checks pass for the reference's own stream (fixture noise_ref)
    for sample in (noise[multi], gold_root["noise_ref"][multi]):
        assert abs(sample.mean() - alpha) < 0.02 and abs(sample.var() - alpha) < 0.04
        assert stats.kstest(sample.astype(np.float64), "gamma", args=(alpha,)).pvalue > 1e-3
    # deterministic in (seed, hash, board, move index); a different seed gives a different stream
    again = engine.rootPolicyApply(packed, moves, off, pri, dirichlet_alpha=alpha, dirichlet_epsilon=eps, noise_seed=99,
                                   position_hash=hashes)
    other = engine.rootPolicyApply(packed, moves, off, pri, dirichlet_alpha=alpha, dirichlet_epsilon=eps, noise_seed=100,
                                   position_hash=hashes)
    assert np.array_equal(again[1], noise) and not np.array_equal(other[1][multi], noise[multi])
    # boards A and B of one position use different salts (node.h:313-314)
    a0, a1, b1 = int(off[0]), int(off[1]), int(off[2])
    m = min(a1 - a0, b1 - a1)
    assert m < 2 or not np.array_equal(noise[a0:a0 + m], noise[a1:a1 + m])
    # disabled when alpha or epsilon is 0 or no hashes are given (node.h:288-290)
    for kw in ({"dirichlet_alpha": 0.0, "dirichlet_epsilon": eps, "position_hash": hashes},
               {"dirichlet_alpha": alpha, "dirichlet_epsilon": 0.0, "position_hash": hashes},
               {"dirichlet_alpha": alpha, "dirichlet_epsilon": eps}):
        assert np.array_equal(engine.rootPolicyApply(packed, moves, off, pri, **kw)[0], pri)


@pytest.mark.gpu
def test_fused_forward_and_root_stages(engine, oracle, gold_root, gold_positions):
    off, packed, moves = gold_positions["offsets"], gold_positions["packed"], gold_positions["moves"]
    u = uniforms_from_draws(gold_root["canonical"], off, str(gold_root["draw_order"]))
    base = engine.evaluatePositions(packed, moves, off)
    assert base is not None
    pri = np.array(base[0], copy=True)
    values = np.array(base[1], copy=True)
    T = 1.6
    assert engine.enqueuePositionsRoot(packed, moves, off, 0, temperature=T, captures=gold_root["captures"], uniforms=u)
    assert engine.synchronizePriors(0) is None  # a root request is collected with synchronizeRoot ...
    assert not engine.enqueuePositions(packed, moves, off, 0)  # ... and still counts as the one request in flight
    got = engine.synchronizeRoot(0)
    assert got is not None
    priors, vals, raw, picks, st = got
    assert np.array_equal(vals, values)
    want = pri.copy()
    for k in range(off.size - 1):
        lo, hi = int(off[k]), int(off[k + 1])
        want[lo:hi] = oracle.apply_temperature(pri[lo:hi], T)
    d = np.abs(np.asarray(priors).view(np.int32).astype(np.int64) - want.view(np.int32))
    assert d.max() <= 1 and (d > 0).mean() < 1e-3
    opicks, ost = oracle_picks(oracle, packed, moves, off, np.asarray(priors), gold_root["captures"], u)
    assert np.array_equal(ost, st) and np.array_equal(opicks[ost >= 0], np.asarray(picks)[ost >= 0])
    assert engine.synchronizeRoot(0) is None  # nothing pending any more
    # a plain request afterwards is unaffected by the root stages
    again = engine.evaluatePositions(packed, moves, off)
    assert np.array_equal(again[0], pri)


@pytest.mark.gpu
def test_root_policy_errors_and_edges(engine, gold_positions, gold_forward):
    off, packed, moves = gold_positions["offsets"], gold_positions["packed"], gold_positions["moves"]
    pri = gold_forward["priors_f32"]
    from hivemind_b200 import _lib

    assert engine.rootPolicyApply(packed, moves, off, pri, temperature=-1.0) is None
    assert "temperature" in _lib.last_error()
    assert not engine.enqueuePositionsRoot(packed, moves, off, 0, temperature=-0.5)
    # empty batch is a no-op; an empty move list is reported, not sampled
    assert engine.rootPolicyApply(packed[:0], moves[:0], np.zeros(1, np.uint32), pri[:0]) is not None
    off2 = np.array([0, 0, 3], np.uint32)
    out = engine.rootPolicyApply(packed[:1], moves[:3], off2, pri[:3], uniforms=np.zeros(3))
    assert out is not None and out[3][0] == _lib.ROOT_EMPTY_LIST
    # all-zero priors on a >= 2 list cannot be sampled (std::discrete_distribution needs a positive sum)
    off3 = np.array([0, 2, 4], np.uint32)
    out = engine.rootPolicyApply(packed[:1], np.array([5, 0, 7, 0], np.uint32), off3, np.zeros(4, np.float32), uniforms=np.zeros(3))
    assert out is not None and out[3][0] == _lib.ROOT_BAD_WEIGHTS


def test_cpp_shim_keeps_a_std_engine_in_step(tmp_path):
    """include/hm_engine.hpp drawRootUniforms / commitRootDraw hand the device the very canonical doubles
    std::discrete_distribution would take from a std::mt19937_64 and leave the engine where the reference
    would (selfplay.cc:363-365,389-390)."""
    import subprocess

    from conftest import ROOT
    from hivemind_b200 import build as hb

    lib = hb.build()
    src = tmp_path / "t.cc"
    src.write_text(r'''
#include <random>
#include <cstdio>
#include "hm_engine.hpp"
int main() {
    for (unsigned long long seed = 1; seed < 200; ++seed)
        for (size_t nA = 1; nA < 4; ++nA)
            for (size_t nB = 1; nB < 4; ++nB)
                for (int resample = 0; resample < 3; ++resample) {
                    std::mt19937_64 ours(seed), theirs(seed);
                    double u[3];
                    Engine::drawRootUniforms(ours, nA, nB, u);
                    std::vector<double> wA(nA, 1.0), wB(nB, 1.0), wJ(resample == 2 ? 1 : 3, 1.0);
                    std::discrete_distribution<size_t> dA(wA.begin(), wA.end()), dB(wB.begin(), wB.end()), dJ(wJ.begin(), wJ.end());
                    std::mt19937_64 probe = theirs;
                    double want[3] = {0, 0, 0};
                    if (nB >= 2) want[1] = std::generate_canonical<double, 53>(probe);
                    if (nA >= 2) want[0] = std::generate_canonical<double, 53>(probe);
                    want[2] = std::generate_canonical<double, 53>(probe);
                    if (u[0] != want[0] || u[1] != want[1] || u[2] != want[2]) { std::puts("uniform mismatch"); return 1; }
                    (void)dB(theirs); (void)dA(theirs);             // g++ order of selfplay.cc:365
                    if (resample) (void)dJ(theirs);                  // one weight: no draw
                    Engine::commitRootDraw(ours, resample == 0 ? HM_ROOT_OK : resample == 1 ? HM_ROOT_RESAMPLED : HM_ROOT_RESAMPLED_FORCED);
                    if (!(ours == theirs)) { std::printf("engine out of step seed %llu nA %zu nB %zu r %d\n", seed, nA, nB, resample); return 2; }
                }
    std::puts("ok");
    return 0;
}
''')
    out = tmp_path / "t"
    cmd = ["g++", "-O1", "-std=c++17", "-I" + os.path.join(ROOT, "include"), str(src), "-L" + os.path.dirname(lib),
           "-lhm_engine", "-Wl,-rpath," + os.path.dirname(lib), "-o", str(out)]
    if os.path.isdir("/usr/local/cuda/include"):
        cmd.insert(4, "-I/usr/local/cuda/include")
    r = subprocess.run(cmd, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    r = subprocess.run([str(out)], capture_output=True, text=True)
    assert r.returncode == 0 and "ok" in r.stdout, r.stdout + r.stderr
