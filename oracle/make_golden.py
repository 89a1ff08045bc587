"""make_golden.py — generates tests/golden/*.npz FROM THE REFERENCE ITSELF.

Runs only where /root/reference exists (this container): the compiled reference C++
(oracle/_ref/libhmref.so, `make -C oracle ref`) for stages P and D, and the reference PyTorch
RiseV3 (imported with the DropPath stub in oracle/timm_stub) for stage F.  The fixtures travel
to the GPU box in git; nothing there reads /root/reference.

    python oracle/make_golden.py
"""
from __future__ import annotations

import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from hivemind_b200 import weights as W  # noqa: E402
from oracle.forward_oracle import deployed_precision_state_dict, reference_forward, reference_model  # noqa: E402
from oracle.pyoracle import POS_DTYPE, Ref, gen_config1  # noqa: E402

GOLD = os.path.join(ROOT, "tests", "golden")
SEED, N = 1234, 256
WEIGHT_SEED = 0


def formula_logits(n_pos: int) -> np.ndarray:
    """Deterministic fp16 policy maps [n_pos, 2, 4672] (same formula in the tests)."""
    i = np.arange(4672, dtype=np.int64)[None, None, :]
    p = np.arange(n_pos, dtype=np.int64)[:, None, None]
    b = np.arange(2, dtype=np.int64)[None, :, None]
    v = ((i * 7 + p * 13 + b * 5) % 41 - 20).astype(np.float32) * 0.25
    return v.astype(np.float16)


def kat_boards(ref: Ref):
    """The positions of the reference's own known-answer tests (engine/tests/test_move_gen.cc)."""
    cases = {}
    b = ref.board()
    cases["start_white"] = (b, 0, 0)                       # :68-83
    b = ref.board()
    for u in ("g1f3", "g8f6", "f3g1"):
        b.push_uci(0, u)
    cases["half_black_tadv"] = (b, 1, 1)                   # :85-103
    b = ref.board()
    b.push_uci(0, "g1f3")
    cases["g1f3_white"] = (b, 0, 0)                        # :105-125
    cases["g1f3_black"] = (b, 1, 0)
    b = ref.board()
    for u in ("g1f3", "g8f6", "f3g1", "f6g8"):
        b.push_uci(0, u)
    cases["repetition_white"] = (b, 0, 0)                  # :145-157
    b = ref.board()
    b.set_fen(0, "7k/ppp2rpP/8/4p1qP/3nP1N1/2NP1P2/PPP2K2/3R3R[Qpp] w - - 0 24")
    cases["low_prior_check"] = (b, 0, 0)                   # :695-729
    b = ref.board()
    b.set("2rq1rk1/pppnb1p1/4p1p1/3pP1pp/4P3/2N1P1B1/PPP2NPP/R2Q1RK1/NN b - - 0 3|"
          "r4rk1/ppp2p1p/4bB1p/8/6b1/2P5/P1PB1PPP/R3R1K1/qbbnnppPB w")
    cases["fast_index_fen"] = (b, 0, 0)                    # :747-767
    return cases


def main():
    os.makedirs(GOLD, exist_ok=True)
    ref = Ref()

    # ---- stage P + D: config-1 positions ------------------------------------------------
    boards, packed, moves, offsets = gen_config1(ref, SEED, N)
    planes_f32 = np.stack([b.planes(g & 1, (g >> 1) & 1, "f32") for g, b in enumerate(boards)])
    planes_f16 = np.stack([b.planes(g & 1, (g >> 1) & 1, "f16") for g, b in enumerate(boards)])
    planes_u8 = np.stack([b.planes(g & 1, (g >> 1) & 1, "u8") for g, b in enumerate(boards)])
    logits = formula_logits(N)
    priors = np.zeros(moves.size, np.float32)
    indices = np.zeros(moves.size, np.int32)
    stm = np.zeros(moves.size, np.uint8)
    for g, b in enumerate(boards):
        for bn in range(2):
            lo, hi = offsets[2 * g + bn], offsets[2 * g + bn + 1]
            priors[lo:hi] = b.priors_f16(bn, logits[g, bn].view(np.uint16), moves[lo:hi])
            stm[lo:hi] = b.stm_black(bn)
            indices[lo:hi] = [ref.policy_index(m, b.stm_black(bn)) for m in moves[lo:hi]]
    # all legal moves (incl. R/B under-promotions) for the index check
    all_moves, all_stm, all_idx = [], [], []
    for b in boards:
        for bn in range(2):
            for m in b.legal(bn):
                all_moves.append(m)
                all_stm.append(b.stm_black(bn))
                all_idx.append(ref.policy_index(m, b.stm_black(bn)))
    regen = "--all" in sys.argv   # committed fixtures are rewritten only on request (python oracle/make_golden.py --all)

    def save(fname, **arrays):
        if regen or not os.path.exists(os.path.join(GOLD, fname)):
            np.savez_compressed(os.path.join(GOLD, fname), **arrays)

    save(
        "positions_seed1234_n256.npz",
        packed=packed.view(np.uint8).reshape(N, 256), moves=moves, offsets=offsets,
        planes_f32=planes_f32, planes_f16=planes_f16, planes_u8=planes_u8,
        priors_formula_f16=priors, policy_index=indices, stm_black=stm,
        all_moves=np.array(all_moves, np.uint32), all_stm=np.array(all_stm, np.uint8),
        all_idx=np.array(all_idx, np.int32),
    )
    normal, drop = ref.policy_tables()
    save("policy_tables.npz", normal=normal.astype(np.int16), drop=drop.astype(np.int16))

    # ---- reference known-answer positions -----------------------------------------------
    kat = {}
    for name, (b, team, tadv) in kat_boards(ref).items():
        kat[name + ".packed"] = b.pack(team, tadv).view(np.uint8).reshape(256)
        kat[name + ".planes_f32"] = b.planes(team, tadv, "f32")
        kat[name + ".planes_f16"] = b.planes(team, tadv, "f16")
        for bn in range(2):
            acts = np.concatenate([b.legal(bn), np.zeros(1, np.uint32)])  # legal + MOVE_NONE (tests :662-664)
            kat[name + f".actions{bn}"] = acts
            kat[name + f".stm{bn}"] = np.array(b.stm_black(bn), np.uint8)
            kat[name + f".uci{bn}"] = np.array([b.move_uci(bn, m) for m in acts])
    save("reference_kat.npz", **kat)

    # ---- stages F and X: reference PyTorch forward on seeded weights -------------------------
    # forward_<name>.npz: the UNMODIFIED reference module in fp32.  forward_<name>_fp16w.npz: the same module with
    # its tensor-core operand weights at the deployed precision (fp16, engine/src/nn/engine.cc:331-354;
    # forward_oracle.deployed_precision_state_dict) and fp32 arithmetic — device minus THIS is kernel error alone
    # (activation rounding, accumulation order), device minus the fp32 fixture adds the weight quantisation.
    fixtures = (
        (W.risev33_spec(), "forward_risev33", True),
        (W.crossboard_spec(), "forward_crossboard", True),
        (W.risev33_spec(shared_policy_trunk=False, value_head="mlp"), "forward_risev33_perboard_mlp", False),
        (W.risev33_spec(value_head="tanh"), "forward_risev33_tanh", False),
    )
    for spec, name, with_deployed in fixtures:
        sd = W.random_state_dict(spec, WEIGHT_SEED)
        variants = [("", sd)]
        if with_deployed:
            variants.append(("_fp16w", deployed_precision_state_dict(sd, spec)))
        for suffix, state in variants:
            fname = name + suffix + ".npz"
            if os.path.exists(os.path.join(GOLD, fname)) and not regen:
                continue
            model = reference_model(state)
            out = reference_forward(model, planes_f32)
            legal_logits = np.zeros(moves.size, np.float32)
            ref_priors = np.zeros(moves.size, np.float32)
            for g, b in enumerate(boards):
                for bn in range(2):
                    lo, hi = offsets[2 * g + bn], offsets[2 * g + bn + 1]
                    pol = out["pi_a" if bn == 0 else "pi_b"][g]
                    idx = indices[lo:hi]
                    legal_logits[lo:hi] = np.where(idx >= 0, pol[np.maximum(idx, 0)], -np.inf)
                    ref_priors[lo:hi] = b.priors_f32(bn, pol, moves[lo:hi])
            save(
                fname,
                weight_seed=np.array(WEIGHT_SEED), value=out["value"], wdl=out["wdl"], moves_left=out["moves_left"],
                legal_logits=legal_logits, priors_f32=ref_priors,
                pi_a_first8=out["pi_a"][:8], pi_b_first8=out["pi_b"][:8],
                param_count=np.array(sum(p.numel() for p in model.parameters())),
            )
    for f in sorted(os.listdir(GOLD)):
        print(f, os.path.getsize(os.path.join(GOLD, f)))


if __name__ == "__main__":
    main()
