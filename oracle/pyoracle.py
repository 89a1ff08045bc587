"""ctypes bindings for the CHECKERS — TEST INFRASTRUCTURE, never imported by hivemind_b200/.

* ``Oracle``  : oracle/liboracle.so, our plain-C restatement (oracle/hm_oracle.c).
* ``Ref``     : oracle/_ref/libhmref.so, the unmodified reference compiled from
  /root/reference by ``make -C oracle ref`` (exists only where it was built and travels to
  the GPU box as a prebuilt file).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs
import this module.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ORACLE_SO = os.path.join(HERE, "liboracle.so")
REF_SO = os.path.join(HERE, "_ref", "libhmref.so")
REF_LEAFBATCH_SO = os.path.join(HERE, "_ref", "libhmref_leafbatch.so")  # reference Board + product LeafBatch harness
REF_D5_SO = os.path.join(HERE, "_ref", "libhmref_d5.so")                # reference searchthread.cc on a scripted Engine

NB_INPUT_VALUES = 4736
NB_POLICY_VALUES = 4672
PACKED_DTYPE = np.dtype(
    [
        ("pieces", "<u8", (12,)),
        ("promoted", "<u8", (2,)),
        ("pocket", "u1", (10,)),
        ("ep_square", "u1"),
        ("flags", "u1"),
        ("last_from", "u1"),
        ("last_to", "u1"),
        ("rule50", "u1"),
        ("repetitions", "u1"),
    ]
)
assert PACKED_DTYPE.itemsize == 128
POS_DTYPE = np.dtype([("board", PACKED_DTYPE, (2,))])
assert POS_DTYPE.itemsize == 256

FLAG_STM_BLACK = 0x01
FLAG_TEAM_BLACK = 0x20
FLAG_TIME_ADVANTAGE = 0x40


class ValueConfig(C.Structure):
    _fields_ = [
        ("enable_wdl_eval", C.c_int),
        ("wdl_value_weight", C.c_float),
        ("moves_left_discount", C.c_float),
        ("draw_contempt", C.c_float),
    ]

    @classmethod
    def default(cls):
        # search_params.h:122-128,148
        return cls(1, 0.25, 0.005, 0.0)


def _ptr(a):
    return a.ctypes.data_as(C.c_void_p)


def build_oracle():
    if not os.path.exists(ORACLE_SO) or os.path.getmtime(ORACLE_SO) < os.path.getmtime(
        os.path.join(HERE, "hm_oracle.c")
    ):
        subprocess.check_call(["make", "-C", HERE, "oracle"], stdout=subprocess.DEVNULL)
    return ORACLE_SO


class Oracle:
    """Plain-C restatement (oracle/hm_oracle.c)."""

    def __init__(self):
        self.lib = C.CDLL(build_oracle())
        L = self.lib
        L.hmo_float_to_half.restype = C.c_uint16
        L.hmo_float_to_half.argtypes = [C.c_float]
        L.hmo_half_to_float.restype = C.c_float
        L.hmo_half_to_float.argtypes = [C.c_uint16]
        L.hmo_float_to_bf16.restype = C.c_uint16
        L.hmo_float_to_bf16.argtypes = [C.c_float]
        L.hmo_policy_index.restype = C.c_int32
        L.hmo_policy_index.argtypes = [C.c_uint32, C.c_int]
        L.hmo_shape_value.restype = C.c_float
        L.hmo_shape_value.argtypes = [C.c_float, C.c_void_p, C.c_void_p, C.c_void_p]
        L.hmo_shape_value_half.restype = C.c_float
        L.hmo_shape_value_half.argtypes = [C.c_uint16, C.c_void_p, C.c_void_p, C.c_void_p]
        for name in ("f32", "f16", "bf16", "u8"):
            getattr(L, f"hmo_encode_planes_{name}").argtypes = [C.c_void_p, C.c_size_t, C.c_void_p]
        L.hmo_normalize_logits.argtypes = [C.c_void_p, C.c_size_t, C.c_void_p]
        L.hmo_priors_f16.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_int, C.c_void_p]
        L.hmo_priors_f32.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_int, C.c_void_p]
        L.hmo_batch_priors_f16.argtypes = [C.c_void_p] * 3 + [C.c_size_t] + [C.c_void_p] * 3
        L.hmo_policy_tables.argtypes = [C.c_void_p, C.c_void_p]
        L.hmo_apply_temperature.argtypes = [C.c_void_p, C.c_size_t, C.c_double]
        L.hmo_dirichlet_mix.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_float]
        L.hmo_discrete_pick_f32.restype = C.c_size_t
        L.hmo_discrete_pick_f32.argtypes = [C.c_void_p, C.c_size_t, C.c_double]
        L.hmo_sample_joint.argtypes = [C.c_void_p] * 4 + [C.c_size_t] + [C.c_void_p] * 3 + [C.c_size_t] + [C.c_void_p] * 3
        L.hmo_mt19937_64_canonical.argtypes = [C.c_uint64, C.c_size_t, C.c_void_p]
        L.hmo_policy_init()

    # ---- root policy stages (selfplay.cc:167-185,301-376; node.h:288-315) ----
    def apply_temperature(self, p, temperature: float) -> np.ndarray:
        out = np.array(p, np.float32, copy=True)
        if not self.lib.hmo_apply_temperature(_ptr(out), out.size, temperature):
            raise ValueError("Policy temperature must be positive")
        return out

    def dirichlet_mix(self, p, noise, epsilon: float) -> np.ndarray:
        out = np.array(p, np.float32, copy=True)
        nz = np.ascontiguousarray(noise, np.float32)
        self.lib.hmo_dirichlet_mix(_ptr(out), _ptr(nz), out.size, epsilon)
        return out

    def discrete_pick(self, p, u: float) -> int:
        a = np.ascontiguousarray(p, np.float32)
        return int(self.lib.hmo_discrete_pick_f32(_ptr(a), a.size, u))

    def sample_joint(self, pos1, moves_a, p_a, cap_a, moves_b, p_b, cap_b, u3):
        """-> (status, pick_a, pick_b) for ONE packed position."""
        pos1 = np.ascontiguousarray(pos1)
        ma, mb = np.ascontiguousarray(moves_a, np.uint32), np.ascontiguousarray(moves_b, np.uint32)
        pa, pb = np.ascontiguousarray(p_a, np.float32), np.ascontiguousarray(p_b, np.float32)
        ca, cb = np.ascontiguousarray(cap_a, np.uint8), np.ascontiguousarray(cap_b, np.uint8)
        u = np.ascontiguousarray(u3, np.float64)
        ia, ib = C.c_uint32(), C.c_uint32()
        st = self.lib.hmo_sample_joint(_ptr(pos1), _ptr(ma), _ptr(pa), _ptr(ca), ma.size, _ptr(mb), _ptr(pb), _ptr(cb),
                                       mb.size, _ptr(u), C.byref(ia), C.byref(ib))
        return int(st), int(ia.value), int(ib.value)

    def canonical(self, seed: int, n: int) -> np.ndarray:
        out = np.zeros(n, np.float64)
        self.lib.hmo_mt19937_64_canonical(seed, n, _ptr(out))
        return out

    def encode_planes(self, pos: np.ndarray, dtype: str = "f32") -> np.ndarray:
        pos = np.ascontiguousarray(pos)
        n = pos.shape[0]
        npdt = {"f32": np.float32, "f16": np.uint16, "bf16": np.uint16, "u8": np.uint8}[dtype]
        out = np.empty((n, 74, 8, 8), npdt)
        getattr(self.lib, f"hmo_encode_planes_{dtype}")(_ptr(pos), n, _ptr(out))
        return out

    def policy_tables(self):
        normal = np.empty((2, 64, 64, 2), np.int32)
        drop = np.empty((2, 64, 8), np.int32)
        self.lib.hmo_policy_tables(_ptr(normal), _ptr(drop))
        return normal, drop

    def policy_index(self, move: int, stm_black: int) -> int:
        return self.lib.hmo_policy_index(int(move), int(stm_black))

    def normalize_logits(self, logits) -> np.ndarray:
        l = np.ascontiguousarray(logits, np.float32)
        out = np.empty_like(l)
        self.lib.hmo_normalize_logits(_ptr(l), l.size, _ptr(out))
        return out

    def priors_f16(self, policy_u16, moves, stm_black) -> np.ndarray:
        p = np.ascontiguousarray(policy_u16, np.uint16)
        m = np.ascontiguousarray(moves, np.uint32)
        out = np.empty(m.size, np.float32)
        self.lib.hmo_priors_f16(_ptr(p), _ptr(m), m.size, int(stm_black), _ptr(out))
        return out

    def priors_f32(self, policy, moves, stm_black) -> np.ndarray:
        p = np.ascontiguousarray(policy, np.float32)
        m = np.ascontiguousarray(moves, np.uint32)
        out = np.empty(m.size, np.float32)
        self.lib.hmo_priors_f32(_ptr(p), _ptr(m), m.size, int(stm_black), _ptr(out))
        return out

    def batch_priors_f16(self, logits_a, logits_b, pos, moves, offsets) -> np.ndarray:
        la = np.ascontiguousarray(logits_a, np.uint16)
        lb = np.ascontiguousarray(logits_b, np.uint16)
        pos = np.ascontiguousarray(pos)
        m = np.ascontiguousarray(moves, np.uint32)
        o = np.ascontiguousarray(offsets, np.uint32)
        out = np.empty(m.size, np.float32)
        self.lib.hmo_batch_priors_f16(_ptr(la), _ptr(lb), _ptr(pos), pos.shape[0], _ptr(m), _ptr(o), _ptr(out))
        return out

    def shape_value(self, value, wdl=None, moves_left=None, cfg=None) -> float:
        cfg = cfg or ValueConfig.default()
        w = None if wdl is None else np.ascontiguousarray(wdl, np.float32)
        ml = None if moves_left is None else np.array([moves_left], np.float32)
        return self.lib.hmo_shape_value(
            C.c_float(value), None if w is None else _ptr(w), None if ml is None else _ptr(ml), C.byref(cfg)
        )

    def shape_value_half(self, value_u16, wdl_u16, ml_u16, cfg=None) -> float:
        cfg = cfg or ValueConfig.default()
        w = np.ascontiguousarray(wdl_u16, np.uint16)
        ml = np.array([ml_u16], np.uint16)
        return self.lib.hmo_shape_value_half(int(value_u16), _ptr(w), _ptr(ml), C.byref(cfg))


def ref_available() -> bool:
    return os.path.exists(REF_SO)


class Ref:
    """The reference itself (oracle/_ref/libhmref.so)."""

    def __init__(self):
        if not ref_available():
            raise FileNotFoundError(f"{REF_SO} not built (make -C oracle ref; needs /root/reference)")
        self.lib = C.CDLL(REF_SO)
        L = self.lib
        vp = C.c_void_p
        L.hmref_board_new.restype = vp
        L.hmref_board_copy.restype = vp
        L.hmref_board_copy.argtypes = [vp]
        L.hmref_board_free.argtypes = [vp]
        L.hmref_board_set.argtypes = [vp, C.c_char_p]
        L.hmref_board_set_fen.argtypes = [vp, C.c_int, C.c_char_p]
        L.hmref_board_fen.argtypes = [vp, C.c_int, C.c_char_p, C.c_int]
        L.hmref_board_push_uci.argtypes = [vp, C.c_int, C.c_char_p]
        L.hmref_board_pop.argtypes = [vp, C.c_int]
        L.hmref_board_make_moves_uci.argtypes = [vp, C.c_char_p, C.c_char_p, vp, vp]
        L.hmref_board_unmake_moves.argtypes = [vp, C.c_uint32, C.c_uint32]
        L.hmref_board_stm.argtypes = [vp, C.c_int]
        L.hmref_board_gives_check.argtypes = [vp, C.c_int, C.c_uint32]
        for n in ("f32", "f16", "u8"):
            getattr(L, f"hmref_board_planes_{n}").argtypes = [vp, C.c_int, C.c_int, vp]
        L.hmref_board_pack.argtypes = [vp, C.c_int, C.c_int, vp]
        L.hmref_board_legal.argtypes = [vp, C.c_int, vp, C.c_int]
        L.hmref_board_search_actions.argtypes = [vp, C.c_int, C.c_int, vp, C.c_int]
        L.hmref_board_move_uci.argtypes = [vp, C.c_int, C.c_uint32, C.c_char_p, C.c_int]
        L.hmref_board_priors_f16.argtypes = [vp, C.c_int, vp, vp, C.c_int, vp]
        L.hmref_board_priors_f32.argtypes = [vp, C.c_int, vp, vp, C.c_int, vp]
        L.hmref_normalize_logits.argtypes = [vp, C.c_int, vp]
        L.hmref_policy_index.argtypes = [C.c_uint32, C.c_int]
        L.hmref_policy_index_of_label.argtypes = [C.c_char_p]
        L.hmref_policy_label.argtypes = [C.c_int, C.c_char_p, C.c_int]
        L.hmref_policy_tables.argtypes = [vp, vp]
        L.hmref_gen_positions.argtypes = [C.c_uint64, C.c_int, C.POINTER(vp)]
        L.hmref_time_planes_f16.restype = C.c_double
        L.hmref_time_planes_f16.argtypes = [C.POINTER(vp), C.c_int, C.c_int, C.c_int]
        L.hmref_time_priors_f16.restype = C.c_double
        L.hmref_time_priors_f16.argtypes = [C.POINTER(vp), C.c_int, vp, C.c_int, C.c_int]
        L.hmref_board_is_capture.argtypes = [vp, C.c_int, C.c_uint32]
        L.hmref_canonical.argtypes = [C.c_uint64, C.c_int, vp]
        L.hmref_apply_temperature.argtypes = [vp, C.c_int, C.c_double]
        L.hmref_sample_joint.argtypes = [vp, vp, vp, C.c_int, vp, vp, vp, C.c_int, C.c_int, C.c_int, C.c_int,
                                         C.c_uint64, vp, vp, vp]
        L.hmref_dirichlet_noise.argtypes = [C.c_uint64, C.c_float, C.c_int, vp]
        L.hmref_dirichlet_mix.argtypes = [vp, vp, C.c_int, C.c_float]
        L.hmref_product_pack.argtypes = [vp, C.c_int, C.c_int, vp]
        L.hmref_product_actions.argtypes = [vp, C.c_int, C.c_int, vp, vp, C.c_int]
        L.hmref_init()
        self._leafbatch = None
        self._d5 = None

    def d5_shape(self, cases_u16, cfg=None) -> np.ndarray:
        """The reference's OWN value shaping (searchthread.cc:569-621, compiled unmodified against a scripted Engine,
        oracle/d5_pin.cc): cases_u16 [n,5] fp16 bit patterns (value, wdl loss/draw/win, moves_left) -> the value the
        reference's one-leaf search backed up into the root."""
        if self._d5 is None:
            self._d5 = C.CDLL(REF_D5_SO)
            self._d5.hmref_d5_shape.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_float, C.c_float, C.c_float, C.c_void_p]
        cfg = cfg or ValueConfig.default()
        cases = np.ascontiguousarray(cases_u16, np.uint16).reshape(-1, 5)
        out = np.zeros(cases.shape[0], np.float32)
        bad = self._d5.hmref_d5_shape(_ptr(cases), cases.shape[0], int(cfg.enable_wdl_eval), cfg.wdl_value_weight,
                                      cfg.draw_contempt, cfg.moves_left_discount, _ptr(out))
        if bad:
            raise RuntimeError(f"{bad} one-leaf searches did not end with exactly one root visit")
        return out

    # -- the product's include/hm_batch.hpp instantiated on the reference Board ----------
    def leafbatch_run(self, boards, blob: bytes, root=None, device: int = 0):
        """LeafBatch.add/submit/collect through the C++ Engine shim -> dict, or a negative error code."""
        n = len(boards)
        cap = n * 2 * 256
        moves, caps = np.zeros(cap, np.uint32), np.zeros(cap, np.uint8)
        offsets, priors = np.zeros(2 * n + 1, np.uint32), np.zeros(cap, np.float32)
        values, picks, status = np.zeros(n, np.float32), np.zeros((n, 2), np.uint32), np.zeros(n, np.int32)
        buf = C.create_string_buffer(blob, len(blob))
        if self._leafbatch is None:
            vp = C.c_void_p
            self._leafbatch = C.CDLL(REF_LEAFBATCH_SO)
            self._leafbatch.hmref_leafbatch_run.argtypes = [C.c_int, vp, C.c_size_t, C.POINTER(vp), C.c_int, vp] + [vp] * 7 + [C.c_int]
        rc = self._leafbatch.hmref_leafbatch_run(device, buf, len(blob), self.handles(boards), n,
                                          C.byref(root) if root is not None else None, _ptr(moves), _ptr(caps),
                                          _ptr(offsets), _ptr(priors), _ptr(values), _ptr(picks), _ptr(status), cap)
        if rc < 0:
            return rc
        return {"moves": moves[:rc], "captures": caps[:rc], "offsets": offsets, "priors": priors[:rc], "values": values,
                "picks": picks, "status": status}

    # -- root policy stages ---------------------------------------------------------
    def canonical(self, seed: int, n: int) -> np.ndarray:
        out = np.zeros(n, np.float64)
        self.lib.hmref_canonical(seed, n, _ptr(out))
        return out

    def apply_temperature(self, p, temperature: float) -> np.ndarray:
        out = np.array(p, np.float32, copy=True)
        self.lib.hmref_apply_temperature(_ptr(out), out.size, temperature)
        return out

    def sample_joint(self, moves_a, p_a, cap_a, moves_b, p_b, cap_b, a_on, b_on, tadv, seed):
        """-> (rc, pick_a, pick_b, draws) from the reference's JointActionCandidate + std::discrete_distribution."""
        ma, mb = np.ascontiguousarray(moves_a, np.uint32), np.ascontiguousarray(moves_b, np.uint32)
        pa, pb = np.ascontiguousarray(p_a, np.float32), np.ascontiguousarray(p_b, np.float32)
        ca, cb = np.ascontiguousarray(cap_a, np.uint8), np.ascontiguousarray(cap_b, np.uint8)
        ia, ib, dr = C.c_uint32(), C.c_uint32(), C.c_int()
        rc = self.lib.hmref_sample_joint(_ptr(ma), _ptr(pa), _ptr(ca), ma.size, _ptr(mb), _ptr(pb), _ptr(cb), mb.size,
                                         int(a_on), int(b_on), int(tadv), seed, C.byref(ia), C.byref(ib), C.byref(dr))
        return int(rc), int(ia.value), int(ib.value), int(dr.value)

    def dirichlet_noise(self, key: int, alpha: float, n: int) -> np.ndarray:
        out = np.zeros(n, np.float32)
        self.lib.hmref_dirichlet_noise(key, alpha, n, _ptr(out))
        return out

    def dirichlet_mix(self, p, noise, epsilon: float) -> np.ndarray:
        out = np.array(p, np.float32, copy=True)
        nz = np.ascontiguousarray(noise, np.float32)
        self.lib.hmref_dirichlet_mix(_ptr(out), _ptr(nz), out.size, epsilon)
        return out

    # -- boards -------------------------------------------------------------------
    def board(self):
        return RefBoard(self, self.lib.hmref_board_new())

    def gen_positions(self, seed: int, n: int):
        arr = (C.c_void_p * n)()
        self.lib.hmref_gen_positions(seed, n, arr)
        return [RefBoard(self, arr[i]) for i in range(n)]

    def handles(self, boards):
        arr = (C.c_void_p * len(boards))()
        for i, b in enumerate(boards):
            arr[i] = b.h
        return arr

    # -- free functions -----------------------------------------------------------
    def normalize_logits(self, logits):
        l = np.ascontiguousarray(logits, np.float32)
        out = np.empty_like(l)
        self.lib.hmref_normalize_logits(_ptr(l), l.size, _ptr(out))
        return out

    def policy_index(self, move, stm_black):
        return self.lib.hmref_policy_index(int(move), int(stm_black))

    def policy_index_of_label(self, label: str):
        return self.lib.hmref_policy_index_of_label(label.encode())

    def policy_label(self, idx):
        buf = C.create_string_buffer(16)
        self.lib.hmref_policy_label(idx, buf, 16)
        return buf.value.decode()

    def policy_tables(self):
        normal = np.empty((2, 64, 64, 2), np.int32)
        drop = np.empty((2, 64, 8), np.int32)
        self.lib.hmref_policy_tables(_ptr(normal), _ptr(drop))
        return normal, drop


class RefBoard:
    def __init__(self, ref: Ref, h):
        self.ref, self.L, self.h = ref, ref.lib, h

    def copy(self):
        return RefBoard(self.ref, self.L.hmref_board_copy(self.h))

    def set(self, fen_pair: str):
        assert self.L.hmref_board_set(self.h, fen_pair.encode())

    def set_fen(self, bn: int, fen: str):
        assert self.L.hmref_board_set_fen(self.h, bn, fen.encode())

    def fen(self, bn):
        buf = C.create_string_buffer(256)
        self.L.hmref_board_fen(self.h, bn, buf, 256)
        return buf.value.decode()

    def push_uci(self, bn: int, uci: str):
        assert self.L.hmref_board_push_uci(self.h, bn, uci.encode()), uci

    def make_moves_uci(self, a: str, b: str):
        ma, mb = C.c_uint32(), C.c_uint32()
        assert self.L.hmref_board_make_moves_uci(self.h, a.encode(), b.encode(), C.byref(ma), C.byref(mb))
        return ma.value, mb.value

    def unmake_moves(self, ma, mb):
        self.L.hmref_board_unmake_moves(self.h, ma, mb)

    def stm_black(self, bn):
        return self.L.hmref_board_stm(self.h, bn)

    def gives_check(self, bn, move):
        return bool(self.L.hmref_board_gives_check(self.h, bn, move))

    def planes(self, team_black: int, tadv: int, dtype="f32"):
        npdt = {"f32": np.float32, "f16": np.uint16, "u8": np.uint8}[dtype]
        out = np.empty((74, 8, 8), npdt)
        getattr(self.L, f"hmref_board_planes_{dtype}")(self.h, int(team_black), int(tadv), _ptr(out))
        return out

    def pack(self, team_black: int, tadv: int):
        out = np.zeros(1, POS_DTYPE)
        self.L.hmref_board_pack(self.h, int(team_black), int(tadv), _ptr(out))
        return out

    def legal(self, bn):
        buf = np.empty(1024, np.uint32)
        k = self.L.hmref_board_legal(self.h, bn, _ptr(buf), 1024)
        return buf[:k].copy()

    def product_pack(self, team_black: int, tadv: int):
        """hm::pack_position of include/hm_batch.hpp on this reference Board."""
        out = np.zeros(1, POS_DTYPE)
        self.L.hmref_product_pack(self.h, int(team_black), int(tadv), _ptr(out))
        return out

    def product_actions(self, bn, team_black):
        """hm::append_actions of include/hm_batch.hpp -> (moves, capture flags)."""
        m, c = np.empty(1024, np.uint32), np.empty(1024, np.uint8)
        k = self.L.hmref_product_actions(self.h, bn, int(team_black), _ptr(m), _ptr(c), 1024)
        return m[:k].copy(), c[:k].copy()

    def is_capture(self, bn, move) -> int:
        return int(self.L.hmref_board_is_capture(self.h, bn, int(move)))

    def search_actions(self, bn, team_black):
        buf = np.empty(1024, np.uint32)
        k = self.L.hmref_board_search_actions(self.h, bn, int(team_black), _ptr(buf), 1024)
        return buf[:k].copy()

    def move_uci(self, bn, move):
        buf = C.create_string_buffer(16)
        self.L.hmref_board_move_uci(self.h, bn, int(move), buf, 16)
        return buf.value.decode()

    def priors_f16(self, bn, policy_u16, moves):
        p = np.ascontiguousarray(policy_u16, np.uint16)
        m = np.ascontiguousarray(moves, np.uint32)
        out = np.empty(m.size, np.float32)
        self.L.hmref_board_priors_f16(self.h, bn, _ptr(p), _ptr(m), m.size, _ptr(out))
        return out

    def priors_f32(self, bn, policy, moves):
        p = np.ascontiguousarray(policy, np.float32)
        m = np.ascontiguousarray(moves, np.uint32)
        out = np.empty(m.size, np.float32)
        self.L.hmref_board_priors_f32(self.h, bn, _ptr(p), _ptr(m), m.size, _ptr(out))
        return out


def gen_config1(ref: Ref, seed: int, n: int):
    """SURVEY.md §8d config-1 generator: returns (boards, packed[n], moves, offsets[2n+1])."""
    boards = ref.gen_positions(seed, n)
    packed = np.zeros(n, POS_DTYPE)
    moves, offsets = [], [0]
    for g, b in enumerate(boards):
        team, tadv = g & 1, (g >> 1) & 1
        packed[g] = b.pack(team, tadv)[0]
        for bn in range(2):
            a = b.search_actions(bn, team)
            moves.append(a)
            offsets.append(offsets[-1] + a.size)
    return boards, packed, np.concatenate(moves).astype(np.uint32), np.array(offsets, np.uint32)
