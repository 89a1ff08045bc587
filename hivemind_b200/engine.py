"""Python mirror of the reference's nn/ ``Engine`` (engine/src/nn/engine.h:43-132) on top of
the C ABI (include/hm_engine.h).  Method names, argument meaning and error behaviour follow the
reference: calls return ``False`` instead of raising, outputs are views of engine-owned pinned
host buffers valid until the worker's next enqueue.
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass
from typing import Optional

import numpy as np

from . import _lib
from ._lib import NB_INPUT_VALUES, NB_POLICY_VALUES, KernelTime, RootPolicy, ValueConfig

BATCH_SIZE = 8          # engine/src/search/search_params.h:26
NUM_SEARCH_THREADS = 4  # engine/src/search/search_params.h:30

PACKED_BOARD = np.dtype(
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
PACKED_POS = np.dtype([("board", PACKED_BOARD, (2,))])
assert PACKED_POS.itemsize == 256


@dataclass
class HalfInferenceOutputs:  # engine.h:45-51
    value: Optional[np.ndarray] = None       # [B] fp16
    policyA: Optional[np.ndarray] = None     # [B,4672] fp16 raw logits
    policyB: Optional[np.ndarray] = None
    wdl: Optional[np.ndarray] = None         # [B,3] fp16 raw logits (loss, draw, win)
    movesLeft: Optional[np.ndarray] = None   # [B] fp16 in [0,1]


def _view(ptr: C.c_void_p, shape, dtype):
    n = int(np.prod(shape))
    buf = (C.c_char * (n * np.dtype(dtype).itemsize)).from_address(ptr.value)
    return np.frombuffer(buf, dtype=dtype, count=n).reshape(shape)


class Engine:
    """``Engine(deviceId, batchSize)`` — engine.h:53."""

    def __init__(self, deviceId: int, batchSize: int = BATCH_SIZE, numWorkers: int = NUM_SEARCH_THREADS):
        self._L = _lib.load()
        self._h = self._L.hm_engine_create(deviceId, batchSize, numWorkers)
        if not self._h:
            raise RuntimeError(f"hm_engine_create failed: {_lib.last_error()}")
        self._batch = batchSize
        self._workers = numWorkers

    def close(self):
        if getattr(self, "_h", None):
            self._L.hm_engine_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---- reference surface ----------------------------------------------------------
    def loadNetwork(self, onnxFile: str, engineFile: str = "") -> bool:  # engine.h:60
        return bool(self._L.hm_engine_load(self._h, onnxFile.encode()))

    def loadNetworkFromMemory(self, blob: bytes) -> bool:
        buf = (C.c_char * len(blob)).from_buffer_copy(blob)
        return bool(self._L.hm_engine_load_memory(self._h, C.cast(buf, C.c_void_p), len(blob)))

    def getBatchSize(self) -> int:  # engine.h:81
        return self._L.hm_engine_batch_size(self._h)

    def enqueueInferenceHalf(self, obs: np.ndarray, workerIndex: int = 0) -> bool:  # engine.h:74
        if obs.dtype not in (np.float16, np.uint16) or obs.size != self._batch * NB_INPUT_VALUES or not obs.flags.c_contiguous:
            return False
        self._keep = obs  # "must remain unchanged until synchronizeInferenceHalf returns"
        return bool(self._L.hm_engine_enqueue_half(self._h, obs.ctypes.data, workerIndex))

    def synchronizeInferenceHalf(self, outputs: HalfInferenceOutputs, workerIndex: int = 0) -> bool:  # engine.h:75
        p = [C.c_void_p() for _ in range(5)]
        ok = self._L.hm_engine_sync_half(self._h, workerIndex, *[C.byref(x) for x in p])
        if not ok:
            return False
        B = self._batch
        outputs.value = _view(p[0], (B,), np.float16)
        outputs.policyA = _view(p[1], (B, NB_POLICY_VALUES), np.float16)
        outputs.policyB = _view(p[2], (B, NB_POLICY_VALUES), np.float16)
        outputs.wdl = _view(p[3], (B, 3), np.float16)
        outputs.movesLeft = _view(p[4], (B,), np.float16)
        return True

    def runInferenceHalf(self, obs: np.ndarray, outputs: HalfInferenceOutputs, workerIndex: int = 0) -> bool:  # engine.h:69
        return self.enqueueInferenceHalf(obs, workerIndex) and self.synchronizeInferenceHalf(outputs, workerIndex)

    def runInference(self, obs, value, piA, piB, wdl, movesLeft, workerIndex: int = 0) -> bool:  # engine.h:66
        arrs = [obs, value, piA, piB, wdl, movesLeft]
        if any(a.dtype != np.float32 or not a.flags.c_contiguous for a in arrs):
            return False
        return bool(self._L.hm_engine_run_float(self._h, *[a.ctypes.data for a in arrs], workerIndex))

    # ---- fused B200 path --------------------------------------------------------------
    def enqueuePositions(self, pos: np.ndarray, moves: np.ndarray, offsets: np.ndarray, workerIndex: int = 0) -> bool:
        pos = np.ascontiguousarray(pos)
        moves = np.ascontiguousarray(moves, np.uint32)
        offsets = np.ascontiguousarray(offsets, np.uint32)
        if pos.dtype.itemsize != 256 or offsets.size != 2 * pos.shape[0] + 1:
            return False
        self._pending_n = (pos.shape[0], int(offsets[-1]))
        return bool(
            self._L.hm_engine_enqueue_positions(
                self._h, pos.ctypes.data, pos.shape[0], moves.ctypes.data, offsets.ctypes.data, workerIndex
            )
        )

    def synchronizePriors(self, workerIndex: int = 0):
        """-> (priors[n_moves] f32, values[n_pos] f32, raw[n_pos,5] f32) or None."""
        p = [C.c_void_p() for _ in range(3)]
        if not self._L.hm_engine_sync_priors(self._h, workerIndex, *[C.byref(x) for x in p]):
            return None
        n_pos, n_moves = self._pending_n
        priors = _view(p[0], (n_moves,), np.float32) if n_moves else np.empty(0, np.float32)
        return priors, _view(p[1], (n_pos,), np.float32), _view(p[2], (n_pos, 5), np.float32)

    def evaluatePositions(self, pos, moves, offsets, workerIndex: int = 0):
        if not self.enqueuePositions(pos, moves, offsets, workerIndex):
            return None
        return self.synchronizePriors(workerIndex)

    # ---- root policy on the device (include/hm_engine.h "root policy"): temperature
    # (selfplay.cc:167-185), root Dirichlet noise (node.h:288-315), joint-action sampling
    # (selfplay.cc:301-376) ----------------------------------------------------------------
    def _root_policy(self, n_pos, n_moves, temperature, dirichlet_alpha, dirichlet_epsilon, noise_seed,
                     position_hash, captures, uniforms):
        keep = []

        def arr(a, dtype, size):
            if a is None:
                return None
            a = np.ascontiguousarray(a, dtype)
            if a.size != size:
                raise ValueError(f"expected {size} elements, got {a.size}")
            keep.append(a)
            return a.ctypes.data

        rp = RootPolicy(float(temperature), float(dirichlet_alpha), float(dirichlet_epsilon), int(noise_seed),
                        arr(position_hash, np.uint64, n_pos), arr(captures, np.uint8, n_moves),
                        arr(uniforms, np.float64, 3 * n_pos))
        return rp, keep

    def rootPolicyApply(self, pos, moves, offsets, priors, temperature=0.0, dirichlet_alpha=0.0, dirichlet_epsilon=0.0,
                        noise_seed=0, position_hash=None, captures=None, uniforms=None):
        """Stand-alone root stages on host priors -> (priors, noise, picks[n,2], status[n]) or None."""
        pos = np.ascontiguousarray(pos)
        moves = np.ascontiguousarray(moves, np.uint32)
        offsets = np.ascontiguousarray(offsets, np.uint32)
        n = pos.shape[0]
        if pos.dtype.itemsize != 256 or offsets.size != 2 * n + 1:
            return None
        out = np.array(priors, np.float32, copy=True)
        n_moves = int(offsets[-1])
        if out.size != n_moves:
            return None
        rp, keep = self._root_policy(n, n_moves, temperature, dirichlet_alpha, dirichlet_epsilon, noise_seed,
                                     position_hash, captures, uniforms)
        noise = np.zeros(n_moves, np.float32)
        picks = np.zeros((n, 2), np.uint32)
        status = np.zeros(n, np.int32)
        ok = self._L.hm_root_policy_apply(self._h, pos.ctypes.data, n, moves.ctypes.data, offsets.ctypes.data,
                                          out.ctypes.data, C.byref(rp), noise.ctypes.data, picks.ctypes.data,
                                          status.ctypes.data)
        del keep
        return (out, noise, picks, status) if ok else None

    def enqueuePositionsRoot(self, pos, moves, offsets, workerIndex: int = 0, **root) -> bool:
        pos = np.ascontiguousarray(pos)
        moves = np.ascontiguousarray(moves, np.uint32)
        offsets = np.ascontiguousarray(offsets, np.uint32)
        if pos.dtype.itemsize != 256 or offsets.size != 2 * pos.shape[0] + 1:
            return False
        self._pending_n = (pos.shape[0], int(offsets[-1]))
        rp, keep = self._root_policy(pos.shape[0], int(offsets[-1]), root.get("temperature", 0.0),
                                     root.get("dirichlet_alpha", 0.0), root.get("dirichlet_epsilon", 0.0),
                                     root.get("noise_seed", 0), root.get("position_hash"), root.get("captures"),
                                     root.get("uniforms"))
        ok = bool(self._L.hm_engine_enqueue_positions_root(self._h, pos.ctypes.data, pos.shape[0], moves.ctypes.data,
                                                           offsets.ctypes.data, C.byref(rp), workerIndex))
        del keep  # the ABI copies every input into engine-owned pinned memory before returning
        return ok

    def synchronizeRoot(self, workerIndex: int = 0):
        """-> (priors, values, raw, picks[n_pos,2] u32, status[n_pos] i32) or None."""
        p = [C.c_void_p() for _ in range(5)]
        if not self._L.hm_engine_sync_root(self._h, workerIndex, *[C.byref(x) for x in p]):
            return None
        n_pos, n_moves = self._pending_n
        priors = _view(p[0], (n_moves,), np.float32) if n_moves else np.empty(0, np.float32)
        return (priors, _view(p[1], (n_pos,), np.float32), _view(p[2], (n_pos, 5), np.float32),
                _view(p[3], (n_pos, 2), np.uint32), _view(p[4], (n_pos,), np.int32))

    def setValueConfig(self, enable_wdl_eval=True, wdl_value_weight=0.25, moves_left_discount=0.005, draw_contempt=0.0):
        cfg = ValueConfig(int(enable_wdl_eval), wdl_value_weight, moves_left_discount, draw_contempt)
        return bool(self._L.hm_engine_set_value_config(self._h, C.byref(cfg)))

    # ---- stand-alone stages -------------------------------------------------------------
    def encodePlanes(self, pos: np.ndarray, dtype: str = "f32") -> np.ndarray:
        code, npdt = {"f32": (0, np.float32), "f16": (1, np.uint16), "bf16": (2, np.uint16), "u8": (3, np.uint8)}[dtype]
        pos = np.ascontiguousarray(pos)
        out = np.empty((pos.shape[0], 74, 8, 8), npdt)
        if not self._L.hm_encode_planes(self._h, pos.ctypes.data, pos.shape[0], out.ctypes.data, code):
            raise RuntimeError(_lib.last_error())
        return out

    def policyPriorsHalf(self, logits_a, logits_b, pos, moves, offsets) -> np.ndarray:
        la = np.ascontiguousarray(logits_a).view(np.uint16)
        lb = np.ascontiguousarray(logits_b).view(np.uint16)
        pos = np.ascontiguousarray(pos)
        moves = np.ascontiguousarray(moves, np.uint32)
        offsets = np.ascontiguousarray(offsets, np.uint32)
        out = np.zeros(moves.size, np.float32)
        if not self._L.hm_policy_priors_half(
            self._h, la.ctypes.data, lb.ctypes.data, pos.ctypes.data, pos.shape[0], moves.ctypes.data,
            offsets.ctypes.data, out.ctypes.data,
        ):
            raise RuntimeError(_lib.last_error())
        return out

    def policyIndices(self, moves, stm_black) -> np.ndarray:
        moves = np.ascontiguousarray(moves, np.uint32)
        stm = np.ascontiguousarray(stm_black, np.uint8)
        out = np.empty(moves.size, np.int32)
        if not self._L.hm_policy_indices(self._h, moves.ctypes.data, stm.ctypes.data, moves.size, out.ctypes.data):
            raise RuntimeError(_lib.last_error())
        return out

    def shapeValuesHalf(self, value, wdl, moves_left) -> np.ndarray:
        v = np.ascontiguousarray(value).view(np.uint16)
        w = np.ascontiguousarray(wdl).view(np.uint16)
        m = np.ascontiguousarray(moves_left).view(np.uint16)
        out = np.empty(v.size, np.float32)
        if not self._L.hm_shape_values_half(self._h, v.ctypes.data, w.ctypes.data, m.ctypes.data, v.size, out.ctypes.data):
            raise RuntimeError(_lib.last_error())
        return out

    # ---- measurement / debug ------------------------------------------------------------
    def benchResident(self, iters: int, workerIndex: int = 0):
        ms, n = C.c_float(), C.c_int()
        if not self._L.hm_engine_bench_resident(self._h, workerIndex, iters, C.byref(ms), C.byref(n)):
            raise RuntimeError(_lib.last_error())
        return ms.value, n.value

    def benchE2E(self, pos, moves, offsets, iters: int, slots: int = 2):
        """The enqueuePositions / synchronizePriors loop driven from C (as the reference's C++ caller would):
        -> (wall ms for all iters incl. H2D/D2H, checksum of the last request's values)."""
        pos = np.ascontiguousarray(pos)
        moves = np.ascontiguousarray(moves, np.uint32)
        offsets = np.ascontiguousarray(offsets, np.uint32)
        ms, chk = C.c_double(), C.c_double()
        if not self._L.hm_engine_bench_e2e(self._h, pos.ctypes.data, pos.shape[0], moves.ctypes.data, offsets.ctypes.data,
                                           iters, slots, C.byref(ms), C.byref(chk)):
            raise RuntimeError(_lib.last_error())
        return ms.value, chk.value

    def benchPlanes(self, pos: np.ndarray, dtype: str = "f16", iters: int = 20) -> float:
        """Device-resident stand-alone plane encode of len(pos) positions -> ms per launch."""
        pos = np.ascontiguousarray(pos)
        ms = C.c_float()
        code = {"f32": _lib.DTYPE_F32, "f16": _lib.DTYPE_F16, "bf16": _lib.DTYPE_BF16, "u8": _lib.DTYPE_U8}[dtype]
        if not self._L.hm_bench_planes(self._h, pos.ctypes.data, pos.shape[0], code, iters, C.byref(ms)):
            raise RuntimeError(_lib.last_error())
        return ms.value / iters

    def benchPriors(self, pos, moves, offsets, iters: int = 20) -> float:
        """Device-resident stand-alone legal-move gather + softmax -> ms per launch."""
        pos = np.ascontiguousarray(pos)
        moves = np.ascontiguousarray(moves, np.uint32)
        offsets = np.ascontiguousarray(offsets, np.uint32)
        ms = C.c_float()
        if not self._L.hm_bench_priors(self._h, pos.ctypes.data, pos.shape[0], moves.ctypes.data, offsets.ctypes.data,
                                       iters, C.byref(ms)):
            raise RuntimeError(_lib.last_error())
        return ms.value / iters

    def benchResidentMulti(self, iters: int, numWorkers: int):
        """`iters` fused passes issued round-robin on `numWorkers` worker streams; device ms."""
        ms, n = C.c_float(), C.c_int()
        if not self._L.hm_engine_bench_resident_multi(self._h, numWorkers, iters, C.byref(ms), C.byref(n)):
            raise RuntimeError(_lib.last_error())
        return ms.value, n.value

    def traceTrunk(self, run: int, workerIndex: int = 0):
        """Timeline of trunk run `run` (CTA 0): dict role -> list of (tag, clock)."""
        buf = np.zeros(3 * 2048, np.uint64)
        cap = C.c_int()
        if not self._L.hm_engine_trace_trunk(self._h, workerIndex, run, buf.ctypes.data, C.byref(cap)):
            raise RuntimeError(_lib.last_error())
        out = {}
        for role in range(3):
            ev = buf[role * cap.value:(role + 1) * cap.value]
            ev = ev[ev != 0]
            out[role] = [(int(e >> np.uint64(48)), int(e & np.uint64(0xFFFFFFFFFFFF))) for e in ev]
        return out

    def profilePass(self, workerIndex: int = 0):
        arr = (KernelTime * 128)()
        n = C.c_int()
        if not self._L.hm_engine_profile_pass(self._h, workerIndex, arr, 128, C.byref(n)):
            raise RuntimeError(_lib.last_error())
        return [(arr[i].name.decode(), arr[i].ms, arr[i].flops, arr[i].bytes) for i in range(n.value)]

    def debugFetch(self, name: str, workerIndex: int = 0) -> np.ndarray:
        n = C.c_size_t()
        probe = np.empty(1, np.float32)
        self._L.hm_engine_debug_fetch(self._h, workerIndex, name.encode(), probe.ctypes.data, 0, C.byref(n))
        out = np.empty(n.value, np.float32)
        if not self._L.hm_engine_debug_fetch(self._h, workerIndex, name.encode(), out.ctypes.data, out.size, C.byref(n)):
            raise RuntimeError(_lib.last_error())
        return out
