"""HVM4 self-play chunk writer — Python mirror of the reference's `ChunkWriter`
(engine/src/tools/selfplay.cc:69-158) over the C ABI in include/hm_chunk.h.

    writer = ChunkWriter(directory, samplesPerChunk, runId)
    writer.append([TrainingSample(...), ...])      # publishes chunk_<runId>_%06d.hvm every samplesPerChunk
    writer.finish()                                 # publishes the remainder

`TrainingSample` carries the fields of selfplay.cc:40-53; `planes` are the 4736 u8 values produced on the
device by `Engine.encodePlanes(..., dtype="u8")` (selfplay.cc:464-476 quantisation), policies are sparse
`(index, probability)` arrays, e.g. from `marginal_policy` (selfplay.cc:400-426).
The files are read back by the reference trainer: src/preprocessing/convert_selfplay_data.py:69-146.
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass, field
from typing import Iterable, Sequence, Tuple

import numpy as np

from . import _lib

NB_INPUT_VALUES = 4736
NB_POLICY_VALUES = 4672


class _CSample(C.Structure):
    _fields_ = [("game_id", C.c_uint64), ("nodes", C.c_uint32), ("macro_ply", C.c_uint16),
                ("moves_left", C.c_uint16), ("team", C.c_uint8), ("has_time_advantage", C.c_uint8),
                ("outcome", C.c_int8), ("wdl", C.c_uint8), ("root_q", C.c_float)]


def _empty_policy() -> Tuple[np.ndarray, np.ndarray]:
    return np.zeros(0, np.uint16), np.zeros(0, np.float32)


@dataclass
class TrainingSample:
    """selfplay.cc:40-53 (same names and defaults)."""
    gameId: int = 0
    nodes: int = 0
    macroPly: int = 0
    movesLeft: int = 0
    team: int = 0
    hasTimeAdvantage: int = 0
    outcome: int = 0
    wdl: int = 1
    rootQ: float = 0.0
    planes: np.ndarray = field(default_factory=lambda: np.zeros(NB_INPUT_VALUES, np.uint8))
    policyA: Tuple[np.ndarray, np.ndarray] = field(default_factory=_empty_policy)
    policyB: Tuple[np.ndarray, np.ndarray] = field(default_factory=_empty_policy)


def _bind(L):
    if getattr(L, "_hm_chunk_bound", False):
        return L
    u16p, f32p, u8p = C.POINTER(C.c_uint16), C.POINTER(C.c_float), C.POINTER(C.c_uint8)
    L.hm_chunk_writer_create.restype = C.c_void_p
    L.hm_chunk_writer_create.argtypes = [C.c_char_p, C.c_size_t, C.c_uint64]
    L.hm_chunk_writer_append.restype = C.c_int
    L.hm_chunk_writer_append.argtypes = [C.c_void_p, C.POINTER(_CSample), u8p, u16p, f32p, C.c_size_t, u16p, f32p, C.c_size_t]
    L.hm_chunk_writer_finish.restype = C.c_int
    L.hm_chunk_writer_finish.argtypes = [C.c_void_p]
    L.hm_chunk_writer_chunks_written.restype = C.c_size_t
    L.hm_chunk_writer_chunks_written.argtypes = [C.c_void_p]
    L.hm_chunk_writer_pending.restype = C.c_size_t
    L.hm_chunk_writer_pending.argtypes = [C.c_void_p]
    L.hm_chunk_writer_destroy.restype = None
    L.hm_chunk_writer_destroy.argtypes = [C.c_void_p]
    L.hm_sparse_policy_from_visits.restype = C.c_size_t
    L.hm_sparse_policy_from_visits.argtypes = [C.POINTER(C.c_int32), C.POINTER(C.c_int64), C.c_size_t, u16p, f32p, C.c_size_t]
    L._hm_chunk_bound = True
    return L


def _policy_arrays(policy) -> Tuple[np.ndarray, np.ndarray]:
    idx, prob = policy
    idx = np.ascontiguousarray(idx, dtype=np.uint16)
    prob = np.ascontiguousarray(prob, dtype=np.float32)
    if idx.shape != prob.shape or idx.ndim != 1:
        raise ValueError("a sparse policy is a pair of equally long 1-D arrays (index, probability)")
    return idx, prob


class ChunkWriter:
    """`ChunkWriter(directory, samplesPerChunk, runId)`, selfplay.cc:71-76.  Failures raise RuntimeError
    (the reference throws std::runtime_error)."""

    def __init__(self, directory, samplesPerChunk: int, runId: int):
        self._L = _bind(_lib.load())
        self._h = self._L.hm_chunk_writer_create(str(directory).encode(), max(0, int(samplesPerChunk)), int(runId))
        if not self._h:
            raise RuntimeError(_lib.last_error())

    def append(self, samples: Iterable[TrainingSample]) -> None:
        u16p, f32p, u8p = C.POINTER(C.c_uint16), C.POINTER(C.c_float), C.POINTER(C.c_uint8)
        for s in samples:
            planes = np.ascontiguousarray(s.planes, dtype=np.uint8).reshape(-1)
            if planes.size != NB_INPUT_VALUES:
                raise ValueError(f"planes must hold {NB_INPUT_VALUES} u8 values")
            ia, pa = _policy_arrays(s.policyA)
            ib, pb = _policy_arrays(s.policyB)
            cs = _CSample(int(s.gameId), int(s.nodes), int(s.macroPly), int(s.movesLeft), int(s.team),
                          int(s.hasTimeAdvantage), int(s.outcome), int(s.wdl), float(s.rootQ))
            ok = self._L.hm_chunk_writer_append(
                self._h, C.byref(cs), planes.ctypes.data_as(u8p),
                ia.ctypes.data_as(u16p), pa.ctypes.data_as(f32p), ia.size,
                ib.ctypes.data_as(u16p), pb.ctypes.data_as(f32p), ib.size)
            if not ok:
                raise RuntimeError(_lib.last_error())

    def finish(self) -> None:
        if not self._L.hm_chunk_writer_finish(self._h):
            raise RuntimeError(_lib.last_error())

    @property
    def chunksWritten(self) -> int:
        return int(self._L.hm_chunk_writer_chunks_written(self._h))

    @property
    def pending(self) -> int:
        return int(self._L.hm_chunk_writer_pending(self._h))

    def close(self) -> None:
        if getattr(self, "_h", None):
            self._L.hm_chunk_writer_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def marginal_policy(policyIndex: Sequence[int], visits: Sequence[int]) -> Tuple[np.ndarray, np.ndarray]:
    """Root-edge visits -> sparse policy of one board, selfplay.cc:400-426 (after its move -> index step)."""
    L = _bind(_lib.load())
    idx = np.ascontiguousarray(policyIndex, dtype=np.int32)
    vis = np.ascontiguousarray(visits, dtype=np.int64)
    if idx.shape != vis.shape:
        raise ValueError("policyIndex and visits must have the same length")
    out_i = np.zeros(max(1, idx.size), np.uint16)
    out_p = np.zeros(max(1, idx.size), np.float32)
    n = L.hm_sparse_policy_from_visits(idx.ctypes.data_as(C.POINTER(C.c_int32)), vis.ctypes.data_as(C.POINTER(C.c_int64)),
                                       idx.size, out_i.ctypes.data_as(C.POINTER(C.c_uint16)),
                                       out_p.ctypes.data_as(C.POINTER(C.c_float)), out_i.size)
    if n == 0:
        raise RuntimeError(_lib.last_error())
    return out_i[:n].copy(), out_p[:n].copy()
