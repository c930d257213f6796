"""ctypes face of oracle/liboracle.so (edge_oracle.c). TEST INFRASTRUCTURE ONLY: importable from tests/,
__graft_entry__.smoke() and bench.py's cpu_baseline / reference legs, never from haploclique_b200/."""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = os.path.join(_HERE, "liboracle.so")

ST_REJECT_COV, ST_REJECT_LAYOUT, ST_WALKED, ST_UNDEFINED = 0, 1, 2, 3


class _Seg(C.Structure):
    _fields_ = [("start", C.c_int32), ("end", C.c_int32), ("ncig", C.c_int32), ("nq", C.c_int32),
                ("cig", C.c_void_p), ("seq", C.c_void_p), ("qual", C.c_void_p)]


class _Read(C.Structure):
    _fields_ = [("single_end", C.c_int32), ("is_clique", C.c_int32), ("ncov", C.c_int32),
                ("cov_first", C.c_int32), ("cov_last", C.c_int32), ("s", _Seg * 2)]


class _Params(C.Structure):
    _fields_ = [("cutoff_cliques", C.c_double), ("cutoff_mixed", C.c_double), ("cutoff_single", C.c_double),
                ("overlap_cliques", C.c_double), ("overlap_single", C.c_double), ("no_prob0", C.c_int32),
                ("simpson_len", C.c_int32), ("simpson", C.c_void_p)]


RESULT_DTYPE = np.dtype([("status", np.int32), ("edge", np.int32), ("gap", np.int32), ("cc", np.int32),
                         ("tc", np.int32), ("layout", np.int32), ("probM", np.float64), ("prob0", np.float64),
                         ("score", np.float64)], align=True)


class _QuasiParams(C.Structure):
    _fields_ = [("cutoff_cliques", C.c_double), ("cutoff_mixed", C.c_double), ("cutoff_single", C.c_double),
                ("overlap_cliques", C.c_double), ("overlap_single", C.c_double), ("frameshift_merge", C.c_int32),
                ("simpson_len", C.c_int32), ("simpson", C.c_void_p)]


QUASI_RESULT_DTYPE = np.dtype([("status", np.int32), ("edge", np.int32), ("n_single", np.int32), ("subnormal", np.int32),
                               ("q", np.float64), ("cutoff", np.float64)], align=True)


def build(force: bool = False) -> str:
    """Compiles the C restatement (gcc). Building the checker is not using it."""
    srcs = [os.path.join(_HERE, f) for f in ("edge_oracle.c", "quasi_oracle.c", "merge_oracle.c", "orc_types.h")]
    if force or not os.path.exists(_LIB) or os.path.getmtime(_LIB) < max(os.path.getmtime(f) for f in srcs):
        subprocess.check_call(["make", "-s", "-C", _HERE, "oracle"])
    return _LIB


_lib = None


def lib():
    global _lib
    if _lib is None:
        build()
        _lib = C.CDLL(_LIB)
        _lib.orc_edge_all.restype = C.c_int64
        assert _lib.orc_sizeof_read() == C.sizeof(_Read)
        assert _lib.orc_sizeof_result() == RESULT_DTYPE.itemsize
        assert _lib.qorc_sizeof_result() == QUASI_RESULT_DTYPE.itemsize
    return _lib


def tables():
    """The two 127x127 match-probability tables (NewEdgeCalculator.cpp:51-74)."""
    eq = np.zeros((127, 127), np.float64)
    un = np.zeros((127, 127), np.float64)
    lib().orc_tables(eq.ctypes.data_as(C.c_void_p), un.ctypes.data_as(C.c_void_p))
    return eq, un


class Oracle:
    """Oracle(batch, **NewEdgeCalculator params). Parameter names follow the reference constructor
    (NewEdgeCalculator.h:58)."""

    def __init__(self, batch, edge_quasi_cutoff=0.99, overlap=0.9, edge_quasi_cutoff_single=0.95,
                 overlap_single=0.6, edge_quasi_cutoff_mixed=0.97, maxPosition=None, noProb0=False,
                 simpson_map=None):
        self.batch = batch
        n = batch.n
        self._reads = (_Read * max(n, 1))()
        cig, seq, qual = batch.cig.ctypes.data, batch.seq.ctypes.data, batch.qual.ctypes.data
        for r in range(n):
            rd = self._reads[r]
            rd.single_end = int(batch.single_end[r])
            rd.is_clique = int(batch.is_clique[r])
            rd.ncov = int(batch.ncov[r])
            rd.cov_first = int(batch.cov_first[r])
            rd.cov_last = int(batch.cov_last[r])
            for s in range(2):
                k = 2 * r + s
                sg = rd.s[s]
                sg.start, sg.end = int(batch.seg_start[k]), int(batch.seg_end[k])
                sg.ncig, sg.nq = int(batch.cig_len[k]), int(batch.seq_len[k])
                sg.cig = cig + int(batch.cig_off[k])
                sg.seq = seq + int(batch.seq_off[k])
                sg.qual = qual + int(batch.seq_off[k])
        if maxPosition is None:
            maxPosition = batch.max_position()
        self._simpson = np.zeros(int(maxPosition) + 1, np.float64)
        for k, v in (simpson_map or {}).items():
            self._simpson[int(k)] = v
        self._par = _Params(edge_quasi_cutoff, edge_quasi_cutoff_mixed, edge_quasi_cutoff_single, overlap,
                            overlap_single, 1 if noProb0 else 0, self._simpson.shape[0],
                            self._simpson.ctypes.data)

    def pairs(self, pi, pj) -> np.ndarray:
        """Scores ordered pairs (ap1 = pi[k] the later arrival, ap2 = pj[k])."""
        pi = np.ascontiguousarray(pi, np.int32)
        pj = np.ascontiguousarray(pj, np.int32)
        out = np.zeros(pi.shape[0], RESULT_DTYPE)
        lib().orc_edge_pairs(self._reads, pi.ctypes.data_as(C.c_void_p), pj.ctypes.data_as(C.c_void_p),
                             C.c_int64(pi.shape[0]), C.byref(self._par), out.ctypes.data_as(C.c_void_p))
        return out

    def edge(self, i: int, j: int) -> bool:
        return bool(self.pairs([i], [j])["edge"][0])

    def all_pairs(self, detail: bool = True):
        """Every ordered pair i>j, row-major over i (CLEVER's order, CLEVER.cpp:218-232)."""
        n = self.batch.n
        npairs = n * (n - 1) // 2
        if detail:
            out = np.zeros(npairs, RESULT_DTYPE)
            edges = lib().orc_edge_all(self._reads, n, C.byref(self._par), out.ctypes.data_as(C.c_void_p), None)
            return int(edges), out
        e = np.zeros(npairs, np.uint8)
        edges = lib().orc_edge_all(self._reads, n, C.byref(self._par), None, e.ctypes.data_as(C.c_void_p))
        return int(edges), e


class QuasiOracle(Oracle):
    """QuasiOracle(batch, **QuasispeciesEdgeCalculator params): quasi_oracle.c, the restatement of the reference's older
    scorer (QuasispeciesEdgeCalculator.cpp:55-426). Parameter names follow its constructor (QuasispeciesEdgeCalculator.h:57)."""

    def __init__(self, batch, edge_quasi_cutoff=0.99, overlap=0.9, frameshift_merge=False, simpson_map=None,
                 edge_quasi_cutoff_single=0.95, overlap_single=0.6, edge_quasi_cutoff_mixed=0.97, maxPosition=None):
        super().__init__(batch, maxPosition=maxPosition, simpson_map=simpson_map)
        self._qpar = _QuasiParams(edge_quasi_cutoff, edge_quasi_cutoff_mixed, edge_quasi_cutoff_single, overlap,
                                  overlap_single, 1 if frameshift_merge else 0, self._simpson.shape[0],
                                  self._simpson.ctypes.data)

    def pairs(self, pi, pj) -> np.ndarray:
        pi = np.ascontiguousarray(pi, np.int32)
        pj = np.ascontiguousarray(pj, np.int32)
        out = np.zeros(pi.shape[0], QUASI_RESULT_DTYPE)
        lib().qorc_edge_pairs(self._reads, pi.ctypes.data_as(C.c_void_p), pj.ctypes.data_as(C.c_void_p),
                              C.c_int64(pi.shape[0]), C.byref(self._qpar), out.ctypes.data_as(C.c_void_p))
        return out


def tri_pairs(n: int):
    """(i, j) index arrays of all ordered pairs i>j in the row-major order all_pairs uses."""
    i, j = np.tril_indices(n, -1)
    return i.astype(np.int32), j.astype(np.int32)


# ---- single-end super-read merge (oracle/merge_oracle.c) -----------------------------------
def compute_entry(b1: int, q1: int, b2: int, q2: int):
    """computeEntry (src/AlignmentRecord.cpp:169-184): (base char, quality char) of one merged column."""
    out = (C.c_uint8 * 2)()
    lib().orc_compute_entry(C.c_uint8(b1), C.c_uint8(q1), C.c_uint8(b2), C.c_uint8(q2), out)
    return out[0], out[1]


def merge_clique_se(batch, members):
    """AlignmentRecord(alignments, id) for single-end members (read indices into `batch`, fold order).
    Returns (status, start1, end1, cigar_unrolled bytes, sequence bytes, quality bytes)."""
    mem = np.ascontiguousarray(members, np.int32)
    cap = int(batch.cig_len[2 * mem].sum()) + 1
    oc, os_, oq = (np.zeros(cap, np.uint8) for _ in range(3))
    info = np.zeros(5, np.int32)
    p = lambda a: a.ctypes.data_as(C.c_void_p)
    lib().orc_merge_clique_se(p(batch.seg_start), p(batch.seg_end), p(batch.cig_off), p(batch.cig_len), p(batch.seq_off),
                              p(batch.seq_len), p(batch.cig), p(batch.seq), p(batch.qual), p(mem), C.c_int32(mem.size),
                              p(oc), p(os_), p(oq), p(info))
    st, s, e, no, ns = (int(x) for x in info)
    return st, s, e, oc[:no].tobytes(), os_[:ns].tobytes(), oq[:ns].tobytes()


def merge_clique(batch, members):
    """AlignmentRecord(alignments, id) for any mix of single-end / paired-end members (read indices into `batch`, fold
    order). Returns (status, single_end, seg1, seg2) with seg = (start, end, cigar_unrolled, sequence, qualities) and
    seg2 = None for a single-end result."""
    mem = np.ascontiguousarray(members, np.int32)
    rows = np.stack([2 * mem, 2 * mem + 1], 1).reshape(-1)
    cap = int(batch.cig_len[rows].sum()) + 16
    oc, os_, oq = (np.zeros(2 * cap, np.uint8) for _ in range(3))
    info = np.zeros(10, np.int32)
    p = lambda a: np.ascontiguousarray(a).ctypes.data_as(C.c_void_p)
    lib().orc_merge_clique(p(batch.single_end), p(batch.seg_start), p(batch.seg_end), p(batch.cig_off), p(batch.cig_len),
                           p(batch.seq_off), p(batch.seq_len), p(batch.cig), p(batch.seq), p(batch.qual), p(mem),
                           C.c_int32(mem.size), C.c_int64(cap), p(oc), p(os_), p(oq), p(info))
    st, single = int(info[0]), bool(info[1])
    if st:
        return st, False, None, None
    segs = []
    for sg in range(1 if single else 2):
        s, e, no, ns = (int(x) for x in info[2 + 4 * sg:6 + 4 * sg])
        segs.append((s, e, oc[sg * cap:sg * cap + no].tobytes(), os_[sg * cap:sg * cap + ns].tobytes(), oq[sg * cap:sg * cap + ns].tobytes()))
    return st, single, segs[0], (None if single else segs[1])
