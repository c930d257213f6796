"""Host-side mirror of the reference's edge-calculator plugin interface over libhcgpu.so.

``EdgeCalculator`` / ``NewEdgeCalculator`` keep the reference's names, constructor argument order and
meaning (``src/EdgeCalculator.h:24-37``, ``src/NewEdgeCalculator.h:58``). The scoring itself happens in the
sm_100a kernels behind the C ABI (``include/hc_edge.h``); there is no CPU scoring path here: without the
built library or without a CUDA device every scoring call raises.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Dict, Optional, Sequence, Union

import numpy as np

from .records import AlignmentRecord, ReadBatch

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("HC_LIBHCGPU") or os.path.join(_HERE, "libhcgpu.so")     # HC_LIBHCGPU: an experiment build of the same library

HC_OK = 0
HC_E_INVALID = -1
KERNEL_AUTO, KERNEL_GENERIC, KERNEL_TILED, KERNEL_TILED2, KERNEL_STREAMED = 0, 1, 2, 3, 4
STATIC_MAX_PAIRS = 16384                                  # include/hc_edge.h: HC_STATIC_MAX_PAIRS
ST_REJECT_COV, ST_REJECT_LAYOUT, ST_WALKED, ST_UNDEFINED = 0, 1, 2, 3


class HcError(RuntimeError):
    pass


class _Params(C.Structure):
    _fields_ = [("edge_quasi_cutoff_cliques", C.c_double), ("edge_quasi_cutoff_mixed", C.c_double),
                ("edge_quasi_cutoff_single", C.c_double), ("min_overlap_cliques", C.c_double),
                ("min_overlap_single", C.c_double), ("no_prob0", C.c_int32), ("simpson_len", C.c_int32),
                ("simpson", C.c_void_p), ("device", C.c_int32), ("reserved", C.c_int32),
                ("ambiguity_rel_tol", C.c_double)]


class _Reads(C.Structure):
    _fields_ = [("n", C.c_int32)] + [(k, C.c_void_p) for k in (
        "single_end", "is_clique", "ncov", "cov_first", "cov_last", "seg_start", "seg_end", "cig_off", "cig_len",
        "seq_off", "seq_len", "cig", "seq", "qual")]


class _Packed(C.Structure):
    _fields_ = [("n_reads", C.c_int32), ("n_qual_codes", C.c_int32), ("qual_chars", C.c_uint8 * 128),
                ("hdr", C.c_void_p), ("ops", C.c_void_p), ("n_ops_total", C.c_int64),
                ("max_ops_per_read", C.c_int32), ("reserved", C.c_int32), ("owner", C.c_void_p)]


class _BandStats(C.Structure):
    _fields_ = [("pairs_total", C.c_int64), ("pairs_scored", C.c_int64), ("pairs_walked", C.c_int64),
                ("edges", C.c_int64), ("sum_cc", C.c_int64), ("sum_tc", C.c_int64), ("undefined", C.c_int64),
                ("ambiguous", C.c_int64), ("kernel_ms", C.c_float), ("launches", C.c_int32)]


PAIR_RESULT_DTYPE = np.dtype([("status", np.int32), ("edge", np.int32), ("gap", np.int32), ("cc", np.int32),
                              ("tc", np.int32), ("ambiguous", np.int32), ("probM", np.float64),
                              ("prob0", np.float64), ("score", np.float64)], align=True)
AMBIGUOUS_DTYPE = np.dtype([("i", np.int32), ("j", np.int32), ("cc", np.int32), ("tc", np.int32),
                            ("probM", np.float64), ("prob0", np.float64), ("score", np.float64),
                            ("cutoff", np.float64), ("edge", np.int32), ("edge_device", np.int32)], align=True)
HDR_DTYPE = np.dtype([("raw", np.int32, 4), ("ext", np.int32, 4), ("ops_off", np.uint32, 2), ("n_ops", np.int32, 2),
                      ("ncov_flags", np.uint32), ("cov_first", np.int32), ("cov_last", np.int32),
                      ("seg_flags", np.uint16, 2)], align=True)

_lib = None


def load_library():
    """Loads libhcgpu.so; fails loudly when the CUDA extension has not been built."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise HcError("haploclique_b200/libhcgpu.so is missing: build it with "
                          "`python -c 'import __graft_entry__ as g; g.build()'` (there is no CPU fallback)")
        L = C.CDLL(LIB_PATH)
        L.hc_last_error.restype = C.c_char_p
        L.hc_last_error.argtypes = [C.c_void_p]
        L.hc_host_alloc.restype = C.c_void_p
        L.hc_host_alloc.argtypes = [C.c_int64]
        L.hc_host_free.argtypes = [C.c_void_p]
        L.hc_create.argtypes = [C.c_void_p, C.c_void_p]
        L.hc_destroy.argtypes = [C.c_void_p]
        L.hc_pack.argtypes = [C.c_void_p, C.c_void_p]
        L.hc_packed_free.argtypes = [C.c_void_p]
        L.hc_upload.argtypes = [C.c_void_p, C.c_void_p]
        L.hc_set_params.argtypes = [C.c_void_p, C.c_void_p]
        L.hc_band_layout.argtypes = [C.c_void_p, C.c_int32, C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p]
        L.hc_score_band.argtypes = [C.c_void_p, C.c_int32, C.c_int32, C.c_void_p, C.c_int64, C.c_void_p, C.c_int32,
                                    C.c_void_p]
        L.hc_score_pairs.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p]
        L.hc_score_pairs_static.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p]
        L.hc_score_band_detail.argtypes = [C.c_void_p, C.c_int32, C.c_int32, C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p]
        L.hc_set_kernel.argtypes = [C.c_void_p, C.c_int32]
        L.hc_last_kernel.argtypes = [C.c_void_p]
        L.hc_measure_peaks.argtypes = [C.c_int32, C.c_void_p, C.c_void_p]
        L.hc_get_tables.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
        L.hc_quasi_score_pairs.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p]
        L.hc_quasi_score_band.argtypes = [C.c_void_p, C.c_void_p, C.c_int32, C.c_int32, C.c_void_p, C.c_int64, C.c_void_p]
        L.hc_partition_rows.argtypes = [C.c_void_p, C.c_void_p, C.c_int32, C.c_int32, C.c_void_p, C.c_void_p]
        L.hc_packed_slice.argtypes = [C.c_void_p, C.c_int32, C.c_int32, C.c_void_p]
        L.hc_multi_create.argtypes = [C.c_void_p, C.c_void_p, C.c_int32, C.c_void_p]
        L.hc_multi_destroy.argtypes = [C.c_void_p]
        L.hc_multi_last_error.restype = C.c_char_p
        L.hc_multi_last_error.argtypes = [C.c_void_p]
        L.hc_multi_n_devices.argtypes = [C.c_void_p]
        L.hc_multi_context.restype = C.c_void_p
        L.hc_multi_context.argtypes = [C.c_void_p, C.c_int32]
        L.hc_multi_set_params.argtypes = [C.c_void_p, C.c_void_p]
        L.hc_multi_upload.argtypes = [C.c_void_p, C.c_void_p]
        L.hc_multi_partition.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
        L.hc_multi_band_layout.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        L.hc_multi_score_band.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_int32, C.c_void_p, C.c_void_p]
        assert C.sizeof(_BandStats) == 72 and HDR_DTYPE.itemsize == 64 and AMBIGUOUS_DTYPE.itemsize == 56
        _lib = L
    return _lib


def _ptr(a: Optional[np.ndarray]):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


class PinnedBuffer:
    """Page-locked host array (cudaHostAlloc through the C ABI)."""

    def __init__(self, n: int, dtype=np.uint32):
        self._L = load_library()
        self.nbytes = max(int(n), 1) * np.dtype(dtype).itemsize
        self._p = self._L.hc_host_alloc(self.nbytes)
        if not self._p:
            raise HcError("cudaHostAlloc failed")
        buf = (C.c_uint8 * self.nbytes).from_address(self._p)
        self.array = np.frombuffer(buf, dtype=dtype, count=int(n))

    def __del__(self):
        try:
            if self._p:
                self.array = None
                self._L.hc_host_free(self._p)
                self._p = None
        except Exception:
            pass


class PackedBatch:
    """Result of hc_pack: 64-byte headers + 16-bit op records, owned by the library."""

    def __init__(self, batch: ReadBatch):
        L = load_library()
        self.batch = batch
        rd = _Reads(batch.n, *[_ptr(np.ascontiguousarray(getattr(batch, k))) for k in (
            "single_end", "is_clique", "ncov", "cov_first", "cov_last", "seg_start", "seg_end", "cig_off", "cig_len",
            "seq_off", "seq_len", "cig", "seq", "qual")])
        self._h = C.c_void_p()
        rc = L.hc_pack(C.byref(rd), C.byref(self._h))
        if rc != HC_OK:
            raise HcError("hc_pack failed (%d): %s" % (rc, L.hc_last_error(None).decode()))
        self.c = C.cast(self._h, C.POINTER(_Packed)).contents

    @property
    def n_reads(self) -> int:
        return self.c.n_reads

    @property
    def n_ops_total(self) -> int:
        return self.c.n_ops_total

    def nbytes(self) -> int:
        return self.c.n_reads * 64 + (self.c.n_ops_total + 64) * 2

    def headers(self) -> np.ndarray:
        buf = (C.c_uint8 * (64 * self.c.n_reads)).from_address(self.c.hdr)
        return np.frombuffer(buf, dtype=HDR_DTYPE).copy()

    def ops(self) -> np.ndarray:
        """The op records of all reads back to back (the per-read alignment padding removed)."""
        if self.c.n_ops_total == 0:
            return np.zeros(0, np.uint16)
        buf = (C.c_uint16 * self.c.n_ops_total).from_address(self.c.ops)
        raw = np.frombuffer(buf, dtype=np.uint16)
        h = self.headers()
        off = h["ops_off"][:, 0].astype(np.int64)
        cnt = h["n_ops"].sum(axis=1).astype(np.int64)
        idx = np.repeat(off - (np.cumsum(cnt) - cnt), cnt) + np.arange(int(cnt.sum()), dtype=np.int64)
        return raw[idx].copy()

    def __del__(self):
        try:
            if self._h:
                load_library().hc_packed_free(self._h)
                self._h = None
        except Exception:
            pass


class Adjacency:
    """Banded adjacency of one scored batch: row i holds one bit per earlier read j in [col0[i], i)."""

    def __init__(self, row_begin: int, row_end: int, col0: np.ndarray, word_off: np.ndarray, words: np.ndarray,
                 stats: Dict[str, float], ambiguous: np.ndarray):
        self.row_begin, self.row_end = row_begin, row_end
        self.col0, self.word_off, self.words = col0, word_off, words
        self.stats, self.ambiguous = stats, ambiguous

    def edge(self, i: int, j: int) -> bool:
        """edgeBetween(read i, read j) for i the later arrival; anything left of the band is a non-overlap."""
        if j >= i:
            raise ValueError("adjacency holds ordered pairs with j < i")
        k = i - self.row_begin
        c0 = int(self.col0[k])
        if j < c0:
            return False
        w = int(self.word_off[k]) + ((j - c0) >> 5)
        return bool((int(self.words[w]) >> ((j - c0) & 31)) & 1)

    def row(self, i: int) -> np.ndarray:
        k = i - self.row_begin
        w0, w1 = int(self.word_off[k]), int(self.word_off[k + 1])
        bits = np.unpackbits(self.words[w0:w1].view(np.uint8), bitorder="little")
        return int(self.col0[k]) + np.nonzero(bits)[0]

    def edges(self):
        """(i, j) arrays of all edges, row-major."""
        ii, jj = [], []
        for i in range(self.row_begin, self.row_end):
            r = self.row(i)
            ii.append(np.full(r.shape[0], i, np.int32))
            jj.append(r.astype(np.int32))
        if not ii:
            return np.zeros(0, np.int32), np.zeros(0, np.int32)
        return np.concatenate(ii), np.concatenate(jj)

    def count(self) -> int:
        return int(np.unpackbits(self.words.view(np.uint8)).sum())


class EdgeCalculator:
    """The reference's pure-virtual plugin API (src/EdgeCalculator.h:24-37)."""

    def edgeBetween(self, ar1, ar2) -> bool:
        raise NotImplementedError

    def getPartnerLengthRange(self, ar):
        raise NotImplementedError

    def setOverlapCliques(self, d: float) -> None:
        pass


class NewEdgeCalculator(EdgeCalculator):
    """GPU-backed drop-in for the reference's NewEdgeCalculator. Constructor arguments in the reference's
    order (src/NewEdgeCalculator.h:58); ``device`` is the only addition."""

    def __init__(self, Q: float = 0.9, edge_quasi_cutoff: float = 0.99, overlap: float = 0.9,
                 frameshift_merge: bool = False, simpson_map: Optional[Dict[int, float]] = None,
                 edge_quasi_cutoff_single: float = 0.95, overlap_single: float = 0.6,
                 edge_quasi_cutoff_mixed: float = 0.97, maxPosition: int = 0, noProb0: bool = False,
                 device: int = 0, ambiguity_rel_tol: float = 1e-12, devices: Optional[Sequence[int]] = None):
        self.L = load_library()
        self._multi = None
        self._ctx = None
        self._aux = None
        self.Q, self.FRAMESHIFT_MERGE = Q, frameshift_merge       # stored, unused by the scorer (as in the reference)
        self._simpson = None
        if simpson_map:
            self._simpson = np.zeros(int(maxPosition) + 1, np.float64)   # SIMPSON_MAP.resize(maxPosition+1) (:85)
            for k, v in simpson_map.items():
                self._simpson[int(k)] = v
        self._par = _Params(edge_quasi_cutoff, edge_quasi_cutoff_mixed, edge_quasi_cutoff_single, overlap,
                            overlap_single, int(bool(noProb0)), 0 if self._simpson is None else self._simpson.shape[0],
                            None if self._simpson is None else self._simpson.ctypes.data, device, 0, ambiguity_rel_tol)
        self._ctx = C.c_void_p()
        self.devices = None if devices is None else [int(d) for d in devices]
        if self.devices is not None:
            # several GPUs behind one handle (hc_multi_*): rows partitioned by pair count, halo from the covered intervals
            dv = np.asarray(self.devices, np.int32)
            self._multi = C.c_void_p()
            rc = self.L.hc_multi_create(C.byref(self._par), _ptr(dv), dv.shape[0], C.byref(self._multi))
            if rc != HC_OK:
                self._multi = None
                raise HcError("hc_multi_create failed (%d): %s" % (rc, self.L.hc_last_error(None).decode()))
            self._ctx = C.c_void_p(self.L.hc_multi_context(self._multi, 0))
        else:
            rc = self.L.hc_create(C.byref(self._par), C.byref(self._ctx))
            if rc != HC_OK:
                raise HcError("hc_create failed (%d): %s" % (rc, self.L.hc_last_error(None).decode()))
        self._packed: Optional[PackedBatch] = None
        self._index: Dict[int, int] = {}
        self._adj: Optional[Adjacency] = None

    def __del__(self):
        try:
            if self._aux:
                self.L.hc_destroy(self._aux)
                self._aux = None
            if self._multi:
                self.L.hc_multi_destroy(self._multi)
                self._multi, self._ctx = None, None
            elif self._ctx:
                self.L.hc_destroy(self._ctx)
                self._ctx = None
        except Exception:
            pass

    def _check(self, rc: int, what: str):
        if rc != HC_OK:
            msg = self.L.hc_multi_last_error(self._multi) if (self._multi and what.startswith("hc_multi")) else self.L.hc_last_error(self._ctx)
            raise HcError("%s failed (%d): %s" % (what, rc, msg.decode()))

    def contexts(self):
        """The per-device contexts (one unless `devices` was given)."""
        if self._multi:
            return [C.c_void_p(self.L.hc_multi_context(self._multi, k)) for k in range(len(self.devices))]
        return [self._ctx]

    def partition(self):
        """(cuts, first) of the resident batch over the devices (hc_multi_partition)."""
        cuts = np.zeros(len(self.devices) + 1, np.int32)
        first = np.zeros(len(self.devices), np.int32)
        self._check(self.L.hc_multi_partition(self._multi, _ptr(cuts), _ptr(first)), "hc_multi_partition")
        return cuts, first

    # -- reference interface ---------------------------------------------------------------------------
    def setOverlapCliques(self, d: float) -> None:            # src/NewEdgeCalculator.cpp:1382
        self._par.min_overlap_cliques = d
        if self._multi:
            self._check(self.L.hc_multi_set_params(self._multi, C.byref(self._par)), "hc_multi_set_params")
        else:
            self._check(self.L.hc_set_params(self._ctx, C.byref(self._par)), "hc_set_params")
        self._adj = None

    def getOverlapCliques(self) -> float:
        return self._par.min_overlap_cliques

    def getPartnerLengthRange(self, ar):                       # the reference asserts false here (:1378-1380)
        raise AssertionError("NewEdgeCalculator::getPartnerLengthRange is not implemented in the reference either")

    def edgeBetween(self, ap1: AlignmentRecord, ap2: AlignmentRecord) -> bool:
        """Answers from the adjacency of the last scoreBatch when both records belong to it (the clique
        finders' calls); any other two records are packed as a 2-read batch and scored on the GPU."""
        i, j = self._index.get(id(ap1)), self._index.get(id(ap2))
        if i is not None and j is not None and self._adj is not None and j < i:
            return self._adj.edge(i, j)
        if i is not None and j is not None and self._packed is not None and not self._multi:
            return bool(self.score_pairs([i], [j])["edge"][0])     # e.g. the reversed order: scored on the resident batch
        # records outside the resident batch (the reference's unit tests call it this way): a two-read batch on a small
        # second context, so that the resident batch stays where it is
        if self._aux is None:
            self._aux = C.c_void_p()
            par = _Params.from_buffer_copy(self._par)
            par.device = self._par.device if self.devices is None else self.devices[0]
            rc = self.L.hc_create(C.byref(par), C.byref(self._aux))
            if rc != HC_OK:
                self._aux = None
                raise HcError("hc_create failed (%d): %s" % (rc, self.L.hc_last_error(None).decode()))
        self.L.hc_set_params(self._aux, C.byref(self._par))
        pair = PackedBatch(ReadBatch.from_records([ap2, ap1]))
        rc = self.L.hc_upload(self._aux, pair._h)
        if rc != HC_OK:
            raise HcError("hc_upload failed (%d): %s" % (rc, self.L.hc_last_error(self._aux).decode()))
        return bool(self._score_pair_on(self._aux, 1, 0))

    def _score_pair_on(self, ctx, i: int, j: int) -> int:
        pi, pj = np.array([i], np.int32), np.array([j], np.int32)
        out = np.zeros(1, PAIR_RESULT_DTYPE)
        rc = self.L.hc_score_pairs(ctx, _ptr(pi), _ptr(pj), 1, _ptr(out))
        if rc != HC_OK:
            raise HcError("hc_score_pairs failed (%d): %s" % (rc, self.L.hc_last_error(ctx).decode()))
        return int(out["edge"][0])

    # -- batch interface (the driver hook of SURVEY.md 8(b)) -------------------------------------------
    def upload(self, reads: Union[ReadBatch, Sequence[AlignmentRecord], PackedBatch]) -> PackedBatch:
        if isinstance(reads, PackedBatch):
            packed = reads
        else:
            if not isinstance(reads, ReadBatch):
                recs = list(reads)
                self._index = {id(r): k for k, r in enumerate(recs)}
                reads = ReadBatch.from_records(recs)
            packed = PackedBatch(reads)
        if self._multi:
            self._check(self.L.hc_multi_upload(self._multi, packed._h), "hc_multi_upload")
        else:
            self._check(self.L.hc_upload(self._ctx, packed._h), "hc_upload")
        self._packed = packed
        self._adj = None
        return packed

    def band_layout(self, row_begin: int = 0, row_end: Optional[int] = None):
        row_end = self._packed.n_reads if row_end is None else row_end
        nrows = row_end - row_begin
        col0 = np.zeros(max(nrows, 1), np.int32)
        off = np.zeros(nrows + 1, np.int64)
        nw = C.c_int64(0)
        if self._multi:
            if (row_begin, row_end) != (0, self._packed.n_reads):
                raise HcError("several devices: the band is scored as a whole (rows are cut by the library)")
            self._check(self.L.hc_multi_band_layout(self._multi, _ptr(col0), _ptr(off), C.byref(nw)), "hc_multi_band_layout")
        else:
            self._check(self.L.hc_band_layout(self._ctx, row_begin, row_end, _ptr(col0), _ptr(off), C.byref(nw)),
                        "hc_band_layout")
        return col0[:nrows], off, int(nw.value)

    def score_band(self, row_begin: int = 0, row_end: Optional[int] = None, out: Optional[np.ndarray] = None,
                   amb_cap: int = 4096, fetch: bool = True) -> Adjacency:
        row_end = self._packed.n_reads if row_end is None else row_end
        col0, off, nw = self.band_layout(row_begin, row_end)
        if fetch and out is None:
            out = np.zeros(max(nw, 1), np.uint32)
        amb = np.zeros(max(amb_cap, 1), AMBIGUOUS_DTYPE)
        st = _BandStats()
        if self._multi:
            per = (_BandStats * len(self.devices))()
            self._check(self.L.hc_multi_score_band(self._multi, _ptr(out) if fetch else None, nw,
                                                   _ptr(amb) if amb_cap > 0 else None, amb_cap, C.byref(st), per), "hc_multi_score_band")
            self.per_device_stats = [{k: getattr(p, k) for k, _ in _BandStats._fields_} for p in per]
        else:
            self._check(self.L.hc_score_band(self._ctx, row_begin, row_end, _ptr(out) if fetch else None, nw,
                                             _ptr(amb) if amb_cap > 0 else None, amb_cap, C.byref(st)), "hc_score_band")
        stats = {k: getattr(st, k) for k, _ in _BandStats._fields_}
        words = out[:nw] if fetch else np.zeros(0, np.uint32)
        return Adjacency(row_begin, row_end, col0, off, words, stats, amb[:min(int(st.ambiguous), amb_cap)].copy())

    def set_kernel(self, mode: int) -> None:
        """KERNEL_AUTO / KERNEL_GENERIC / KERNEL_TILED / KERNEL_TILED2 / KERNEL_STREAMED (include/hc_edge.h: hc_set_kernel)."""
        for ctx in self.contexts():
            self._check(self.L.hc_set_kernel(ctx, mode), "hc_set_kernel")

    def last_kernel(self) -> int:
        return int(self.L.hc_last_kernel(self._ctx))

    def score_band_detail(self, row_begin: int = 0, row_end: Optional[int] = None):
        """Tiled kernel with per-slot detail: returns (Adjacency, pi, pj, detail) where detail[k] holds the
        values of the ordered pair (pi[k], pj[k]) -- every band slot with j < i, row-major."""
        row_end = self._packed.n_reads if row_end is None else row_end
        col0, off, nw = self.band_layout(row_begin, row_end)
        words = np.zeros(max(nw, 1), np.uint32)
        det = np.zeros(max(nw, 1) * 32, PAIR_RESULT_DTYPE)
        st = _BandStats()
        self._check(self.L.hc_score_band_detail(self._ctx, row_begin, row_end, _ptr(words), _ptr(det), nw, C.byref(st)),
                    "hc_score_band_detail")
        stats = {k: getattr(st, k) for k, _ in _BandStats._fields_}
        adj = Adjacency(row_begin, row_end, col0, off, words[:nw], stats, np.zeros(0, AMBIGUOUS_DTYPE))
        nrows = row_end - row_begin
        wcount = np.diff(off).astype(np.int64)
        row_of_word = np.repeat(np.arange(nrows, dtype=np.int64), wcount)
        slot = np.arange(nw * 32, dtype=np.int64)
        w = slot >> 5
        i = row_begin + row_of_word[w]
        j = col0[row_of_word[w]].astype(np.int64) + (w - off[row_of_word[w]]) * 32 + (slot & 31)
        keep = j < i
        return adj, i[keep].astype(np.int32), j[keep].astype(np.int32), det[:nw * 32][keep]

    def scoreBatch(self, reads) -> Adjacency:
        """One call per clique-finder iteration: pack, upload, score the whole band, fetch the adjacency."""
        self.upload(reads)
        self._adj = self.score_band()
        return self._adj

    def score_pairs(self, pi, pj) -> np.ndarray:
        pi = np.ascontiguousarray(pi, np.int32)
        pj = np.ascontiguousarray(pj, np.int32)
        out = np.zeros(pi.shape[0], PAIR_RESULT_DTYPE)
        self._check(self.L.hc_score_pairs(self._ctx, _ptr(pi), _ptr(pj), pi.shape[0], _ptr(out)), "hc_score_pairs")
        return out

    def score_pairs_static(self, pi, pj) -> np.ndarray:
        """hc_score_pairs_static: the same results from a static device scratch array, no device allocation (at most
        STATIC_MAX_PAIRS pairs per call)."""
        pi = np.ascontiguousarray(pi, np.int32)
        pj = np.ascontiguousarray(pj, np.int32)
        out = np.zeros(pi.shape[0], PAIR_RESULT_DTYPE)
        self._check(self.L.hc_score_pairs_static(self._ctx, _ptr(pi), _ptr(pj), pi.shape[0], _ptr(out)), "hc_score_pairs_static")
        return out

    def tables(self):
        eq = np.zeros((127, 127), np.float64)
        un = np.zeros((127, 127), np.float64)
        self._check(self.L.hc_get_tables(self._ctx, _ptr(eq), _ptr(un)), "hc_get_tables")
        return eq, un


class _QuasiParams(C.Structure):
    _fields_ = [("edge_quasi_cutoff_cliques", C.c_double), ("edge_quasi_cutoff_mixed", C.c_double),
                ("edge_quasi_cutoff_single", C.c_double), ("min_overlap_cliques", C.c_double),
                ("min_overlap_single", C.c_double), ("frameshift_merge", C.c_int32), ("reserved", C.c_int32),
                ("ambiguity_rel_tol", C.c_double)]


QUASI_RESULT_DTYPE = np.dtype([("status", np.int32), ("edge", np.int32), ("n_single", np.int32), ("flags", np.int32),
                               ("q", np.float64), ("cutoff", np.float64)], align=True)


QUASI_AMBIGUOUS, QUASI_SUBNORMAL = 1, 2


class QuasispeciesEdgeCalculator(NewEdgeCalculator):
    """GPU-backed drop-in for the reference's older scorer QuasispeciesEdgeCalculator (constructor arguments in the
    reference's order, src/QuasispeciesEdgeCalculator.h:57; ``maxPosition`` sizes the dense copy of simpson_map,
    ``device`` as in NewEdgeCalculator). Same batch interface; scores with hc_quasi_score_band / hc_quasi_score_pairs."""

    def __init__(self, Q: float = 0.9, edge_quasi_cutoff: float = 0.99, overlap: float = 0.9,
                 frameshift_merge: bool = False, simpson_map: Optional[Dict[int, float]] = None,
                 edge_quasi_cutoff_single: float = 0.95, overlap_single: float = 0.6,
                 edge_quasi_cutoff_mixed: float = 0.97, maxPosition: int = 0, device: int = 0,
                 ambiguity_rel_tol: float = 1e-12):
        super().__init__(Q, edge_quasi_cutoff, overlap, frameshift_merge, simpson_map, edge_quasi_cutoff_single,
                         overlap_single, edge_quasi_cutoff_mixed, maxPosition, False, device, ambiguity_rel_tol)
        self._qpar = _QuasiParams(edge_quasi_cutoff, edge_quasi_cutoff_mixed, edge_quasi_cutoff_single, overlap,
                                  overlap_single, int(bool(frameshift_merge)), 0, ambiguity_rel_tol)

    def setOverlapCliques(self, d: float) -> None:
        self._qpar.min_overlap_cliques = d
        self._adj = None

    def edgeBetween(self, ap1: AlignmentRecord, ap2: AlignmentRecord) -> bool:
        if ap1.name == ap2.name:                               # src/QuasispeciesEdgeCalculator.cpp:56
            return True
        return super().edgeBetween(ap1, ap2)

    def score_band(self, row_begin: int = 0, row_end: Optional[int] = None, out: Optional[np.ndarray] = None,
                   amb_cap: int = 0, fetch: bool = True) -> Adjacency:
        row_end = self._packed.n_reads if row_end is None else row_end
        col0, off, nw = self.band_layout(row_begin, row_end)
        if fetch and out is None:
            out = np.zeros(max(nw, 1), np.uint32)
        st = _BandStats()
        self._check(self.L.hc_quasi_score_band(self._ctx, C.byref(self._qpar), row_begin, row_end,
                                               _ptr(out) if fetch else None, nw, C.byref(st)), "hc_quasi_score_band")
        stats = {k: getattr(st, k) for k, _ in _BandStats._fields_}
        words = out[:nw] if fetch else np.zeros(0, np.uint32)
        return Adjacency(row_begin, row_end, col0, off, words, stats, np.zeros(0, AMBIGUOUS_DTYPE))

    def score_band_detail(self, row_begin: int = 0, row_end: Optional[int] = None):
        raise HcError("QuasispeciesEdgeCalculator: per-slot detail comes from score_pairs")

    def _score_pair_on(self, ctx, i: int, j: int) -> int:
        pi, pj = np.array([i], np.int32), np.array([j], np.int32)
        out = np.zeros(1, QUASI_RESULT_DTYPE)
        rc = self.L.hc_quasi_score_pairs(ctx, C.byref(self._qpar), _ptr(pi), _ptr(pj), 1, _ptr(out))
        if rc != HC_OK:
            raise HcError("hc_quasi_score_pairs failed (%d): %s" % (rc, self.L.hc_last_error(ctx).decode()))
        return int(out["edge"][0])

    def score_pairs(self, pi, pj) -> np.ndarray:
        pi = np.ascontiguousarray(pi, np.int32)
        pj = np.ascontiguousarray(pj, np.int32)
        out = np.zeros(pi.shape[0], QUASI_RESULT_DTYPE)
        self._check(self.L.hc_quasi_score_pairs(self._ctx, C.byref(self._qpar), _ptr(pi), _ptr(pj), pi.shape[0], _ptr(out)),
                    "hc_quasi_score_pairs")
        return out


def partition_rows(batch: ReadBatch, n_parts: int):
    """hc_partition_rows: (cuts, first) -- part p scores rows [cuts[p], cuts[p+1]) and holds reads [first[p], cuts[p+1])."""
    L = load_library()
    cuts = np.zeros(n_parts + 1, np.int32)
    first = np.zeros(n_parts, np.int32)
    cf, cl = np.ascontiguousarray(batch.cov_first, np.int32), np.ascontiguousarray(batch.cov_last, np.int32)
    rc = L.hc_partition_rows(_ptr(cf), _ptr(cl), batch.n, n_parts, _ptr(cuts), _ptr(first))
    if rc != HC_OK:
        raise HcError("hc_partition_rows failed (%d): %s" % (rc, L.hc_last_error(None).decode()))
    return cuts, first


def pack_only(batch: ReadBatch) -> PackedBatch:
    """hc_pack without a device (host C++ packing only)."""
    return PackedBatch(batch)


def measure_peaks(device: int = 0):
    """(shared-memory load GB/s, DMUL/s) of `device`, measured now by the library's probes."""
    L = load_library()
    a, b = C.c_double(0), C.c_double(0)
    rc = L.hc_measure_peaks(device, C.byref(a), C.byref(b))
    if rc != HC_OK:
        raise HcError("hc_measure_peaks failed (%d): %s" % (rc, L.hc_last_error(None).decode()))
    return a.value, b.value


def device_count() -> int:
    return int(load_library().hc_device_count())


# ---- super-read construction (include/hc_edge.h hc_merger_* / hc_merge_*) ------------------------------
MERGED_RECORD_DTYPE = np.dtype([("status", np.int32), ("single_end", np.int32), ("start", np.int32, 2), ("end", np.int32, 2),
                                ("n_ops", np.int32, 2), ("n_seq", np.int32, 2)])
HC_MERGE_OK, HC_MERGE_UNDEFINED, HC_MERGE_OVERRUN = 0, 1, 2


class MergedCliques:
    """Result of Merger.merge: per clique the super-read's header (`info`) and its unrolled CIGAR(s), bases and quality
    chars, compacted back to back (clique-major, segment 1 then segment 2)."""

    def __init__(self, info, cig, seq, qual, kernel_ms, folds):
        self.info, self.cig, self.seq, self.qual, self.kernel_ms, self.folds = info, cig, seq, qual, kernel_ms, folds
        ok = info["status"] == HC_MERGE_OK
        nseg = np.where(info["single_end"] != 0, 1, 2)
        no = np.where(ok, info["n_ops"][:, 0] + np.where(nseg == 2, info["n_ops"][:, 1], 0), 0).astype(np.int64)
        ns = np.where(ok, info["n_seq"][:, 0] + np.where(nseg == 2, info["n_seq"][:, 1], 0), 0).astype(np.int64)
        self.ops_at = np.concatenate([[0], np.cumsum(no)])
        self.seq_at = np.concatenate([[0], np.cumsum(ns)])

    def __len__(self):
        return self.info.shape[0]

    def segments(self, c: int):
        """(single_end, seg1, seg2) with seg = (start, end, cigar, sequence, qualities), seg2 None for a single-end result;
        None where the reference's merge is undefined (info['status'] != 0)."""
        r = self.info[c]
        if r["status"] != HC_MERGE_OK:
            return None
        oa, sa = int(self.ops_at[c]), int(self.seq_at[c])
        out = []
        for sg in range(1 if r["single_end"] else 2):
            no, ns = int(r["n_ops"][sg]), int(r["n_seq"][sg])
            out.append((int(r["start"][sg]), int(r["end"][sg]), self.cig[oa:oa + no].tobytes(), self.seq[sa:sa + ns].tobytes(),
                        self.qual[sa:sa + ns].tobytes()))
            oa += no
            sa += ns
        return bool(r["single_end"]), out[0], (None if r["single_end"] else out[1])

    def record(self, c: int, clique_id: Optional[int] = None) -> Optional[AlignmentRecord]:
        """The super-read as an AlignmentRecord named like the reference names it (src/AlignmentRecord.cpp:246)."""
        s = self.segments(c)
        if s is None:
            return None
        single, a, b = s
        rec = AlignmentRecord("Clique_%d" % (c if clique_id is None else clique_id), single, a[0], a[1], a[2], a[3], a[4])
        if not single:
            rec.start2, rec.end2, rec.cigar2_unrolled, rec.sequence2, rec.qualities2 = b
        return rec


class Merger:
    """hc_merger: grow-only device buffers, one stream; set_reads once per finder round, merge per batch of cliques."""

    def __init__(self, device: int = 0):
        self.L = load_library()
        self.L.hc_merger_create.argtypes = [C.c_int32, C.c_void_p]
        self.L.hc_merger_destroy.argtypes = [C.c_void_p]
        self.L.hc_merger_last_error.restype = C.c_char_p
        self.L.hc_merger_last_error.argtypes = [C.c_void_p]
        self.L.hc_merger_set_reads.argtypes = [C.c_void_p, C.c_void_p]
        self.L.hc_merge_run.argtypes = [C.c_void_p, C.c_int32] + [C.c_void_p] * 6
        self.L.hc_merge_fetch.argtypes = [C.c_void_p] * 4
        self.L.hc_merge_fold_counts.argtypes = [C.c_void_p] * 3
        self._h = C.c_void_p()
        rc = self.L.hc_merger_create(C.c_int32(device), C.byref(self._h))
        if rc != HC_OK:
            self._h = None
            raise HcError("hc_merger_create failed (%d): %s" % (rc, self.L.hc_last_error(None).decode()))
        self._keep = None

    def __del__(self):
        try:
            if self._h:
                self.L.hc_merger_destroy(self._h)
                self._h = None
        except Exception:
            pass

    def _check(self, rc, what):
        if rc != HC_OK:
            raise HcError("%s failed (%d): %s" % (what, rc, self.L.hc_merger_last_error(self._h).decode()))

    def set_reads(self, reads: ReadBatch):
        self._keep = [np.ascontiguousarray(getattr(reads, k)) for k in (
            "single_end", "is_clique", "ncov", "cov_first", "cov_last", "seg_start", "seg_end", "cig_off", "cig_len",
            "seq_off", "seq_len", "cig", "seq", "qual")]
        rd = _Reads(reads.n, *[_ptr(a) for a in self._keep])
        self._check(self.L.hc_merger_set_reads(self._h, C.byref(rd)), "hc_merger_set_reads")

    def merge(self, cliques: Sequence[Sequence[int]]) -> MergedCliques:
        n = len(cliques)
        member_off = np.zeros(n + 1, np.int64)
        member_off[1:] = np.cumsum([len(c) for c in cliques])
        members = np.ascontiguousarray(np.concatenate([np.asarray(c, np.int32) for c in cliques]) if n else np.zeros(0, np.int32))
        info = np.zeros(max(n, 1), MERGED_RECORD_DTYPE)
        to, ts, ms = C.c_int64(0), C.c_int64(0), C.c_float(0.0)
        self._check(self.L.hc_merge_run(self._h, C.c_int32(n), _ptr(member_off), _ptr(members), _ptr(info), C.byref(to), C.byref(ts),
                                        C.byref(ms)), "hc_merge_run")
        cig = np.zeros(max(int(to.value), 1), np.uint8)
        seq = np.zeros(max(int(ts.value), 1), np.uint8)
        qual = np.zeros(max(int(ts.value), 1), np.uint8)
        self._check(self.L.hc_merge_fetch(self._h, _ptr(cig), _ptr(seq), _ptr(qual)), "hc_merge_fetch")
        a, b = C.c_int64(0), C.c_int64(0)
        self.L.hc_merge_fold_counts(self._h, C.byref(a), C.byref(b))
        return MergedCliques(info[:n], cig[:int(to.value)], seq[:int(ts.value)], qual[:int(ts.value)], float(ms.value),
                             {"column_space": int(a.value), "walked": int(b.value)})


def merge_cliques(reads: ReadBatch, cliques: Sequence[Sequence[int]], device: int = 0) -> MergedCliques:
    """AlignmentRecord(alignments, clique_id) (src/AlignmentRecord.cpp:211-249) for a batch of cliques on the GPU:
    `cliques[c]` lists read indices into `reads` in the order of the reference's `alignments` vector."""
    m = Merger(device)
    m.set_reads(reads)
    return m.merge(cliques)
