"""ctypes face of oracle/_ref/libhcref.so: the UNMODIFIED reference (compiled by oracle/Makefile from
/root/reference where it lies) behind oracle/ref_api.cpp. TEST INFRASTRUCTURE ONLY.

Nothing here reads /root/reference at run time; only `build()` does, and only when it is present
(this container). On the GPU box the prebuilt oracle/_ref/ that travelled with the snapshot is used."""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = os.path.join(_HERE, "_ref", "libhcref.so")
_LIB_GPU = os.path.join(_HERE, "_ref", "libhcref_gpu.so")   # + integration/GpuEdgeCalculator, linked with libhcgpu.so
CLI_REF = os.path.join(_HERE, "_ref", "haploclique_ref")
CLI_GPU = os.path.join(_HERE, "_ref", "haploclique_gpu")
REFERENCE_ROOT = "/root/reference"


def available() -> bool:
    return os.path.exists(_LIB)


def gpu_adapter_available() -> bool:
    return os.path.exists(_LIB_GPU)


def build() -> bool:
    if os.path.exists(os.path.join(REFERENCE_ROOT, "src", "NewEdgeCalculator.cpp")):
        subprocess.check_call(["make", "-s", "-j8", "-C", _HERE, "ref"])
        if os.path.exists(os.path.join(_HERE, "..", "haploclique_b200", "libhcgpu.so")):
            subprocess.check_call(["make", "-s", "-j8", "-C", _HERE, "gpu"])
    return available()


_lib = None
_lib_gpu = None


def _declare(L):
    L.hcref_new.restype = C.c_void_p
    L.hcref_error.restype = C.c_char_p
    L.hcref_score_all.restype = C.c_long
    L.hcref_time_pairs.restype = C.c_double
    L.hcref_max_position.restype = C.c_uint
    return L


def lib_gpu():
    """libhcref_gpu.so: the same face plus hcref_iterate_gpu / hcref_edge_gpu (the reference's finders driven by the
    product's GpuEdgeCalculator)."""
    global _lib_gpu
    if _lib_gpu is None:
        if not gpu_adapter_available():
            raise RuntimeError("oracle/_ref/libhcref_gpu.so missing: run `make -C oracle gpu` where /root/reference exists")
        _lib_gpu = _declare(C.CDLL(_LIB_GPU))
    return _lib_gpu


def lib():
    global _lib
    if _lib is None:
        if not available():
            raise RuntimeError("oracle/_ref/libhcref.so missing: run `make -C oracle ref` where /root/reference exists")
        L = C.CDLL(_LIB)
        L.hcref_new.restype = C.c_void_p
        L.hcref_error.restype = C.c_char_p
        L.hcref_score_all.restype = C.c_long
        L.hcref_time_pairs.restype = C.c_double
        L.hcref_max_position.restype = C.c_uint
        for name in ("hcref_free", "hcref_error", "hcref_load_bam", "hcref_add_alignment", "hcref_finish_load",
                     "hcref_restore_fixture", "hcref_n", "hcref_max_position", "hcref_record_info",
                     "hcref_record_data", "hcref_set_params", "hcref_edge", "hcref_prefilter",
                     "hcref_pair_detail", "hcref_score_all", "hcref_time_pairs", "hcref_iterate"):
            getattr(L, name).argtypes = None
        _lib = L
    return _lib


def _p(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None else None


class Reference:
    """Holds reference AlignmentRecord objects and one NewEdgeCalculator."""

    def __init__(self, gpu_adapter: bool = False):
        self.L = lib_gpu() if gpu_adapter else lib()
        self.h = C.c_void_p(self.L.hcref_new())

    def __del__(self):
        try:
            self.L.hcref_free(self.h)
        except Exception:
            pass

    # -- loading ------------------------------------------------------------------------------------
    def load_bam(self, path: str) -> int:
        n = self.L.hcref_load_bam(self.h, path.encode())
        if n < 0:
            raise RuntimeError(self.L.hcref_error(self.h).decode())
        return n

    def bam_writer_open(self, path: str, ref_name: str, ref_len: int) -> None:
        if self.L.hcref_bam_writer_open(self.h, path.encode(), ref_name.encode(), C.c_int(ref_len)) != 0:
            raise RuntimeError(self.L.hcref_error(self.h).decode())

    def bam_writer_close(self) -> None:
        self.L.hcref_bam_writer_close(self.h)

    def restore_fixture(self, path: str) -> int:
        return self.L.hcref_restore_fixture(self.h, path.encode())

    def add_batch(self, batch, sort: bool = False) -> int:
        """Feeds a ReadBatch as BAM alignments through the reference's own constructors
        (AlignmentRecord(BamAlignment), pairWith for the second segment). Names are made unique per
        read unless the batch carries names; a 'Clique' marker in the name is preserved."""
        for r in range(batch.n):
            base = batch.names[r] if batch.names else ("Clique_%d" % r if batch.is_clique[r] else "read_%d" % r)
            nm = ("%s#%d" % (base, r)).encode()
            for s in range(1 if batch.single_end[r] else 2):
                k = 2 * r + s
                c0, cl = int(batch.cig_off[k]), int(batch.cig_len[k])
                s0, sl = int(batch.seq_off[k]), int(batch.seq_len[k])
                self.L.hcref_add_alignment(self.h, nm, C.c_int(int(batch.seg_start[k]) - 1),
                                           batch.cig[c0:c0 + cl].tobytes(), C.c_int(cl),
                                           batch.seq[s0:s0 + sl].tobytes(), batch.qual[s0:s0 + sl].tobytes(),
                                           C.c_int(sl))
            if batch.single_end[r]:
                # close the record now so arrival order is the batch order
                self.L.hcref_finish_load(self.h, 0)
        return self.L.hcref_finish_load(self.h, 1 if sort else 0)

    @property
    def n(self) -> int:
        return self.L.hcref_n(self.h)

    def max_position(self) -> int:
        return int(self.L.hcref_max_position(self.h))

    def to_batch(self, check_cov: bool = True):
        """Dumps the reference's records into our ReadBatch (the packing input). check_cov=False: records restored from the
        reference's unit-test fixture files carry the cov_pos rows stored in the file, which need not be what
        coveredPositions() would build; the reference's values are then taken as they are."""
        from haploclique_b200.records import AlignmentRecord, ReadBatch
        recs = []
        info = (C.c_int * 12)()
        name = C.create_string_buffer(512)
        covs = []
        for i in range(self.n):
            self.L.hcref_record_info(self.h, i, info, name, 512)
            se, s1, e1, s2, e2, nc1, nq1, nc2, nq2, ncov, cf, cl = list(info)
            bufs = [C.create_string_buffer(max(k, 1)) for k in (nc1, nq1, nq1, nc2, nq2, nq2)]
            self.L.hcref_record_data(self.h, i, *bufs)
            rec = AlignmentRecord(name.value.decode(), bool(se), s1, e1, bufs[0].raw[:nc1], bufs[1].raw[:nq1], bufs[2].raw[:nq1])
            if not se:
                rec.start2, rec.end2 = s2, e2
                rec.cigar2_unrolled, rec.sequence2, rec.qualities2 = bufs[3].raw[:nc2], bufs[4].raw[:nq2], bufs[5].raw[:nq2]
            recs.append(rec)
            covs.append((ncov, cf, cl))
        b = ReadBatch.from_records(recs)
        cov = np.array(covs, np.int32).reshape(-1, 3)
        # coveredPositions restated in records.py must agree with the reference's cov_pos
        if check_cov:
            assert (b.ncov == cov[:, 0]).all() and (b.cov_first == cov[:, 1]).all() and (b.cov_last == cov[:, 2]).all()
        else:
            b.ncov[:], b.cov_first[:], b.cov_last[:] = cov[:, 0], cov[:, 1], cov[:, 2]
        return b

    # -- calculator ---------------------------------------------------------------------------------
    def set_params(self, Q=0.9, edge_quasi_cutoff=0.99, overlap=0.9, frameshift_merge=False, simpson_map=None,
                   edge_quasi_cutoff_single=0.95, overlap_single=0.6, edge_quasi_cutoff_mixed=0.97,
                   maxPosition=None, noProb0=False):
        sm = simpson_map or {}
        pos = np.array(list(sm.keys()), np.int32)
        val = np.array(list(sm.values()), np.float64)
        self.L.hcref_set_params(self.h, C.c_double(Q), C.c_double(edge_quasi_cutoff), C.c_double(overlap),
                                C.c_int(int(frameshift_merge)), _p(pos), _p(val), C.c_int(len(sm)),
                                C.c_double(edge_quasi_cutoff_single), C.c_double(overlap_single),
                                C.c_double(edge_quasi_cutoff_mixed),
                                C.c_long(-1 if maxPosition is None else int(maxPosition)), C.c_int(int(noProb0)))

    def edge(self, i: int, j: int) -> bool:
        return bool(self.L.hcref_edge(self.h, i, j))

    # -- super-read construction ------------------------------------------------
    def compute_entry(self, b1: int, q1: int, b2: int, q2: int):
        out = C.create_string_buffer(2)
        self.L.hcref_compute_entry(C.c_char(bytes([b1])), C.c_char(bytes([q1])), C.c_char(bytes([b2])), C.c_char(bytes([q2])), out)
        return out.raw[0], out.raw[1]

    def merge_clique(self, members, clique_id: int = 0):
        """AlignmentRecord(alignments, clique_id) over records[members]; returns (single_end, start1, end1, cigar, seq, qual)."""
        mem = np.ascontiguousarray(members, np.int32)
        i = self.L.hcref_merge_clique(self.h, _p(mem), C.c_int(mem.size), C.c_uint(clique_id))
        info = (C.c_int * 12)()
        name = C.create_string_buffer(512)
        self.L.hcref_record_info(self.h, i, info, name, 512)
        se, s1, e1, _, _, nc1, nq1 = list(info)[:7]
        bufs = [C.create_string_buffer(max(k, 1)) for k in (nc1, nq1, nq1, 1, 1, 1)]
        self.L.hcref_record_data(self.h, i, *bufs)
        return bool(se), s1, e1, bufs[0].raw[:nc1], bufs[1].raw[:nq1], bufs[2].raw[:nq1]

    def merge_clique_full(self, members, clique_id: int = 0):
        """AlignmentRecord(alignments, clique_id) over records[members], any mix of single-end / paired-end members;
        returns (single_end, seg1, seg2) with seg = (start, end, cigar, seq, qual), seg2 None for a single-end result."""
        mem = np.ascontiguousarray(members, np.int32)
        i = self.L.hcref_merge_clique(self.h, _p(mem), C.c_int(mem.size), C.c_uint(clique_id))
        info = (C.c_int * 12)()
        name = C.create_string_buffer(512)
        self.L.hcref_record_info(self.h, i, info, name, 512)
        se, s1, e1, s2, e2, nc1, nq1, nc2, nq2 = list(info)[:9]
        bufs = [C.create_string_buffer(max(k, 1)) for k in (nc1, nq1, nq1, nc2, nq2, nq2)]
        self.L.hcref_record_data(self.h, i, *bufs)
        seg1 = (s1, e1, bufs[0].raw[:nc1], bufs[1].raw[:nq1], bufs[2].raw[:nq1])
        seg2 = None if se else (s2, e2, bufs[3].raw[:nc2], bufs[4].raw[:nq2], bufs[5].raw[:nq2])
        return bool(se), seg1, seg2

    def score_all(self):
        n = self.n
        npairs = n * (n - 1) // 2
        out = {"edge": np.zeros(npairs, np.uint8), "pre": np.zeros(npairs, np.uint8), "gap": np.zeros(npairs, np.uint8),
               "cc": np.zeros(npairs, np.int32), "tc": np.zeros(npairs, np.int32),
               "probM": np.ones(npairs, np.float64), "prob0": np.zeros(npairs, np.float64)}
        edges = self.L.hcref_score_all(self.h, _p(out["edge"]), _p(out["pre"]), _p(out["gap"]), _p(out["cc"]),
                                       _p(out["tc"]), _p(out["probM"]), _p(out["prob0"]))
        return int(edges), out

    def quasi_pairs(self, pi, pj, edge_quasi_cutoff=0.99, overlap=0.9, frameshift_merge=False, simpson_map=None,
                    edge_quasi_cutoff_single=0.95, overlap_single=0.6, edge_quasi_cutoff_mixed=0.97):
        """QuasispeciesEdgeCalculator on an explicit pair list: (q = computeOverlap(...), edgeBetween(...)) per pair."""
        pi = np.ascontiguousarray(pi, np.int32)
        pj = np.ascontiguousarray(pj, np.int32)
        sm = simpson_map or {}
        pos = np.array(list(sm.keys()), np.int32)
        val = np.array(list(sm.values()), np.float64)
        q = np.zeros(pi.shape[0], np.float64)
        e = np.zeros(pi.shape[0], np.uint8)
        self.L.hcref_quasi_pairs.restype = C.c_long
        self.L.hcref_quasi_pairs(self.h, _p(pi), _p(pj), C.c_long(pi.shape[0]), C.c_double(edge_quasi_cutoff),
                                 C.c_double(overlap), C.c_int(int(frameshift_merge)), _p(pos), _p(val), C.c_int(len(sm)),
                                 C.c_double(edge_quasi_cutoff_single), C.c_double(overlap_single),
                                 C.c_double(edge_quasi_cutoff_mixed), _p(q), _p(e))
        return q, e

    def quasi_gpu_pairs(self, pi, pj, edge_quasi_cutoff=0.99, overlap=0.9, frameshift_merge=False, simpson_map=None,
                        edge_quasi_cutoff_single=0.95, overlap_single=0.6, edge_quasi_cutoff_mixed=0.97):
        """The same pairs through integration/'s GpuQuasispeciesEdgeCalculator (needs gpu_adapter=True and a device)."""
        pi = np.ascontiguousarray(pi, np.int32)
        pj = np.ascontiguousarray(pj, np.int32)
        sm = simpson_map or {}
        pos = np.array(list(sm.keys()), np.int32)
        val = np.array(list(sm.values()), np.float64)
        e = np.zeros(pi.shape[0], np.uint8)
        self.L.hcref_quasi_gpu_pairs.restype = C.c_long
        n = self.L.hcref_quasi_gpu_pairs(self.h, _p(pi), _p(pj), C.c_long(pi.shape[0]), C.c_double(edge_quasi_cutoff),
                                         C.c_double(overlap), C.c_int(int(frameshift_merge)), _p(pos), _p(val), C.c_int(len(sm)),
                                         C.c_double(edge_quasi_cutoff_single), C.c_double(overlap_single),
                                         C.c_double(edge_quasi_cutoff_mixed), _p(e))
        if n < 0:
            raise RuntimeError("hcref_quasi_gpu_pairs failed (%d): %s" % (n, self.L.hcref_error(self.h).decode()))
        return e

    def time_pairs(self, pi, pj, threads: int = 1):
        pi = np.ascontiguousarray(pi, np.int32)
        pj = np.ascontiguousarray(pj, np.int32)
        edges = C.c_long(0)
        e = np.zeros(pi.shape[0], np.uint8)
        secs = self.L.hcref_time_pairs(self.h, _p(pi), _p(pj), C.c_long(pi.shape[0]), C.c_int(threads),
                                       C.byref(edges), _p(e))
        return float(secs), int(edges.value), e

    def iterate(self, ct: int, stdev: float, use_bk=False, filter_singletons=False, significance=0.0, logfile=""):
        """One pass of the CLI main loop (haploclique.cpp:349-372). Returns (n_next, ct, edges, converged, stdev)."""
        state = (C.c_int * 3)(ct, 0, 0)
        sd = C.c_double(stdev)
        n = self.L.hcref_iterate(self.h, C.c_int(int(use_bk)), C.c_int(int(filter_singletons)),
                                 C.c_double(significance), state, C.byref(sd), logfile.encode())
        return n, state[0], state[1], bool(state[2]), sd.value

    def iterate_gpu(self, ct: int, stdev: float, use_bk=False, filter_singletons=False, significance=0.0, logfile=""):
        """The same pass with GpuEdgeCalculator (integration/) plugged into the reference's finder; needs a CUDA device.
        Returns (n_next, ct, edges counted by the finder, converged, stdev, edges in the GPU adjacency)."""
        state = (C.c_int * 3)(ct, 0, 0)
        sd = C.c_double(stdev)
        ge = C.c_long(0)
        n = self.L.hcref_iterate_gpu(self.h, C.c_int(int(use_bk)), C.c_int(int(filter_singletons)),
                                     C.c_double(significance), state, C.byref(sd), logfile.encode(), C.byref(ge))
        if n < 0:
            raise RuntimeError("hcref_iterate_gpu failed (%d): %s" % (n, self.L.hcref_error(self.h).decode()))
        return n, state[0], state[1], bool(state[2]), sd.value, int(ge.value)

    def edge_gpu(self, i: int, j: int) -> bool:
        r = self.L.hcref_edge_gpu(self.h, i, j)
        if r < 0:
            raise RuntimeError("hcref_edge_gpu failed: %s" % self.L.hcref_error(self.h).decode())
        return bool(r)
