"""Host-side mirror of v2m::variant_graph (include/vcf2multialign/variant_graph.hh:36-79) as flat numpy arrays,
and its device-resident counterpart behind libv2m_cuda.so."""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass, field

import numpy as np

from . import _lib
from ._lib import PLOIDY_MAX, V2MError, check, lib, ptr


@dataclass
class VariantGraph:
    """Field names follow the reference struct; alt_edge_labels is flattened (label_offsets + label_chars), the two
    path matrices are the raw word arrays of lb::bit_matrix::values() (column-major, LSB-first)."""
    reference_positions: np.ndarray
    aligned_positions: np.ndarray
    alt_edge_targets: np.ndarray
    alt_edge_count_csum: np.ndarray
    label_offsets: np.ndarray
    label_chars: bytes
    paths_by_edge_and_chrom_copy: np.ndarray | None   # rows = h_pad copies, cols = e_pad edges
    paths_by_chrom_copy_and_edge: np.ndarray | None   # rows = e_pad edges, cols = h_pad copies
    h_pad: int
    e_pad: int
    sample_names: list
    ploidy_csum: np.ndarray

    @classmethod
    def from_arrays(cls, a: dict) -> "VariantGraph":
        u64 = lambda k: np.ascontiguousarray(a[k], dtype=np.uint64)
        return cls(
            u64("reference_positions"), u64("aligned_positions"), u64("alt_edge_targets"), u64("alt_edge_count_csum"),
            u64("label_offsets"), bytes(a["label_chars"]),
            None if a.get("paths_ec") is None else u64("paths_ec"),
            None if a.get("paths_ce") is None else u64("paths_ce"),
            int(a["h_pad"]), int(a["e_pad"]), list(a["sample_names"]), np.ascontiguousarray(a["ploidy_csum"], dtype=np.uint32))

    def node_count(self) -> int:
        return int(self.reference_positions.size)

    def edge_count(self) -> int:
        return int(self.alt_edge_targets.size)

    def total_chromosome_copies(self) -> int:
        return int(self.ploidy_csum[-1]) if self.ploidy_csum.size else 0

    def sample_ploidy(self, sample_idx: int) -> int:
        return int(self.ploidy_csum[1 + sample_idx] - self.ploidy_csum[sample_idx])

    def aligned_length(self) -> int:
        return int(self.aligned_positions[-1])


class DeviceGraph:
    """v2m_graph_t: the graph + reference resident in HBM (uploads once; K1 derives the missing path orientation)."""

    def __init__(self, graph: VariantGraph, ref_seq, d_paths_by_edge_and_chrom_copy: int | None = None):
        """d_paths_by_edge_and_chrom_copy: optional DEVICE pointer to the edge-major path words (e.g. from synth.path_bits_device);
        the host arrays of `graph` are then not consulted for paths."""
        self.graph = graph
        # ref_seq None: the graph goes up without the reference (enough for find_cut_positions / find_matchings); attach_reference later
        ref = None if ref_seq is None else (np.frombuffer(bytes(ref_seq), dtype=np.uint8) if not isinstance(ref_seq, np.ndarray) else np.ascontiguousarray(ref_seq, dtype=np.uint8))
        self.ref_len = 0 if ref is None else int(ref.size)
        labels = np.frombuffer(graph.label_chars, dtype=np.uint8) if len(graph.label_chars) else np.zeros(1, np.uint8)
        v = _lib.GraphView()
        v.node_count, v.edge_count = graph.node_count(), graph.edge_count()
        v.reference_positions, v.aligned_positions = ptr(graph.reference_positions), ptr(graph.aligned_positions)
        v.alt_edge_targets, v.alt_edge_count_csum = ptr(graph.alt_edge_targets), ptr(graph.alt_edge_count_csum)
        v.label_offsets, v.label_chars = ptr(graph.label_offsets), ptr(labels)
        if d_paths_by_edge_and_chrom_copy is not None:
            v.paths_by_edge_and_chrom_copy, v.paths_by_chrom_copy_and_edge, v.paths_on_device = d_paths_by_edge_and_chrom_copy, None, 1
        else:
            v.paths_by_edge_and_chrom_copy = ptr(graph.paths_by_edge_and_chrom_copy)
            v.paths_by_chrom_copy_and_edge = ptr(graph.paths_by_chrom_copy_and_edge)
        v.h_pad, v.e_pad = graph.h_pad, graph.e_pad
        v.total_chromosome_copies = graph.total_chromosome_copies()
        v.ref_seq, v.ref_len = ptr(ref), self.ref_len
        h = C.c_void_p()
        check(lib().v2m_cuda_graph_upload(C.byref(v), C.byref(h)))
        self._h = h

    def close(self):
        if getattr(self, "_h", None):
            lib().v2m_cuda_graph_free(self._h)
            self._h = None

    def attach_reference(self, ref_seq):
        ref = np.frombuffer(bytes(ref_seq), dtype=np.uint8) if not isinstance(ref_seq, np.ndarray) else np.ascontiguousarray(ref_seq, dtype=np.uint8)
        check(lib().v2m_cuda_graph_attach_reference(self._h, ptr(ref), ref.size))
        self.ref_len = int(ref.size)

    __del__ = close

    def aligned_length(self) -> int:
        out = C.c_uint64()
        check(lib().v2m_cuda_graph_aligned_length(self._h, C.byref(out)))
        return int(out.value)

    def paths_by_chrom_copy_and_edge(self) -> np.ndarray:
        out = np.zeros(self.graph.h_pad * self.graph.e_pad // 64, np.uint64)
        if out.size:
            check(lib().v2m_cuda_graph_download_paths_by_chrom_copy_and_edge(self._h, ptr(out)))
        return out

    def template(self) -> bytes:
        out = np.zeros(max(self.aligned_length(), 1), np.uint8)
        check(lib().v2m_cuda_graph_download_template(self._h, ptr(out)))
        return out[: self.aligned_length()].tobytes()

    def plan_haplotypes(self, copy_idx) -> "Plan":
        idx = np.ascontiguousarray(copy_idx, dtype=np.uint32)
        p = C.c_void_p()
        check(lib().v2m_cuda_plan_haplotypes(self._h, ptr(idx), idx.size, C.byref(p)))
        return Plan(self, p)

    def plan_founders(self, cut_positions, assigned_samples) -> "Plan":
        """assigned_samples: segments × founders (logical indexing); passed on in the reference's column-major order."""
        cuts = np.ascontiguousarray(cut_positions, dtype=np.uint64)
        m = np.asarray(assigned_samples, dtype=np.uint32)
        if m.ndim != 2 or m.shape[0] != cuts.size - 1:
            raise V2MError("assigned_samples must be segments × founders with segments = len(cut_positions) - 1")
        flat = np.asfortranarray(m).T.reshape(-1)              # no copy when m is column-major already (find_matchings returns it so)
        p = C.c_void_p()
        check(lib().v2m_cuda_plan_founders(self._h, ptr(cuts), cuts.size - 1, ptr(flat), m.shape[1], C.byref(p)))
        return Plan(self, p)

    def find_matchings(self, cut_positions, founder_count: int, keep_ref_edges: bool = False):
        """founder_sequence_greedy_output::find_matchings: K3 + host greedy inside the library. Returns the segments × founders
        matrix of assigned chromosome copies (logical indexing), or None where the reference returns false."""
        cuts = np.ascontiguousarray(cut_positions, dtype=np.uint64)
        n_seg = max(cuts.size - 1, 0)
        assigned = np.full(max(n_seg * founder_count, 1), PLOIDY_MAX, np.uint32)
        found = C.c_int(0)
        check(lib().v2m_cuda_find_matchings(self._h, ptr(cuts), cuts.size, founder_count, int(keep_ref_edges), ptr(assigned), C.byref(found)))
        if not found.value:
            return None
        return assigned[: n_seg * founder_count].reshape(founder_count, n_seg).T          # column-major like lb::matrix: plan_founders takes it as is

    def find_cut_positions(self, min_distance: int = 0):
        """find_initial_cut_positions_lambda_min (find_cut_positions.cc:93-211): K4 on the device, recurrence + trace-back on
        the host inside the library. Returns (cut position node ids, score), or (None, CUT_POSITION_SCORE_MAX) on failure."""
        cuts = np.zeros(max(int(self.graph.reference_positions.size), 1), np.uint64)
        n = C.c_uint64(0)
        score = C.c_uint32(0)
        found = C.c_int(0)
        check(lib().v2m_cuda_find_cut_positions(self._h, int(min_distance), ptr(cuts), cuts.size, C.byref(n), C.byref(score), C.byref(found)))
        if not found.value:
            return None, int(score.value)
        return cuts[: n.value].copy(), int(score.value)

    def last_cut_positions_timing(self):
        ms = (C.c_float * 4)()
        counts = (C.c_uint64 * 6)()
        check(lib().v2m_cuda_last_cut_positions_timing(ms, counts))
        return dict(local_ms=ms[0], compose_ms=ms[1], lists_ms=ms[2], total_host_ms=ms[3], chunks=counts[0], candidates=counts[1],
                    edges=counts[2], list_pairs=counts[3], cap=counts[4], attempts=counts[5])

    def segment_classes(self, cut_positions, want_rhs_eq: bool = True):
        """K3: per-cut class data for find_matchings. Returns a list (one entry per cut ≥ 1) of dicts with
        distinct, first_class, first_is_ref, rhs_eq (or None) and joined (k × 3: lhs_rep, rhs_rep, size; permutation order)."""
        cuts = np.ascontiguousarray(cut_positions, dtype=np.uint64)
        n = cuts.size
        H = self.graph.total_chromosome_copies()
        rhs = np.zeros(max((n - 1) * H, 1), np.uint32) if want_rhs_eq else None
        distinct = np.zeros(max(n - 1, 1), np.uint32)
        first = np.zeros(max(n - 1, 1), np.uint32)
        is_ref = np.zeros(max(n - 1, 1), np.uint8)
        offs = np.zeros(max(n, 1), np.uint64)
        cap = max(4 * H, 1024)
        while True:
            joined = np.zeros(3 * cap, np.uint32)
            st = lib().v2m_cuda_segment_classes(self._h, ptr(cuts), n, ptr(rhs), ptr(distinct), ptr(first), ptr(is_ref), ptr(offs), ptr(joined), cap)
            if st == 0:
                break
            if st == 1 and n and int(offs[n - 1]) > cap:
                cap = int(offs[n - 1])
                continue
            check(st)
        out = []
        for c in range(1, n):
            lo, hi = int(offs[c - 1]), int(offs[c])
            out.append(dict(distinct=int(distinct[c - 1]), first_class=int(first[c - 1]), first_is_ref=bool(is_ref[c - 1]),
                            rhs_eq=None if rhs is None else rhs[(c - 1) * H : c * H].copy(),
                            joined=joined[3 * lo : 3 * hi].reshape(-1, 3).copy()))
        return out


class Plan:
    """v2m_plan_t: a set of output rows with their taken-edge bits resolved on the device (K2a)."""

    def __init__(self, dgraph: DeviceGraph, handle):
        self.dgraph = dgraph
        self._p = handle

    def close(self):
        if getattr(self, "_p", None):
            lib().v2m_cuda_plan_free(self._p)
            self._p = None

    __del__ = close

    def row_count(self) -> int:
        out = C.c_uint32()
        check(lib().v2m_cuda_plan_row_count(self._p, C.byref(out)))
        return int(out.value)

    def taken(self) -> np.ndarray:
        n, We = self.row_count(), self.dgraph.graph.e_pad // 64
        out = np.zeros(max(n * We, 1), np.uint64)
        if n * We:
            check(lib().v2m_cuda_plan_download_taken(self._p, ptr(out)))
        return out[: n * We].reshape(n, We)

    def row_lengths(self, unaligned: bool = False) -> np.ndarray:
        out = np.zeros(max(self.row_count(), 1), np.uint64)
        check(lib().v2m_cuda_plan_row_lengths(self._p, int(unaligned), ptr(out)))
        return out[: self.row_count()]

    def emit_rows(self, row_begin=0, row_end=None, col_begin=0, col_end=None) -> np.ndarray:
        """Aligned rows as a (rows × columns) uint8 array (host memory)."""
        row_end = self.row_count() if row_end is None else row_end
        col_end = self.dgraph.aligned_length() if col_end is None else col_end
        out = np.zeros((max(row_end - row_begin, 0), max(col_end - col_begin, 0)), np.uint8)
        if out.size:
            check(lib().v2m_cuda_emit_rows(self._p, row_begin, row_end, col_begin, col_end, ptr(out), out.shape[1]))
        return out

    def emit_rows_unaligned(self, row_begin=0, row_end=None):
        row_end = self.row_count() if row_end is None else row_end
        lens = self.row_lengths(True)[row_begin:row_end]
        offs = np.zeros(max(lens.size, 1), np.uint64)
        if lens.size:
            offs[: lens.size] = np.concatenate([[0], np.cumsum(lens)[:-1]])
        buf = np.zeros(max(int(lens.sum()), 1), np.uint8)
        if lens.size:
            check(lib().v2m_cuda_emit_rows_unaligned(self._p, row_begin, row_end, ptr(offs), ptr(buf)))
        return [buf[int(o) : int(o) + int(l)].tobytes() for o, l in zip(offs[: lens.size], lens)]

    def write_a2m(self, ids, fd: int, unaligned: bool = False, row_done=None):
        blob = b"".join(i.encode() + b"\0" for i in ids)
        cb = _lib.ROW_CALLBACK(lambda r, ctx: row_done(r)) if row_done else C.cast(None, _lib.ROW_CALLBACK)
        check(lib().v2m_cuda_write_a2m(self._p, blob, int(unaligned), fd, cb, None))


GT_NULL = 0xFFFF
EDGE_MAX = 0xFFFFFFFFFFFFFFFF


def paths_from_genotypes(gt, alt_offsets, edges_by_alt, h_pad: int, e_pad: int, record_ref_pos=None, edge_target_pos=None):
    """The "Paths" loop of build_variant_graph (variant_graph.cc:379-425) on the device (K5): gt is records × chromosome copies
    (GT allele per output row; GT_NULL = '.'), ALT a of record r has edge edges_by_alt[alt_offsets[r] + a - 1] (EDGE_MAX: none).
    Returns (words of paths_by_edge_and_chrom_copy, overlaps) with overlaps a k × 2 array of (record, row) in the
    reference's reporting order, or None when the positions needed to detect them were not given."""
    gt = np.ascontiguousarray(gt, dtype=np.uint16)
    R, Hc = (gt.shape if gt.ndim == 2 else (0, 0))
    off = np.ascontiguousarray(alt_offsets, dtype=np.uint64)
    eba = np.ascontiguousarray(edges_by_alt, dtype=np.uint64)
    out = np.zeros(max(h_pad * e_pad // 64, 1), np.uint64)
    want = record_ref_pos is not None and edge_target_pos is not None
    rp = np.ascontiguousarray(record_ref_pos, dtype=np.uint64) if want else None
    tp = np.ascontiguousarray(edge_target_pos, dtype=np.uint64) if want else None
    n_edges = int(tp.size) if want else 0
    cap = max(1024, int(R))          # more overlaps than records is unusual; the call reports the needed count if it happens
    while True:
        ov = np.zeros(2 * cap, np.uint64)
        n = C.c_uint64(0)
        st = lib().v2m_cuda_paths_from_genotypes(ptr(gt), R, Hc, ptr(off), ptr(eba) if eba.size else None, ptr(rp), ptr(tp), n_edges, h_pad, e_pad,
                                                  ptr(out), ptr(ov) if want else None, cap if want else 0, C.byref(n) if want else None)
        if st != 0 and want and n.value > cap:
            cap = int(n.value)
            continue
        check(st)
        break
    return out[: h_pad * e_pad // 64], (ov[: 2 * n.value].reshape(-1, 2) if want else None)


def paths_from_genotypes_both(gt, alt_offsets, edges_by_alt, h_pad: int, e_pad: int):
    """K5 with both orientations born on the device (v2m_cuda_paths_from_genotypes_both): returns the words of
    paths_by_edge_and_chrom_copy and of paths_by_chrom_copy_and_edge."""
    gt = np.ascontiguousarray(gt, dtype=np.uint16)
    R, Hc = (gt.shape if gt.ndim == 2 else (0, 0))
    off = np.ascontiguousarray(alt_offsets, dtype=np.uint64)
    eba = np.ascontiguousarray(edges_by_alt, dtype=np.uint64)
    n_bytes = max(h_pad * e_pad // 8, 16)
    L = lib()
    d_ec, d_ce = L.v2m_cuda_device_alloc(n_bytes), L.v2m_cuda_device_alloc(n_bytes)
    if not d_ec or not d_ce:
        raise V2MError("device_alloc failed")
    try:
        check(L.v2m_cuda_paths_from_genotypes_both(ptr(gt), R, Hc, ptr(off), ptr(eba) if eba.size else None, None, None, 0, h_pad, e_pad, d_ec, d_ce, None, 0, None))
        ec, ce = np.zeros(max(h_pad * e_pad // 64, 1), np.uint64), np.zeros(max(h_pad * e_pad // 64, 1), np.uint64)
        if h_pad * e_pad:
            check(L.v2m_cuda_memcpy_d2h(ptr(ec), d_ec, h_pad * e_pad // 8))
            check(L.v2m_cuda_memcpy_d2h(ptr(ce), d_ce, h_pad * e_pad // 8))
    finally:
        L.v2m_cuda_device_free(d_ec)
        L.v2m_cuda_device_free(d_ce)
    return ec[: h_pad * e_pad // 64], ce[: h_pad * e_pad // 64]
