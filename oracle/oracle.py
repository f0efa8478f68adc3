"""ORACLE — TEST INFRASTRUCTURE ONLY.

ctypes wrapper around oracle/_build/liboracle.so (the CPU restatement of the reference's
aligned-sequence materialisation path, oracle/v2m_oracle.hh). Only tests/, __graft_entry__.smoke()
and bench.py's cpu_baseline / --impl reference legs may import this module; the product package
(vcf2multialign_b200/) never does.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "_build", "liboracle.so")
PLOIDY_MAX = 0xFFFFFFFF

_lib = None


def build(force: bool = False) -> str:
    src_newer = (not os.path.exists(_LIB_PATH)) or any(
        os.path.getmtime(os.path.join(_HERE, f)) > os.path.getmtime(_LIB_PATH) for f in ("capi.cc", "v2m_oracle.hh")
    )
    if force or src_newer:
        subprocess.run(["make", "-C", _HERE], check=True, capture_output=True)
    return _LIB_PATH


def lib():
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(_LIB_PATH):
        build()
    L = C.CDLL(_LIB_PATH)
    vp, u64, u32, i32, cp = C.c_void_p, C.c_uint64, C.c_uint32, C.c_int, C.c_char_p
    L.orc_last_error.restype = cp
    L.orc_transpose.argtypes = [vp, u64, u64, vp]
    L.orc_transpose8x8.restype = u64
    L.orc_transpose8x8.argtypes = [u64, i32]
    L.orc_graph_from_files.restype = vp
    L.orc_graph_from_files.argtypes = [cp, cp, cp]
    L.orc_graph_from_arrays.restype = vp
    L.orc_graph_from_arrays.argtypes = [u64, u64, vp, vp, vp, vp, vp, vp, vp, vp, u64, u64, u32, vp, cp, vp, u64]
    L.orc_graph_free.argtypes = [vp]
    L.orc_graph_dims.argtypes = [vp, vp]
    L.orc_time_paths_loop.restype = C.c_double
    L.orc_time_paths_loop.argtypes = [vp, u64, C.c_uint32, vp, vp, vp, vp, u64, u64, vp, vp]
    L.orc_graph_trace_dims.restype = None
    L.orc_graph_trace_dims.argtypes = [vp, vp]
    L.orc_graph_trace_copy.restype = None
    L.orc_graph_trace_copy.argtypes = [vp] * 7
    L.orc_graph_copy.argtypes = [vp] * 12
    L.orc_graph_overlap.restype = C.c_longlong
    L.orc_graph_overlap.argtypes = [vp, u64, vp, vp, u64]
    L.orc_output_sequence.restype = C.c_longlong
    L.orc_output_sequence.argtypes = [vp, u32, cp, i32, vp, u64]
    L.orc_haplotype_output_a2m.restype = C.c_longlong
    L.orc_haplotype_output_a2m.argtypes = [vp, cp, i32, i32, vp, u64]
    L.orc_haplotype_output_separate.restype = C.c_longlong
    L.orc_haplotype_output_separate.argtypes = [vp, cp, i32, i32, i32, vp, u64]
    L.orc_founder_new.restype = vp
    L.orc_founder_new.argtypes = [cp, i32, i32, i32]
    L.orc_founder_free.argtypes = [vp]
    L.orc_founder_find_cut_positions.argtypes = [vp, vp, u64]
    L.orc_founder_set_cut_positions.argtypes = [vp, vp, u64]
    L.orc_founder_cut_count.restype = u64
    L.orc_founder_cut_count.argtypes = [vp]
    L.orc_founder_score.restype = u32
    L.orc_founder_score.argtypes = [vp]
    L.orc_founder_cuts.argtypes = [vp, vp]
    L.orc_founder_find_matchings.argtypes = [vp, vp, u32]
    L.orc_founder_assigned.argtypes = [vp, vp, vp, vp]
    L.orc_founder_set_assigned.argtypes = [vp, u64, u64, vp]
    L.orc_founder_recorded_cuts.restype = u64
    L.orc_founder_recorded_cuts.argtypes = [vp]
    L.orc_founder_recorded.argtypes = [vp, u64, vp, vp, vp]
    L.orc_founder_output_a2m.restype = C.c_longlong
    L.orc_founder_output_a2m.argtypes = [vp, vp, vp, u64]
    L.orc_founder_output_separate.restype = C.c_longlong
    L.orc_founder_output_separate.argtypes = [vp, vp, i32, vp, u64]
    L.orc_time_rows.restype = C.c_double
    L.orc_time_rows.argtypes = [vp, u32, u32, i32, i32, vp]
    L.orc_time_transpose.restype = C.c_double
    L.orc_time_transpose.argtypes = [vp, u64, u64, vp]
    L.orc_synth_path_lines.restype = None
    L.orc_synth_path_lines.argtypes = [u32, u32, u32, u64, u64, vp, vp, vp, vp, i32]
    _lib = L
    return L


class OracleError(RuntimeError):
    pass


def _err():
    return OracleError(lib().orc_last_error().decode())


def _ptr(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def _opt(s):
    return None if s is None else s.encode()


def _two_pass(fn):
    """Call fn(dst, cap) → size twice: size query, then fill."""
    n = fn(None, 0)
    if n < 0:
        raise _err()
    buf = np.empty(max(n, 1), dtype=np.uint8)
    n2 = fn(_ptr(buf), n)
    if n2 != n:
        raise _err()
    return buf[:n].tobytes()


def _unpack_files(blob: bytes):
    out, i = [], 0
    while i < len(blob):
        j = blob.index(b"\0", i)
        name = blob[i:j].decode()
        ln = int.from_bytes(blob[j + 1 : j + 9], "little")
        out.append((name, blob[j + 9 : j + 9 + ln]))
        i = j + 9 + ln
    return out


def transpose(words: np.ndarray, rows: int, cols: int) -> np.ndarray:
    """v2m::transpose_matrix on the raw word array of a rows×cols column-major bit matrix."""
    words = np.ascontiguousarray(words, dtype=np.uint64)
    assert words.size == rows * cols // 64
    if cols == 0:
        return np.zeros(0, dtype=np.uint64)
    dst = np.zeros(words.size, dtype=np.uint64)
    if lib().orc_transpose(_ptr(words), rows, cols, _ptr(dst)):
        raise _err()
    return dst


def synth_path_lines(seed: int, first_copy: int, n_copies: int, e_pad: int, edge_site, edge_alt, edge_nalts, threads: int = 1) -> np.ndarray:
    """Measurement infrastructure (not the reference's algorithm): copy-major path lines of the synthetic workloads for a few
    chromosome copies, the CPU mirror of vcf2multialign_b200/csrc/synth.cu."""
    site = np.ascontiguousarray(edge_site, np.uint32)
    alt = np.ascontiguousarray(edge_alt, np.uint8)
    nalts = np.ascontiguousarray(edge_nalts, np.uint8)
    out = np.zeros(max(n_copies * (e_pad // 64), 1), np.uint64)
    lib().orc_synth_path_lines(seed & 0xFFFFFFFF, first_copy, n_copies, site.size, e_pad // 64, _ptr(site), _ptr(alt), _ptr(nalts), _ptr(out), threads)
    return out[: n_copies * (e_pad // 64)]


def transpose8x8(word: int, terms: bool = False) -> int:
    return int(lib().orc_transpose8x8(word, int(terms)))


class Graph:
    """Handle to an oracle-side variant_graph + reference sequence."""

    def __init__(self, handle):
        if not handle:
            raise _err()
        self._h = handle
        self._arrays = None

    @classmethod
    def from_files(cls, fasta: str, vcf: str, chr_id: str = "1") -> "Graph":
        return cls(lib().orc_graph_from_files(fasta.encode(), vcf.encode(), chr_id.encode()))

    @classmethod
    def from_arrays(cls, a: dict) -> "Graph":
        """`a` uses the key names of arrays() below (paths_ce optional)."""
        c = lambda k, dt: np.ascontiguousarray(a[k], dtype=dt)
        refpos, alnpos = c("reference_positions", np.uint64), c("aligned_positions", np.uint64)
        targets, csum = c("alt_edge_targets", np.uint64), c("alt_edge_count_csum", np.uint64)
        loff = c("label_offsets", np.uint64)
        lch = np.ascontiguousarray(np.frombuffer(bytes(a["label_chars"]), dtype=np.uint8))
        pec = c("paths_ec", np.uint64)
        pce = None if a.get("paths_ce") is None else c("paths_ce", np.uint64)
        pl = c("ploidy_csum", np.uint32)
        ref = np.ascontiguousarray(np.frombuffer(bytes(a["ref"]), dtype=np.uint8)) if not isinstance(a["ref"], np.ndarray) else c("ref", np.uint8)
        names = "".join(n + "\n" for n in a["sample_names"]).encode()
        h = lib().orc_graph_from_arrays(
            refpos.size, targets.size, _ptr(refpos), _ptr(alnpos), _ptr(targets), _ptr(csum), _ptr(loff),
            _ptr(lch) if lch.size else None, _ptr(pec), _ptr(pce), int(a["h_pad"]), int(a["e_pad"]),
            len(a["sample_names"]), _ptr(pl), names, _ptr(ref), ref.size)
        g = cls(h)
        g._keep = (refpos, alnpos, targets, csum, loff, lch, pec, pce, pl, ref)
        return g

    def __del__(self):
        try:
            if self._h:
                lib().orc_graph_free(self._h)
                self._h = None
        except Exception:
            pass

    def dims(self):
        d = np.zeros(14, dtype=np.uint64)
        lib().orc_graph_dims(self._h, _ptr(d))
        return [int(x) for x in d]

    def arrays(self) -> dict:
        if self._arrays is not None:
            return self._arrays
        d = self.dims()
        N, E, hp, ep, H, S, LC, RL = d[:8]
        ce_rows, ce_cols, names_len = d[10], d[11], d[12]
        a = dict(
            reference_positions=np.zeros(N, np.uint64), aligned_positions=np.zeros(N, np.uint64),
            alt_edge_targets=np.zeros(E, np.uint64), alt_edge_count_csum=np.zeros(N + 1, np.uint64),
            label_offsets=np.zeros(E + 1, np.uint64), label_chars=np.zeros(LC, np.uint8),
            paths_ec=np.zeros(hp * ep // 64, np.uint64), paths_ce=np.zeros(ce_rows * ce_cols // 64, np.uint64),
            ploidy_csum=np.zeros(S + 1, np.uint32), ref=np.zeros(RL, np.uint8),
        )
        names = np.zeros(max(names_len, 1), np.uint8)
        lib().orc_graph_copy(
            self._h, _ptr(a["reference_positions"]), _ptr(a["aligned_positions"]), _ptr(a["alt_edge_targets"]),
            _ptr(a["alt_edge_count_csum"]), _ptr(a["label_offsets"]), _ptr(a["label_chars"]), _ptr(a["paths_ec"]),
            _ptr(a["paths_ce"]), _ptr(a["ploidy_csum"]), _ptr(names), _ptr(a["ref"]))
        a["sample_names"] = [s for s in names[:names_len].tobytes().decode().split("\n") if s != ""] if names_len else []
        a["h_pad"], a["e_pad"], a["total_copies"] = hp, ep, H
        a["label_chars"] = a["label_chars"].tobytes()
        self._arrays = a
        return a

    def genotype_trace(self) -> dict:
        """What the "Paths" loop of build_variant_graph consumed (variant_graph.cc:379-425), flat: gt[records][copies],
        alt_offsets, edges_by_alt, ref_pos per record, target reference position per edge, reported overlaps (record, row)."""
        d = np.zeros(6, np.uint64)
        lib().orc_graph_trace_dims(self._h, _ptr(d))
        R, Hc, n_alts, n_edges, n_ov = (int(x) for x in d[:5])
        t = dict(gt=np.zeros(max(R * Hc, 1), np.uint16), alt_offsets=np.zeros(R + 1, np.uint64), edges_by_alt=np.zeros(max(n_alts, 1), np.uint64),
                 ref_pos=np.zeros(max(R, 1), np.uint64), edge_target_pos=np.zeros(max(n_edges, 1), np.uint64), overlaps=np.zeros(max(2 * n_ov, 1), np.uint64))
        lib().orc_graph_trace_copy(self._h, _ptr(t["gt"]), _ptr(t["alt_offsets"]), _ptr(t["edges_by_alt"]), _ptr(t["ref_pos"]),
                                   _ptr(t["edge_target_pos"]), _ptr(t["overlaps"]))
        t["gt"] = t["gt"][: R * Hc].reshape(R, Hc) if Hc else np.zeros((R, 0), np.uint16)
        t["edges_by_alt"], t["ref_pos"], t["edge_target_pos"] = t["edges_by_alt"][:n_alts], t["ref_pos"][:R], t["edge_target_pos"][:n_edges]
        t["overlaps"] = t["overlaps"][: 2 * n_ov].reshape(n_ov, 2)
        t["records"], t["copies"] = R, Hc
        return t

    def labels(self):
        a = self.arrays()
        off, ch = a["label_offsets"], a["label_chars"]
        return [ch[int(off[i]) : int(off[i + 1])].decode() for i in range(len(off) - 1)]

    def overlaps(self):
        out = []
        for i in range(self.dims()[8]):
            nums = np.zeros(4, np.uint64)
            buf = np.zeros(4096, np.uint8)
            n = lib().orc_graph_overlap(self._h, i, _ptr(nums), _ptr(buf), buf.size)
            var_id, sample = buf[:n].tobytes().decode().split("\t")
            out.append(dict(lineno=int(nums[0]), ref_pos=int(nums[1]), chrom_copy=int(nums[2]), gt=int(nums[3]), var_id=var_id, sample=sample))
        return out

    # --- rows ---
    def output_sequence(self, copy_idx: int, fasta_id: str | None = None, unaligned: bool = False) -> bytes:
        return _two_pass(lambda dst, cap: lib().orc_output_sequence(self._h, copy_idx, _opt(fasta_id), int(unaligned), dst, cap))

    def haplotype_output_a2m(self, chrom_id=None, output_ref=True, unaligned=False) -> bytes:
        return _two_pass(lambda dst, cap: lib().orc_haplotype_output_a2m(self._h, _opt(chrom_id), int(output_ref), int(unaligned), dst, cap))

    def haplotype_output_separate(self, chrom_id=None, output_ref=True, unaligned=False, include_header=True):
        return _unpack_files(_two_pass(lambda dst, cap: lib().orc_haplotype_output_separate(
            self._h, _opt(chrom_id), int(output_ref), int(unaligned), int(include_header), dst, cap)))

    # --- timing (bench only) ---
    def time_rows(self, first_copy: int, n_rows: int, threads: int = 1, bulk: bool = False):
        bases = C.c_uint64(0)
        s = lib().orc_time_rows(self._h, first_copy, n_rows, threads, int(bulk), C.byref(bases))
        return s, int(bases.value)


class FounderOutput:
    """founder_sequence_greedy_output of the oracle (founder_sequence_greedy_output.cc)."""

    def __init__(self, graph: Graph, chrom_id=None, output_ref=True, keep_ref_edges=False, unaligned=False):
        self.graph = graph
        self._f = lib().orc_founder_new(_opt(chrom_id), int(output_ref), int(keep_ref_edges), int(unaligned))

    def __del__(self):
        try:
            if self._f:
                lib().orc_founder_free(self._f)
                self._f = None
        except Exception:
            pass

    def find_cut_positions(self, min_distance: int) -> bool:
        r = lib().orc_founder_find_cut_positions(self._f, self.graph._h, min_distance)
        if r < 0:
            raise _err()
        return bool(r)

    def set_cut_positions(self, cuts):
        c = np.ascontiguousarray(cuts, dtype=np.uint64)
        lib().orc_founder_set_cut_positions(self._f, _ptr(c), c.size)

    def cut_positions(self) -> np.ndarray:
        n = lib().orc_founder_cut_count(self._f)
        c = np.zeros(n, np.uint64)
        if n:
            lib().orc_founder_cuts(self._f, _ptr(c))
        return c

    def score(self) -> int:
        return int(lib().orc_founder_score(self._f))

    def find_matchings(self, founder_count: int) -> bool:
        r = lib().orc_founder_find_matchings(self._f, self.graph._h, founder_count)
        if r < 0:
            raise _err()
        return bool(r)

    def assigned_samples(self) -> np.ndarray:
        """segments × founders array (logical indexing; the reference stores it column-major)."""
        rows, cols = C.c_uint64(0), C.c_uint64(0)
        lib().orc_founder_assigned(self._f, C.byref(rows), C.byref(cols), None)
        d = np.zeros(rows.value * cols.value, np.uint32)
        lib().orc_founder_assigned(self._f, C.byref(rows), C.byref(cols), _ptr(d))
        return d.reshape(cols.value, rows.value).T.copy()

    def set_assigned_samples(self, m: np.ndarray):
        m = np.asarray(m, dtype=np.uint32)
        flat = np.ascontiguousarray(m.T).reshape(-1)
        lib().orc_founder_set_assigned(self._f, m.shape[0], m.shape[1], _ptr(flat))

    def recorded_classes(self):
        """Per reached cut: dict(distinct, first_class, first_is_ref, rhs_eq[H], joined[(lhs, rhs, size)…] in permutation order)."""
        H = self.graph.dims()[4]
        out = []
        for i in range(lib().orc_founder_recorded_cuts(self._f)):
            meta = np.zeros(4, np.uint32)
            lib().orc_founder_recorded(self._f, i, _ptr(meta), None, None)
            rhs = np.zeros(H, np.uint32)
            joined = np.zeros(3 * max(int(meta[3]), 1), np.uint32)
            lib().orc_founder_recorded(self._f, i, _ptr(meta), _ptr(rhs), _ptr(joined))
            out.append(dict(distinct=int(meta[0]), first_class=int(meta[1]), first_is_ref=bool(meta[2]), rhs_eq=rhs,
                            joined=joined[: 3 * int(meta[3])].reshape(-1, 3).copy()))
        return out

    def output_a2m(self) -> bytes:
        return _two_pass(lambda dst, cap: lib().orc_founder_output_a2m(self._f, self.graph._h, dst, cap))

    def output_separate(self, include_header=True):
        return _unpack_files(_two_pass(lambda dst, cap: lib().orc_founder_output_separate(self._f, self.graph._h, int(include_header), dst, cap)))
