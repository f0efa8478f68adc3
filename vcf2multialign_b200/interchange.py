"""Graph / cut-position files of the reference (--input-graph, --output-graph, --input-cut-positions, --output-cut-positions):
cereal PortableBinary archives of variant_graph::serialize (include/vcf2multialign/variant_graph.hh:197-209) and
founder_sequence_greedy_output::cut_positions::serialize (include/vcf2multialign/output.hh:133-139). Python twin of
vcf2multialign_b200/host/v2m_cereal.hh (see there for the encoding); checked against archives written by the real cereal
headers (tests/golden/cereal/)."""
from __future__ import annotations

import io
import struct
import sys

import numpy as np

from .graph import VariantGraph


class ArchiveError(ValueError):
    pass


class _Reader:
    def __init__(self, f):
        self.f = f
        flag = self._raw(1)[0]
        if flag > 1:
            raise ArchiveError("portable binary archive: bad endianness flag")
        self.bo = "<" if flag == 1 else ">"

    def _raw(self, n):
        b = self.f.read(n)
        if len(b) != n:
            raise ArchiveError("portable binary archive: truncated input")
        return b

    def scalar(self, fmt):
        return struct.unpack(self.bo + fmt, self._raw(struct.calcsize(fmt)))[0]

    def count(self, element_bytes=1):
        n = self.scalar("Q")
        if n * max(element_bytes, 1) > 1 << 40:
            raise ArchiveError("portable binary archive: implausible element count")
        return n

    def vector(self, dtype):
        dt = np.dtype(dtype)
        n = self.count(dt.itemsize)
        a = np.frombuffer(self._raw(n * dt.itemsize), dtype=dt.newbyteorder(self.bo)) if n else np.zeros(0, dt)
        return a.astype(dt, copy=True)

    def string(self):
        return self._raw(self.count()).decode()

    def strings(self):
        return [self.string() for _ in range(self.count(8))]

    def bit_matrix(self):
        """→ (words, rows, cols): int_vector (words, element count) + stride."""
        words = self.vector(np.uint64)
        bits, stride = self.scalar("Q"), self.scalar("Q")
        if stride == 0 or bits % stride or words.size < (bits + 63) // 64:
            raise ArchiveError("portable binary archive: inconsistent bit matrix")
        return words[: (bits + 63) // 64].copy(), (stride if bits else 0), (bits // stride if bits else 0)


class _Writer:
    def __init__(self, f):
        self.f = f
        self.f.write(bytes([1 if sys.byteorder == "little" else 0]))

    def scalar(self, fmt, v):
        self.f.write(struct.pack("=" + fmt, v))

    def vector(self, a, dtype):
        a = np.ascontiguousarray(a, dtype=dtype)
        self.scalar("Q", a.size)
        self.f.write(a.tobytes())

    def string(self, s):
        b = s.encode()
        self.scalar("Q", len(b))
        self.f.write(b)

    def strings(self, ss):
        self.scalar("Q", len(ss))
        for s in ss:
            self.string(s)

    def bit_matrix(self, words, rows, cols):
        self.vector(np.zeros(0, np.uint64) if words is None else words, np.uint64)
        self.scalar("Q", rows * cols)
        self.scalar("Q", rows if rows else 1)


def load_variant_graph(src) -> VariantGraph:
    f = open(src, "rb") if isinstance(src, (str, bytes)) and not isinstance(src, io.IOBase) else src
    try:
        r = _Reader(f)
        r.scalar("I")                                   # cereal_class_version of variant_graph (0)
        refpos, alnpos = r.vector(np.uint64), r.vector(np.uint64)
        targets, csum = r.vector(np.uint64), r.vector(np.uint64)
        labels = r.strings()
        ce, ce_rows, ce_cols = r.bit_matrix()           # paths_by_chrom_copy_and_edge: rows = edges, cols = chromosome copies
        ec, ec_rows, ec_cols = r.bit_matrix()           # paths_by_edge_and_chrom_copy: rows = chromosome copies, cols = edges
        names = r.strings()
        ploidy = r.vector(np.uint32)
    finally:
        if f is not src:
            f.close()
    if not (alnpos.size == refpos.size and csum.size == refpos.size + 1 and len(labels) == targets.size and ploidy.size == len(names) + 1):
        raise ArchiveError("variant graph archive: array sizes do not agree")
    if (ce_rows, ce_cols) != (ec_cols, ec_rows):
        raise ArchiveError("variant graph archive: the two path matrices are not transposes in shape")
    chars = "".join(labels).encode()
    offs = np.concatenate([[0], np.cumsum([len(s.encode()) for s in labels])]).astype(np.uint64)
    return VariantGraph(refpos, alnpos, targets, csum, offs, chars, ec if ec.size else np.zeros(0, np.uint64), ce if ce.size else np.zeros(0, np.uint64),
                        int(ec_rows), int(ec_cols), names, ploidy)


def save_variant_graph(dst, g: VariantGraph, paths_by_chrom_copy_and_edge=None):
    """paths_by_chrom_copy_and_edge: the copy-major words if g does not carry them (e.g. DeviceGraph.paths_by_chrom_copy_and_edge())."""
    ce = g.paths_by_chrom_copy_and_edge if paths_by_chrom_copy_and_edge is None else paths_by_chrom_copy_and_edge
    if ce is None or g.paths_by_edge_and_chrom_copy is None:
        raise ArchiveError("both path matrices are needed to write a graph archive")
    f = open(dst, "xb") if isinstance(dst, (str, bytes)) else dst          # 'x': the reference refuses to overwrite
    try:
        w = _Writer(f)
        w.scalar("I", 0)
        w.vector(g.reference_positions, np.uint64)
        w.vector(g.aligned_positions, np.uint64)
        w.vector(g.alt_edge_targets, np.uint64)
        w.vector(g.alt_edge_count_csum, np.uint64)
        off = g.label_offsets
        w.strings([g.label_chars[int(off[i]) : int(off[i + 1])].decode() for i in range(len(off) - 1)])
        w.bit_matrix(ce, g.e_pad if g.h_pad else 0, g.h_pad if g.e_pad else 0)
        w.bit_matrix(g.paths_by_edge_and_chrom_copy, g.h_pad if g.e_pad else 0, g.e_pad if g.h_pad else 0)
        w.strings(list(g.sample_names))
        w.vector(g.ploidy_csum, np.uint32)
    finally:
        if f is not dst:
            f.close()


def load_cut_positions(src):
    """→ (cut positions, minimum distance, score)."""
    f = open(src, "rb") if isinstance(src, (str, bytes)) else src
    try:
        r = _Reader(f)
        r.scalar("I")
        md = r.scalar("Q")
        cuts = r.vector(np.uint64)
        score = r.scalar("I")
    finally:
        if f is not src:
            f.close()
    return cuts, int(md), int(score)


def save_cut_positions(dst, cuts, min_distance: int, score: int):
    f = open(dst, "xb") if isinstance(dst, (str, bytes)) else dst
    try:
        w = _Writer(f)
        w.scalar("I", 0)
        w.scalar("Q", int(min_distance))
        w.vector(cuts, np.uint64)
        w.scalar("I", int(score))
    finally:
        if f is not dst:
            f.close()
