"""Python mirror of the reference's output interfaces for the materialisation path, on top of libv2m_cuda.so:

    transpose_matrix                  include/vcf2multialign/transpose_matrix.hh:14
    output_delegate                   include/vcf2multialign/output.hh:26-38
    haplotype_output                  include/vcf2multialign/output.hh:71-78, libvcf2multialign/haplotype_output.cc
    founder_sequence_greedy_output    include/vcf2multialign/output.hh:81-130, libvcf2multialign/founder_sequence_greedy_output.cc

Same names, argument meaning and error behaviour (dst files are created exclusively, output.cc:41 / file_handling.cc:81-94).
The authoritative drop-in is the C++ adapter layer in vcf2multialign_b200/host/; this module exists so that the parity
tests read like the reference's own tests.
"""
from __future__ import annotations

import os

import numpy as np

from ._lib import PLOIDY_MAX, V2MError, check, lib, ptr
from .graph import DeviceGraph, VariantGraph


def transpose_matrix(words: np.ndarray, rows: int, cols: int) -> np.ndarray:
    """v2m::transpose_matrix on the word array of a rows × cols lb::bit_matrix; returns the words of the cols × rows result."""
    words = np.ascontiguousarray(words, dtype=np.uint64)
    if rows % 64 or cols % 64:
        raise V2MError("transpose_matrix: dimensions must be multiples of 64 (transpose_matrix.cc:53-54)")
    if words.size != rows * cols // 64:
        raise V2MError("transpose_matrix: word count does not match the dimensions")
    if cols == 0:
        return np.zeros(0, np.uint64)
    dst = np.zeros(words.size, np.uint64)
    check(lib().v2m_cuda_transpose_bits(ptr(words), rows, cols, ptr(dst)))
    return dst


class output_delegate:
    """output.hh:26-38; all hooks are called synchronously, in row order, from the calling thread."""

    def will_handle_sample(self, sample: str, sample_idx: int, chr_copy_idx: int): pass
    def will_handle_founder_sequence(self, idx: int): pass
    def handled_sequences(self, sequence_count: int): pass
    def handled_node(self, node: int): pass


class output:
    """output.hh:41-68 (no --pipe in the Python mirror; the C++ adapter has it)."""

    def __init__(self, chromosome_id, should_output_reference: bool, should_output_unaligned: bool, delegate: output_delegate | None = None):
        self.m_chromosome_id = chromosome_id
        self.m_should_output_reference = bool(should_output_reference)
        self.m_should_output_unaligned = bool(should_output_unaligned)
        self.m_delegate = delegate or output_delegate()

    # -- helpers shared by both subclasses --
    def _rows(self, dgraph: DeviceGraph):
        """→ (plan, [(a2m id, separate-file stem, before-hook, after-hook)…])"""
        raise NotImplementedError

    def _a2m_id(self, tail: str) -> str:
        return (self.m_chromosome_id + "\t" if self.m_chromosome_id else "") + tail

    def _file_name(self, stem: str, should_include_fasta_header: bool) -> str:
        name = (self.m_chromosome_id + "." if self.m_chromosome_id else "") + stem
        if should_include_fasta_header:
            name += ".fa" if self.m_should_output_unaligned else ".a2m"
        return name

    def output_a2m(self, ref_seq, graph, dst):
        """output_a2m(ref_seq, graph, stream | dst_name) — haplotype_output.cc:38-82 / founder_…cc:515-550 / output.cc:47-76.
        `graph` may be a VariantGraph (uploaded here) or an already resident DeviceGraph."""
        dgraph = graph if isinstance(graph, DeviceGraph) else DeviceGraph(graph, ref_seq)
        plan, rows = self._rows(dgraph)
        own_fd = isinstance(dst, (str, bytes, os.PathLike))
        fd = os.open(dst, os.O_RDWR | os.O_CREAT | os.O_EXCL, 0o644) if own_fd else dst.fileno()
        if not own_fd:
            dst.flush()
        try:
            # will_handle_* fires before each row and handled_sequences after it, as in the reference loop; the rows
            # of one pinned block are produced together, so "before" hooks of a block run when its first row is due.
            state = {"next": 0}

            def fire_before(upto):
                while state["next"] <= upto and state["next"] < len(rows):
                    before = rows[state["next"]][2]
                    if before:
                        before()
                    state["next"] += 1

            def row_done(r):
                fire_before(r)
                after = rows[r][3]
                if after:
                    after()

            plan.write_a2m([r[0] for r in rows], fd, self.m_should_output_unaligned, row_done)
        finally:
            if own_fd:
                os.close(fd)
            plan.close()

    def output_separate(self, ref_seq, graph, should_include_fasta_header: bool, dst_dir: str = "."):
        """One file per row; the file name doubles as the FASTA identifier whatever should_include_fasta_header says, and
        there is no trailing newline (output.cc:36-42 quirk)."""
        dgraph = graph if isinstance(graph, DeviceGraph) else DeviceGraph(graph, ref_seq)
        plan, rows = self._rows(dgraph)
        try:
            n = plan.row_count()
            block = max(1, min(n, (64 << 20) // max(dgraph.aligned_length(), 1)))
            for r0 in range(0, n, block):
                r1 = min(n, r0 + block)
                if self.m_should_output_unaligned:
                    seqs = plan.emit_rows_unaligned(r0, r1)
                else:
                    arr = plan.emit_rows(r0, r1)
                    seqs = [arr[i].tobytes() for i in range(r1 - r0)]
                for r, seq in zip(range(r0, r1), seqs):
                    _, stem, before, _after = rows[r]
                    if before:
                        before()
                    name = self._file_name(stem, should_include_fasta_header)
                    fd = os.open(os.path.join(dst_dir, name), os.O_RDWR | os.O_CREAT | os.O_EXCL, 0o644)
                    with os.fdopen(fd, "wb") as f:
                        f.write(b">" + name.encode() + b"\n" + seq)
        finally:
            plan.close()


class haplotype_output(output):
    def _row_table(self, g: VariantGraph):
        """→ (chromosome copy index per row, [(a2m id, separate-file stem, before-hook, after-hook)…]); no device involved."""
        d = self.m_delegate
        copy_idx, rows = [], []
        if self.m_should_output_reference:
            copy_idx.append(PLOIDY_MAX)
            rows.append((self._a2m_id("REF"), "REF", None, lambda: d.handled_sequences(1)))
        k = 0
        for s, name in enumerate(g.sample_names):
            for c in range(g.sample_ploidy(s)):
                copy_idx.append(int(g.ploidy_csum[s]) + c)
                # haplotype_output.cc:46, 78-79: seq_count starts at 1 even with --omit-reference, so the k-th sample row reports k + 2
                rows.append((self._a2m_id(f"{name}-{1 + c}"), f"{name}.{1 + c}",
                             (lambda name=name, s=s, c=c: d.will_handle_sample(name, s, c)), (lambda k=k: d.handled_sequences(k + 2))))
                k += 1
        return copy_idx, rows

    def _rows(self, dgraph: DeviceGraph):
        copy_idx, rows = self._row_table(dgraph.graph)
        return dgraph.plan_haplotypes(np.array(copy_idx, np.uint32)), rows


class founder_sequence_greedy_output(output):
    """output.hh:81-130. Cut positions come from find_cut_positions (K4 + host recurrence, find_cut_positions.cc:93-211) or are
    supplied the way --input-cut-positions supplies them (set_cut_positions ~ load_cut_positions)."""

    def __init__(self, chromosome_id, should_output_reference: bool, should_keep_ref_edges: bool, should_output_unaligned: bool,
                 delegate: output_delegate | None = None):
        super().__init__(chromosome_id, should_output_reference, should_output_unaligned, delegate)
        self.m_should_keep_ref_edges = bool(should_keep_ref_edges)
        self.m_cut_positions = np.zeros(0, np.uint64)
        self.m_assigned_samples = None                    # segments × founders (logical indexing)
        self.m_min_distance = 0
        self.m_score = 0

    def cut_positions(self):
        return self.m_cut_positions

    def set_cut_positions(self, cuts):
        self.m_cut_positions = np.ascontiguousarray(cuts, dtype=np.uint64)

    def assigned_samples(self):
        return self.m_assigned_samples

    def max_segmentation_height(self):
        return self.m_score

    def find_cut_positions(self, graph, minimum_distance: int, ref_seq=None) -> bool:
        """output.hh:123, founder_sequence_greedy_output.cc:139-152. `graph`: DeviceGraph, or VariantGraph + ref_seq."""
        dgraph = graph if isinstance(graph, DeviceGraph) else DeviceGraph(graph, ref_seq)
        cuts, score = dgraph.find_cut_positions(minimum_distance)
        if cuts is None:
            return False
        self.m_cut_positions = cuts
        self.m_min_distance = int(minimum_distance)
        self.m_score = score
        return True

    def find_matchings(self, graph, founder_count: int, ref_seq=None) -> bool:
        """founder_sequence_greedy_output.cc:154-512. `graph`: DeviceGraph, or VariantGraph + ref_seq."""
        dgraph = graph if isinstance(graph, DeviceGraph) else DeviceGraph(graph, ref_seq)
        m = dgraph.find_matchings(self.m_cut_positions, founder_count, self.m_should_keep_ref_edges)
        if m is None:
            return False
        self.m_assigned_samples = m
        return True

    def _rows(self, dgraph: DeviceGraph):
        d = self.m_delegate
        m = self.m_assigned_samples
        n_founders = m.shape[1]
        rows = [(self._a2m_id(str(1 + f)), str(1 + f), (lambda f=f: d.will_handle_founder_sequence(f)), (lambda f=f: d.handled_sequences(2 + f)))
                for f in range(n_founders)]
        if self.m_should_output_reference:
            # The REF row goes first (founder_…cc:519-531): a founder column assigned PLOIDY_MAX everywhere follows REF edges only.
            m = np.concatenate([np.full((m.shape[0], 1), PLOIDY_MAX, np.uint32), m], axis=1)
            rows = [(self._a2m_id("REF"), "REF", None, lambda: d.handled_sequences(1))] + rows
        return dgraph.plan_founders(self.m_cut_positions, m), rows
