"""Seeded synthetic workloads of the shapes BASELINE.json names: random ACGT reference, SNV / insertion /
deletion sites with some tri-allelic ones, and LD-structured haplotypes (mosaic of 64 founder haplotypes, 1/f-like
allele frequencies, sparse private mutations). The variant_graph arrays are produced directly — the VCF text of a
5,008-haplotype chromosome would be tens of GB — following the node / aligned-position rules of build_variant_graph
(libvcf2multialign/variant_graph.cc:157-179, 299-366, 437-443); `to_vcf_text` writes the same data as VCF + FASTA for
small shapes so that the direct construction can be checked against a graph built from text.

Path bits are a pure function of (seed, chromosome copy, site), evaluated either by numpy (`path_bits_numpy`, small
shapes) or on the device by v2m_cuda_synth_paths (same function, see csrc/synth.cu)."""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass

import numpy as np

from ._lib import check, lib, ptr
from .graph import VariantGraph

BASES = np.frombuffer(b"ACGT", dtype=np.uint8)
N_FOUNDERS = 64
BLOCK_SHIFT = 13          # founder switches every 8192 sites (≈1.2e-4 per site)


@dataclass
class Workload:
    name: str
    graph: VariantGraph
    ref: np.ndarray            # uint8
    seed: int
    edge_site: np.ndarray      # uint32 [E]
    edge_alt: np.ndarray       # uint8  [E] 0-based ALT number within the site
    site_pos: np.ndarray       # 0-based reference positions
    site_ref_len: np.ndarray
    site_alts: list | None     # only kept for small shapes (VCF text)


SHAPES = {
    # name: (ref_len, n_sites, n_samples)
    "c1": (1_000_000, 2_000, 10),
    "c2": (51_000_000, 1_100_000, 2_504),
    "c4": (249_000_000, 6_000_000, 2_504),
    "tiny": (20_000, 600, 6),
    "small": (300_000, 9_000, 40),
}


def _hash32(x):
    x = np.asarray(x, dtype=np.uint64) & np.uint64(0xFFFFFFFF)
    x ^= x >> np.uint64(16)
    x = (x * np.uint64(0x7FEB352D)) & np.uint64(0xFFFFFFFF)
    x ^= x >> np.uint64(15)
    x = (x * np.uint64(0x846CA68B)) & np.uint64(0xFFFFFFFF)
    x ^= x >> np.uint64(16)
    return x


def _mix(seed, a, b):
    """hash32(hash32(a ^ seed) + b) on uint32 lanes — mirrored by mix() in csrc/synth.cu."""
    return _hash32((_hash32(np.asarray(a, np.uint64) ^ np.uint64(seed & 0xFFFFFFFF)) + np.asarray(b, np.uint64)) & np.uint64(0xFFFFFFFF))


def allele_numpy(seed, h, site, site_n_alts):
    """Allele (0 = REF, 1…) of chromosome copy h at site `site` (arrays broadcast)."""
    h = np.asarray(h, np.uint64)
    site = np.asarray(site, np.uint64)
    thr = np.uint64(65536) >> (np.uint64(2) + _mix(seed, site, 11) % np.uint64(10))          # site frequency 2^-2 … 2^-11
    off = _mix(seed, h, 23) & np.uint64((1 << BLOCK_SHIFT) - 1)
    block = (site + off) >> np.uint64(BLOCK_SHIFT)
    founder = _mix(seed, h * np.uint64(0x9E3779B1) + block, 37) & np.uint64(N_FOUNDERS - 1)
    fh = _mix(seed, site * np.uint64(N_FOUNDERS) + founder, 41)
    nonref = (fh & np.uint64(0xFFFF)) < thr
    alt = np.where(np.asarray(site_n_alts) > 1, (fh >> np.uint64(16)) & np.uint64(1), 0).astype(np.uint64)
    mut = (_mix(seed, h * np.uint64(0x85EBCA6B) + site, 53) & np.uint64(1023)) == 0
    nonref2 = nonref ^ mut
    alt = np.where(mut & ~nonref, 0, alt)
    return np.where(nonref2, alt + np.uint64(1), 0).astype(np.uint32)


def path_bits_numpy(w: Workload) -> np.ndarray:
    """paths_by_edge_and_chrom_copy words (edge-major lines) for small shapes."""
    g = w.graph
    H, E = g.total_chromosome_copies(), g.edge_count()
    Wh = g.h_pad // 64
    n_alts = np.bincount(w.edge_site, minlength=len(w.site_pos)).astype(np.uint32)
    words = np.zeros(g.e_pad * Wh, np.uint64)
    hs = np.arange(H, dtype=np.uint64)
    for e in range(E):
        s = int(w.edge_site[e])
        al = allele_numpy(w.seed, hs, np.uint64(s), n_alts[s])
        bits = (al == np.uint32(1 + int(w.edge_alt[e])))
        idx = np.nonzero(bits)[0]
        np.bitwise_or.at(words, e * Wh + idx // 64, np.uint64(1) << (idx % 64).astype(np.uint64))
    return words


def path_bits_device(w: Workload, d_dst) -> None:
    """Fills device memory d_dst (e_pad × h_pad/64 words) with paths_by_edge_and_chrom_copy (K-synth)."""
    g = w.graph
    n_alts = np.bincount(w.edge_site, minlength=len(w.site_pos)).astype(np.uint8)
    edge_nalts = np.ascontiguousarray(n_alts[w.edge_site])
    check(lib().v2m_cuda_synth_paths(d_dst, g.h_pad, g.e_pad, g.total_chromosome_copies(), g.edge_count(), w.seed & 0xFFFFFFFF,
                                      ptr(np.ascontiguousarray(w.edge_site, np.uint32)), ptr(np.ascontiguousarray(w.edge_alt, np.uint8)), ptr(edge_nalts)))


def make_workload(name: str, seed: int = 1, ref_len=None, n_sites=None, n_samples=None, keep_alts=None, with_paths="auto") -> Workload:
    L, S, n_samp = SHAPES[name] if name in SHAPES else (ref_len, n_sites, n_samples)
    L, S, n_samp = ref_len or L, n_sites or S, n_samples or n_samp
    rng = np.random.default_rng(seed)
    ref = BASES[rng.integers(0, 4, L, dtype=np.uint8)]

    # Sites: kind 0 = SNV, 1 = insertion, 2 = deletion; 5 % get a second ALT.
    kind = rng.choice(3, S, p=[0.85, 0.075, 0.075]).astype(np.uint8)
    indel = np.minimum(rng.geometric(0.35, S), 10).astype(np.int64)
    ref_len_site = np.where(kind == 2, 1 + indel, 1).astype(np.int64)
    tri = rng.random(S) < 0.05
    # Non-overlapping REF spans: gap ≥ previous REF length, the rest geometric so that the sites fill the reference.
    slack = L - 40 - int(ref_len_site.sum())
    if slack <= S:
        raise ValueError("too many sites for this reference length")
    extra = rng.geometric(min(1.0, S / slack), S).astype(np.int64)
    pos = np.cumsum(np.concatenate([[1], ref_len_site[:-1]]) + extra)
    if pos[-1] + ref_len_site[-1] >= L - 1:                       # rescale the random part if it overshot
        scale = (L - 40 - ref_len_site.sum()) / extra.sum()
        extra = np.maximum(1, (extra * scale * 0.98).astype(np.int64))
        pos = np.cumsum(np.concatenate([[1], ref_len_site[:-1]]) + extra)
    assert pos[-1] + ref_len_site[-1] < L

    # ALT labels, flattened. alt 0 per site, plus alt 1 for tri-allelic sites.
    ins_len = np.where(kind == 1, indel, 0)
    alt_len0 = (1 + ins_len).astype(np.int64)                     # SNV: 1, insertion: 1 + k, deletion: 1
    indel2 = np.where(indel < 10, indel + 1, indel - 1)
    alt_len1 = np.where(kind == 1, 1 + indel2, 1).astype(np.int64)
    n_alts = 1 + tri.astype(np.int64)
    E = int(n_alts.sum())
    edge_site = np.repeat(np.arange(S, dtype=np.uint32), n_alts)
    first_edge = np.concatenate([[0], np.cumsum(n_alts)[:-1]])
    edge_alt = (np.arange(E) - first_edge[edge_site]).astype(np.uint8)
    edge_len = np.where(edge_alt == 0, alt_len0[edge_site], alt_len1[edge_site]).astype(np.int64)
    label_offsets = np.concatenate([[0], np.cumsum(edge_len)]).astype(np.uint64)
    labels = BASES[rng.integers(0, 4, int(edge_len.sum()), dtype=np.uint8)]
    # First label base: SNV → a base different from REF (and from the other ALT); indels → the REF base.
    ref_base_code = np.searchsorted(BASES, ref[pos[edge_site]])
    first = np.where(kind[edge_site] == 0, BASES[(ref_base_code + 1 + edge_alt) % 4], ref[pos[edge_site]])
    labels[label_offsets[:-1].astype(np.int64)] = first
    # a tri-allelic deletion's second ALT would equal the first ("same single base"): make it a 2-base label instead
    # (still shorter than or equal to the REF span since deletions have ref_len ≥ 2).
    # (handled through alt_len1 == 1 for deletions: labels differ only if bases differ; give it the next base)
    second_del = (kind[edge_site] == 2) & (edge_alt == 1)
    labels[label_offsets[:-1].astype(np.int64)[second_del]] = BASES[(ref_base_code[second_del] + 1) % 4]

    # Nodes (variant_graph.cc:299-308, 157-179, 437-443): one per distinct position among 0, site starts, site ends, L.
    ends = pos + ref_len_site
    node_pos = np.unique(np.concatenate([[0], pos, ends, [L]])).astype(np.int64)
    N = node_pos.size
    src_node = np.searchsorted(node_pos, pos)
    tgt_node = np.searchsorted(node_pos, ends)
    # Aligned positions: REF distance everywhere, plus at each site's target node the excess of its longest ALT label
    # over the REF span (aln = max(aln + dist, aln_src + |alt|)); spans do not overlap, so the excesses simply add up.
    max_alt = np.where(tri, np.maximum(alt_len0, alt_len1), alt_len0)
    excess = np.maximum(0, max_alt - ref_len_site)
    bump = np.zeros(N, np.int64)
    np.add.at(bump, tgt_node, excess)
    aln = node_pos + np.cumsum(bump)
    csum = np.zeros(N + 1, np.uint64)
    np.add.at(csum, src_node + 1, n_alts.astype(np.uint64))
    csum = np.cumsum(csum).astype(np.uint64)

    H = 2 * n_samp
    h_pad = (H + 63) // 64 * 64
    e_pad = (E + 63) // 64 * 64
    g = VariantGraph(
        node_pos.astype(np.uint64), aln.astype(np.uint64), tgt_node[edge_site].astype(np.uint64), csum, label_offsets, labels.tobytes(),
        None, None, h_pad, e_pad, [f"HG{i:05d}" for i in range(n_samp)], (2 * np.arange(n_samp + 1)).astype(np.uint32))
    keep = (S <= 20000) if keep_alts is None else keep_alts
    w = Workload(name, g, ref, seed, edge_site, edge_alt, pos, ref_len_site, None)
    if keep:
        lab = labels.tobytes()
        w.site_alts = [[lab[int(label_offsets[e]) : int(label_offsets[e + 1])].decode() for e in range(int(first_edge[s]), int(first_edge[s] + n_alts[s]))] for s in range(S)]
    if with_paths is True or (with_paths == "auto" and E * H <= 40_000_000):
        g.paths_by_edge_and_chrom_copy = path_bits_numpy(w)
    return w


def to_vcf_text(w: Workload, chrom: str = "1"):
    """(fasta_text, vcf_text) of a small workload; genotypes phased diploid."""
    assert w.site_alts is not None
    g = w.graph
    H = g.total_chromosome_copies()
    ref = w.ref.tobytes().decode()
    fasta = f">{chrom}\n" + "\n".join(ref[i : i + 80] for i in range(0, len(ref), 80)) + "\n"
    out = ["##fileformat=VCFv4.2", '##FORMAT=<ID=GT,Number=1,Type=String,Description="Genotype">',
           "#CHROM\tPOS\tID\tREF\tALT\tQUAL\tFILTER\tINFO\tFORMAT\t" + "\t".join(g.sample_names)]
    hs = np.arange(H, dtype=np.uint64)
    for s, alts in enumerate(w.site_alts):
        al = allele_numpy(w.seed, hs, np.uint64(s), len(alts))
        p, rl = int(w.site_pos[s]), int(w.site_ref_len[s])
        gts = "\t".join(f"{al[2 * i]}|{al[2 * i + 1]}" for i in range(H // 2))
        out.append(f"{chrom}\t{p + 1}\ts{s}\t{ref[p : p + rl]}\t{','.join(alts)}\t.\tPASS\t.\tGT\t{gts}")
    return fasta, "\n".join(out) + "\n"
