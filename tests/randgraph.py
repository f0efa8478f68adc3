"""Random small VCF + FASTA inputs for parity tests (test infrastructure). The text goes through the oracle's
restated builder (oracle/v2m_oracle.hh: build_variant_graph), so graphs contain everything the reference builder can
produce: multi-allelic sites, insertions, deletions, <DEL>, unhandled SV ALTs, '*', '.', null alleles, overlapping
ALTs within one chromosome copy, several records at one position."""
from __future__ import annotations

import os

import numpy as np

from oracle import oracle as orc

BASES = "ACGT"


def write_random_inputs(dirpath, seed, ref_len=200, n_records=30, n_samples=5, ploidy=2, overlap=True, sv=True,
                        max_indel=6, alt_freq=0.3, long_del_prob=0.0, chrom="1", many_alts_prob=0.0):
    rng = np.random.default_rng(seed)
    ref = "".join(rng.choice(list(BASES), ref_len))
    fa = os.path.join(dirpath, f"r{seed}.fa")
    vcf = os.path.join(dirpath, f"r{seed}.vcf")
    with open(fa, "w") as f:
        f.write(f">{chrom} test\n")
        for i in range(0, ref_len, 60):
            f.write(ref[i : i + 60] + "\n")
    positions = np.sort(rng.integers(0, ref_len, n_records))
    if not overlap:
        positions = np.unique(positions)
    lines = ["##fileformat=VCFv4.2", '##FORMAT=<ID=GT,Number=1,Type=String,Description="Genotype">',
             "#CHROM\tPOS\tID\tREF\tALT\tQUAL\tFILTER\tINFO\tFORMAT\t" + "\t".join(f"S{i}" for i in range(n_samples))]
    prev_end = 0
    for k, pos in enumerate(positions):
        pos = int(pos)
        if not overlap and pos < prev_end:
            continue
        kind = rng.random()
        if rng.random() < long_del_prob:
            ref_span = int(rng.integers(20, 80))
        elif kind < 0.55:
            ref_span = 1
        else:
            ref_span = int(rng.integers(1, max_indel + 1))
        ref_span = max(1, min(ref_span, ref_len - pos))
        ref_allele = ref[pos : pos + ref_span]
        n_alts = 1 + int(rng.random() < 0.3) + int(rng.random() < 0.1)
        if many_alts_prob and rng.random() < many_alts_prob:
            n_alts = int(rng.integers(5, 10))
        alts = []
        for _ in range(n_alts):
            r = rng.random()
            if sv and r < 0.06:
                alts.append("<DEL>")
            elif sv and r < 0.09:
                alts.append(rng.choice(["<INS>", "<CNV:TEST>", "<INV>", "<DUP>"]))
            elif sv and r < 0.11:
                alts.append("*")
            elif sv and r < 0.12 and n_alts == 1:
                alts.append(".")
            else:
                alt_len = 1 if kind < 0.55 and r < 0.8 else int(rng.integers(1, max_indel + 1))
                a = "".join(rng.choice(list(BASES), alt_len))
                if a == ref_allele:
                    a = ("A" if ref_allele[0] != "A" else "C") + a[1:]
                alts.append(a)
        gts = []
        for s in range(n_samples):
            parts = []
            for c in range(ploidy):
                r = rng.random()
                if r < alt_freq:
                    parts.append(str(int(rng.integers(1, len(alts) + 1))))
                elif r < alt_freq + 0.02:
                    parts.append(".")
                else:
                    parts.append("0")
            gts.append(("|" if rng.random() < 0.8 else "/").join(parts))
        lines.append(f"{chrom}\t{pos + 1}\tv{k}\t{ref_allele}\t{','.join(alts)}\t.\tPASS\t.\tGT\t" + "\t".join(gts))
        prev_end = pos + ref_span
    with open(vcf, "w") as f:
        f.write("\n".join(lines) + "\n")
    return fa, vcf


def random_oracle_graph(dirpath, seed, **kw):
    fa, vcf = write_random_inputs(dirpath, seed, **kw)
    return orc.Graph.from_files(fa, vcf, kw.get("chrom", "1"))


# ---- numpy models of the data-parallel reformulation (SURVEY §7 C1–C6, appendix A) ------------------------------
def bit(words, line_words, line, i):
    return (int(words[line * line_words + i // 64]) >> (i % 64)) & 1


def model_taken_edges(a, h):
    """Appendix A.2: edges the walk of chromosome copy h takes."""
    E = len(a["alt_edge_targets"])
    We = a["e_pad"] // 64
    csum = a["alt_edge_count_csum"]
    src = np.repeat(np.arange(len(csum) - 1), np.diff(csum).astype(np.int64))
    cur, taken = 0, []
    for e in range(E):
        if bit(a["paths_ce"], We, h, e) and src[e] >= cur:
            taken.append(e)
            cur = int(a["alt_edge_targets"][e])
    return taken


def model_template(a):
    """Appendix A.3 with no taken edges: the REF row."""
    ref = a["ref"].tobytes()
    rp, ap = a["reference_positions"], a["aligned_positions"]
    out = bytearray()
    for n in range(len(rp) - 1):
        piece = ref[int(rp[n]) : int(rp[n + 1])]
        out += piece + b"-" * (int(ap[n + 1]) - int(ap[n]) - len(piece))
    return bytes(out)


def model_row(a, taken, template=None):
    """Appendix A.3: template patched by the taken edges."""
    row = bytearray(template if template is not None else model_template(a))
    csum = a["alt_edge_count_csum"]
    src = np.repeat(np.arange(len(csum) - 1), np.diff(csum).astype(np.int64))
    ap = a["aligned_positions"]
    off, ch = a["label_offsets"], a["label_chars"]
    for e in taken:
        a0, a1 = int(ap[src[e]]), int(ap[int(a["alt_edge_targets"][e])])
        label = ch[int(off[e]) : int(off[e + 1])]
        row[a0:a1] = label + b"-" * (a1 - a0 - len(label))
    return bytes(row)


def model_edge_classes(a):
    """K2a's static edge classes as graph_upload computes them: (simple, sibling) bit words (uint64 per 64 edges)."""
    csum = np.asarray(a["alt_edge_count_csum"], np.int64)
    tgt = np.asarray(a["alt_edge_targets"], np.int64)
    We = a["e_pad"] // 64
    simple = [(1 << 64) - 1] * We
    sibling = [0] * We
    reach = 0
    for n in range(len(csum) - 1):
        lb, rb = int(csum[n]), int(csum[n + 1])
        for e in range(lb, rb):
            k = e - lb
            if reach <= n and 1 <= k <= 3 and e % 64 >= k:
                sibling[e // 64] |= 1 << (e % 64)
            elif reach > n or k >= 1:
                simple[e // 64] &= ~(1 << (e % 64))
        for e in range(lb, rb):
            reach = max(reach, int(tgt[e]))
    return simple, sibling


def model_taken_edges_by_classes(a, h):
    """The walk as resolve_rows_kernel runs it, word by word: simple and sibling edges by bit arithmetic, complex edges by the
    walk's rule against the target of the last taken edge (the taken edge with the largest index)."""
    We = a["e_pad"] // 64
    M = (1 << 64) - 1
    csum = a["alt_edge_count_csum"]
    src = np.repeat(np.arange(len(csum) - 1), np.diff(csum).astype(np.int64))
    tgt = a["alt_edge_targets"]
    simple, sibling = model_edge_classes(a)
    taken, last = [], -1
    for w in range(We):
        S = int(a["paths_ce"][h * We + w])
        m1 = sibling[w]
        m2 = m1 & (m1 << 1) & M
        m3 = m2 & (m1 << 2) & M
        fast = S & ~((((S << 1) & m1) | ((S << 2) & m2) | ((S << 3) & m3)) & M)
        tk = fast
        if S & ~simple[w] & M:
            tk, cur = 0, last
            for b in range(64):
                if not (S >> b) & 1:
                    continue
                e = 64 * w + b
                if (simple[w] >> b) & 1:
                    take = bool((fast >> b) & 1)
                else:
                    take = int(src[e]) >= (int(tgt[cur]) if cur >= 0 else 0)
                if take:
                    tk |= 1 << b
                    cur = e
        for b in range(64):
            if (tk >> b) & 1:
                taken.append(64 * w + b)
                last = 64 * w + b
    return taken
