"""The synthetic workload generator (vcf2multialign_b200/synth.py) builds variant_graph arrays directly; at small
shapes the same data written as VCF + FASTA text and pushed through the oracle's restated build_variant_graph must give
the identical graph (nodes, aligned positions, edges, labels, path bits)."""
import numpy as np
import pytest

from oracle import oracle as orc
from vcf2multialign_b200 import synth


@pytest.mark.parametrize("seed,n_sites", [(1, 600), (2, 900), (3, 50)])
def test_direct_graph_equals_graph_built_from_text(tmp_path, seed, n_sites):
    w = synth.make_workload("tiny", seed=seed, n_sites=n_sites)
    fasta, vcf = synth.to_vcf_text(w)
    (tmp_path / "s.fa").write_text(fasta)
    (tmp_path / "s.vcf").write_text(vcf)
    og = orc.Graph.from_files(str(tmp_path / "s.fa"), str(tmp_path / "s.vcf"), "1")
    a = og.arrays()
    g = w.graph
    assert np.array_equal(a["reference_positions"], g.reference_positions)
    assert np.array_equal(a["aligned_positions"], g.aligned_positions)
    assert np.array_equal(a["alt_edge_targets"], g.alt_edge_targets)
    assert np.array_equal(a["alt_edge_count_csum"], g.alt_edge_count_csum)
    assert np.array_equal(a["label_offsets"], g.label_offsets)
    assert a["label_chars"] == g.label_chars
    assert (a["h_pad"], a["e_pad"]) == (g.h_pad, g.e_pad)
    assert np.array_equal(a["paths_ec"], g.paths_by_edge_and_chrom_copy)
    assert a["sample_names"] == g.sample_names
    assert np.array_equal(a["ploidy_csum"], g.ploidy_csum)
    assert og.overlaps() == []
    assert a["ref"].tobytes() == w.ref.tobytes()


def test_workload_statistics():
    w = synth.make_workload("small", seed=1)
    g = w.graph
    H, E = g.total_chromosome_copies(), g.edge_count()
    bits = np.unpackbits(g.paths_by_edge_and_chrom_copy.view(np.uint8), bitorder="little").reshape(g.e_pad, g.h_pad)[:E, :H]
    density = bits.mean()
    assert 0.01 < density < 0.15, density
    # LD structure: far fewer distinct haplotype patterns over a short window than haplotypes
    window = bits[:64].T
    assert len({r.tobytes() for r in window}) < H
    assert int(g.aligned_positions[-1]) > int(g.reference_positions[-1])


def test_cpu_mirror_of_the_path_bit_generator():
    """bench.py's CPU legs make the path lines of sampled chromosome copies with oracle.synth_path_lines (C++ mirror of
    csrc/synth.cu); it must agree with the numpy model the small shapes are built with."""
    from oracle import oracle as orc
    w = synth.make_workload("small", seed=4, with_paths=True)
    g = w.graph
    H, We = g.total_chromosome_copies(), g.e_pad // 64
    n_alts = np.bincount(w.edge_site, minlength=len(w.site_pos)).astype(np.uint8)
    ce = orc.transpose(g.paths_by_edge_and_chrom_copy, g.h_pad, g.e_pad)
    got = orc.synth_path_lines(w.seed, 0, H, g.e_pad, w.edge_site, w.edge_alt, n_alts[w.edge_site], threads=3)
    assert np.array_equal(got, ce[: H * We])
    part = orc.synth_path_lines(w.seed, 7, 5, g.e_pad, w.edge_site, w.edge_alt, n_alts[w.edge_site], threads=1)
    assert np.array_equal(part, ce[7 * We : 12 * We])
