"""Pins the CPU oracle (oracle/v2m_oracle.hh) against the reference's own golden vectors:
tests/variant_graph.cc:249-338, tests/founder_sequences.cc:118-188, tests/transpose_matrix.cc:188-279."""
import os

import numpy as np
import pytest

from oracle import oracle as orc


def _set_bit(words, rows, y, x):
    i = x * rows + y
    words[i // 64] |= np.uint64(1) << np.uint64(i % 64)


def test_transpose8x8_forms_agree():
    rng = np.random.default_rng(0)
    for w in [0, 1, 1 << 63, 0x8040201008040201, 0xFFFFFFFFFFFFFFFF] + [int(x) for x in rng.integers(0, 2**63, 200, dtype=np.uint64)]:
        assert orc.transpose8x8(w) == orc.transpose8x8(w, terms=True)
        assert orc.transpose8x8(orc.transpose8x8(w)) == w
    # bit (8r + c) ↔ bit (8c + r)
    for r in range(8):
        for c in range(8):
            assert orc.transpose8x8(1 << (8 * r + c)) == 1 << (8 * c + r)


def test_transpose_known_answers(expected):
    for kat in expected["transpose_kat"]:
        rows, cols = kat["rows"], kat["cols"]
        y, x = kat["set"]
        src = np.zeros(rows * cols // 64, np.uint64)
        exp = np.zeros(rows * cols // 64, np.uint64)
        _set_bit(src, rows, y, x)
        _set_bit(exp, cols, x, y)
        assert np.array_equal(orc.transpose(src, rows, cols), exp)


@pytest.mark.parametrize("seed", range(12))
def test_transpose_property(seed):
    """tests/transpose_matrix.cc:254-279: random word dims, count/3 unique set positions."""
    rng = np.random.default_rng(seed)
    h, w = (int(v) for v in rng.integers(1, 9, 2))
    rows, cols = 64 * h, 64 * w
    n = rows * cols
    pos = rng.choice(n, n // 3, replace=False)
    y, x = pos // cols, pos % cols
    src = np.zeros(n // 64, np.uint64)
    exp = np.zeros(n // 64, np.uint64)
    np.bitwise_or.at(src, (x * rows + y) // 64, np.uint64(1) << ((x * rows + y) % 64).astype(np.uint64))
    np.bitwise_or.at(exp, (y * cols + x) // 64, np.uint64(1) << ((y * cols + x) % 64).astype(np.uint64))
    got = orc.transpose(src, rows, cols)
    assert np.array_equal(got, exp)
    assert np.array_equal(orc.transpose(got, cols, rows), src)


def test_transpose_empty():
    assert orc.transpose(np.zeros(0, np.uint64), 64, 0).size == 0


def test_variant_graph_goldens(expected, golden_dir):
    for case in expected["variant_graph"]:
        d = os.path.join(golden_dir, "variant-graph")
        g = orc.Graph.from_files(os.path.join(d, case["fa"]), os.path.join(d, case["vcf"]), "1")
        a = g.arrays()
        labels = g.labels()
        ref = a["ref"].tobytes().decode()
        nodes = case["nodes"]
        assert len(a["reference_positions"]) == len(nodes), case["vcf"]
        for node, pos, aln, ref_label, edges in nodes:
            assert int(a["reference_positions"][node]) == pos
            assert int(a["aligned_positions"][node]) == aln
            if node + 1 < len(nodes):
                assert ref[pos : int(a["reference_positions"][node + 1])] == ref_label
            lb, rb = int(a["alt_edge_count_csum"][node]), int(a["alt_edge_count_csum"][node + 1])
            got = [[int(a["alt_edge_targets"][e]), labels[e]] for e in range(lb, rb)]
            assert got == edges, (case["vcf"], node)
        got_ov = [{k: o[k] for k in ("sample", "chrom_copy", "ref_pos", "var_id", "gt")} for o in g.overlaps()]
        for o in got_ov:
            assert o in case["overlaps"]
        # the two path matrices are transposes of each other and padded to ×64
        assert a["h_pad"] % 64 == 0 and a["e_pad"] % 64 == 0
        assert np.array_equal(orc.transpose(a["paths_ec"], a["h_pad"], a["e_pad"]), a["paths_ce"])


def test_founder_goldens(expected, golden_dir):
    for case in expected["founder_sequences"]:
        d = os.path.join(golden_dir, "founder-sequences")
        g = orc.Graph.from_files(os.path.join(d, case["fa"]), os.path.join(d, case["vcf"]), "1")
        f = orc.FounderOutput(g, None, True, False, False)
        assert f.find_cut_positions(0)
        assert f.cut_positions().tolist() == case["cuts"], case
        assert f.find_matchings(2)
        m = f.assigned_samples()
        rows = case["rows"]
        exp = np.array(case["assigned_colmajor"], np.uint32).reshape(-1, rows).T
        assert np.array_equal(m, exp), (case["vcf"], m, exp)
        assert f.output_a2m().decode() == case["a2m"], case["vcf"]


def test_survey_worked_values(golden_dir):
    """SURVEY.md §8c worked values for founder test-1 (second cut)."""
    d = os.path.join(golden_dir, "founder-sequences")
    g = orc.Graph.from_files(os.path.join(d, "test-1.fa"), os.path.join(d, "test-1.vcf"), "1")
    a = g.arrays()
    assert a["reference_positions"].tolist() == [0, 3, 5, 8, 10, 13]
    assert a["aligned_positions"].tolist() == [0, 4, 6, 10, 12, 16]
    assert a["alt_edge_count_csum"].tolist() == [0, 3, 3, 6, 6, 9, 9]
    assert a["alt_edge_targets"].tolist() == [1, 1, 1, 3, 3, 3, 5, 5, 5]
    f = orc.FounderOutput(g)
    assert f.find_cut_positions(0) and f.find_matchings(2)
    rec = f.recorded_classes()
    assert rec[1]["rhs_eq"].tolist() == [6, 6, 6, 3, 4, 5, 6, 6, 6, 3, 6, 6, 6, 5]
    assert [tuple(r) for r in rec[1]["joined"].tolist()] == [(3, 6, 3), (0, 6, 3), (1, 6, 1), (2, 6, 2), (3, 3, 1), (0, 3, 1), (3, 4, 1), (3, 5, 2)]


def test_haplotype_output_matches_rows(golden_dir):
    d = os.path.join(golden_dir, "variant-graph")
    g = orc.Graph.from_files(os.path.join(d, "test-4.fa"), os.path.join(d, "test-4.vcf"), "1")
    a = g.arrays()
    text = g.haplotype_output_a2m("chrT", True, False)
    lines = text.split(b"\n")
    assert lines[-1] == b""
    assert lines[0] == b">chrT\tREF"
    A = int(a["aligned_positions"][-1])
    H = a["total_copies"]
    assert len(lines) == 2 * (H + 1) + 1
    for i in range(H + 1):
        assert len(lines[2 * i + 1]) == A
    assert lines[2] == b">chrT\tSAMPLE1-1"
    assert lines[1] == g.output_sequence(orc.PLOIDY_MAX)
    assert lines[3] == g.output_sequence(0)
    # unaligned rows are the aligned ones without gaps
    for h in range(H):
        assert g.output_sequence(h, None, True) == g.output_sequence(h).replace(b"-", b"")
    files = g.haplotype_output_separate("chrT", True, False, True)
    assert files[0][0] == "chrT.REF.a2m" and files[0][1] == b">chrT.REF.a2m\n" + lines[1]
    assert files[1][0] == "chrT.SAMPLE1.1.a2m"
    files = g.haplotype_output_separate(None, False, True, False)
    assert files[0][0] == "SAMPLE1.1" and files[0][1].startswith(b">SAMPLE1.1\n")
