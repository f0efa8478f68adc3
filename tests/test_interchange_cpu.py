"""Graph / cut-position interchange against archives written by the real cereal headers of the reference
checkout (tests/golden/cereal/, made by tests/golden/make_cereal_golden.sh → oracle/cereal_golden.cc). Host only."""
import io
import os
import subprocess

import numpy as np
import pytest

from oracle import oracle as orc

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CASES = ["founder-sequences/test-1", "founder-sequences/test-2", "founder-sequences/test-3", "founder-sequences/test-4",
         "variant-graph/test-2", "variant-graph/test-3", "variant-graph/test-4"]


def _paths(golden_dir, case):
    name = case.replace("/", "_")
    return (os.path.join(golden_dir, case + ".fa"), os.path.join(golden_dir, case + ".vcf"),
            os.path.join(golden_dir, "cereal", name + ".graph"), os.path.join(golden_dir, "cereal", name + ".cuts"))


@pytest.fixture(scope="module")
def cereal_test_binary():
    subprocess.run(["make", "-C", os.path.join(ROOT, "tests", "host"), "_build/cereal_test"], check=True, capture_output=True)
    return os.path.join(ROOT, "tests", "host", "_build", "cereal_test")


@pytest.mark.parametrize("case", CASES)
def test_cxx_archives_match_cereal(cereal_test_binary, golden_dir, case):
    fa, vcf, graph, cuts = _paths(golden_dir, case)
    r = subprocess.run([cereal_test_binary, fa, vcf, "0", graph, cuts], capture_output=True, text=True)
    assert r.returncode == 0 and "PASSED" in r.stdout, r.stdout + r.stderr


def test_cxx_cut_positions_with_minimum_distance(cereal_test_binary, golden_dir):
    fa, vcf, graph, _ = _paths(golden_dir, "founder-sequences/test-1")
    cuts = os.path.join(golden_dir, "cereal", "founder-sequences_test-1.md3.cuts")
    r = subprocess.run([cereal_test_binary, fa, vcf, "3", graph, cuts], capture_output=True, text=True)
    assert r.returncode == 0 and "PASSED" in r.stdout, r.stdout + r.stderr


@pytest.mark.parametrize("case", CASES)
def test_python_archives_match_cereal(golden_dir, case):
    from vcf2multialign_b200 import interchange as ic
    fa, vcf, graph_p, cuts_p = _paths(golden_dir, case)
    og = orc.Graph.from_files(fa, vcf, "1")
    a = og.arrays()
    g = ic.load_variant_graph(graph_p)
    for k in ("reference_positions", "aligned_positions", "alt_edge_targets", "alt_edge_count_csum"):
        assert np.array_equal(getattr(g, k), a[k]), k
    assert np.array_equal(g.label_offsets, a["label_offsets"]) and g.label_chars == a["label_chars"]
    assert np.array_equal(g.paths_by_edge_and_chrom_copy, a["paths_ec"]) and np.array_equal(g.paths_by_chrom_copy_and_edge, a["paths_ce"])
    assert (g.h_pad, g.e_pad) == (a["h_pad"], a["e_pad"]) and g.sample_names == a["sample_names"]
    assert np.array_equal(g.ploidy_csum, a["ploidy_csum"])
    buf = io.BytesIO()
    ic.save_variant_graph(buf, g)
    assert buf.getvalue() == open(graph_p, "rb").read()

    f = orc.FounderOutput(og)
    assert f.find_cut_positions(0)
    cuts, md, score = ic.load_cut_positions(cuts_p)
    assert cuts.tolist() == f.cut_positions().tolist() and md == 0 and score == f.score()
    buf = io.BytesIO()
    ic.save_cut_positions(buf, cuts, md, score)
    assert buf.getvalue() == open(cuts_p, "rb").read()


def test_python_rejects_broken_archives(golden_dir, tmp_path):
    from vcf2multialign_b200 import interchange as ic
    _, _, graph_p, cuts_p = _paths(golden_dir, "founder-sequences/test-1")
    data = open(graph_p, "rb").read()
    with pytest.raises(ic.ArchiveError):
        ic.load_variant_graph(io.BytesIO(data[: len(data) // 2]))
    with pytest.raises(ic.ArchiveError):
        ic.load_variant_graph(io.BytesIO(b"\x07" + data[1:]))
    dst = tmp_path / "c.cuts"
    ic.save_cut_positions(str(dst), [0, 3], 5, 2)
    with pytest.raises(FileExistsError):
        ic.save_cut_positions(str(dst), [0, 3], 5, 2)          # the reference opens with CREATE: an existing file is an error
    assert ic.load_cut_positions(str(dst))[1:] == (5, 2)


@pytest.mark.skipif(not os.path.exists("/root/reference/lib/cereal/include/cereal/cereal.hpp"), reason="reference checkout (cereal headers) not present")
def test_goldens_are_what_cereal_writes_now(golden_dir, tmp_path):
    """Regenerates one golden pair with the real cereal headers and compares with the committed files."""
    subprocess.run(["make", "-C", os.path.join(ROOT, "oracle"), "_ref/cereal_golden"], check=True, capture_output=True)
    fa, vcf, graph_p, cuts_p = _paths(golden_dir, "founder-sequences/test-4")
    g2, c2 = str(tmp_path / "g"), str(tmp_path / "c")
    subprocess.run([os.path.join(ROOT, "oracle", "_ref", "cereal_golden"), fa, vcf, "0", g2, c2], check=True)
    assert open(g2, "rb").read() == open(graph_p, "rb").read() and open(c2, "rb").read() == open(cuts_p, "rb").read()
