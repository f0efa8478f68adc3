"""CPU property tests of the data-parallel reformulation the kernels implement (SURVEY §7 claims C1–C3):
template + sparse patches with the 'src >= last target' rule reproduce the oracle's graph walk byte for byte."""
import numpy as np
import pytest

from oracle import oracle as orc
from tests import randgraph as rg


@pytest.mark.parametrize("seed", range(40))
def test_rows_equal_template_plus_patches(tmp_path, seed):
    g = rg.random_oracle_graph(str(tmp_path), seed, ref_len=150 + 7 * seed, n_records=10 + seed, n_samples=4,
                               ploidy=1 + seed % 3, long_del_prob=0.05 if seed % 4 == 0 else 0.0)
    a = g.arrays()
    tmpl = rg.model_template(a)
    assert tmpl == g.output_sequence(orc.PLOIDY_MAX)
    assert len(tmpl) == int(a["aligned_positions"][-1])
    for h in range(a["total_copies"]):
        taken = rg.model_taken_edges(a, h)
        assert rg.model_row(a, taken, tmpl) == g.output_sequence(h), (seed, h)


def test_overlaps_are_exercised(tmp_path):
    n = 0
    for seed in range(40):
        g = rg.random_oracle_graph(str(tmp_path), seed, ref_len=150 + 7 * seed, n_records=10 + seed, n_samples=4, ploidy=1 + seed % 3)
        n += len(g.overlaps())
    assert n > 20


@pytest.mark.parametrize("seed", range(36))
def test_edge_classes_need_no_walk_state(tmp_path, seed):
    """K2a (resolve_rows_kernel) decides set bits of simple and sibling edges without loading anything: check its word
    arithmetic and the complex-edge fallback against the walk, on graphs with many multi-allelic sites and overlaps."""
    g = rg.random_oracle_graph(str(tmp_path), 100 + seed, ref_len=200 + 9 * seed, n_records=20 + 3 * seed, n_samples=5,
                               ploidy=1 + seed % 2, long_del_prob=0.08 if seed % 2 else 0.0, many_alts_prob=0.3 if seed % 3 == 0 else 0.0,
                               alt_freq=0.5 if seed % 4 == 0 else 0.3)
    a = g.arrays()
    simple, sibling = rg.model_edge_classes(a)
    E = len(a["alt_edge_targets"])
    assert any(simple) and (seed % 3 != 0 or any(sibling))
    for h in range(a["total_copies"]):
        assert rg.model_taken_edges_by_classes(a, h) == rg.model_taken_edges(a, h), (seed, h)
