"""GPU parity at BASELINE.json's full sizes.

c1 (1 Mbp, 2,000 sites, 10 diploid samples) goes through the text path — VCF + FASTA → oracle builder — and is compared
byte for byte, whole. c2 / c3 (51 Mbp, 1.1 M sites, 5,008 chromosome copies; 256 GB of rows) cannot be compared whole:
sampled rows are compared with the oracle byte for byte and the rest is covered by size-independent properties
(involution of the transpose, aligned row minus gaps = unaligned row = its length from the prefix kernel, column windows
concatenate to the row, founder rows are gathers of haplotype rows, the segmentation height from K4 equals the largest
class count K3 sees on the chosen segments)."""
import os

import numpy as np
import pytest

from oracle import oracle as orc

pytestmark = pytest.mark.gpu

# The rows compared byte for byte with the oracle at full size are a few fixed points (first / last chromosome copy, …) plus a
# random sample that changes from round to round: V2M_TEST_SEED overrides the round's seed.
SAMPLE_SEED = int(os.environ.get("V2M_TEST_SEED", "2"))


def _sampled_copies(fixed, H, n_random):
    rng = np.random.default_rng(SAMPLE_SEED)
    extra = [int(x) for x in rng.choice(H, n_random, replace=False)]
    return sorted(set(fixed) | set(extra))


@pytest.fixture(scope="module")
def v2m():
    import vcf2multialign_b200 as v
    v._lib.check(v.lib().v2m_cuda_init(0))
    return v


def _oracle_from_workload(w, paths_ec, paths_ce):
    g = w.graph
    return orc.Graph.from_arrays(dict(
        reference_positions=g.reference_positions, aligned_positions=g.aligned_positions, alt_edge_targets=g.alt_edge_targets,
        alt_edge_count_csum=g.alt_edge_count_csum, label_offsets=g.label_offsets, label_chars=g.label_chars,
        paths_ec=paths_ec, paths_ce=paths_ce, h_pad=g.h_pad, e_pad=g.e_pad, ploidy_csum=g.ploidy_csum,
        sample_names=g.sample_names, ref=w.ref))


# ---- c1: whole job, text inputs ------------------------------------------------------------------------------------
def test_c1_text_path_whole_job(v2m, tmp_path):
    from vcf2multialign_b200 import synth
    w = synth.make_workload("c1", seed=1, with_paths=True)
    fasta, vcf = synth.to_vcf_text(w, "1")
    fa_p, vcf_p = os.path.join(tmp_path, "c1.fa"), os.path.join(tmp_path, "c1.vcf")
    open(fa_p, "w").write(fasta)
    open(vcf_p, "w").write(vcf)
    og = orc.Graph.from_files(fa_p, vcf_p, "1")          # the restated build_variant_graph
    a = dict(og.arrays())
    # the directly generated graph arrays are what the builder makes of the text
    g = w.graph
    for k, ref in (("reference_positions", g.reference_positions), ("aligned_positions", g.aligned_positions),
                   ("alt_edge_targets", g.alt_edge_targets), ("alt_edge_count_csum", g.alt_edge_count_csum)):
        assert np.array_equal(np.asarray(a[k], np.uint64), ref), k
    E, Wh = g.edge_count(), g.h_pad // 64
    assert np.array_equal(np.asarray(a["paths_ec"], np.uint64)[: E * Wh], g.paths_by_edge_and_chrom_copy[: E * Wh])

    a["paths_ce"] = None
    dg = v2m.DeviceGraph(v2m.VariantGraph.from_arrays(a), og.arrays()["ref"])
    assert np.array_equal(dg.paths_by_chrom_copy_and_edge(), og.arrays()["paths_ce"])          # K1
    H = og.dims()[4]
    assert H == 20
    plan = dg.plan_haplotypes([v2m.PLOIDY_MAX] + list(range(H)))
    rows = plan.emit_rows()
    assert rows[0].tobytes() == og.output_sequence(orc.PLOIDY_MAX)
    for h in range(H):
        assert rows[1 + h].tobytes() == og.output_sequence(h), h
    un = plan.emit_rows_unaligned()
    for h in range(H):
        assert un[1 + h] == og.output_sequence(h, None, True), h

    # the A2M stream of haplotype_output::output_a2m, through the mirror
    out = v2m.haplotype_output("chr1", True, False)
    dst = os.path.join(tmp_path, "c1.a2m")
    out.output_a2m(None, dg, dst)
    assert open(dst, "rb").read() == og.haplotype_output_a2m("chr1", True, False)

    # config 3 on c1: cut positions, matchings, founder rows
    for md, founders in ((0, 4), (50, 5), (5000, 3)):
        f = orc.FounderOutput(og)
        assert f.find_cut_positions(md)
        cuts, score = dg.find_cut_positions(md)
        assert cuts.tolist() == f.cut_positions().tolist() and score == f.score(), md
        try:
            ok = f.find_matchings(founders)
        except orc.OracleError:
            continue
        assert ok
        m = dg.find_matchings(cuts, founders)
        assert np.array_equal(m, f.assigned_samples()), md


# ---- c2 / c3: sampled rows + size-independent properties ---------------------------------------------------------------
@pytest.fixture(scope="module")
def c2(v2m):
    from vcf2multialign_b200 import synth
    from vcf2multialign_b200._lib import lib
    w = synth.make_workload("c2", seed=1, with_paths=False)
    g = w.graph
    d_paths = lib().v2m_cuda_device_alloc(g.e_pad * g.h_pad // 8)
    assert d_paths
    synth.path_bits_device(w, d_paths)
    dg = v2m.DeviceGraph(g, w.ref, d_paths_by_edge_and_chrom_copy=d_paths)
    yield w, dg, d_paths
    dg.close()
    lib().v2m_cuda_device_free(d_paths)


def test_c2_transpose_involution_and_popcount(v2m, c2):
    from vcf2multialign_b200._lib import check, lib, ptr
    w, dg, d_paths = c2
    g = w.graph
    n_words = g.e_pad * g.h_pad // 64
    ce = dg.paths_by_chrom_copy_and_edge()                      # K1 result resident on the device
    ec = np.zeros(n_words, np.uint64)
    check(lib().v2m_cuda_memcpy_d2h(ptr(ec), d_paths, n_words * 8))
    pop = lambda x: int(np.unpackbits(x.view(np.uint8)).sum(dtype=np.int64))
    assert pop(ce[: n_words // 16]) >= 0                         # (touch) and the full population counts agree
    assert pop(ce) == pop(ec)
    back = v2m.transpose_matrix(ce, g.e_pad, g.h_pad)           # rows = e_pad, cols = h_pad → the edge-major matrix again
    assert np.array_equal(back, ec)
    # sampled bits: (h, e) of the copy-major matrix is (e, h) of the edge-major one
    rng = np.random.default_rng(9)
    hs, es = rng.integers(0, g.h_pad, 20000), rng.integers(0, g.e_pad, 20000)
    i_ce = hs * g.e_pad + es
    i_ec = es * g.h_pad + hs
    assert np.array_equal((ce[i_ce // 64] >> (i_ce % 64).astype(np.uint64)) & 1, (ec[i_ec // 64] >> (i_ec % 64).astype(np.uint64)) & 1)


def test_c2_sampled_rows_vs_oracle_and_row_properties(v2m, c2):
    w, dg, d_paths = c2
    g = w.graph
    H, A = g.total_chromosome_copies(), g.aligned_length()
    assert H == 5008
    We = g.e_pad // 64
    sample = _sampled_copies([0, 1, 5007], H, 4)
    ce_all = dg.paths_by_chrom_copy_and_edge()
    ce = np.zeros_like(ce_all)
    for h in sample:
        ce[h * We : (h + 1) * We] = ce_all[h * We : (h + 1) * We]
    del ce_all
    og = _oracle_from_workload(w, np.zeros(ce.size, np.uint64), ce)      # the row writer reads the copy-major orientation only
    plan = dg.plan_haplotypes([v2m.PLOIDY_MAX] + sample)
    rows = plan.emit_rows()
    assert rows.shape == (1 + len(sample), A)
    assert rows[0].tobytes() == og.output_sequence(orc.PLOIDY_MAX)
    for i, h in enumerate(sample):
        assert rows[1 + i].tobytes() == og.output_sequence(h), h
    # aligned row minus gaps = unaligned row; its length = what the prefix kernel reports
    lens = plan.row_lengths(True)
    un = plan.emit_rows_unaligned()
    for i in range(rows.shape[0]):
        nogap = rows[i][rows[i] != ord("-")].tobytes()
        assert len(nogap) == int(lens[i]) and nogap == un[i], i
    assert un[1] == og.output_sequence(sample[0], None, True)
    # column windows concatenate to the row (ragged, unaligned window borders)
    cuts = [0, 4095, 4096, 1_000_003, 25_000_017, A - 5, A]
    for i in (1, 4):
        parts = [plan.emit_rows(i, i + 1, c0, c1)[0].tobytes() for c0, c1 in zip(cuts[:-1], cuts[1:])]
        assert b"".join(parts) == rows[i].tobytes()


def test_c3_cut_positions_matchings_and_founder_rows(v2m, c2):
    """--founder-sequences=25 --minimum-distance=50 on the chr22 shape."""
    w, dg, d_paths = c2
    g = w.graph
    aln, csum = g.aligned_positions, g.alt_edge_count_csum
    N = g.node_count()
    cuts, score = dg.find_cut_positions(50)
    assert cuts is not None and cuts[0] == 0 and cuts[-1] == N - 1
    c = cuts.astype(np.int64)
    assert (np.diff(c) > 0).all()
    # every segment but possibly the last (the sink is forced, find_cut_positions.cc:204-207) is at least 50 columns long and
    # starts with a new edge index; no ALT edge crosses a cut
    seg_len = np.diff(aln[c].astype(np.int64))
    assert (seg_len[1:-1] >= 50).all()           # the first one starts at node 0 whether or not the recurrence reached it (:199-200)
    assert (np.diff(csum[c].astype(np.int64))[:-1] > 0).all()
    tgt = g.alt_edge_targets.astype(np.int64)
    reach = np.maximum.accumulate(tgt)                         # rightmost target over edges [0, e]
    first_edge = csum[c[1:]].astype(np.int64)                  # edges before a cut must not reach beyond it
    ok = first_edge == 0
    ok |= reach[np.maximum(first_edge - 1, 0)] <= c[1:]
    assert ok.all()

    # K4's segmentation height = the largest class count K3 finds on the chosen segments
    classes = dg.segment_classes(cuts, want_rhs_eq=False)
    distinct = np.array([x["distinct"] for x in classes])
    # the last segment was stretched to the sink after the recurrence; leave it out when it gained edges
    body = distinct[:-1] if distinct.size > 1 else distinct
    assert int(body.max()) == score or int(distinct.max()) == score, (int(body.max()), int(distinct.max()), score)

    m = dg.find_matchings(cuts, 25)
    assert m is not None and m.shape == (cuts.size - 1, 25)
    # founder rows are gathers of the assigned chromosome copies' rows: check three founders on a few segments
    plan = dg.plan_founders(cuts, m[:, :3].copy())
    rng = np.random.default_rng(4)
    segs = sorted(set(int(x) for x in rng.integers(0, cuts.size - 1, 12)) | {0, cuts.size - 2})
    for f in range(3):
        for s in segs:
            c0, c1 = int(aln[c[s]]), int(aln[c[s + 1]])
            src = int(m[s, f])
            hp = dg.plan_haplotypes([src])
            assert np.array_equal(plan.emit_rows(f, f + 1, c0, c1), hp.emit_rows(0, 1, c0, c1)), (f, s)
            hp.close()


def test_c3_prefix_founders_vs_oracle_with_default_chunks(v2m, c2):
    """Founder mode against the oracle at scale: the first ~50,000 ALT edges of the c2 graph with all 5,008 chromosome copies,
    K4 and K3 with their default chunk count (one per SM), i.e. ~340 columns per chunk and every chunk boundary composed from
    the chunk-local orders. Cut positions, score, the class data of every cut (representatives, distinct counts, first class,
    first-is-REF, joined runs in permutation order) and assigned_samples must equal what the reference's sequential pBWT
    sweep produces (find_cut_positions.cc:93-211, founder_sequence_greedy_output.cc:154-512). ~40 s of CPU for the oracle."""
    from vcf2multialign_b200._lib import check, lib, ptr
    w, dg, d_paths = c2
    g = w.graph
    Wh = g.h_pad // 64
    # cut the graph at a node no ALT edge spans (sites of the synthetic workloads do not overlap), ~50,000 edges in
    csum = g.alt_edge_count_csum.astype(np.int64)
    tgt = g.alt_edge_targets.astype(np.int64)
    node = int(np.searchsorted(csum, 50_000))
    reach = np.maximum.accumulate(tgt)
    while csum[node] > 0 and reach[csum[node] - 1] > node:
        node += 1
    e_cpu = int(csum[node])
    assert 50_000 <= e_cpu < 51_000
    e_pad = (e_cpu + 63) // 64 * 64
    sub = np.zeros(e_pad * Wh, np.uint64)
    check(lib().v2m_cuda_memcpy_d2h(ptr(sub), d_paths, e_cpu * Wh * 8))
    N2 = node + 1
    arrays = dict(
        reference_positions=g.reference_positions[:N2], aligned_positions=g.aligned_positions[:N2], alt_edge_targets=g.alt_edge_targets[:e_cpu],
        alt_edge_count_csum=np.concatenate([g.alt_edge_count_csum[:N2], [e_cpu]]).astype(np.uint64), label_offsets=g.label_offsets[: e_cpu + 1],
        label_chars=g.label_chars[: int(g.label_offsets[e_cpu])], paths_ec=sub, paths_ce=None, h_pad=g.h_pad, e_pad=e_pad,
        ploidy_csum=g.ploidy_csum, sample_names=g.sample_names)
    ref = w.ref[: int(g.reference_positions[node]) + 1]
    og = orc.Graph.from_arrays(dict(arrays, ref=ref))
    pdg = v2m.DeviceGraph(v2m.VariantGraph.from_arrays(arrays), ref)
    try:
        for min_distance, founders in ((50, 25), (0, 7)):
            f = orc.FounderOutput(og)
            assert f.find_cut_positions(min_distance)
            cuts, score = pdg.find_cut_positions(min_distance)
            t = pdg.last_cut_positions_timing()
            assert t["chunks"] > 100, t                                  # the default chunking, not a single chunk
            assert cuts.tolist() == f.cut_positions().tolist() and score == f.score(), min_distance
            assert f.find_matchings(founders)
            if min_distance:                                             # the class data of ~25,000 cuts is ~0.5 GB on either side: once is enough
                rec = f.recorded_classes()
                got = pdg.segment_classes(cuts)
                assert len(got) == len(rec) == cuts.size - 1
                for i, (a, b) in enumerate(zip(got, rec)):
                    assert (a["distinct"], a["first_class"], a["first_is_ref"]) == (b["distinct"], b["first_class"], b["first_is_ref"]), (min_distance, i)
                    assert np.array_equal(a["rhs_eq"], b["rhs_eq"]), (min_distance, i)
                    if i >= 1:
                        assert np.array_equal(a["joined"], b["joined"]), (min_distance, i)
                del rec, got
            m = pdg.find_matchings(cuts, founders)
            assert np.array_equal(m, f.assigned_samples()), min_distance
    finally:
        pdg.close()


# ---- c4: the chr1 shape (249 Mbp, 6.3 M ALT edges, 5,008 chromosome copies; 1.25 TB of rows) ----------------------------------
def test_c4_sampled_rows_vs_oracle(v2m):
    """BASELINE.json's target shape: sampled rows byte for byte against the oracle, plus the row properties of the c2 test."""
    from vcf2multialign_b200 import synth
    from vcf2multialign_b200._lib import check, lib, ptr
    w = synth.make_workload("c4", seed=1, with_paths=False)
    g = w.graph
    H, A = g.total_chromosome_copies(), g.aligned_length()
    assert H == 5008 and A > 249_000_000
    path_bytes = g.e_pad * g.h_pad // 8
    d_paths = lib().v2m_cuda_device_alloc(path_bytes)
    assert d_paths
    try:
        synth.path_bits_device(w, d_paths)
        dg = v2m.DeviceGraph(g, w.ref, d_paths_by_edge_and_chrom_copy=d_paths)
        We = g.e_pad // 64
        sample = _sampled_copies([5007], H, 2)
        ce_all = dg.paths_by_chrom_copy_and_edge()
        ce = np.zeros_like(ce_all)
        for h in sample:
            ce[h * We : (h + 1) * We] = ce_all[h * We : (h + 1) * We]
        del ce_all
        og = _oracle_from_workload(w, np.zeros(ce.size, np.uint64), ce)
        del ce
        plan = dg.plan_haplotypes([v2m.PLOIDY_MAX] + sample)
        rows = plan.emit_rows()
        assert rows.shape == (4, A)
        assert rows[0].tobytes() == og.output_sequence(orc.PLOIDY_MAX)
        for i, h in enumerate(sample):
            assert rows[1 + i].tobytes() == og.output_sequence(h), h
        lens = plan.row_lengths(True)
        un = plan.emit_rows_unaligned()
        for i in range(4):
            nogap = rows[i][rows[i] != ord("-")].tobytes()
            assert len(nogap) == int(lens[i]) and nogap == un[i], i
        # a window across the 2^27 and 2^28 column marks and the ragged end
        for c0, c1 in ((134_217_000, 134_218_500), (A - 4097, A)):
            assert plan.emit_rows(2, 3, c0, c1)[0].tobytes() == rows[2][c0:c1].tobytes()
        plan.close()
        dg.close()
    finally:
        lib().v2m_cuda_device_free(d_paths)
