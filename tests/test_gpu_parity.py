"""GPU parity tests: the CUDA path (through the C ABI of libv2m_cuda.so) against the CPU oracle, byte for byte."""
import os

import numpy as np
import pytest

from oracle import oracle as orc
from tests import randgraph as rg

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def v2m():
    import vcf2multialign_b200 as v
    v._lib.check(v.lib().v2m_cuda_init(0))
    return v


def _set_bit(words, rows, y, x):
    i = x * rows + y
    words[i // 64] |= np.uint64(1) << np.uint64(i % 64)


def _device_graph(v2m, og, with_ce=False):
    a = dict(og.arrays())
    if not with_ce:
        a["paths_ce"] = None            # let K1 derive it on the device, as variant_graph.cc:453 does on the host
    g = v2m.VariantGraph.from_arrays(a)
    return v2m.DeviceGraph(g, og.arrays()["ref"]), g


# ---- K1 ------------------------------------------------------------------------------------------------------
def test_transpose_known_answers(v2m, expected):
    for kat in expected["transpose_kat"]:
        rows, cols = kat["rows"], kat["cols"]
        y, x = kat["set"]
        src = np.zeros(rows * cols // 64, np.uint64)
        exp = np.zeros(rows * cols // 64, np.uint64)
        _set_bit(src, rows, y, x)
        _set_bit(exp, cols, x, y)
        assert np.array_equal(v2m.transpose_matrix(src, rows, cols), exp)


@pytest.mark.parametrize("seed", range(10))
def test_transpose_property_vs_oracle(v2m, seed):
    """tests/transpose_matrix.cc:254-279 (random word dims, a third of the bits set), checked against the oracle."""
    rng = np.random.default_rng(seed)
    h, w = (int(x) for x in rng.integers(1, 12, 2))
    rows, cols = 64 * h, 64 * w
    src = rng.integers(0, 2**63, rows * cols // 64, dtype=np.uint64) & rng.integers(0, 2**63, rows * cols // 64, dtype=np.uint64)
    got = v2m.transpose_matrix(src, rows, cols)
    assert np.array_equal(got, orc.transpose(src, rows, cols))
    assert np.array_equal(v2m.transpose_matrix(got, cols, rows), src)


# Both kernels: tall tiles are chosen for rows <= 512 or (cols / 8) % 16 != 0, the 512-line tiles otherwise.
@pytest.mark.parametrize("rows,cols", [(64, 64), (64, 4096), (4096, 64), (5056, 2112), (1088, 1600), (960, 576), (2112, 5056),
                                       (1024, 256), (5056, 2176), (576, 1152), (4096, 128), (640, 3200), (512, 2048), (448, 1280)])
def test_transpose_tile_edges(v2m, rows, cols):
    """Shapes around the 512-line × 1024-row tile borders, incl. the c2/c4 row count (5056 = 4·1024 + 960)."""
    rng = np.random.default_rng(rows * 7 + cols)
    src = rng.integers(0, 2**64 - 1, rows * cols // 64, dtype=np.uint64)
    got = v2m.transpose_matrix(src, rows, cols)
    assert np.array_equal(got, orc.transpose(src, rows, cols))


@pytest.mark.parametrize("rows,cols", [(64, 64), (64, 192), (128, 320), (192, 128), (256, 1088), (320, 64), (512, 448), (1024, 192), (2048, 704),
                                       (5056, 2112), (5056, 64), (8192, 384), (14080, 192)])
def test_transpose_more_shapes(v2m, rows, cols):
    """Ragged last tiles in both directions, the 5,008-haplotype line length, the 8,192-copy limit and beyond."""
    rng = np.random.default_rng(rows * 13 + cols)
    src = rng.integers(0, 2**64 - 1, rows * cols // 64, dtype=np.uint64)
    got = v2m.transpose_matrix(src, rows, cols)
    assert np.array_equal(got, orc.transpose(src, rows, cols))


@pytest.mark.parametrize("rows,cols", [(1024, 8192), (1088, 3264), (576, 5184), (64, 1024), (512, 64), (5056, 11328), (1280, 20480 + 448), (2112, 1024 * 7 + 64),
                                       (640, 1024 * 33 + 8 * 64 * 3), (1216, 1024 * 5 + 192)])
@pytest.mark.parametrize("kernel", ["2", "0", "1", "3", "6", "8", "9", "12"])
def test_transpose_every_kernel_on_ragged_shapes(v2m, monkeypatch, rows, cols, kernel):
    """Every K1 kernel (V2M_K1_TALL = 0: 512 x 1024 tiles, 1: tall tiles, 2: sheared, 3: wide with two sub-tiles per thread,
    6: half-size tiles, 8: persistent streaming kernel with cp.async, 9: bulk stores where lines allow them, 12: 1024 x 256 tiles) on shapes with short last bands / last tiles and destination
    lines at every 8-byte phase of a 32-byte sector."""
    monkeypatch.setenv("V2M_K1_TALL", kernel)
    rng = np.random.default_rng(rows * 7 + cols)
    src = rng.integers(0, 2**64, rows * cols // 64, dtype=np.uint64) & rng.integers(0, 2**64, rows * cols // 64, dtype=np.uint64)
    got = v2m.transpose_matrix(src, rows, cols)
    assert np.array_equal(got, orc.transpose(src, rows, cols)), (rows, cols, kernel)


def test_transpose_large_involution(v2m):
    """Bigger than the oracle finishes quickly: T(T(x)) == x plus population count and a sampled bit check."""
    rows, cols = 5056, 64 * 3000
    rng = np.random.default_rng(5)
    src = rng.integers(0, 2**64 - 1, rows * cols // 64, dtype=np.uint64) & rng.integers(0, 2**64 - 1, rows * cols // 64, dtype=np.uint64)
    t = v2m.transpose_matrix(src, rows, cols)
    assert np.array_equal(v2m.transpose_matrix(t, cols, rows), src)
    ys, xs = rng.integers(0, rows, 5000), rng.integers(0, cols, 5000)
    si = xs * rows + ys
    ti = ys * cols + xs
    assert np.array_equal((src[si // 64] >> (si % 64).astype(np.uint64)) & 1, (t[ti // 64] >> (ti % 64).astype(np.uint64)) & 1)


def test_transpose_rejects_bad_dims(v2m):
    with pytest.raises(v2m.V2MError):
        v2m.transpose_matrix(np.zeros(3, np.uint64), 96, 2)
    assert v2m.transpose_matrix(np.zeros(0, np.uint64), 64, 0).size == 0


# ---- K2 on the reference's fixtures ------------------------------------------------------------------------------
def test_rows_on_reference_fixtures(v2m, expected, golden_dir):
    for sub, key in (("variant-graph", "variant_graph"), ("founder-sequences", "founder_sequences")):
        for case in expected[key]:
            d = os.path.join(golden_dir, sub)
            og = orc.Graph.from_files(os.path.join(d, case["fa"]), os.path.join(d, case["vcf"]), "1")
            dg, g = _device_graph(v2m, og)
            a = og.arrays()
            assert np.array_equal(dg.paths_by_chrom_copy_and_edge(), a["paths_ce"])
            assert dg.template() == og.output_sequence(orc.PLOIDY_MAX)
            H = a["total_copies"]
            plan = dg.plan_haplotypes([orc.PLOIDY_MAX] + list(range(H)))
            rows = plan.emit_rows()
            assert rows[0].tobytes() == og.output_sequence(orc.PLOIDY_MAX)
            for h in range(H):
                assert rows[1 + h].tobytes() == og.output_sequence(h), (case["vcf"], h)


def test_founder_rows_on_reference_goldens(v2m, expected, golden_dir):
    """Exact A2M text of tests/founder_sequences.cc:118-188 from the GPU row writer, given the golden cuts and matchings."""
    for case in expected["founder_sequences"]:
        d = os.path.join(golden_dir, "founder-sequences")
        og = orc.Graph.from_files(os.path.join(d, case["fa"]), os.path.join(d, case["vcf"]), "1")
        dg, g = _device_graph(v2m, og)
        assigned = np.array(case["assigned_colmajor"], np.uint32).reshape(-1, case["rows"]).T
        plan = dg.plan_founders(case["cuts"], assigned)
        rows = plan.emit_rows()
        text = b">REF\n" + dg.template() + b"\n"
        for f in range(rows.shape[0]):
            text += b">" + str(1 + f).encode() + b"\n" + rows[f].tobytes() + b"\n"
        assert text.decode() == case["a2m"], case["vcf"]


# ---- K2 on random graphs -------------------------------------------------------------------------------------------
@pytest.mark.parametrize("seed", range(24))
def test_rows_random_graphs(v2m, tmp_path, seed):
    og = rg.random_oracle_graph(str(tmp_path), seed, ref_len=300 + 211 * seed, n_records=20 + 9 * seed, n_samples=3 + seed % 5,
                                ploidy=1 + seed % 3, long_del_prob=0.08 if seed % 3 == 0 else 0.0)
    a = og.arrays()
    dg, g = _device_graph(v2m, og, with_ce=bool(seed % 2))
    H = a["total_copies"]
    plan = dg.plan_haplotypes(list(range(H)) + [orc.PLOIDY_MAX])
    # K2a: taken-edge bits against the numpy model of the walk
    taken = plan.taken()
    for h in range(H):
        exp = np.zeros(a["e_pad"] // 64, np.uint64)
        for e in rg.model_taken_edges(a, h):
            exp[e // 64] |= np.uint64(1) << np.uint64(e % 64)
        assert np.array_equal(taken[h], exp), (seed, h)
    assert not taken[H].any()
    # K2b
    rows = plan.emit_rows()
    for h in range(H):
        assert rows[h].tobytes() == og.output_sequence(h), (seed, h)
    assert rows[H].tobytes() == og.output_sequence(orc.PLOIDY_MAX)


def test_rows_multi_tile_and_column_windows(v2m, tmp_path):
    """A graph several 4096-column tiles long with long deletions that cross tile borders; arbitrary column windows."""
    og = rg.random_oracle_graph(str(tmp_path), 1234, ref_len=40000, n_records=2500, n_samples=6, ploidy=2, long_del_prob=0.04)
    a = og.arrays()
    dg, g = _device_graph(v2m, og)
    H = a["total_copies"]
    A = int(a["aligned_positions"][-1])
    assert A > 9 * 4096
    plan = dg.plan_haplotypes(list(range(H)))
    full = [og.output_sequence(h) for h in range(H)]
    rows = plan.emit_rows()
    for h in range(H):
        assert rows[h].tobytes() == full[h], h
    rng = np.random.default_rng(0)
    for _ in range(12):
        c0, c1 = sorted(int(x) for x in rng.integers(0, A + 1, 2))
        r0, r1 = sorted(int(x) for x in rng.integers(0, H + 1, 2))
        win = plan.emit_rows(r0, r1, c0, c1)
        for i, h in enumerate(range(r0, r1)):
            assert win[i].tobytes() == full[h][c0:c1], (h, c0, c1)


def test_rows_many_rows_blocks(v2m, tmp_path):
    """More rows than one CTA row block (64) so that blockIdx.y > 0 is exercised, with repeated and REF rows."""
    og = rg.random_oracle_graph(str(tmp_path), 77, ref_len=9000, n_records=700, n_samples=40, ploidy=2)
    a = og.arrays()
    dg, g = _device_graph(v2m, og)
    H = a["total_copies"]
    idx = list(range(H)) + [orc.PLOIDY_MAX] + list(range(H - 1, -1, -1)) + [3, 3, orc.PLOIDY_MAX]
    plan = dg.plan_haplotypes(idx)
    rows = plan.emit_rows()
    cache = {}
    for i, h in enumerate(idx):
        if h not in cache:
            cache[h] = og.output_sequence(h)
        assert rows[i].tobytes() == cache[h], (i, h)


def test_rows_dense_tiles_fall_back_to_global_arrays(v2m, tmp_path):
    """Tiles with more ALT edges, label bytes and taken-edge words than emit_rows stages in shared memory (256 edge records,
    2 KiB of labels, 4 words per row): the row warps then read the global arrays. Many rows, so that several row chunks and
    short last chunks occur, plus ragged column windows over the dense tiles."""
    og = rg.random_oracle_graph(str(tmp_path), 4321, ref_len=6000, n_records=2600, n_samples=45, ploidy=2, max_indel=9, alt_freq=0.12,
                                many_alts_prob=0.6, long_del_prob=0.01)
    a = og.arrays()
    A = int(a["aligned_positions"][-1])
    E = len(a["alt_edge_targets"])
    assert E > 3 * 256 * ((A + 4095) // 4096), (E, A)          # every tile is far over the edge cap on average
    assert int(a["label_offsets"][-1]) > 3 * 2048 * ((A + 4095) // 4096)
    dg, g = _device_graph(v2m, og)
    H = a["total_copies"]
    idx = list(range(H)) + [orc.PLOIDY_MAX]
    plan = dg.plan_haplotypes(idx)
    full = [og.output_sequence(h) for h in idx]
    rows = plan.emit_rows()
    for i in range(len(idx)):
        assert rows[i].tobytes() == full[i], i
    rng = np.random.default_rng(5)
    for _ in range(8):
        c0, c1 = sorted(int(x) for x in rng.integers(0, A + 1, 2))
        r0, r1 = sorted(int(x) for x in rng.integers(0, len(idx) + 1, 2))
        win = plan.emit_rows(r0, r1, c0, c1)
        for i, r in enumerate(range(r0, r1)):
            assert win[i].tobytes() == full[r][c0:c1], (r, c0, c1)


def test_rows_windows_under_16_bytes_in_last_tile(v2m):
    """Rows whose window leaves fewer than 16 bytes in its last 4096-column tile hand the TMA engine nothing for that tile
    (the bytes go out lane by lane). With more units than SMs, so that CTAs draw several units, and more than two rows per
    row warp in the last-tile unit, the row buffers are reused right after such rows: their reuse must still wait for the
    bulk store that last read them (emit_rows commits one bulk group per row, empty or not)."""
    from vcf2multialign_b200 import synth
    w = synth.make_workload("tail", seed=5, ref_len=700_000, n_sites=4000, n_samples=32, with_paths=True)
    g = w.graph
    for tail in (1, 7, 15):
        # move the sink so that A % 4096 == tail: the last node only ends the reference, nothing else refers to its position
        A0 = int(g.aligned_positions[-1])
        d = (tail - A0) % 4096
        ref = np.concatenate([w.ref, synth.BASES[np.arange(d) % 4]])
        g.reference_positions = g.reference_positions.copy(); g.reference_positions[-1] += np.uint64(d)
        g.aligned_positions = g.aligned_positions.copy(); g.aligned_positions[-1] += np.uint64(d)
        A = int(g.aligned_positions[-1])
        assert A % 4096 == tail and A // 4096 > 160
        ec = g.paths_by_edge_and_chrom_copy
        og = orc.Graph.from_arrays(dict(
            reference_positions=g.reference_positions, aligned_positions=g.aligned_positions, alt_edge_targets=g.alt_edge_targets,
            alt_edge_count_csum=g.alt_edge_count_csum, label_offsets=g.label_offsets, label_chars=g.label_chars,
            paths_ec=ec, paths_ce=orc.transpose(ec, g.h_pad, g.e_pad), h_pad=g.h_pad, e_pad=g.e_pad, ploidy_csum=g.ploidy_csum,
            sample_names=g.sample_names, ref=ref))
        dg = v2m.DeviceGraph(g, ref)
        H = g.total_chromosome_copies()
        idx = [orc.PLOIDY_MAX] + list(range(H))                 # 65 rows: two row chunks, up to 4 rows per row warp
        plan = dg.plan_haplotypes(idx)
        full = [og.output_sequence(h) for h in idx]
        for _ in range(3):                                      # the hazard is a race: give it more than one chance
            rows = plan.emit_rows()
            for i in range(len(idx)):
                assert rows[i].tobytes() == full[i], (tail, i)
        # column windows that end 1 … 15 bytes into a tile, starting in the tile before (or further back)
        for c1 in (4096 * 100 + 3, 4096 * 161 + 15, 4096 * 37 + 9):
            for c0 in (c1 - 5000, 0):
                win = plan.emit_rows(0, len(idx), c0, c1)
                for i in range(len(idx)):
                    assert win[i].tobytes() == full[i][c0:c1], (tail, i, c0, c1)
        plan.close()
        dg.close()
        w.ref = ref


def test_founder_rows_random(v2m, tmp_path):
    """Founder rows given the oracle's cut positions and matchings (K2c = compose + K2a + K2b)."""
    done = 0
    for seed in range(30):
        og = rg.random_oracle_graph(str(tmp_path), 500 + seed, ref_len=600, n_records=60, n_samples=6, ploidy=2, overlap=False, sv=False)
        f = orc.FounderOutput(og)
        if not f.find_cut_positions(0):
            continue
        try:
            if not f.find_matchings(3):
                continue
        except orc.OracleError:
            continue    # inputs on which the reference itself asserts / does not terminate (SURVEY §7 hard parts)
        dg, g = _device_graph(v2m, og)
        plan = dg.plan_founders(f.cut_positions(), f.assigned_samples())
        rows = plan.emit_rows()
        exp = f.output_a2m().split(b"\n")
        for k in range(rows.shape[0]):
            assert rows[k].tobytes() == exp[2 * (k + 1) + 1], (seed, k)
        done += 1
    assert done >= 10


# ---- whole-stream driver ---------------------------------------------------------------------------------------------
@pytest.mark.parametrize("chrom,output_ref", [(None, True), ("chr7", True), ("chr7", False)])
def test_haplotype_output_a2m_stream(v2m, tmp_path, chrom, output_ref):
    og = rg.random_oracle_graph(str(tmp_path), 4242, ref_len=5000, n_records=400, n_samples=7, ploidy=2)
    a = og.arrays()
    g = v2m.VariantGraph.from_arrays(a)
    events = []

    class delegate(v2m.output_delegate):
        def will_handle_sample(self, sample, sample_idx, chr_copy_idx): events.append(("will", sample, sample_idx, chr_copy_idx))
        def handled_sequences(self, n): events.append(("done", n))

    out = v2m.haplotype_output(chrom, output_ref, False, delegate())
    dst = str(tmp_path / f"out-{chrom}-{output_ref}.a2m")
    out.output_a2m(a["ref"], g, dst)
    with open(dst, "rb") as f:
        got = f.read()
    assert got == og.haplotype_output_a2m(chrom, output_ref, False)
    # delegate protocol: seq_count starts at 1 even without the REF row (haplotype_output.cc:46, 78-79)
    dones = [e[1] for e in events if e[0] == "done"]
    H = a["total_copies"]
    assert dones == (list(range(1, H + 2)) if output_ref else list(range(2, H + 2)))
    assert [e[1:] for e in events if e[0] == "will"] == [(n, s, c) for s, n in enumerate(a["sample_names"]) for c in range(2)]
    with pytest.raises(FileExistsError):
        out.output_a2m(a["ref"], g, dst)    # O_CREAT|O_EXCL as lb::open_file_for_writing(CREATE)


def test_write_a2m_sinks(v2m, tmp_path):
    """v2m_cuda_write_a2m on a regular file opened read-write (filled through a shared mapping by several threads, starting at
    the descriptor's position and leaving it at the end), on write-only and O_APPEND descriptors and on a pipe (sequential
    writes): the same bytes every time."""
    import threading
    og = rg.random_oracle_graph(str(tmp_path), 777, ref_len=30000, n_records=900, n_samples=9, ploidy=2, long_del_prob=0.02)
    dg, g = _device_graph(v2m, og)
    H = og.dims()[4]
    plan = dg.plan_haplotypes([v2m.PLOIDY_MAX] + list(range(H)))
    ids = ["REF"] + [f"row{h}" for h in range(H)]
    exp = b"".join(b">" + i.encode() + b"\n" + (og.output_sequence(orc.PLOIDY_MAX if k == 0 else k - 1)) + b"\n" for k, i in enumerate(ids))
    for unaligned in (False, True):
        e = exp if not unaligned else b"".join(b">" + i.encode() + b"\n" + og.output_sequence(orc.PLOIDY_MAX if k == 0 else k - 1, None, True) + b"\n" for k, i in enumerate(ids))
        # regular file, not at offset 0: read-write (mapped) and write-only (sequential)
        for mode in (os.O_RDWR, os.O_WRONLY):
            p1 = str(tmp_path / f"a{int(unaligned)}{mode}.a2m")
            fd = os.open(p1, mode | os.O_CREAT, 0o644)
            os.write(fd, b"prefix\n")
            done = []
            plan.write_a2m(ids, fd, unaligned, row_done=done.append)
            assert os.lseek(fd, 0, os.SEEK_CUR) == 7 + len(e)
            os.write(fd, b"tail")
            os.close(fd)
            assert open(p1, "rb").read() == b"prefix\n" + e + b"tail"
            assert done == list(range(H + 1))
        # overwriting the middle of a longer file keeps what follows
        p3 = str(tmp_path / f"c{int(unaligned)}.a2m")
        open(p3, "wb").write(b"x" * (len(e) + 100))
        fd = os.open(p3, os.O_RDWR)
        os.lseek(fd, 10, os.SEEK_SET)
        plan.write_a2m(ids, fd, unaligned)
        os.close(fd)
        assert open(p3, "rb").read() == b"x" * 10 + e + b"x" * 90
        # O_APPEND
        p2 = str(tmp_path / f"b{int(unaligned)}.a2m")
        open(p2, "wb").write(b"head\n")
        fd = os.open(p2, os.O_WRONLY | os.O_APPEND)
        plan.write_a2m(ids, fd, unaligned)
        os.close(fd)
        assert open(p2, "rb").read() == b"head\n" + e
        # pipe
        r, w_ = os.pipe()
        chunks = []
        t = threading.Thread(target=lambda: [chunks.append(b) for b in iter(lambda: os.read(r, 1 << 20), b"")])
        t.start()
        plan.write_a2m(ids, w_, unaligned)
        os.close(w_)
        t.join()
        os.close(r)
        assert b"".join(chunks) == e


def test_haplotype_output_separate(v2m, tmp_path):
    og = rg.random_oracle_graph(str(tmp_path), 99, ref_len=3000, n_records=200, n_samples=4, ploidy=2)
    a = og.arrays()
    g = v2m.VariantGraph.from_arrays(a)
    for include_header in (True, False):
        d = tmp_path / f"sep{int(include_header)}"
        d.mkdir()
        v2m.haplotype_output("c", True, False).output_separate(a["ref"], g, include_header, str(d))
        exp = og.haplotype_output_separate("c", True, False, include_header)
        assert sorted(os.listdir(d)) == sorted(n for n, _ in exp)
        for name, data in exp:
            assert (d / name).read_bytes() == data, name


# ---- --unaligned -----------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("seed", range(12))
def test_unaligned_rows_random_graphs(v2m, tmp_path, seed):
    """sequence_writer.cc:80 with should_output_unaligned: no gap fill; row lengths and bytes against the oracle walk."""
    og = rg.random_oracle_graph(str(tmp_path), 900 + seed, ref_len=400 + 977 * seed, n_records=30 + 60 * seed, n_samples=3 + seed % 4,
                                ploidy=1 + seed % 2, long_del_prob=0.06 if seed % 2 else 0.0)
    a = og.arrays()
    dg, g = _device_graph(v2m, og)
    H = a["total_copies"]
    idx = list(range(H)) + [orc.PLOIDY_MAX]
    plan = dg.plan_haplotypes(idx)
    exp = [og.output_sequence(h, None, True) for h in idx]
    assert [int(x) for x in plan.row_lengths(True)] == [len(e) for e in exp]
    assert [int(x) for x in plan.row_lengths(False)] == [int(a["aligned_positions"][-1])] * len(idx)
    got = plan.emit_rows_unaligned()
    for i, h in enumerate(idx):
        assert got[i] == exp[i], (seed, h)
    if H >= 3:
        part = plan.emit_rows_unaligned(1, 3)
        assert part == exp[1:3]


def test_unaligned_multi_tile(v2m, tmp_path):
    og = rg.random_oracle_graph(str(tmp_path), 4321, ref_len=30000, n_records=2000, n_samples=5, ploidy=2, long_del_prob=0.03)
    a = og.arrays()
    dg, g = _device_graph(v2m, og)
    H = a["total_copies"]
    plan = dg.plan_haplotypes(list(range(H)))
    got = plan.emit_rows_unaligned()
    for h in range(H):
        assert got[h] == og.output_sequence(h, None, True), h


def test_haplotype_output_unaligned_stream_and_separate(v2m, tmp_path):
    og = rg.random_oracle_graph(str(tmp_path), 555, ref_len=6000, n_records=500, n_samples=5, ploidy=2)
    a = og.arrays()
    g = v2m.VariantGraph.from_arrays(a)
    dst = str(tmp_path / "u.fa")
    v2m.haplotype_output("chrU", True, True).output_a2m(a["ref"], g, dst)
    with open(dst, "rb") as f:
        assert f.read() == og.haplotype_output_a2m("chrU", True, True)
    d = tmp_path / "usep"
    d.mkdir()
    v2m.haplotype_output(None, False, True).output_separate(a["ref"], g, True, str(d))
    exp = og.haplotype_output_separate(None, False, True, True)
    assert sorted(os.listdir(d)) == sorted(n for n, _ in exp)
    for name, data in exp:
        assert (d / name).read_bytes() == data, name


# ---- K3: per-cut classes and find_matchings ---------------------------------------------------------------------------
def _check_classes(v2m, og, dg, founders=2, label=""):
    """K3 against the class data the oracle's pBWT sweep records at every cut; returns the oracle FounderOutput (or None)."""
    f = orc.FounderOutput(og)
    if not f.find_cut_positions(0):
        return None
    cuts = f.cut_positions()
    oracle_failed = False
    try:
        f.find_matchings(founders)
    except orc.OracleError:
        oracle_failed = True            # the class data of every cut reached before the failure is still recorded
    rec = f.recorded_classes()
    got = dg.segment_classes(cuts)
    assert len(got) == len(cuts) - 1
    H = og.dims()[4]
    for i, r in enumerate(rec):
        g = got[i]
        assert g["distinct"] == r["distinct"], (label, i)
        assert g["first_class"] == r["first_class"], (label, i)
        assert g["first_is_ref"] == r["first_is_ref"], (label, i)
        assert np.array_equal(g["rhs_eq"], r["rhs_eq"]), (label, i)
        if i >= 1:
            assert np.array_equal(g["joined"], r["joined"]), (label, i, g["joined"], r["joined"])
        else:
            # single-segment form (founder_…cc:474-509): one run per class, lhs = PLOIDY_MAX
            assert g["joined"].shape[0] == r["distinct"] and int(g["joined"][:, 2].sum()) == H
            assert (g["joined"][:, 0] == v2m.PLOIDY_MAX).all()
            assert sorted(g["joined"][:, 1].tolist()) == sorted(set(r["rhs_eq"].tolist()))
    return None if oracle_failed else f


def test_segment_classes_on_reference_goldens(v2m, expected, golden_dir):
    for case in expected["founder_sequences"]:
        d = os.path.join(golden_dir, "founder-sequences")
        og = orc.Graph.from_files(os.path.join(d, case["fa"]), os.path.join(d, case["vcf"]), "1")
        dg, g = _device_graph(v2m, og)
        f = _check_classes(v2m, og, dg, 2, case["vcf"])
        assert f is not None and f.cut_positions().tolist() == case["cuts"]


def test_find_matchings_on_reference_goldens(v2m, expected, golden_dir):
    """cut positions → assigned_samples → A2M text of tests/founder_sequences.cc:118-188, through the reference-shaped mirror."""
    for case in expected["founder_sequences"]:
        d = os.path.join(golden_dir, "founder-sequences")
        og = orc.Graph.from_files(os.path.join(d, case["fa"]), os.path.join(d, case["vcf"]), "1")
        dg, g = _device_graph(v2m, og)
        out = v2m.founder_sequence_greedy_output(None, True, False, False)
        out.set_cut_positions(case["cuts"])
        assert out.find_matchings(dg, 2)
        exp = np.array(case["assigned_colmajor"], np.uint32).reshape(-1, case["rows"]).T
        assert np.array_equal(out.assigned_samples(), exp), (case["vcf"], out.assigned_samples(), exp)
        import tempfile
        with tempfile.TemporaryDirectory() as td:
            dst = os.path.join(td, "out.a2m")
            out.output_a2m(None, dg, dst)
            assert open(dst).read() == case["a2m"], case["vcf"]


@pytest.mark.parametrize("chunks", ["1", "2", "3", "5", None])
def test_segment_classes_random_graphs(v2m, tmp_path, chunks, monkeypatch):
    if chunks is None:
        monkeypatch.delenv("V2M_K3_CHUNKS", raising=False)
    else:
        monkeypatch.setenv("V2M_K3_CHUNKS", chunks)
    done = 0
    for seed in range(12):
        og = rg.random_oracle_graph(str(tmp_path), 900 + seed, ref_len=900, n_records=90 + 10 * seed, n_samples=4 + 3 * seed, ploidy=2,
                                    overlap=bool(seed % 2), sv=bool(seed % 3), alt_freq=0.15 + 0.03 * seed)
        dg, g = _device_graph(v2m, og)
        _check_classes(v2m, og, dg, 3, f"seed {seed} chunks {chunks}")
        done += 1
    assert done == 12


@pytest.mark.parametrize("n_samples,chunks", [(520, "4"), (1100, "6"), (2504, None), (4300, "5"), (5500, None), (8192, None), (8192, "3")])
def test_segment_classes_many_copies(v2m, n_samples, chunks, monkeypatch):
    """More chromosome copies than one warp-word per warp (KW 2, 3 and the 5,008-copy shape of c3), LD-structured bits; beyond
    8,192 copies (1000 Genomes + HGDP has ~8,400) up to the 16,384 the sweep holds: KW 10, 12 and 16."""
    from vcf2multialign_b200 import synth
    if chunks is None:
        monkeypatch.delenv("V2M_K3_CHUNKS", raising=False)
    else:
        monkeypatch.setenv("V2M_K3_CHUNKS", chunks)
    w = synth.make_workload("custom", seed=5, ref_len=120_000, n_sites=2_400, n_samples=n_samples, with_paths=True)
    g = w.graph
    og = orc.Graph.from_arrays(dict(
        reference_positions=g.reference_positions, aligned_positions=g.aligned_positions, alt_edge_targets=g.alt_edge_targets,
        alt_edge_count_csum=g.alt_edge_count_csum, label_offsets=g.label_offsets, label_chars=g.label_chars,
        paths_ec=g.paths_by_edge_and_chrom_copy, paths_ce=None, h_pad=g.h_pad, e_pad=g.e_pad, ploidy_csum=g.ploidy_csum,
        sample_names=g.sample_names, ref=w.ref))
    dg = v2m.DeviceGraph(g, w.ref)
    f = _check_classes(v2m, og, dg, 25, f"synth {n_samples}")
    assert f is not None
    m = dg.find_matchings(f.cut_positions(), 25)
    assert np.array_equal(m, f.assigned_samples())


def test_find_matchings_random_graphs(v2m, tmp_path):
    ok = failed = 0
    for seed in range(40):
        og = rg.random_oracle_graph(str(tmp_path), 1300 + seed, ref_len=700, n_records=70, n_samples=3 + seed % 9, ploidy=2,
                                    overlap=bool(seed % 2), sv=False, alt_freq=0.25)
        f = orc.FounderOutput(og, None, True, bool(seed % 4 == 0), False)
        if not f.find_cut_positions(seed % 3 * 10):
            continue
        founders = 1 + seed % 5
        dg, g = _device_graph(v2m, og)
        try:
            assert f.find_matchings(founders)
        except orc.OracleError:
            # inputs on which the reference itself asserts or does not terminate: the library reports them
            with pytest.raises(v2m.V2MError):
                dg.find_matchings(f.cut_positions(), founders, bool(seed % 4 == 0))
            failed += 1
            continue
        m = dg.find_matchings(f.cut_positions(), founders, bool(seed % 4 == 0))
        assert np.array_equal(m, f.assigned_samples()), (seed, m, f.assigned_samples())
        ok += 1
    assert ok >= 20


def test_find_matchings_degenerate(v2m, golden_dir):
    d = os.path.join(golden_dir, "founder-sequences")
    og = orc.Graph.from_files(os.path.join(d, "test-1.fa"), os.path.join(d, "test-1.vcf"), "1")
    dg, g = _device_graph(v2m, og)
    assert dg.find_matchings([0], 2) is None                    # fewer than two cut positions (founder_…cc:162-163)
    N = g.node_count()
    # a single segment: the trivial case (founder_…cc:474-509)
    f = orc.FounderOutput(og)
    f.set_cut_positions([0, N - 1])
    assert f.find_matchings(3)
    assert np.array_equal(dg.find_matchings([0, N - 1], 3), f.assigned_samples())
    with pytest.raises(v2m.V2MError):
        dg.segment_classes([0, 2, 2, N - 1])                    # cut positions must increase
    with pytest.raises(v2m.V2MError):
        dg.segment_classes([1, N - 1])                          # first cut must be node 0


# ---- K4: cut positions (find_cut_positions.cc:93-211) --------------------------------------------------------------------
def _check_cuts(og, dg, min_distance, label=""):
    f = orc.FounderOutput(og)
    ok = f.find_cut_positions(min_distance)
    cuts, score = dg.find_cut_positions(min_distance)
    if not ok:
        assert cuts is None, label
        return None
    assert cuts is not None, label
    assert cuts.tolist() == f.cut_positions().tolist(), (label, cuts.tolist(), f.cut_positions().tolist())
    assert score == f.score(), (label, score, f.score())
    return cuts


def test_cut_positions_on_reference_goldens(v2m, expected, golden_dir):
    """The cut positions of tests/founder_sequences.cc:118-188 (minimum distance 0), then the whole chain
    find_cut_positions -> find_matchings -> output_a2m through the reference-shaped mirror."""
    for case in expected["founder_sequences"]:
        d = os.path.join(golden_dir, "founder-sequences")
        og = orc.Graph.from_files(os.path.join(d, case["fa"]), os.path.join(d, case["vcf"]), "1")
        dg, g = _device_graph(v2m, og)
        cuts = _check_cuts(og, dg, 0, case["vcf"])
        assert cuts.tolist() == case["cuts"]
        out = v2m.founder_sequence_greedy_output(None, True, False, False)
        assert out.find_cut_positions(dg, 0)
        assert out.find_matchings(dg, 2)
        import tempfile
        with tempfile.TemporaryDirectory() as td:
            dst = os.path.join(td, "out.a2m")
            out.output_a2m(None, dg, dst)
            assert open(dst).read() == case["a2m"], case["vcf"]


@pytest.mark.parametrize("chunks,cap", [("1", None), ("2", "1"), ("3", "2"), ("5", "3"), (None, None), (None, "1")])
def test_cut_positions_random_graphs(v2m, tmp_path, chunks, cap, monkeypatch):
    """Seeded random graphs (overlapping deletions, multi-allelic sites), several minimum distances; odd chunk counts and tiny
    list capacities force the chunk-boundary composition, the old-index extraction and the capacity retry."""
    for name, val in (("V2M_K3_CHUNKS", chunks), ("V2M_K4_CAP", cap)):
        if val is None:
            monkeypatch.delenv(name, raising=False)
        else:
            monkeypatch.setenv(name, val)
    for seed in range(14):
        og = rg.random_oracle_graph(str(tmp_path), 2100 + seed, ref_len=900, n_records=60 + 12 * seed, n_samples=1 + 3 * seed, ploidy=1 + seed % 2,
                                    overlap=bool(seed % 2), sv=bool(seed % 3), alt_freq=0.1 + 0.03 * seed)
        dg, g = _device_graph(v2m, og)
        for md in (0, 7, 40, 400, 5000):
            _check_cuts(og, dg, md, f"seed {seed} md {md} chunks {chunks} cap {cap}")


@pytest.mark.parametrize("n_samples,chunks,cap", [(520, "4", None), (1100, "7", "4"), (2504, None, None), (4300, None, None), (6900, "7", "4"), (8192, None, None)])
def test_cut_positions_many_copies(v2m, n_samples, chunks, cap, monkeypatch):
    """KW 2, 3 and the 5,008-copy shape of c3 on LD-structured bits, and 8,600 / 13,800 / 16,384 copies (KW 10, 14, 16)."""
    from vcf2multialign_b200 import synth
    for name, val in (("V2M_K3_CHUNKS", chunks), ("V2M_K4_CAP", cap)):
        if val is None:
            monkeypatch.delenv(name, raising=False)
        else:
            monkeypatch.setenv(name, val)
    w = synth.make_workload("custom", seed=6, ref_len=120_000, n_sites=2_400, n_samples=n_samples, with_paths=True)
    g = w.graph
    og = orc.Graph.from_arrays(dict(
        reference_positions=g.reference_positions, aligned_positions=g.aligned_positions, alt_edge_targets=g.alt_edge_targets,
        alt_edge_count_csum=g.alt_edge_count_csum, label_offsets=g.label_offsets, label_chars=g.label_chars,
        paths_ec=g.paths_by_edge_and_chrom_copy, paths_ce=None, h_pad=g.h_pad, e_pad=g.e_pad, ploidy_csum=g.ploidy_csum,
        sample_names=g.sample_names, ref=w.ref))
    dg = v2m.DeviceGraph(g, w.ref)
    for md in (0, 50, 1000):
        cuts = _check_cuts(og, dg, md, f"synth {n_samples} md {md}")
        assert cuts is not None


def test_founder_mode_refuses_more_than_16384_copies(v2m):
    """The sweep keeps the state of all copies in one CTA's shared memory: 16,386 copies are refused with a clear message (row
    emission on the same graph is unaffected)."""
    from vcf2multialign_b200 import synth
    w = synth.make_workload("custom", seed=7, ref_len=20_000, n_sites=300, n_samples=8193, with_paths=True)
    dg = v2m.DeviceGraph(w.graph, w.ref)
    with pytest.raises(v2m.V2MError, match="16384"):
        dg.find_cut_positions(0)
    with pytest.raises(v2m.V2MError, match="16384"):
        dg.segment_classes(np.array([0, 5, 9], np.uint64))
    rows = dg.plan_haplotypes([0, 16385]).emit_rows()
    assert rows.shape[0] == 2


def test_cut_positions_degenerate(v2m, tmp_path):
    """No ALT edges at all: the only candidate is node 0 and the result is the sink alone (find_cut_positions.cc:199-207)."""
    og = rg.random_oracle_graph(str(tmp_path), 77, ref_len=200, n_records=12, n_samples=3, ploidy=2, alt_freq=0.0)
    dg, g = _device_graph(v2m, og)
    _check_cuts(og, dg, 0, "no edges")
    _check_cuts(og, dg, 10, "no edges md 10")


# ---- K5: path bits from parsed GT fields (variant_graph.cc:379-425) ---------------------------------------------------------
def _check_ingest(v2m, og, label=""):
    a, t = og.arrays(), og.genotype_trace()
    words, ov = v2m.paths_from_genotypes(t["gt"], t["alt_offsets"], t["edges_by_alt"], a["h_pad"], a["e_pad"], t["ref_pos"], t["edge_target_pos"])
    assert np.array_equal(words, a["paths_ec"]), label
    assert ov.tolist() == t["overlaps"].tolist(), (label, ov.tolist(), t["overlaps"].tolist())
    words2, none = v2m.paths_from_genotypes(t["gt"], t["alt_offsets"], t["edges_by_alt"], a["h_pad"], a["e_pad"])
    assert none is None and np.array_equal(words2, words)
    return len(ov)


def test_ingest_genotypes_on_reference_fixtures(v2m, expected, golden_dir):
    """The path matrices behind the node / edge tables of tests/variant_graph.cc:249-338 and the founder fixtures, rebuilt on the
    device from the GT values the (restated) builder saw."""
    import glob
    n = 0
    for vcf in sorted(glob.glob(os.path.join(golden_dir, "*", "*.vcf"))):
        fa = vcf[:-4] + ".fa"
        if not os.path.exists(fa):
            fa = os.path.join(os.path.dirname(vcf), "test-1.fa")
        og = orc.Graph.from_files(fa, vcf, "1")
        _check_ingest(v2m, og, vcf)
        n += 1
    assert n >= 8


def test_ingest_genotypes_random_graphs(v2m, tmp_path):
    """Multi-allelic sites, unhandled SV ALTs (no edge), null alleles, overlapping ALTs within a chromosome copy (reported),
    more than 1,024 chromosome copies (several warp blocks per record) and more than 4 ALTs per record."""
    overlaps = 0
    for seed in range(10):
        og = rg.random_oracle_graph(str(tmp_path), 3100 + seed, ref_len=1200, n_records=80 + 30 * seed, n_samples=2 + 70 * seed, ploidy=1 + seed % 3,
                                    overlap=bool(seed % 2), sv=bool(seed % 3 != 1), alt_freq=0.05 + 0.04 * seed, many_alts_prob=0.1)
        overlaps += _check_ingest(v2m, og, f"seed {seed}")
    assert overlaps > 0


@pytest.mark.parametrize("block_records", [None, "7", "64"])
def test_ingest_genotypes_both_orientations(v2m, tmp_path, expected, golden_dir, monkeypatch, block_records):
    """K5 with paths_by_chrom_copy_and_edge born next to paths_by_edge_and_chrom_copy (variant_graph.cc:379-425 + :453): both equal
    the oracle builder's matrices, with the record blocks forced small so that ranges of final edge columns are transposed
    block by block, word-straddling edges included."""
    if block_records:
        monkeypatch.setenv("V2M_K5_BLOCK_RECORDS", block_records)
    import glob
    graphs = []
    for vcf in sorted(glob.glob(os.path.join(golden_dir, "*", "*.vcf"))):
        fa = vcf[:-4] + ".fa"
        graphs.append(orc.Graph.from_files(fa if os.path.exists(fa) else os.path.join(os.path.dirname(vcf), "test-1.fa"), vcf, "1"))
    graphs += [rg.random_oracle_graph(str(tmp_path), 700 + i, ref_len=3000 + 500 * i, n_records=300 + 150 * i, n_samples=20 + 30 * i, ploidy=2, many_alts_prob=0.1)
               for i in range(3)]
    for k, og in enumerate(graphs):
        t = og.genotype_trace()
        a = og.arrays()
        ec, ce = v2m.paths_from_genotypes_both(t["gt"], t["alt_offsets"], t["edges_by_alt"], a["h_pad"], a["e_pad"])
        assert np.array_equal(ec, a["paths_ec"]), k
        assert np.array_equal(ce, a["paths_ce"]), k


def test_ingest_genotypes_rejects_bad_allele(v2m):
    gt = np.array([[0, 2, 0, 0]], np.uint16)            # ALT 2 of a record with one ALT
    with pytest.raises(v2m.V2MError):
        v2m.paths_from_genotypes(gt, [0, 1], [0], 64, 64)
    words, _ = v2m.paths_from_genotypes(np.array([[0, 1, v2m.GT_NULL, 1]], np.uint16), [0, 1], [v2m.EDGE_MAX], 64, 64)
    assert not words.any()                               # the only ALT has no edge


def test_ingest_genotypes_vector_kernel_allele_check(v2m):
    """H % 8 == 0 takes the 128-bit-load kernel, whose range check is a packed maximum of value + 1 (mod 2^16): every bad value
    at every position of a lane's 8 values and in every load of a warp is caught, and 0, NULL and the last ALT pass."""
    H, n_alts = 2056, 3                                  # three warps per record, the last one partial
    ok = np.zeros((2, H), np.uint16)
    ok[0, ::3] = v2m.GT_NULL
    ok[0, 1::3] = n_alts
    ok[1, :] = v2m.GT_NULL
    ok[1, 5::7] = 1
    words, _ = v2m.paths_from_genotypes(ok, [0, n_alts, 2 * n_alts], list(range(2 * n_alts)), 2112, 64)
    w = np.asarray(words).reshape(-1)                    # edge-major lines of 2112 / 64 words
    bits = np.unpackbits(w.view(np.uint8), bitorder="little").reshape(64, 2112)
    assert (bits[n_alts - 1, :H] == (ok[0] == n_alts)).all() and (bits[n_alts, :H] == (ok[1] == 1)).all()
    assert bits.sum() == (ok[0] == n_alts).sum() + (ok[1] == 1).sum()
    for pos in (0, 1, 6, 7, 255, 256, 1023, 1024, 1531, 2048, 2055):
        for bad in (n_alts + 1, 0x7fff, 0x8000, 0xfffe):
            gt = ok.copy()
            gt[1, pos] = bad
            with pytest.raises(v2m.V2MError):
                v2m.paths_from_genotypes(gt, [0, n_alts, 2 * n_alts], list(range(2 * n_alts)), 2112, 64)


def test_ingest_genotypes_alt_numbers_beyond_half_precision(v2m):
    """ALT numbers up to 0x7bff are matched with half-precision compares (exact for positive finite patterns, subnormals
    included), larger ones — infinities and NaNs as binary16 — with integer compares: both sides of the switch, and the
    values that only differ from them in the sign bit or the low bit, land in the right edge columns."""
    alts = [1, 2, 0x03ff, 0x0400, 0x3c00, 0x7bfe, 0x7bff, 0x7c00, 0x7c01, 0x8000, 0x8001, 0xbc00, 0xfc00, 0xfffe]
    n_alts = 0xfffe
    gt = np.zeros((1, 64), np.uint16)
    gt[0, : 3 * len(alts)] = np.repeat(np.array(alts, np.uint16), 3)
    gt[0, 60:] = v2m.GT_NULL
    eba = np.full(n_alts, v2m.EDGE_MAX, np.uint64)
    for k, a in enumerate(alts):
        eba[a - 1] = k
    words, _ = v2m.paths_from_genotypes(gt, [0, n_alts], eba, 64, 64)
    bits = np.unpackbits(np.asarray(words).view(np.uint8), bitorder="little").reshape(64, 64)
    want = np.zeros((64, 64), np.uint8)
    for k, a in enumerate(alts):
        want[k, 3 * k: 3 * k + 3] = 1
    assert (bits == want).all()


# ---- reference-format graph files → device rows --------------------------------------------------------------
def test_rows_from_reference_format_graph_files(v2m, expected, golden_dir):
    """A graph archive as a stock vcf2multialign writes it (--output-graph; here written by the real cereal headers) is loaded,
    uploaded and materialised: founder A2M text identical to the reference's goldens, cut positions identical to the file's."""
    from vcf2multialign_b200 import interchange as ic
    for case in expected["founder_sequences"]:
        stem = "founder-sequences_" + case["vcf"][:-4]
        gp = os.path.join(golden_dir, "cereal", stem + ".graph")
        if case["fa"][:-3] != case["vcf"][:-4]:
            continue                                     # same VCF over another reference: a different graph (the sink moves)
        g = ic.load_variant_graph(gp)
        ref = orc.Graph.from_files(os.path.join(golden_dir, "founder-sequences", case["fa"]), os.path.join(golden_dir, "founder-sequences", case["vcf"]), "1").arrays()["ref"]
        dg = v2m.DeviceGraph(g, ref)
        file_cuts, md, score = ic.load_cut_positions(os.path.join(golden_dir, "cereal", stem + ".cuts"))
        cuts, sc = dg.find_cut_positions(md)
        assert cuts.tolist() == file_cuts.tolist() == case["cuts"] and sc == score
        out = v2m.founder_sequence_greedy_output(None, True, False, False)
        out.set_cut_positions(file_cuts)
        assert out.find_matchings(dg, 2)
        import tempfile
        with tempfile.TemporaryDirectory() as td:
            dst = os.path.join(td, "out.a2m")
            out.output_a2m(None, dg, dst)
            assert open(dst).read() == case["a2m"], case["vcf"]


# ---- C++ host adapters (vcf2multialign_b200/host/v2m_b200.hh) ------------------------------------------------------------
@pytest.mark.parametrize("fa,vcf", [("founder-sequences/test-1.fa", "founder-sequences/test-1.vcf"), ("founder-sequences/test-1-2.fa", "founder-sequences/test-1.vcf"),
                                    ("founder-sequences/test-2.fa", "founder-sequences/test-2.vcf"), ("founder-sequences/test-3.fa", "founder-sequences/test-3.vcf"),
                                    ("founder-sequences/test-4.fa", "founder-sequences/test-4.vcf"), ("variant-graph/test-2.fa", "variant-graph/test-2.vcf"),
                                    ("variant-graph/test-4.fa", "variant-graph/test-4.vcf")])
def test_cxx_host_adapters(v2m, expected, golden_dir, tmp_path, fa, vcf):
    """transpose_matrix, output_sequence, haplotype_output, founder_sequence_greedy_output with the reference's signatures,
    byte-compared with the oracle inside tests/host/adapter_test.cc (incl. delegate call order, O_EXCL, --pipe, output_separate)."""
    import subprocess
    exe = os.path.join(os.path.dirname(__file__), "host", "_build", "adapter_test")
    if not os.path.exists(exe):
        subprocess.run(["make", "-C", os.path.join(os.path.dirname(__file__), "host")], check=True)
    r = subprocess.run([exe, os.path.join(golden_dir, fa), os.path.join(golden_dir, vcf), str(tmp_path), "2"], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0 and "ALL OK" in r.stdout, r.stdout[-3000:] + r.stderr[-2000:]
    if fa.startswith("founder-sequences"):
        case = [c for c in expected["founder_sequences"] if c["vcf"] == os.path.basename(vcf) and c["fa"] == os.path.basename(fa)]
        assert case and case[0]["a2m"] in r.stdout      # the founder A2M text the test program echoes is the reference's golden


# ---- row-range sharding with the GPU as the row producer ------------------------------------------------------------
@pytest.mark.parametrize("unaligned", [False, True])
def test_sharded_stream_gpu_rows(v2m, tmp_path, unaligned):
    """Three 'ranks' (run one after the other in this process, no collective needed) each write their row range."""
    from vcf2multialign_b200 import sharding
    og = rg.random_oracle_graph(str(tmp_path), 4242, ref_len=5000, n_records=500, n_samples=7, ploidy=2, long_del_prob=0.02)
    dg, g = _device_graph(v2m, og)
    out = v2m.haplotype_output("c", True, unaligned)
    dst = str(tmp_path / "sharded.a2m")
    world = 3
    if unaligned:
        # every rank needs all lengths: without a process group, hand them over as gather_lengths would
        plan = dg.plan_haplotypes(np.array(out._row_table(g)[0], np.uint32))
        all_lengths = plan.row_lengths(True)

        class FakeDist:
            def barrier(self): pass
            def all_gather_object(self, parts, local):
                for r in range(world):
                    b, e = sharding.row_range(len(all_lengths), r, world)
                    parts[r] = [int(x) for x in all_lengths[b:e]]
        dist = FakeDist()
    else:
        dist = None
    for rank in range(world):
        sharding.sharded_haplotype_output_a2m(out, g, dg, dst, rank, world, dist, block_rows=4)
    assert open(dst, "rb").read() == og.haplotype_output_a2m("c", True, unaligned)


def test_shard_graph_rows_on_device(v2m, tmp_path):
    """A rank's slab of the path matrix (sharding.shard_graph) uploaded on its own gives that rank's rows of the cohort."""
    from vcf2multialign_b200 import sharding
    og = rg.random_oracle_graph(str(tmp_path), 99, ref_len=9000, n_records=800, n_samples=90, ploidy=2, long_del_prob=0.03)
    a = dict(og.arrays())
    a["paths_ce"] = None
    g = v2m.VariantGraph.from_arrays(a)
    H = g.total_chromosome_copies()
    for world in (2, 3):
        for rank in range(world):
            b, e = sharding.row_range(H + 1, rank, world)                     # rows: REF, then every copy
            rows = [orc.PLOIDY_MAX if r == 0 else r - 1 for r in range(b, e)]
            copies = [c for c in rows if c != orc.PLOIDY_MAX]
            sub, off = sharding.shard_graph(g, min(copies), max(copies) + 1)
            dg = v2m.DeviceGraph(sub, og.arrays()["ref"])
            plan = dg.plan_haplotypes([c if c == orc.PLOIDY_MAX else c - off for c in rows])
            out = plan.emit_rows()
            for i, c in enumerate(rows):
                assert out[i].tobytes() == og.output_sequence(c), (world, rank, c)
            plan.close()
            dg.close()


def test_graph_without_reference_serves_the_sweeps(v2m, tmp_path):
    """find_cut_positions(graph, …) / find_matchings(graph, …) take no reference sequence (output.hh:123, 126): a graph uploaded
    without one runs both; rows need v2m_cuda_graph_attach_reference first and are then identical."""
    og = rg.random_oracle_graph(str(tmp_path), 31, ref_len=4000, n_records=400, n_samples=12, ploidy=2)
    a = dict(og.arrays())
    a["paths_ce"] = None
    g = v2m.VariantGraph.from_arrays(a)
    dg = v2m.DeviceGraph(g, None)
    f = orc.FounderOutput(og)
    assert f.find_cut_positions(10)
    cuts, score = dg.find_cut_positions(10)
    assert cuts.tolist() == f.cut_positions().tolist() and score == f.score()
    try:
        ok = f.find_matchings(3)
    except orc.OracleError:
        ok = False
    if ok:
        assert np.array_equal(dg.find_matchings(cuts, 3), f.assigned_samples())
    plan = dg.plan_haplotypes([0, orc.PLOIDY_MAX])
    with pytest.raises(v2m.V2MError):
        plan.emit_rows()
    dg.attach_reference(og.arrays()["ref"])
    rows = plan.emit_rows()
    assert rows[0].tobytes() == og.output_sequence(0) and rows[1].tobytes() == og.output_sequence(orc.PLOIDY_MAX)
    un = plan.emit_rows_unaligned()
    assert un[0] == og.output_sequence(0, None, True)


def test_device_writers_stay_inside_their_buffers(v2m, tmp_path):
    """compute-sanitizer is closed on this GPU pool, so the writers' bounds are checked with guard bytes instead: emit_rows
    (bulk stores + ragged tails, aligned and unaligned windows) must not touch the bytes before the buffer, after it, or
    between a row's last column and the next row; the three transpose kernels must not touch what surrounds their destination."""
    import ctypes as C
    from vcf2multialign_b200._lib import check, lib, ptr
    L = lib()
    GUARD = 4096
    og = rg.random_oracle_graph(str(tmp_path), 4242, ref_len=30000, n_records=2200, n_samples=35, ploidy=2, long_del_prob=0.03)
    a = dict(og.arrays()); a["paths_ce"] = None
    g = v2m.VariantGraph.from_arrays(a)
    dg = v2m.DeviceGraph(g, og.arrays()["ref"])
    H, A = g.total_chromosome_copies(), g.aligned_length()
    idx = list(range(H)) + [orc.PLOIDY_MAX]
    plan = dg.plan_haplotypes(idx)
    full = [og.output_sequence(h) for h in idx]
    for c0, c1 in ((0, A), (4096 * 2 + 16, A - 3), (16, 4096 * 3 + 5), (4096, 4096 * 2)):
        width = c1 - c0
        pitch = int(L.v2m_cuda_row_pitch(width))
        n = len(idx)
        total = 2 * GUARD + n * pitch
        host = np.full(total, 0xEE, np.uint8)
        d = L.v2m_cuda_device_alloc(total)
        assert d
        check(L.v2m_cuda_memcpy_h2d(d, ptr(host), total))
        check(L.v2m_cuda_emit_rows_device(plan._p, 0, n, c0, c1, C.c_void_p(d + GUARD), pitch, None))
        check(L.v2m_cuda_synchronize())
        check(L.v2m_cuda_memcpy_d2h(ptr(host), d, total))
        L.v2m_cuda_device_free(d)
        assert (host[:GUARD] == 0xEE).all() and (host[GUARD + n * pitch:] == 0xEE).all(), (c0, c1)
        body = host[GUARD : GUARD + n * pitch].reshape(n, pitch)
        assert (body[:, width:] == 0xEE).all(), (c0, c1)
        for i in range(n):
            assert body[i, :width].tobytes() == full[i][c0:c1], (c0, c1, i)
    plan.close(); dg.close()

    rng = np.random.default_rng(8)
    for kernel in ("0", "1", "2"):
        for rows, cols in ((1088, 3264), (576, 5184), (1280, 20928)):
            os.environ["V2M_K1_TALL"] = kernel
            try:
                src = rng.integers(0, 2**64, rows * cols // 64, dtype=np.uint64)
                nbytes = rows * cols // 8
                host = np.full(2 * GUARD + nbytes, 0xEE, np.uint8)
                d_src, d = L.v2m_cuda_device_alloc(nbytes), L.v2m_cuda_device_alloc(host.size)
                check(L.v2m_cuda_memcpy_h2d(d_src, ptr(src), nbytes))
                check(L.v2m_cuda_memcpy_h2d(d, ptr(host), host.size))
                check(L.v2m_cuda_transpose_bits_device(d_src, rows, cols, C.c_void_p(d + GUARD), None))
                check(L.v2m_cuda_synchronize())
                check(L.v2m_cuda_memcpy_d2h(ptr(host), d, host.size))
                L.v2m_cuda_device_free(d_src); L.v2m_cuda_device_free(d)
                assert (host[:GUARD] == 0xEE).all() and (host[GUARD + nbytes:] == 0xEE).all(), (kernel, rows, cols)
                assert np.array_equal(host[GUARD : GUARD + nbytes].view(np.uint64), orc.transpose(src, rows, cols)), (kernel, rows, cols)
            finally:
                os.environ.pop("V2M_K1_TALL", None)
