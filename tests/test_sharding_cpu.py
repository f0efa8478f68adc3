"""N > 1 host logic on CPU: two gloo processes write one A2M stream by row range (vcf2multialign_b200/sharding.py).
The GPU row producer is replaced by the oracle (row_source) — the sharding, offsets, barriers and hooks are what is tested."""
import os
import socket
import sys

import numpy as np
import pytest

from oracle import oracle as orc
from tests import randgraph as rg
from vcf2multialign_b200 import sharding
from vcf2multialign_b200.graph import VariantGraph
from vcf2multialign_b200.output import haplotype_output, output_delegate


def test_row_range_partitions():
    for n in (0, 1, 5, 5009, 64):
        for world in (1, 2, 3, 4, 8):
            ranges = [sharding.row_range(n, r, world) for r in range(world)]
            assert ranges[0][0] == 0 and ranges[-1][1] == n
            assert all(ranges[i][1] == ranges[i + 1][0] for i in range(world - 1))
            sizes = [e - b for b, e in ranges]
            assert max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        sharding.row_range(4, 2, 2)


def test_row_offsets():
    off = sharding.row_offsets(["REF", "a\tS1-1"], [10, 10])
    assert off.tolist() == [0, 1 + 3 + 1 + 10 + 1, 16 + 1 + 6 + 1 + 10 + 1]


class OracleRows:
    def __init__(self, og):
        self.og = og

    def lengths(self, copy_idx, unaligned):
        return [len(self.og.output_sequence(int(c), None, unaligned)) for c in copy_idx]

    def rows(self, copy_idx, unaligned):
        return [self.og.output_sequence(int(c), None, unaligned) for c in copy_idx]


class CountingDelegate(output_delegate):
    def __init__(self):
        self.events = []

    def will_handle_sample(self, sample, sample_idx, chr_copy_idx):
        self.events.append(("S", sample, sample_idx, chr_copy_idx))

    def handled_sequences(self, n):
        self.events.append(("H", n))


def _worker(rank, world, port, workdir, unaligned, omit_ref):
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        og = rg.random_oracle_graph(workdir, 77, ref_len=400, n_records=50, n_samples=5, ploidy=2)
        g = VariantGraph.from_arrays(og.arrays())
        d = CountingDelegate()
        out = haplotype_output("chrS", not omit_ref, unaligned, d)
        dst = os.path.join(workdir, f"out_{int(unaligned)}_{int(omit_ref)}.a2m")
        sharding.sharded_haplotype_output_a2m(out, g, None, dst, rank, world, dist, block_rows=3, row_source=OracleRows(og))
        n_rows = g.total_chromosome_copies() + (0 if omit_ref else 1)
        b, e = sharding.row_range(n_rows, rank, world)
        handled = [ev[1] for ev in d.events if ev[0] == "H"]
        first_sample_row = 0 if omit_ref else 1
        expect = [1 if (not omit_ref and r == 0) else (r - first_sample_row) + 2 for r in range(b, e)]
        assert handled == expect, (rank, handled, expect)
        if rank == 0:
            with open(dst, "rb") as f:
                got = f.read()
            assert got == og.haplotype_output_a2m("chrS", not omit_ref, unaligned)
            # a second run must refuse to overwrite (O_EXCL), on rank 0 where the file is created
            with pytest.raises(FileExistsError):
                sharding.ShardedA2MWriter(dst, ["x"], [1], 0, 1).__enter__()
    finally:
        dist.destroy_process_group()


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


@pytest.mark.parametrize("unaligned,omit_ref", [(False, False), (True, False), (False, True)])
def test_two_rank_gloo_stream(tmp_path, unaligned, omit_ref):
    import torch.multiprocessing as mp
    for attempt in range(3):        # the probed port can be taken by someone else before the ranks bind it
        try:
            mp.spawn(_worker, args=(2, _free_port(), str(tmp_path), unaligned, omit_ref), nprocs=2, join=True)
            break
        except Exception as exc:
            if attempt == 2 or "address" not in str(exc).lower():
                raise


def test_single_process_stream(tmp_path):
    og = rg.random_oracle_graph(str(tmp_path), 78, ref_len=300, n_records=40, n_samples=3, ploidy=2)
    g = VariantGraph.from_arrays(og.arrays())
    out = haplotype_output(None, True, False)
    dst = str(tmp_path / "single.a2m")
    sharding.sharded_haplotype_output_a2m(out, g, None, dst, row_source=OracleRows(og))
    assert open(dst, "rb").read() == og.haplotype_output_a2m(None, True, False)


def test_shard_graph_rows_equal_cohort_rows():
    """sharding.shard_graph: a rank's slab of the path matrices gives the same rows (checked with the oracle's row writer)."""
    import numpy as np
    from oracle import oracle as orc
    from vcf2multialign_b200 import sharding, synth
    w = synth.make_workload("x", seed=11, ref_len=30_000, n_sites=900, n_samples=100, with_paths=True)
    g = w.graph
    g.paths_by_chrom_copy_and_edge = orc.transpose(g.paths_by_edge_and_chrom_copy, g.h_pad, g.e_pad)

    def oracle_graph(vg):
        return orc.Graph.from_arrays(dict(
            reference_positions=vg.reference_positions, aligned_positions=vg.aligned_positions, alt_edge_targets=vg.alt_edge_targets,
            alt_edge_count_csum=vg.alt_edge_count_csum, label_offsets=vg.label_offsets, label_chars=vg.label_chars,
            paths_ec=vg.paths_by_edge_and_chrom_copy, paths_ce=vg.paths_by_chrom_copy_and_edge, h_pad=vg.h_pad, e_pad=vg.e_pad,
            ploidy_csum=vg.ploidy_csum, sample_names=vg.sample_names, ref=w.ref))

    full = oracle_graph(g)
    for begin, end in ((0, 200), (70, 131), (128, 192), (199, 200), (63, 65)):
        sub, off = sharding.shard_graph(g, begin, end)
        assert off % 64 == 0 and off <= begin and sub.h_pad % 64 == 0 and sub.h_pad <= g.h_pad
        assert np.array_equal(sub.paths_by_chrom_copy_and_edge, orc.transpose(sub.paths_by_edge_and_chrom_copy, sub.h_pad, sub.e_pad))
        og = oracle_graph(sub)
        for c in {begin, (begin + end) // 2, end - 1}:
            assert og.output_sequence(c - off) == full.output_sequence(c), (begin, end, c)


def _take_blocks(path, n_rows, block_rows, out_q):
    from vcf2multialign_b200 import sharding
    q = sharding.RowBlockQueue(path, n_rows, block_rows)
    got = []
    while True:
        b = q.next()
        if b is None:
            break
        got.append(b)
    q.close()
    out_q.put(got)


def test_row_block_queue_hands_every_block_out_once(tmp_path):
    """sharding.RowBlockQueue: several processes draw row blocks from one counter; every block is taken exactly once."""
    import multiprocessing as mp
    from vcf2multialign_b200 import sharding
    path = str(tmp_path / "queue")
    q0 = sharding.RowBlockQueue(path, 5009, 8)
    q0.reset()
    ctx = mp.get_context("spawn")
    out = ctx.Queue()
    procs = [ctx.Process(target=_take_blocks, args=(path, 5009, 8, out)) for _ in range(3)]
    for p in procs:
        p.start()
    taken = [b for _ in procs for b in out.get(timeout=120)]
    for p in procs:
        p.join(timeout=60)
    assert sorted(taken) == [(b, min(5009, b + 8)) for b in range(0, 5009, 8)]
    assert q0.next() is None
    q0.reset()
    assert q0.next() == (0, 8)
    q0.close()
