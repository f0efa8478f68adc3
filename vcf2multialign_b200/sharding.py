"""Row-range sharding of one A2M stream over the GPUs of a box: rows are independent, so rank r owns a
contiguous range of the output rows, produces them on its own GPU and writes them at their final byte offsets of the
shared output file with pwrite(2). No data-path collective: the only exchange is the control-plane barrier after rank 0
has created the file (and, for --unaligned, an all-gather of the per-row lengths — a few KB).

The byte layout is the one haplotype_output::output_a2m produces (haplotype_output.cc:38-82): '>' id '\\n' row '\\n' per row.
"""
from __future__ import annotations

import os

import numpy as np


def row_range(n_rows: int, rank: int, world: int) -> tuple[int, int]:
    """Contiguous, balanced: the first n_rows % world ranks get one row more."""
    if world < 1 or not 0 <= rank < world:
        raise ValueError("rank / world out of range")
    base, extra = divmod(n_rows, world)
    begin = rank * base + min(rank, extra)
    return begin, begin + base + (1 if rank < extra else 0)


def shard_graph(graph, copy_begin: int, copy_end: int):
    """The variant graph as one rank needs it: same nodes, edges and labels, the path matrices cut down to the 64-copy words
    that cover chromosome copies [copy_begin, copy_end). Returns (sub_graph, copy_offset): chromosome copy c of the cohort is
    copy c - copy_offset of sub_graph. A row of the output depends on its own line of paths_by_chrom_copy_and_edge only
    (sequence_writer.cc:51-67 tests bit (edge, copy)), so the shard's rows are the cohort's rows — and a rank uploads
    (copy_end - copy_begin) / H of the path bits instead of all of them."""
    from .graph import VariantGraph
    H = graph.total_chromosome_copies()
    if not 0 <= copy_begin < copy_end <= H:
        raise ValueError("chromosome copy range out of bounds")
    w0, w1 = copy_begin // 64, (copy_end + 63) // 64
    Wh, We = graph.h_pad // 64, graph.e_pad // 64
    ec = ce = None
    if graph.paths_by_edge_and_chrom_copy is not None:
        ec = np.ascontiguousarray(graph.paths_by_edge_and_chrom_copy.reshape(graph.e_pad, Wh)[:, w0:w1]).reshape(-1)
    if graph.paths_by_chrom_copy_and_edge is not None:
        ce = np.ascontiguousarray(graph.paths_by_chrom_copy_and_edge[64 * w0 * We : 64 * w1 * We])
    n_copies = min(H, 64 * w1) - 64 * w0
    sub = VariantGraph(
        graph.reference_positions, graph.aligned_positions, graph.alt_edge_targets, graph.alt_edge_count_csum, graph.label_offsets,
        graph.label_chars, ec, ce, 64 * (w1 - w0), graph.e_pad, [f"copy{64 * w0 + i}" for i in range(n_copies)],
        np.arange(n_copies + 1, dtype=np.uint32))
    return sub, 64 * w0


class RowBlockQueue:
    """Row blocks handed out to the ranks of one box as they ask for them (control plane only: a counter in a small file that
    every rank can open, e.g. under /dev/shm, advanced under an fcntl lock — no collective, nothing on the data path).

    Why: the host link is not the same for every GPU of a box (on this pool's 8-GPU boxes a plain D2H copy runs at 8 – 18 GB/s
    per GPU when all eight copy at once, profiles/r02_d2h_ceiling.txt), so with equal static row ranges the whole stream waits
    for the rank behind the slowest link. Every aligned row has the same length and its place in the A2M stream is known up
    front (row_offsets), so any rank can produce any block."""

    def __init__(self, path: str, n_rows: int, block_rows: int):
        import fcntl
        self._fcntl = fcntl
        self.path, self.n_rows, self.block_rows = path, int(n_rows), int(block_rows)
        self.n_blocks = (self.n_rows + self.block_rows - 1) // self.block_rows
        self.fd = os.open(path, os.O_RDWR | os.O_CREAT, 0o600)

    def reset(self):
        """Called by ONE rank before the others start taking blocks (callers put a barrier after it)."""
        os.pwrite(self.fd, (0).to_bytes(8, "little"), 0)

    def next(self):
        """The next block's row range (begin, end), or None when all blocks have been taken."""
        f = self._fcntl
        f.lockf(self.fd, f.LOCK_EX, 8, 0)
        try:
            raw = os.pread(self.fd, 8, 0)
            b = int.from_bytes(raw, "little") if len(raw) == 8 else 0
            if b >= self.n_blocks:
                return None
            os.pwrite(self.fd, (b + 1).to_bytes(8, "little"), 0)
        finally:
            f.lockf(self.fd, f.LOCK_UN, 8, 0)
        begin = b * self.block_rows
        return begin, min(self.n_rows, begin + self.block_rows)

    def close(self):
        if self.fd is not None:
            os.close(self.fd)
            self.fd = None


def replicate_path_matrix(dist, device_index: int, words, n_words: int):
    """The edge-major path matrix on every GPU of the box with ONE host-to-device pass over it: every rank uploads its 1 / N
    slice of the host words and the ranks exchange the slices over NVLink (one NCCL all-gather — replication of the job's
    input, not a collective on the row data path; SURVEY §5 names it as the alternative to N full uploads through the shared
    host link). `words`: the host matrix (every rank has it; only the rank's own slice is read). Returns a torch int64 tensor
    on cuda:device_index whose data_ptr() goes into DeviceGraph(d_paths_by_edge_and_chrom_copy=…)."""
    import torch
    dev = torch.device("cuda", device_index)
    rank, world = dist.get_rank(), dist.get_world_size()
    per = (int(n_words) + world - 1) // world
    full = torch.empty(per * world, dtype=torch.int64, device=dev)
    lo, hi = min(n_words, rank * per), min(n_words, (rank + 1) * per)
    part = full[rank * per : (rank + 1) * per]
    if hi > lo:
        part[: hi - lo].copy_(torch.from_numpy(np.ascontiguousarray(words[lo:hi]).view(np.int64)))
    dist.all_gather_into_tensor(full, part.clone())
    torch.cuda.synchronize(dev)
    return full[: int(n_words)]


def row_offsets(ids, lengths) -> np.ndarray:
    """Byte offset of every row's '>' in the stream, plus the total size as the last entry."""
    sizes = np.array([1 + len(i.encode()) + 1 + int(n) + 1 for i, n in zip(ids, lengths)], np.int64)
    out = np.zeros(len(sizes) + 1, np.int64)
    np.cumsum(sizes, out=out[1:])
    return out


def _pwrite_all(fd: int, data, offset: int):
    view = memoryview(data)
    while len(view):
        n = os.pwrite(fd, view, offset)
        view = view[n:]
        offset += n


class ShardedA2MWriter:
    """Writes this rank's rows of an A2M stream into the shared file `path`.

    ids: the FASTA identifiers of ALL rows (every rank computes the same list from the graph);
    lengths: byte length of every row (aligned: all equal to aligned_positions.back());
    dist: an initialised torch.distributed module (any backend) or None for a single process."""

    def __init__(self, path: str, ids, lengths, rank: int = 0, world: int = 1, dist=None):
        if len(ids) != len(lengths):
            raise ValueError("ids and lengths differ in length")
        self.path, self.ids, self.rank, self.world, self.dist = path, list(ids), rank, world, dist
        self.lengths = [int(x) for x in lengths]
        self.offsets = row_offsets(self.ids, self.lengths)
        self.begin, self.end = row_range(len(self.ids), rank, world)
        self.fd = None

    def __enter__(self):
        if self.rank == 0:
            # exclusive create, as lb::open_file_for_writing(…, CREATE) does (output.cc:70)
            fd = os.open(self.path, os.O_WRONLY | os.O_CREAT | os.O_EXCL, 0o644)
            os.ftruncate(fd, int(self.offsets[-1]))
            self.fd = fd
        if self.dist is not None and self.world > 1:
            self.dist.barrier()
        if self.rank != 0:
            self.fd = os.open(self.path, os.O_WRONLY)
        return self

    def write_rows(self, first_row: int, rows):
        """rows: iterable of bytes-like row bodies for rows first_row, first_row + 1, … (all within this rank's range)."""
        r = first_row
        for body in rows:
            if not self.begin <= r < self.end:
                raise ValueError(f"row {r} is outside rank {self.rank}'s range [{self.begin}, {self.end})")
            if len(body) != self.lengths[r]:
                raise ValueError(f"row {r}: {len(body)} bytes, expected {self.lengths[r]}")
            head = b">" + self.ids[r].encode() + b"\n"
            off = int(self.offsets[r])
            _pwrite_all(self.fd, head, off)
            _pwrite_all(self.fd, body, off + len(head))
            _pwrite_all(self.fd, b"\n", off + len(head) + len(body))
            r += 1

    def __exit__(self, *exc):
        if self.fd is not None:
            os.close(self.fd)
            self.fd = None
        if self.dist is not None and self.world > 1 and exc[0] is None:
            self.dist.barrier()
        return False


def gather_lengths(local_lengths, n_rows: int, rank: int, world: int, dist=None) -> list[int]:
    """--unaligned: every rank knows only its own rows' lengths; all-gather them (control plane, 8 bytes per row)."""
    begin, end = row_range(n_rows, rank, world)
    if len(local_lengths) != end - begin:
        raise ValueError("local_lengths does not match this rank's row range")
    if dist is None or world == 1:
        return [int(x) for x in local_lengths]
    parts = [None] * world
    dist.all_gather_object(parts, [int(x) for x in local_lengths])
    return [x for p in parts for x in p]


def sharded_haplotype_output_a2m(out, graph, dgraph, path: str, rank: int = 0, world: int = 1, dist=None, block_rows: int = 64, row_source=None):
    """haplotype_output::output_a2m over `world` GPUs. `out`: vcf2multialign_b200.output.haplotype_output; `graph`: the
    VariantGraph; `dgraph`: this rank's DeviceGraph (None only when row_source is given). row_source — an object with
    lengths(copy_idx, unaligned) and rows(copy_idx, unaligned) — replaces the GPU producer in the CPU (gloo) tests."""
    plan_idx, rows = out._row_table(graph)
    n = len(rows)
    begin, end = row_range(n, rank, world)
    ids = [r[0] for r in rows]
    unaligned = out.m_should_output_unaligned
    plan = None
    if row_source is None:
        plan = dgraph.plan_haplotypes(np.asarray(plan_idx[begin:end], np.uint32))
    try:
        if unaligned:
            local = plan.row_lengths(True) if plan is not None else row_source.lengths(plan_idx[begin:end], True)
            lengths = gather_lengths(local, n, rank, world, dist)
        else:
            lengths = [graph.aligned_length()] * n
        with ShardedA2MWriter(path, ids, lengths, rank, world, dist) as w:
            for r0 in range(begin, end, block_rows):
                r1 = min(end, r0 + block_rows)
                if plan is None:
                    bodies = row_source.rows(plan_idx[r0:r1], unaligned)
                elif unaligned:
                    bodies = plan.emit_rows_unaligned(r0 - begin, r1 - begin)
                else:
                    arr = plan.emit_rows(r0 - begin, r1 - begin)
                    bodies = [arr[i] for i in range(r1 - r0)]
                for r in range(r0, r1):
                    if rows[r][2]:
                        rows[r][2]()
                w.write_rows(r0, bodies)
                for r in range(r0, r1):
                    if rows[r][3]:
                        rows[r][3]()
    finally:
        if plan is not None:
            plan.close()
