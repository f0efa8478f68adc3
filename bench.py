#!/usr/bin/env python3
"""bench.py — A2M aligned Gbase/s of the row-materialisation hot path (BASELINE.json metric) on B200.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload c4|c2|c1|c5] [--scaling strong|weak] [--impl reference]

Workload (default): c4, the chr1-shaped job north_star's target sentence names — 249 Mbp reference, 6 M variant sites, 5,008
haplotypes + the REF row = 5,009 rows × A ≈ 250.3 M aligned columns ≈ 1.25 TB of A2M rows per pass. One "step" = one pass of
the hot path over the rank's rows: K2a (resolve_rows: which ALT edges each row follows, sequence_writer.cc:44-67) + K2b
(emit_rows: the bytes, sequence_writer.cc:70-83), the row loop of haplotype_output.cc:61-81.

N > 1 (torchrun): ONE job, strong scaling — rank r owns rows sharding.row_range(5009, r, N) of the same graph (same seed on
every rank); rows shard by haplotype range with no collective on the data path (NCCL carries the barrier and the max of the
device times only). `--scaling weak` gives every rank its own 5,008-haplotype cohort instead (round 1's measurement).

`value` is device-resident (rows written into an HBM ring of two row blocks that is overwritten); `e2e` runs the same job
through the C ABI with host buffers: graph arrays uploaded from host memory, K1, K2a, K2b, every row copied into pinned host
memory, all inside the timed region, and is reported as a fraction of the host link's own ceiling — plain cudaMemcpyAsync
device-to-host copies, measured live on the same box with all ranks copying at once (`e2e.peak`). At N > 1 the e2e leg
hands the job's rows out in blocks from a shared counter (sharding.RowBlockQueue; still no collective): the GPUs of a box do
not all get the same share of the host link, and static equal ranges would wait for the slowest one.

`--workload c5` prints the transpose_matrix microbench line (K1 on a 5056 × 6,000,000 bit matrix, GB/s against the HBM
roofline) instead. `--impl reference` times the reference's CPU implementation of the path — its restatement in oracle/
(the reference itself needs Boost/Ragel, absent from this image) — on all host cores, on a bounded row sample of the same
workload.
"""
from __future__ import annotations

import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "a2m_aligned_gbase_per_s"
UNIT = "Gbase/s"


def measured_peak_gbs():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


class ClockSampler:
    """SM clock / throttle reasons sampled during the timed region (B200_PROFILING.md recipe): NVML in-process every 10 ms
    (the timed region of a short run is ~0.1 s), nvidia-smi every 0.2 s if NVML cannot be loaded."""
    Q = "clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"
    NAMES = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]

    def __init__(self, index: int):
        self.index, self.samples, self.stop, self.thread = index, [], threading.Event(), None
        self.nvml = None
        try:
            import pynvml
            pynvml.nvmlInit()
            # NVML enumerates physical devices; honour CUDA_VISIBLE_DEVICES when it lists indices
            vis = os.environ.get("CUDA_VISIBLE_DEVICES", "")
            phys = index
            if vis and all(x.strip().isdigit() for x in vis.split(",")):
                phys = int(vis.split(",")[index])
            self.handle = pynvml.nvmlDeviceGetHandleByIndex(phys)
            self.sm_max = float(pynvml.nvmlDeviceGetMaxClockInfo(self.handle, pynvml.NVML_CLOCK_SM))
            self.nvml = pynvml
        except Exception:
            self.nvml = None

    def _run_nvml(self):
        n = self.nvml
        masks = [n.nvmlClocksThrottleReasonHwSlowdown, n.nvmlClocksThrottleReasonHwThermalSlowdown,
                 n.nvmlClocksThrottleReasonSwThermalSlowdown, n.nvmlClocksThrottleReasonSwPowerCap]
        while not self.stop.is_set():
            try:
                sm = float(n.nvmlDeviceGetClockInfo(self.handle, n.NVML_CLOCK_SM))
                r = int(n.nvmlDeviceGetCurrentClocksThrottleReasons(self.handle))
                self.samples.append([str(sm), str(self.sm_max)] + ["Active" if r & m else "Not Active" for m in masks])
            except Exception:
                pass
            self.stop.wait(0.01)

    def _run_smi(self):
        while not self.stop.is_set():
            try:
                out = subprocess.run(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-i", str(self.index)],
                                     capture_output=True, text=True, timeout=5).stdout.strip()
                if out:
                    self.samples.append([x.strip() for x in out.split(",")])
            except Exception:
                pass
            self.stop.wait(0.2)

    def __enter__(self):
        self.thread = threading.Thread(target=self._run_nvml if self.nvml else self._run_smi, daemon=True)
        self.thread.start()
        return self

    def __exit__(self, *a):
        self.stop.set()
        self.thread.join(timeout=6)

    def summary(self):
        sm = [float(s[0]) for s in self.samples if s and s[0].replace(".", "").isdigit()]
        mx = [float(s[1]) for s in self.samples if len(s) > 1 and s[1].replace(".", "").isdigit()]
        reasons = sorted({self.NAMES[i] for s in self.samples if len(s) >= 6 for i in range(4) if s[2 + i].lower().startswith("active")})
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": reasons, "samples": len(self.samples),
                "source": "nvml" if self.nvml else "nvidia-smi"}


# stdout carries the one JSON line and nothing else: the process's fd 1 is pointed at stderr for the whole run (NCCL prints
# its version banner on fd 1 from C code when NCCL_DEBUG is set on the box), the line goes out through the saved descriptor.
_JSON_FD = None


def claim_stdout():
    global _JSON_FD
    if _JSON_FD is None:
        sys.stdout.flush()
        _JSON_FD = os.dup(1)
        os.dup2(2, 1)


def emit_line(line):
    data = (json.dumps(line) + "\n").encode()
    if _JSON_FD is None:
        sys.stdout.write(data.decode())
        sys.stdout.flush()
    else:
        sys.stdout.flush()
        os.write(_JSON_FD, data)


def dist_setup(n_gpus):
    rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
    dist = None
    if world > 1:
        import torch
        import torch.distributed as dist_mod
        torch.cuda.set_device(local)
        dist_mod.init_process_group("nccl", device_id=torch.device("cuda", local))
        dist = dist_mod
    return rank, world, local, dist


def barrier(dist, local):
    if dist is not None:
        import torch
        dist.barrier(device_ids=[local])
        torch.cuda.synchronize()


def sum_over_ranks(dist, local, x: float) -> float:
    if dist is None:
        return x
    import torch
    t = torch.tensor([x], dtype=torch.float64, device=torch.device("cuda", local))
    dist.all_reduce(t, op=dist.ReduceOp.SUM)
    return float(t.item())


def max_over_ranks(dist, local, x: float) -> float:
    if dist is None:
        return x
    import torch
    t = torch.tensor([x], dtype=torch.float64, device=torch.device("cuda", local))
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def build_workload(name, rank, seed=1, scaling="strong"):
    from vcf2multialign_b200 import synth
    w = synth.make_workload(name, seed=seed, with_paths=False)
    if scaling == "weak":
        # rank r owns chromosome copies [r·H, (r+1)·H) of an N×H cohort: same sites, different haplotypes.
        w.seed = (seed + 7919 * rank) & 0xFFFFFFFF
    return w


def edge_nalts(w):
    n_alts = np.bincount(w.edge_site, minlength=len(w.site_pos)).astype(np.uint8)
    return np.ascontiguousarray(n_alts[w.edge_site])


def oracle_subcohort(w, first_copy, n_copies, lines):
    """The oracle's graph restricted to chromosome copies [first_copy, first_copy + n_copies) (n_copies even): the row writer
    reads only a copy's own line of paths_by_chrom_copy_and_edge, so the sub-cohort's rows are the cohort's rows."""
    from oracle import oracle as orc
    g = w.graph
    We = g.e_pad // 64
    h_pad = (n_copies + 63) // 64 * 64
    ce = np.zeros(h_pad * We, np.uint64)
    ce[: n_copies * We] = lines[: n_copies * We]
    n_samples = n_copies // 2
    return orc.Graph.from_arrays(dict(
        reference_positions=g.reference_positions, aligned_positions=g.aligned_positions, alt_edge_targets=g.alt_edge_targets,
        alt_edge_count_csum=g.alt_edge_count_csum, label_offsets=g.label_offsets, label_chars=g.label_chars,
        paths_ec=np.zeros(g.e_pad * (h_pad // 64), np.uint64),      # the edge-major orientation is not read by the row writer
        paths_ce=ce, h_pad=h_pad, e_pad=g.e_pad, ploidy_csum=(2 * np.arange(n_samples + 1)).astype(np.uint32),
        sample_names=g.sample_names[first_copy // 2 : first_copy // 2 + n_samples], ref=w.ref))


def cpu_reference_arm(args):
    """--impl reference: the CPU path (oracle port) with all host threads on a bounded sample of the workload's rows."""
    rank = int(os.environ.get("RANK", 0))
    if rank != 0:
        return 0
    from oracle import oracle as orc
    w = build_workload(args.workload, 0)
    g = w.graph
    cores = os.cpu_count() or 1
    # Path bits of the sampled chromosome copies only, from the CPU mirror of the generator (no GPU library in this arm).
    n_rows = 2 * cores
    lines = orc.synth_path_lines(w.seed, 0, n_rows, g.e_pad, w.edge_site, w.edge_alt, edge_nalts(w), threads=cores)
    og = oracle_subcohort(w, 0, n_rows, lines)
    A = g.aligned_length()
    for _ in range(args.warmup):
        og.time_rows(0, cores, cores, bulk=False)
    t = 0.0
    bases = 0
    for _ in range(args.steps):
        s, b = og.time_rows(0, n_rows, cores, bulk=False)
        t += s
        bases += b
    value = bases / t / 1e9
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": 1e3 * t / args.steps, "higher_is_better": True, "scaling": args.scaling, "vs_baseline": None, "dtype": "u8", "data": "synthetic",
        "config": {"workload": workload_name(args.workload, g, A)},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "port",
                         "sample": f"{n_rows} haplotype rows x {A} columns per step, row-parallel over {cores} threads, faithful I/O shape (4 KiB buffer, per-char gaps)"},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    emit_line(line)
    return 0


def workload_name(name, g, A):
    return f"{name}: synthetic {int(g.reference_positions[-1])} bp reference, {g.edge_count()} ALT edges, {g.total_chromosome_copies()} haplotypes + REF row, A={A} aligned columns"


def transpose_microbench(args):
    """--workload c5: transpose_matrix (K1) on a 5056 × 6,000,000 bit matrix (transpose_matrix.cc:41-109), GB/s vs the HBM roofline."""
    rank, world, local, dist = dist_setup(args.gpus)
    from vcf2multialign_b200._lib import check, lib, ptr
    L = lib()
    check(L.v2m_cuda_init(local))
    rows, cols = 5056, 6_000_000
    n_words = rows * cols // 64
    rng = np.random.default_rng(3 + rank)
    words = rng.integers(0, 2**64, n_words, dtype=np.uint64)
    for _ in range(4):                                         # density ≈ 3 %; the traffic does not depend on it
        words &= rng.integers(0, 2**64, n_words, dtype=np.uint64)
    d_src, d_dst = L.v2m_cuda_device_alloc(n_words * 8), L.v2m_cuda_device_alloc(n_words * 8)
    if not d_src or not d_dst:
        raise RuntimeError("device_alloc failed")
    check(L.v2m_cuda_memcpy_h2d(d_src, ptr(words), n_words * 8))
    ms = C.c_float()
    check(L.v2m_cuda_time_transpose_bits(d_src, rows, cols, d_dst, max(args.warmup, 3), C.byref(ms)))
    launches0 = L.v2m_cuda_launch_count()
    barrier(dist, local)
    with ClockSampler(local) as clocks:
        check(L.v2m_cuda_time_transpose_bits(d_src, rows, cols, d_dst, args.steps, C.byref(ms)))
    barrier(dist, local)
    launches = L.v2m_cuda_launch_count() - launches0
    ms_step = max_over_ranks(dist, local, float(ms.value))
    nbytes = 2 * n_words * 8
    gbs = world * nbytes / (ms_step * 1e-3) / 1e9
    peak, peak_src = measured_peak_gbs()
    # involution on the device result (parity with the oracle on this kernel: tests/test_gpu_parity.py)
    d_back = L.v2m_cuda_device_alloc(n_words * 8)
    check(L.v2m_cuda_transpose_bits_device(d_dst, cols, rows, d_back, None))
    back = np.zeros(n_words, np.uint64)
    check(L.v2m_cuda_memcpy_d2h(ptr(back), d_back, n_words * 8))
    ok = bool(np.array_equal(back, words))
    if rank == 0:
        emit_line({
            "metric": "transpose_matrix_gb_per_s", "value": gbs, "unit": "GB/s", "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
            "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u64", "data": "synthetic",
            "config": {"workload": f"c5: transpose_matrix microbench, {rows} x {cols} path-edge bit matrix, density ~3 %", "replicas": world,
                       "l2": "no flush needed: each launch reads and writes 3.79 GB (>> 126 MB L2)"},
            "roofline": {"bound": "hbm", "kernel": "transpose_bits_kernel", "achieved": nbytes / (float(ms.value) * 1e-3) / 1e9, "peak": peak, "unit": "GB/s",
                         "frac": nbytes / (float(ms.value) * 1e-3) / 1e9 / peak, "traffic": None, "peak_source": peak_src, "algorithmic_bytes_per_launch": nbytes},
            "cpu_baseline": None, "e2e": None, "gpu_launches": int(launches), "clocks": clocks.summary(), "involution_ok": ok,
        })
    if dist is not None:
        dist.destroy_process_group()
    return 0 if ok else 1


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--workload", default="c4")
    ap.add_argument("--scaling", default="strong", choices=["strong", "weak"],
                    help="N > 1: strong = one job's rows sharded over the ranks (default); weak = one cohort per rank")
    ap.add_argument("--impl", default="b200")
    ap.add_argument("--block-rows", type=int, default=0,
                    help="rows per emit_rows launch at most (the HBM ring holds two blocks); 0: up to 512 rows, at most 32 GiB per block")
    ap.add_argument("--e2e-steps", type=int, default=1)
    ap.add_argument("--cpu-seconds", type=float, default=15.0)
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-kernels", action="store_true", help="skip the side measurements of K1 / K2a in the `kernels` block")
    args = ap.parse_args()
    claim_stdout()
    args.warmup = max(args.warmup, 3) if args.impl != "reference" else args.warmup

    if args.impl == "reference":
        return cpu_reference_arm(args)
    if args.workload == "c5":
        return transpose_microbench(args)

    rank, world, local, dist = dist_setup(args.gpus)
    import vcf2multialign_b200 as v2m
    from vcf2multialign_b200 import sharding, synth
    from vcf2multialign_b200._lib import check, lib, ptr
    L = lib()
    check(L.v2m_cuda_init(local))

    # ---- workload + residency (untimed) -----------------------------------------------------------------------------
    w = build_workload(args.workload, rank, scaling=args.scaling)
    g = w.graph
    H, A = g.total_chromosome_copies(), g.aligned_length()
    path_bytes = g.e_pad * g.h_pad // 8
    d_paths = L.v2m_cuda_device_alloc(path_bytes)
    if not d_paths:
        raise v2m.V2MError("device_alloc for path bits failed")
    synth.path_bits_device(w, d_paths)
    dgraph = v2m.DeviceGraph(g, w.ref, d_paths_by_edge_and_chrom_copy=d_paths)
    rows_all = np.concatenate([[v2m.PLOIDY_MAX], np.arange(H)]).astype(np.uint32)      # haplotype_output.cc:48-81: REF row, then every copy
    if args.scaling == "strong":
        row_begin, row_end = sharding.row_range(rows_all.size, rank, world)
        total_rows = int(rows_all.size)
    else:
        row_begin, row_end = 0, int(rows_all.size)
        total_rows = world * int(rows_all.size)
    rows_idx = np.ascontiguousarray(rows_all[row_begin:row_end])
    plan = dgraph.plan_haplotypes(rows_idx)
    n_rows = int(rows_idx.size)
    pitch = int(L.v2m_cuda_row_pitch(A))      # rows start on a 128-byte line, as in the library's own buffers
    max_block = args.block_rows if args.block_rows > 0 else min(512, max(64, (32 << 30) // pitch // 64 * 64))
    n_launches = max(1, (n_rows + max_block - 1) // max_block)
    block = (n_rows + n_launches - 1) // n_launches            # launches sized to the shard: no short tail launch
    ring = [L.v2m_cuda_device_alloc(block * pitch) for _ in range(2)]
    if not all(ring):
        raise v2m.V2MError("device_alloc for the row ring failed")

    def one_step(ev=None):
        if ev:
            check(L.v2m_cuda_event_record(ev[0], None))
        check(L.v2m_cuda_plan_resolve(plan._p, None))                                   # K2a
        if ev:
            check(L.v2m_cuda_event_record(ev[1], None))
        k = 0
        for r0 in range(0, n_rows, block):                                               # K2b
            check(L.v2m_cuda_emit_rows_device(plan._p, r0, min(n_rows, r0 + block), 0, A, ring[k], pitch, None))
            k ^= 1

    for _ in range(args.warmup):
        one_step()
    check(L.v2m_cuda_synchronize())

    # ---- timed: device resident -----------------------------------------------------------------------------------
    events = [[L.v2m_cuda_event_create() for _ in range(2)] for _ in range(args.steps)]
    e_end = L.v2m_cuda_event_create()
    launches0 = L.v2m_cuda_launch_count()
    barrier(dist, local)
    with ClockSampler(local) as clocks:
        for s in range(args.steps):
            one_step(events[s])
        check(L.v2m_cuda_event_record(e_end, None))
        ms = C.c_float()
        check(L.v2m_cuda_event_elapsed_ms(events[0][0], e_end, C.byref(ms)))
        check(L.v2m_cuda_synchronize())
    barrier(dist, local)
    launches = L.v2m_cuda_launch_count() - launches0

    def elapsed(a, b):
        x = C.c_float()
        check(L.v2m_cuda_event_elapsed_ms(a, b, C.byref(x)))
        return float(x.value)

    resolve_ms = sum(elapsed(events[s][0], events[s][1]) for s in range(args.steps)) / args.steps
    emit_ms = sum(elapsed(events[s][1], events[s + 1][0] if s + 1 < args.steps else e_end) for s in range(args.steps)) / args.steps
    ms_total = max_over_ranks(dist, local, float(ms.value))
    ms_per_step = ms_total / args.steps
    value = total_rows * A / (ms_per_step * 1e-3) / 1e9

    # Dominant kernel = emit_rows_kernel: algorithmic bytes = rows × A written, over its own share of the step (events around it).
    peak, peak_src = measured_peak_gbs()
    achieved = n_rows * A / (emit_ms * 1e-3) / 1e9            # this rank's GB/s
    traffic = None
    try:
        with open(os.path.join(ROOT, "profiles", "emit_rows_traffic.json")) as f:
            tj = json.load(f)
            if tj.get("workload") == args.workload and tj.get("rows_per_launch") == min(block, n_rows):
                traffic = tj.get("dram_bytes_per_launch")
    except Exception:
        pass
    roofline = {"bound": "hbm", "kernel": "emit_rows_kernel", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                "traffic": traffic, "peak_source": peak_src,
                "note": "peak is the measured copy (read + write) bandwidth; a write-only stream reaches 7.38 TB/s on this GPU (cudaMemset, profiles/r01_write_ceiling.txt), so frac of this write-only kernel can exceed 1",
                "algorithmic_bytes_per_launch": min(block, n_rows) * A,
                "mean_algorithmic_bytes_per_launch": n_rows * A / n_launches,
                "launches_per_step": n_launches, "avg_launch_ms": emit_ms / n_launches,
                "step_frac": n_rows * A / (float(ms.value) / args.steps * 1e-3) / 1e9 / peak}      # K2a + K2b against the same peak

    # ---- side measurements: the other kernels of the path on this workload's own shapes ---------------------------------
    kernels = {"emit_rows_kernel": {"ms_per_step": emit_ms, "gbs": achieved, "frac": achieved / peak, "bytes": n_rows * A},
               "resolve_rows_kernel": {"ms_per_step": resolve_ms, "gbs": 2 * n_rows * g.e_pad / 8 / (resolve_ms * 1e-3) / 1e9,
                                       "frac": 2 * n_rows * g.e_pad / 8 / (resolve_ms * 1e-3) / 1e9 / peak, "bytes": 2 * n_rows * g.e_pad // 8,
                                       "note": "rows x E_pad/8 path bits read + the same taken bits written; inside the timed step"}}
    if not args.no_kernels:
        d_scratch = L.v2m_cuda_device_alloc(path_bytes)
        if d_scratch:
            t = C.c_float()
            check(L.v2m_cuda_time_transpose_bits(d_paths, g.h_pad, g.e_pad, d_scratch, 3, C.byref(t)))
            check(L.v2m_cuda_time_transpose_bits(d_paths, g.h_pad, g.e_pad, d_scratch, 10, C.byref(t)))
            # the launcher's choice for this shape (csrc/transpose_bits.cu): the wide kernel where destination lines are 8 mod 16 bytes long
            k1_name = "transpose_bits_wide_kernel" if (g.e_pad // 8) % 16 and g.h_pad > 512 else "transpose_bits_tall_kernel" if g.h_pad <= 512 else "transpose_bits_kernel"
            kernels[k1_name] = {"shape": [int(g.h_pad), int(g.e_pad)], "ms": float(t.value), "gbs": 2 * path_bytes / (t.value * 1e-3) / 1e9,
                                                "frac": 2 * path_bytes / (t.value * 1e-3) / 1e9 / peak, "bytes": 2 * path_bytes,
                                                "note": "K1 on this workload's path matrix (once per graph, outside the step); --workload c5 for the 5056 x 6,000,000 microbench"}
            L.v2m_cuda_device_free(d_scratch)

    # ---- timed: end to end through host buffers ---------------------------------------------------------------------
    e2e = None
    if not args.no_e2e:
        e2e_block = int(max(1, min(64, n_rows, (2 << 30) // pitch)))
        h_out = L.v2m_cuda_host_alloc(e2e_block * pitch)
        if not h_out:
            raise v2m.V2MError("pinned host allocation failed")
        # Host copies of everything the C-ABI uploads (path bits fetched back once, untimed).
        paths_ec = np.zeros(path_bytes // 8, np.uint64)
        check(L.v2m_cuda_memcpy_d2h(ptr(paths_ec), d_paths, path_bytes))
        g.paths_by_edge_and_chrom_copy = paths_ec
        h2d = int(paths_ec.nbytes + w.ref.nbytes + g.reference_positions.nbytes + g.aligned_positions.nbytes + g.alt_edge_targets.nbytes
                  + g.alt_edge_count_csum.nbytes + g.label_offsets.nbytes + len(g.label_chars) + rows_idx.nbytes)

        # N > 1, strong scaling: the rows of the ONE job are handed out in blocks as the ranks ask for them (sharding.RowBlockQueue:
        # a counter under /dev/shm, control plane only). The host link is not the same for every GPU of a box — with all eight
        # copying at once a plain D2H copy runs at 8 to 18 GB/s per GPU here — so equal static row ranges leave the stream waiting
        # for the rank behind the slowest link (profiles/r02_n_bench_c4_n8_static_shards.json: 88 Gbase/s). Any rank can produce
        # any row block, so every rank uploads the whole graph and plans every row.
        dynamic = args.scaling == "strong" and world > 1
        step_rows = rows_all if dynamic else rows_idx
        n_step_rows = int(step_rows.size)
        queue = None
        if dynamic:
            queue = sharding.RowBlockQueue(f"/dev/shm/v2m_bench_rows_{os.environ.get('MASTER_PORT', '0')}", n_step_rows, e2e_block)
            h2d += int(rows_all.nbytes - rows_idx.nbytes) - int(paths_ec.nbytes) * (world - 1) // world
        setup_s, rows_done = [], []

        def e2e_step(row_limit=None):
            t_step = time.perf_counter()
            if dynamic:
                # one pass of the path matrix over the host link per box: every rank uploads 1 / N of it and the ranks exchange the
                # slices over NVLink; everything else of the graph (0.6 GB) goes up from every rank's own host arrays
                t_paths = sharding.replicate_path_matrix(dist, local, paths_ec, path_bytes // 8)
                if row_limit is not None:                         # warm-up step: the replica is the host matrix (wrapping 64-bit sums agree)
                    assert int(t_paths.sum().item()) == int(paths_ec.view(np.int64).sum()), "replicated path matrix differs from the host matrix"
                dg = v2m.DeviceGraph(g, w.ref, d_paths_by_edge_and_chrom_copy=t_paths.data_ptr())
                del t_paths
            else:
                dg = v2m.DeviceGraph(g, w.ref)                    # H2D of the graph + K1 transpose + template
            pl = dg.plan_haplotypes(step_rows)                    # K2a
            setup_s.append(time.perf_counter() - t_step)
            done = 0
            if row_limit is not None or not dynamic:
                for r0 in range(0, row_limit or n_step_rows, e2e_block):                                   # K2b → pinned host
                    check(L.v2m_cuda_emit_rows(pl._p, r0, min(n_step_rows, r0 + e2e_block), 0, A, h_out, pitch))
                    done += min(n_step_rows, r0 + e2e_block) - r0
            else:
                while True:
                    blk = queue.next()
                    if blk is None:
                        break
                    check(L.v2m_cuda_emit_rows(pl._p, blk[0], blk[1], 0, A, h_out, pitch))
                    done += blk[1] - blk[0]
            rows_done.append(done)
            pl.close()
            dg.close()

        e2e_step(min(n_step_rows, 2 * e2e_block))                 # warm-up: every allocation, page-locking and code path of a step, two blocks of rows
        t_e2e = 0.0
        for _ in range(args.e2e_steps):
            if dynamic:
                barrier(dist, local)
                if rank == 0:
                    queue.reset()
            barrier(dist, local)
            t0 = time.perf_counter()
            e2e_step()
            check(L.v2m_cuda_synchronize())
            t_e2e += max_over_ranks(dist, local, time.perf_counter() - t0)
        barrier(dist, local)
        if queue is not None:
            queue.close()
        d2h = int(rows_done[-1] * A)
        # The host link's own ceiling on this box: plain D2H copies of the same row blocks (256 MiB pieces, two streams), every rank
        # copying at once for ~3 s; the ceiling is the sum of the ranks' rates.
        copy_bytes = int(min(e2e_block, block) * pitch)
        secs = C.c_double()
        barrier(dist, local)
        check(L.v2m_cuda_time_d2h_copies(h_out, ring[0], copy_bytes, 256 << 20, 2, C.byref(secs)))      # calibration, all ranks at once
        iters = int(max(2, min(400, 3.0 / max(secs.value / 2, 1e-4))))
        barrier(dist, local)
        check(L.v2m_cuda_time_d2h_copies(h_out, ring[0], copy_bytes, 256 << 20, iters, C.byref(secs)))
        rank_gbs = copy_bytes * iters / float(secs.value) / 1e9
        link_peak = sum_over_ranks(dist, local, rank_gbs)
        barrier(dist, local)
        e2e_value = total_rows * A * args.e2e_steps / t_e2e / 1e9
        e2e = {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
               "steps": args.e2e_steps, "s_per_step": t_e2e / args.e2e_steps, "setup_s_per_step": float(np.mean(setup_s[1:])) if len(setup_s) > 1 else None,
               "peak": link_peak, "peak_unit": "GB/s", "frac": e2e_value / link_peak,
               "peak_source": f"measured live on this box: plain cudaMemcpyAsync D2H into pinned memory, {copy_bytes} B per pass in 256 MiB pieces on two streams, "
                              f"~3 s per rank, all {world} rank(s) copying at once, sum of the ranks' rates (rank 0: {rank_gbs:.1f} GB/s); 1 byte crosses the link per base",
               "row_assignment": (f"dynamic: blocks of {e2e_block} rows drawn from a shared counter by {world} ranks (rank 0 took {rows_done[-1]} rows)" if dynamic
                                  else "static: this rank's rows in order"),
               "note": "setup_s_per_step = graph upload (host arrays" + ("; every rank uploads 1 / N of the path matrix and the slices are exchanged over NVLink" if dynamic else "") + ") + K1 + K2a; then K2b with the rows copied into pinned host memory; full job per step; byte counts are rank 0's"}
        L.v2m_cuda_host_free(h_out)
        g.paths_by_edge_and_chrom_copy = None

    # ---- CPU baseline (rank 0, N = 1 only): oracle port, 1 thread, bounded row sample ---------------------------------
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu:
        We = g.e_pad // 64
        n_cpu_rows = 8
        lines = np.zeros(n_cpu_rows * We, np.uint64)
        check(L.v2m_cuda_graph_download_path_lines(dgraph._h, 0, n_cpu_rows, ptr(lines)))
        og = oracle_subcohort(w, 0, n_cpu_rows, lines)
        s1, b1 = og.time_rows(0, 1, 1, bulk=False)
        rows_for_budget = int(max(1, min(n_cpu_rows, args.cpu_seconds / max(s1, 1e-3))))
        s, b = og.time_rows(0, rows_for_budget, 1, bulk=False)
        sb, bb = og.time_rows(0, rows_for_budget, 1, bulk=True)
        cpu = {"value": b / s / 1e9, "unit": UNIT, "cores": 1, "kind": "port",
               "sample": f"{rows_for_budget} haplotype rows x {A} columns, oracle output_sequence, 1 thread, faithful I/O shape (4 KiB buffer, per-char gaps)",
               "bulk_io_value": bb / sb / 1e9}
        del og

    if rank == 0:
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_per_step,
            "higher_is_better": True, "scaling": args.scaling, "vs_baseline": None, "dtype": "u8", "data": "synthetic",
            "config": {"workload": workload_name(args.workload, g, A), "total_rows": int(total_rows), "rows_per_rank": int(n_rows), "block_rows": int(block),
                       "step": "K2a resolve_rows + K2b emit_rows over this rank's rows (plan resident, path bits resident)",
                       "sharding": ("one job: rows sharded by contiguous haplotype-row range per rank, no collective" if args.scaling == "strong"
                                    else "one 5,008-haplotype cohort per rank, no collective"),
                       "l2": "no flush needed: each step writes rows x A bytes (>> 126 MB L2) and re-reads < 10 GB"},
            "roofline": roofline, "kernels": kernels, "cpu_baseline": cpu, "e2e": e2e, "gpu_launches": int(launches), "clocks": clocks.summary(),
        }
        emit_line(line)
    if dist is not None:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
