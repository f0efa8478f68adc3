#!/usr/bin/env python3
"""bench.py — A2M aligned Gbase/s of the row-materialisation hot path (BASELINE.json metric) on B200.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload c2] [--impl reference]

One "step" = one pass of the hot path (K2b emit_rows) over the rank's whole row set: the REF row + every chromosome copy of
the chr22-shaped synthetic workload (c2: 51 Mbp, 1.1 M sites, 5,008 haplotypes → 5,009 rows × A ≈ 51.2 M columns ≈ 256 GB
written per step). `value` is device-resident (rows written into an HBM ring of row blocks that is overwritten); `e2e` runs
the same job through the C-ABI with host buffers: graph arrays uploaded from host memory, rows copied back into pinned
host memory, all inside the timed region. N > 1 (torchrun): every rank owns the 5,008 haplotype rows of its own slice of
an N×5,008-haplotype cohort over the same sites — rows shard by haplotype range, no collective on the data path.

`--impl reference` times the reference's CPU implementation of the path — its restatement in oracle/ (the reference
itself needs Boost/Ragel, absent from this image) — on all host cores, on a bounded row sample of the same workload.
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


def max_over_ranks(dist, local, x: float) -> float:
    if dist is None:
        return x
    import torch
    t = torch.tensor([x], dtype=torch.float64, device=torch.device("cuda", local))
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def build_workload(name, rank, seed=1):
    from vcf2multialign_b200 import synth
    w = synth.make_workload(name, seed=seed, with_paths=False)
    # rank r owns chromosome copies [r·H, (r+1)·H) of the cohort: same sites, different haplotypes.
    w.seed = (seed + 7919 * rank) & 0xFFFFFFFF
    return w


def oracle_graph_for(w, paths_ec, paths_ce):
    from oracle import oracle as orc
    g = w.graph
    return orc.Graph.from_arrays(dict(
        reference_positions=g.reference_positions, aligned_positions=g.aligned_positions, alt_edge_targets=g.alt_edge_targets,
        alt_edge_count_csum=g.alt_edge_count_csum, label_offsets=g.label_offsets, label_chars=g.label_chars,
        paths_ec=paths_ec, paths_ce=paths_ce, h_pad=g.h_pad, e_pad=g.e_pad, ploidy_csum=g.ploidy_csum,
        sample_names=g.sample_names, ref=w.ref))


def cpu_reference_arm(args):
    """--impl reference: the CPU path (oracle port) with all host threads on a bounded sample of the workload's rows."""
    rank = int(os.environ.get("RANK", 0))
    if rank != 0:
        return 0
    from vcf2multialign_b200 import synth
    w = build_workload(args.workload, 0)
    g = w.graph
    cores = os.cpu_count() or 1
    # Path bits for the sampled chromosome copies only (copy-major lines), from the numpy model of the generator.
    n_rows = 2 * cores
    We = g.e_pad // 64
    n_alts = np.bincount(w.edge_site, minlength=len(w.site_pos)).astype(np.uint32)
    ce = np.zeros(g.h_pad * We, np.uint64)
    sites = w.edge_site.astype(np.uint64)
    for h in range(n_rows):
        al = synth.allele_numpy(w.seed, np.uint64(h), sites, n_alts[w.edge_site])
        idx = np.nonzero(al == (1 + w.edge_alt.astype(np.uint32)))[0]
        np.bitwise_or.at(ce, h * We + idx // 64, np.uint64(1) << (idx % 64).astype(np.uint64))
    og = oracle_graph_for(w, np.zeros(g.h_pad * We, np.uint64), ce)   # edge-major orientation is not used by the row writer
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
        "ms_per_step": 1e3 * t / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
        "config": {"workload": workload_name(args.workload, g, A)},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "port",
                         "sample": f"{n_rows} haplotype rows x {A} columns per step, row-parallel over {cores} threads, faithful I/O shape (4 KiB buffer, per-char gaps)"},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    emit_line(line)
    return 0


def workload_name(name, g, A):
    return f"{name}: synthetic {int(g.reference_positions[-1])} bp reference, {g.edge_count()} ALT edges, {g.total_chromosome_copies()} haplotypes + REF row, A={A} aligned columns"


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--workload", default="c2")
    ap.add_argument("--impl", default="b200")
    ap.add_argument("--block-rows", type=int, default=0,
                    help="rows per emit_rows launch (the HBM ring holds two blocks); 0: up to 512 rows, at most 32 GiB per block")
    ap.add_argument("--e2e-steps", type=int, default=1)
    ap.add_argument("--cpu-seconds", type=float, default=15.0)
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-cpu", action="store_true")
    args = ap.parse_args()
    claim_stdout()
    args.warmup = max(args.warmup, 3) if args.impl != "reference" else args.warmup

    if args.impl == "reference":
        return cpu_reference_arm(args)

    rank, world, local, dist = dist_setup(args.gpus)
    import vcf2multialign_b200 as v2m
    from vcf2multialign_b200 import synth
    from vcf2multialign_b200._lib import check, lib, ptr
    L = lib()
    check(L.v2m_cuda_init(local))

    # ---- workload + residency (untimed) -----------------------------------------------------------------------------
    w = build_workload(args.workload, rank)
    g = w.graph
    H, A = g.total_chromosome_copies(), g.aligned_length()
    path_bytes = g.e_pad * g.h_pad // 8
    d_paths = L.v2m_cuda_device_alloc(path_bytes)
    if not d_paths:
        raise v2m.V2MError("device_alloc for path bits failed")
    synth.path_bits_device(w, d_paths)
    dgraph = v2m.DeviceGraph(g, w.ref, d_paths_by_edge_and_chrom_copy=d_paths)
    rows_idx = np.concatenate([[v2m.PLOIDY_MAX], np.arange(H)]).astype(np.uint32)
    plan = dgraph.plan_haplotypes(rows_idx)
    n_rows = rows_idx.size
    pitch = (A + 127) // 128 * 128      # rows of the device ring start on a 128-byte line (as the library's own buffers do)
    if pitch % 4096 == 0:
        pitch += 128
    block_rows = args.block_rows if args.block_rows > 0 else min(512, max(64, (32 << 30) // pitch // 64 * 64))
    block = min(block_rows, n_rows)
    ring = [L.v2m_cuda_device_alloc(block * pitch) for _ in range(2)]
    if not all(ring):
        raise v2m.V2MError("device_alloc for the row ring failed")

    def one_step():
        k = 0
        for r0 in range(0, n_rows, block):
            check(L.v2m_cuda_emit_rows_device(plan._p, r0, min(n_rows, r0 + block), 0, A, ring[k], pitch, None))
            k ^= 1

    for _ in range(args.warmup):
        one_step()
    check(L.v2m_cuda_synchronize())

    # ---- timed: device resident -----------------------------------------------------------------------------------
    e0, e1 = L.v2m_cuda_event_create(), L.v2m_cuda_event_create()
    launches0 = L.v2m_cuda_launch_count()
    barrier(dist, local)
    with ClockSampler(local) as clocks:
        check(L.v2m_cuda_event_record(e0, None))
        for _ in range(args.steps):
            one_step()
        check(L.v2m_cuda_event_record(e1, None))
        ms = C.c_float()
        check(L.v2m_cuda_event_elapsed_ms(e0, e1, C.byref(ms)))
        check(L.v2m_cuda_synchronize())
    barrier(dist, local)
    launches = L.v2m_cuda_launch_count() - launches0
    ms_total = max_over_ranks(dist, local, float(ms.value))
    ms_per_step = ms_total / args.steps
    bases_per_step_all = world * n_rows * A
    value = bases_per_step_all / (ms_per_step * 1e-3) / 1e9

    # Dominant kernel = emit_rows_kernel (the only kernel in the step): algorithmic bytes = rows × A written per launch set.
    peak, peak_src = measured_peak_gbs()
    kernel_launches_per_step = (n_rows + block - 1) // block
    achieved = n_rows * A / (float(ms.value) / args.steps * 1e-3) / 1e9     # this rank's GB/s
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
                "note": "peak is the measured copy (read + write) bandwidth; a write-only stream reaches 7.38 TB/s on this GPU (cudaMemset, profiles/r01_write_ceiling.txt), so frac of this write-only kernel can exceed 1", "algorithmic_bytes_per_launch": min(block, n_rows) * A,
                "mean_algorithmic_bytes_per_launch": n_rows * A / kernel_launches_per_step,      # the last launch of a step is shorter
                "launches_per_step": kernel_launches_per_step, "avg_launch_ms": float(ms.value) / args.steps / kernel_launches_per_step}

    # ---- timed: end to end through host buffers ---------------------------------------------------------------------
    e2e = None
    if not args.no_e2e:
        e2e_block = min(64, n_rows)
        h_out = L.v2m_cuda_host_alloc(e2e_block * pitch)
        if not h_out:
            raise v2m.V2MError("pinned host allocation failed")
        # Host copies of everything the C-ABI uploads (path bits fetched back once, untimed).
        paths_ec = np.zeros(path_bytes // 8, np.uint64)
        check(L.v2m_cuda_memcpy_d2h(ptr(paths_ec), d_paths, path_bytes))
        g.paths_by_edge_and_chrom_copy = paths_ec
        h2d = int(paths_ec.nbytes + w.ref.nbytes + g.reference_positions.nbytes + g.aligned_positions.nbytes + g.alt_edge_targets.nbytes
                  + g.alt_edge_count_csum.nbytes + g.label_offsets.nbytes + len(g.label_chars) + rows_idx.nbytes)
        d2h = int(n_rows * A)

        def e2e_step():
            dg = v2m.DeviceGraph(g, w.ref)                       # H2D of the graph + K1 transpose + template
            pl = dg.plan_haplotypes(rows_idx)                     # K2a
            for r0 in range(0, n_rows, e2e_block):                # K2b → pinned host
                check(L.v2m_cuda_emit_rows(pl._p, r0, min(n_rows, r0 + e2e_block), 0, A, h_out, pitch))
            pl.close()
            dg.close()

        e2e_step()                                                # warm-up
        barrier(dist, local)
        t0 = time.perf_counter()
        for _ in range(args.e2e_steps):
            e2e_step()
        check(L.v2m_cuda_synchronize())
        t_e2e = max_over_ranks(dist, local, time.perf_counter() - t0)
        barrier(dist, local)
        e2e = {"value": world * n_rows * A * args.e2e_steps / t_e2e / 1e9, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
               "steps": args.e2e_steps, "s_per_step": t_e2e / args.e2e_steps,
               "note": "graph upload (host arrays) + K1 + K2a + K2b with rows copied into pinned host memory; full job per step"}
        L.v2m_cuda_host_free(h_out)
        g.paths_by_edge_and_chrom_copy = None

    # ---- CPU baseline (rank 0, N = 1 only): oracle port, 1 thread, bounded row sample ---------------------------------
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu:
        We = g.e_pad // 64
        n_cpu_rows = 8
        ce_all = dgraph.paths_by_chrom_copy_and_edge()
        ce = np.zeros_like(ce_all)
        ce[: n_cpu_rows * We] = ce_all[: n_cpu_rows * We]
        del ce_all
        og = oracle_graph_for(w, np.zeros(ce.size, np.uint64), ce)   # the row writer reads the copy-major orientation only
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
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
            "config": {"workload": workload_name(args.workload, g, A), "rows_per_rank": int(n_rows), "block_rows": int(block),
                       "sharding": "haplotype-row range per rank, no collective", "l2": "no flush needed: each step writes rows x A bytes (>> 126 MB L2) and re-reads < 1 GB"},
            "roofline": roofline, "cpu_baseline": cpu, "e2e": e2e, "gpu_launches": int(launches), "clocks": clocks.summary(),
        }
        emit_line(line)
    if dist is not None:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
