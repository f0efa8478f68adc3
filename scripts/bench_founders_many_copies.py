#!/usr/bin/env python3
"""Founder-mode sweeps (K4 + K3) against the number of chromosome copies, up to the 16,384 the sweep holds (measurement
infrastructure): device time per edge column for KW = 4 ... 16 words per warp on the same LD-structured synthetic sites.

    python scripts/bench_founders_many_copies.py [--sites 20000] [--samples 2048,4096,4300,6144,8192]

One JSON line per cohort size: K4 passes (local / compose / lists), K3 passes, microseconds per edge column per chunk."""
import argparse
import ctypes as C
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--sites", type=int, default=20000)
    ap.add_argument("--samples", default="2048,4096,4300,6144,8192")
    ap.add_argument("--founders", type=int, default=25)
    ap.add_argument("--min-distance", type=int, default=50)
    args = ap.parse_args()
    import vcf2multialign_b200 as v2m
    from vcf2multialign_b200 import synth
    from vcf2multialign_b200._lib import check, lib
    L = lib()
    check(L.v2m_cuda_init(0))
    for n_samples in [int(x) for x in args.samples.split(",")]:
        w = synth.make_workload("custom", seed=11, ref_len=50 * args.sites, n_sites=args.sites, n_samples=n_samples, with_paths=True)
        g = w.graph
        dg = v2m.DeviceGraph(g, w.ref)
        H, E = g.total_chromosome_copies(), g.edge_count()
        best = None
        for _ in range(3):
            t0 = time.perf_counter()
            cuts, score = dg.find_cut_positions(args.min_distance)
            dt = time.perf_counter() - t0
            t = dg.last_cut_positions_timing()
            if best is None or dt < best[0]:
                best = (dt, t, cuts, score)
        dt, t, cuts, score = best
        t0 = time.perf_counter()
        dg.find_matchings(cuts, args.founders)
        dm = time.perf_counter() - t0
        ms = (C.c_float * 3)()
        cnt = (C.c_uint64 * 4)()
        L.v2m_cuda_last_segment_classes_timing(ms, cnt)
        k4 = t["local_ms"] + t["compose_ms"] + t["lists_ms"]
        k3 = ms[0] + ms[1] + ms[2]
        chunks = max(int(t["chunks"]), 1)
        print(json.dumps({"copies": H, "words_per_warp": -(-((H + 31) // 32) // 32), "edges": E, "cuts": int(cuts.size), "score": int(score),
                          "find_cut_positions_s": round(dt, 4), "k4_local_ms": round(t["local_ms"], 3), "k4_compose_ms": round(t["compose_ms"], 3),
                          "k4_lists_ms": round(t["lists_ms"], 3), "k4_device_ms": round(k4, 3), "chunks": chunks,
                          "k4_us_per_edge_column_per_chunk": round(1e3 * t["local_ms"] / (E / chunks), 4),
                          "find_matchings_s": round(dm, 4), "k3_device_ms": round(k3, 3)}), flush=True)
        del dg


if __name__ == "__main__":
    main()
