#!/usr/bin/env python3
"""The rows -> pinned host memory pipeline of v2m_cuda_emit_rows alone (no graph upload, no plan): GB/s against the plain-copy
ceiling of the same box, for a slice of the c4 (default) rows. V2M_STAGE_MIB sets the device staging budget per buffer.

    python scripts/bench_e2e_rows.py [--workload c4] [--rows 400] [--block 8]"""
import argparse
import ctypes as C
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="c4")
    ap.add_argument("--rows", type=int, default=400)
    ap.add_argument("--block", type=int, default=8)
    args = ap.parse_args()
    import vcf2multialign_b200 as v2m
    from vcf2multialign_b200 import synth
    from vcf2multialign_b200._lib import check, lib
    L = lib()
    check(L.v2m_cuda_init(0))
    w = synth.make_workload(args.workload, seed=1, with_paths=False)
    g = w.graph
    A = g.aligned_length()
    d_paths = L.v2m_cuda_device_alloc(g.e_pad * g.h_pad // 8)
    synth.path_bits_device(w, d_paths)
    dg = v2m.DeviceGraph(g, w.ref, d_paths_by_edge_and_chrom_copy=d_paths)
    n = args.rows
    plan = dg.plan_haplotypes(np.arange(n, dtype=np.uint32))
    pitch = int(L.v2m_cuda_row_pitch(A))
    h_out = L.v2m_cuda_host_alloc(args.block * pitch)
    d_src = L.v2m_cuda_device_alloc(args.block * pitch)
    secs = C.c_double()
    check(L.v2m_cuda_time_d2h_copies(h_out, d_src, args.block * pitch, 256 << 20, 2, C.byref(secs)))
    iters = max(2, int(3.0 / (secs.value / 2)))
    check(L.v2m_cuda_time_d2h_copies(h_out, d_src, args.block * pitch, 256 << 20, iters, C.byref(secs)))
    ceiling = args.block * pitch * iters / secs.value / 1e9

    def run():
        for r0 in range(0, n, args.block):
            check(L.v2m_cuda_emit_rows(plan._p, r0, min(n, r0 + args.block), 0, A, h_out, pitch))
    run()
    t0 = time.perf_counter()
    run()
    dt = time.perf_counter() - t0
    print(json.dumps({"workload": args.workload, "rows": n, "block_rows": args.block, "stage_mib": os.environ.get("V2M_STAGE_MIB", "256"),
                      "gbs": n * A / dt / 1e9, "ceiling_gbs": ceiling, "frac": n * A / dt / 1e9 / ceiling}))


if __name__ == "__main__":
    main()
