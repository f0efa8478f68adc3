#!/usr/bin/env python3
"""emit_rows at few rows per launch (what the host-facing drivers issue when a row block is sized by bytes: 256 MiB is one row
of c4): device time per launch and GB/s for 1 … 128 rows of the c4 (default) or c2 shape.

    python scripts/bench_small_launch.py [--workload c4] [--rows 1,2,4,8,16,32,64,128]"""
import argparse
import ctypes as C
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="c4")
    ap.add_argument("--rows", default="1,2,4,8,16,32,64,128")
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
    rows = [int(x) for x in args.rows.split(",")]
    plan = dg.plan_haplotypes(np.arange(max(rows), dtype=np.uint32))
    pitch = int(L.v2m_cuda_row_pitch(A))
    d_out = L.v2m_cuda_device_alloc(max(rows) * pitch)
    out = {"workload": args.workload, "A": A, "lib": os.environ.get("V2M_CUDA_LIB", "default"), "launches": []}
    for n in rows:
        ms = C.c_float()
        check(L.v2m_cuda_time_emit_rows(plan._p, 0, n, d_out, pitch, 3, C.byref(ms)))
        check(L.v2m_cuda_time_emit_rows(plan._p, 0, n, d_out, pitch, max(3, min(50, 400 // n)), C.byref(ms)))
        out["launches"].append({"rows": n, "ms": ms.value, "gbs": n * A / (ms.value * 1e-3) / 1e9})
    print(json.dumps(out))


if __name__ == "__main__":
    main()
