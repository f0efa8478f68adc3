#!/usr/bin/env python3
"""The whole-stream driver (haplotype_output::output_a2m → v2m_cuda_write_a2m) on c2 rows: bytes reaching a file descriptor.

    python scripts/bench_stream.py [--rows 256] [--tmpfs-rows 48]

Three sinks: /dev/null (write(2) costs nothing: device + host link only), a file on tmpfs (/dev/shm: write(2) is a page-cache
copy on the calling thread) and --unaligned to /dev/null. Reported as Gbase/s of sequence bytes (headers excluded)."""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--rows", type=int, default=256)
    ap.add_argument("--tmpfs-rows", type=int, default=48)
    args = ap.parse_args()
    import vcf2multialign_b200 as v2m
    from vcf2multialign_b200 import synth
    from vcf2multialign_b200._lib import check, lib
    L = lib()
    check(L.v2m_cuda_init(0))
    w = synth.make_workload("c2", seed=1, with_paths=False)
    g = w.graph
    A = g.aligned_length()
    d_paths = L.v2m_cuda_device_alloc(g.e_pad * g.h_pad // 8)
    synth.path_bits_device(w, d_paths)
    dg = v2m.DeviceGraph(g, w.ref, d_paths_by_edge_and_chrom_copy=d_paths)
    out = {"workload": f"c2 rows, A = {A}"}

    def run(n_rows, path, unaligned, excl=False):
        plan = dg.plan_haplotypes(np.arange(n_rows, dtype=np.uint32))
        ids = [f"S{r // 2}-{1 + r % 2}" for r in range(n_rows)]
        lens = plan.row_lengths(unaligned)
        fd = os.open(path, (os.O_RDWR | os.O_CREAT | os.O_TRUNC) if excl else os.O_WRONLY, 0o644)
        try:
            t0 = time.perf_counter()
            plan.write_a2m(ids, fd, unaligned)
            dt = time.perf_counter() - t0
        finally:
            os.close(fd)
        plan.close()
        return {"rows": n_rows, "seconds": dt, "gbase_per_s": float(lens.sum()) / dt / 1e9}

    run(8, "/dev/null", False)                                   # warm-up
    out["dev_null_aligned"] = run(args.rows, "/dev/null", False)
    out["dev_null_unaligned"] = run(args.rows, "/dev/null", True)
    if args.tmpfs_rows and os.path.isdir("/dev/shm"):
        p = "/dev/shm/v2m_stream_bench.a2m"
        try:
            st = os.statvfs("/dev/shm")
            if st.f_bavail * st.f_frsize > 2 * args.tmpfs_rows * A:
                out["tmpfs_aligned"] = run(args.tmpfs_rows, p, False, excl=True)
        finally:
            if os.path.exists(p):
                os.unlink(p)
    print(json.dumps(out))


if __name__ == "__main__":
    main()
