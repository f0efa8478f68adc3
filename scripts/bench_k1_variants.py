#!/usr/bin/env python3
"""K1 kernels side by side on the shapes the path hands over (measurement infrastructure).

    python scripts/bench_k1_variants.py [--kernels default,0,1,3] [--iters 20]

V2M_K1_TALL picks the kernel (unset: the launcher's choice; 0: 512 x 1024 tiles; 1: tall tiles; 2: sheared; 3: wide).
Random bits made on the device (the kernels do no data-dependent work); time = mean of --iters launches, CUDA events inside
v2m_cuda_time_transpose_bits. One JSON line per (shape, kernel)."""
import argparse
import ctypes as C
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

SHAPES = [
    ("c5", 5056, 6_000_000),
    ("c4", 5056, 6_300_480),
    ("aligned", 5056, 6_000_128),
    ("c2", 5056, 1_155_008),
    ("c4 back (copy-major -> edge-major)", 6_300_480, 5056),
]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--kernels", default="default,0,1,3")
    ap.add_argument("--iters", type=int, default=20)
    ap.add_argument("--shapes", default="")
    ap.add_argument("--ld", default="0", help="V2M_K1_LD values to try (load flavour: 0 nc, 1 cg, 2 nc + L1::no_allocate, 3 lu)")
    ap.add_argument("--modes", default="0", help="V2M_K1_MODE values (kernel 6 ablations: 1 no stores, 2 no loads, 3 neither)")
    ap.add_argument("--env", default="", help="extra NAME=VALUE,... set for the whole run")
    args = ap.parse_args()
    import torch
    from vcf2multialign_b200._lib import check, lib
    L = lib()
    check(L.v2m_cuda_init(0))
    try:
        peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]
    except Exception:
        peak = 6650.0
    n_max = max(r * c // 64 for _, r, c in SHAPES)
    src = torch.randint(-2**62, 2**62, (n_max,), dtype=torch.int64, device="cuda")
    dst = torch.empty(n_max, dtype=torch.int64, device="cuda")
    torch.cuda.synchronize()
    wanted = [s for s in args.shapes.split(",") if s]
    for kv in [x for x in args.env.split(",") if x]:
        os.environ[kv.split("=")[0]] = kv.split("=")[1]
    for name, rows, cols in SHAPES:
        if wanted and name.split()[0] not in wanted:
            continue
        for k, ld, mode in [(k, ld, m) for k in args.kernels.split(",") for ld in args.ld.split(",") for m in args.modes.split(",")]:
            os.environ["V2M_K1_LD"] = ld
            os.environ["V2M_K1_MODE"] = mode
            if k == "default":
                os.environ.pop("V2M_K1_TALL", None)
            else:
                os.environ["V2M_K1_TALL"] = k
            ms = C.c_float()
            check(L.v2m_cuda_time_transpose_bits(src.data_ptr(), rows, cols, dst.data_ptr(), 3, C.byref(ms)))
            check(L.v2m_cuda_time_transpose_bits(src.data_ptr(), rows, cols, dst.data_ptr(), args.iters, C.byref(ms)))
            gbs = 2 * rows * cols / 8 / (ms.value * 1e-3) / 1e9
            print(json.dumps({"shape": name, "rows": rows, "cols": cols, "kernel": k, "ld": ld, "mode": mode, "ms": round(ms.value, 4), "gbs": round(gbs, 1), "frac": round(gbs / peak, 4)}), flush=True)


if __name__ == "__main__":
    main()
