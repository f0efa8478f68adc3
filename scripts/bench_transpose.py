#!/usr/bin/env python3
"""Config 5: transpose_matrix microbench — rows x cols path-edge bit matrix, GB/s against the HBM roofline.

    python scripts/bench_transpose.py [--rows 5056] [--cols 6000000] [--density 0.03] [--iters 20] [--cpu-cols 65536]

Algorithmic bytes = 2 * rows * cols / 8 (every bit read once and written once). The CPU figure is the oracle's
restatement of transpose_matrix.cc on a column sample of the same matrix (1 thread, as the reference)."""
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
    ap.add_argument("--rows", type=int, default=5056)
    ap.add_argument("--cols", type=int, default=6_000_000)
    ap.add_argument("--density", type=float, default=0.03)
    ap.add_argument("--iters", type=int, default=20)
    ap.add_argument("--cpu-cols", type=int, default=65536)
    args = ap.parse_args()
    from vcf2multialign_b200._lib import check, lib, ptr
    L = lib()
    check(L.v2m_cuda_init(0))
    rows, cols = args.rows, args.cols
    assert rows % 64 == 0 and cols % 64 == 0
    n_words = rows * cols // 64
    rng = np.random.default_rng(3)
    # density-d random bits: AND / OR of a few uniform words gets close enough; exact density is irrelevant to the traffic
    k = max(1, int(round(-np.log2(max(args.density, 1e-6)))))
    words = rng.integers(0, 2**64, n_words, dtype=np.uint64)
    for _ in range(k - 1):
        words &= rng.integers(0, 2**64, n_words, dtype=np.uint64)
    d_src, d_dst = L.v2m_cuda_device_alloc(n_words * 8), L.v2m_cuda_device_alloc(n_words * 8)
    check(L.v2m_cuda_memcpy_h2d(d_src, ptr(words), n_words * 8))
    ms = C.c_float()
    check(L.v2m_cuda_time_transpose_bits(d_src, rows, cols, d_dst, 3, C.byref(ms)))
    check(L.v2m_cuda_time_transpose_bits(d_src, rows, cols, d_dst, args.iters, C.byref(ms)))
    gbs = 2 * n_words * 8 / (ms.value * 1e-3) / 1e9
    # involution check on the device result: T(T(x)) == x
    d_back = L.v2m_cuda_device_alloc(n_words * 8)
    check(L.v2m_cuda_transpose_bits_device(d_dst, cols, rows, d_back, None))
    back = np.zeros(n_words, np.uint64)
    check(L.v2m_cuda_memcpy_d2h(ptr(back), d_back, n_words * 8))
    assert np.array_equal(back, words), "transpose involution failed"
    try:
        peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]
    except Exception:
        peak = 6650.0
    out = {"kernel": "transpose_bits_kernel", "rows": rows, "cols": cols, "bytes_per_launch": 2 * n_words * 8, "ms": ms.value, "gbs": gbs,
           "peak_gbs": peak, "frac": gbs / peak, "involution_ok": True}
    if args.cpu_cols:
        from oracle import oracle as orc
        cc = min(args.cpu_cols, cols) // 64 * 64
        sub = words[: rows * cc // 64]
        t = C.c_double()
        dst = np.zeros(sub.size, np.uint64)
        secs = orc.lib().orc_time_transpose(sub.ctypes.data_as(C.c_void_p), rows, cc, dst.ctypes.data_as(C.c_void_p))
        # parity on the sample: the first cc source columns are the first cc/64 words of every destination line
        got = np.zeros(n_words, np.uint64)
        check(L.v2m_cuda_memcpy_d2h(ptr(got), d_dst, n_words * 8))
        got_sub = got.reshape(rows, cols // 64)[:, : cc // 64].reshape(-1)
        assert np.array_equal(got_sub, dst), "GPU transpose differs from the oracle on the sample"
        out["cpu_oracle"] = {"cols": cc, "seconds": secs, "gbs": 2 * sub.size * 8 / secs / 1e9, "threads": 1, "sample_identical": True}
    print(json.dumps(out))


if __name__ == "__main__":
    main()
