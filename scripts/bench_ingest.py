#!/usr/bin/env python3
"""K5 (genotype ingestion) timing: the "Paths" loop of build_variant_graph on a chr22-shaped block of records.

    python scripts/bench_ingest.py [--records 200000] [--copies 5008] [--iters 10] [--cpu-records 20000]

Synthetic GT block: 5 % of the records tri-allelic, ALT allele frequency ~3 %, 0.1 % null alleles; every ALT has an edge
(edge indices in record order). Algorithmic bytes of ingest_genotypes_kernel = 2 · records · copies read + E · H_pad / 8
written. The CPU figure is the oracle's restatement of the loop, 1 thread (as the reference), on a prefix of the block; the
device result on that prefix must be identical."""
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
    ap.add_argument("--records", type=int, default=200_000)
    ap.add_argument("--copies", type=int, default=5008)
    ap.add_argument("--iters", type=int, default=10)
    ap.add_argument("--cpu-records", type=int, default=20_000)
    args = ap.parse_args()
    import vcf2multialign_b200 as v2m
    from vcf2multialign_b200._lib import check, lib, ptr
    L = lib()
    check(L.v2m_cuda_init(0))
    R, H = args.records, args.copies
    rng = np.random.default_rng(11)
    n_alts = 1 + (rng.random(R) < 0.05).astype(np.int64)
    alt_offsets = np.concatenate([[0], np.cumsum(n_alts)]).astype(np.uint64)
    E = int(alt_offsets[-1])
    edges_by_alt = np.arange(E, dtype=np.uint64)
    # GT block, generated in slabs to bound host memory for the random draws
    gt = np.zeros((R, H), np.uint16)
    for r0 in range(0, R, 20_000):
        r1 = min(R, r0 + 20_000)
        u = rng.random((r1 - r0, H), dtype=np.float32)
        a = (u < 0.03).astype(np.uint16)
        a[(u < 0.01) & (n_alts[r0:r1, None] > 1)] = 2
        a[u > 0.999] = v2m.GT_NULL
        gt[r0:r1] = a
    ref_pos = np.cumsum(rng.integers(1, 90, R)).astype(np.uint64)
    edge_rec = np.repeat(np.arange(R), n_alts)
    edge_target_pos = (ref_pos[edge_rec] + rng.integers(1, 12, E).astype(np.uint64)).astype(np.uint64)
    h_pad, e_pad = (H + 63) // 64 * 64, (E + 63) // 64 * 64

    # device-resident kernel time
    d_gt = L.v2m_cuda_device_alloc(gt.nbytes)
    d_off = L.v2m_cuda_device_alloc(alt_offsets.nbytes)
    d_eba = L.v2m_cuda_device_alloc(edges_by_alt.nbytes)
    d_paths = L.v2m_cuda_device_alloc(h_pad * e_pad // 8)
    assert d_gt and d_off and d_eba and d_paths
    check(L.v2m_cuda_memcpy_h2d(d_gt, ptr(gt), gt.nbytes))
    check(L.v2m_cuda_memcpy_h2d(d_off, ptr(alt_offsets), alt_offsets.nbytes))
    check(L.v2m_cuda_memcpy_h2d(d_eba, ptr(edges_by_alt), edges_by_alt.nbytes))
    ms = C.c_float()
    check(L.v2m_cuda_time_ingest_genotypes(d_gt, R, H, d_off, d_eba, h_pad, e_pad, d_paths, 3, C.byref(ms)))
    check(L.v2m_cuda_time_ingest_genotypes(d_gt, R, H, d_off, d_eba, h_pad, e_pad, d_paths, args.iters, C.byref(ms)))
    alg_bytes = 2 * R * H + E * h_pad // 8
    try:
        peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]
    except Exception:
        peak = 6650.0
    gbs = alg_bytes / (ms.value * 1e-3) / 1e9
    out = {"kernel": "ingest_genotypes_kernel", "records": R, "copies": H, "edges": E, "algorithmic_bytes": alg_bytes, "ms": ms.value, "gbs": gbs,
           "peak_gbs": peak, "frac": gbs / peak, "genotypes_per_s": R * H / (ms.value * 1e-3)}
    dev_words = np.zeros(h_pad * e_pad // 64, np.uint64)
    check(L.v2m_cuda_memcpy_d2h(ptr(dev_words), d_paths, dev_words.nbytes))
    for d in (d_gt, d_off, d_eba, d_paths):
        L.v2m_cuda_device_free(d)

    # through the C ABI with host buffers (H2D of the GT block inside), with overlap detection
    # (twice: the first call page-locks the two staging buffers, which the calling thread keeps afterwards)
    times = []
    for _ in range(2):
        t0 = time.perf_counter()
        words, ov = v2m.paths_from_genotypes(gt, alt_offsets, edges_by_alt, h_pad, e_pad, ref_pos, edge_target_pos)
        times.append(time.perf_counter() - t0)
        assert np.array_equal(words, dev_words)
    dt = times[1]
    out["e2e"] = {"seconds": dt, "first_call_seconds": times[0], "genotypes_per_s": R * H / dt, "h2d_bytes": int(gt.nbytes), "overlaps": int(len(ov))}

    # both orientations born on the device (v2m_cuda_paths_from_genotypes_both) against the edge-major one alone, host GT block in,
    # matrices left on the device; the copy-major result must be the transpose of the edge-major one
    d_ec, d_ce = L.v2m_cuda_device_alloc(h_pad * e_pad // 8), L.v2m_cuda_device_alloc(h_pad * e_pad // 8)
    t_one, t_both = [], []
    for _ in range(3):
        t0 = time.perf_counter()
        check(L.v2m_cuda_paths_from_genotypes_device(ptr(gt), R, H, ptr(alt_offsets), ptr(edges_by_alt), None, None, 0, h_pad, e_pad, d_ec, None, 0, None))
        t_one.append(time.perf_counter() - t0)
        t0 = time.perf_counter()
        check(L.v2m_cuda_paths_from_genotypes_both(ptr(gt), R, H, ptr(alt_offsets), ptr(edges_by_alt), None, None, 0, h_pad, e_pad, d_ec, d_ce, None, 0, None))
        t_both.append(time.perf_counter() - t0)
    ce = np.zeros(h_pad * e_pad // 64, np.uint64)
    check(L.v2m_cuda_memcpy_d2h(ptr(ce), d_ce, ce.nbytes))
    d_t = L.v2m_cuda_device_alloc(h_pad * e_pad // 8)
    check(L.v2m_cuda_transpose_bits_device(d_ec, h_pad, e_pad, d_t, None))
    ce_k1 = np.zeros_like(ce)
    check(L.v2m_cuda_memcpy_d2h(ptr(ce_k1), d_t, ce.nbytes))
    assert np.array_equal(ce, ce_k1), "copy-major matrix from K5 differs from K1 of the edge-major one"
    ms_k1 = C.c_float()
    check(L.v2m_cuda_time_transpose_bits(d_ec, h_pad, e_pad, d_t, 5, C.byref(ms_k1)))
    out["both_orientations"] = {"edge_major_only_s": min(t_one), "both_s": min(t_both), "k1_whole_matrix_ms": ms_k1.value,
                                "device_resident_estimate_ms": ms.value + ms_k1.value,
                                "device_resident_estimate_gbs": (alg_bytes + E * h_pad // 8) / ((ms.value + ms_k1.value) * 1e-3) / 1e9,
                                "identical_to_k1": True}
    for d in (d_ec, d_ce, d_t):
        L.v2m_cuda_device_free(d)

    if args.cpu_records:
        from oracle import oracle as orc
        rc = min(args.cpu_records, R)
        ec = int(alt_offsets[rc])
        ep = (ec + 63) // 64 * 64
        dst = np.zeros(h_pad * ep // 64, np.uint64)
        n_ov = C.c_uint64(0)
        secs = orc.lib().orc_time_paths_loop(gt[:rc].ctypes.data_as(C.c_void_p), rc, H, alt_offsets.ctypes.data_as(C.c_void_p), edges_by_alt.ctypes.data_as(C.c_void_p),
                                             ref_pos.ctypes.data_as(C.c_void_p), edge_target_pos.ctypes.data_as(C.c_void_p), h_pad, ep,
                                             dst.ctypes.data_as(C.c_void_p), C.byref(n_ov))
        Wh = h_pad // 64
        assert np.array_equal(dst[: ec * Wh], dev_words[: ec * Wh]), "device path bits differ from the oracle on the prefix"
        assert int(n_ov.value) == int((ov[:, 0] < rc).sum()), "overlap counts differ on the prefix"
        out["cpu_oracle"] = {"records": rc, "seconds": secs, "genotypes_per_s": rc * H / secs, "threads": 1, "prefix_identical": True}
    print(json.dumps(out))


if __name__ == "__main__":
    main()
