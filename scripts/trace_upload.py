import sys, time, numpy as np
sys.path.insert(0, '.')
import vcf2multialign_b200 as v2m
from vcf2multialign_b200 import synth
from vcf2multialign_b200._lib import check, lib, ptr
L = lib(); check(L.v2m_cuda_init(0))
w = synth.make_workload("c4", seed=1, with_paths=False); g = w.graph
nb = g.e_pad * g.h_pad // 8
d = L.v2m_cuda_device_alloc(nb); synth.path_bits_device(w, d)
ec = np.zeros(nb // 8, np.uint64); check(L.v2m_cuda_memcpy_d2h(ptr(ec), d, nb)); L.v2m_cuda_device_free(d)
g.paths_by_edge_and_chrom_copy = ec
rows = np.concatenate([[v2m.PLOIDY_MAX], np.arange(g.total_chromosome_copies())]).astype(np.uint32)
for i in range(3):
    t0 = time.perf_counter(); dg = v2m.DeviceGraph(g, w.ref); t1 = time.perf_counter(); pl = dg.plan_haplotypes(rows); t2 = time.perf_counter()
    print(f"DeviceGraph {t1 - t0:.3f} s, plan {t2 - t1:.3f} s", file=sys.stderr)
    pl.close(); dg.close()
