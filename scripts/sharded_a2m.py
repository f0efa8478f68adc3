#!/usr/bin/env python3
"""One A2M stream written by several GPUs: haplotype_output::output_a2m for the c1 workload, rows sharded by
range over the ranks of a torchrun launch, every rank writing its rows at their final offsets of one file; rank 0 then
compares the file byte for byte with the CPU oracle (test infrastructure) and prints one JSON line.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P scripts/sharded_a2m.py [--unaligned]
"""
import argparse
import json
import os
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--unaligned", action="store_true")
    ap.add_argument("--out", default="/dev/shm/v2m_sharded_c1.a2m")
    args = ap.parse_args()
    rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
    dist = None
    if world > 1:
        import torch
        import torch.distributed as dist_mod
        torch.cuda.set_device(local)
        dist_mod.init_process_group("nccl", device_id=torch.device("cuda", local))
        dist = dist_mod
    import vcf2multialign_b200 as v2m
    from vcf2multialign_b200 import sharding, synth
    v2m._lib.check(v2m.lib().v2m_cuda_init(local))
    w = synth.make_workload("c1", seed=1, with_paths=True)
    g = w.graph
    dg = v2m.DeviceGraph(g, w.ref)
    out = v2m.haplotype_output("chr1", True, args.unaligned)
    if rank == 0 and os.path.exists(args.out):
        os.unlink(args.out)
    if dist is not None:
        dist.barrier()
    t0 = time.perf_counter()
    sharding.sharded_haplotype_output_a2m(out, g, dg, args.out, rank, world, dist, block_rows=4)
    dt = time.perf_counter() - t0
    if rank == 0:
        from oracle import oracle as orc
        fasta, vcf = synth.to_vcf_text(w, "1")
        with tempfile.TemporaryDirectory() as d:
            open(os.path.join(d, "a.fa"), "w").write(fasta)
            open(os.path.join(d, "a.vcf"), "w").write(vcf)
            og = orc.Graph.from_files(os.path.join(d, "a.fa"), os.path.join(d, "a.vcf"), "1")
        exp = og.haplotype_output_a2m("chr1", True, args.unaligned)
        got = open(args.out, "rb").read()
        os.unlink(args.out)
        print(json.dumps({"workload": "c1", "ranks": world, "unaligned": args.unaligned, "bytes": len(got), "identical_to_oracle": got == exp, "seconds": dt}))
        if got != exp:
            sys.exit(1)
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
