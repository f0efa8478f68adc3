#!/bin/bash
# Kernel experiments: bench value of emit_rows for every built variant library (vcf2multialign_b200/variants/*.so).
#   scripts/exp_emit.sh [block_rows ...]
blocks=${@:-128}
units=${UNITS:-0}
for lib in vcf2multialign_b200/libv2m_cuda.so vcf2multialign_b200/variants/*.so; do
  [ -f "$lib" ] || continue
  for b in $blocks; do for ur in $units; do
    V2M_K2_UNIT_ROWS=$ur V2M_CUDA_LIB=$PWD/$lib timeout 300 python bench.py --steps 3 --warmup 3 --no-e2e --no-cpu --block-rows $b 2>/dev/null \
      | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$(basename $lib)', 'block', $b, 'unit', $ur, round(d['value'],1), round(d['roofline']['avg_launch_ms'],4))"
  done; done
done
