#!/bin/bash
# GPU-box driver used with gpurun: scripts/gpu_round.sh <stage>...   (outputs under gpurun_out/)
#   test      pytest -m gpu
#   bench     full bench.py line
#   quick     bench.py without e2e / cpu legs
#   launches  ncu launch list of the quick bench
#   full:<re> ncu --set full of kernels matching <re> in the quick bench (NCU_SKIP launches of that kernel skipped, default 130)
mkdir -p gpurun_out
QUICK="python bench.py --steps 1 --warmup 3 --no-e2e --no-cpu --no-kernels"
for stage in "$@"; do
  case "$stage" in
    test) python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log; tail -15 gpurun_out/pytest_gpu.log ;;
    bench) python bench.py --steps 3 --warmup 3 > gpurun_out/bench.log 2>&1; echo "bench exit $?" >> gpurun_out/bench.log; tail -3 gpurun_out/bench.log ;;
    quick) $QUICK ${QUICK_ARGS} > gpurun_out/quick.log 2>&1; echo "quick exit $?" >> gpurun_out/quick.log; tail -3 gpurun_out/quick.log ;;
    launches) $QUICK > gpurun_out/plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv $QUICK > gpurun_out/ncu_launches.log 2>&1 ;;
    full:*) re="${stage#full:}"; $QUICK > gpurun_out/plain2.log 2>&1 && ncu --set full --clock-control none --import-source on -k "regex:$re" -s ${NCU_SKIP:-130} -c 1 -f -o "gpurun_out/prof_$re" $QUICK > gpurun_out/ncu_full_$re.log 2>&1; tail -3 gpurun_out/ncu_full_$re.log ;;
    *) echo "unknown stage $stage" ;;
  esac
done
