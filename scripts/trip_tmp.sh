#!/bin/bash
# scratch trip: stem split-TF32 parity margins + one-wave grids of the BatchNorm kernels (in-situ kernel sums)
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
O=gpurun_out/$1
python scripts/debug_mc700.py > ${O}_mc700.log 2>&1; cat ${O}_mc700.log | cut -c1-250
python -m pytest tests/test_memops_gpu.py -x -q -m gpu -k "stem or small_cin" 2>&1 | tail -2
python bench.py --workload mc --no-cpu-baseline --no-gpu-eager > ${O}_bench_mc.json 2> ${O}_bench_mc.err; tail -c 300 ${O}_bench_mc.json; tail -3 ${O}_bench_mc.err
for cfg in "8 16 16" "2 16 16" "8 3 16" "8 16 2" "2 3 2" "3 4 3" "4 6 4"; do
  set -- $cfg
  echo "== VU_BN_MULT_SUMS=$1 APPLY=$2 BWD=$3"
  VU_BN_MULT_SUMS=$1 VU_BN_MULT_APPLY=$2 VU_BN_MULT_BWD=$3 python scripts/train_kineto.py 2>&1 | grep "span\|channel_sums_kernel<1>\|bn_bwd_apply\|bn_apply"
done > ${O}_bn_grid.log 2>&1
cat ${O}_bn_grid.log
