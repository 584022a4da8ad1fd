#!/bin/bash
# scratch trip: fast concat kernel vs generic at the three decoder shapes of a 16-patch x 10-sample micro-batch
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
O=gpurun_out/$1
python -m pytest tests/test_memops_gpu.py -x -q -m gpu > ${O}_pytest_memops.txt 2>&1; tail -3 ${O}_pytest_memops.txt
for env in 0 1; do
  for shape in "160 10 128 256 128 64 32" "160 10 64 128 256 64 64" "160 10 32 64 512 128 128"; do
    if [ $env = 1 ]; then VU_NO_FAST_CONCAT=1 python scripts/prof_concat.py $shape 4 0; else python scripts/prof_concat.py $shape 4 0; fi
  done
done > ${O}_concat.log 2>&1
cat ${O}_concat.log
python bench.py --workload mc --no-cpu-baseline --no-gpu-eager > ${O}_bench_mc.json 2> ${O}_bench_mc.err; tail -c 600 ${O}_bench_mc.json; tail -3 ${O}_bench_mc.err
