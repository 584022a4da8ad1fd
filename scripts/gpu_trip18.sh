#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
for pol in 0 1 2; do
 for b in 8 16; do
  VU_BN_POLICY=$pol timeout 600 python bench.py --steps 6 --warmup 3 --no-cpu-baseline --batch $b --layers gpurun_out/layers_pol${pol}_b$b.csv > gpurun_out/trip18_pol${pol}_b$b.log 2>&1
  echo "pol $pol b $b: $(tail -1 gpurun_out/trip18_pol${pol}_b$b.log | cut -c1-90)"
 done
done
