#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
for shp in "16 32 32 256 256 3" "16 128 128 64 64 3" "16 256 256 224 64 3" "16 64 64 128 128 3" "16 16 16 512 512 3"; do
  python scripts/prof_wgrad.py $shp 6 >> gpurun_out/trip25_wgrad.log 2>&1
done
cat gpurun_out/trip25_wgrad.log
timeout 600 ncu --set full --clock-control none --import-source on -k regex:conv_wgrad -s 2 -c 1 -f -o gpurun_out/r1c_wgrad_256_32sq python scripts/prof_wgrad.py 16 32 32 256 256 3 4 > gpurun_out/trip25_ncu1.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:conv_wgrad -s 2 -c 1 -f -o gpurun_out/r1c_wgrad_64_128sq python scripts/prof_wgrad.py 16 128 128 64 64 3 4 > gpurun_out/trip25_ncu2.log 2>&1
ls -la gpurun_out/*.ncu-rep
