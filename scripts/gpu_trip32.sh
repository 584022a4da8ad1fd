#!/bin/bash
# round-1c evidence: full GPU tests, bench (MC + train), launch lists, ncu captures of the training hot kernels
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q --timeout 600 2>&1 | tail -6 > gpurun_out/r1c_tests.log
tail -3 gpurun_out/r1c_tests.log
timeout 900 python bench.py --steps 10 --warmup 3 > gpurun_out/r1c_bench.json 2> gpurun_out/r1c_bench.err
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r1c_bench_reference.json 2>> gpurun_out/r1c_bench.err
timeout 900 ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none --csv \
   --log-file gpurun_out/r1c_train_launches.csv python scripts/train_prof_step.py > gpurun_out/r1c_ncu_train.log 2>&1
python scripts/summarize_launches.py gpurun_out/r1c_train_launches.csv > gpurun_out/r1c_train_launch_summary.txt 2>/dev/null
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/r1c_mc_launches.csv \
   python bench.py --steps 1 --warmup 3 --no-cpu-baseline --workload mc > gpurun_out/r1c_ncu_mc.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:conv_wgrad_row3 -s 2 -c 1 -f -o gpurun_out/r1c_wgrad_row3_224to64_256sq python scripts/prof_wgrad.py 16 256 256 224 64 3 4 > gpurun_out/r1c_ncu1.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:conv_wgrad_kernel -s 2 -c 1 -f -o gpurun_out/r1c_wgrad_256to256_32sq python scripts/prof_wgrad.py 16 32 32 256 256 3 4 > gpurun_out/r1c_ncu2.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:bn_bwd_apply -s 20 -c 1 -f -o gpurun_out/r1c_bn_bwd_apply python scripts/train_prof_step.py 16 512 > gpurun_out/r1c_ncu3.log 2>&1
head -c 600 gpurun_out/r1c_bench.json; echo; head -20 gpurun_out/r1c_train_launch_summary.txt
