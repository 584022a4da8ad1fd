#!/bin/bash
# Round-1 evidence trip: GPU tests, smoke, default bench (+cpu baseline), reference arm, layer table,
# ncu launch list of a short bench, full ncu capture of the dominant kernel on its two heaviest layers.
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
rm -f gpurun_out/parity_report.txt
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/smi.txt 2>&1
timeout 1500 python -m pytest tests -m gpu -q --timeout 600 2>&1 | tail -40 > gpurun_out/trip7_tests.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/trip7_smoke.log 2>&1
timeout 900 python bench.py --steps 10 --warmup 3 --layers gpurun_out/layers_b8.csv > gpurun_out/trip7_bench.log 2>&1
echo "bench rc=$?" >> gpurun_out/trip7_bench.log
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/trip7_bench_ref.log 2>&1
timeout 600 python bench.py --steps 1 --warmup 3 --batch 2 --no-cpu-baseline > gpurun_out/trip7_plain.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -s 300 -c 400 --csv \
    --log-file gpurun_out/r1_launches_b2.csv python bench.py --steps 1 --warmup 3 --batch 2 --no-cpu-baseline > gpurun_out/trip7_ncu.log 2>&1
timeout 300 python scripts/prof_conv.py 16 256 256 224 64 3 4 > gpurun_out/trip7_conv64.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:conv_igemm -s 2 -c 1 \
    -o gpurun_out/r1_conv64 -f python scripts/prof_conv.py 16 256 256 224 64 3 4 > gpurun_out/trip7_ncu64.log 2>&1
timeout 300 python scripts/prof_conv.py 16 32 32 800 512 3 4 > gpurun_out/trip7_conv512.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:conv_igemm -s 2 -c 1 \
    -o gpurun_out/r1_conv512 -f python scripts/prof_conv.py 16 32 32 800 512 3 4 > gpurun_out/trip7_ncu512.log 2>&1
tail -8 gpurun_out/trip7_tests.log; tail -5 gpurun_out/trip7_smoke.log; tail -3 gpurun_out/trip7_bench.log; cat gpurun_out/trip7_conv64.log gpurun_out/trip7_conv512.log
ls -la gpurun_out
