#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
rm -f gpurun_out/parity_report.txt
timeout 1500 python -m pytest tests -m gpu -q --timeout 600 2>&1 | tail -40 > gpurun_out/trip6_tests.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/trip6_smoke.log 2>&1
# launch list of a short bench (batch 2)
timeout 600 python bench.py --steps 1 --warmup 3 --batch 2 --no-cpu-baseline > gpurun_out/trip6_plain.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -s 300 -c 400 --csv \
    --log-file gpurun_out/r1_launches_b2.csv python bench.py --steps 1 --warmup 3 --batch 2 --no-cpu-baseline > gpurun_out/trip6_ncu.log 2>&1
# full capture of the top kernel on its heaviest layer (decoder_blocks.3.conv1: 224->64 @256^2)
timeout 300 python scripts/prof_conv.py 16 256 256 224 64 3 4 > gpurun_out/trip6_conv64.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:conv_igemm -s 2 -c 1 \
    -o gpurun_out/r1_conv64 -f python scripts/prof_conv.py 16 256 256 224 64 3 4 > gpurun_out/trip6_ncu64.log 2>&1
timeout 300 python scripts/prof_conv.py 16 32 32 800 512 3 4 > gpurun_out/trip6_conv512.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:conv_igemm -s 2 -c 1 \
    -o gpurun_out/r1_conv512 -f python scripts/prof_conv.py 16 32 32 800 512 3 4 > gpurun_out/trip6_ncu512.log 2>&1
tail -8 gpurun_out/trip6_tests.log; tail -5 gpurun_out/trip6_smoke.log; cat gpurun_out/trip6_conv64.log gpurun_out/trip6_conv512.log
ls -la gpurun_out
