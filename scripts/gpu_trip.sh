#!/bin/bash
# One parametrised evidence script for every GPU trip of a round (replaces the per-trip scripts of round 1).
#   gpurun --timeout 1500 -- 'bash scripts/gpu_trip.sh <tag> <step> [<step> ...]'
# Every step writes gpurun_out/<tag>_<step>.* ; numbers printed under ncu are never bench values.
# Steps:
#   pytest            python -m pytest tests -m gpu -x -q
#   smoke             __graft_entry__.smoke()
#   bench             python bench.py (default workload: MC line + train object)
#   bench_mc / bench_train / bench_train_nostream (VU_TRAIN_STREAMS=0) / bench_unet / bench_cfg1 / bench_cfg4 / bench_cfg5
#   bench_ref         python bench.py --impl reference
#   ncu_train_list    ncu launch list (+ tensor pipe %, DRAM / L2 bytes) of one training step, serial streams
#   ncu_mc_list       ncu launch list of one MC step
#   ncu_full:<regex>:<script args...>   ncu --set full of kernels matching <regex> in `python scripts/<args>` (':' = space)
#   run:<script args...>                python scripts/<args> (':' = space), output to <tag>_run_<n>.log
#   sanitize:<tool>:<pytest -k expr>    compute-sanitizer --tool <tool> over the selected GPU tests
#   dist:<N>:<workload>                 torchrun on N GPUs: bench.py --gpus N --workload <workload>   (gpurun --gpus N)
#   distrun:<N>:<script args...>        torchrun on N GPUs: scripts/<args>
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
TAG=$1; shift
O=gpurun_out/$TAG
METRICS=gpu__time_duration.sum,sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed,dram__bytes_read.sum,dram__bytes_write.sum,lts__t_bytes.sum,l1tex__m_xbar2l1tex_read_bytes.sum
n=0
for step in "$@"; do
  n=$((n+1))
  echo "=== $step"
  case "$step" in
    pytest) timeout 1500 python -m pytest tests -m gpu -x -q > ${O}_pytest.txt 2>&1; tail -3 ${O}_pytest.txt ;;
    smoke) timeout 600 python -c "import __graft_entry__ as g; g.smoke()" > ${O}_smoke.log 2>&1; tail -3 ${O}_smoke.log ;;
    bench) timeout 1500 python bench.py > ${O}_bench.json 2> ${O}_bench.err; tail -c 1500 ${O}_bench.json ;;
    bench_mc) timeout 900 python bench.py --workload mc --no-cpu-baseline > ${O}_bench_mc.json 2> ${O}_bench_mc.err; tail -c 900 ${O}_bench_mc.json ;;
    bench_train) timeout 900 python bench.py --workload train --no-cpu-baseline > ${O}_bench_train.json 2> ${O}_bench_train.err; tail -c 1200 ${O}_bench_train.json ;;
    bench_train_nostream) VU_TRAIN_STREAMS=0 timeout 900 python bench.py --workload train --no-cpu-baseline > ${O}_bench_train_nostream.json 2> ${O}_bench_train_nostream.err; tail -c 1200 ${O}_bench_train_nostream.json ;;
    bench_unet|bench_cfg1|bench_cfg4|bench_cfg5) w=${step#bench_}; timeout 1200 python bench.py --workload $w > ${O}_bench_$w.json 2> ${O}_bench_$w.err; tail -c 1200 ${O}_bench_$w.json ;;
    bench_ref) timeout 1200 python bench.py --impl reference --steps 3 --warmup 1 > ${O}_bench_reference.json 2> ${O}_bench_reference.err; tail -c 600 ${O}_bench_reference.json ;;
    ncu_train_list)
      VU_TRAIN_STREAMS=0 timeout 1500 ncu --metrics $METRICS --clock-control none --profile-from-start off --csv \
        --log-file ${O}_train_launches.csv python scripts/train_prof_step.py > ${O}_ncu_train.log 2>&1
      python scripts/summarize_launches.py ${O}_train_launches.csv > ${O}_train_launch_summary.txt 2>&1; python scripts/launch_table.py ${O}_train_launches.csv > ${O}_train_launch_table.txt 2>&1; head -30 ${O}_train_launch_summary.txt ;;
    ncu_mc_list)
      timeout 1200 ncu --metrics $METRICS --clock-control none --csv --log-file ${O}_mc_launches.csv \
        python bench.py --steps 1 --warmup 3 --batch 32 --no-cpu-baseline --no-gpu-eager --workload mc > ${O}_ncu_mc.log 2>&1
      python scripts/mc_launch_summary.py ${O}_mc_launches.csv > ${O}_mc_launch_summary.txt 2>&1; head -16 ${O}_mc_launch_summary.txt ;;
    ncu_full:*)
      IFS=':' read -r _ regex rest <<< "$step"; args=${rest//:/ }
      timeout 1200 ncu --set full --clock-control none --import-source on -k regex:$regex -c 3 -f -o ${O}_full_$n \
        python scripts/$args > ${O}_ncu_full_$n.log 2>&1
      python scripts/ncu_summary.py ${O}_full_$n.ncu-rep > ${O}_full_$n.summary.txt 2>&1; head -40 ${O}_full_$n.summary.txt ;;
    run:*)
      args=${step#run:}; args=${args//:/ }
      timeout 1200 python scripts/$args > ${O}_run_$n.log 2>&1; tail -25 ${O}_run_$n.log ;;
    sanitize:*)
      IFS=':' read -r _ tool expr <<< "$step"
      timeout 2400 compute-sanitizer --tool $tool --print-limit 20 python -m pytest tests -m gpu -x -q -k "$expr" \
        > ${O}_sanitize_$tool.txt 2>&1; tail -15 ${O}_sanitize_$tool.txt ;;
    dist:*)
      IFS=':' read -r _ ng w <<< "$step"
      timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node $ng --master-addr 127.0.0.1 --master-port $((29600+n)) \
        bench.py --gpus $ng --workload $w > ${O}_bench_${w}_${ng}gpu.json 2> ${O}_bench_${w}_${ng}gpu.err
      grep '^{' ${O}_bench_${w}_${ng}gpu.json | tail -c 1500; tail -3 ${O}_bench_${w}_${ng}gpu.err ;;
    distrun:*)
      IFS=':' read -r _ ng rest <<< "$step"; args=${rest//:/ }
      timeout 1200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $ng --master-addr 127.0.0.1 --master-port $((29700+n)) \
        scripts/$args > ${O}_distrun_${ng}gpu_$n.log 2>&1; tail -25 ${O}_distrun_${ng}gpu_$n.log ;;
    *) echo "unknown step $step" ;;
  esac
done
