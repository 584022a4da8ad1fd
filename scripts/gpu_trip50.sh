#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_model_gpu.py -m gpu -q --timeout 600 -x -k "evaluate or ensemble or predict_paths" 2>&1 | tail -15
grep -i "evaluate\|ensemble" gpurun_out/parity_report.txt | tail -8
