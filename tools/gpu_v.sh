#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
O=gpurun_out; mkdir -p $O
timeout 120 tools/fp32_peak > $O/v_fp32_peak.json 2>&1
nvidia-smi --query-gpu=clocks.sm,clocks.max.sm,power.draw --format=csv,noheader >> $O/v_fp32_peak.json
timeout 1500 python -m pytest tests -m gpu -x -q > $O/v_pytest.log 2>&1
tail -n 6 $O/v_pytest.log; cat $O/v_fp32_peak.json
