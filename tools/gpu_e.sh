#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
O=gpurun_out; mkdir -p $O
timeout 1200 python -m pytest tests/test_gpu_parity.py tests/test_synth.py -m gpu -x -q -k "compact or hbm or built_tree or global_tree" > $O/e_pytest.log 2>&1; echo "pytest rc=$?" >> $O/e_pytest.log
timeout 900 python tools/sweep_c5.py 10000000 100000 4 full,compact "8388608,1024,16" "8388608,768,16" "8388608,512,16" "8388608,768,24" "8388608,512,24" > $O/e_sweep.log 2>&1
tail -n 5 $O/e_pytest.log; cat $O/e_sweep.log
