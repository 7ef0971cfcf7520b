#!/bin/bash
mkdir -p gpurun_out
run() { echo "=== $*"; env "$@" python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "default and (worm_fused_steps or config_a_fused_steps)" 2>&1 | grep -E "passed|failed|differs|Error" | cut -c1-400; }
run SPH_DIAG=0
