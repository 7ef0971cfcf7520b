#!/bin/bash
# two-GPU tests (peer-memory links and NCCL transports between two processes) + the 2-GPU bench line
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_slab_multigpu.py -m gpu -x -q > gpurun_out/r02_v7_gputests_2gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r02_v7_gputests_2gpu.log
