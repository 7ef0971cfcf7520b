#!/bin/bash
mkdir -p gpurun_out
timeout 420 python -m pytest tests -m gpu -x -q > gpurun_out/b14_tests.log 2>&1; echo "pytest exit $?" >> gpurun_out/b14_tests.log; tail -3 gpurun_out/b14_tests.log
timeout 120 python tools/late_profile.py dambreak_1m 1200
timeout 60 python tools/late_profile.py dambreak_1m 0
timeout 60 python tools/late_profile.py config_a 0
