#!/bin/bash
# Final confirmation on ONE B200: GPU tests, smoke(), the default bench line, the reference arm. Outputs under gpurun_out/.
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r02_${TAG:-v4}_gputests.log 2>&1; echo "pytest exit $?" >> gpurun_out/r02_${TAG:-v4}_gputests.log
tail -3 gpurun_out/r02_${TAG:-v4}_gputests.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/r02_${TAG:-v4}_smoke.log 2>&1; echo "smoke rc=$?"
[ -n "$SKIP_REF" ] || timeout 400 python bench.py --impl reference > gpurun_out/r02_${TAG:-v4}_bench_reference.json 2> gpurun_out/r02_${TAG:-v4}_bench_reference.err; echo "reference rc=$?"
timeout 600 python bench.py > gpurun_out/r02_${TAG:-v4}_bench_n1.json 2> gpurun_out/r02_${TAG:-v4}_bench_n1.err; echo "bench rc=$?"
tail -2 gpurun_out/r02_${TAG:-v4}_smoke.log
