#!/usr/bin/env bash
# full GPU test-suite (no -x) + smoke; logs under gpurun_out/<tag>_*
TAG=${1:-run}
timeout 1500 python -m pytest tests -m gpu -q --timeout 600 > gpurun_out/${TAG}_pytest_gpu.log 2>&1
echo "pytest exit $?" >> gpurun_out/${TAG}_pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${TAG}_smoke.log 2>&1
echo "smoke exit $?" >> gpurun_out/${TAG}_smoke.log
grep -E "passed|failed|FAILED|exit" gpurun_out/${TAG}_pytest_gpu.log | tail -20; tail -2 gpurun_out/${TAG}_smoke.log
