#!/bin/bash
# round 2, GPU call R (1 GPU): full GPU suite + smoke on the final tree
mkdir -p gpurun_out
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r_smoke.log 2>&1; echo "smoke rc=$?"; tail -1 gpurun_out/r_smoke.log
timeout 1500 python -m pytest tests -m gpu -q -p no:cacheprovider > gpurun_out/r_pytest.log 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/r_pytest.log
