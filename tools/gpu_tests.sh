#!/bin/bash
# all GPU tests + smoke, nothing else
timeout 900 python -m pytest tests -m gpu -q > gpurun_out/pytest_all.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/pytest_all.log | cut -c1-300
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke_all.log 2>&1; echo "smoke rc=$?"; tail -1 gpurun_out/smoke_all.log
