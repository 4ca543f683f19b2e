#!/bin/bash
# GEGLU epilogue bring-up: its tests under a short timeout (every barrier wait in the kernel is bounded)
mkdir -p gpurun_out
O=gpurun_out
python -c "import __graft_entry__ as g; g.build()" > $O/n_build.log 2>&1
timeout 300 python -m pytest tests/test_gpu_geglu.py -x -q -m gpu -s > $O/n_pytest_geglu.log 2>&1; echo "pytest geglu rc=$?"
tail -40 $O/n_pytest_geglu.log
