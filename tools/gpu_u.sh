#!/bin/bash
mkdir -p gpurun_out
O=gpurun_out
python -c "import __graft_entry__ as g; g.build()" > $O/u_build.log 2>&1
timeout 600 python -m pytest tests/test_gpu_model.py -x -q -m gpu -k "merge" > $O/u_pytest_merge.log 2>&1; echo "pytest merge rc=$?"
tail -12 $O/u_pytest_merge.log
