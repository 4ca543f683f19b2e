#!/bin/bash
# graphed-vs-eager determinism check three times (prints the measured worst parameter error), then the whole GPU suite
mkdir -p gpurun_out
O=gpurun_out
python -c "import __graft_entry__ as g; g.build()" > $O/m_build.log 2>&1
for i in 1 2 3; do timeout 300 python -m pytest tests/test_gpu_model.py -x -q -m gpu -s -k "graphed_training_step" 2>&1 | grep -E "graphed vs eager|passed|failed|Error" ; done
timeout 1700 python -m pytest tests -x -q -m gpu > $O/m_pytest.log 2>&1; echo "pytest rc=$?"
tail -3 $O/m_pytest.log
