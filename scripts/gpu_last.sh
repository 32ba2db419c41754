#!/bin/bash
mkdir -p gpurun_out
timeout 120 python -m pytest tests/test_gpu_trainer.py -m gpu -q -x --timeout 100 --timeout-method thread 2>&1 | tail -n 2
timeout 100 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -n 1
