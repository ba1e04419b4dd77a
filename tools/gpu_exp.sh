#!/bin/bash
# One GPU-box visit for layout experiments: A/B of the N = 8192 search layouts (bit-identical check + timing), then the
# GPU parity tests with the default layout.  usage: tools/gpu_exp.sh <tag> [layouts...]
tag=${1:-exp}; shift
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/${tag}_smi.txt 2>&1
timeout 900 python tools/staged_check.py "$@" > gpurun_out/${tag}_ab.txt 2>&1; echo "ab rc=$?"
cat gpurun_out/${tag}_ab.txt
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/${tag}_pytest.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/${tag}_pytest.log
tail -5 gpurun_out/${tag}_pytest.log
