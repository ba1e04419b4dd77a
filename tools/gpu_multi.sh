#!/bin/bash
# multi-GPU visit: tools/gpu_multi.sh <N>  (runtime tests that need two devices, then the sharded bench at N GPUs)
n=${1:-2}
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_runtime.py -m gpu -x -q 2>&1 | tail -5
timeout 1200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $n --steps 3 --warmup 3 > gpurun_out/bench_n$n.json 2> gpurun_out/bench_n$n.err; echo "bench rc=$?"
tail -3 gpurun_out/bench_n$n.err | cut -c1-400; cat gpurun_out/bench_n$n.json | cut -c1-2500
