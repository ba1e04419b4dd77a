mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/t1_pytest.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/t1_pytest.log
timeout 900 python bench.py > gpurun_out/t1_bench.json 2> gpurun_out/t1_bench.err; echo "bench rc=$?"; tail -3 gpurun_out/t1_bench.err; cat gpurun_out/t1_bench.json
