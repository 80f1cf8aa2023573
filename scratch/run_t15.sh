#!/bin/bash
mkdir -p gpurun_out
timeout 60 python scratch/dbg9.py > gpurun_out/t15_dbg.log 2>&1; echo "dbg rc=$?"
timeout 300 python -m pytest tests/test_gpu_fused.py -q -x > gpurun_out/t15_pytest_fused.log 2>&1; echo "pytest fused rc=$?"
tail -5 gpurun_out/t15_pytest_fused.log
timeout 400 python -m pytest tests -m gpu -q -x --deselect tests/test_gpu_fused.py > gpurun_out/t15_pytest.log 2>&1; echo "pytest rest rc=$?"
tail -5 gpurun_out/t15_pytest.log
timeout 300 python bench.py --no-cpu > gpurun_out/t15_bench.json 2> gpurun_out/t15_bench.err; echo "bench rc=$?"
cat gpurun_out/t15_dbg.log | tail -8
cat gpurun_out/t15_bench.json | cut -c1-400
