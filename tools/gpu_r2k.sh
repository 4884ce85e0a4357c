#!/bin/bash
# round 2, pass k (final tree): GPU parity suite, smoke(), default-workload bench line
mkdir -p gpurun_out
timeout 150 python -m pytest tests -m gpu -q > gpurun_out/r02_k_gpu_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/r02_k_gpu_pytest.log
grep -E "^FAILED|passed|failed|pytest exit" gpurun_out/r02_k_gpu_pytest.log | cut -c1-300
timeout 60 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/r02_k_smoke.log 2>&1; tail -n 1 gpurun_out/r02_k_smoke.log
timeout 100 python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/r02_k_bench_c5.json 2> gpurun_out/r02_k_bench_c5.err; echo "bench exit $?"; cut -c1-160 gpurun_out/r02_k_bench_c5.json
