#!/bin/bash
# round 2, GPU call A: full GPU test suite + the three bench workloads (no ncu)
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.max.sm,power.limit --format=csv > gpurun_out/a_gpu.txt 2>&1
timeout 900 python -m pytest tests -m gpu -q -rf -p no:cacheprovider > gpurun_out/a_pytest.log 2>&1
echo "pytest rc=$?" >> gpurun_out/a_pytest.log
tail -40 gpurun_out/a_pytest.log
timeout 600 python bench.py --steps 10 --warmup 3 > gpurun_out/a_bench_cfg4.json 2> gpurun_out/a_bench_cfg4.err; echo "cfg4 rc=$?"
timeout 600 python bench.py --workload cfg3 --steps 10 > gpurun_out/a_bench_cfg3_bf16.json 2> gpurun_out/a_bench_cfg3_bf16.err; echo "cfg3 rc=$?"
timeout 600 python bench.py --workload cfg3 --dtype f32 --steps 3 --seqs-per-gpu 16 --no-cpu-baseline > gpurun_out/a_bench_cfg3_f32.json 2> gpurun_out/a_bench_cfg3_f32.err; echo "cfg3 f32 rc=$?"
timeout 600 python bench.py --workload cfg5 --steps 5 --no-cpu-baseline > gpurun_out/a_bench_cfg5.json 2> gpurun_out/a_bench_cfg5.err; echo "cfg5 rc=$?"
tail -c 1500 gpurun_out/a_bench_cfg4.json; tail -5 gpurun_out/a_bench_cfg4.err
