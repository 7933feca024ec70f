set -x
python -m pytest tests -m gpu -x -q > gpurun_out/gputest_s18.log 2>&1; tail -2 gpurun_out/gputest_s18.log
python bench.py > gpurun_out/r02b_bench_c3_1gpu.json 2> gpurun_out/bench_s18.err; echo rc=$?
python bench.py --config C5 --no-cpu-baseline > gpurun_out/r02b_bench_c5_1gpu.json 2>/dev/null
python bench.py --config C4 --no-cpu-baseline > gpurun_out/r02b_bench_c4_1gpu.json 2>/dev/null
python bench.py --config C2 --no-cpu-baseline > gpurun_out/r02b_bench_c2_1gpu.json 2>/dev/null
python bench.py --config C1 --no-cpu-baseline > gpurun_out/r02b_bench_c1_1gpu.json 2>/dev/null
python bench.py --config B32 --no-cpu-baseline > gpurun_out/r02b_bench_c3_b32.json 2>/dev/null
ncu --metrics gpu__time_duration.sum --clock-control none -c 4000 --csv --log-file gpurun_out/r02b_launches_c3.csv python bench.py --steps 2 --warmup 3 --graph 0 --no-cpu-baseline --no-e2e > gpurun_out/ncu_launch.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:gemm -o gpurun_out/r02b_gemm python tests/bench_ops.py gemm_once > gpurun_out/ncu_gemm.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:"attn|tn_reduce" -o gpurun_out/r02b_other python tests/bench_ops.py other_once > gpurun_out/ncu_other.log 2>&1
ls -la gpurun_out | tail -15
