# Round-end evidence run (one GPU box): full GPU test suite, headline bench line, ncu --set full of one C3 layer's GEMM launches
# (-> tools/ncu_traffic.py -> profiles/gemm_traffic.json).  Outputs under gpurun_out/.
set -x
python -m pytest tests -m gpu -x -q > gpurun_out/gputest_final.log 2>&1; tail -2 gpurun_out/gputest_final.log
python bench.py > gpurun_out/r02b_bench_c3_final.json 2> gpurun_out/bench_final.err; echo rc=$?
ncu --set full --clock-control none --import-source on -k regex:gemm -o gpurun_out/r02b_gemm python tests/bench_ops.py gemm_once > gpurun_out/ncu_gemm.log 2>&1
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -3
