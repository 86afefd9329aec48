set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests -q -m gpu > gpurun_out/g_pytest_1gpu.log 2>&1; echo "rc=$?" >> gpurun_out/g_pytest_1gpu.log
timeout 900 python bench.py > gpurun_out/g_bench_cfg2.json 2> gpurun_out/g_bench_cfg2.err; echo "rc=$?" >> gpurun_out/g_bench_cfg2.err
timeout 300 python profiles/solve_table.py cfg5 > gpurun_out/g_solve_cfg5.json 2> gpurun_out/g_solve.err
timeout 120 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/g_smoke.log 2>&1
tail -3 gpurun_out/g_pytest_1gpu.log; tail -c 400 gpurun_out/g_bench_cfg2.json; cat gpurun_out/g_solve_cfg5.json; tail -2 gpurun_out/g_smoke.log
