set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests -q -m gpu > gpurun_out/c_pytest_1gpu.log 2>&1; echo "rc=$?" >> gpurun_out/c_pytest_1gpu.log
timeout 900 python bench.py > gpurun_out/c_bench_cfg2.json 2> gpurun_out/c_bench_cfg2.err; echo "rc=$?" >> gpurun_out/c_bench_cfg2.err
timeout 300 python profiles/solve_table.py cfg2 > gpurun_out/c_solve.json 2> gpurun_out/c_solve.err
timeout 300 python profiles/solve_table.py cfg5 > gpurun_out/c_solve_cfg5.json 2>> gpurun_out/c_solve.err
timeout 900 ncu --set full --clock-control none -k regex:k_solve_.*256 --launch-skip 82 --launch-count 120 -o /tmp/solve_cap python profiles/solve_table.py cfg2 > gpurun_out/c_ncu.log 2>&1
ncu -i /tmp/solve_cap.ncu-rep --page raw --csv 2>/dev/null | python profiles/ncu_full_summary.py "k_solve_fwd256 launches 82..139 (root, first to last superstep) and k_solve_bwd256 launches 0..61 of the first solve, config 2" > gpurun_out/c_ncu_solve.json 2>> gpurun_out/c_ncu.log
tail -3 gpurun_out/c_pytest_1gpu.log; tail -c 600 gpurun_out/c_bench_cfg2.json; cat gpurun_out/c_solve.json gpurun_out/c_solve_cfg5.json; ls -la gpurun_out/c_ncu_solve.json
