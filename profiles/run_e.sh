set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_complex.py -q -m gpu -k "sweeps or multi_rhs or hermitian or complex or single_precision or full_size_properties" > gpurun_out/e_pytest.log 2>&1; echo "rc=$?" >> gpurun_out/e_pytest.log
timeout 300 python profiles/solve_trace.py cfg2 > gpurun_out/e_trace2.json 2> gpurun_out/e_trace2.err
timeout 300 python profiles/solve_table.py cfg2 > gpurun_out/e_solve.json 2> gpurun_out/e_solve.err
CHOLMOD_B200_SOLVE_W256=0 timeout 300 python profiles/solve_table.py cfg2 > gpurun_out/e_solve_now256.json 2>> gpurun_out/e_solve.err
tail -3 gpurun_out/e_pytest.log; cat gpurun_out/e_trace2.json gpurun_out/e_solve.json gpurun_out/e_solve_now256.json
