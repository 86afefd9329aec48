set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_complex.py -q -m gpu -k "sweeps or multi_rhs or hermitian or complex or single_precision or full_size_properties" > gpurun_out/f_pytest.log 2>&1; echo "rc=$?" >> gpurun_out/f_pytest.log
timeout 300 python profiles/solve_trace.py cfg2 > gpurun_out/f_trace.json 2> gpurun_out/f_trace.err
timeout 300 python profiles/solve_table.py cfg2 > gpurun_out/f_solve_pdl.json 2> gpurun_out/f_solve.err
CHOLMOD_B200_SOLVE_PDL=0 timeout 300 python profiles/solve_table.py cfg2 > gpurun_out/f_solve_nopdl.json 2>> gpurun_out/f_solve.err
timeout 300 python profiles/solve_table.py cfg2 4 > gpurun_out/f_solve_4rhs.json 2>> gpurun_out/f_solve.err
tail -3 gpurun_out/f_pytest.log; cat gpurun_out/f_trace.json gpurun_out/f_solve_pdl.json gpurun_out/f_solve_nopdl.json gpurun_out/f_solve_4rhs.json
