set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_complex.py -q -m gpu -k "graph or sweeps or multi_rhs or hermitian or complex or single_precision" > gpurun_out/d_pytest.log 2>&1; echo "rc=$?" >> gpurun_out/d_pytest.log
timeout 300 python profiles/solve_table.py cfg2 > gpurun_out/d_solve.json 2> gpurun_out/d_solve.err
tail -3 gpurun_out/d_pytest.log; cat gpurun_out/d_solve.json
