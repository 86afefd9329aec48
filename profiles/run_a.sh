set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "sweeps or multi_rhs or hermitian or single or full_size_properties or golden" > gpurun_out/a_pytest_solve.log 2>&1; echo "rc=$?" >> gpurun_out/a_pytest_solve.log
for i in 1 2 3 4 5 6; do timeout 300 python -m pytest tests/test_gpu_dropin2.py -x -q -m gpu -k "memory_settings" 2>&1 | tail -15 >> gpurun_out/a_threads.log; done
timeout 400 python -m pytest tests/test_gpu_complex.py tests/test_gpu_dropin2.py tests/test_gpu_dropin.py -x -q -m gpu > gpurun_out/a_dropin.log 2>&1; echo "rc=$?" >> gpurun_out/a_dropin.log
timeout 300 python profiles/solve_table.py cfg2 > gpurun_out/a_solve256.json 2> gpurun_out/a_solve256.err
CHOLMOD_B200_SOLVE_COLS=64 timeout 300 python profiles/solve_table.py cfg2 > gpurun_out/a_solve64.json 2> gpurun_out/a_solve64.err
timeout 300 python profiles/solve_table.py cfg2 4 > gpurun_out/a_solve256_4rhs.json 2>> gpurun_out/a_solve256.err
tail -3 gpurun_out/a_pytest_solve.log; tail -2 gpurun_out/a_dropin.log; grep -c passed gpurun_out/a_threads.log; cat gpurun_out/a_solve256.json gpurun_out/a_solve64.json gpurun_out/a_solve256_4rhs.json
