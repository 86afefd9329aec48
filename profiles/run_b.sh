set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "graph or sweeps or multi_rhs or not_posdef" > gpurun_out/b_pytest.log 2>&1; echo "rc=$?" >> gpurun_out/b_pytest.log
timeout 400 python -m pytest tests/test_gpu_complex.py -x -q -m gpu > gpurun_out/b_complex.log 2>&1; echo "rc=$?" >> gpurun_out/b_complex.log
timeout 600 python bench.py --workload cfg5 --no-cpu-baseline > gpurun_out/b_cfg5_graph.json 2> gpurun_out/b_cfg5_graph.err
CHOLMOD_B200_GRAPH=0 timeout 600 python bench.py --workload cfg5 --no-cpu-baseline --no-e2e > gpurun_out/b_cfg5_plain.json 2> gpurun_out/b_cfg5_plain.err
CHOLMOD_B200_GRAPH=1 timeout 600 python bench.py --workload cfg2 --no-cpu-baseline > gpurun_out/b_cfg2_graph.json 2> gpurun_out/b_cfg2_graph.err
timeout 300 python profiles/solve_table.py cfg2 > gpurun_out/b_solve256.json 2> gpurun_out/b_solve256.err
tail -3 gpurun_out/b_pytest.log; tail -2 gpurun_out/b_complex.log
for f in b_cfg5_graph b_cfg5_plain b_cfg2_graph; do python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/$f.json").read().strip().splitlines()[-1])
    print("$f", d["ms_per_step"], d.get("e2e",{}).get("ms_per_step"), d.get("first_call_ms"), d["gpu_launches"], d.get("check"))
except Exception as e:
    print("$f", "ERR", e)
PY
done
cat gpurun_out/b_solve256.json
