mkdir -p gpurun_out/r2n
python -m pytest tests -m gpu -q -x > gpurun_out/r2n/pytest.log 2>&1; echo "pytest exit $?"; tail -5 gpurun_out/r2n/pytest.log
python bench.py --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/r2n/bench.log 2> gpurun_out/r2n/bench.err; echo "bench exit $?"; tail -c 1500 gpurun_out/r2n/bench.err
python - <<PY
import json
l=[x for x in open("gpurun_out/r2n/bench.log") if x.startswith("{")]
if l:
    d=json.loads(l[-1]); print(d["ms_per_step"], d["detail"], d["roofline"]["frac"], d["e2e"]["value"], d["e2e"]["value_incremental"])
    for k,v in d.get("other_workloads",{}).items(): print(k, v["ms_per_step"], v["value"], v["step_frac_of_peak"], v["dominant_kernel"], v["kernel_ms_per_step"])
PY
