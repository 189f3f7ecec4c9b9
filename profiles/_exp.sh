mkdir -p gpurun_out/r2o
python -m pytest tests -m gpu -q -x > gpurun_out/r2o/pytest.log 2>&1; echo "pytest exit $?"; tail -8 gpurun_out/r2o/pytest.log
python bench.py --steps 20 --warmup 5 --workload 4lane_lidar_ogm_dagm_256 --no-cpu-baseline > gpurun_out/r2o/bench.log 2> gpurun_out/r2o/bench.err; echo "bench exit $?"; tail -c 1000 gpurun_out/r2o/bench.err
python - <<PY
import json
l=[x for x in open("gpurun_out/r2o/bench.log") if x.startswith("{")]
if l:
    d=json.loads(l[-1]); print(d["ms_per_step"], d["roofline"]["frac"], d["e2e"])
PY
