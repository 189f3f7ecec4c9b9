mkdir -p gpurun_out/r2j
python -m pytest tests/test_gpu_parity.py -m gpu -q -x > gpurun_out/r2j/pytest.log 2>&1; echo "pytest exit $?"; tail -5 gpurun_out/r2j/pytest.log
for v in base r96 r64b5; do
  if [ $v = base ]; then unset SMARTS_B200_LIB; else export SMARTS_B200_LIB=$PWD/profiles/_bin/lib_$v.so; fi
  python bench.py --steps 20 --warmup 5 --e2e-steps 1 --no-cpu-baseline > gpurun_out/r2j/bench_$v.log 2>&1; echo "bench $v exit $?"
  python - <<PY
import json
l=[x for x in open("gpurun_out/r2j/bench_$v.log") if x.startswith("{")]
if l:
    d=json.loads(l[-1]); print(d["ms_per_step"], d["config"]["kernel_ms_per_step"], d["roofline"]["frac"])
else:
    print(open("gpurun_out/r2j/bench_$v.log").read()[-600:])
PY
done
