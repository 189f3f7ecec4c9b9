mkdir -p gpurun_out/r2f
python -m pytest tests -m gpu -q > gpurun_out/r2f/pytest.log 2>&1; echo "pytest exit $?"; tail -25 gpurun_out/r2f/pytest.log
python bench.py --steps 20 --warmup 5 > gpurun_out/r2f/bench.log 2>&1; echo "bench exit $?"; tail -3 gpurun_out/r2f/bench.log
