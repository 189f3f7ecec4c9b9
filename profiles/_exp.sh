mkdir -p gpurun_out/r2d
python -m pytest tests -m gpu -x -q > gpurun_out/r2d/pytest.log 2>&1; echo "pytest exit $?"; tail -15 gpurun_out/r2d/pytest.log
python bench.py --no-cpu-baseline --e2e-steps 1 | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('timed', d['ms_per_step'], d['config']['kernel_ms_per_step'])"
