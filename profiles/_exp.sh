mkdir -p gpurun_out/r2q
python -m pytest tests -m gpu -q > gpurun_out/r2q/pytest.log 2>&1; echo "pytest exit $?"; tail -30 gpurun_out/r2q/pytest.log
