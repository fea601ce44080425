set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_dropin_gpu.py -m gpu -x -q -k "resume_from or python_fn or output_times or barrier_is_bounded or evolve_small_cases or run_callback" > gpurun_out/r02_pytest_d.log 2>&1
tail -5 gpurun_out/r02_pytest_d.log
