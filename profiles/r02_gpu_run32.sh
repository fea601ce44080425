set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
O=gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_dropin_gpu.py -m gpu -x -q -k "evolve_small_cases or interaction or exact_diagonalisation or python_fn or apply_matches" > $O/r02_pytest_phase.log 2>&1; tail -3 $O/r02_pytest_phase.log
