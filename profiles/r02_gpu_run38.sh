set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_dropin_gpu.py -m gpu -x -q -k "not oracle and not full_protocol and not large and not tiled and not sharded and not config3_statuses" > gpurun_out/r02_pytest_engine_policy.log 2>&1; tail -3 gpurun_out/r02_pytest_engine_policy.log
