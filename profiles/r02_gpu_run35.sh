set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "sharded" > gpurun_out/r02_pytest_shard3.log 2>&1; tail -3 gpurun_out/r02_pytest_shard3.log
