set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 470 python -m pytest tests -m gpu -x -q > gpurun_out/r02_pytest_gpu_final4.log 2>&1; tail -3 gpurun_out/r02_pytest_gpu_final4.log
