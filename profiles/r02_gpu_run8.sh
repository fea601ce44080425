set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "tiled or sharded_apply or large_apply or large_streamed" > gpurun_out/r02_pytest_tma.log 2>&1
tail -5 gpurun_out/r02_pytest_tma.log
for n in 25 24 26; do timeout 300 python profiles/prof_apply.py $n > gpurun_out/r02_hpsi_tma_n$n.txt 2>&1; tail -1 gpurun_out/r02_hpsi_tma_n$n.txt; done
