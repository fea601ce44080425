set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
O=gpurun_out
timeout 200 python profiles/prof_kernels.py c2 > $O/r02_c2_subrhs.txt 2>&1; tail -1 $O/r02_c2_subrhs.txt
BQA_SMALL_THREADS=1024 timeout 200 python profiles/prof_kernels.py c2 > $O/r02_c2_subrhs_1024.txt 2>&1; tail -1 $O/r02_c2_subrhs_1024.txt
BQA_SMALL_THREADS=256 timeout 200 python profiles/prof_kernels.py c2 > $O/r02_c2_subrhs_256.txt 2>&1; tail -1 $O/r02_c2_subrhs_256.txt
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_dropin_gpu.py -m gpu -x -q -k "config2 or config_cases or blockade or all_engines or evolve_small_cases or interaction" > $O/r02_pytest_sub.log 2>&1
tail -3 $O/r02_pytest_sub.log
