set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 300 python profiles/prof_kernels.py three_level 8 148 > gpurun_out/r02_three_level_8.txt 2>&1; tail -1 gpurun_out/r02_three_level_8.txt
timeout 300 python profiles/prof_kernels.py three_level 6 296 > gpurun_out/r02_three_level_6.txt 2>&1; tail -1 gpurun_out/r02_three_level_6.txt
