set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 300 python profiles/check_fused_step.py 148 0.5 > gpurun_out/r02_check_tmem.txt 2>&1
BQA_NO_TMEM=1 timeout 300 python profiles/check_fused_step.py 148 0.5 >> gpurun_out/r02_check_tmem.txt 2>&1
timeout 300 python profiles/check_fused_step.py 40 2.0 >> gpurun_out/r02_check_tmem.txt 2>&1
timeout 300 compute-sanitizer --tool memcheck python profiles/check_fused_step.py 8 0.05 > gpurun_out/r02_sanitizer_tmem.txt 2>&1
