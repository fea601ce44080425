set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 900 ncu --set full --clock-control none --import-source on -k regex:evolve_small -c 1 -o gpurun_out/r02_evolve_small_148_fused -f python profiles/prof_evolve.py cta 148 0.5 > gpurun_out/ncu_f2.log 2>&1
BQA_SMALL_THREADS=1024 timeout 300 python profiles/prof_evolve.py cta 148 0.5 > gpurun_out/r02_prof_1024.txt 2>&1
timeout 300 python profiles/prof_evolve.py cta 148 0.5 > gpurun_out/r02_prof_512.txt 2>&1
