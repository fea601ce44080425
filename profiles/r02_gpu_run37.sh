set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 400 python profiles/prof_small_ip.py > gpurun_out/r02_small_ip_engines.txt 2>&1; tail -8 gpurun_out/r02_small_ip_engines.txt
