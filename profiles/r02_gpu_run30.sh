set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
BQA_SMALL_THREADS=1024 timeout 400 python profiles/c3_pictures.py 512 > gpurun_out/r02_c3_pictures_1024.json 2>&1; tail -1 gpurun_out/r02_c3_pictures_1024.json
