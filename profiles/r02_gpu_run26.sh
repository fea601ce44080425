set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 400 python profiles/c3_pictures.py 512 > gpurun_out/r02_c3_pictures.json 2>&1; tail -1 gpurun_out/r02_c3_pictures.json
