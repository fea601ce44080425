set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
O=gpurun_out
run() { tag=$1; shift; env "$@" timeout 200 python profiles/prof_kernels.py c1 > $O/r02_c1b_$tag.txt 2>&1; tail -1 $O/r02_c1b_$tag.txt; }
run cs1_k2 X=1
run cs2_k2 BQA_CLUSTER_SIZE=2
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "cluster or evolve_small_cases or all_engines" > $O/r02_pytest_cluster.log 2>&1
tail -3 $O/r02_pytest_cluster.log
timeout 200 python profiles/prof_evolve.py cluster 296 0.5 9 > $O/r02_prof_cluster_9atoms.txt 2>&1; tail -1 $O/r02_prof_cluster_9atoms.txt
timeout 200 python profiles/prof_evolve.py cluster 296 0.5 11 > $O/r02_prof_cluster_11atoms.txt 2>&1; tail -1 $O/r02_prof_cluster_11atoms.txt
