set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
O=gpurun_out
timeout 200 python profiles/prof_kernels.py c1 > $O/r02_c1c.txt 2>&1; tail -1 $O/r02_c1c.txt
timeout 200 python profiles/prof_evolve.py cluster 296 0.5 9 > $O/r02_prof_cluster_9atoms_c.txt 2>&1; tail -1 $O/r02_prof_cluster_9atoms_c.txt
timeout 200 python profiles/prof_evolve.py cluster 296 0.5 11 > $O/r02_prof_cluster_11atoms_c.txt 2>&1; tail -1 $O/r02_prof_cluster_11atoms_c.txt
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "cluster or evolve_small_cases or all_engines" > $O/r02_pytest_cluster2.log 2>&1; tail -2 $O/r02_pytest_cluster2.log
