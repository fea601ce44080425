set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
O=gpurun_out
run() { tag=$1; shift; env "$@" timeout 200 python profiles/prof_kernels.py c1 > $O/r02_c1_$tag.txt 2>&1; tail -1 $O/r02_c1_$tag.txt; }
run cs1_k2 X=1
run cs1_k1 BQA_FAST_K1=1
run cs2_k2 BQA_CLUSTER_SIZE=2
run cs2_k1 BQA_CLUSTER_SIZE=2 BQA_FAST_K1=1
run cs4_k1 BQA_CLUSTER_SIZE=4
timeout 600 ncu --set full --clock-control none -f -k "regex:scan_|sample_search" --launch-skip 9 -c 3 -o $O/r02_ncu_sampling python profiles/prof_kernels.py reductions 25 > $O/ncu_s.log 2>&1
ncu -i $O/r02_ncu_sampling.ncu-rep --page raw --csv > $O/r02_ncu_sampling_raw.csv 2>/dev/null; rm -f $O/r02_ncu_sampling.ncu-rep
