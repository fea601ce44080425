set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
O=gpurun_out
timeout 300 python profiles/run_config4.py 0.02 200 ip 1e-7 1e-14 compare > $O/r02_c4_compare_0p02.json 2>&1; tail -2 $O/r02_c4_compare_0p02.json
timeout 900 python profiles/run_config4.py 1.0 1000 ip > $O/r02_c4_full_ip.json 2>&1; tail -1 $O/r02_c4_full_ip.json
