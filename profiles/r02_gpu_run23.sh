set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
O=gpurun_out
timeout 300 python profiles/run_config4.py 0.02 200 s > $O/r02_c4_s_0p02.json 2>&1; tail -1 $O/r02_c4_s_0p02.json
timeout 300 python profiles/run_config4.py 0.02 200 ip > $O/r02_c4_ip_0p02.json 2>&1; tail -1 $O/r02_c4_ip_0p02.json
timeout 300 python profiles/run_config4.py 0.02 200 s 1e-10 1e-7 > $O/r02_c4_s_tol2_0p02.json 2>&1; tail -1 $O/r02_c4_s_tol2_0p02.json
