set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
O=gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "interaction or streamed" > $O/r02_pytest_ip.log 2>&1; tail -3 $O/r02_pytest_ip.log
timeout 300 python profiles/run_config4.py 0.02 200 ip 1e-7 1e-14 compare > $O/r02_c4_compare_0p02_b.json 2>&1; tail -2 $O/r02_c4_compare_0p02_b.json
timeout 900 python profiles/run_config4.py 1.0 1000 ip > $O/r02_c4_full_ip_b.json 2>&1; tail -1 $O/r02_c4_full_ip_b.json
