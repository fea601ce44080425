set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
O=gpurun_out
timeout 400 python profiles/c3_pictures.py 512 > $O/r02_c3_pictures_c.json 2>&1; tail -1 $O/r02_c3_pictures_c.json
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_dropin_gpu.py -m gpu -x -q -k "interaction or config_cases or dopri5 or run_batch or resume or output_times" > $O/r02_pytest_ip3.log 2>&1; tail -3 $O/r02_pytest_ip3.log
