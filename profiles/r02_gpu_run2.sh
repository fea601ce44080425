set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 300 python profiles/check_fused_step.py 148 0.5 > gpurun_out/r02_check_fused.txt 2>&1
timeout 300 python profiles/check_fused_step.py 40 2.0 >> gpurun_out/r02_check_fused.txt 2>&1
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "output_times_after or barrier_is_bounded or config3_statuses or evolve_small_cases or all_engines" > gpurun_out/r02_pytest_b.log 2>&1
timeout 300 python profiles/c3_knife_gpu.py 440 951 1783 > gpurun_out/r02_c3_knife_gpu.txt 2>&1
timeout 300 python profiles/c3_status_gpu.py > gpurun_out/r02_c3_status_gpu.json 2> gpurun_out/r02_c3_status_gpu.err
timeout 600 python bench.py --steps 2 --warmup 3 --no-small --cpu-tasks 2 > gpurun_out/r02_bench_n1_b.json 2> gpurun_out/r02_bench_n1_b.err
