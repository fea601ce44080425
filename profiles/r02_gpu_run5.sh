set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r02_pytest_gpu_c.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r02_pytest_gpu_c.log
timeout 300 python profiles/c3_status_gpu.py > gpurun_out/r02_c3_status_gpu.json 2> gpurun_out/r02_c3_status_gpu.err
timeout 600 python bench.py --steps 3 --warmup 3 > gpurun_out/r02_bench_n1_c.json 2> gpurun_out/r02_bench_n1_c.err
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r02_launches_bench_512tasks.csv python bench.py --steps 2 --warmup 1 --tasks 512 --no-hpsi --no-small --cpu-tasks 2 > gpurun_out/ncu_l.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:evolve_small -c 1 -o gpurun_out/r02_evolve_small_148_tmem -f python profiles/prof_evolve.py cta 148 0.5 > gpurun_out/ncu_f3.log 2>&1
