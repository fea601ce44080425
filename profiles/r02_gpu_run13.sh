set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
O=gpurun_out
timeout 2400 python -m pytest tests -m gpu -x -q > $O/r02_pytest_gpu_final.log 2>&1
tail -5 $O/r02_pytest_gpu_final.log
timeout 200 python profiles/prof_kernels.py reductions 25 > $O/r02_prof_reductions2.txt 2>&1; tail -1 $O/r02_prof_reductions2.txt
timeout 900 python bench.py > $O/r02_bench_n1_b.json 2> $O/r02_bench_n1_b.err; tail -c 3000 $O/r02_bench_n1_b.json
timeout 600 python bench.py --impl reference --steps 1 --warmup 1 > $O/r02_bench_ref_b.json 2> $O/r02_bench_ref_b.err; tail -c 800 $O/r02_bench_ref_b.json
python -c "import __graft_entry__ as g; g.smoke()" > $O/r02_smoke.txt 2>&1; tail -2 $O/r02_smoke.txt
