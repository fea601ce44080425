set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
O=gpurun_out
timeout 2400 python -m pytest tests -m gpu -x -q > $O/r02_pytest_gpu_final3.log 2>&1
tail -3 $O/r02_pytest_gpu_final3.log
timeout 900 python bench.py > $O/r02_bench_n1_d.json 2> $O/r02_bench_n1_d.err; tail -c 700 $O/r02_bench_n1_d.json; tail -2 $O/r02_bench_n1_d.err
python -c "import __graft_entry__ as g; g.smoke()" > $O/r02_smoke.txt 2>&1; tail -1 $O/r02_smoke.txt
