set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
O=gpurun_out
rm -f $O/*.ncu-rep
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "sampling or full_protocol or run_batch or gate_run" > $O/r02_pytest_sampling.log 2>&1
tail -3 $O/r02_pytest_sampling.log
timeout 300 python profiles/prof_kernels.py reductions 25 > $O/r02_prof_reductions.txt 2>&1; tail -1 $O/r02_prof_reductions.txt
NCU="ncu --set full --clock-control none -f"
exp() { # raw csv + details text of a report, then drop the report (gpurun_out is capped at 64 MiB)
  ncu -i $O/$1.ncu-rep --page raw --csv > $O/$1_raw.csv 2>/dev/null
  ncu -i $O/$1.ncu-rep --page details > $O/$1_details.txt 2>/dev/null
  rm -f $O/$1.ncu-rep
}
timeout 600 $NCU -k "regex:norm2|finish_sum|scan_|sample_search" --launch-skip 8 -c 5 -o $O/r02_ncu_reductions python profiles/prof_kernels.py reductions 25 > $O/ncu_b.log 2>&1; exp r02_ncu_reductions
timeout 900 $NCU -k regex:lg_ --launch-skip 60 -c 48 -o $O/r02_ncu_streamed python profiles/prof_kernels.py streamed 25 > $O/ncu_c.log 2>&1; exp r02_ncu_streamed
timeout 900 $NCU --import-source on -k regex:evolve_ -c 1 -o $O/r02_ncu_c1 python profiles/prof_kernels.py c1 > $O/ncu_d.log 2>&1
ncu -i $O/r02_ncu_c1.ncu-rep --page source --csv > $O/r02_ncu_c1_source.csv 2>/dev/null; exp r02_ncu_c1
timeout 900 $NCU --import-source on -k regex:evolve_ -c 1 -o $O/r02_ncu_c2 python profiles/prof_kernels.py c2 > $O/ncu_e.log 2>&1
ncu -i $O/r02_ncu_c2.ncu-rep --page source --csv > $O/r02_ncu_c2_source.csv 2>/dev/null; exp r02_ncu_c2
ls -la $O; du -sh $O
