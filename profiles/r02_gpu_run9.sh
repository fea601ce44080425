set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
O=gpurun_out
timeout 400 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "tiled_apply or sharded_apply" > $O/r02_pytest_tma2.log 2>&1
tail -3 $O/r02_pytest_tma2.log
for n in 25 26; do timeout 300 python profiles/prof_apply.py $n > $O/r02_hpsi_tma2_n$n.txt 2>&1; tail -1 $O/r02_hpsi_tma2_n$n.txt; done
timeout 300 python profiles/prof_kernels.py reductions 25 > $O/r02_prof_reductions.txt 2>&1; tail -1 $O/r02_prof_reductions.txt
timeout 300 python profiles/prof_kernels.py streamed 25 > $O/r02_prof_streamed.txt 2>&1; tail -1 $O/r02_prof_streamed.txt
timeout 300 python profiles/prof_kernels.py c1 > $O/r02_prof_c1.txt 2>&1; tail -1 $O/r02_prof_c1.txt
timeout 300 python profiles/prof_kernels.py c2 > $O/r02_prof_c2.txt 2>&1; tail -1 $O/r02_prof_c2.txt
NCU="ncu --set full --clock-control none --import-source on -f"
timeout 600 $NCU -k regex:apply_tiled --launch-skip 6 -c 2 -o $O/r02_apply_tma python profiles/prof_apply.py 25 > $O/ncu_a.log 2>&1
timeout 600 $NCU -k "regex:norm2|finish_sum|scan_|sample_search" --launch-skip 8 -c 6 -o $O/r02_reductions python profiles/prof_kernels.py reductions 25 > $O/ncu_b.log 2>&1
timeout 900 ncu --section SpeedOfLight --section MemoryWorkloadAnalysis --section WarpStateStats --section LaunchStats --clock-control none -f -k regex:lg_ --launch-skip 60 -c 48 -o $O/r02_streamed python profiles/prof_kernels.py streamed 25 > $O/ncu_c.log 2>&1
timeout 900 $NCU -k regex:evolve_ -c 1 -o $O/r02_c1 python profiles/prof_kernels.py c1 > $O/ncu_d.log 2>&1
timeout 900 $NCU -k regex:evolve_ -c 1 -o $O/r02_c2 python profiles/prof_kernels.py c2 > $O/ncu_e.log 2>&1
ls -la $O
