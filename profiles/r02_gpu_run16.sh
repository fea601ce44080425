set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
O=gpurun_out
timeout 900 ncu --set full --import-source on --clock-control none -f -k regex:evolve_ -c 1 -o $O/c1 python profiles/prof_kernels.py c1 > $O/ncu_c1.log 2>&1
ncu -i $O/c1.ncu-rep --page source --csv --print-source sass,cuda > /tmp/c1_src.csv 2>/dev/null
python profiles/ncu_lines.py /tmp/c1_src.csv 90 > $O/r02_ncu_c1_lines.txt 2>&1
ncu -i $O/c1.ncu-rep --page source --csv > /tmp/c1_sass.csv 2>/dev/null
python profiles/ncu_stalls.py /tmp/c1_sass.csv 40 > $O/r02_ncu_c1_sass_stalls.txt 2>&1
rm -f $O/c1.ncu-rep
head -50 $O/r02_ncu_c1_lines.txt
