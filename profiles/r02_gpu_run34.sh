set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
O=gpurun_out
rm -f $O/*.ncu-rep
timeout 600 ncu --set full --clock-control none -f -k regex:evolve_ -c 1 -o $O/c2 python profiles/prof_kernels.py c2 > $O/ncu_c2b.log 2>&1
ncu -i $O/c2.ncu-rep --page raw --csv > $O/r02_ncu_c2b_raw.csv 2>/dev/null; rm -f $O/c2.ncu-rep
cat > /tmp/ip_stream.py <<'PY'
import sys, os, json, time
sys.path.insert(0, os.getcwd())
import numpy as np
from bloqade_analog_b200 import _lib, ir, programs
from bloqade_analog_b200.hamiltonian import RydbergHamiltonianCodeGen
prog = programs.adiabatic_program(ir.square_sites(5, 5, lattice_spacing=5.0))
h = RydbergHamiltonianCodeGen().emit(prog)
opts = _lib.make_opts("dop853", 1e-7, 1e-14, interaction_picture=True, engine="streamed")
out, st = h.plan.evolve_batch(h.record, np.array([0, h.record.size], dtype=np.int64), np.array([[0.02]]), opts)
print(int(st.reshape(-1)[0]["nfcn"]))
PY
timeout 600 ncu --section SpeedOfLight --section MemoryWorkloadAnalysis --section LaunchStats --clock-control none -f -k regex:lg_ --launch-skip 40 -c 60 -o $O/ip python /tmp/ip_stream.py > $O/ncu_ip.log 2>&1
ncu -i $O/ip.ncu-rep --page raw --csv > $O/r02_ncu_streamed_ip_raw.csv 2>/dev/null; rm -f $O/ip.ncu-rep
ls -la $O | tail -5
