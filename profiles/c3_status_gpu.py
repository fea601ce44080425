"""All 4096 C3 tasks on the GPU against tests/golden/c3_status.json (SciPy): statuses, RHS-count agreement.
   python profiles/c3_status_gpu.py > gpurun_out/r02_c3_status_gpu.json"""
import json, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
from bloqade_analog_b200 import programs
from bloqade_analog_b200.batch import run_programs

gold = json.load(open(os.path.join(ROOT, "tests", "golden", "c3_status.json")))
progs = [programs.config3_task(d, t) for d, t in programs.config3_grid(64, 64)]
np.random.seed(3)
t0 = time.time()
res = run_programs(progs, 4, raise_on_error=False)
dt = time.time() - t0
status = np.array([r.status for r in res]); ref = np.array(gold["status"])
nf = np.array([r.nfcn + r.naccpt + r.nrejct for r in res]); calls = np.array(gold["calls"])
ok = (status == 1) & (ref == 1)
same = nf[ok] == calls[ok] - 1
rel = np.abs(nf[ok] - (calls[ok] - 1)) / calls[ok]
print(json.dumps({"seconds": dt, "mismatching_tasks": [int(k) for k in np.nonzero(status != ref)[0]],
                  "gpu_failed": [int(k) for k in np.nonzero(status != 1)[0]], "scipy_failed": [int(k) for k in np.nonzero(ref != 1)[0]],
                  "rhs_count_identical_fraction": float(same.mean()), "rhs_total_ratio_gpu_over_scipy": float(nf[ok].sum() / (calls[ok] - 1).sum()),
                  "rhs_count_rel_diff_median": float(np.median(rel)), "rhs_count_max_rel_diff": float(rel.max()),
                  "rhs_count_rel_diff_p99": float(np.quantile(rel, 0.99))}))
