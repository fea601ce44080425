"""Integrator outcome of EVERY task of BASELINE configs[2] (the 4096-task C3 sweep) from SciPy itself: the
reference's integrator (scipy.integrate.ode dop853, atol=1e-7, rtol=1e-14: src/bloqade/analog/emulate/ir/
state_vector.py:589-600) driven by the oracle's per-term restatement of the Hamiltonian (oracle/emulator.py).

Writes tests/golden/c3_status.json: per task {status, calls} where `calls` is the number of RHS evaluations SciPy
made.  SciPy's C translation re-evaluates the RHS once after every accepted step (at x+h) and once after every
rejected one (at x), which the device controller does not, so device nfcn + naccpt + nrejct == calls - 1 when
the step sequences are identical.  tests/test_gpu_parity.py::test_config3_statuses_match_scipy checks all 4096.

   python tests/golden/make_c3_status.py [--workers 8]        (~15 min on 8 cores)
"""
import argparse
import json
import multiprocessing as mp
import os
import sys
import time

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)


def one(k):
    from scipy.integrate import ode

    from bloqade_analog_b200 import programs
    from oracle import emulator as oe

    d, t = programs.config3_grid(64, 64)[k]
    h = oe.Hamiltonian(programs.config3_task(d, t))
    psi0 = np.zeros(h.dim, dtype=np.complex128)
    psi0[0] = 1.0
    calls = [0]

    def f(time_, y):
        calls[0] += 1
        return h.rhs(time_, y.view(np.complex128)).view(np.float64)

    s = ode(f)
    s.set_initial_value(psi0.view(np.float64).copy())
    s.set_integrator("dop853", atol=1e-7, rtol=1e-14, nsteps=2**31 - 1)
    y = s.integrate(h.duration)
    p = np.abs(y.view(np.complex128)) ** 2
    return k, int(s.get_return_code()), calls[0], float(p.sum()), float(p[0])


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--workers", type=int, default=os.cpu_count())
    ap.add_argument("--tasks", type=int, default=4096)
    args = ap.parse_args()
    import scipy

    t0 = time.time()
    # longest tasks first so the pool drains evenly
    order = sorted(range(args.tasks), key=lambda k: -(k % 64))
    out = {}
    with mp.get_context("fork").Pool(args.workers) as pool:
        for k, status, calls, norm2, p0 in pool.imap_unordered(one, order, chunksize=4):
            out[k] = {"status": status, "calls": calls, "norm2": norm2, "p0": p0}
            if len(out) % 256 == 0:
                print(len(out), "tasks", round(time.time() - t0), "s", flush=True)
    doc = {"scipy": scipy.__version__, "solver": "dop853", "atol": 1e-7, "rtol": 1e-14,
           "grid": "programs.config3_grid(64, 64), task = i_det * 64 + i_dur",
           "status": [out[k]["status"] for k in range(args.tasks)],
           "calls": [out[k]["calls"] for k in range(args.tasks)],
           "norm2": [out[k]["norm2"] for k in range(args.tasks)],
           "p0": [out[k]["p0"] for k in range(args.tasks)]}
    with open(os.path.join(HERE, "c3_status.json"), "w") as fp:
        json.dump(doc, fp)
    print("done", round(time.time() - t0), "s; non-OK:", [k for k in range(args.tasks) if out[k]["status"] != 1])
