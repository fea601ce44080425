"""Small driver for ncu: a few standalone fused H(t) psi applies on a 2^n full-space state.
   python profiles/prof_apply.py [n_atoms]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import bench
n = int(sys.argv[1]) if len(sys.argv) > 1 else 25
peaks, kind = bench.measured_peaks()
print(bench.hpsi_roofline(torch, 0, peaks, kind, n, reps=5))
