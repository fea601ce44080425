"""Diagnostic: status of selected C3 sweep tasks on every engine (stiffness detection is knife-edge on these)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import bench
from bloqade_analog_b200.batch import evolve_programs
progs = bench.sweep_programs(4096, 0)
ids = [124, 188, 437, 1143, 60, 125]
for eng in ("cta", "cluster", "streamed"):
    for env in ("", "1"):
        if env and eng != "cluster": continue
        if env: os.environ["BQA_NO_FAST"] = "1"
        states, st = evolve_programs([progs[i] for i in ids], engine=eng)
        os.environ.pop("BQA_NO_FAST", None)
        print(eng, "nofast" if env else "", [(i, int(s["status"]), int(s["naccpt"]), int(s["nrejct"])) for i, s in zip(ids, st)])
