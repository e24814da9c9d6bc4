import os, sys, json
sys.path.insert(0, os.getcwd())
import torch
import bench
from gato_python_b200 import Plan
r = bench.single_solve_latency(torch, Plan, 0, reps=1000)
print(os.environ.get("GATO_B200_LIB", "production")[-20:], r["kernel_us_median"], r["kernel_us_min"], r["bracket_us_median"], r["iters"])
