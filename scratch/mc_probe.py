import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import optiland_pr_b200 as ol
from optiland_pr_b200 import samples
ol.backend.set_device("cuda:0"); ol.backend.set_precision("float64")
import bench
from optiland_pr_b200.tolerancing import MonteCarloSpot
lens = samples.EvenAsphereDoubleGauss(); lens.surface_group.record = "none"
for ch in (1, 2, 4, 8):
    MonteCarloSpot.draw_chunks = ch
    r = bench.bench_monte_carlo(torch, lens, 12)["mc_tolerancing_c5"]
    print(ch, r["seconds"], r.get("seconds_runs"), r["seconds_host_draws"], r["rms_spot_mm_mean_std"])
