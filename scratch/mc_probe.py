import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import optiland_pr_b200 as ol
from optiland_pr_b200 import samples
ol.backend.set_device("cuda:0"); ol.backend.set_precision("float64")
import bench
lens = samples.EvenAsphereDoubleGauss(); lens.surface_group.record = "none"
for _ in range(4):
    r = bench.bench_monte_carlo(torch, lens, 12)["mc_tolerancing_c5"]
    print(r["seconds"], r["seconds_host_draws"], r["path"])
import cProfile, pstats
from optiland_pr_b200.tolerancing import DistributionSampler, MonteCarloSpot, Perturbation
