#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests/test_gpu_grads.py tests/test_gpu_grads_bx.py tests/test_gpu_golden.py -x -q > gpurun_out/pytest_bx.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_bx.log
python scratch/prof_bwd.py 16777216 > gpurun_out/prof_bwd16.log 2>&1; cat gpurun_out/prof_bwd16.log
python scratch/prof_fx.py > gpurun_out/prof_fx.log 2>&1; cat gpurun_out/prof_fx.log
timeout 600 ncu --set full --clock-control none --import-source on -k regex:trace_bx_kernel -s 1 -c 1 -f -o gpurun_out/prof_bx \
  python scratch/prof_bwd.py > gpurun_out/ncu_bx.log 2>&1; echo "ncu bx rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:trace_fx_kernel -s 2 -c 1 -f -o gpurun_out/prof_fxrec \
  python scratch/prof_bwd.py > gpurun_out/ncu_fxrec.log 2>&1; echo "ncu fxrec rc=$?"
