#!/bin/bash
# one GPU visit: parity tests, bench, ncu launch list, ncu --set full of the dominant kernel(s)
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/smoke.log
python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"; tail -c 600 gpurun_out/bench.err
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref rc=$?"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv \
  python bench.py --steps 5 --warmup 3 --no-cpu-baseline --extras 0 > gpurun_out/ncu_list.log 2>&1; echo "ncu list rc=$?"
python scratch/prof_fx.py > gpurun_out/prof_fx.log 2>&1 && \
timeout 600 ncu --set full --clock-control none --import-source on -k regex:trace_fx_kernel -s 3 -c 1 -f -o gpurun_out/prof_fx \
  python scratch/prof_fx.py > gpurun_out/ncu_fx.log 2>&1; echo "ncu fx rc=$?"; cat gpurun_out/prof_fx.log
python scratch/prof_bwd.py > gpurun_out/prof_bwd.log 2>&1 && \
timeout 600 ncu --set full --clock-control none --import-source on -k regex:trace_bx_kernel -s 1 -c 1 -f -o gpurun_out/prof_bx \
  python scratch/prof_bwd.py > gpurun_out/ncu_bwd.log 2>&1; echo "ncu bx rc=$?"; cat gpurun_out/prof_bwd.log
