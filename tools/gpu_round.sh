#!/bin/bash
# One GPU-box pass: parity tests, bench line, ncu launch lists and full captures (run under gpurun).
set -u
mkdir -p gpurun_out
K='regex:dce_|msrcr_|down4_|decim4_|fir_|up_bc_'
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
python bench.py --steps 5 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"
python tools/prof_run.py dce 16 > gpurun_out/plain_dce.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -k "$K" -c 40 --csv --log-file gpurun_out/launches_dce.csv python tools/prof_run.py dce 16 > gpurun_out/ncu_dce_l.log 2>&1
python tools/prof_run.py msrcr 8 > gpurun_out/plain_msrcr.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -k "$K" -c 60 --csv --log-file gpurun_out/launches_msrcr.csv python tools/prof_run.py msrcr 8 > gpurun_out/ncu_msrcr_l.log 2>&1
if [ "${FULL:-1}" = "1" ]; then
PROF_WARMUP=0 ncu --set full --clock-control none --import-source on -k "$K" -c 12 -f -o gpurun_out/prof_dce python tools/prof_run.py dce 16 > gpurun_out/ncu_dce_f.log 2>&1
PROF_WARMUP=0 ncu --set full --clock-control none --import-source on -k "$K" -c 30 -f -o gpurun_out/prof_msrcr python tools/prof_run.py msrcr 8 > gpurun_out/ncu_msrcr_f.log 2>&1
fi
cat gpurun_out/plain_dce.log gpurun_out/plain_msrcr.log
tail -3 gpurun_out/pytest_gpu.log
echo done
