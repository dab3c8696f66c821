mkdir -p gpurun_out
K='regex:dce_'
python tools/prof_run.py dce 16 > gpurun_out/plain_dce.log 2>&1 &&
PROF_WARMUP=0 ncu --set full --clock-control none --import-source on -k "$K" -c 4 -f -o gpurun_out/prof_dce python tools/prof_run.py dce 16 > gpurun_out/ncu_dce_f.log 2>&1
cat gpurun_out/plain_dce.log; tail -2 gpurun_out/ncu_dce_f.log
