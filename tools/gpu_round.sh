#!/bin/bash
# One GPU-box pass for the evidence under profiles/ (run under gpurun): launch lists (ncu gpu__time_duration only) and
# `--set full` captures of one Zero-DCE call (16 x 1080p) and one MSRCR call (8 x 4K, fp32 and uint8 images).
#   R=r03 tools/gpu_round.sh
set -u
R="${R:-r03}"
mkdir -p gpurun_out
K='regex:dce_|msrcr_|down4_|decim4_|fir_|up_bc_'
python tools/prof_run.py dce 16 > gpurun_out/plain_dce.log 2>&1 || { echo "plain dce run failed"; exit 1; }
python tools/prof_run.py msrcr 8 > gpurun_out/plain_msrcr.log 2>&1 || { echo "plain msrcr run failed"; exit 1; }
cat gpurun_out/plain_dce.log gpurun_out/plain_msrcr.log
# launch lists: warm-up call + one call
ncu --metrics gpu__time_duration.sum --clock-control none -k "$K" -c 40 --csv --log-file gpurun_out/${R}_ncu_launches_dce_16x1080p.csv python tools/prof_run.py dce 16 > gpurun_out/ncu_dce_l.log 2>&1
ncu --metrics gpu__time_duration.sum --clock-control none -k "$K" -c 60 --csv --log-file gpurun_out/${R}_ncu_launches_msrcr_8x4k.csv python tools/prof_run.py msrcr 8 > gpurun_out/ncu_msrcr_l.log 2>&1
# full captures of the second (warm) call only: prof_run prints the launches per call (4 for dce, 14 for msrcr)
PROF_WARMUP=1 ncu --set full --clock-control none --import-source on -k "$K" -s 4 -c 4 -f -o gpurun_out/${R}_prof_dce python tools/prof_run.py dce 16 > gpurun_out/ncu_dce_f.log 2>&1
PROF_WARMUP=1 ncu --set full --clock-control none --import-source on -k "$K" -s 14 -c 14 -f -o gpurun_out/${R}_prof_msrcr python tools/prof_run.py msrcr 8 > gpurun_out/ncu_msrcr_f.log 2>&1
PROF_WARMUP=1 ncu --set full --clock-control none -k "$K" -s 14 -c 14 -f -o gpurun_out/${R}_prof_msrcr_u8 python tools/prof_run.py msrcr_u8 8 > gpurun_out/ncu_msrcr_u8_f.log 2>&1
# config 1 (fp32-accurate split path, one 1080p frame): 8 launches per call (pack + 7 layers)
python tools/prof_run.py dce32 1 > gpurun_out/plain_dce32.log 2>&1 && cat gpurun_out/plain_dce32.log
ncu --metrics gpu__time_duration.sum --clock-control none -k "$K" -c 16 --csv --log-file gpurun_out/${R}_ncu_launches_dce32_1x1080p.csv python tools/prof_run.py dce32 1 > gpurun_out/ncu_dce32_l.log 2>&1
PROF_WARMUP=1 ncu --set full --clock-control none -k "$K" -s 8 -c 8 -f -o gpurun_out/${R}_prof_dce32 python tools/prof_run.py dce32 1 > gpurun_out/ncu_dce32_f.log 2>&1
for n in dce msrcr msrcr_u8 dce32; do
  ncu -i gpurun_out/${R}_prof_$n.ncu-rep --page raw --csv > gpurun_out/${R}_raw_$n.csv 2>/dev/null
done
ls -la gpurun_out/*.ncu-rep
echo done
