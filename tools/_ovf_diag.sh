L=low_light_image_enhancement_b200/libllie.so
cp $L /tmp/good.so
cp gpurun_ab/libllie_e8.so $L
python tools/stress_dce.py 2>&1 | tail -34
python bench.py --workload c2 --steps 5 --no-e2e --no-cpu --no-also 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('E8', d['ms_per_step'], d['roofline']['kernel_ms_per_step'])"
cp /tmp/good.so $L
python bench.py --workload c2 --steps 5 --no-e2e --no-cpu --no-also 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('HEAD', d['ms_per_step'], d['roofline']['kernel_ms_per_step'])"
