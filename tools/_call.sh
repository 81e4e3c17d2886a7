mkdir -p gpurun_out
L=gpurun_out/r02_call25.log
: > $L
run() { echo "--- $1" >> $L; env $1 MNENC_BENCH_SKIP_EXTRAS=2 timeout 200 python bench.py --steps 10 --warmup 3 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('value',d['value'],'e2e',d['e2e']['value'],'mp_ms',d['roofline']['ms_per_launch'],'lat',d['single_image_latency_ms'])" >> $L 2>&1; }
run "MNENC_LANES=6"
run "MNENC_LANES=8"
run "MNENC_LANES=12"
run "MNENC_FRAME_GROUP=2"
run "MNENC_FRAME_GROUP=2 MNENC_LANES=4"
run "MNENC_FRAME_GROUP=4 MNENC_LANES=4"
cat $L
echo "=== failure-injection test" >> $L
timeout 200 python -m pytest tests/test_gpu_parity.py -x -q -k "failed_encode" 2>&1 | tail -n 3 >> $L
tail -n 4 $L
