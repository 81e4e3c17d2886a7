mkdir -p gpurun_out
L=gpurun_out/r02_call28.log
: > $L
for lanes in 6 9 12 16; do
  echo "--- MNENC_LANES=$lanes, 192 frames" >> $L
  MNENC_LANES=$lanes BATCH_FRAMES=192 BATCH_SKIP_PARITY=1 timeout 120 python tools/batch_check.py 2>&1 | tail -n 1 >> $L
done
echo "--- MNENC_LANES=12 MNENC_FRAME_GROUP=8" >> $L
MNENC_LANES=12 MNENC_FRAME_GROUP=8 BATCH_FRAMES=192 BATCH_SKIP_PARITY=1 timeout 120 python tools/batch_check.py 2>&1 | tail -n 1 >> $L
echo "--- MNENC_LANES=16 MNENC_FRAME_GROUP=6" >> $L
MNENC_LANES=16 MNENC_FRAME_GROUP=6 BATCH_FRAMES=192 BATCH_SKIP_PARITY=1 timeout 120 python tools/batch_check.py 2>&1 | tail -n 1 >> $L
cat $L
