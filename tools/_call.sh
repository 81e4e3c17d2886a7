mkdir -p gpurun_out
L=gpurun_out/r02_call17.log
: > $L
echo "=== batch check" >> $L
timeout 300 python tools/batch_check.py >> $L 2>&1; echo "rc=$?" >> $L
echo "=== per-frame path for comparison (MNENC_FRAME_GROUP=1)" >> $L
MNENC_FRAME_GROUP=1 timeout 300 python tools/batch_check.py 2>&1 | tail -n 1 >> $L
echo "=== groups of 4 / 16" >> $L
MNENC_FRAME_GROUP=4 timeout 300 python tools/batch_check.py 2>&1 | tail -n 1 >> $L
MNENC_FRAME_GROUP=16 timeout 300 python tools/batch_check.py 2>&1 | tail -n 1 >> $L
echo "=== 16 MP timing with the restored decide pass" >> $L
timeout 200 python tools/mp_check.py --sizes 64x64 --q 3.5 2>&1 | tail -n 3 >> $L
tail -n 40 $L
