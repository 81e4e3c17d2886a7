mkdir -p gpurun_out
L=gpurun_out/r02_call20.log
: > $L
echo "=== multi-GPU context, one process (60 s limit)" >> $L
timeout 60 python tools/multi_check.py 2 >> $L 2>&1; echo "rc=$?" >> $L
tail -n 30 $L
