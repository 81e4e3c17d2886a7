mkdir -p gpurun_out
L=gpurun_out/r02_call7.log
: > $L
echo "=== wave check" >> $L
MNENC_WAVE_TRACE=1 timeout 300 python tools/mp_check.py 2>&1 | grep -v "wave round" >> $L; echo "rc=$?" >> $L
echo "=== ncu launch list, wave 16MP q3.5 (2 encodes)" >> $L
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/r02_wave_launches_16mp_c.csv python tools/profile_encode.py 2 >> $L 2>&1; echo "rc=$?" >> $L
tail -n 60 $L
