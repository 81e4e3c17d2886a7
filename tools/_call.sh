mkdir -p gpurun_out
L=gpurun_out/r02_call3.log
: > $L
echo "=== ncu launch list, wave 16MP q3.5 (2 encodes)" >> $L
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/r02_wave_launches_16mp.csv python tools/profile_encode.py 2 >> $L 2>&1; echo "rc=$?" >> $L
echo "=== ncu launch list, wave 512x512 q8" >> $L
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/r02_wave_launches_512_q8.csv python tools/mp_check.py --sizes 512x512 --q 8 --time 0 >> $L 2>&1; echo "rc=$?" >> $L
tail -n 30 $L
