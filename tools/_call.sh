mkdir -p gpurun_out
L=gpurun_out/r02_call27.log
: > $L
echo "=== wave check" >> $L
timeout 90 python tools/mp_check.py --sizes 64x64,200x136 --q 3.5 --time 0 >> $L 2>&1; rc=$?; echo "rc=$rc" >> $L
if [ $rc -ne 0 ]; then tail -n 30 $L; exit 0; fi
timeout 200 python tools/mp_check.py 2>&1 >> $L; echo "rc=$?" >> $L
echo "=== ncu launch list" >> $L
timeout 200 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/r02_wave_launches_16mp_h.csv python tools/profile_encode.py 2 >> $L 2>&1; echo "rc=$?" >> $L
python tools/agg_launches.py gpurun_out/r02_wave_launches_16mp_h.csv | grep -i "wave\|matching" >> $L 2>&1
grep -v "==WARNING\|==PROF" $L | tail -n 40
