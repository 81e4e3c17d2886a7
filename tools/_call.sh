mkdir -p gpurun_out
L=gpurun_out/r02_call11.log
: > $L
echo "=== wave check (small first, short limits)" >> $L
timeout 90 python tools/mp_check.py --sizes 64x64 --q 3.5 --time 0 >> $L 2>&1; rc=$?; echo "rc=$rc" >> $L
if [ $rc -ne 0 ]; then tail -n 30 $L; exit 0; fi
MNENC_WAVE_TRACE=1 timeout 200 python tools/mp_check.py 2>&1 | grep -v "wave round" >> $L; echo "rc=$?" >> $L
echo "=== ncu launch list, wave 16MP q3.5 (2 encodes)" >> $L
timeout 200 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/r02_wave_launches_16mp_e.csv python tools/profile_encode.py 2 >> $L 2>&1; echo "rc=$?" >> $L
echo "=== ncu full: base iteration 1, decide iteration 1, detail iteration 2 (of the 2nd encode)" >> $L
timeout 300 ncu --set full --import-source on --clock-control none -k regex:waveBaseKernel --launch-skip 15 --launch-count 1 -o gpurun_out/r02_wave_base_v5 -f python tools/profile_encode.py 2 >> $L 2>&1; echo "rc=$?" >> $L
timeout 300 ncu --set full --import-source on --clock-control none -k regex:waveDecideKernel --launch-skip 15 --launch-count 1 -o gpurun_out/r02_wave_decide_v5 -f python tools/profile_encode.py 2 >> $L 2>&1; echo "rc=$?" >> $L
timeout 300 ncu --set full --import-source on --clock-control none -k regex:waveDetailKernel --launch-skip 14 --launch-count 1 -o gpurun_out/r02_wave_detail_v5 -f python tools/profile_encode.py 2 >> $L 2>&1; echo "rc=$?" >> $L
grep -v "^wave candidates\|==WARNING\|==PROF" $L | tail -n 50
grep "^wave candidates" $L | tail -n 1
