mkdir -p gpurun_out
L=gpurun_out/r02_call19.log
: > $L
echo "=== smoke" >> $L
timeout 120 python -c "import __graft_entry__ as g; g.smoke()" >> $L 2>&1; echo "rc=$?" >> $L
echo "=== DRAM traffic of the matching-pursuit launches (ncu metrics list)" >> $L
timeout 300 ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/r02_mp_traffic.csv python tools/profile_encode.py 2 >> $L 2>&1; echo "rc=$?" >> $L
python tools/mp_traffic_summary.py gpurun_out/r02_mp_traffic.csv gpurun_out/r02_mp_ncu_summary.json >> $L 2>&1
echo "=== compute-sanitizer memcheck (200x136, q 3.5 and 8)" >> $L
timeout 600 compute-sanitizer --tool memcheck --error-exitcode 9 python tools/mp_check.py --sizes 200x136 --q 3.5,8 --time 0 2>&1 | tail -n 8 >> $L; echo "rc=${PIPESTATUS[0]}" >> $L
echo "=== compute-sanitizer racecheck (64x64, q 3.5)" >> $L
timeout 600 compute-sanitizer --tool racecheck --error-exitcode 9 python tools/mp_check.py --sizes 64x64 --q 3.5 --time 0 2>&1 | tail -n 8 >> $L; echo "rc=${PIPESTATUS[0]}" >> $L
echo "=== decode bench (defaults)" >> $L
timeout 200 python tools/decode_bench.py 2>&1 | tail -n 12 >> $L
echo "=== ncu full: base / detail (v6), iteration 1 of the 2nd encode" >> $L
timeout 300 ncu --set full --import-source on --clock-control none -k regex:waveBaseKernel --launch-skip 15 --launch-count 1 -o gpurun_out/r02_wave_base_v6 -f python tools/profile_encode.py 2 > /dev/null 2>&1; echo "rc=$?" >> $L
timeout 300 ncu --set full --import-source on --clock-control none -k regex:waveDetailKernel --launch-skip 14 --launch-count 1 -o gpurun_out/r02_wave_detail_v6 -f python tools/profile_encode.py 2 > /dev/null 2>&1; echo "rc=$?" >> $L
grep -v "==WARNING\|==PROF" $L | tail -n 70
