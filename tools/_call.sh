mkdir -p gpurun_out
L=gpurun_out/r02_call14.log
: > $L
echo "=== ncu full: decide iteration 1, bin fill iteration 2 (of the 2nd encode)" >> $L
timeout 300 ncu --set full --import-source on --clock-control none -k regex:waveDecideKernel --launch-skip 15 --launch-count 1 -o gpurun_out/r02_wave_decide_v6 -f python tools/profile_encode.py 2 >> $L 2>&1; echo "rc=$?" >> $L
timeout 300 ncu --set full --import-source on --clock-control none -k regex:waveBinFillKernel --launch-skip 14 --launch-count 1 -o gpurun_out/r02_wave_binfill_v6 -f python tools/profile_encode.py 2 >> $L 2>&1; echo "rc=$?" >> $L
echo "=== pytest -m gpu" >> $L
timeout 1500 python -m pytest tests/ -x -q -m gpu >> $L 2>&1; echo "rc=$?" >> $L
grep -v "==WARNING\|==PROF" $L | tail -n 40
