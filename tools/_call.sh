mkdir -p gpurun_out
L=gpurun_out/r02_call22.log
: > $L
echo "=== batch check" >> $L
timeout 300 python tools/batch_check.py 2>&1 | tail -n 3 >> $L
echo "=== pytest -m gpu" >> $L
timeout 1500 python -m pytest tests/ -x -q -m gpu 2>&1 | tail -n 6 >> $L
echo "=== bench N=1" >> $L
( time timeout 900 python bench.py ) > gpurun_out/r02_bench_n1_b.json 2>> $L; echo "rc=$?" >> $L
tail -n 1 gpurun_out/r02_bench_n1_b.json | cut -c1-300 >> $L
tail -n 40 $L
