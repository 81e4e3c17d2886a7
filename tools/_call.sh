mkdir -p gpurun_out
L=gpurun_out/r02_call29.log
: > $L
echo "=== smoke" >> $L
timeout 120 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -n 2 >> $L
echo "=== pytest -m gpu" >> $L
timeout 1500 python -m pytest tests/ -x -q -m gpu 2>&1 | tail -n 6 >> $L
echo "=== bench N=1" >> $L
( time timeout 900 python bench.py ) > gpurun_out/r02_bench_n1_c.json 2>> $L; echo "rc=$?" >> $L
tail -n 1 gpurun_out/r02_bench_n1_c.json | cut -c1-300 >> $L
tail -n 30 $L
