mkdir -p gpurun_out
L=gpurun_out/r02_call24.log
: > $L
lscpu | grep -E "Model name|^CPU\(s\)" >> $L
echo "=== stripes over NCCL, 2 processes" >> $L
timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29517 tools/stripe_check.py 2>&1 | grep -v "OMP_NUM\|^\*\*\*\|^$" >> $L; echo "rc=$?" >> $L
echo "=== multi-GPU tests" >> $L
timeout 200 python -m pytest tests/test_gpu_stripes.py -x -q -k "multi_gpu or in_library or nccl" 2>&1 | tail -n 3 >> $L
echo "=== bench N=2" >> $L
( time timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus 2 --steps 10 --warmup 3 ) > gpurun_out/r02_bench_n2_b.json 2>> $L; echo "rc=$?" >> $L
tail -n 1 gpurun_out/r02_bench_n2_b.json | cut -c1-200 >> $L
grep -v "OMP_NUM\|^\*\*\*\|^$" $L | tail -n 40
