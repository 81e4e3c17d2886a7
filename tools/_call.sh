mkdir -p gpurun_out
L=gpurun_out/r02_call18.log
: > $L
echo "=== multi-GPU context, one process (60 s limit)" >> $L
timeout 60 python tools/multi_check.py 2 >> $L 2>&1; echo "rc=$?" >> $L
echo "=== bench N=2" >> $L
( time timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus 2 --steps 10 --warmup 3 ) > gpurun_out/r02_bench_n2.json 2>> $L; echo "rc=$?" >> $L
tail -n 1 gpurun_out/r02_bench_n2.json | cut -c1-6000 >> $L
echo "=== batch check (1 GPU)" >> $L
timeout 200 python tools/batch_check.py 2>&1 | tail -n 2 >> $L
tail -c 7000 $L
