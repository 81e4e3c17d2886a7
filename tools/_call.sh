mkdir -p gpurun_out
L=gpurun_out/r02_call21.log
: > $L
nvidia-smi -L | wc -l >> $L; nproc >> $L
echo "=== multi-GPU context, one process, 8 GPUs (90 s limit)" >> $L
timeout 90 python tools/multi_check.py 8 >> $L 2>&1; echo "rc=$?" >> $L
for N in 8 4 1; do
  echo "=== bench N=$N" >> $L
  if [ $N -eq 1 ]; then
    ( time MNENC_BENCH_SKIP_EXTRAS=1 timeout 400 python bench.py --steps 10 --warmup 3 ) > gpurun_out/r02_scale_n$N.json 2>> $L; echo "rc=$?" >> $L
  else
    ( time timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2953$N bench.py --gpus $N --steps 10 --warmup 3 ) > gpurun_out/r02_scale_n$N.json 2>> $L; echo "rc=$?" >> $L
  fi
  tail -n 1 gpurun_out/r02_scale_n$N.json | cut -c1-200 >> $L
done
grep -v "OMP_NUM_THREADS\|^\*\*\*\|^$" $L | tail -n 60
