mkdir -p gpurun_out
L=gpurun_out/r02_call30.log
: > $L
timeout 300 python -m pytest tests/test_gpu_parity.py tests/test_cli.py -x -q -m gpu -k "built or builtin or failed_encode" 2>&1 | tail -n 8 >> $L
cat $L
