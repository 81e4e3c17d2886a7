#!/bin/bash
# First GPU-box call of the next round: everything that was written after round 1's GPU minutes ran out,
# verified and A/B-measured in one go. Usage:
#   gpurun --timeout 900 -- 'bash tools/next_round_first_call.sh'
# Writes gpurun_out/r02_first_call.log (+ the ncu report of the reconstruct kernel). Each step has its own timeout.
mkdir -p gpurun_out
L=gpurun_out/r02_first_call.log
: > "$L"
step() { echo "=== $1" | tee -a "$L"; shift; "$@" >> "$L" 2>&1; echo "rc=$?" | tee -a "$L"; }

# 1. the variants must give the reference decoder's pixels before any timing counts
step "decode tests, default"            timeout 120 python -m pytest tests/test_gpu_decode.py -x -q
step "decode tests, MNENC_RECON=2"      env MNENC_RECON=2 timeout 120 python -m pytest tests/test_gpu_decode.py -x -q
step "decode tests, parallel reader"    env MNENC_DECODE_THREADS=8 timeout 120 python -m pytest tests/test_gpu_decode.py -x -q

step "encode tests, MNENC_PACK=2"       env MNENC_PACK=2 timeout 300 python -m pytest tests/test_gpu_parity.py tests/test_gpu_stripes.py -x -q
step "encode tests, MNENC_PACK=3"       env MNENC_PACK=3 timeout 300 python -m pytest tests/test_gpu_parity.py tests/test_gpu_stripes.py -x -q
step "CLI sweep (-g), first GPU run"    timeout 120 python -m pytest tests/test_zz_cli_sweep.py -x -q -m gpu

# 2. A/B of the decoder (same box, back to back): serial vs parallel reader, kernel variants, lanes of the batch path
step "decode bench, default"            timeout 90 python tools/decode_bench.py --cpu 0
step "decode bench, MNENC_RECON=2"      env MNENC_RECON=2 timeout 90 python tools/decode_bench.py --cpu 0
step "decode bench, 8 reader threads"   env MNENC_DECODE_THREADS=8 timeout 90 python tools/decode_bench.py --cpu 0
step "decode bench, 12 lanes"           env MNENC_LANES=12 timeout 90 python tools/decode_bench.py --cpu 0
step "decode bench, 16 lanes"           env MNENC_LANES=16 timeout 90 python tools/decode_bench.py --cpu 0

# 3. encoder after the host table builder rework (host stage = ms[4] of mnenc_last_timing; bench line for the record)
step "bench.py"                         timeout 240 python bench.py --steps 3 --warmup 3

step "bench.py, MNENC_PACK=2"           env MNENC_PACK=2 timeout 240 python bench.py --steps 3 --warmup 3
step "bench.py, MNENC_PACK=3"           env MNENC_PACK=3 timeout 240 python bench.py --steps 3 --warmup 3

# 4. ncu of the decode kernels: one lane, no batch pass worth mentioning under the profiler
step "ncu reconstruct"                  env MNENC_LANES=1 timeout 150 ncu --set full --clock-control none --import-source on \
    -k regex:reconstructKernel -c 1 -f -o gpurun_out/r02_decode_reconstruct python tools/decode_bench.py --steps 1 --warmup 1 --cpu 0 --batch 0
step "ncu gather"                       env MNENC_LANES=1 timeout 150 ncu --set full --clock-control none --import-source on \
    -k regex:gatherKernel -c 1 -f -o gpurun_out/r02_decode_gather python tools/decode_bench.py --steps 1 --warmup 1 --cpu 0 --batch 0
tail -n 60 "$L"
