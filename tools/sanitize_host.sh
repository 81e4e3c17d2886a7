#!/bin/bash
# Host side of libmnenc.so (table builder, .mn reader, stripe planner) under AddressSanitizer + UBSan, driven by the CPU tests.
# The CUDA objects are reused as built (make -C imageexperiments_b200/csrc first); no GPU needed. Round 1: clean.
set -e
ROOT=$(cd "$(dirname "$0")/.." && pwd)
SRC=$ROOT/imageexperiments_b200/csrc
OUT=${TMPDIR:-/tmp}/mnenc_asan
mkdir -p "$OUT"
for f in entropy_host stripe_host decode_host; do
    g++ -std=c++20 -O1 -g -fno-omit-frame-pointer -fsanitize=address,undefined -ffp-contract=off -fPIC -c "$SRC/$f.cpp" -o "$OUT/$f.o"
done
GCCLIB=$(dirname "$(gcc -print-file-name=libasan.so)")
nvcc -gencode arch=compute_100a,code=sm_100a -shared -o "$OUT/libmnenc_asan.so" \
    $(ls "$SRC"/build/*.o | grep -v -E '/(entropy_host|stripe_host|decode_host)\.o$') "$OUT"/*.o \
    -Xlinker --exclude-libs,ALL -L"$GCCLIB" -lasan -lubsan
cd "$ROOT"
LD_PRELOAD="$GCCLIB/libasan.so $GCCLIB/libubsan.so" ASAN_OPTIONS=detect_leaks=0 UBSAN_OPTIONS=print_stacktrace=1 \
    MNENC_LIB="$OUT/libmnenc_asan.so" python -m pytest tests/test_decode_host.py tests/test_host_planner.py tests/test_stripe_host.py \
    tests/test_quant_tables.py tests/test_abi.py -x -q -s 2>&1 | tee "$OUT/run.log" | tail -n 3
echo "sanitizer reports: $(grep -c -E 'runtime error|ERROR: AddressSanitizer' "$OUT/run.log" || true)"
