#!/bin/bash
# tools/asan_host.sh -- the host-side parsers (csrc/nrrd.cpp, csv.cpp, hostside.cpp) built with AddressSanitizer +
# UndefinedBehaviorSanitizer and linked into a copy of the library; tests/test_host.py (reader differential tests, 94
# header cases, random headers, CSV error texts, huge-size headers) then runs against that build.
# compute-sanitizer is closed on this pool, so this is the memory check the parsers get.  CPU only.
set -e
cd "$(dirname "$0")/.."
B=/tmp/vr_asan; mkdir -p $B
make -s -j8 -C volumetric-renderer_b200
for f in nrrd csv hostside; do
  g++ -O1 -g -std=c++17 -fPIC -ffp-contract=off -fsanitize=address,undefined -fno-sanitize-recover=undefined -fno-omit-frame-pointer -c volumetric-renderer_b200/csrc/$f.cpp -o $B/$f.o
done
OBJS=$(ls volumetric-renderer_b200/build/*.o | grep -v -E '/(nrrd|csv|hostside)\.o')
nvcc -gencode arch=compute_100a,code=sm_100a -shared -o $B/libvr_b200_asan.so $OBJS $B/nrrd.o $B/csv.o $B/hostside.o -Xlinker -lasan -Xlinker -lubsan
ASAN=$(gcc -print-file-name=libasan.so)
LD_PRELOAD="$ASAN $(gcc -print-file-name=libstdc++.so)" ASAN_OPTIONS=detect_leaks=0:abort_on_error=1 UBSAN_OPTIONS=halt_on_error=1 VR_LIB_PATH=$B/libvr_b200_asan.so \
  python -m pytest tests/test_host.py -x -q -m "not gpu" 2>&1 | tail -5
