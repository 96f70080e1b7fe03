#!/bin/bash
# Index-safety check without a GPU: the kernel source compiled for the host (tests/host_emulation) is rebuilt with
# AddressSanitizer and the emulation tests are run against it; the host arrays have exactly the device arrays'
# sizes, so an out-of-bounds tile / row / lane index of any phase function aborts the run.
# (compute-sanitizer is not available on the GPU pool.)
set -e
cd "$(dirname "$0")/.."
lib=tests/host_emulation/libmcfd_emu.so
cp $lib /tmp/libmcfd_emu.keep 2>/dev/null || true
g++ -std=c++17 -O1 -g -fPIC -shared -ffp-contract=off -fsanitize=address -fno-omit-frame-pointer \
    -o $lib tests/host_emulation/emu.cpp
LD_PRELOAD=$(g++ -print-file-name=libasan.so) ASAN_OPTIONS=detect_leaks=0 \
    python -m pytest tests/test_host_logic.py -x -q -k "kernel_source or weno or energy or viscous or cell_viscosity"
rc=$?
[ -f /tmp/libmcfd_emu.keep ] && cp /tmp/libmcfd_emu.keep $lib && touch $lib
exit $rc
