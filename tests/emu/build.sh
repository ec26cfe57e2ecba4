#!/bin/sh
# TEST INFRASTRUCTURE: builds tests/emu/libfk_emu.so (CPU run of the kernels' per-thread code).
set -e
here=$(cd "$(dirname "$0")" && pwd)
src=$here/../../tinyfk_b200/csrc
${CXX:-g++} -std=c++17 -O2 -fPIC -ffp-contract=off -Wall -shared -Wl,-Bsymbolic -Wl,--exclude-libs,ALL \
  -o "$here/libfk_emu.so" "$here/emu_driver.cpp" "$src/model.cpp" "$src/plan.cpp" "$src/xml_lite.cpp" "$src/disk_cache.cpp" -lm
