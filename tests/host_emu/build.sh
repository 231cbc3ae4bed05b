#!/bin/sh
# builds the host emulation of the env kernels (test infrastructure; see pf_host_emu.cpp)
set -e
here="$(cd "$(dirname "$0")" && pwd)"
mkdir -p "$here/_build"
g++ -std=c++17 -O2 -ffp-contract=off -fPIC -shared -Wall -Wno-unknown-pragmas "$here/pf_host_emu.cpp" -o "$here/_build/libpf_host_emu.so"
