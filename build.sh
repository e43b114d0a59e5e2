#!/usr/bin/env bash
# Builds raytracing-hw_b200/librtb.so (sm_100a kernels + C ABI + C++ host front end) and build/solution (the CLI).
set -e
here="$(cd "$(dirname "${BASH_SOURCE[0]}")" && pwd)"
make -C "$here/raytracing-hw_b200" -j"$(nproc)"
