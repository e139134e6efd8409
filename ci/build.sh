#!/bin/bash
# CI build step (reference: ci/build.sh): configure + build everything for sm_100a. Needs no GPU.
set -euo pipefail
cd "$(dirname "$0")/.."
cmake -S . -B build -G Ninja -DCMAKE_BUILD_TYPE="${BUILD_TYPE:-Release}"
ninja -C build
