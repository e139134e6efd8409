#!/bin/bash
# Environment of the CI steps (reference: ci/env.sh exported the CMake / CUDA / MPI prefixes it had installed).
# Nothing is installed here: the image brings CUDA 12.9, CMake and Ninja; this only makes them visible.
export CUDA_HOME="${CUDA_HOME:-/usr/local/cuda}"
export PATH="$CUDA_HOME/bin:$PATH"
export CMAKE_BUILD_TYPE="${CMAKE_BUILD_TYPE:-Release}"
export STENCIL_LOG_LEVEL="${STENCIL_LOG_LEVEL:-2}"
