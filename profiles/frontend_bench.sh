#!/bin/bash
# builds and runs profiles/frontend_bench.cpp against the in-tree libqpolicy.so
set -e
cd "$(dirname "$0")/.."
LIBDIR=scala_akka_tensorflow_sharetrade_helper_b200
mkdir -p gpurun_out
g++ -O2 -std=c++17 -pthread -Iinclude profiles/frontend_bench.cpp -o gpurun_out/frontend_bench -L$LIBDIR -lqpolicy -Wl,-rpath,$PWD/$LIBDIR
gpurun_out/frontend_bench "$@"
