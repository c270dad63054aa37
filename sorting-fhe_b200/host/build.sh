#!/bin/bash
# Build libsortfhe_host.so: the sorter layer (host C++17) linked against the engine's C-ABI.
set -e
cd "$(dirname "$0")"
g++ -O2 -std=c++17 -fPIC -fopenmp -shared -o ../libsortfhe_host.so sign.cpp capi_host.cpp -L.. -lckks_b200 -Wl,-rpath,'$ORIGIN'
echo "built ../libsortfhe_host.so"
