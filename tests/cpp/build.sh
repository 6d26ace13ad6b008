#!/bin/sh
# tests/cpp/build.sh -- compile the C++ host-mirror test against libnnbvh.so and the oracle.
set -e
HERE="$(cd "$(dirname "$0")" && pwd)"
ROOT="$HERE/../.."
make -s -C "$ROOT/oracle" oracle
g++ -std=c++17 -O2 -ffp-contract=off -o "$HERE/host_mirror_test" "$HERE/host_mirror_test.cpp" \
    -L"$ROOT/nn_bvh_b200" -lnnbvh -L"$ROOT/oracle/_build" -lnnbvh_oracle \
    -Wl,-rpath,'$ORIGIN/../../nn_bvh_b200' -Wl,-rpath,'$ORIGIN/../../oracle/_build' -pthread
echo "built $HERE/host_mirror_test"
g++ -std=c++17 -O2 -ffp-contract=off -fPIC -shared -I"$ROOT/include" \
    -o "$HERE/libwavefront_shim.so" "$HERE/wavefront_shim.cpp" \
    -L"$ROOT/nn_bvh_b200" -lnnbvh -Wl,-rpath,'$ORIGIN/../../nn_bvh_b200' -pthread
echo "built $HERE/libwavefront_shim.so"
g++ -std=c++17 -O2 -ffp-contract=off -fPIC -shared -w -I"$HERE/cuda_stub" \
    -o "$HERE/libdevice_twin.so" "$HERE/device_twin.cpp"
echo "built $HERE/libdevice_twin.so"
