#!/bin/bash
# Builds libvulcan.so (the reference's C++ classes over the C ABI) and the C++ drop-in test.
set -e
ROOT="$(cd "$(dirname "$0")/../.." && pwd)"
CUDA="${CUDA_HOME:-/usr/local/cuda}"
OUT="$ROOT/vulcan_b200"
SRC="$ROOT/vulcan_b200/cxx/vulcan.cpp $ROOT/vulcan_b200/cxx/pipeline.cpp $ROOT/vulcan_b200/cxx/sharded.cpp $ROOT/vulcan_b200/cxx/sharded_pipeline.cpp"
LIB="$OUT/libvulcan.so"
FLAGS="-std=c++14 -O2 -fPIC -Wall -DNDEBUG -I$ROOT/include -I$CUDA/include"
if [ ! -f "$LIB" ] || [ "$ROOT/vulcan_b200/cxx/vulcan.cpp" -nt "$LIB" ] || [ "$ROOT/vulcan_b200/cxx/pipeline.cpp" -nt "$LIB" ] || [ "$ROOT/vulcan_b200/cxx/sharded.cpp" -nt "$LIB" ] || [ "$ROOT/vulcan_b200/cxx/sharded_pipeline.cpp" -nt "$LIB" ] || [ -n "$(find "$ROOT/include" -newer "$LIB" -name '*.h*' | head -1)" ]; then
  g++ $FLAGS -shared -o "$LIB" $SRC -L"$OUT" -lvulcan_b200 -L"$CUDA/lib64" -lcudart \
      -Wl,-rpath,'$ORIGIN' -Wl,-rpath,"$CUDA/lib64"
fi
TEST_SRC="$ROOT/tests/cxx/test_dropin.cpp"
TEST_BIN="$ROOT/tests/cxx/_bin/test_dropin"
if [ -f "$TEST_SRC" ]; then
  mkdir -p "$ROOT/tests/cxx/_bin"
  if [ ! -f "$TEST_BIN" ] || [ "$TEST_SRC" -nt "$TEST_BIN" ] || [ "$LIB" -nt "$TEST_BIN" ]; then
    g++ $FLAGS -o "$TEST_BIN" "$TEST_SRC" -L"$OUT" -lvulcan -lvulcan_b200 -L"$CUDA/lib64" -lcudart \
        -Wl,-rpath,'$ORIGIN/../../../vulcan_b200' -Wl,-rpath,"$CUDA/lib64"
  fi
fi
