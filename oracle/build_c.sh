#!/usr/bin/env bash
# Builds the C restatement of the oracle (CPU baseline) into oracle/_build/.
# The reference itself is Python only, so there is nothing to compile into oracle/_ref/.
set -euo pipefail
HERE="$(cd "$(dirname "${BASH_SOURCE[0]}")" && pwd)"
mkdir -p "$HERE/_build"
gcc -O3 -march=x86-64-v2 -fopenmp -fno-fast-math -ffp-contract=off -fPIC -shared \
    "$HERE/apex_oracle.c" -o "$HERE/_build/liboracle.so" -lm
echo "built $HERE/_build/liboracle.so"
