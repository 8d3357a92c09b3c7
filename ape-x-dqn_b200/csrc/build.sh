#!/usr/bin/env bash
# Builds libapexb200.so (the C-ABI library) in-tree for sm_100a.
set -euo pipefail
HERE="$(cd "$(dirname "${BASH_SOURCE[0]}")" && pwd)"
ROOT="$(cd "$HERE/../.." && pwd)"
OUT="$HERE/../libapexb200.so"
NVCC="${NVCC:-nvcc}"
"$NVCC" -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 \
    -Xcompiler -fPIC -shared \
    -I"$ROOT/include" -I"$HERE" ${APEX_NVCC_EXTRA:-} \
    "$HERE/replay.cu" "$HERE/tree.cu" "$HERE/gather.cu" "$HERE/td.cu" "$HERE/publish.cu" "$HERE/ingest.cu" "$HERE/pack.cu" \
    -o "$OUT"
echo "built $OUT"
