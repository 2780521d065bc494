#!/usr/bin/env bash
# Builds libmiocodec_b200.so (sm_100a only) in-tree: csrc/*.cu -> miocodec_b200/libmiocodec_b200.so
set -euo pipefail
HERE="$(cd "$(dirname "${BASH_SOURCE[0]}")" && pwd)"
SRC="$HERE/csrc"
OUT="$HERE/miocodec_b200"
OBJ="$HERE/build"
NVCC=${NVCC:-/usr/local/cuda/bin/nvcc}
mkdir -p "$OBJ"
FLAGS=(-gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC -Xcompiler -O2 -Xptxas -v)
pids=()
for f in gemm_simt gemm3 attention elementwise istft engine stream capi noalloc; do
  if [ ! -f "$OBJ/$f.o" ] || [ "$SRC/$f.cu" -nt "$OBJ/$f.o" ] || [ -n "$(find "$SRC" -name '*.h' -newer "$OBJ/$f.o" -o -name '*.cuh' -newer "$OBJ/$f.o" | head -1)" ] || [ "$HERE/../include/miocodec_b200.h" -nt "$OBJ/$f.o" ]; then
    ( "$NVCC" "${FLAGS[@]}" -c "$SRC/$f.cu" -o "$OBJ/$f.o" > "$OBJ/$f.log" 2>&1 || { cat "$OBJ/$f.log"; exit 1; } ) &
    pids+=($!)
  fi
done
g++ -O2 -std=c++17 -fPIC -c "$SRC/gguf_reader.cpp" -o "$OBJ/gguf_reader.o"
for p in "${pids[@]}"; do wait "$p"; done
"$NVCC" -gencode arch=compute_100a,code=sm_100a -shared -o "$OUT/libmiocodec_b200.so" "$OBJ"/gemm_simt.o "$OBJ"/gemm3.o "$OBJ"/attention.o "$OBJ"/elementwise.o "$OBJ"/istft.o "$OBJ"/engine.o "$OBJ"/stream.o "$OBJ"/capi.o "$OBJ"/noalloc.o "$OBJ"/gguf_reader.o -cudart static
echo "built $OUT/libmiocodec_b200.so"
# C++ drop-in layer (reference-named entry points on top of the C ABI) + its test program
g++ -O2 -std=c++17 -fPIC -shared -I"$HERE/shim" "$HERE/shim/miocodec_shim.cpp" "$HERE/shim/batch_tts.cpp" -o "$OUT/libmiocodec_shim.so" \
    -L"$OUT" -lmiocodec_b200 -Wl,-rpath,'$ORIGIN'
g++ -O2 -std=c++17 -I"$HERE/shim" "$HERE/../tests/shim_demo.cpp" -o "$OUT/shim_demo" -L"$OUT" -lmiocodec_shim -lmiocodec_b200 \
    -Wl,-rpath,'$ORIGIN'
g++ -O2 -std=c++17 -I"$HERE/shim" "$HERE/../tests/batch_tts_demo.cpp" -o "$OUT/batch_tts_demo" -L"$OUT" -lmiocodec_shim -lmiocodec_b200 \
    -Wl,-rpath,'$ORIGIN'
echo "built $OUT/libmiocodec_shim.so, $OUT/shim_demo and $OUT/batch_tts_demo"
