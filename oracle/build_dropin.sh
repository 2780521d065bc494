#!/usr/bin/env bash
# TEST INFRASTRUCTURE ONLY (the drop-in proof, VERDICT r1 "what's missing" item 2).
#
# Builds the reference's OWN executables from its unmodified sources where they lie under /root/reference:
#   miotts                   (src/main.cpp,               CMakeLists.txt:17-37)
#   miotts-stream-compare    (examples/stream-compare.cpp)
#   miotts-stream-benchmark  (examples/stream-benchmark.cpp)
# twice:
#   oracle/_ref/bin/<name>                                   linked the stock way: the reference's codec + ISTFT + ggml
#                                                            (all inside oracle/_ref/libmio_ref_v3.so)
#   miotts-llama.cpp_b200/miocodec_b200/dropin/<name>        the same caller sources (main / example, test-to-speech.cpp,
#                                                            token-parser.cpp, wav-writer.cpp, text-normalize.cpp) compiled
#                                                            against the REFERENCE's headers and linked against
#                                                            libmiocodec_shim.so + libmiocodec_b200.so instead of
#                                                            src/miocodec.cpp + src/istft.cpp + ggml
# libllama is replaced in both by oracle/stub_llama.cpp (token generation is out of scope; --skip-llm never touches it,
# and the LLM branch gets scripted codes).  tests/test_gpu_dropin.py runs both sets on the GPU box and compares them.
set -euo pipefail
REF=${MIO_REFERENCE_ROOT:-/root/reference}
HERE="$(cd "$(dirname "${BASH_SOURCE[0]}")" && pwd)"
ROOT="$(dirname "$HERE")"
PKG="$ROOT/miotts-llama.cpp_b200/miocodec_b200"
if [ ! -d "$REF/src" ]; then echo "reference not present at $REF: keeping prebuilt drop-in binaries" >&2; exit 0; fi
GG="$REF/third_party/llama.cpp/ggml"
LL="$REF/third_party/llama.cpp/include"
INC=(-I"$REF/src" -I"$LL" -I"$GG/include")
FP=(-std=gnu++17 -O3 -DNDEBUG)          # the reference's Release flags for first-party code: no -march (CMakeLists.txt sets none)
mkdir -p "$HERE/_ref/bin" "$PKG/dropin" "$HERE/_ref/obj_dropin"
O="$HERE/_ref/obj_dropin"
pids=()
for f in test-to-speech token-parser wav-writer text-normalize main; do
  g++ "${FP[@]}" "${INC[@]}" -c "$REF/src/$f.cpp" -o "$O/$f.o" & pids+=($!)
done
for f in stream-compare stream-benchmark; do
  g++ "${FP[@]}" "${INC[@]}" -c "$REF/examples/$f.cpp" -o "$O/$f.o" & pids+=($!)
done
g++ "${FP[@]}" "${INC[@]}" -c "$HERE/stub_llama.cpp" -o "$O/stub_llama.o" & pids+=($!)
for p in "${pids[@]}"; do wait "$p"; done
CALLERS=("$O/test-to-speech.o" "$O/token-parser.o" "$O/wav-writer.o" "$O/text-normalize.o" "$O/stub_llama.o")
link() {   # name, main object
  # ours: callers + shim (exports the reference's miocodec_* / istft* symbols) -- THE link line INTEGRATION.md quotes
  g++ -o "$PKG/dropin/$1" "$2" "${CALLERS[@]}" -L"$PKG" -lmiocodec_shim -lmiocodec_b200 -Wl,-rpath,'$ORIGIN/..' -lpthread -lm
  # stock: the library already holds test-to-speech / token-parser / wav-writer / text-normalize / stub llama objects
  g++ -o "$HERE/_ref/bin/$1" "$2" -L"$HERE/_ref" -lmio_ref_v3 -Wl,-rpath,'$ORIGIN/..' -lpthread -lm
}
link miotts "$O/main.o"
link miotts-stream-compare "$O/stream-compare.o"
link miotts-stream-benchmark "$O/stream-benchmark.o"
rm -rf "$O"
echo "built $PKG/dropin/{miotts,miotts-stream-compare,miotts-stream-benchmark} and $HERE/_ref/bin/*"
