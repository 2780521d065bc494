#!/usr/bin/env bash
# TEST INFRASTRUCTURE ONLY.
# Builds the UNMODIFIED reference hot path (src/miocodec.cpp + src/istft.cpp) together
# with the vendored ggml base + CPU backend, compiled directly with gcc/g++ from the
# sources where they lie under /root/reference (no cmake, nothing copied), plus our
# extern "C" wrapper oracle/ref_capi.cpp, into oracle/_ref/libmio_ref_<variant>.so.
#
# Two ISA variants are produced so the same prebuilt files work on whatever host CPU
# the GPU box has: "v3" (x86-64-v3: AVX2+FMA+F16C) and "v4" (x86-64-v4: AVX-512).
# oracle/ref.py picks the best one the running CPU supports.
set -euo pipefail
REF=${MIO_REFERENCE_ROOT:-/root/reference}
HERE="$(cd "$(dirname "${BASH_SOURCE[0]}")" && pwd)"
OUT="$HERE/_ref"
GG="$REF/third_party/llama.cpp/ggml"
if [ ! -d "$GG/src" ]; then echo "reference not present at $REF (expected on the GPU box): keeping prebuilt oracle/_ref" >&2; exit 0; fi
mkdir -p "$OUT"

C_SRCS=(ggml.c ggml-alloc.c ggml-quants.c ggml-cpu/ggml-cpu.c ggml-cpu/quants.c ggml-cpu/arch/x86/quants.c)
CXX_SRCS=(ggml.cpp ggml-backend.cpp ggml-backend-reg.cpp ggml-backend-dl.cpp ggml-opt.cpp ggml-threading.cpp gguf.cpp
          ggml-cpu/ggml-cpu.cpp ggml-cpu/repack.cpp ggml-cpu/hbm.cpp ggml-cpu/traits.cpp
          ggml-cpu/amx/amx.cpp ggml-cpu/amx/mmq.cpp ggml-cpu/binary-ops.cpp ggml-cpu/unary-ops.cpp
          ggml-cpu/vec.cpp ggml-cpu/ops.cpp ggml-cpu/arch/x86/repack.cpp ggml-cpu/arch/x86/cpu-feats.cpp
          ggml-cpu/llamafile/sgemm.cpp)
DEFS=(-DGGML_VERSION=\"0.9.5\" -DGGML_COMMIT=\"079feab\" -D_GNU_SOURCE -D_XOPEN_SOURCE=600
      -DGGML_USE_CPU -DGGML_USE_LLAMAFILE -DGGML_CPU_REPACK -DNDEBUG)
INCS=(-I"$GG/include" -I"$GG/src" -I"$GG/src/ggml-cpu")

build_variant() {
  local tag=$1; shift
  local arch=("$@")
  local obj="$OUT/obj_$tag"
  mkdir -p "$obj"
  local pids=()
  for s in "${C_SRCS[@]}"; do
    o="$obj/$(echo "$s" | tr '/' '_').o"
    gcc -std=gnu11 -O3 -fPIC "${arch[@]}" "${DEFS[@]}" "${INCS[@]}" -c "$GG/src/$s" -o "$o" &
    pids+=($!)
  done
  for s in "${CXX_SRCS[@]}"; do
    o="$obj/$(echo "$s" | tr '/' '_').o"
    g++ -std=gnu++17 -O3 -fPIC "${arch[@]}" "${DEFS[@]}" "${INCS[@]}" -c "$GG/src/$s" -o "$o" &
    pids+=($!)
  done
  # the reference hot path itself + our C wrapper.  The reference's CMakeLists.txt gives its first-party sources NO
  # -march flag (only ggml is built "native", GGML_NATIVE): they are compiled for baseline x86-64, i.e. without FMA
  # contraction.  That is visible in the bits of irfft's accumulation (src/istft.cpp:57-62) and of the crossfade
  # (src/test-to-speech.cpp:388-395), so the first-party files are compiled the same way here (Release: -O3 -DNDEBUG).
  FP=(-std=gnu++17 -O3 -DNDEBUG -fPIC)
  g++ "${FP[@]}" "${INCS[@]}" -I"$REF/src" -c "$REF/src/miocodec.cpp" -o "$obj/ref_miocodec.o" & pids+=($!)
  g++ "${FP[@]}" -I"$REF/src" -c "$REF/src/istft.cpp" -o "$obj/ref_istft.o" & pids+=($!)
  # the steps either side of the path: token parser and WAV writer, also unmodified
  g++ "${FP[@]}" -I"$REF/src" -c "$REF/src/token-parser.cpp" -o "$obj/ref_token_parser.o" & pids+=($!)
  g++ "${FP[@]}" -I"$REF/src" -c "$REF/src/wav-writer.cpp" -o "$obj/ref_wav_writer.o" & pids+=($!)
  # the streaming policy (emit_range / maybe_emit live inside TestToSpeech::synthesize_stream_profiled): the reference's
  # orchestrator, unmodified, against a scripted stand-in for libllama (oracle/stub_llama.cpp; token generation is out
  # of scope, the stub only feeds the codes the "LLM" would emit)
  LL="$REF/third_party/llama.cpp/include"
  g++ "${FP[@]}" "${INCS[@]}" -I"$LL" -I"$REF/src" -c "$REF/src/test-to-speech.cpp" -o "$obj/ref_test_to_speech.o" & pids+=($!)
  g++ "${FP[@]}" -I"$REF/src" -c "$REF/src/text-normalize.cpp" -o "$obj/ref_text_normalize.o" & pids+=($!)
  g++ "${FP[@]}" "${INCS[@]}" -I"$LL" -c "$HERE/stub_llama.cpp" -o "$obj/stub_llama.o" & pids+=($!)
  g++ "${FP[@]}" -I"$REF/src" -c "$HERE/ref_capi.cpp" -o "$obj/ref_capi.o" & pids+=($!)
  for p in "${pids[@]}"; do wait "$p"; done
  g++ -shared -o "$OUT/libmio_ref_$tag.so" "$obj"/*.o -lpthread -lm -ldl
  rm -rf "$obj"
  echo "built $OUT/libmio_ref_$tag.so"
}

build_variant v3 -march=x86-64-v3
build_variant v4 -march=x86-64-v4
