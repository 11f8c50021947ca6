#!/usr/bin/env bash
# TEST INFRASTRUCTURE ONLY -- builds the REFERENCE's own CPU kernels for the
# paged-decode path into oracle/_ref/ (git-ignored, travels to the GPU box).
#
# Sources are compiled where they lie under /root/reference (never copied):
#   csrc/cpu/attention.cpp   paged_attention_v1 / paged_attention_v2  (CPU)
#   csrc/cpu/cache.cpp       reshape_and_cache                        (CPU)
# plus oracle/ref_binding.cpp (our 3 op registrations, private namespace).
#
# Flags follow the reference's cmake/cpu_extension.cmake:28-40 (-mf16c,
# -fopenmp, -DVLLM_CPU_EXTENSION) and :107-135 (ISA flags).  The reference's
# cmake picks the ISA from /proc/cpuinfo of the build host; the GPU box may
# have a different host CPU, so three variants are built and the loader
# (oracle/ref_cpu.py) picks the best one the running host supports:
#   avx512bf16 : -mavx512f -mavx512vl -mavx512bw -mavx512dq -mavx512bf16 -mavx512vnni
#   avx512     : -mavx512f -mavx512vl -mavx512bw -mavx512dq
#   avx2       : -mavx2 -mfma
# The reference's own build system (cmake + oneDNN etc.) is NOT run.
set -euo pipefail
HERE="$(cd "$(dirname "${BASH_SOURCE[0]}")" && pwd)"
REF="${REFERENCE_ROOT:-/root/reference}"
OUT="$HERE/_ref"
PY="${PYTHON:-python}"

if [ ! -d "$REF/csrc/cpu" ]; then
  echo "build_ref.sh: $REF/csrc/cpu not present; keeping prebuilt $OUT" >&2
  exit 0
fi
mkdir -p "$OUT"

TORCH_DIR="$($PY -c 'import torch,os;print(os.path.dirname(torch.__file__))')"
PY_INC="$($PY -c 'import sysconfig;print(sysconfig.get_paths()["include"])')"
ABI="$($PY -c 'import torch;print(int(torch._C._GLIBCXX_USE_CXX11_ABI))')"

COMMON=(-std=c++17 -O3 -fPIC -mf16c -fopenmp -DVLLM_CPU_EXTENSION
        -D_GLIBCXX_USE_CXX11_ABI="$ABI" -DNDEBUG -w
        -I"$REF/csrc" -I"$REF/csrc/cpu"
        -I"$TORCH_DIR/include" -I"$TORCH_DIR/include/torch/csrc/api/include"
        -I"$PY_INC")
LINK=(-shared -fopenmp -L"$TORCH_DIR/lib" -Wl,-rpath,"$TORCH_DIR/lib"
      -ltorch -ltorch_cpu -lc10)

build_variant() {
  local name="$1"; shift
  local isa=("$@")
  local so="$OUT/libvllm_ref_cpu_${name}.so"
  local stamp="$OUT/.stamp_${name}"
  local key
  key="$(cat "$REF/csrc/cpu/attention.cpp" "$REF/csrc/cpu/cache.cpp" \
            "$REF/csrc/cpu/cpu_types_x86.hpp" "$HERE/ref_binding.cpp" \
            "$HERE/build_ref.sh" | sha1sum | cut -d' ' -f1)-$TORCH_DIR"
  if [ -f "$so" ] && [ -f "$stamp" ] && [ "$(cat "$stamp")" = "$key" ]; then
    echo "build_ref.sh: $so up to date"
    return 0
  fi
  local tmp; tmp="$(mktemp -d)"
  g++ "${COMMON[@]}" "${isa[@]}" -c "$REF/csrc/cpu/attention.cpp" -o "$tmp/attention.o" &
  g++ "${COMMON[@]}" "${isa[@]}" -c "$REF/csrc/cpu/cache.cpp" -o "$tmp/cache.o" &
  g++ "${COMMON[@]}" "${isa[@]}" -c "$HERE/ref_binding.cpp" -o "$tmp/binding.o" &
  wait
  g++ "$tmp/attention.o" "$tmp/cache.o" "$tmp/binding.o" "${LINK[@]}" -o "$so"
  rm -rf "$tmp"
  echo "$key" > "$stamp"
  echo "build_ref.sh: built $so"
}

build_variant avx512bf16 -mavx512f -mavx512vl -mavx512bw -mavx512dq -mavx512bf16 -mavx512vnni &
build_variant avx512 -mavx512f -mavx512vl -mavx512bw -mavx512dq &
build_variant avx2 -mavx2 -mfma &
wait
ls -la "$OUT"
