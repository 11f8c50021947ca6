#!/usr/bin/env bash
# TEST INFRASTRUCTURE ONLY -- builds the REFERENCE's own CUDA kernels for the
# paged-decode path, unmodified, for sm_100a into oracle/_ref/libvllm_ref_cuda.so
# (git-ignored, travels to the GPU box).  Oracle tier O1.
#
# Sources are compiled where they lie under /root/reference (never copied):
#   csrc/attention/paged_attention_v1.cu   (all template instances, ~7 min)
#   csrc/attention/paged_attention_v2.cu
#   csrc/cache_kernels.cu
#   csrc/attention/merge_attn_states.cu
# plus oracle/ref_cuda_binding.cpp (op registrations, private namespace).
# Flags follow the reference's CMakeLists.txt (-DENABLE_FP8,
# --expt-relaxed-constexpr, c++17).  The reference's cmake is NOT run.
set -euo pipefail
HERE="$(cd "$(dirname "${BASH_SOURCE[0]}")" && pwd)"
REF="${REFERENCE_ROOT:-/root/reference}"
OUT="$HERE/_ref"
PY="${PYTHON:-python}"
NVCC="${NVCC:-/usr/local/cuda/bin/nvcc}"

if [ ! -d "$REF/csrc/attention" ]; then
  echo "build_ref_cuda.sh: $REF/csrc/attention not present; keeping prebuilt $OUT" >&2
  exit 0
fi
mkdir -p "$OUT"
TORCH_DIR="$($PY -c 'import torch,os;print(os.path.dirname(torch.__file__))')"
PY_INC="$($PY -c 'import sysconfig;print(sysconfig.get_paths()["include"])')"
ABI="$($PY -c 'import torch;print(int(torch._C._GLIBCXX_USE_CXX11_ABI))')"

SO="$OUT/libvllm_ref_cuda.so"
STAMP="$OUT/.stamp_cuda"
KEY="$(cat "$REF/csrc/attention/"*.cu "$REF/csrc/attention/"*.cuh "$REF/csrc/attention/"*.h \
          "$REF/csrc/cache_kernels.cu" "$REF/csrc/quantization/fp8/nvidia/quant_utils.cuh" \
          "$HERE/ref_cuda_binding.cpp" "$HERE/build_ref_cuda.sh" 2>/dev/null | sha1sum | cut -d' ' -f1)-$TORCH_DIR"
if [ -f "$SO" ] && [ -f "$STAMP" ] && [ "$(cat "$STAMP")" = "$KEY" ]; then
  echo "build_ref_cuda.sh: $SO up to date"; exit 0
fi

INC=(-I"$REF/csrc" -I"$TORCH_DIR/include" -I"$TORCH_DIR/include/torch/csrc/api/include" -I"$PY_INC")
NVFLAGS=(-std=c++17 -O2 -gencode arch=compute_100a,code=sm_100a -DENABLE_FP8
         --expt-relaxed-constexpr --expt-extended-lambda -Xcompiler -fPIC -w
         -D_GLIBCXX_USE_CXX11_ABI="$ABI" -DNDEBUG
         -U__CUDA_NO_HALF_OPERATORS__ -U__CUDA_NO_HALF_CONVERSIONS__
         -U__CUDA_NO_BFLOAT16_CONVERSIONS__ -U__CUDA_NO_HALF2_OPERATORS__ --threads 4)
TMP="$(mktemp -d)"
"$NVCC" "${NVFLAGS[@]}" "${INC[@]}" -c "$REF/csrc/attention/paged_attention_v1.cu" -o "$TMP/pa_v1.o" &
"$NVCC" "${NVFLAGS[@]}" "${INC[@]}" -c "$REF/csrc/attention/paged_attention_v2.cu" -o "$TMP/pa_v2.o" &
"$NVCC" "${NVFLAGS[@]}" "${INC[@]}" -c "$REF/csrc/cache_kernels.cu" -o "$TMP/cache.o" &
"$NVCC" "${NVFLAGS[@]}" "${INC[@]}" -c "$REF/csrc/attention/merge_attn_states.cu" -o "$TMP/merge.o" &
g++ -std=c++17 -O2 -fPIC -w -D_GLIBCXX_USE_CXX11_ABI="$ABI" "${INC[@]}" \
    -c "$HERE/ref_cuda_binding.cpp" -o "$TMP/binding.o" &
wait
g++ -shared "$TMP/pa_v1.o" "$TMP/pa_v2.o" "$TMP/cache.o" "$TMP/merge.o" "$TMP/binding.o" \
    -L"$TORCH_DIR/lib" -Wl,-rpath,"$TORCH_DIR/lib" -L/usr/local/cuda/lib64 \
    -ltorch -ltorch_cpu -ltorch_cuda -lc10 -lc10_cuda -lcudart -o "$SO"
rm -rf "$TMP"
echo "$KEY" > "$STAMP"
echo "build_ref_cuda.sh: built $SO"
