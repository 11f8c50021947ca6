#!/bin/bash
# Build copies of the package whose decode kernel lacks one ingredient of its main loop
# (B200PA_ABLATE in csrc/paged_attention_stream.cu; head size 128 / bf16 query instantiations
# only) into _ab/<label>/vllm_main_modify_b200, for tools/powerprobe.py and tools/kbench.py:
#   B200PA_PKG_ROOT=_ab/no_cvt B200PA_AB_LABEL=no_cvt python tools/powerprobe.py --tune 100
# (the ablation switches sit in the per-warp stream kernel: --tune 100 + i selects it with its
# launch shape i; without it head size 128 with several kv heads takes the whole-block kernel)
# Results of such a build are WRONG by construction; nothing but timing tools may load it.
#   tools/ablate.sh no_cvt:1 no_mma:2 no_ldmatrix:4 no_exp:8 only_copy:15
set -eu
cd "$(dirname "$0")/.."
PKG=vllm_main_modify_b200
python -m $PKG.build >/dev/null
for spec in "$@"; do
  label=${spec%%:*}; mask=${spec##*:}
  dst=_ab/$label/$PKG
  mkdir -p $dst/_build
  cp $PKG/*.py $dst/
  cp $PKG/_build/*.o $dst/_build/
  /usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo \
    -Xcompiler -fPIC -ccbin /usr/bin/g++ --expt-relaxed-constexpr -Wno-deprecated-gpu-targets \
    -DB200PA_TESTING=1 -DSTREAM_PART=0 -DB200PA_ABLATE=$mask \
    -c $PKG/csrc/paged_attention_stream.cu -o $dst/_build/paged_attention_stream.o
  /usr/local/cuda/bin/nvcc -shared -o $dst/libb200pa.so $dst/_build/*.o -cudart static \
    -ccbin /usr/bin/g++ -gencode arch=compute_100a,code=sm_100a
  cp $PKG/_b200pa_torch.so $dst/     # rpath $ORIGIN: binds to the libb200pa.so beside it
  rm -rf $dst/_build
  echo "built $dst (ablate mask $mask)"
done
