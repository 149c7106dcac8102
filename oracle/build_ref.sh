#!/usr/bin/env bash
# TEST INFRASTRUCTURE ONLY.
# Builds the reference's own CUDA extension (`vren`, /root/reference/models/csrc) for
# sm_100a as the python module `vren_ref`, from the sources WHERE THEY LIE (nothing is
# copied into this repo), into oracle/_ref/ (git-ignored, travels to the GPU box).
# The only accommodation is oracle/ref_shim.h, force-included so that the reference's
# AT_DISPATCH_*(x.type(), ...) call sites compile against torch 2.11.
# Used by tests/ (GPU-side oracle pinning the CPU restatement + integer parity) and by
# bench.py's optional "gpu_anchor" leg. Never imported by the product package.
set -euo pipefail
HERE="$(cd "$(dirname "$0")" && pwd)"
REF=${REF:-/root/reference/models/csrc}
OUT="$HERE/_ref"
mkdir -p "$OUT/obj"
if [ ! -d "$REF" ]; then echo "reference not present at $REF; keeping prebuilt $OUT" ; exit 0; fi
PY=${PYTHON:-python}
read -r TORCH_INC1 TORCH_INC2 TORCH_LIB PY_INC EXT <<<"$($PY - <<'EOF'
import sysconfig, os, torch
r = os.path.dirname(torch.__file__)
print(r + "/include", r + "/include/torch/csrc/api/include", r + "/lib",
      sysconfig.get_paths()["include"], sysconfig.get_config_var("EXT_SUFFIX"))
EOF
)"
TARGET="$OUT/vren_ref$EXT"
if [ -f "$TARGET" ] && [ "$TARGET" -nt "$REF/raymarching.cu" ] && [ "$TARGET" -nt "$HERE/ref_shim.h" ]; then
  echo "up to date: $TARGET"; exit 0
fi
COMMON=(-I"$REF/include" -I"$TORCH_INC1" -I"$TORCH_INC2" -I"$PY_INC" -I/usr/local/cuda/include
        -include "$HERE/ref_shim.h" -DTORCH_EXTENSION_NAME=vren_ref -DTORCH_API_INCLUDE_EXTENSION_H
        -D_GLIBCXX_USE_CXX11_ABI=1 -std=c++17)
NVFLAGS=(-O2 -gencode arch=compute_100a,code=sm_100a --expt-relaxed-constexpr
         -D__CUDA_NO_HALF_OPERATORS__ -D__CUDA_NO_HALF_CONVERSIONS__
         -D__CUDA_NO_HALF2_OPERATORS__ -D__CUDA_NO_BFLOAT16_CONVERSIONS__
         -Xcompiler -fPIC)
pids=()
for f in intersection raymarching volumerendering losses; do
  nvcc "${COMMON[@]}" "${NVFLAGS[@]}" -c "$REF/$f.cu" -o "$OUT/obj/$f.o" &
  pids+=($!)
done
g++ "${COMMON[@]}" -O2 -fPIC -c "$REF/binding.cpp" -o "$OUT/obj/binding.o" &
pids+=($!)
for p in "${pids[@]}"; do wait "$p"; done
g++ -shared -o "$TARGET" "$OUT"/obj/*.o -L"$TORCH_LIB" -L/usr/local/cuda/lib64 \
    -lc10 -ltorch_cpu -ltorch -ltorch_python -lc10_cuda -ltorch_cuda -lcudart \
    -Wl,-rpath,"$TORCH_LIB"
echo "built $TARGET"
