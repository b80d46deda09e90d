#!/usr/bin/env bash
# TEST INFRASTRUCTURE ONLY.  Builds the UNMODIFIED reference `droid_backends` CUDA extension
# (src/droid.cpp, src/droid_kernels.cu, src/correlation_kernels.cu, src/altcorr_kernel.cu under
# /root/reference) for sm_100 into oracle/_ref/droid_backends_ref*.so, compiling the sources where they
# lie (nothing is copied into this repo).  The reference's own setup.py is NOT run: it has no sm_100
# target (setup.py:20-27).  The only accommodation is the force-included oracle/ref_compat.h.
# The result cannot execute without a GPU; it travels to the GPU box with gpurun and is the GPU-side
# oracle / "bar to beat" there (tests/golden/make_golden_gpu.py).
set -euo pipefail
HERE="$(cd "$(dirname "$0")" && pwd)"
REF="${REFERENCE_ROOT:-/root/reference}"
OUT="$HERE/_ref"
[ -d "$REF/src" ] || { echo "no reference at $REF; nothing to build"; exit 0; }
mkdir -p "$OUT/obj"
PY=python
TORCH_INC=$($PY -c 'import torch,os;print(os.path.join(os.path.dirname(torch.__file__),"include"))')
TORCH_LIB=$($PY -c 'import torch,os;print(os.path.join(os.path.dirname(torch.__file__),"lib"))')
PY_INC=$($PY -c 'import sysconfig;print(sysconfig.get_paths()["include"])')
SUFFIX=$($PY -c 'import sysconfig;print(sysconfig.get_config_var("EXT_SUFFIX"))')
NAME=droid_backends_ref
INC="-I$TORCH_INC -I$TORCH_INC/torch/csrc/api/include -I$PY_INC -I$REF/thirdparty/eigen -I/usr/local/cuda/include"
DEFS="-DTORCH_EXTENSION_NAME=$NAME -DTORCH_API_INCLUDE_EXTENSION_H -D_GLIBCXX_USE_CXX11_ABI=1"
NVFLAGS="-O3 -std=c++17 -gencode arch=compute_100,code=sm_100 --expt-relaxed-constexpr -Xcompiler -fPIC \
 -D__CUDA_NO_HALF_OPERATORS__ -D__CUDA_NO_HALF_CONVERSIONS__ -D__CUDA_NO_HALF2_OPERATORS__ -w"
pids=()
for f in droid_kernels correlation_kernels altcorr_kernel; do
  if [ ! -f "$OUT/obj/$f.o" ]; then
    nvcc $NVFLAGS $DEFS $INC -include "$HERE/ref_compat.h" -c "$REF/src/$f.cu" -o "$OUT/obj/$f.o" &
    pids+=($!)
  fi
done
if [ ! -f "$OUT/obj/droid.o" ]; then
  g++ -O2 -std=c++17 -fPIC -w $DEFS $INC -include "$HERE/ref_compat.h" -c "$REF/src/droid.cpp" -o "$OUT/obj/droid.o" &
  pids+=($!)
fi
for p in "${pids[@]:-}"; do [ -n "$p" ] && wait "$p"; done
g++ -shared -o "$OUT/$NAME$SUFFIX" "$OUT"/obj/*.o -L"$TORCH_LIB" -L/usr/local/cuda/lib64 \
  -lc10 -ltorch_cpu -ltorch -ltorch_python -lc10_cuda -ltorch_cuda -lcudart -Wl,-rpath,"$TORCH_LIB"
echo "built $OUT/$NAME$SUFFIX"
