#!/usr/bin/env bash
# TEST INFRASTRUCTURE ONLY.  Builds the reference's vendored lietorch extension (`lietorch_backends`,
# /root/reference/thirdparty/lietorch/lietorch/src/{lietorch.cpp,lietorch_cpu.cpp,lietorch_gpu.cu}) for the host CPU
# and sm_100 into oracle/_ref/lietorch_backends*.so, compiling the sources where they lie.  The reference's Python
# callers of this path (droid_slam/depth_video.py, factor_graph.py, geom/*.py) need it; with it the UNMODIFIED Python
# reference runs against either back end (tests/test_reference_callers.py, bench.py cpu_baseline kind "reference").
# lietorch's own setup.py is NOT run (targets sm_60..75 only).  Same single accommodation as build_ref.sh: the
# force-included oracle/ref_compat.h (the DeprecatedTypeProperties overload torch >= 2 dropped).
set -euo pipefail
HERE="$(cd "$(dirname "$0")" && pwd)"
REF="${REFERENCE_ROOT:-/root/reference}"
LT="$REF/thirdparty/lietorch"
OUT="$HERE/_ref"
[ -d "$LT/lietorch/src" ] || { echo "no lietorch at $LT; nothing to build"; exit 0; }
mkdir -p "$OUT/obj_lt"
PY=python
TORCH_INC=$($PY -c 'import torch,os;print(os.path.join(os.path.dirname(torch.__file__),"include"))')
TORCH_LIB=$($PY -c 'import torch,os;print(os.path.join(os.path.dirname(torch.__file__),"lib"))')
PY_INC=$($PY -c 'import sysconfig;print(sysconfig.get_paths()["include"])')
SUFFIX=$($PY -c 'import sysconfig;print(sysconfig.get_config_var("EXT_SUFFIX"))')
NAME=lietorch_backends
INC="-I$TORCH_INC -I$TORCH_INC/torch/csrc/api/include -I$PY_INC -I$LT/lietorch/include -I$LT/eigen -I$REF/thirdparty/eigen -I/usr/local/cuda/include"
DEFS="-DTORCH_EXTENSION_NAME=$NAME -DTORCH_API_INCLUDE_EXTENSION_H -D_GLIBCXX_USE_CXX11_ABI=1"
NVFLAGS="-O2 -std=c++17 -gencode arch=compute_100,code=sm_100 --expt-relaxed-constexpr -Xcompiler -fPIC \
 -D__CUDA_NO_HALF_OPERATORS__ -D__CUDA_NO_HALF_CONVERSIONS__ -D__CUDA_NO_HALF2_OPERATORS__ -w"
pids=()
if [ ! -f "$OUT/obj_lt/lietorch_gpu.o" ]; then
  nvcc $NVFLAGS $DEFS $INC -include "$HERE/ref_compat.h" -c "$LT/lietorch/src/lietorch_gpu.cu" -o "$OUT/obj_lt/lietorch_gpu.o" &
  pids+=($!)
fi
for f in lietorch lietorch_cpu; do
  if [ ! -f "$OUT/obj_lt/$f.o" ]; then
    g++ -O2 -std=c++17 -fPIC -fopenmp -w $DEFS $INC -include "$HERE/ref_compat.h" -c "$LT/lietorch/src/$f.cpp" -o "$OUT/obj_lt/$f.o" &
    pids+=($!)
  fi
done
for p in "${pids[@]:-}"; do [ -n "$p" ] && wait "$p"; done
g++ -shared -o "$OUT/$NAME$SUFFIX" "$OUT"/obj_lt/*.o -L"$TORCH_LIB" -L/usr/local/cuda/lib64 \
  -lc10 -ltorch_cpu -ltorch -ltorch_python -lc10_cuda -ltorch_cuda -lcudart -lgomp -Wl,-rpath,"$TORCH_LIB"
echo "built $OUT/$NAME$SUFFIX"
