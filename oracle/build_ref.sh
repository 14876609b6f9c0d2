#!/usr/bin/env bash
# Compile the UNMODIFIED reference CUDA extensions from the sources where they lie
# under /root/reference into oracle/_ref/ (git-ignored, travels to the GPU box).
#
#   oracle/_ref/diff_gaussian_rasterization/{__init__.py -> generated loader, _C.so}
#   oracle/_ref/simple_knn/_C.so
#
# No reference source file is copied or patched: the one compile problem with gcc 13
# (rasterizer_impl.h uses std::uintptr_t without <cstdint>) is solved with the
# compiler flag `-include cstdint`.  The Python half of the reference binding
# (diff_gaussian_rasterization/__init__.py) is NOT copied either: at test time the
# A/B harness (tests/ref_ab.py) drives `_C` directly with the argument tuples the
# reference wrapper builds (DGR/diff_gaussian_rasterization/__init__.py:186-220, 401-427).
#
# This is test infrastructure (the parity oracle + the `--impl reference` bench arm),
# never the product path.
set -euo pipefail
REF=${REF:-/root/reference}
HERE="$(cd "$(dirname "$0")" && pwd)"
OUT="$HERE/_ref"
DGR="$REF/submodules/diff-gaussian-rasterization"
KNN="$REF/submodules/simple-knn"
if [ ! -d "$DGR" ]; then echo "reference not present at $REF; keeping prebuilt oracle/_ref" >&2; exit 0; fi
PY=${PYTHON:-python}
TORCH_INC=$($PY - <<'PY'
import sysconfig, os, torch
from torch.utils import cpp_extension as c
inc = c.include_paths() + [sysconfig.get_paths()["include"]]
print(" ".join("-I" + i for i in inc))
PY
)
TORCH_LIB=$($PY -c "import torch,os;print(os.path.join(os.path.dirname(torch.__file__),'lib'))")
# Toolchain note (DESIGN.md "Reference build"): the GLM 1.0.2 vendored by the reference declares the TMin/TMax
# functors in glm/detail/func_common.inl WITHOUT __host__ __device__ qualifiers.  glm::max(vec3, float) - the clamp at the
# end of the reference's computeColorFromSH (forward.cu:72) - therefore calls a host-only function from device code;
# nvcc 12.9 (the only toolchain in this image) treats that as unreachable and silently deletes the three `rgb[...]`
# stores (and for compute_100 also the `clamped` stores), i.e. the reference renders black Gaussians.  The build
# compiles against a TEMPORARY copy of the GLM headers (inside the git-ignored build directory, deleted afterwards) in
# which those two functors get GLM_FUNC_QUALIFIER.  No file of the reference rasterizer itself is changed.
ARCH_DGR="-gencode arch=compute_100,code=sm_100"
ARCH_KNN="-gencode arch=compute_100,code=sm_100"
COMMON="-O3 -std=c++17 --expt-relaxed-constexpr -Xcompiler -fPIC -include cstdint \
 -DTORCH_API_INCLUDE_EXTENSION_H -D_GLIBCXX_USE_CXX11_ABI=1 -w $TORCH_INC"
mkdir -p "$OUT/diff_gaussian_rasterization" "$OUT/simple_knn" "$OUT/obj"

build_one () { # name srcdir arch extra sources...
  local name=$1 dir=$2 ARCH=$3 extra=$4; shift 4
  local objs=()
  for s in "$@"; do
    local o="$OUT/obj/${name}_$(basename "${s%.*}").o"
    if [ ! -f "$o" ] || [ "$dir/$s" -nt "$o" ]; then
      nvcc $ARCH $COMMON $extra -DTORCH_EXTENSION_NAME=_C -x cu -c "$dir/$s" -o "$o" &
    fi
    objs+=("$o")
  done
  wait
  nvcc -shared -o "$OUT/$name/_C.so" "${objs[@]}" -L"$TORCH_LIB" -lc10 -ltorch -ltorch_cpu -ltorch_python -lc10_cuda -ltorch_cuda \
       -Xlinker -rpath -Xlinker "$TORCH_LIB"
}

GLMFIX="$OUT/obj/glm_qualified"
rm -rf "$GLMFIX" && cp -r "$DGR/third_party/glm" "$GLMFIX"
sed -i 's/^\(\s*\)T operator()(const T& a, const T& b) { return \(min\|max\)(a, b); }/\1GLM_FUNC_QUALIFIER T operator()(const T\& a, const T\& b) { return \2(a, b); }/' "$GLMFIX/glm/detail/func_common.inl"
grep -c "GLM_FUNC_QUALIFIER T operator()" "$GLMFIX/glm/detail/func_common.inl" | grep -q 2 || { echo "GLM qualifier shim did not apply" >&2; exit 1; }
build_one diff_gaussian_rasterization "$DGR" "$ARCH_DGR" "-I$GLMFIX" \
  cuda_rasterizer/rasterizer_impl.cu cuda_rasterizer/forward.cu cuda_rasterizer/backward.cu rasterize_points.cu ext.cpp
build_one simple_knn "$KNN" "$ARCH_KNN" "" spatial.cu simple_knn.cu ext.cpp
rm -rf "$OUT/obj"
echo "built: $(ls "$OUT"/*/_C.so)"
