#!/usr/bin/env bash
# Compile the UNMODIFIED reference CUDA extensions from the sources where they lie
# under /root/reference into oracle/_ref/ (git-ignored, travels to the GPU box).
#
#   oracle/_ref/diff_gaussian_rasterization/{_C.so, __init__.py = the reference's own Python wrapper}
#   oracle/_ref/simple_knn/{_C.so, __init__.py (empty package marker)}
#   oracle/_ref/pyref/{tools/*.py, scene/gaussian_model.py, scene/appearance_network.py, gaussian_renderer/__init__.py}
#                      = the reference's own Python modules of the hot path (render(), GaussianModel, the loss and
#                        normal utilities), staged byte for byte so that they travel to the GPU box, where
#                        /root/reference does not exist.  scene/__init__.py is an EMPTY generated package marker (the
#                        reference's own pulls in the dataset readers), and pyref/_stubs/ holds generated placeholder
#                        modules for two third-party imports this image lacks (plyfile, pytorch3d.ops) — neither is
#                        touched by the functions on the hot path.
#
# Everything under oracle/_ref/ is git-ignored: no reference source enters the repository's history.  No reference
# rasterizer source is patched: the one compile problem with gcc 13 (rasterizer_impl.h uses std::uintptr_t without
# <cstdint>) is solved with the compiler flag `-include cstdint`; the GLM qualifier shim is described below.
#
# This is test infrastructure (the parity oracle + the `--impl reference` bench arm), never the product path.
# Variant build: NUM_DIST=1 NUM_DEPTH2=2 bash oracle/build_ref.sh builds the distortion variant of the rasterizer
# (DGR/cuda_rasterizer/config.h:17-18) into oracle/_ref/diff_gaussian_rasterization_d${NUM_DIST:-0}_${NUM_DEPTH2:-0}/.
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

stage_python () {
  # the reference's own Python half of the path, copied verbatim (git-ignored destination)
  cp "$DGR/diff_gaussian_rasterization/__init__.py" "$OUT/diff_gaussian_rasterization/__init__.py"
  : > "$OUT/simple_knn/__init__.py"
  local PYREF="$OUT/pyref"
  rm -rf "$PYREF" && mkdir -p "$PYREF/tools" "$PYREF/scene" "$PYREF/gaussian_renderer" "$PYREF/_stubs/pytorch3d"
  cp "$REF"/tools/*.py "$PYREF/tools/"
  cp "$REF/scene/gaussian_model.py" "$REF/scene/appearance_network.py" "$REF/scene/cameras.py" "$PYREF/scene/"
  : > "$PYREF/scene/__init__.py"
  cp "$REF/gaussian_renderer/__init__.py" "$PYREF/gaussian_renderer/__init__.py"
  cat > "$PYREF/_stubs/plyfile.py" <<'STUB'
"""Generated placeholder (oracle/build_ref.sh): `plyfile` is not installed in this image.  scene/gaussian_model.py imports
PlyData / PlyElement for save_ply / load_ply only; nothing on the rasterizer hot path calls them."""


class _Missing:
    def __init__(self, *a, **k):
        raise ImportError("plyfile is not available in this image")

    @classmethod
    def read(cls, *a, **k):
        raise ImportError("plyfile is not available in this image")

    @classmethod
    def describe(cls, *a, **k):
        raise ImportError("plyfile is not available in this image")


PlyData = PlyElement = _Missing
STUB
  : > "$PYREF/_stubs/pytorch3d/__init__.py"
  cat > "$PYREF/_stubs/pytorch3d/ops.py" <<'STUB'
"""Generated placeholder (oracle/build_ref.sh): `pytorch3d` is not installed in this image.  tools/denoise_pcd.py imports
ball_query / knn_points for point-cloud outlier removal only; nothing on the rasterizer hot path calls them."""


def ball_query(*a, **k):
    raise ImportError("pytorch3d is not available in this image")


def knn_points(*a, **k):
    raise ImportError("pytorch3d is not available in this image")
STUB
}
stage_python
if [ "${STAGE_ONLY:-0}" = "1" ]; then echo "staged: $OUT/pyref"; rm -rf "$OUT/obj"; exit 0; fi

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
if [ -n "${NUM_DIST:-}" ] || [ -n "${NUM_DEPTH2:-}" ]; then
  # distortion variant (README.md:150-155): config.h gives NUM_DIST / NUM_DEPTH2 no #ifndef guard, so the build compiles
  # against a temporary copy of cuda_rasterizer/ in which those two #define lines carry the requested values
  VAR="$OUT/obj/dgr_variant"
  rm -rf "$VAR" && mkdir -p "$VAR" "$OUT/diff_gaussian_rasterization_d${NUM_DIST:-0}_${NUM_DEPTH2:-0}"
  cp -r "$DGR/cuda_rasterizer" "$DGR/rasterize_points.cu" "$DGR/rasterize_points.h" "$DGR/ext.cpp" "$VAR/"
  sed -i "s/^#define NUM_DIST .*/#define NUM_DIST ${NUM_DIST:-0}/; s/^#define NUM_DEPTH2 .*/#define NUM_DEPTH2 ${NUM_DEPTH2:-0}/" "$VAR/cuda_rasterizer/config.h"
  grep -E "^#define NUM_(DIST|DEPTH2) " "$VAR/cuda_rasterizer/config.h" >&2
  build_one diff_gaussian_rasterization_d${NUM_DIST:-0}_${NUM_DEPTH2:-0} "$VAR" "$ARCH_DGR" "-I$GLMFIX" \
    cuda_rasterizer/rasterizer_impl.cu cuda_rasterizer/forward.cu cuda_rasterizer/backward.cu rasterize_points.cu ext.cpp
  rm -rf "$OUT/obj"
  cp "$DGR/diff_gaussian_rasterization/__init__.py" "$OUT/diff_gaussian_rasterization_d${NUM_DIST:-0}_${NUM_DEPTH2:-0}/__init__.py"
  echo "built: $OUT/diff_gaussian_rasterization_d${NUM_DIST:-0}_${NUM_DEPTH2:-0}/_C.so"
  exit 0
fi
build_one diff_gaussian_rasterization "$DGR" "$ARCH_DGR" "-I$GLMFIX" \
  cuda_rasterizer/rasterizer_impl.cu cuda_rasterizer/forward.cu cuda_rasterizer/backward.cu rasterize_points.cu ext.cpp
build_one simple_knn "$KNN" "$ARCH_KNN" "" spatial.cu simple_knn.cu ext.cpp
rm -rf "$OUT/obj"
echo "built: $(ls "$OUT"/*/_C.so)"
