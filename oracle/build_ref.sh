#!/bin/bash
# TEST INFRASTRUCTURE.  Compiles the REFERENCE's own rasterization kernels (vendored gsplat 1.5.3)
# from the sources WHERE THEY LIE under /root/reference, for sm_100a, with the reference's own
# --use_fast_math (cpp/third_party/CMakeLists.txt:42), into oracle/_ref/gsplat_ref.so (git-ignored,
# travels to the GPU box).  Needs a GPU to RUN, so it is a GPU-side second oracle and the "existing
# kernel to beat" (SURVEY F5, 8d(4)); the CPU oracle stays the portable checker.
set -e
HERE="$(cd "$(dirname "$0")" && pwd)"
GS=/root/reference/cpp/third_party/gsplat/gsplat/cuda
[ -d "$GS/csrc" ] || { echo "reference not present: nothing to build"; exit 0; }
OUT="$HERE/_ref"; mkdir -p "$OUT"
SO="$OUT/gsplat_ref.so"
if [ -f "$SO" ] && [ "$SO" -nt "$HERE/ref_gsplat/binding.cpp" ] && [ "$SO" -nt "$0" ]; then exit 0; fi
PY=${PYTHON:-python}
TINC=$($PY -c "from torch.utils.cpp_extension import include_paths; print(' '.join('-I'+p for p in include_paths()))")
TLIB=$($PY -c "import torch,os; print(os.path.join(os.path.dirname(torch.__file__),'lib'))")
PINC=$($PY -c "import sysconfig; print('-I'+sysconfig.get_paths()['include'])")
ABI=$($PY -c "import torch; print(int(torch._C._GLIBCXX_USE_CXX11_ABI))")
NVCC=${NVCC:-/usr/local/cuda/bin/nvcc}
FLAGS="-O3 --use_fast_math -std=c++17 -gencode arch=compute_100a,code=sm_100a --expt-relaxed-constexpr \
 -ccbin /usr/bin/g++ -Xcompiler -fPIC -I$GS/include -I$GS/csrc -I$GS/csrc/third_party/glm $TINC $PINC \
 -D_GLIBCXX_USE_CXX11_ABI=$ABI -DTORCH_EXTENSION_NAME=gsplat_ref -DTORCH_API_INCLUDE_EXTENSION_H -w"
# the two rasterize kernel files take ~70 s each (19 channel-count instantiations): keep their objects
[ -f "$OUT/rast_fwd.o" ] || $NVCC $FLAGS -c "$GS/csrc/RasterizeToPixels3DGSFwd.cu" -o "$OUT/rast_fwd.o" &
[ -f "$OUT/rast_bwd.o" ] || $NVCC $FLAGS -c "$GS/csrc/RasterizeToPixels3DGSBwd.cu" -o "$OUT/rast_bwd.o" &
# projection (with the opacity-aware extent, ProjectionEWA3DGSFused.cu:164-184) and the tile intersection:
# the kernels, their CUB sort and the reference's own host wrapper intersect_tile / intersect_offset (Intersect.cpp)
[ -f "$OUT/proj.o" ] || $NVCC $FLAGS -c "$GS/csrc/ProjectionEWA3DGSFused.cu" -o "$OUT/proj.o" &
[ -f "$OUT/isect_tile.o" ] || $NVCC $FLAGS -c "$GS/csrc/IntersectTile.cu" -o "$OUT/isect_tile.o" &
[ -f "$OUT/isect_host.o" ] || $NVCC $FLAGS -x cu -c "$GS/csrc/Intersect.cpp" -o "$OUT/isect_host.o" &
$NVCC $FLAGS -x cu -c "$HERE/ref_gsplat/binding.cpp" -o "$OUT/binding.o" &
wait
/usr/bin/g++ -shared -o "$SO" "$OUT/binding.o" "$OUT/rast_fwd.o" "$OUT/rast_bwd.o" "$OUT/proj.o" "$OUT/isect_tile.o" \
  "$OUT/isect_host.o" -L"$TLIB" -Wl,-rpath,"$TLIB" \
  -ltorch -ltorch_cpu -ltorch_cuda -lc10 -lc10_cuda -ltorch_python -L/usr/local/cuda/lib64 -lcudart
rm -f "$OUT/binding.o"
echo "built $SO"
