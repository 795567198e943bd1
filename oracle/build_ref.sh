#!/usr/bin/env bash
# Build the UNMODIFIED reference (keiserlab/rad's vendored usearch fork 2.17.2) python binding
# from the sources where they lie under /root/reference, outputs ONLY into oracle/_ref/ (git-ignored).
# Test infrastructure / CPU baseline only -- never imported by the product path (rad_b200/).
# Flags follow SURVEY.md §8c: SimSIMD header-only, OpenMP executor, no -ffast-math (IEEE f64 divide).
set -euo pipefail
HERE="$(cd "$(dirname "${BASH_SOURCE[0]}")" && pwd)"
REF="${RAD_REFERENCE_ROOT:-/root/reference}/usearch"
OUT="$HERE/_ref/usearch"
PY="${PYTHON:-python3}"
if [ ! -d "$REF/include/usearch" ]; then
  echo "build_ref: reference sources not present at $REF (expected on the GPU box); skipping" >&2
  exit 0
fi
EXT="$($PY -c 'import sysconfig;print(sysconfig.get_config_var("EXT_SUFFIX"))')"
mkdir -p "$OUT"
# RAD itself is pure Python above the index: its UNMODIFIED modules are staged next to the binding (git-ignored, like
# the rest of oracle/_ref) so that the GPU box -- which has no /root/reference -- can run the reference's traverser
# and services over a GPU-built index (tests/test_rad_gpu_e2e.py)
RAD_SRC="${RAD_REFERENCE_ROOT:-/root/reference}/rad"
if [ -d "$RAD_SRC" ]; then
  mkdir -p "$HERE/_ref/rad"
  cp "$RAD_SRC"/*.py "$HERE/_ref/rad"/
fi
if [ -f "$OUT/compiled$EXT" ] && [ "$OUT/compiled$EXT" -nt "$REF/python/lib.cpp" ] && [ -z "${FORCE:-}" ]; then
  echo "build_ref: $OUT/compiled$EXT up to date"; exit 0
fi
# the pure-python half of the binding is installed (copied) next to the compiled module, like pip would
cp "$REF"/python/usearch/*.py "$OUT"/
MARCH="${RAD_REF_MARCH:--march=x86-64-v3 -mtune=native}"
g++ -O3 $MARCH -std=c++17 -shared -fPIC -w "$REF/python/lib.cpp" \
  -I"$REF/include" -I"$REF/fp16/include" -I"$REF/simsimd/include" -I"$REF/stringzilla/include" \
  $($PY -m pybind11 --includes) \
  -DUSEARCH_USE_SIMSIMD=1 -DSIMSIMD_DYNAMIC_DISPATCH=0 -DUSEARCH_USE_FP16LIB=1 -DUSEARCH_USE_OPENMP=1 -fopenmp \
  -o "$OUT/compiled$EXT"
echo "build_ref: built $OUT/compiled$EXT"
