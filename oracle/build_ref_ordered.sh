#!/bin/sh
# Like build_ref.sh, but with oracle/ref_order_shim.h force-included: the reference core with the Clipper's
# pointer-ordered sets ordered by object id instead (see the shim's header comment). Output: oracle/_ref_ordered/
# (git-ignored). Only needed to regenerate tests/golden/clipper_spread.npz in the build container.
set -e
HERE="$(cd "$(dirname "$0")" && pwd)"
R=${REFERENCE_ROOT:-/root/reference}/packages/mettagrid/cpp
[ -d "$R" ] || { echo "reference tree not present: $R" >&2; exit 3; }
mkdir -p "$HERE/_ref_ordered"
OUT="$HERE/_ref_ordered/mettagrid_c$(python3-config --extension-suffix)"
g++ -std=c++20 -O2 -DNDEBUG -fPIC -shared -fvisibility=hidden -include "$HERE/ref_order_shim.h" \
    -DNPY_NO_DEPRECATED_API=NPY_1_7_API_VERSION $(python3 -m pybind11 --includes) \
    -I"$R/include/mettagrid" -I"$R" "$R/bindings/mettagrid_c.cpp" $(find "$R/src" -name '*.cpp') "$HERE/ref_order_shim.cpp" -o "$OUT"
echo "$OUT"
