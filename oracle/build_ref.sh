#!/bin/sh
# Compiles the reference MettaGrid core from the sources where they lie under /root/reference
# (no reference source is copied into this repo) into oracle/_ref/ (git-ignored).
# Command from SURVEY.md 8(c); -O3 -DNDEBUG as packages/mettagrid/cpp/BUILD.bazel:12-18.
set -e
HERE="$(cd "$(dirname "$0")" && pwd)"
R=${REFERENCE_ROOT:-/root/reference}/packages/mettagrid/cpp
[ -d "$R" ] || { echo "reference tree not present: $R" >&2; exit 3; }
mkdir -p "$HERE/_ref"
OUT="$HERE/_ref/mettagrid_c$(python3-config --extension-suffix)"
g++ -std=c++20 -O3 -DNDEBUG -fPIC -shared -fvisibility=hidden \
    -DNPY_NO_DEPRECATED_API=NPY_1_7_API_VERSION $(python3 -m pybind11 --includes) \
    -I"$R/include/mettagrid" -I"$R" "$R/bindings/mettagrid_c.cpp" $(find "$R/src" -name '*.cpp') -o "$OUT"
echo "$OUT"
