#!/bin/sh
# Builds tools/make_golden_cameras.cpp against the reference's own src/Camera.cpp (read in place) and
# writes tests/golden/cameras.json.  Needs /root/reference; run from the repo root.
set -e
REF=${REF:-/root/reference}
OUT=$(mktemp -d)
g++ -std=c++20 -O1 -D_USE_MATH_DEFINES -I "$REF/src" -I "$REF/lib" -I "$REF/lib/json/include" \
    tools/make_golden_cameras.cpp "$REF/src/Camera.cpp" -o "$OUT/camgold"
"$OUT/camgold" tests/golden/cameras.json
rm -rf "$OUT"
