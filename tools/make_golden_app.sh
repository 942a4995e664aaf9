#!/bin/sh
# Builds tools/make_golden_app.cpp against the reference's own src/Project.h, src/Camera.cpp, nlohmann json and glm
# (read in place) and writes tests/golden/app_project.json.  Needs /root/reference; run from the repo root.
set -e
REF=${REF:-/root/reference}
OUT=$(mktemp -d)
g++ -std=c++20 -O1 -D_USE_MATH_DEFINES -I "$REF/src" -I "$REF/lib" -I "$REF/lib/json/include" \
    tools/make_golden_app.cpp "$REF/src/Camera.cpp" -o "$OUT/appgold"
"$OUT/appgold" tests/golden/app_project.json
rm -rf "$OUT"
