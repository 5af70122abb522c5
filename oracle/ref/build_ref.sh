#!/bin/bash
# TEST INFRASTRUCTURE ONLY. Builds oracle/_ref/libatray_ref.so from the reference sources where they lie
# (/root/reference/Source). A scratch copy is made under a mktemp dir only because one header needs a
# layout-preserving patch for g++ (aabb.h:7-16: anonymous struct with non-trivial members inside a union is
# an MSVC extension). Nothing but the built .so and the staged mesh assets land in oracle/_ref/ (git-ignored).
set -euo pipefail
HERE="$(cd "$(dirname "$0")" && pwd)"
REF="${ATRAY_REFERENCE:-/root/reference}"
OUT="$HERE/../_ref"
if [ ! -d "$REF/Source" ]; then echo "build_ref: $REF/Source not present; keeping prebuilt oracle/_ref" >&2; exit 0; fi
mkdir -p "$OUT/assets"
STAGE="$(mktemp -d /tmp/atray_ref_stage.XXXXXX)"
trap 'rm -rf "$STAGE"' EXIT
cp -r "$REF/Source" "$STAGE/Source"
python3 - "$STAGE/Source/engine/renderer/aabb.h" <<'PY'
import re, sys
p = sys.argv[1]
s = open(p).read()
s2 = re.sub(r"union\s*\{\s*vec3f bounds\[2\];\s*struct\s*\{.*?vec3f min;\s*vec3f max;\s*\};\s*\};",
            "vec3f min;\n    vec3f max;", s, count=1, flags=re.S)
assert s2 != s, "aabb.h union pattern not found"
n = s2.count("bb.bounds[")
s2 = s2.replace("bb.bounds[", "(&bb.min)[")
assert n >= 12, n
open(p, "w").write(s2)
PY
g++ -std=c++17 -O2 -ffp-contract=off -fPIC -shared -fno-semantic-interposition -pthread -w \
    -I "$HERE/overlay" -I "$HERE" -I "$STAGE/Source" \
    "$HERE/ref_harness.cpp" -o "$OUT/libatray_ref.so"
# second build: the reference compiled WITHOUT `#define USE_KD_TREE` (renderer.h:8) — get_intersection_data's brute-force
# branch (renderer.cpp:58-82), the oracle of atray_scene_use_kd_tree(ctx, 0)
sed -i 's|^#define USE_KD_TREE|// #define USE_KD_TREE  (no-tree build)|' "$STAGE/Source/engine/renderer/renderer.h"
! grep -q '^#define USE_KD_TREE' "$STAGE/Source/engine/renderer/renderer.h"
g++ -std=c++17 -O2 -ffp-contract=off -fPIC -shared -fno-semantic-interposition -pthread -w \
    -I "$HERE/overlay" -I "$HERE" -I "$STAGE/Source" \
    "$HERE/ref_harness.cpp" -o "$OUT/libatray_ref_notree.so"
# mesh inputs for the parity tests travel beside the .so (data, not source; git-ignored)
cp "$REF/Assets/Cube.obj" "$REF/Assets/Deer.obj" "$REF/Assets/Monkey.obj" "$REF/Assets/Simple.obj" "$OUT/assets/"
echo "built $OUT/libatray_ref.so"
