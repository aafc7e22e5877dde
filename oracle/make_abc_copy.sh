#!/bin/sh
# Builds the L-ABC variant of the reference: copies
# /root/reference/src into a scratch build dir (never into the repo history:
# oracle/_ref/ is git-ignored) and applies the 8-line bug-fix patch with sed.
#   F3: Intersect(..., vec3 out_normal) -> vec3& out_normal   (5 sites)
#   F2: posi.x = ...[1] / [2]          -> posi.y / posi.z     (2 lines)
#   F4: PathTracing(nextRay, data, rand) -> (..., randomGene) (1 line)
# usage: make_abc_copy.sh <reference_src_dir> <out_dir>
set -e
SRC="$1"; OUT="$2"
mkdir -p "$OUT"
cp "$SRC"/Ray.h "$SRC"/Renderer.h "$SRC"/Renderer.cpp "$SRC"/SettingData.h "$SRC"/SettingData.cpp "$OUT"/
sed -i 's/double& out_dis, vec3 out_normal)/double\& out_dis, vec3\& out_normal)/' "$OUT"/SettingData.h "$OUT"/SettingData.cpp
sed -i 's/posi\.x = it_object\.value()\["00 position"\]\[1\]/posi.y = it_object.value()["00 position"][1]/; s/posi\.x = it_object\.value()\["00 position"\]\[2\]/posi.z = it_object.value()["00 position"][2]/' "$OUT"/SettingData.cpp
sed -i 's/PathTracing(nextRay, data, rand)/PathTracing(nextRay, data, randomGene)/' "$OUT"/Renderer.cpp
# verify: exactly the 8 expected lines differ
N=$(cat "$OUT"/SettingData.h "$OUT"/SettingData.cpp "$OUT"/Renderer.cpp | grep -c -E 'vec3& out_normal\)|posi\.[yz] = it_object|PathTracing\(nextRay, data, randomGene\)')
if [ "$N" != "8" ]; then echo "abc patch: expected 8 patched lines, got $N" >&2; exit 1; fi
