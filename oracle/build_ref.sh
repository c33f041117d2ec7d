#!/usr/bin/env bash
# TEST INFRASTRUCTURE ONLY (see oracle/README.md).
# Compiles the UNMODIFIED reference programs hw1..hw8 straight from /root/reference
# with g++ (no cmake, no source edits, no copies) into oracle/_ref/hwN.
# Flags follow the reference's CMake "Release" build (hwN/CMakeLists.txt): -O3 -DNDEBUG -std=c++17,
# OpenMP for hw5..hw8.  `-include array` works around color.h using std::array without the
# include (hw5..hw8 fail on GCC 13 otherwise).
set -euo pipefail
REF=${B200RT_REFERENCE:-/root/reference}
HERE="$(cd "$(dirname "$0")" && pwd)"
OUT="$HERE/_ref"
if [ ! -d "$REF/hw8/src" ]; then
  echo "build_ref: $REF not present; keeping prebuilt oracle/_ref as is" >&2
  exit 0
fi
mkdir -p "$OUT"
CXX=${B200RT_CXX:-/usr/bin/g++}   # NOT $CXX: the image exports CXX=/opt/gcc/bin/g++, which has no libgomp.spec
for n in 1 2 3 4 5 6 7 8; do
  src="$REF/hw$n/src"
  flags="-O3 -DNDEBUG -std=c++17 -I$src/include"
  [ "$n" -eq 1 ] && flags="-O2 $flags -fopenmp"
  [ "$n" -ge 5 ] && flags="$flags -fopenmp -include array"
  [ "$n" -ge 6 ] && flags="$flags -I$REF/rapidjson/include"
  [ "$n" -ge 8 ] && flags="$flags -I$REF/stb"
  if [ ! -x "$OUT/hw$n" ] || [ "$0" -nt "$OUT/hw$n" ]; then
    echo "build_ref: hw$n" >&2
    $CXX $flags "$src"/*.cpp -o "$OUT/hw$n" 2> "$OUT/hw$n.log" || { tail -5 "$OUT/hw$n.log" >&2; exit 1; }
  fi
done
echo "build_ref: ok -> $OUT" >&2
# Large example SCENES (data, not sources) for tools/config_sweep.py: BASELINE configs C4 name hw6/hw7 examples that are
# too big for tests/scenes; they are staged next to the binaries (git-ignored, but they travel to the GPU box).
mkdir -p "$OUT/scenes/hw6" "$OUT/scenes/hw7"
for f in practice6_1 practice6_2; do
  for e in gltf bin; do [ -f "$REF/hw6/examples/$f.$e" ] && cp -u "$REF/hw6/examples/$f.$e" "$OUT/scenes/hw6/"; done
done
for f in practice7_1 practice7_2 practice7_3 practice7_4; do
  for e in gltf bin; do [ -f "$REF/hw7/examples/$f.$e" ] && cp -u "$REF/hw7/examples/$f.$e" "$OUT/scenes/hw7/"; done
done
echo "build_ref: example scenes staged -> $OUT/scenes" >&2
