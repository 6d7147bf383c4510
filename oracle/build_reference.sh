#!/bin/bash
# Test infrastructure: builds the UNMODIFIED Mitsuba 2 reference (scalar_rgb + scalar_spectral, CLI + plugins, no Python,
# no GUI) from a scratch copy of /root/reference and installs its `dist/` under oracle/_ref/mitsuba/ (git-ignored, travels
# to the GPU box with the snapshot).  Recipe = SURVEY.md section 8(c).  No reference source is copied into the repository:
# the copy lives under $WORK (default /tmp/mts_ref_build) and only built binaries + the reference's run-time data files
# (dist/data: srgb.coeff, ior tables) land in oracle/_ref/.
#
# Local edits applied to the COPY (none to the renderer's algorithms):
#   1. mitsuba.conf: default variant scalar_rgb (enabled: scalar_rgb, scalar_spectral = the template's own list)
#   2. ext/CMakeLists.txt: build the vendored libpng/libjpeg on Linux too (the image has no system png/jpeg headers)
#   3. compiler flags: -fpermissive -Wno-changes-meaning (vendored TBB 2019 vs GCC 13), rpath-link for the vendored libs
# Usage: oracle/build_reference.sh [jobs]      (log: oracle/_ref/mitsuba/build.log)
set -e
HERE="$(cd "$(dirname "$0")" && pwd)"
REF="${REF:-/root/reference}"
WORK="${WORK:-/tmp/mts_ref_build}"
OUT="$HERE/_ref/mitsuba"
JOBS="${1:-6}"
if [ ! -d "$REF/src/librender" ]; then echo "reference tree not present at $REF: keeping prebuilt $OUT (if any)"; exit 0; fi
mkdir -p "$OUT"
LOG="$OUT/build.log"
: > "$LOG"
if [ ! -d "$WORK/src" ]; then
  mkdir -p "$WORK"
  cp -r "$REF" "$WORK/src"
  chmod -R u+w "$WORK/src"
fi
cd "$WORK/src"
if [ ! -f mitsuba.conf ]; then
  cp resources/mitsuba.conf.template mitsuba.conf
  sed -i 's/"default": "scalar_spectral"/"default": "scalar_rgb"/' mitsuba.conf
fi
sed -i 's/^if (WIN32 OR APPLE)$/if (TRUE)/' ext/CMakeLists.txt
B="$WORK/src/build"
mkdir -p "$B"
cd "$B"
cmake -GNinja .. -DCMAKE_POLICY_VERSION_MINIMUM=3.5 -DMTS_ENABLE_PYTHON=OFF -DMTS_ENABLE_GUI=OFF -DCMAKE_BUILD_TYPE=Release \
  -DCMAKE_CXX_FLAGS="-fpermissive -Wno-changes-meaning -w" \
  -DCMAKE_EXE_LINKER_FLAGS="-Wl,-rpath-link,$B/ext_build/libpng -Wl,-rpath-link,$B/ext_build/libjpeg -Wl,--allow-shlib-undefined" \
  >> "$LOG" 2>&1
ninja -j "$JOBS" >> "$LOG" 2>&1
# install: dist/ + the vendored png/jpeg shared objects next to it
rm -rf "$OUT/dist"
cp -rL "$B/dist" "$OUT/dist"
find "$B/ext_build" -name 'lib*.so*' \( -name '*png*' -o -name '*jpeg*' \) -exec cp -L {} "$OUT/dist/" \; 2>/dev/null || true
strip --strip-unneeded "$OUT"/dist/*.so* "$OUT"/dist/plugins/*.so "$OUT"/dist/mitsuba 2>/dev/null || true
echo "built: $(ls "$OUT/dist" | wc -l) entries, $(du -sh "$OUT/dist" | cut -f1)" | tee -a "$LOG"
