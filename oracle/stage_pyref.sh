#!/usr/bin/env bash
# TEST INFRASTRUCTURE ONLY.  /root/reference does not exist on the GPU box, and a Python reference "cannot travel" --
# except through oracle/_ref/, which is git-ignored (never part of this repo's history) but shipped by gpurun.  This
# script stages UNMODIFIED copies of the reference's Python callers of the hot path there:
#   oracle/_ref/pyref/droid_slam/   <- /root/reference/droid_slam/**/*.py        (modules/corr.py, depth_video.py, ...)
#   oracle/_ref/pyref/lietorch/     <- /root/reference/thirdparty/lietorch/lietorch/*.py   (needs lietorch_backends:
#                                      oracle/build_lietorch.sh)
# so that tests/test_reference_callers.py and bench.py's cpu_baseline can run the reference's own Python against either
# back end (oracle/pyref.py).  Nothing is patched on disk; a sha256 manifest of the staged files is written next to them.
set -euo pipefail
HERE="$(cd "$(dirname "$0")" && pwd)"
REF="${REFERENCE_ROOT:-/root/reference}"
OUT="$HERE/_ref/pyref"
[ -d "$REF/droid_slam" ] || { echo "no reference at $REF; nothing to stage"; exit 0; }
rm -rf "$OUT"
mkdir -p "$OUT/droid_slam/modules" "$OUT/droid_slam/geom" "$OUT/lietorch"
cp "$REF"/droid_slam/*.py "$OUT/droid_slam/"
cp "$REF"/droid_slam/modules/*.py "$OUT/droid_slam/modules/"
cp "$REF"/droid_slam/geom/*.py "$OUT/droid_slam/geom/"
cp "$REF"/thirdparty/lietorch/lietorch/*.py "$OUT/lietorch/"
(cd "$OUT" && find . -name '*.py' | sort | xargs sha256sum > MANIFEST.sha256)
echo "staged $(find "$OUT" -name '*.py' | wc -l) files under $OUT"
