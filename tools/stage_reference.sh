#!/bin/bash
# Stage the reference's OWN scripts, unmodified, under baseline/_ref/ (git-ignored, NOT gpurun-ignored: it travels to
# the GPU box, where /root/reference does not exist).  Nothing is copied into the tracked tree.
# usage: bash tools/stage_reference.sh   (in the build container, where /root/reference is mounted)
set -e
ROOT=$(cd "$(dirname "$0")/.." && pwd)
SRC=${1:-/root/reference}
[ -d "$SRC/experiments" ] || { echo "no reference at $SRC"; exit 1; }
mkdir -p "$ROOT/baseline/_ref"
rm -rf "$ROOT/baseline/_ref/experiments" "$ROOT/baseline/_ref/utils"
cp -r "$SRC/experiments" "$SRC/utils" "$ROOT/baseline/_ref/"
( cd "$SRC" && find experiments utils -type f -name '*.py' | sort | xargs sha256sum ) > "$ROOT/baseline/_ref/SHA256SUMS"
echo "staged $(wc -l < "$ROOT/baseline/_ref/SHA256SUMS") reference files under baseline/_ref/"
