#!/bin/bash
# usage (on the GPU box): tools/ab_test.sh "<tune_volpath.py args>" variant1 variant2 ...
# Times the in-tree library, then every acceleratedvolrenderer_b200/variants/libavr_<variant>.so swapped
# in its place, then the in-tree library again (drift check) — all on the same box.
set -e
cd "$(dirname "$0")/.."
args=$1; shift
lib=acceleratedvolrenderer_b200/libavr_b200.so
cp $lib /tmp/libavr_base.so
echo "== base"; python tools/tune_volpath.py $args 
for v in "$@"; do
  cp acceleratedvolrenderer_b200/variants/libavr_$v.so $lib
  echo "== $v"; python tools/tune_volpath.py $args 
done
cp /tmp/libavr_base.so $lib
echo "== base again"; python tools/tune_volpath.py $args 
