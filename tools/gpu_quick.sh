#!/bin/bash
# Quick GPU iteration: tests, then whatever commands follow (each logged to gpurun_out/<tag>_<i>.log).
# usage: gpurun --timeout 900 -- 'bash tools/gpu_quick.sh <tag> "cmd1" "cmd2" ...'
tag=$1; shift
out=gpurun_out; mkdir -p $out
timeout 1500 python -m pytest tests -m gpu -x -q --durations=15 > $out/pytest_$tag.log 2>&1; echo "pytest exit $?"; tail -25 $out/pytest_$tag.log
i=0
for c in "$@"; do
  i=$((i+1))
  echo "== $c" > $out/${tag}_$i.log
  timeout 600 bash -c "$c" >> $out/${tag}_$i.log 2>&1; echo "cmd $i exit $?"
  tail -25 $out/${tag}_$i.log
done
