#!/bin/bash
# Drop-in boundary check: run the reference's UNMODIFIED scripts (staged by tools/stage_reference.sh
# under baseline/_ref/) against the b200ad `diffusers` shim (compat/ on PYTHONPATH) on the GPU box, and keep the
# transcripts under gpurun_out/boundary_*.txt (copied to profiles/ when committed).
# usage: gpurun --timeout 1500 -- 'bash tools/run_boundary.sh <tag>'
tag=${1:-r02}
ROOT=$(cd "$(dirname "$0")/.." && pwd)
out=$ROOT/gpurun_out; mkdir -p $out
export PYTHONPATH=$ROOT/compat:$ROOT:$PYTHONPATH
cd $ROOT/baseline/_ref || { echo "baseline/_ref missing: run tools/stage_reference.sh first"; exit 1; }
sha256sum -c SHA256SUMS > $out/boundary_sha_$tag.txt 2>&1; echo "reference files unmodified: exit $?"
run() {  # name, command...
  local name=$1; shift
  local t0=$(date +%s.%N)
  ( "$@" ) > $out/boundary_${name}_$tag.txt 2>&1
  local rc=$?
  echo "== $name exit $rc in $(python -c "import time,sys; print(round(time.time()-float(sys.argv[1]),1))" $t0) s" | tee -a $out/boundary_${name}_$tag.txt
  tail -12 $out/boundary_${name}_$tag.txt
}
run 02 timeout 600 python experiments/02_architecture_inspection.py
run 03 timeout 900 python experiments/03_trace_forward_pass.py
# 05: one video of the sweep's CFG phase (6 runs of 25 steps), stdin supplies the "Press Enter" newline (05:403)
run 05 bash -c "echo | timeout 1200 python experiments/05_grid_search_ablation.py --video portrait --phase cfg --output $out/ref05_$tag"
run 01 timeout 1200 python experiments/01_baseline_generation.py
ls $out/ref05_$tag | head -20 >> $out/boundary_05_$tag.txt
find $out/ref05_$tag outputs -name '*.png' | wc -l | sed 's/^/png frames written: /' | tee -a $out/boundary_05_$tag.txt
# keep the traces (text), drop the images (gpurun_out is capped at 64 MiB)
cp outputs/03_traces/unet_trace.txt $out/boundary_03_unet_trace_$tag.txt 2>/dev/null
find $out/ref05_$tag -name '*.png' -delete; find $out/ref05_$tag -name '*.gif' -delete
rm -rf outputs
