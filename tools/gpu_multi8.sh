#!/bin/bash
# 8-GPU measurements (one call, 8x GPU-minutes): bench.py as 4 CFG-split pairs and as 8 replicas, the reference's FULL
# 78-run sweep (2130 denoise steps) with VAE decode + PNG/GIF export in both layouts, and BASELINE config 5 (32 f 768^2).
# usage: gpurun --gpus 8 --timeout 1500 -- 'bash tools/gpu_multi8.sh <tag>'
tag=$1; N=${2:-8}
out=gpurun_out; mkdir -p $out
nproc; free -g | head -2
run() { python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $1 "${@:2}"; }
last() { python - "$1" <<'PY'
import json, sys
try:
    d = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    keep = ("metric", "value", "n_gpus", "ms_per_step", "pair_latents_identical", "steps_per_s", "videos_per_hour", "seconds",
            "sweep_runs", "denoise_steps", "layout", "export", "host_cores", "sum_over_workers")
    print({k: d[k] for k in keep if k in d}, {k: d[k].get("value") for k in ("e2e",) if k in d},
          d.get("config", {}).get("layout"))
except Exception as e:
    print("no JSON line:", e)
PY
}
for layout in replicas pairs; do
  timeout 400 bash -c "$(declare -f run); N=$N run 29551 bench.py --gpus $N --steps 100 --warmup 10 --layout $layout --no-profile" \
      > $out/bench_${tag}_n${N}_$layout.json 2> $out/bench_${tag}_n${N}_$layout.err; echo "bench $layout exit $?"
  tail -c 300 $out/bench_${tag}_n${N}_$layout.err; last $out/bench_${tag}_n${N}_$layout.json
done
for mode in "replicas:--no-pairs" "pairs:"; do
  name=${mode%%:*}; flag=${mode#*:}; d=/tmp/sweep_${tag}_$name
  rm -rf $d
  timeout 600 bash -c "$(declare -f run); N=$N run 29553 -m b200ad.sweep --output $d $flag" \
      > $out/sweep_${tag}_n${N}_$name.json 2> $out/sweep_${tag}_n${N}_$name.err
  echo "sweep [$name] exit $?"; tail -c 300 $out/sweep_${tag}_n${N}_$name.err; last $out/sweep_${tag}_n${N}_$name.json
  echo "  markers: $(find $d -name config.json | wc -l), png: $(find $d -name '*.png' | wc -l), gif: $(find $d -name '*.gif' | wc -l)"
  cp $d/manifest.json $out/sweep_${tag}_n${N}_${name}_manifest.json 2>/dev/null
  rm -rf $d
done
timeout 500 bash -c "$(declare -f run); N=$N run 29555 bench.py --gpus $N --steps 20 --warmup 5 --frames 32 --latent 96 --layout replicas --no-profile" \
    > $out/bench_${tag}_config5_n${N}.json 2> $out/bench_${tag}_config5_n${N}.err; echo "config5 exit $?"
tail -c 300 $out/bench_${tag}_config5_n${N}.err; last $out/bench_${tag}_config5_n${N}.json
