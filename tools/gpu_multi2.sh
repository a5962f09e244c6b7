#!/bin/bash
# N-GPU measurements: bench.py as CFG-split pairs and as independent replicas, BASELINE config 3, and the reference's
# sweep (first LIMIT runs) with VAE decode + frame export, overlapped and synchronous.
# usage: gpurun --gpus N --timeout 1800 -- 'bash tools/gpu_multi2.sh <tag> <N> [LIMIT]'
tag=$1; N=$2; LIMIT=${3:-12}
out=gpurun_out; mkdir -p $out
run() { python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $1 "${@:2}"; }
last() { python - "$1" <<'PY'
import json, sys
try:
    d = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    keep = ("value", "n_gpus", "ms_per_step", "pair_latents_identical", "steps_per_s", "videos_per_hour", "seconds",
            "sweep_runs", "denoise_steps", "layout", "export", "sum_over_workers")
    print({k: d[k] for k in keep if k in d}, {k: d[k].get("value") for k in ("e2e",) if k in d},
          d.get("config", {}).get("layout"), d.get("pair_check"))
except Exception as e:
    print("no JSON line:", e)
PY
}
for layout in pairs replicas; do
  timeout 600 bash -c "$(declare -f run); N=$N run 29541 bench.py --gpus $N --steps 100 --warmup 10 --layout $layout --no-profile" \
      > $out/bench_${tag}_n${N}_$layout.json 2> $out/bench_${tag}_n${N}_$layout.err; echo "bench $layout exit $?"
  tail -c 400 $out/bench_${tag}_n${N}_$layout.err; last $out/bench_${tag}_n${N}_$layout.json
done
if [ "$N" = "2" ]; then
  timeout 600 bash -c "$(declare -f run); N=$N run 29542 bench.py --gpus 2 --steps 60 --warmup 10 --num-steps 15 --guidance 9.0 --no-profile" \
      > $out/bench_${tag}_config3_n2.json 2> $out/bench_${tag}_config3_n2.err; echo "config3 exit $?"
  last $out/bench_${tag}_config3_n2.json
fi
modes=("pairs overlapped:" "replicas overlapped:--no-pairs" "pairs synchronous:--sync-export")
[ -n "$SKIP_SYNC" ] && modes=("pairs overlapped:" "replicas overlapped:--no-pairs")     # the stall comparison is not re-measured
for mode in "${modes[@]}"; do
  name=${mode%%:*}; flag=${mode#*:}; d=$out/sweep_${tag}_$(echo $name | tr ' ' '_')
  rm -rf $d
  timeout 900 bash -c "$(declare -f run); N=$N run 29543 -m b200ad.sweep --output $d --limit $LIMIT $flag" \
      > $out/sweep_${tag}_n${N}_$(echo $name | tr ' ' '_').json 2> $out/sweep_${tag}_n${N}_$(echo $name | tr ' ' '_').err
  echo "sweep [$name] exit $?"; last $out/sweep_${tag}_n${N}_$(echo $name | tr ' ' '_').json
  echo "  markers: $(find $d -name config.json | wc -l), png: $(find $d -name '*.png' | wc -l), gif: $(find $d -name '*.gif' | wc -l)"
  rm -rf $d
done
