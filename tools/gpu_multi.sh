#!/bin/bash
# Multi-GPU check: GPU tests on GPU 0, then bench.py under torchrun at N ranks (and N=1 for the same build).
# usage: gpurun --gpus N --timeout 1200 -- 'bash tools/gpu_multi.sh <tag> <N>'
tag=$1; N=$2
out=gpurun_out; mkdir -p $out
timeout 600 python -m pytest tests -m gpu -x -q > $out/pytest_$tag.log 2>&1; echo "pytest exit $?"; tail -3 $out/pytest_$tag.log
timeout 600 python bench.py --no-cpu-baseline > $out/bench_${tag}_n1.json 2> $out/bench_${tag}_n1.err; echo "bench n1 exit $?"
tail -c 300 $out/bench_${tag}_n1.err
for n in $(seq 2 2 $N); do
  if [ $n -eq 2 ] || [ $n -eq 4 ] || [ $n -eq 8 ]; then
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29533 \
      bench.py --gpus $n > $out/bench_${tag}_n$n.json 2> $out/bench_${tag}_n$n.err; echo "bench n$n exit $?"
  tail -c 600 $out/bench_${tag}_n$n.err
  python -c "
import json,sys
d=json.loads(open('$out/bench_${tag}_n$n.json').read().strip().splitlines()[-1])
print({k:d[k] for k in ('value','n_gpus','ms_per_step','e2e','clocks')})"
  fi
done
python -c "
import json,sys
d=json.loads(open('$out/bench_${tag}_n1.json').read().strip().splitlines()[-1])
print({k:d[k] for k in ('value','n_gpus','ms_per_step','e2e')})
for k,v in sorted(d['families'].items(), key=lambda kv:-kv[1]['ms'])[:10]: print(k, v['launches'], v['ms'], v['tflops'], v['gbs'])"
