#!/bin/bash
# One gpurun call: GPU tests, the bench line, the ncu launch list and full captures of the top kernels.
# usage: gpurun --timeout 1800 -- 'bash tools/gpu_check.sh <tag>'
tag=${1:-r01}
out=gpurun_out
mkdir -p $out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $out/smi_$tag.txt 2>&1
if [ "$3" != "nopytest" ]; then
timeout 1800 python -m pytest tests -m gpu -x -q > $out/pytest_$tag.log 2>&1; echo "pytest exit $?" | tee -a $out/pytest_$tag.log
tail -3 $out/pytest_$tag.log
fi
timeout 600 python bench.py > $out/bench_$tag.json 2> $out/bench_$tag.err; echo "bench exit $?"
tail -c 600 $out/bench_$tag.json
if [ "$2" != "noncu" ]; then
timeout 300 python tools/profile_step.py > $out/plain_$tag.log 2>&1 || { echo 'plain profile_step failed'; tail -5 $out/plain_$tag.log; exit 1; }
L=$(grep -o 'launches_per_step=[0-9]*' $out/plain_$tag.log | cut -d= -f2) &&
B=$(grep -o 'launches_before_first_step=[0-9]*' $out/plain_$tag.log | cut -d= -f2) &&
timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none --csv \
    --log-file $out/launches_$tag.csv python tools/profile_step.py > $out/ncu_$tag.log 2>&1
echo "ncu list exit $?"
for spec in gemm_tc_kernel:20:8 attn_fa_kernel:0:1 attn_tc_kernel:0:2 temporal_attn:0:1 gn_apply_persist:2:2 gn_finalize_colstats:2:2 layernorm:0:2; do
  k=${spec%%:*}; r=${spec#*:}; sk=${r%%:*}; c=${r##*:}
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:$k \
      -s $sk -c $c -o $out/full_${k}_$tag -f python tools/profile_step.py >> $out/ncufull_$tag.log 2>&1
  echo "ncu full $k exit $?"
  ncu -i $out/full_${k}_$tag.ncu-rep --page raw --csv > $out/full_${k}_$tag.csv 2>/dev/null
done
# gpurun copies back at most 64 MiB: drop the largest reports first (the raw csv stays)
while [ $(du -sm $out | cut -f1) -gt 55 ]; do f=$(ls -S $out/*.ncu-rep 2>/dev/null | head -1); [ -z "$f" ] && break; rm -f "$f"; done
du -sh $out
fi
# microbenchmarks (standalone binaries built on the box)
if [ -f tools/ubench/ex2.cu ]; then
  nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o /tmp/ex2 tools/ubench/ex2.cu && /tmp/ex2 > $out/ubench_ex2_$tag.txt 2>&1
fi
