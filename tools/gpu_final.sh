#!/bin/bash
# Round-end check on one GPU: whole GPU suite, determinism stress, smoke(), the bench line, the ncu launch list.
# usage: gpurun --timeout 1500 -- 'bash tools/gpu_final.sh <tag>'
tag=${1:-final}
out=gpurun_out; mkdir -p $out
timeout 900 python -m pytest tests -m gpu -x -q > $out/pytest_$tag.log 2>&1; echo "pytest exit $?"; tail -2 $out/pytest_$tag.log
timeout 300 python tools/stress_determinism.py --rounds 6 > $out/stress_$tag.log 2>&1; echo "stress exit $?"; tail -1 $out/stress_$tag.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > $out/smoke_$tag.log 2>&1; echo "smoke exit $?"; tail -1 $out/smoke_$tag.log
timeout 600 python bench.py > $out/bench_$tag.json 2> $out/bench_$tag.err; echo "bench exit $?"
timeout 300 python tools/profile_step.py > $out/plain_$tag.log 2>&1 || { echo 'plain profile_step failed'; tail -5 $out/plain_$tag.log; exit 1; }
timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none --csv \
    --log-file $out/launches_$tag.csv python tools/profile_step.py > $out/ncu_$tag.log 2>&1
echo "ncu list exit $?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:temporal_attn -s 0 -c 1 -o $out/full_temporal_attn_$tag -f \
    python tools/profile_step.py >> $out/ncufull_$tag.log 2>&1
ncu -i $out/full_temporal_attn_$tag.ncu-rep --page raw --csv > $out/full_temporal_attn_$tag.csv 2>/dev/null; echo "ncu full temporal exit $?"
rm -f $out/full_temporal_attn_$tag.ncu-rep
