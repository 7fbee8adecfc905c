#!/bin/bash
# Evidence of one round on the GPU box (run under gpurun): tests, bench lines, parity report, ncu launch list and one
# `--set full` capture of the kernels of one block.  Everything lands in gpurun_out/<tag>_*; summaries are made here
# afterwards with scripts/launch_summary.py / scripts/ncu_summary.py and copied to profiles/.
#   usage: scripts/evidence.sh r2v1 [quick]
tag=${1:-r2}
out=gpurun_out
mkdir -p $out
python -m pytest tests -m gpu -x -q > $out/${tag}_pytest_gpu.log 2>&1; tail -2 $out/${tag}_pytest_gpu.log
python bench.py --impl reference --steps 3 --warmup 1 > $out/${tag}_bench_reference.json 2> $out/${tag}_bench_reference.err
python bench.py --steps 100 --warmup 20 > $out/${tag}_bench_b64.json 2> $out/${tag}_bench_b64.err; tail -c 400 $out/${tag}_bench_b64.err
if [ "$2" != "quick" ]; then
  python scripts/parity_report.py --out $out/${tag}_parity_report.json > $out/${tag}_parity_report.log 2>&1; tail -1 $out/${tag}_parity_report.log
fi
# launch list: init + 2 un-graphed steps (same command run plain first, as the profiling recipe asks)
python scripts/one_step.py 64 2 > $out/${tag}_plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none --csv \
    --log-file $out/${tag}_launches_b64.csv python scripts/one_step.py 64 2 > $out/${tag}_ncu_launches.log 2>&1
# full capture of the kernels of block 1: fc2 of block 0 (prompt-fold producer), prompt_mid, qkv, attention, proj, fc1, fc2
python scripts/one_step.py 64 1 > $out/${tag}_plain1.log 2>&1 && \
ncu --set full --clock-control none --import-source on -s 11 -c 7 -f -o $out/${tag}_top python scripts/one_step.py 64 1 > $out/${tag}_ncu_full.log 2>&1
ls -la $out | grep ${tag}
