#!/bin/bash
# One GPU-box pass that regenerates everything under profiles/ for a tag (default "c"):
#   gpurun --timeout 1500 -- 'bash tools/gpu_round.sh c'
# Order matters: every ncu pass runs only after the same command has exited 0 without ncu, and a
# number printed under ncu is never a bench value (the bench lines come from the plain runs).
TAG=${1:-c}
O=gpurun_out
mkdir -p $O
python -m pytest tests -m gpu -x -q > $O/pytest_gpu_$TAG.log 2>&1; echo "pytest rc=$?"; tail -3 $O/pytest_gpu_$TAG.log
python __graft_entry__.py smoke > $O/smoke_$TAG.log 2>&1; echo "smoke rc=$?"; tail -2 $O/smoke_$TAG.log
python bench.py > $O/bench_$TAG.json 2> $O/bench_$TAG.err; echo "bench rc=$?"; cat $O/bench_$TAG.json
python bench.py --impl reference --steps 100 --warmup 3 > $O/bench_ref_$TAG.json 2> $O/bench_ref_$TAG.err; echo "ref rc=$?"
python bench_micro.py > $O/micro_$TAG.jsonl 2> $O/micro_$TAG.err; echo "micro rc=$?"; cat $O/micro_$TAG.jsonl
# launch list of the bench command (cold-cache, serialised per-launch times: shares, not absolutes)
python bench.py --steps 30 --warmup 5 --no-cpu-baseline > $O/plain_$TAG.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -s 80 -c 500 --csv --log-file $O/launches_$TAG.csv \
    python bench.py --steps 30 --warmup 5 --no-cpu-baseline > $O/ncu_$TAG.log 2>&1; echo "ncu launches rc=$?"
# full capture of the two dominant kernels of the frame chain
python bench.py --steps 10 --warmup 3 --no-cpu-baseline > $O/plain2_$TAG.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:"klt_track|pnp_chunk" -s 24 -c 6 -o $O/prof_$TAG -f \
    python bench.py --steps 10 --warmup 3 --no-cpu-baseline > $O/ncu_full_$TAG.log 2>&1; echo "ncu full rc=$?"
ncu -i $O/prof_$TAG.ncu-rep --page raw --csv > $O/prof_${TAG}_raw.csv 2>/dev/null
# detector kernels
if [ -n "$GFTT" ]; then
python bench_micro.py --only gftt > $O/plain_gftt_$TAG.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:"shi_" -c 4 -o $O/prof_gftt_$TAG -f \
    python bench_micro.py --only gftt > $O/ncu_gftt_$TAG.log 2>&1; echo "ncu gftt rc=$?"
ncu -i $O/prof_gftt_$TAG.ncu-rep --page raw --csv > $O/prof_gftt_${TAG}_raw.csv 2>/dev/null
fi
ls -la $O | tail -30
