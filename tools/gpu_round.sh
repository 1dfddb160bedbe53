#!/bin/bash
# One GPU-box pass that regenerates everything under profiles/ for a tag (default "a"):
#   gpurun --timeout 2400 -- 'bash tools/gpu_round.sh a'
# Order matters: every ncu pass runs only after the same command has exited 0 without ncu, and a
# number printed under ncu is never a bench value (the bench lines come from the plain runs).
TAG=${1:-a}
O=gpurun_out
mkdir -p $O
rm -f $O/r2_parity_$TAG.json
VO_PARITY_REPORT=$O/r2_parity_$TAG.json python -m pytest tests -m gpu -x -q -s > $O/r2_pytest_gpu_$TAG.log 2>&1; echo "pytest rc=$?"; tail -3 $O/r2_pytest_gpu_$TAG.log
python __graft_entry__.py smoke > $O/r2_smoke_$TAG.log 2>&1; echo "smoke rc=$?"; tail -2 $O/r2_smoke_$TAG.log
python bench.py > $O/r2_bench_$TAG.json 2> $O/r2_bench_$TAG.err; echo "bench rc=$?"; cat $O/r2_bench_$TAG.json
python bench.py --steps 20 --warmup 5 > $O/r2_bench_s20_$TAG.json 2> $O/r2_bench_s20_$TAG.err; echo "bench s20 rc=$?"; cat $O/r2_bench_s20_$TAG.json
python bench.py --impl reference --steps 100 --warmup 3 > $O/r2_bench_ref_$TAG.json 2> $O/r2_bench_ref_$TAG.err; echo "ref rc=$?"; cat $O/r2_bench_ref_$TAG.json
for C in parking malaga pnp_stress; do
  python bench.py --config $C --steps 100 > $O/r2_bench_${C}_$TAG.json 2> $O/r2_bench_${C}_$TAG.err; echo "$C rc=$?"; cat $O/r2_bench_${C}_$TAG.json
  python bench.py --config $C --impl reference --steps 60 --warmup 3 > $O/r2_bench_ref_${C}_$TAG.json 2> $O/r2_bench_ref_${C}_$TAG.err; echo "$C ref rc=$?"
done
python bench_micro.py > $O/r2_micro_$TAG.jsonl 2> $O/r2_micro_$TAG.err; echo "micro rc=$?"; cat $O/r2_micro_$TAG.jsonl
# launch list of the bench command (cold-cache, serialised per-launch times: shares, not absolutes)
python bench.py --steps 30 --warmup 5 --no-cpu-baseline > $O/plain_$TAG.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -s 300 -c 600 --csv --log-file $O/r2_launches_$TAG.csv \
    python bench.py --steps 30 --warmup 5 --no-cpu-baseline > $O/ncu_$TAG.log 2>&1; echo "ncu launches rc=$?"
# full capture of the dominant kernels of the frame chain
python bench.py --steps 10 --warmup 3 --no-cpu-baseline > $O/plain2_$TAG.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:"klt_track|pnp_chunk|pnp_" -s 60 -c 8 -o $O/r2_prof_$TAG -f \
    python bench.py --steps 10 --warmup 3 --no-cpu-baseline > $O/ncu_full_$TAG.log 2>&1; echo "ncu full rc=$?"
ncu -i $O/r2_prof_$TAG.ncu-rep --page raw --csv > $O/r2_prof_${TAG}_raw.csv 2>/dev/null
# the saturated KLT launch (262 144 tracks) with the stall breakdown
python bench_micro.py --only klt > $O/plain_klt_$TAG.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:"klt_track" -s 2 -c 1 -o $O/r2_prof_kltsat_$TAG -f \
    python bench_micro.py --only klt --reps 2 > $O/ncu_kltsat_$TAG.log 2>&1; echo "ncu klt sat rc=$?"
ncu -i $O/r2_prof_kltsat_$TAG.ncu-rep --page raw --csv > $O/r2_prof_kltsat_${TAG}_raw.csv 2>/dev/null
if [ -n "$GFTT" ]; then
python bench_micro.py --only gftt > $O/plain_gftt_$TAG.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:"shi_" -c 4 -o $O/r2_prof_gftt_$TAG -f \
    python bench_micro.py --only gftt > $O/ncu_gftt_$TAG.log 2>&1; echo "ncu gftt rc=$?"
ncu -i $O/r2_prof_gftt_$TAG.ncu-rep --page raw --csv > $O/r2_prof_gftt_${TAG}_raw.csv 2>/dev/null
fi
ls -la $O | tail -40
