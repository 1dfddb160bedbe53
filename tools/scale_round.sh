#!/bin/bash
# N = 1, 2, 4, 8 on ONE 8-GPU box, both arms:  gpurun --gpus 8 -- 'bash tools/scale_round.sh <tag>'
TAG=${1:-d}; O=gpurun_out; mkdir -p $O
python bench.py --steps 100 --warmup 10 > $O/r2_bench_n1_samebox_$TAG.json 2> $O/r2_bench_n1_samebox_$TAG.err; echo "n1 rc=$?"
for N in 2 4 8; do
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 295$N bench.py --gpus $N --steps 100 --warmup 10 > $O/r2_bench_n${N}_$TAG.json 2> $O/r2_bench_n${N}_$TAG.err; echo "n$N rc=$?"
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 296$N bench.py --impl reference --gpus $N --steps 60 --warmup 3 > $O/r2_bench_ref_n${N}_$TAG.json 2> $O/r2_bench_ref_n${N}_$TAG.err; echo "ref n$N rc=$?"
done
python - <<'PY'
import json,glob,sys
tag=sys.argv[1] if len(sys.argv)>1 else 'd'
for f in sorted(glob.glob('gpurun_out/r2_bench_*n*_%s.json'%tag)):
    try:
        d=json.load(open(f)); print(f, d['n_gpus'], round(d['value'],1), round(d['e2e']['value'],1), d.get('pipelined_poses_equal_synchronous'))
    except Exception as e: print(f, 'ERR', e)
PY
