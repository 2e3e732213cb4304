#!/bin/bash
# Multi-GPU box call: tests restricted to the given world sizes, the sweep, and bench.py for every part.
# Usage: tools/gpu_multi.sh N [tests|sweep|bench|all]
set -u
cd "$(dirname "$0")/.."
N=${1:-2}; WHAT=${2:-all}
OUT=gpurun_out; mkdir -p $OUT
run() { local name=$1 t=$2; shift 2; echo "=== $name"; timeout $t "$@" > $OUT/$name.log 2>&1; echo "exit=$? ($name)"; tail -n ${TAILN:-3} $OUT/$name.log | cut -c1-${CUT:-400}; }
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
if [[ $WHAT == tests || $WHAT == all ]]; then
  # the parametrisations of THIS world size (and the un-parametrised 2-rank tests when N == 2)
  K="$N]"; [[ $N == 2 ]] && K="2] or 2-16 or bf16_ddp or checkpoint"
  [[ $N == 4 ]] && K="4] or 4-64 or 4-16"
  [[ $N == 8 ]] && K="8-32 or nvls_multicast_reduce_matches_nccl_baseline[8] or random_arrival_order[8] or after_backward"
  TAILN=6 run t_multi_w$N 1500 python -m pytest tests/test_multigpu.py -x -q -s -rs -p no:cacheprovider -k "$K"
fi
if [[ $WHAT == sweep || $WHAT == all ]]; then
  TAILN=14 CUT=200 run sweep_w$N 900 $TR --master-port 29541 tools/allreduce_sweep.py
fi
if [[ $WHAT == bench || $WHAT == all ]]; then
  p=29600
  for part in ${PARTS:-3 2a 2b}; do
    for impl in ${IMPLS:-ours reference}; do
      p=$((p+1))
      CUT=1500 TAILN=1 run bench_w${N}_p${part}_${impl} 600 $TR --master-port $p bench.py --gpus $N --steps ${STEPS:-50} --warmup 10 --part $part --impl $impl
    done
  done
fi
if [[ $WHAT == bench || $WHAT == all ]]; then
  # the other transport: unicast P2P where "auto" picked the switch (3+ GPUs), NVLS where it picked unicast (2 GPUs)
  ALT=0; ALTNAME=p2p; [[ $N == 2 ]] && { ALT=1; ALTNAME=nvls; }
  CUT=1500 TAILN=1 run bench_w${N}_p3_ours_$ALTNAME 600 env DDPB_NVLS=$ALT $TR --master-port 29650 bench.py --gpus $N --steps ${STEPS:-50} --warmup 10 --part 3
  TAILN=3 run timeline_w$N 300 $TR --master-port 29651 tools/step_timeline.py --out $OUT/step_timeline_w$N.txt
fi
echo done
