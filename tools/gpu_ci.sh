#!/bin/bash
# One GPU-box call.  Each risky group runs in its own process so a trapped kernel cannot take the
# other groups down.  Usage: tools/gpu_ci.sh [stage ...]; stages: kernels engine bench kbench ref ncu
set -u
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
OUT=gpurun_out
STAGES="${*:-kernels engine bench}"
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw,memory.total --format=csv > $OUT/gpu.csv 2>&1
: > $OUT/summary.log
run() { # name, timeout, cmd...
  local name=$1 t=$2; shift 2
  echo "=== $name" | tee -a $OUT/summary.log
  timeout $t "$@" > $OUT/$name.log 2>&1
  echo "exit=$? ($name)" | tee -a $OUT/summary.log
  tail -n ${TAILN:-6} $OUT/$name.log | tee -a $OUT/summary.log
}
for st in $STAGES; do
case $st in
kernels)
  run t_basic 600 python -m pytest tests/test_kernels_gpu.py -q -s -k "not conv_tc" -p no:cacheprovider
  run t_tc 900 python -m pytest tests/test_kernels_gpu.py -q -s -k "conv_tc" -p no:cacheprovider ;;
engine)
  TAILN=25 run t_engine 900 python -m pytest tests/test_engine_gpu.py -q -s -p no:cacheprovider ;;
multigpu)
  TAILN=25 run t_multigpu 900 python -m pytest tests/test_multigpu.py -q -s -p no:cacheprovider ;;
smoke)
  run smoke 300 python __graft_entry__.py --smoke ;;
zz)
  run t_zz_alone 300 python -m pytest tests/test_zz_no_alloc.py -q -p no:cacheprovider ;;
full)
  TAILN=15 run t_full 1500 python -m pytest tests -x -q -m gpu -p no:cacheprovider ;;
bench)
  run bench_ours 600 python bench.py --steps 50 --warmup 10
  run bench_ours_nograph 600 python bench.py --steps 50 --warmup 10 --no-graph ;;
bench2)
  run bench2_p2p 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus 2 --steps 50 --warmup 10
  run bench2_nvls 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29536 bench.py --gpus 2 --steps 50 --warmup 10 --comm nvls
  run bench2_nccl 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29534 bench.py --gpus 2 --steps 50 --warmup 10 --comm nccl
  run bench2_ref 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29535 bench.py --gpus 2 --steps 40 --warmup 10 --impl reference ;;
benchN)
  # scaling runs on N GPUs of one box: ours (auto -> NVLS when available), P2P unicast, NCCL baseline, reference
  for n in ${NLIST:-4 8}; do
    port=$((29600 + n))
    run benchN_${n}_auto 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $port bench.py --gpus $n --steps 50 --warmup 10
    run benchN_${n}_p2p 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((port+10)) bench.py --gpus $n --steps 50 --warmup 10 --comm p2p
    run benchN_${n}_nccl 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((port+20)) bench.py --gpus $n --steps 50 --warmup 10 --comm nccl
    run benchN_${n}_ref 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((port+30)) bench.py --gpus $n --steps 40 --warmup 10 --impl reference
  done ;;
sweep)
  TAILN=20 run sweep 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node ${NGPU:-2} --master-addr 127.0.0.1 --master-port 29541 tools/allreduce_sweep.py ;;
cli)
  # the main.py-style entry points, one process per GPU, started by hand like the reference
  ( python -m cs744_ddp_b200.parts.part3 --master 127.0.0.1 --num-nodes 2 --rank 1 --synthetic --dtype bf16 --iters 60 --train-size 20480 --test-size 1024 --port 6591 > $OUT/cli_part3_rank1.log 2>&1 & )
  run cli_part3_rank0 600 python -m cs744_ddp_b200.parts.part3 --master 127.0.0.1 --num-nodes 2 --rank 0 --synthetic --dtype bf16 --iters 60 --train-size 20480 --test-size 1024 --port 6591
  sleep 2
  ( python -m cs744_ddp_b200.parts.part2a --master 127.0.0.1 --num-nodes 2 --rank 1 --synthetic --iters 40 --train-size 20480 --test-size 512 --port 6592 > $OUT/cli_part2a_rank1.log 2>&1 & )
  run cli_part2a_rank0 600 python -m cs744_ddp_b200.parts.part2a --master 127.0.0.1 --num-nodes 2 --rank 0 --synthetic --iters 40 --train-size 20480 --test-size 512 --port 6592
  sleep 2 ;;
kbench)
  TAILN=80 run kernel_bench 600 python tools/kernel_bench.py ;;
ref)
  run bench_ref 900 python bench.py --impl reference --steps 40 --warmup 10 ;;
ncu)
  # launch list of two training steps (only after the same command exited 0 without ncu)
  python bench.py --steps 2 --warmup 3 --no-graph > $OUT/ncu_plain.log 2>&1 &&
  timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none --cache-control none -s ${NCU_LIST_SKIP:-200} -c ${NCU_LIST_COUNT:-60} --csv \
      --log-file $OUT/launches.csv python bench.py --steps 2 --warmup 3 --no-graph > $OUT/ncu_run.log 2>&1
  echo "ncu exit=$?" | tee -a $OUT/summary.log ;;
ncufull)
  python bench.py --steps 2 --warmup 3 --no-graph > $OUT/ncu_plain.log 2>&1 &&
  timeout 1200 ncu --set full --clock-control none --import-source on -k regex:"${NCU_KERNEL:-conv3x3_tc}" -s ${NCU_SKIP:-30} -c ${NCU_COUNT:-3} \
      -o $OUT/prof_${NCU_TAG:-conv} -f python bench.py --steps 2 --warmup 3 --no-graph > $OUT/ncufull_run.log 2>&1
  echo "ncufull exit=$?" | tee -a $OUT/summary.log ;;
ncustep)
  # ncu --set full over ONE whole eager training step (about 50 launches: every kernel of the step once); the raw
  # page is exported to CSV on the box and the report itself is dropped when it would not fit the 64 MiB merge limit
  python bench.py --steps 2 --warmup 3 --no-graph > $OUT/ncu_plain.log 2>&1 &&
  timeout 1200 ncu --set full --clock-control none --kernel-name-base mangled -k regex:ddpb -s ${NCU_SKIP:-150} -c ${NCU_COUNT:-50} \
      -o $OUT/prof_step -f python bench.py --steps 2 --warmup 3 --no-graph > $OUT/ncustep_run.log 2>&1
  echo "ncustep exit=$?" | tee -a $OUT/summary.log
  ncu -i $OUT/prof_step.ncu-rep --page raw --csv > $OUT/ncu_step_raw.csv 2>/dev/null
  ls -la $OUT/prof_step.ncu-rep | tee -a $OUT/summary.log
  [ $(stat -c %s $OUT/prof_step.ncu-rep 2>/dev/null || echo 0) -gt 40000000 ] && rm -f $OUT/prof_step.ncu-rep ;;
benchjson)
  python bench.py --steps 200 --warmup 20 2>/dev/null | tail -1 > $OUT/bench_n1_ours.json; cut -c1-400 $OUT/bench_n1_ours.json
  python bench.py --steps 100 --warmup 10 --dtype fp32 2>/dev/null | tail -1 > $OUT/bench_n1_ours_fp32.json; cut -c1-400 $OUT/bench_n1_ours_fp32.json
  python bench.py 2>/dev/null | tail -1 > $OUT/bench_n1_ours_defaults.json; cut -c1-600 $OUT/bench_n1_ours_defaults.json
  python bench.py --impl reference --steps 40 --warmup 10 2>/dev/null | tail -1 > $OUT/bench_n1_reference.json; cut -c1-400 $OUT/bench_n1_reference.json ;;
esac
done
grep -h "\[num\]\|\[eng\]" $OUT/t_*.log 2>/dev/null | sort | uniq > $OUT/numerics.log
echo done
