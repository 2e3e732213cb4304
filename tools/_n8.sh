set -u
mkdir -p gpurun_out
N=$1
run() { timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $1 bench.py --gpus $N --steps 50 --warmup 10 "${@:3}" > gpurun_out/$2.log 2>&1; echo "exit=$? $2"; grep '^{' gpurun_out/$2.log | cut -c1-260; }
run 29811 benchN_${N}_auto
run 29812 benchN_${N}_ref --impl reference
