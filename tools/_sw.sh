for c in 148 111 74 37 222; do
  echo "== L0 wgrad CTAs $c"
  DDPB_WGRAD0_CTAS=$c python bench.py --steps 50 --warmup 10 2>&1 | grep '^{' | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['ms_per_step'], d['config']['timeline']['backward_compute_end_ms'], d['config']['timeline']['eager_step_ms'])"
done
