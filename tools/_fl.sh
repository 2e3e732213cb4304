timeout 600 python -m pytest tests/test_kernels_gpu.py tests/test_engine_gpu.py -x -q -p no:cacheprovider -k "first_layer or bf16_tensor_core or trainer_end_to_end or window_loss" 2>&1 | tail -4
for v in 0 1; do echo "bulk=$v: $(DDPB_FL_BULK=$v timeout 300 python bench.py --steps 200 --warmup 20 2>/dev/null | tail -1 | grep -o '"ms_per_step": [0-9.]*' | head -2 | tr '\n' ' ')"; done
python tools/step_timeline.py --out gpurun_out/tl1_flbulk.txt > /dev/null 2>&1; grep "first_layer\|STEP" gpurun_out/tl1_flbulk.txt
