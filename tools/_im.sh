timeout 600 python -m pytest tests/test_kernels_gpu.py tests/test_engine_gpu.py -x -q -p no:cacheprovider -k "im2col or first_layer or bf16_tensor_core or trainer_end_to_end" 2>&1 | tail -4
for v in 0 1; do echo "rows8=$v: $(DDPB_IM2COL_ROWS8=$v timeout 300 python bench.py --steps 200 --warmup 20 2>/dev/null | tail -1 | grep -o '"ms_per_step": [0-9.]*' | head -2 | tr '\n' ' ')"; done
python tools/step_timeline.py --out gpurun_out/tl1_im2col.txt > /dev/null 2>&1; head -6 gpurun_out/tl1_im2col.txt
