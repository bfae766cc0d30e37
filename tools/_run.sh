mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_model_gpu.py -q -s -k per_layer > gpurun_out/s2d_layer.log 2>&1
grep -E "layer|passed|failed|Error" gpurun_out/s2d_layer.log | tail -90
