timeout 300 python -m pytest tests/test_kernels_gpu.py -x -q -k "grouped_weight_gradients_in_the_model" 2>&1 | grep -v Warning | grep -E "^E|assert|Error" | head -20
