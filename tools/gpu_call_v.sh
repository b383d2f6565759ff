#!/bin/bash
mkdir -p gpurun_out
timeout 120 python -m pytest tests/test_gpu_configs.py -m gpu -x -q -k "product_training_loop" 2>&1 | tail -15
