#!/bin/bash
python -m pytest tests/test_gpu_kernels.py -m gpu -x -q -k pair_class 2>&1 | tail -30
