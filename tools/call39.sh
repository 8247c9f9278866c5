#!/bin/bash
python -m pytest tests/test_cce.py tests/test_dap.py tests/test_crc.py -m gpu -x -q 2>&1 | tail -3
