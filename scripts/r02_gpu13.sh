#!/bin/bash
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_loop.py -m gpu -x -q -k "u8 or prebin" 2>&1 | tail -30
