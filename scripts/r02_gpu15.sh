#!/bin/bash
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_loop.py -m gpu -x -q -k "forest or sparse" 2>&1 | tail -30
