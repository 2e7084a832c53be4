#!/bin/bash
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -x -q -m gpu -k "categorical" 2>&1 | tail -15
timeout 2400 python -m pytest tests -q -m gpu 2>&1 | tail -8
