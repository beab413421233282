#!/bin/bash
# Launcher in the spirit of the reference's cbasics.sh (ref:cbasics.sh:1-4: pick the GPUs, run main.py) - no conda here.
# Usage: ./cbasics.sh [main.py flags...]      e.g. ./cbasics.sh --epochs 2 --gpus 4
CUDA_VISIBLE_DEVICES=${CUDA_VISIBLE_DEVICES:-0,1,2,3} python3 "$(dirname "$0")/main.py" "$@"
