#!/bin/bash
python scripts/gpu_debug.py 1 2 36 12 0 0 1 3 0 2>&1 | tail -1 | cut -c1-160
python scripts/gpu_debug.py 3 12 36 12 16 0 1 200 1e-4 2>&1 | tail -1 | cut -c1-160
