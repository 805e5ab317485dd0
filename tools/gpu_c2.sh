#!/bin/bash
timeout 300 python -m pytest tests -m gpu -x -q -k "csrmm or shard" 2>&1 | tail -2
timeout 300 python tools/time_kernels.py --cases c2,c2l,c5l,c5 2>&1 | cut -c1-200
