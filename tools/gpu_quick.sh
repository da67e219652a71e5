#!/bin/bash
set -u
timeout 900 python -m pytest tests/test_gpu_many_mask.py -q 2>&1 | tail -4
