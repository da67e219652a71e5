#!/bin/bash
set -u
mkdir -p gpurun_out
echo "== alone, fresh process"
timeout 600 python -m pytest tests/test_zz_no_alloc.py -q -x -m gpu 2>&1 | tail -4
echo "== after the whole GPU suite"
timeout 2400 python -m pytest tests -q -x -m gpu 2>&1 | tail -6
