#!/bin/bash
set -u
timeout 300 python tools/ab_permute.py 2>&1 | cut -c1-50,150-400
