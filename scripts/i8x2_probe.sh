#!/bin/bash
# first contact with the CTA-pair exact kernel: tiny case under a short timeout, then tests, then timing
set -x
RB_EXACT_KERNEL=i8x2 timeout 120 python scripts/exact_probe.py 1000000 512 1 2>&1 | tail -3
echo "small exit $?"
