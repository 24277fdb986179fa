#!/bin/bash
# postponed-leaf threshold sweep on the C2 soup (exact and fast BVH4 layouts)
for L in 1 4 8 12 16 20 24; do
  echo "== PTB_LEAF_MIN=$L"
  PTB_LEAF_MIN=$L python scratch/r02_trace_check.py soup4 2>&1 | cut -c1-200
done
for R in 1 2 8 12; do
  echo "== PTB_REFILL_MIN=$R"
  PTB_REFILL_MIN=$R python scratch/r02_trace_check.py soup4 2>&1 | cut -c1-200
done
