#!/bin/bash
for i in 1 2; do
for g in 0 1; do echo "C4_SPLIT=$g"; C4_SPLIT=$g python tools/c4_breakdown.py 2>&1 | tail -1; done
done
