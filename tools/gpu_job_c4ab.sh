#!/bin/bash
for i in 1 2; do
for g in 0 1; do echo "MIO_WAVE_GROW=$g"; MIO_WAVE_GROW=$g python tools/c4_breakdown.py 2>&1 | tail -1; done
done
