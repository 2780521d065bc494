#!/usr/bin/env bash
# Runs on the GPU box (under gpurun): launch list of one bench step + `ncu --set full` captures of the kernels that
# matter, converted to CSV on the box so that only small files travel back.
# usage: bash tools/profile.sh <tag> [what...]     what in {list, attn, istft, gn, ln, gemm}; default: all
set -u
TAG=${1:-rX}; shift || true
WHAT=${*:-list attn istft gn ln gemm}
OUT=gpurun_out/$TAG
mkdir -p "$OUT"
BENCH="python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-extra"
NCU="ncu --clock-control none"

full() {   # name regex skip count [import-source]
  local name=$1 rx=$2 skip=$3 cnt=$4 src=${5:-on}
  $NCU --set full --import-source $src -k regex:"$rx" -s "$skip" -c "$cnt" -f -o "$OUT/$name" $BENCH > "$OUT/$name.log" 2>&1
  ncu -i "$OUT/$name.ncu-rep" --page raw --csv > "$OUT/$name.raw.csv" 2>/dev/null
  if [ "$src" = on ]; then ncu -i "$OUT/$name.ncu-rep" --page source --csv > "$OUT/$name.source.csv" 2>/dev/null; fi
  # keep the report only if it is small enough to travel
  local sz; sz=$(stat -c %s "$OUT/$name.ncu-rep" 2>/dev/null || echo 0)
  if [ "$sz" -gt 12000000 ]; then rm -f "$OUT/$name.ncu-rep"; fi
}

for w in $WHAT; do
  case $w in
    list)  $NCU --metrics gpu__time_duration.sum --csv --log-file "$OUT/launches.csv" python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-extra > "$OUT/launches.log" 2>&1 ;;
    attn)  full attn "attention" 19 2 ;;        # one prenet-layer and one decoder-layer launch of the 2nd step
    istft) full istft "istft" 1 1 ;;
    gn)    full gn "gn_" 30 4 ;;
    ln)    full ln "layernorm" 40 3 ;;
    gemm)  full gemm "gemm" 74 74 off ;;        # one whole step of GEMM launches, no source
  esac
done
ls -la "$OUT"
