# ncu --set full of the first GEMMs of a prenet layer and a decoder layer (QKV, Wo, gate/up, down) of the second bench step
set -u
OUT=gpurun_out/${1:-r1r}; mkdir -p $OUT
for spec in "pre 74 4" "dec 104 4"; do
  set -- $spec
  case " ${ONLY:-pre dec} " in *" $1 "*) ;; *) continue ;; esac
  ncu --clock-control none --set full --import-source off -k regex:gemm_tc3 -s $2 -c $3 -f -o $OUT/g_$1 python bench.py --steps 1 --warmup 1 --no-cpu-baseline > $OUT/g_$1.log 2>&1
  ncu -i $OUT/g_$1.ncu-rep --page raw --csv > $OUT/g_$1.raw.csv 2>/dev/null
  rm -f $OUT/g_$1.ncu-rep
  python tools/ncu_table.py $OUT/g_$1.raw.csv | cut -c1-150
done
