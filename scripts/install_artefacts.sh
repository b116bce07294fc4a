#!/bin/bash
# copies the artefacts of one scripts/gpu_r2j.sh visit from gpurun_out/ into profiles/ as r02_<tag>_*:  scripts/install_artefacts.sh <tag> [old tag to drop]
set -e
T=$1; OLD=$2; G=gpurun_out
[ -n "$OLD" ] && rm -f profiles/r02_${OLD}_bench.json profiles/r02_${OLD}_bench_ref.json profiles/r02_${OLD}_launches_bench.csv profiles/r02_${OLD}_ncu_full_*.txt profiles/r02_${OLD}_dyn_sass_hist_*.txt
cp $G/bench_$T.json profiles/r02_${T}_bench.json
cp $G/bench_ref_$T.json profiles/r02_${T}_bench_ref.json
cp $G/launches_$T.csv profiles/r02_${T}_launches_bench.csv
for k in fx14 fx10 f32c3 gos12; do python scripts/ncu_summary.py $G/${k}_${T}_raw.csv > profiles/r02_${T}_ncu_full_$k.txt; done
python scripts/ncu_dyn_hist.py $G/fx14_${T}_source.csv $((32768*16384)) > profiles/r02_${T}_dyn_sass_hist_fx14.txt
python scripts/ncu_dyn_hist.py $G/fx10_${T}_source.csv $((65536*1024)) > profiles/r02_${T}_dyn_sass_hist_fx10.txt
python scripts/ncu_dyn_hist.py $G/f32c3_${T}_source.csv $((32768*2048)) > profiles/r02_${T}_dyn_sass_hist_f32c3.txt
python scripts/ncu_dyn_hist.py $G/gos12_${T}_source.csv $((16384*4096)) > profiles/r02_${T}_dyn_sass_hist_gos12.txt
[ -n "$OLD" ] && sed -i "s/r02_${OLD}_/r02_${T}_/g" BASELINE.md DESIGN.md
ls profiles | grep -c "r02_${T}_"
