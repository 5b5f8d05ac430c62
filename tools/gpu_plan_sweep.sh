#!/bin/bash
for pl in "" "24:8,4,12" "24:6,6,12" "24:9,7,8" "24:8,7,9" "24:7,5,12"; do
  B2S_NTT_PLAN="$pl" python tools/lde_sweep.py 24 2>&1 | tail -1
done
for pl in "" "22:8,6,8" "22:6,4,12" "22:5,5,12" "22:8,5,9"; do
  B2S_NTT_PLAN="$pl" python tools/lde_sweep.py 22 2>&1 | tail -1
done
for pl in "" "20:9,11" "20:7,13" "20:4,4,12"; do
  B2S_NTT_PLAN="$pl" python tools/lde_sweep.py 20 2>&1 | tail -1
done
for pl in "" "18:6,12" "18:9,9" "18:5,13"; do
  B2S_NTT_PLAN="$pl" python tools/lde_sweep.py 18 2>&1 | tail -1
done
for pl in "" "14:6,8" "14:4,10" "14:5,9"; do
  B2S_NTT_PLAN="$pl" python tools/lde_sweep.py 14 2>&1 | tail -1
done
for g in 16777216 33554432 134217728 268435456; do
  B2S_LDE_GROUP_ELEMS=$g python tools/lde_sweep.py 24 2>&1 | tail -1
done
