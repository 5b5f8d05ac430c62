#!/bin/bash
for pl in "" "24:8,4,12" "24:4,8,12" "24:7,5,12" "24:8,8,8" "24:9,3,12" "24:6,5,13" "24:8,3,13" "24:8,6,10"; do
  B2S_NTT_PLAN="$pl" python tools/lde_sweep.py 24 2>&1 | tail -1
done
for pl in "" "20:8,12" "20:7,13" "20:4,4,12" "20:8,4,8" "20:9,11" "20:6,6,8" "20:10,10"; do
  B2S_NTT_PLAN="$pl" python tools/lde_sweep.py 20 2>&1 | tail -1
done
for pl in "" "16:8,8" "16:4,12" "16:6,10" "16:3,13" "16:4,4,8"; do
  B2S_NTT_PLAN="$pl" python tools/lde_sweep.py 16 2>&1 | tail -1
done
