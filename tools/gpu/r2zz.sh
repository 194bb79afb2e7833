#!/bin/bash
# which form of the path kernel for a half frame, now that the shading is shorter (pool form gained more)
for g in 2 1; do
  for f in default pool; do
    E=""; [ $f = pool ] && E="RTC_PATHS=pool"
    env $E timeout 60 python tools/rank_emul.py --gpus $g --frames 5 2>&1 | tail -n 2 | cut -c1-75 | tr '\n' '|'; echo " <- n$g $f"
  done
done
