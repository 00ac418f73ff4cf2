#!/bin/bash
# Same-box comparison of the two-launch receive path (overlap 0) with the merged persistent launch (overlap 2):
# whole lora_b200_demod_batch time at bench.py's per-SF batch sizes, best of 5.
cd "$(dirname "$0")/.."
pk() { case $1 in 7) echo 65536;; 8) echo 65536;; 9) echo 65536;; 10) echo 65536;; 11) echo 32768;; 12) echo 16384;; esac; }
for sf in ${SFS:-7 8 9 10 11 12}; do
  for ov in 0 2; do
    best=$(python scripts/profile_demod.py --sf $sf --packets $(pk $sf) --reps 5 --overlap $ov ${LIB:+--lib $LIB} | grep -o "whole demod_batch [0-9.]* ms" | grep -o "[0-9.]*" | sort -n | head -1)
    echo "sf=$sf overlap=$ov whole demod_batch $best ms"
  done
done
