#!/bin/bash
# Where does the merged persistent launch start to pay?  Whole demod_batch, best of 7, two-launch (0) vs merged (2).
cd "$(dirname "$0")/.."
for cfg in "7 1024" "7 4096" "7 16384" "9 512" "9 2048" "9 8192" "10 512" "10 2048" "11 384" "11 1024" "11 4096" "12 320" "12 640" "12 1280" "12 2560"; do
  set -- $cfg
  for ov in 0 2; do
    best=$(python scripts/profile_demod.py --sf $1 --packets $2 --reps 7 --overlap $ov | grep -o "whole demod_batch [0-9.]* ms" | grep -o "[0-9.]*" | sort -n | head -1)
    echo "sf=$1 packets=$2 overlap=$ov $best ms"
  done
done
