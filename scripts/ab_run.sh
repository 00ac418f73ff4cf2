#!/bin/bash
# Same-box A/B of the working-tree library against variants under ab/: scripts/ab_run.sh NAME [NAME ...]
# Prints the best demod-kernel rate of 5 repetitions per SF for the in-tree library ("tree") and each variant.
cd "$(dirname "$0")/.."
pk() { case $1 in 7) echo 65536;; 8) echo 65536;; 9) echo 32768;; 10) echo 16384;; 11) echo 8192;; 12) echo 4096;; esac; }
for sf in ${SFS:-7 8 9 10 11 12}; do
  for v in tree "$@"; do
    lib=""; [ "$v" != tree ] && lib="--lib ab/$v/liblora_phy_b200.so"
    best=$(python scripts/profile_demod.py --sf $sf --packets $(pk $sf) --reps 5 --no-overlap $lib ${MODE:+--mode $MODE} | grep -o "[0-9.]* Gsamples/s" | sort -n | tail -1)
    echo "sf=$sf $v $best"
  done
done
