#!/bin/bash
# Packets per warp of the chirp pipeline (32 = one recurrence per lane, as in round 1; 16, 8 = more warps per batch).
cd "$(dirname "$0")/.."
for sf in 7 10 12; do
  for ppw in 32 16 8; do
    echo "ppw=$ppw $(LORA_B200_PPW=$ppw python scripts/profile_companions.py --sf $sf --which modulate --packets 65536 2>&1 | grep '^modulate')"
    echo "ppw=$ppw $(LORA_B200_PPW=$ppw python scripts/profile_companions.py --sf $sf --which frame --packets 32768 2>&1 | grep '^frame')"
  done
done
