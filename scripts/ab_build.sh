#!/bin/bash
# Builds a variant of liblora_phy_b200.so for same-box A/B timing:
#   scripts/ab_build.sh NAME [-DMACRO=value ...]   ->  ab/NAME/liblora_phy_b200.so
# Only one source (demod_fast.cu, or csrc/$AB_SRC.cu) is recompiled with the extra flags; the other objects come from the normal build.  Compare with
#   python scripts/profile_demod.py --lib ab/NAME/liblora_phy_b200.so ...
# in ONE gpurun call next to the baseline library: clocks under the power cap differ by 5 % between boxes.
set -e
name=$1; shift
root=$(cd "$(dirname "$0")/.." && pwd)
pkg=$root/lora-sdr-lightweight-standalone-library-_b200
make -s -C "$pkg" liblora_phy_b200.so
mkdir -p "$root/ab/$name"
src=${AB_SRC:-demod_fast}   # which csrc/*.cu gets the extra flags
nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC -I"$pkg/csrc" -I"$root/include" "$@" \
     -c "$pkg/csrc/$src.cu" -o "$root/ab/$name/$src.o"
objs=$(ls "$pkg"/build/*.o | grep -v "/$src.o" | grep -v "/ref_\|/drv_")
nvcc -gencode arch=compute_100a,code=sm_100a -shared -o "$root/ab/$name/liblora_phy_b200.so" $objs "$root/ab/$name/$src.o"
echo "$root/ab/$name/liblora_phy_b200.so"
