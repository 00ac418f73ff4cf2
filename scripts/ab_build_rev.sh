#!/bin/bash
# Builds liblora_phy_b200.so with demod_fast.cu (and the headers it includes) taken from a git revision, for same-box
# A/B timing of the working tree against a committed state:
#   scripts/ab_build_rev.sh NAME [REV=HEAD]   ->  ab/NAME/liblora_phy_b200.so
set -e
name=$1; rev=${2:-HEAD}
root=$(cd "$(dirname "$0")/.." && pwd)
pkg=$root/lora-sdr-lightweight-standalone-library-_b200
rel=lora-sdr-lightweight-standalone-library-_b200/csrc
make -s -C "$pkg" liblora_phy_b200.so
tmp=$(mktemp -d)
mkdir -p "$tmp/csrc" "$root/ab/$name"
for f in $(git -C "$root" ls-tree --name-only "$rev" "$rel/"); do git -C "$root" show "$rev:$f" > "$tmp/csrc/$(basename "$f")"; done
nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC -I"$tmp/csrc" -I"$root/include" \
     -c "$tmp/csrc/demod_fast.cu" -o "$root/ab/$name/demod_fast.o"
objs=$(ls "$pkg"/build/*.o | grep -v demod_fast.o | grep -v "/ref_\|/drv_")
nvcc -gencode arch=compute_100a,code=sm_100a -shared -o "$root/ab/$name/liblora_phy_b200.so" $objs "$root/ab/$name/demod_fast.o"
rm -rf "$tmp"
echo "$root/ab/$name/liblora_phy_b200.so"
