# Round-2 evidence run (one gpurun call): same-box A/Bs, then the ncu captures DESIGN.md / profiles/README.md cite.
# Every ncu command runs only after the same command line has exited 0 without ncu in this call.
set -x
cd "$(dirname "$0")/.."
O=gpurun_out
python -m pytest tests -m gpu -x -q 2>&1 | tail -4
# --- A/B: modulator old vs new, twiddle-preload variants of the demod kernel
for sf in 7 12; do
  python scripts/profile_companions.py --sf $sf --which modulate 2>&1 | grep "^modulate" | sed 's/^/new /'
  python scripts/profile_companions.py --sf $sf --which modulate --lib ab/modold/liblora_phy_b200.so 2>&1 | grep "^modulate" | sed 's/^/old /'
done | tee $O/ab_modulator.txt
SFS="7 8 9 10" scripts/ab_run.sh pre78 pre710 2>&1 | tee $O/ab3.txt
# --- launch list of the bench command
python bench.py --steps 2 --warmup 1 --no-extras --no-perf-test > $O/bench_short.json 2> $O/bench_short.err && \
  ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/r2_ncu_launches_bench.csv \
      python bench.py --steps 2 --warmup 1 --no-extras --no-perf-test > $O/ncu_launch.log 2>&1
# --- full captures
NCU="ncu --set full --clock-control none --import-source on"
pk() { case $1 in 7) echo 65536;; 9) echo 32768;; 12) echo 4096;; esac; }
for sf in 7 9 12; do
  python scripts/profile_demod.py --sf $sf --packets $(pk $sf) --no-overlap > $O/p$sf.log 2>&1 && \
    $NCU -k regex:demod_blocked -s 2 -c 1 -o $O/r2_prof_sf${sf}_fast -f python scripts/profile_demod.py --sf $sf --packets $(pk $sf) --no-overlap > $O/ncu$sf.log 2>&1
done
python scripts/profile_companions.py --sf 12 --which modulate,frame,chain --packets 16384 > $O/pc12.log 2>&1 && \
  $NCU -k regex:"modulate_kernel|chain_decode_kernel" -c 4 -o $O/r2_prof_companions -f python scripts/profile_companions.py --sf 12 --which modulate,frame,chain --packets 16384 --reps 1 > $O/ncupc.log 2>&1
python scripts/profile_companions.py --which stream7 > $O/ps7.log 2>&1 && \
  $NCU -k regex:stream_detect -s 1 -c 1 -o $O/r2_prof_stream7 -f python scripts/profile_companions.py --which stream7 --reps 1 > $O/ncups7.log 2>&1
cat $O/p7.log $O/p9.log $O/p12.log $O/pc12.log $O/ps7.log | grep -v "^{" | tail -20
