# Round-2 evidence run (one gpurun call): bench lines, then the ncu captures DESIGN.md / profiles/README.md cite.
# Every ncu command runs only after the same command line has exited 0 without ncu in this call.  The .ncu-rep files
# are summarised ON the box (scripts/ncu_summary.py, the SASS page gzipped) and deleted: gpurun_out/ is capped at 64 MiB.
set -x
cd "$(dirname "$0")/.."
O=gpurun_out
mkdir -p $O
# --- the bench lines (never under ncu)
python bench.py > $O/r2_bench_1gpu.json 2> $O/r2_bench_1gpu.err; echo bench rc=$?
python bench.py --impl reference --steps 2 --warmup 1 > $O/r2_bench_reference_arm.json 2> $O/r2_bench_reference_arm.err; echo ref rc=$?
# --- launch list of the bench command
python bench.py --steps 2 --warmup 1 --no-extras --no-perf-test > $O/bench_short.json 2> $O/bench_short.err && \
  ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/r2_ncu_launches_bench.csv \
      python bench.py --steps 2 --warmup 1 --no-extras --no-perf-test > $O/ncu_launch.log 2>&1
# --- full captures, summarised here
NCU="ncu --set full --clock-control none --import-source on"
digest() { # $1 = report stem, $2 = summary name
  python scripts/ncu_summary.py $O/$1.ncu-rep > $O/$2_summary.txt 2>&1
  ncu -i $O/$1.ncu-rep --page source --csv --print-source sass 2>/dev/null | gzip -9 > $O/$2_sass.csv.gz
  rm -f $O/$1.ncu-rep
}
pk() { case $1 in 7) echo 65536;; 9) echo 32768;; 12) echo 4096;; esac; }
for sf in 7 9 12; do
  python scripts/profile_demod.py --sf $sf --packets $(pk $sf) --overlap 0 > $O/p$sf.log 2>&1 && \
    $NCU -k regex:demod_blocked -s 2 -c 1 -o $O/prof_sf${sf} -f python scripts/profile_demod.py --sf $sf --packets $(pk $sf) --overlap 0 > $O/ncu$sf.log 2>&1
  digest prof_sf${sf} r2_ncu_sf${sf}_fast
done
# the merged persistent launch (library default): scan + estimate items and demod items in one kernel
python scripts/profile_demod.py --sf 12 --packets 8192 --overlap 2 > $O/pm12.log 2>&1 && \
  $NCU -k regex:receive_merged -s 1 -c 1 -o $O/prof_m12 -f python scripts/profile_demod.py --sf 12 --packets 8192 --overlap 2 > $O/ncum12.log 2>&1
digest prof_m12 r2_ncu_sf12_merged
python scripts/profile_companions.py --sf 12 --which modulate,frame,chain --packets 16384 > $O/pc12.log 2>&1 && \
  $NCU -k regex:"modulate_kernel|chain_decode_kernel" -c 3 -o $O/prof_comp -f python scripts/profile_companions.py --sf 12 --which modulate,frame,chain --packets 16384 --reps 1 > $O/ncupc.log 2>&1
digest prof_comp r2_ncu_companions
python scripts/profile_companions.py --which stream7 > $O/ps7.log 2>&1 && \
  $NCU -k regex:stream_detect -s 1 -c 1 -o $O/prof_s7 -f python scripts/profile_companions.py --which stream7 --reps 1 > $O/ncups7.log 2>&1
digest prof_s7 r2_ncu_stream7
rm -f $O/*.ncu-rep
du -sh $O; ls -la $O | tail -30
head -c 1500 $O/r2_bench_1gpu.json; echo; head -c 600 $O/r2_bench_reference_arm.json
