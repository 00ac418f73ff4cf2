set -x
timeout 900 python -m pytest tests -q -m gpu 2>&1 | tail -3
timeout 300 python __graft_entry__.py smoke 2>&1 | tail -1
timeout 900 python bench.py > gpurun_out/bench_r1.json 2> gpurun_out/bench_r1.err; echo bench rc=$?
timeout 900 python bench.py --impl reference > gpurun_out/bench_r1_ref.json 2> gpurun_out/bench_r1_ref.err; echo ref rc=$?
timeout 900 python bench.py --steps 2 --warmup 1 > gpurun_out/bench_short.json 2>&1 && timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/r1_launches.csv python bench.py --steps 2 --warmup 1 > gpurun_out/ncu_launch.log 2>&1
NCU="ncu --set full --clock-control none --import-source on"
timeout 200 python scripts/profile_demod.py --sf 12 --packets 4096 --no-overlap > gpurun_out/p12.log 2>&1 && timeout 600 $NCU -k regex:demod_blocked -s 2 -c 1 -o gpurun_out/prof_sf12_fast -f python scripts/profile_demod.py --sf 12 --packets 4096 --no-overlap > gpurun_out/ncu12.log 2>&1
timeout 600 $NCU -k regex:prepass_blocked -s 1 -c 1 -o gpurun_out/prof_sf12_prepass -f python scripts/profile_demod.py --sf 12 --packets 4096 --no-overlap > gpurun_out/ncu12p.log 2>&1
timeout 200 python scripts/profile_demod.py --sf 9 --packets 32768 --no-overlap > gpurun_out/p9.log 2>&1 && timeout 600 $NCU -k regex:demod_blocked -s 2 -c 1 -o gpurun_out/prof_sf9_fast -f python scripts/profile_demod.py --sf 9 --packets 32768 --no-overlap > gpurun_out/ncu9.log 2>&1
timeout 200 python scripts/profile_stream.py --sf 10 --streams 2048 --frames 3 --symbols 30 > gpurun_out/ps10.log 2>&1 && timeout 600 $NCU -k regex:stream_detect -s 1 -c 1 -o gpurun_out/prof_sf10_stream -f python scripts/profile_stream.py --sf 10 --streams 2048 --frames 3 --symbols 30 > gpurun_out/ncups10.log 2>&1
cat gpurun_out/p12.log gpurun_out/p9.log gpurun_out/ps10.log | tail -12
