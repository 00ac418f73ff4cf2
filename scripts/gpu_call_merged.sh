python -m pytest tests/test_gpu_parity.py -x -q --tb=short -k "merged" 2>&1 | tail -30
for cfg in "0 2.0" "0 3.0" "0 4.0" "0 6.0"; do
  set -- $cfg
  echo "== chunk=$1 lead=$2"
  LORA_B200_MERGED_CHUNK=$1 LORA_B200_MERGED_LEAD=$2 SFS="${SFS_SWEEP:-7 9 10 12}" scripts/ab_merged.sh 2>&1 | grep "overlap=2"
done | tee gpurun_out/r2_ab_merged4.txt
