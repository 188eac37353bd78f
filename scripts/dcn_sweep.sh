#!/bin/bash
# sweeps the internal tuning knobs of dcnv2_forward (env vars read by the C library)
for cfg in "" "CTDET_DCN_MMA_SLEEP=32" "CTDET_DCN_MMA_SLEEP=32 CTDET_DCN_LD_SLEEP=0" "CTDET_DCN_LD_SLEEP=0" "CTDET_DCN_LD_SLEEP=20 CTDET_DCN_MMA_SLEEP=100" "CTDET_DCN_STAGES=6 CTDET_DCN_LD_SLEEP=0" "CTDET_DCN_STAGES=5 CTDET_DCN_LD_SLEEP=0 CTDET_DCN_MMA_SLEEP=32"; do
  echo "== $cfg"
  env $cfg timeout 120 python scripts/dcn_bench.py 32 10 2>&1 | tail -4
done
