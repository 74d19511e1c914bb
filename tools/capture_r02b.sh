#!/bin/bash
# ncu evidence of the round (run on the GPU box): launch lists of one MSM per configuration and
# `ncu --set full` captures of the accumulation kernel; only CSV exports come back (gpurun_out/).
OUT=gpurun_out/capture_r02b
mkdir -p $OUT
for a in "0 16 alt_bn128_g1" "0 20 alt_bn128_g1" "1 20 alt_bn128_g2" "2 22 bls12_377_g1" "3 19 bw6_761_g1" "4 20 bls12_377_g2"; do
  set -- $a
  timeout 250 ncu --metrics gpu__time_duration.sum --clock-control none --csv \
      --log-file $OUT/launches_$3_2p$2_r02b.csv python tools/run_once.py $1 $2 > /dev/null 2>&1
  echo "launch list $3 2^$2: $(grep -c k_ $OUT/launches_$3_2p$2_r02b.csv) rows"
done
for a in "1 20 alt_bn128_g2" "2 22 bls12_377_g1" "4 20 bls12_377_g2"; do
  set -- $a
  timeout 400 ncu --set full --clock-control none --import-source on -k regex:k_accum_first \
      --launch-skip 1 -c 1 -f -o /tmp/cap_$3 python tools/run_once.py $1 $2 > /dev/null 2>&1
  ncu -i /tmp/cap_$3.ncu-rep --page raw --csv > $OUT/ncu_accum_$3_2p$2_r02b_raw.csv 2>/dev/null
  echo "capture $3 2^$2: $(wc -l < $OUT/ncu_accum_$3_2p$2_r02b_raw.csv) lines"
done
