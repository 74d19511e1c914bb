#!/bin/bash
# `ncu --set full` captures of the accumulation kernel on the final build of the round (run on the
# GPU box); only the CSV raw pages come back, named where bench.py looks for them.
OUT=gpurun_out/capture_r02d
mkdir -p $OUT
for a in "0 24 alt_bn128_g1" "1 20 alt_bn128_g2" "3 19 bw6_761_g1" "4 20 bls12_377_g2"; do
  set -- $a
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_accum_first \
      --launch-skip 1 -c 1 -f -o /tmp/cap_$3 python tools/run_once.py $1 $2 > /dev/null 2>&1
  ncu -i /tmp/cap_$3.ncu-rep --page raw --csv > $OUT/ncu_accum_$3_2p$2_r02d_raw.csv 2>/dev/null
  echo "capture $3 2^$2: $(wc -l < $OUT/ncu_accum_$3_2p$2_r02d_raw.csv) lines"
done
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv \
    --log-file $OUT/launches_r02d.csv python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-strong > /dev/null 2>&1
echo "launch list: $(grep -c k_ $OUT/launches_r02d.csv) rows"
