#!/bin/bash
# The round's bench lines, one JSON file each under gpurun_out/lines_$TAG/ (run on the GPU box).
TAG=${1:-r02}
OUT=gpurun_out/lines_$TAG
mkdir -p $OUT
run() { name=$1; shift; timeout 900 python bench.py "$@" > $OUT/$name.json 2> $OUT/$name.err || echo "FAILED $name"; tail -c 400 $OUT/$name.json; echo; }
run bench_${TAG}_2p24 
run bench_${TAG}_2p16 --log2n 16 --no-cpu-baseline
run bench_${TAG}_2p20 --log2n 20 --no-cpu-baseline
run bench_${TAG}_2p26 --log2n 26 --no-cpu-baseline --steps 3
run bench_${TAG}_bls12_377_g1_2p22 --group bls12_377_g1 --log2n 22 --no-cpu-baseline
run bench_${TAG}_alt_bn128_g2_2p20 --group alt_bn128_g2 --log2n 20 --no-cpu-baseline
run bench_${TAG}_bw6_761_g1_2p19 --group bw6_761_g1 --log2n 19 --no-cpu-baseline
run bench_${TAG}_bls12_377_g2_2p20 --group bls12_377_g2 --log2n 20 --no-cpu-baseline
run bench_${TAG}_bls12_381_g1_2p22 --group bls12_381_g1 --log2n 22 --no-cpu-baseline
run bench_${TAG}_bls12_381_g2_2p20 --group bls12_381_g2 --log2n 20 --no-cpu-baseline
run bench_${TAG}_batch_exp --op batch_exp
run bench_${TAG}_batch_to_special --op batch_to_special
run bench_${TAG}_reference --impl reference --steps 2 --warmup 1
