#!/bin/bash
# One profiling pass of the benched build on a B200 box (run through gpurun from the repository root):
#   bash tools/profile_round.sh r2
# Leaves in gpurun_out/: the bench lines, the ncu launch list of the bench command, one `ncu --set full` capture
# per kernel of interest (first launch, at the benched size) with its raw / source pages as CSV.
# tools/summarize_profiles.py turns them into the summaries kept under profiles/.
tag=${1:-r2}
out=gpurun_out
mkdir -p $out
python bench.py --steps 20 --warmup 3 > $out/${tag}_bench_n1.json 2> $out/${tag}_bench_n1.err
python bench.py --impl reference --steps 5 --warmup 1 > $out/${tag}_bench_reference.json 2>> $out/${tag}_bench_n1.err
small="python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-configs --no-xtc"
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $out/${tag}_launches.csv $small > $out/${tag}_launches.log 2>&1
for k in mdmat_tile_kernel pair_stats_kernel prep_slots_kernel; do
  ncu --set full --clock-control none --import-source on -k regex:$k -c 1 -o $out/${tag}_$k $small --no-stage3 > $out/${tag}_$k.log 2>&1
done
for k in gemm3x_pair_kernel pearson_sums_kernel sym_block_matvec_kernel operand_pairs_kernel; do
  ncu --set full --clock-control none --import-source on -k regex:$k -c 1 -o $out/${tag}_$k $small > $out/${tag}_$k.log 2>&1
done
ncu --set full --clock-control none --import-source on -k regex:xtc_decode_kernel -c 1 -o $out/${tag}_xtc_decode_kernel python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-configs --no-stage3 > $out/${tag}_xtc.log 2>&1
for f in $out/${tag}_*.ncu-rep; do
  ncu -i $f --page raw --csv > ${f%.ncu-rep}_raw.csv 2>/dev/null
done
ls -la $out | tail -n 40
