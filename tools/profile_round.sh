#!/bin/bash
# One profiling pass of the benched build on a B200 box (run through gpurun from the repository root):
#   bash tools/profile_round.sh r2 [kernel ...]
# Leaves in gpurun_out/: the bench lines, the ncu launch list of the bench command, one `ncu --set full` capture
# per kernel of interest (first launch, at the benched size).  tools/summarize_profiles.py (run afterwards, where
# gpurun_out/ has been brought back) turns them into the summaries kept under profiles/.
tag=${1:-r2}
shift
kernels=${@:-mdmat_tile_kernel pair_stats_kernel prep_slots_kernel gemm3x_pair_kernel pearson_sums_kernel sym_block_matvec_kernel operand_pairs_kernel xtc_walk_kernel xtc_groups_kernel}
out=gpurun_out
mkdir -p $out
python bench.py --steps 20 --warmup 3 > $out/${tag}_bench_n1.json 2> $out/${tag}_bench_n1.err
python bench.py --impl reference --steps 5 --warmup 1 > $out/${tag}_bench_reference.json 2>> $out/${tag}_bench_n1.err
small="python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-configs --no-xtc"
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $out/${tag}_launches.csv $small > $out/${tag}_launches.log 2>&1
for k in $kernels; do
  case $k in
    mdmat_tile_kernel|pair_stats_kernel|prep_slots_kernel|xtc_walk_kernel|xtc_groups_kernel) extra=--no-stage3 ;;
    *) extra= ;;
  esac
  ncu --set full --clock-control none --import-source on -k regex:$k -c 1 -o $out/${tag}_$k $small $extra > $out/${tag}_$k.log 2>&1
done
ls -la $out | tail -n 30
