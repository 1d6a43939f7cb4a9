# dev helper for one gpurun call: gathers the evidence that goes under profiles/ (bench line, launch list, ncu full
# capture of the cost / aggregation kernels).  usage: bash tests/collect_profiles.sh <tag> [mode] [what]
# mode = reference (default: the headline configuration) | integer | nofilter (bench.py --mode; secondary lines)
# what = all (default) | noncu
tag=${1:-vX}
mode=${2:-reference}
what=${3:-all}
out=gpurun_out
mkdir -p $out
python bench.py --mode $mode --no-cpu-baseline --no-configs $BENCH_EXTRA > $out/r02_${tag}_bench_line.json 2> $out/r02_${tag}_bench.err || exit 1

[ "$what" = noncu ] && exit 0
# launch list of one bench step (serialised, cold cache: shares only)
ncu --metrics gpu__time_duration.sum --clock-control none -s 300 -c 40 --csv --log-file $out/r02_${tag}_launches_bench_kitti128.csv \
    python bench.py --mode $mode --steps 1 --no-cpu-baseline --no-configs > $out/ncu_launch.log 2>&1
# full capture of the cost / aggregation kernels of one warm frame
ncu --set full --clock-control none --import-source on -k regex:"sgm_sweep|hpair|vfilter|hfilter|hamming" -s 15 -c 5 -o $out/r02_${tag}_agg \
    python bench.py --mode $mode --steps 1 --no-cpu-baseline --no-configs > $out/ncu_full.log 2>&1
tail -2 $out/ncu_full.log
