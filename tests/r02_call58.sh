# ncu --set full of the cost / aggregation kernels of one warm frame (v30), after the same command ran clean without ncu.
timeout 300 python bench.py --steps 1 --no-cpu-baseline --no-configs > gpurun_out/pre_ncu.log 2>&1; echo "plain rc=$?"
ncu --set full --clock-control none --import-source on -k regex:"sgm_sweep|hpair|vfilter|hfilter|hamming" -s 15 -c 5 -o gpurun_out/r02_v30_agg \
    python bench.py --steps 1 --no-cpu-baseline --no-configs > gpurun_out/ncu_full.log 2>&1; echo "ncu rc=$?"
tail -2 gpurun_out/ncu_full.log
