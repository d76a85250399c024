# final evidence run of the round: plain bench (q5 + q11 + CPU baseline), reference arm, launch lists, full captures
python bench.py --steps 5 --warmup 3 > gpurun_out/r2i_bench.json 2> gpurun_out/r2i_bench.err
python bench.py --impl reference --steps 1 --warmup 1 > gpurun_out/r2i_reference.json 2> gpurun_out/r2i_reference.err
ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file gpurun_out/r2i_q5_launches.csv python bench.py --lanes 1 --steps 1 --warmup 1 --no-cpu --no-verify --no-q11 > gpurun_out/r2i_ncu1.log 2>&1
ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file gpurun_out/r2i_q11_launches.csv python bench.py --workload config3 --steps 1 --warmup 1 --no-cpu --no-verify > gpurun_out/r2i_ncu2.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:"k_sort_hist|k_sort_scan|k_sort_scatter|k_probe_sorted" -s 7 -c 7 -o gpurun_out/r2i_match -f python bench.py --lanes 1 --steps 1 --warmup 1 --no-cpu --no-verify --no-q11 > gpurun_out/r2i_ncu3.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:"k_write2|k_tilebits2|k_split_acc|k_ctx_hist_lit|k_cluster_batch|k_h10_tree_hot" -c 12 -o gpurun_out/r2i_backend -f python bench.py --lanes 1 --steps 1 --warmup 0 --no-cpu --no-verify --no-q11 > gpurun_out/r2i_ncu4.log 2>&1
ls -la gpurun_out/r2i_*
