# evidence run: GPU tests, plain bench (q5 + q11 + CPU baseline), reference arm, launch lists, full captures
TAG=${1:-r2r}
python bench.py --steps 5 --warmup 3 > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err
python bench.py --impl reference --steps 1 --warmup 1 > gpurun_out/${TAG}_reference.json 2> gpurun_out/${TAG}_reference.err
ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file gpurun_out/${TAG}_q5_launches.csv python bench.py --lanes 1 --steps 1 --warmup 1 --no-cpu --no-verify --no-q11 > gpurun_out/${TAG}_ncu1.log 2>&1
ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file gpurun_out/${TAG}_q11_launches.csv python bench.py --workload config3 --steps 1 --warmup 1 --no-cpu --no-verify > gpurun_out/${TAG}_ncu2.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:"k_sort_hist|k_sort_scan|k_sort_scatter|k_probe_sorted" -s 7 -c 7 -o gpurun_out/${TAG}_match -f python bench.py --lanes 1 --steps 1 --warmup 1 --no-cpu --no-verify --no-q11 > gpurun_out/${TAG}_ncu3.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:"k_write2|k_tilebits2|k_parse" -c 3 -o gpurun_out/${TAG}_backend -f python bench.py --lanes 1 --steps 1 --warmup 0 --no-cpu --no-verify --no-q11 > gpurun_out/${TAG}_ncu4.log 2>&1
ls -la gpurun_out/${TAG}_*
du -sh gpurun_out
bash tools/run_q11lanes.sh ${TAG}
