set -x
python -m pytest tests -m gpu -x -q 2>&1 | tail -6 > gpurun_out/r1t_tests.log
cat gpurun_out/r1t_tests.log
python bench.py --steps 3 --warmup 3 --lanes 1 --no-cpu > gpurun_out/r1t_bench_l1.json 2> gpurun_out/r1t_bench_l1.err
python bench.py --steps 3 --warmup 3 --lanes 4 --group-shards 64 --no-cpu --no-q11 > gpurun_out/r1t_bench_l4g64.json 2> gpurun_out/r1t_bench_l4g64.err
# ncu: full capture of k_zop_dp (64 MiB, q11) with source, and of k_probe_sorted at the bench size (traffic)
ncu --set full --clock-control none --import-source on -k regex:k_zop_dp -s 2 -c 1 -o gpurun_out/r1t_zop -f python bench.py --workload config3 --size-mb 64 --steps 1 --warmup 1 --no-cpu --no-verify > gpurun_out/r1t_ncu_zop.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:"k_probe_sorted|k_write2" -s 2 -c 2 -o gpurun_out/r1t_q5 -f python bench.py --lanes 1 --steps 1 --warmup 1 --no-cpu --no-verify --no-q11 > gpurun_out/r1t_ncu_q5.log 2>&1
ls -la gpurun_out/*.ncu-rep
