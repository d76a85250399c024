ncu --set full --clock-control none --import-source on -k regex:"k_write2|k_tilebits2|k_mb_compact" -c 3 -o gpurun_out/r2m_writer -f python bench.py --lanes 1 --steps 1 --warmup 0 --no-cpu --no-verify --no-q11 > gpurun_out/r2m_ncu.log 2>&1
ls -la gpurun_out/r2m_writer.ncu-rep
