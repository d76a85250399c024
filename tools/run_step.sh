python -m pytest tests -m gpu -x -q 2>&1 | tail -6 > gpurun_out/r2a_tests.log
cat gpurun_out/r2a_tests.log
for cfg in "1 0" "0 0"; do
  set -- $cfg
  python bench.py --steps 3 --warmup 3 --no-q11 --no-cpu --lanes $1 --group-shards $2 > gpurun_out/r2a_l$1_g$2.json 2> gpurun_out/r2a_l$1_g$2.err
done
for f in gpurun_out/r2a_l*.json; do python - "$f" <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print(sys.argv[1], round(d["value"],3), round(d["e2e"]["value"],3), round(d["ms_per_step"],1), d["gpu_launches"], d["config"]["verified_roundtrip"], d["config"]["compressed_ratio"])
    print("  ", d["roofline"]["launch_ms"])
except Exception as e: print(sys.argv[1], "ERR", e)
PY
done
ncu --metrics gpu__time_duration.sum --clock-control none -c 1200 --csv --log-file gpurun_out/r2a_q5_launches.csv python bench.py --lanes 1 --steps 1 --warmup 1 --no-cpu --no-verify --no-q11 > gpurun_out/r2a_ncu.log 2>&1
