TAG=${1:-r2c}
python -m pytest tests -m gpu -x -q 2>&1 | tail -6 > gpurun_out/${TAG}_tests.log
cat gpurun_out/${TAG}_tests.log
for cfg in "1 0" "0 0"; do
  set -- $cfg
  python bench.py --steps 3 --warmup 3 --no-q11 --no-cpu --lanes $1 --group-shards $2 > gpurun_out/${TAG}_l$1_g$2.json 2> gpurun_out/${TAG}_l$1_g$2.err
done
for f in gpurun_out/${TAG}_l*.json; do python - "$f" <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print(sys.argv[1], round(d["value"],3), round(d["e2e"]["value"],3), round(d["ms_per_step"],1), d["gpu_launches"], d["config"]["verified_roundtrip"], d["config"]["compressed_ratio"])
    print("  ", d["roofline"]["launch_ms"])
except Exception as e: print(sys.argv[1], "ERR", e)
PY
done
