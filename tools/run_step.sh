# quick check of a change: the GPU tests named by $2 (default: all), then the q5 bench with library defaults
TAG=${1:-r2t}
python -m pytest tests -m gpu -x -q ${2:+-k "$2"} 2>&1 | tail -6 > gpurun_out/${TAG}_tests.log
cat gpurun_out/${TAG}_tests.log
for i in 1 2; do
python bench.py --steps 5 --warmup 3 --no-q11 --no-cpu > gpurun_out/${TAG}_run$i.json 2> gpurun_out/${TAG}_run$i.err
done
for f in gpurun_out/${TAG}_run*.json; do python - "$f" <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print(sys.argv[1], round(d["value"],3), round(d["e2e"]["value"],3), round(d["ms_per_step"],1), d["gpu_launches"], d["config"]["verified_roundtrip"], d["config"]["compressed_ratio"])
except Exception as e: print(sys.argv[1], "ERR", e)
PY
done
