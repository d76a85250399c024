python -m pytest tests/test_sharded.py -m gpu -x -q 2>&1 | tail -8 > gpurun_out/r1s_tests.log
for cfg in "1 32" "2 32" "3 32" "4 32" "3 16" "3 64" "6 16"; do
  set -- $cfg
  python bench.py --steps 3 --warmup 3 --no-q11 --no-cpu --no-verify --lanes $1 --group-shards $2 > gpurun_out/r1s_l$1_g$2.json 2> gpurun_out/r1s_l$1_g$2.err
done
cat gpurun_out/r1s_tests.log
for f in gpurun_out/r1s_l*.json; do python - "$f" <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print(sys.argv[1], round(d["value"],3), round(d["e2e"]["value"],3), d["ms_per_step"], d["gpu_launches"])
except Exception as e: print(sys.argv[1], "ERR", e)
PY
done
