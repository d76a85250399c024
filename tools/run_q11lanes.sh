# q11 (config 3) with lanes: does the H10 stage of one group hide behind the Zopfli passes of another?
TAG=${1:-r2s}
for cfg in "2 32" "4 16" "2 16" "3 22" "8 8"; do
  set -- $cfg
  python bench.py --workload config3 --steps 2 --warmup 3 --no-cpu --lanes $1 --group-shards $2 > gpurun_out/${TAG}_q11_l$1_g$2.json 2> gpurun_out/${TAG}_q11_l$1_g$2.err
done
for f in gpurun_out/${TAG}_q11_l*.json; do python - "$f" <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print(sys.argv[1], round(d["value"],4), round(d["e2e"]["value"],4), round(d["ms_per_step"],1), d["gpu_launches"], d["config"]["verified_roundtrip"], d["config"]["compressed_ratio"])
except Exception as e: print(sys.argv[1], "ERR", e)
PY
done
