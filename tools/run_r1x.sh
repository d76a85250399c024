for w in 1 2 4; do
  BGPU_LIB_PATH=$PWD/variants/libv_w$w.so python bench.py --steps 2 --warmup 3 --no-q11 --no-cpu --lanes 1 > gpurun_out/r1x_w$w.json 2> gpurun_out/r1x_w$w.err
  python - gpurun_out/r1x_w$w.json <<'PY'
import json,sys
d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
print(sys.argv[1], round(d["value"],3), round(d["ms_per_step"],1), d["config"]["verified_roundtrip"], "k_probe", d["roofline"]["launch_ms"]["k_probe"])
PY
done
