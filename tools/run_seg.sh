for kb in 32 40 48 50 56 64; do
  python bench.py --workload config3 --steps 2 --warmup 3 --no-cpu --segment-kb $kb > gpurun_out/r2l_q11_s$kb.json 2> gpurun_out/r2l_q11_s$kb.err
  python - gpurun_out/r2l_q11_s$kb.json <<'PY'
import json,sys
d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
print(sys.argv[1], round(d["value"],4), round(d["ms_per_step"],1), d["config"]["verified_roundtrip"], d["config"]["compressed_ratio"], "dp", d["roofline"]["launch_ms"]["k_zop_dp1"], d["roofline"]["launch_ms"]["k_zop_dp2"])
PY
done
for kb in 8 12 20 24 36; do
  python bench.py --steps 2 --warmup 3 --no-cpu --no-q11 --lanes 1 --segment-kb $kb > gpurun_out/r2l_q5_s$kb.json 2> gpurun_out/r2l_q5_s$kb.err
  python - gpurun_out/r2l_q5_s$kb.json <<'PY'
import json,sys
d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
print(sys.argv[1], round(d["value"],4), round(d["ms_per_step"],1), d["config"]["verified_roundtrip"], d["config"]["compressed_ratio"], "parse", d["roofline"]["launch_ms"]["k_parse"])
PY
done
