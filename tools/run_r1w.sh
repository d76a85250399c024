python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "match_candidates or commands_bit_exact or stream" 2>&1 | tail -4 > gpurun_out/r1w_tests.log
cat gpurun_out/r1w_tests.log
python bench.py --steps 3 --warmup 3 --no-q11 --no-cpu --lanes 1 > gpurun_out/r1w_l1.json 2> gpurun_out/r1w_l1.err
python - <<'PY'
import json
d=json.loads(open("gpurun_out/r1w_l1.json").read().strip().splitlines()[-1])
print(round(d["value"],3), round(d["e2e"]["value"],3), round(d["ms_per_step"],1), d["config"]["verified_roundtrip"], d["config"]["compressed_ratio"])
print("  ", d["roofline"]["launch_ms"])
PY
