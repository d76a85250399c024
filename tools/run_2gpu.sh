TAG=${1:-r2x}
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 5 --warmup 3 > gpurun_out/${TAG}_bench_2gpu.json 2> gpurun_out/${TAG}_bench_2gpu.err
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --impl reference --gpus 2 --steps 1 --warmup 1 > gpurun_out/${TAG}_ref_2gpu.json 2> gpurun_out/${TAG}_ref_2gpu.err
tail -c 300 gpurun_out/${TAG}_bench_2gpu.err
python - gpurun_out/${TAG}_bench_2gpu.json <<'PY'
import json,sys
d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
print(d["n_gpus"], round(d["value"],3), round(d["e2e"]["value"],3), round(d["ms_per_step"],1), d["config"]["verified_roundtrip"], d["q11"]["value"], d["q11"]["e2e"]["value"])
PY
tail -c 400 gpurun_out/${TAG}_ref_2gpu.json
