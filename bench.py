#!/usr/bin/env python
"""bench.py — input GB/s of the Brotli encoder hot path (BASELINE.json metric).

A step = one pass of the whole encoder (match finding -> parse -> block split / clustering -> histograms -> Huffman ->
bit writer) over one batch of synthetic input.  Default workload: BASELINE config 2 — q5, GENERIC, lgwin 22 over
1 GiB of synthetic mixed text/binary, 256 independent 4 MiB shards, on one B200.  With --gpus N (torchrun) each rank
encodes its own 1 GiB (weak scaling, no collective on the data path).

  value         device-resident throughput (input already in HBM), whole job
  e2e           the same through bgpu_encode with pinned HOST buffers (H2D + kernels + D2H + stitch in the timed region)
  e2e_pageable  the same call on ordinary (pageable) host memory, as a Node Uint8Array is
  q11           BASELINE config 3 (q11 TEXT lgwin 24, 256 MiB), measured the same way in the same run
  cpu_baseline / cpu_lines   the reference's CPU encoder (oracle/ C++ restatement; no JS runtime on the box):
                single thread and all host threads (4 shards per thread); Google's C encoder as a context line
  size_delta    compressed size vs the restatement, per 4 MiB shard type
  extra         config 1 (enc-ttf, q11, one call: latency), config 5 (FONT q11 batch of WOFF2-glyf-like blobs through
                bgpu_encode_batch), config 4 (q9 lgwin 24: ONE input over all N GPUs through one bgpu_encode call of a
                bgpu_ctx_create_multi context, one stitched stream, decoded and compared)
  --impl reference   the reference arm: the restatement on all host threads, bounded sample per step
"""
import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

UNIT = "GB/s"

# BASELINE.json configs that fit one GPU.  size_mb is per GPU (weak scaling: config 4 is 8 GiB over 8 GPUs).
WORKLOADS = {
    "config2": dict(quality=5, lgwin=22, mode=0, size_mb=1024, gen="mixed", metric="input_GBps_q5",
                    label="q5 GENERIC lgwin22, {mb} MiB synthetic mixed text/binary per GPU (BASELINE config 2), {shards} independent 4 MiB shards, H5 hash-chain matcher depth 16"),
    "config3": dict(quality=11, lgwin=24, mode=1, size_mb=256, gen="text", metric="input_GBps_q11",
                    label="q11 TEXT lgwin24, {mb} MiB synthetic English-like text per GPU (BASELINE config 3), {shards} independent 4 MiB shards, H10 binary-tree matcher + 2-pass Zopfli"),
    "config4": dict(quality=9, lgwin=24, mode=0, size_mb=1024, gen="mixed", metric="input_GBps_q9",
                    label="q9 GENERIC lgwin24, {mb} MiB synthetic mixed corpus per GPU (BASELINE config 4: 8 GiB over 8 GPUs), {shards} independent 4 MiB shards, H5 hash-chain matcher depth 256"),
}


def _text_segment(i, seed_offset, n):
    from brotli_lib_b200 import workloads as W
    return W.text(n, 0x7E570002 + i + seed_offset)


def gen_data(kind, nbytes, seed_offset):
    """Synthetic workload, one 4 MiB segment per shard, generated with a process pool (numpy generators are single-threaded)."""
    import numpy as np
    from multiprocessing import get_context
    from brotli_lib_b200 import workloads as W
    fn = W.mixed_segment if kind == "mixed" else _text_segment
    nseg = (nbytes + W.SEG - 1) // W.SEG
    out = np.empty(nbytes, dtype=np.uint8)
    procs = max(1, min(os.cpu_count() or 1, 32, nseg))
    args = [(i, seed_offset, min(W.SEG, nbytes - i * W.SEG)) for i in range(nseg)]
    if procs == 1 or nseg < 4:
        segs = [fn(*a) for a in args]
    else:
        with get_context("fork").Pool(procs) as pool:
            segs = pool.starmap(fn, args, chunksize=1)
    for i, seg in enumerate(segs):
        out[i * W.SEG:i * W.SEG + seg.size] = seg
    return out


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md)."""
    Q = "clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index):
        self.rows, self.proc, self.index = [], None, index

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "200"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], None, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[0]))
                mx = float(r[1])
                for k, nme in enumerate(names):
                    if r[2 + k].lower().startswith("active"):
                        reasons.add(nme)
            except Exception:
                pass
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": mx, "reasons": sorted(reasons), "samples": len(sm)}


def host_threads():
    try:
        return len(os.sched_getaffinity(0))
    except Exception:
        return os.cpu_count() or 1


def sample_shards(wl, count, shard_bytes):
    """`count` independent shards of the workload's generator, shard i of segment type i mod 8 for the mixed workloads
    (so any multiple of 8 shards has the workload's type mix), each the first shard_bytes of its 4 MiB segment."""
    import numpy as np
    from brotli_lib_b200 import workloads as W
    fn = W.mixed_segment if wl["gen"] == "mixed" else _text_segment
    return np.concatenate([fn(i, 0, shard_bytes) for i in range(count)])


def time_oracle(sample, quality, lgwin, mode, threads, shard_bytes=1 << 22):
    """The oracle (C++ restatement of brotli-lib's encoder) over `sample` cut into independent shards, `threads` host
    threads pulling shards from one queue.  Returns (GB/s, per-shard compressed sizes, seconds)."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import _oracle as O
    t = time.perf_counter()
    sizes = O.encode_shards_sizes(sample, shard_bytes, quality, lgwin, mode, threads)
    dt = time.perf_counter() - t
    return sample.size / dt / 1e9, sizes, dt


def cpu_lines(wl, threads):
    """line A (the restatement) single-thread and on every host thread with 4 shards per thread,
    line B (the real brotliEncode under Node) when a `node` binary exists, line C (Google's C encoder, context only)."""
    import shutil
    q, lgwin, mode = wl["quality"], wl["lgwin"], wl["mode"]
    shard = (1 << 20) if q < 10 else (1 << 18)
    one = sample_shards(wl, 8 if wl["gen"] == "mixed" else 2, shard)
    g1, _, t1 = time_oracle(one, q, lgwin, mode, 1, shard)
    n_many = 4 * threads
    many = sample_shards(wl, n_many, shard)
    gN, _, tN = time_oracle(many, q, lgwin, mode, threads, shard)
    lines = {
        "A_single_thread": {"value": g1, "unit": UNIT, "threads": 1, "kind": "port",
                            "sample": f"{one.size >> 20} MiB = {one.size // shard} shards of {shard >> 10} KiB" + (" (one of each of the 8 segment types)" if wl["gen"] == "mixed" else ""), "seconds": round(t1, 1)},
        "A_all_threads": {"value": gN, "unit": UNIT, "threads": min(threads, n_many), "kind": "port",
                          "sample": f"{many.size >> 20} MiB = {n_many} shards of {shard >> 10} KiB (4 per host thread, one work queue)", "seconds": round(tN, 1)},
    }
    node = shutil.which("node")
    lines["B_node"] = {"node": node, "note": "the reference's own TypeScript needs a resolver hook and --experimental-transform-types (SURVEY.md section 8b); not wired"} if node \
        else "node: unavailable on this box (command -v node fails; deno / bun / qjs absent too: profiles/r4a_gpu_box_js_runtime_probe.txt)"
    try:
        sys.path.insert(0, os.path.join(ROOT, "tests"))
        import _oracle as O
        t = time.perf_counter()
        csz = sum(len(O.ref_encode(one[o:o + shard], q, lgwin, 0)) for o in range(0, one.size, shard))
        dt = time.perf_counter() - t
        lines["C_google_c_encoder"] = {"value": one.size / dt / 1e9, "unit": UNIT, "threads": 1, "library": "system libbrotlienc (1.1.0)", "sample": lines["A_single_thread"]["sample"],
                                       "compressed_ratio": round(csz / one.size, 4), "note": "context only: a different encoder (lazy matching, static dictionary), never a parity target"}
    except Exception as e:   # pragma: no cover
        lines["C_google_c_encoder"] = f"unavailable: {e}"
    return lines


def size_sample(wl):
    """The shards of the ratio accounting: real 4 MiB shards, one of each segment type."""
    import numpy as np
    from brotli_lib_b200 import workloads as W
    if wl["gen"] == "mixed":
        return np.concatenate([W.mixed_segment(i, 0) for i in range(8)]), "one 4 MiB shard of each of the 8 segment types"
    return np.concatenate([_text_segment(i, 0, W.SEG) for i in range(8)]), "8 text shards of 4 MiB"


def size_delta(sample, ref_sizes, quality, lgwin, mode, device, tune=None):
    """SURVEY.md section 8(d) ratio accounting: the shards of the CPU sample, each as its own stream, through the
    GPU library vs the reference restatement at the same quality / lgwin: delta = (M_gpu - M_ref) / M_ref."""
    import brotli_lib_b200 as B
    ctx = B.Context(device)
    if tune:
        ctx.set_tuning(**tune)
    items = [sample[o:o + (1 << 22)] for o in range(0, sample.size, 1 << 22)]
    gpu = [len(s) for s in ctx.encode_batch(items, quality, lgwin, mode)]
    ctx.close()
    ref = [int(x) for x in ref_sizes]
    d = [(g - r) / r for g, r in zip(gpu, ref)]
    return {"shards": len(d), "total": round((sum(gpu) - sum(ref)) / sum(ref), 5), "max": round(max(d), 5), "mean": round(sum(d) / len(d), 5),
            "gate_total": 0.005, "per_shard": [round(x, 5) for x in d], "reference": "oracle/ restatement, same shards as independent streams"}


# ALGORITHMIC (compulsory) HBM bytes of each stage per launch: SURVEY.md §8(d) / DESIGN.md §4.
# n = input bytes, M = output bytes, S = sum of H10 matches.  Command-record terms (16 C) are left out
# (C is not known on the host), so the figures are lower bounds.  A stage groups the kernels that
# §8(d) judges as one unit (K1+K2+K3 = key sort + probe; K4 = H10 sort + short scan + tree).
STAGE_OF = {"k_sort_pass0": "match_finding", "k_sort_pass1": "match_finding", "k_sort_pass2": "match_finding", "k_probe": "match_finding",
            "k_h10_sort": "h10_matches", "k_h10_short": "h10_matches", "k_h10_tree": "h10_matches", "k_h10_tree_hot": "h10_matches",
            "k_parse": "parse", "k_stitch": "parse",
            "k_zop_model1": "zopfli", "k_zop_dp1": "zopfli", "k_zop_model2": "zopfli", "k_zop_dp2": "zopfli",
            "k_hist": "histograms", "k_huff": "huffman", "k_tilebits": "bit_writer", "k_offsets": "bit_writer", "k_write": "bit_writer",
            "k_mb_compact": "histograms", "k_ctx_hist": "histograms", "k_split": "block_split", "k_cluster_blocks": "clustering",
            "k_cluster_ctx": "clustering", "k_huff_hdr": "huffman", "k_write2": "bit_writer"}
STAGE_KERNELS = {"match_finding": "k_sort_hist/scan/scatter x2 + k_probe_sorted", "h10_matches": "k_sort_* x3 + k_h10_short/runs/tree/tree_hot",
                 "parse": "k_parse<8> + k_stitch", "zopfli": "k_zop_model/litcost/dp (+ k_stitch, k_hist between passes)",
                 "histograms": "k_mb_counts/scan/compact + k_ctx_hist", "block_split": "k_split_table/acc/dp/remap/hist, k_split_blocks, k_block_hist",
                 "clustering": "k_cluster_batch/gather/final/remap/reindex", "huffman": "k_huff_mb + k_mb_header",
                 "bit_writer": "k_tile_model + k_tilebits2 + k_offsets2 + k_write2"}
ALGO_BYTES = {
    "match_finding": lambda n, M, S: 9.0 * n,                # K1+K2+K3: N + 8N
    "h10_matches": lambda n, M, S: n + 8.0 * S,              # K4: N + 8 * matches
    "parse": lambda n, M, S: 9.0 * n,                        # K5: N + 8N (+16C)
    "zopfli": lambda n, M, S: 2 * (9.0 * n + 8.0 * S),       # K6, two passes: N + 8N + 8 * matches (+16C) each
    "histograms": lambda n, M, S: 1.0 * n,                   # K7: L (+16C + 4AH), L <= N
    "block_split": lambda n, M, S: 1.0 * n,                  # K8: the symbol streams once
    "clustering": lambda n, M, S: 0.0,                       # K9: 4A(H_in + H_out): from the device counters (BGPU_F_MODEL_STATS), see measure()
    "huffman": lambda n, M, S: 0.0,                          # K10: 7AH: from the device counters
    "bit_writer": lambda n, M, S: n + M,                     # K11: L + M (+16C + 3AH)
}


def stage_of(name, quality):
    if name == "k_stitch" and quality >= 10:
        return "zopfli"      # at q10/11 the stitch joins Zopfli segments; there is no greedy parse stage
    return STAGE_OF.get(name)


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured"
        except Exception:
            pass
    return 6650.0, "fallback"


def ncu_traffic(stage, n):
    """DRAM bytes per step of the stage's kernels from the committed `ncu --set full` capture
    (profiles/traffic.json: dram__bytes_read.sum + dram__bytes_write.sum), only for the input size it was taken at."""
    try:
        t = json.load(open(os.path.join(ROOT, "profiles", "traffic.json"))).get(stage)
        if t and int(t["input_bytes"]) == int(n):
            return int(t["dram_bytes"])
    except Exception:
        pass
    return None


def run_reference(args, wl, rank, world):
    """The reference's CPU implementation of the path (the oracle restatement: no JS runtime exists on the box) with all
    the host threads, 4 shards per thread from one work queue.  A step is a bounded sample of the arm's workload — the
    same generators, quality, window and mode, shard size chosen from a calibration pass so that the K + W steps take
    about 100 s — because the full workload would take about a minute per step on these cores."""
    if rank != 0:
        return
    threads = host_threads()
    q, lgwin, mode = wl["quality"], wl["lgwin"], wl["mode"]
    n_shards = 4 * threads
    probe = sample_shards(wl, n_shards, 1 << 16)
    rate, _, _ = time_oracle(probe, q, lgwin, mode, threads, 1 << 16)
    budget = max(1.0, min(20.0, 100.0 / max(1, args.steps + args.warmup)))
    shard = 1 << 16
    while shard < (1 << 22) and 2 * shard * n_shards / (rate * 1e9) <= budget:
        shard *= 2
    sample = sample_shards(wl, n_shards, shard)
    what = f"{sample.size >> 20} MiB per step = {n_shards} independent shards of {shard >> 10} KiB (4 per host thread, types cycling as in the workload)"
    for _ in range(args.warmup):
        time_oracle(sample, q, lgwin, mode, threads, shard)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        time_oracle(sample, q, lgwin, mode, threads, shard)
    wall = time.perf_counter() - t0
    value = sample.size * args.steps / wall / 1e9
    line = {
        "impl": "reference", "metric": wl["metric"], "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": wall / args.steps * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u8",
        "data": "synthetic",
        "config": {"workload": wl["label"].format(mb=wl["size_mb"], shards=wl["size_mb"] // 4)},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": min(threads, n_shards), "kind": "port",
                         "sample": "oracle/ C++ restatement of brotli-lib (no JS runtime on this box), " + what},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line))

def measure(args, wl, ctx, torch, dist, dev, rank, world, local_rank, steps, warmup, verify, tune):
    """Device-resident and end-to-end timing of one workload on this rank; returns the JSON object (rank 0) or None."""
    import numpy as np
    n = wl["size_mb"] << 20
    q, lgwin, mode = wl["quality"], wl["lgwin"], wl["mode"]
    t_gen = time.perf_counter()
    host = gen_data(wl["gen"], n, rank * 0x100)
    t_gen = time.perf_counter() - t_gen
    h_in = torch.empty(n, dtype=torch.uint8, pin_memory=True)
    h_in.numpy()[:] = host
    d_in = torch.zeros(n + 4096, dtype=torch.uint8, device=dev)
    d_in[:n].copy_(h_in, non_blocking=False)
    L = ctx._L
    cap = n + (n >> 3) + (1 << 20)
    h_out = torch.empty(cap, dtype=torch.uint8, pin_memory=True)

    def barrier():
        torch.cuda.synchronize()
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    def step_device():
        return ctx.encode_device(d_in.data_ptr(), n, q, lgwin, mode)

    def step_e2e():
        ln = C.c_size_t(0)
        rc = L.bgpu_encode(ctx._h, h_in.data_ptr(), n, q, lgwin, mode, h_out.data_ptr(), cap, C.byref(ln))
        if rc:
            raise RuntimeError(f"bgpu_encode rc={rc}: {L.bgpu_last_cuda_error().decode()}")
        return ln.value

    # all work of the context goes to torch's current stream so that torch.cuda.Event brackets it
    stream = torch.cuda.current_stream()
    ctx.set_stream(stream.cuda_stream)
    # ---- device-resident timing (CUDA events on the launching stream)
    for _ in range(warmup):
        step_device()
    sampler = ClockSampler(local_rank)
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    sampler.start()
    ev0.record(stream)
    out_size = 0
    for _ in range(steps):
        out_size, kms, _ = step_device()
    ev1.record(stream)
    barrier()
    t_dev = ev0.elapsed_time(ev1) * 1e-3
    clocks = sampler.stop()
    launches = ctx.launch_count() * steps
    n_matches = ctx.counter(0)
    # ---- end to end (pinned host buffers; H2D + kernels + D2H + stitch inside the timed region)
    step_e2e()
    barrier()
    ev0.record(stream)
    for _ in range(steps):
        e2e_size = step_e2e()
    ev1.record(stream)
    barrier()
    t_e2e = ev0.elapsed_time(ev1) * 1e-3
    # ---- the same call on pageable host memory (what a Node Uint8Array is): the driver stages the copies
    import numpy as np
    pg_out = np.empty(cap, dtype=np.uint8)

    def step_e2e_pageable():
        ln = C.c_size_t(0)
        rc = L.bgpu_encode(ctx._h, host.ctypes.data, n, q, lgwin, mode, pg_out.ctypes.data, cap, C.byref(ln))
        if rc:
            raise RuntimeError(f"bgpu_encode rc={rc}: {L.bgpu_last_cuda_error().decode()}")
        return ln.value

    step_e2e_pageable()
    pg_steps = max(1, min(steps, 3))
    barrier()
    ev0.record(stream)
    for _ in range(pg_steps):
        step_e2e_pageable()
    ev1.record(stream)
    barrier()
    t_pg = ev0.elapsed_time(ev1) * 1e-3 / pg_steps
    del pg_out
    # ---- per-kernel durations: in the timed regions above the groups of shards run on several CUDA
    # streams (lanes) and kernels of different groups overlap, so a kernel's own duration is taken from
    # one-lane (in-order) passes of the same work, CUDA events around every launch family
    ctx.set_tuning(**dict(tune, lanes=1, flags=(tune.get("flags") or 0) | 8))   # 8 = BGPU_F_MODEL_STATS: table bytes of the modelling stages
    step_device()
    stage_ms = {}
    stage_steps = max(1, min(steps, 2))
    barrier()
    ev0.record(stream)
    for _ in range(stage_steps):
        step_device()
        for name, ms in ctx.timings():
            stage_ms[name] = stage_ms.get(name, 0.0) + ms
        model_bytes = [ctx.counter(1) + ctx.counter(2), ctx.counter(3)]
    ev1.record(stream)
    barrier()
    t_one_lane = ev0.elapsed_time(ev1) * 1e-3 / stage_steps
    # ---- on-device verification of the device-resident result at full size (BGPU_F_VERIFY: every shard decoded by the
    # library's own device decoder and compared with the input in HBM; no device->host copy, no CPU decode)
    device_verified, t_verify = None, None
    if verify:
        ctx.set_tuning(**dict(tune, flags=(tune.get("flags") or 0) | 16))
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        try:
            step_device()
            device_verified = ctx.counter(4) == 0
        except Exception as e:
            device_verified = f"FAILED: {e}"
        torch.cuda.synchronize()
        t_verify = time.perf_counter() - t0
    ctx.set_tuning(**tune)
    times = torch.tensor([t_dev, t_e2e, t_pg], dtype=torch.float64, device=dev)
    if dist is not None:
        dist.all_reduce(times, op=dist.ReduceOp.MAX)
    t_dev, t_e2e, t_pg = times.tolist()
    # ---- verification outside the timed region (rank 0): the stream decodes to the input
    verified = None
    if rank == 0 and verify:
        sys.path.insert(0, os.path.join(ROOT, "tests"))
        import _oracle as O
        verified = O.decode(bytes(h_out.numpy()[:e2e_size]), n) == host.tobytes()
    del d_in, h_in, h_out
    if rank != 0:
        return None
    M = out_size
    peak, which = peaks()
    per_kernel, per_stage = {}, {}
    for name, ms in stage_ms.items():
        avg = ms / stage_steps
        per_kernel[name] = round(per_kernel.get(name, 0.0) + avg, 3)
        st = stage_of(name, q)
        if st:
            per_stage[st] = per_stage.get(st, 0.0) + avg
    stages = {}
    for st, avg in per_stage.items():
        ab = ALGO_BYTES[st](float(n), float(M), float(n_matches))
        if st == "clustering":
            ab = float(model_bytes[0])
        elif st == "huffman":
            ab = float(model_bytes[1])
        stages[st] = {"ms": round(avg, 3), "algo_bytes": int(ab), "algo_GBps": round(ab / (avg * 1e-3) / 1e9, 1) if avg > 0 else None,
                      "frac": round(ab / (avg * 1e-3) / 1e9 / peak, 4) if avg > 0 else None}
    top = max(stages, key=lambda k: stages[k]["ms"])
    achieved = stages[top]["algo_GBps"] or 0.0
    return {
        "metric": wl["metric"], "value": n * world * steps / t_dev / 1e9, "unit": UNIT, "n_gpus": world, "steps": steps,
        "warmup": warmup, "ms_per_step": t_dev / steps * 1e3, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "u8", "data": "synthetic",
        "config": {"workload": wl["label"].format(mb=wl["size_mb"], shards=(n + (1 << 22) - 1) >> 22),
                   "timing": "CUDA events on the launching stream (lanes fork from / join to it), max over ranks",
                   "lanes": tune.get("lanes") or "library default (4 lanes, groups of shards/4 clamped to 16..64, at every quality >= 1; at q10/11 a lane defers its H10 capacity check to the host loop)",
                   "stage_timing": f"roofline.stages / launch_ms: one-lane in-order passes after the timed region ({round(t_one_lane * 1e3, 1)} ms per step)",
                   "l2": f"input ({wl['size_mb']} MiB) and every intermediate array exceed the 126 MB L2; no flush needed",
                   "compressed_ratio": round(M / n, 4), "verified_roundtrip": verified, "verified_on_device": device_verified,
                   "verify_on_device_s": round(t_verify, 2) if t_verify else None, "datagen_s": round(t_gen, 1),
                   "h10_matches_per_position": round(n_matches / n, 3) if n_matches else None},
        "e2e": {"value": n * world * steps / t_e2e / 1e9, "unit": UNIT, "h2d_bytes_per_step": n, "d2h_bytes_per_step": int(e2e_size), "host_memory": "pinned"},
        "e2e_pageable": {"value": n * world / t_pg / 1e9, "unit": UNIT, "steps": pg_steps, "host_memory": "pageable (numpy arrays, as a Node Uint8Array)",
                         "h2d_bytes_per_step": n, "d2h_bytes_per_step": int(e2e_size)},
        "gpu_launches": launches,
        "clocks": clocks,
        "roofline": {"bound": "hbm", "kernel": top, "kernels": STAGE_KERNELS.get(top, top), "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": round(achieved / peak, 4),
                     "traffic": ncu_traffic(top, n), "peak_source": which, "whole_step_frac": round((n + M) * steps / t_dev / 1e9 / peak, 4),
                     "stages": stages, "launch_ms": per_kernel},
    }


def cpu_block(wl, threads, device, tune):
    """cpu_baseline (all host threads), the three BASELINE.md lines, and the per-shard size accounting."""
    lines = cpu_lines(wl, threads)
    a = lines["A_all_threads"]
    base = {"value": a["value"], "unit": UNIT, "cores": a["threads"], "kind": "port",
            "sample": f"oracle/ C++ restatement of brotli-lib (no JS runtime on this box), {a['sample']}, {a['seconds']} s; single thread: {lines['A_single_thread']['value']:.5f} GB/s"}
    sample, what = size_sample(wl)
    _, sizes, _ = time_oracle(sample, wl["quality"], wl["lgwin"], wl["mode"], threads)
    sd = size_delta(sample, sizes, wl["quality"], wl["lgwin"], wl["mode"], device, tune)
    sd["sample"] = what
    return base, lines, sd


def extra_config1(ctx, torch):
    """BASELINE config 1: brotliEncode q11 lgwin22 GENERIC on bench/fixtures/enc-ttf.bin (312 668 bytes): ONE shard, so the
    figure is the latency of one call (host buffers in and out), not a throughput the GPU is sized for."""
    path = os.path.join(ROOT, "tests", "golden", "data", "enc-ttf.bin")
    data = open(path, "rb").read()
    for _ in range(2):
        out = ctx.encode(data, 11, 22, 0)
    ts = []
    for _ in range(5):
        torch.cuda.synchronize()
        t = time.perf_counter()
        out = ctx.encode(data, 11, 22, 0)
        ts.append(time.perf_counter() - t)
    ts.sort()
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import _oracle as O
    ok = O.decode(out, len(data)) == data
    t = time.perf_counter()
    ref = O.encode(data, 11, 22, 0)
    t_cpu = time.perf_counter() - t
    return {"workload": "brotliEncode q11 lgwin22 GENERIC on bench/fixtures/enc-ttf.bin (312668 bytes), one call = one shard", "ms_per_call_median": round(ts[2] * 1e3, 2),
            "ms_per_call_min": round(ts[0] * 1e3, 2), "input_GBps": len(data) / ts[2] / 1e9, "compressed_bytes": len(out), "oracle_bytes": len(ref),
            "size_delta": round((len(out) - len(ref)) / len(ref), 5), "verified_roundtrip": ok, "launches": ctx.launch_count(),
            "stages_ms": {k: round(v, 2) for k, v in ctx.timings() if v >= 0.05},
            "cpu_restatement_ms_single_thread": round(t_cpu * 1e3, 1)}


def extra_config5(ctx, torch, n_blobs):
    """BASELINE config 5: FONT mode q11 batch of WOFF2-transformed glyf-like blobs (4-300 KB each) through ONE
    bgpu_encode_batch call; every blob is its own stream."""
    from brotli_lib_b200 import workloads as W
    blobs = [b.tobytes() for b in W.font_blobs(n_blobs)]
    total = sum(len(b) for b in blobs)
    ctx.encode_batch(blobs[: max(8, n_blobs // 20)], 11, 22, 2)
    torch.cuda.synchronize()
    times = []
    for _ in range(2):   # the first full-size call also grows the context's device buffers (cudaMalloc); both are reported
        t = time.perf_counter()
        outs = ctx.encode_batch(blobs, 11, 22, 2)
        times.append(time.perf_counter() - t)
    dt = min(times)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import _oracle as O
    check = list(range(0, n_blobs, max(1, n_blobs // 50)))
    ok = all(O.decode(outs[i], len(blobs[i])) == blobs[i] for i in check)
    return {"workload": f"FONT q11 lgwin22 batch of {n_blobs} WOFF2-glyf-like blobs, log-uniform 4-300 KiB (BASELINE config 5 is 10000 blobs; --config5-blobs sets the count), one bgpu_encode_batch call, host buffers",
            "blobs": n_blobs, "input_bytes": total, "seconds": round(dt, 3), "seconds_all": [round(x, 3) for x in times], "input_GBps": total / dt / 1e9, "blobs_per_s": round(n_blobs / dt, 1),
            "compressed_ratio": round(sum(len(o) for o in outs) / total, 4), "verified_roundtrip": ok, "verified_blobs": len(check), "launches": ctx.launch_count()}


def extra_config4(torch, devices, mb_per_gpu, steps):
    """BASELINE config 4 as north_star describes it: ONE input, cut into one part per GPU, each part compressed on its
    own GPU, stitched on the host into a single valid stream — one bgpu_encode call on a bgpu_ctx_create_multi context
    (one host thread per device), host buffers in and out.  Wall-clock of the call (it returns when the stream is complete)."""
    import numpy as np
    import brotli_lib_b200 as B
    wl = WORKLOADS["config4"]
    n = (mb_per_gpu << 20) * len(devices)
    host = gen_data(wl["gen"], n, 0)
    ctx = B.Context(devices=devices)
    L = ctx._L
    cap = int(L.bgpu_max_compressed_size(n))
    out = np.empty(cap, dtype=np.uint8)
    ln = C.c_size_t(0)

    def call():
        rc = L.bgpu_encode(ctx._h, host.ctypes.data, n, wl["quality"], wl["lgwin"], wl["mode"], out.ctypes.data, cap, C.byref(ln))
        if rc:
            raise RuntimeError(f"bgpu_encode (multi) rc={rc}: {L.bgpu_last_cuda_error().decode()}")
    call()
    ts = []
    for _ in range(max(1, steps)):
        t = time.perf_counter()
        call()
        ts.append(time.perf_counter() - t)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import _oracle as O
    t = time.perf_counter()
    ok = O.decode(out[: ln.value].tobytes(), n) == host.tobytes()
    t_dec = time.perf_counter() - t
    launches = ctx.launch_count()
    ctx.close()
    best = min(ts)
    return {"workload": f"q9 GENERIC lgwin24, ONE {n >> 20} MiB input ({mb_per_gpu} MiB per GPU; BASELINE config 4 is 8 GiB over 8 GPUs: --config4-mb-per-gpu 1024) over {len(devices)} GPU(s), one stitched stream",
            "n_gpus": len(devices), "input_bytes": n, "e2e_GBps": n / best / 1e9, "seconds_best": round(best, 3), "seconds_all": [round(x, 3) for x in ts],
            "host_memory": "pageable", "compressed_ratio": round(ln.value / n, 4), "verified_roundtrip": ok, "decode_seconds_cpu": round(t_dec, 1), "launches": launches,
            "limiter": "one pageable host buffer feeds every device (driver-staged H2D per worker thread) and each part crosses the host once more in the final memcpy; DESIGN.md section 5"}


def run_gpu(args, wl, rank, world, local_rank):
    import torch
    import brotli_lib_b200 as B
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
        # a host-side group for the final wait: while rank 0 drives ALL GPUs through one multi-device context (extra.config4)
        # the other ranks must not sit in an NCCL barrier — its kernel spins on their GPU and time-slices with rank 0's work
        # there (measured: 2 GPUs 0.353 s instead of 0.168 s for the same call)
        host_group = dist.new_group(backend="gloo")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    ctx = B.Context(local_rank)
    tune = {"segment_bytes": args.segment_kb << 10, "lanes": args.lanes, "group_shards": args.group_shards, "flags": args.flags}
    ctx.set_tuning(**tune)
    warm = max(args.warmup, 3)
    line = measure(args, wl, ctx, torch, dist, dev, rank, world, local_rank, args.steps, warm, not args.no_verify, tune)
    ctx.close()
    torch.cuda.empty_cache()
    q11 = None
    tune11 = {"lanes": args.q11_lanes, "group_shards": args.q11_group_shards}
    if args.workload == "config2" and not args.no_q11:
        # the metric names q5 AND q11: BASELINE config 3 measured in the same run, reported beside the headline
        ctx = B.Context(local_rank)
        w3 = dict(WORKLOADS["config3"])
        if args.q11_size_mb:
            w3["size_mb"] = args.q11_size_mb
        ctx.set_tuning(**tune11)
        q11 = measure(args, w3, ctx, torch, dist, dev, rank, world, local_rank, min(args.steps, 3), 3, not args.no_verify, tune11)
        ctx.close()
        torch.cuda.empty_cache()
    torch.cuda.synchronize()
    if dist is not None:
        dist.barrier(group=host_group)     # every rank's GPU is idle from here on; rank 0 may use all of them
    if rank == 0:
        if q11 is not None:
            line["q11"] = {k: q11[k] for k in ("metric", "value", "unit", "steps", "warmup", "ms_per_step", "config", "e2e", "e2e_pageable", "gpu_launches", "roofline")}
        if world == 1 and not args.no_cpu:
            threads = host_threads()
            line["cpu_baseline"], line["cpu_lines"], line["size_delta"] = cpu_block(wl, threads, local_rank, tune)
            if q11 is not None:
                line["q11"]["cpu_baseline"], line["q11"]["cpu_lines"], line["q11"]["size_delta"] = cpu_block(WORKLOADS["config3"], threads, local_rank, tune11)
        if not args.no_extra and args.workload == "config2":
            extra = {}
            try:
                if world == 1:
                    ectx = B.Context(local_rank)
                    extra["config1"] = extra_config1(ectx, torch)
                    extra["config5"] = extra_config5(ectx, torch, args.config5_blobs)
                    ectx.close()
                n_dev = min(world, torch.cuda.device_count())
                extra["config4"] = extra_config4(torch, list(range(n_dev)), args.config4_mb_per_gpu, 2)
            except Exception as e:   # an extra leg never takes the headline line down
                extra["error"] = f"{type(e).__name__}: {e}"
            line["extra"] = extra
        print(json.dumps(line))
    if dist is not None:
        dist.barrier(group=host_group)
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours")
    ap.add_argument("--workload", default="config2", choices=sorted(WORKLOADS))
    ap.add_argument("--size-mb", type=int, default=0, help="per-GPU input size override")
    ap.add_argument("--quality", type=int, default=0, help="quality override")
    ap.add_argument("--lgwin", type=int, default=0)
    ap.add_argument("--segment-kb", type=int, default=0, help="parse segment size override (0 = library default)")
    ap.add_argument("--lanes", type=int, default=0, help="CUDA streams encoding groups of shards at once (0 = library default)")
    ap.add_argument("--group-shards", type=int, default=0, help="shards per lane group (0 = library default)")
    ap.add_argument("--q11-lanes", type=int, default=0)
    ap.add_argument("--q11-group-shards", type=int, default=0)
    ap.add_argument("--flags", type=int, default=0, help="bgpu_tuning.flags (BGPU_F_*) for A/B runs")
    ap.add_argument("--no-verify", action="store_true")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-q11", action="store_true", help="skip the config-3 (q11) measurement that accompanies the default run")
    ap.add_argument("--q11-size-mb", type=int, default=0)
    ap.add_argument("--no-extra", action="store_true", help="skip the config 1 / 4 / 5 legs")
    ap.add_argument("--config5-blobs", type=int, default=2000, help="blobs of the config-5 batch leg (BASELINE: 10000)")
    ap.add_argument("--config4-mb-per-gpu", type=int, default=256, help="input MiB per GPU of the stitched multi-GPU leg (BASELINE: 1024)")
    args = ap.parse_args()
    wl = dict(WORKLOADS[args.workload])
    if args.size_mb:
        wl["size_mb"] = args.size_mb
    if args.quality:
        wl["quality"] = args.quality
        wl["metric"] = f"input_GBps_q{args.quality}"
        wl["label"] = f"[--quality {args.quality} override: " + ("SimpleHasher one-slot table + trivial metablocks" if args.quality < 5 else
                                                               "hash-chain matcher" if args.quality < 10 else "H10 + Zopfli") + "] " + wl["label"]
    if args.lgwin:
        wl["lgwin"] = args.lgwin
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference(args, wl, rank, world)
    else:
        run_gpu(args, wl, rank, world, local_rank)


if __name__ == "__main__":
    main()
