#!/usr/bin/env python
"""bench.py - throughput of the salz compression hot path on B200.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference]
                    [--workload english256|dna1g|single512|random512|mixed] [--level L] [--mib M]

A "step" is one pass of the whole hot path (suffix array -> PLCP/LCP -> PSV/NSV
+ match lengths -> optimal parse -> bit/byte emission) over the workload's
blocks.  Default workload = BASELINE.json configs[1]: 256 MiB + 4096 B of
synthetic English-like text cut into 1 MiB blocks (CLI level 5), per GPU.

  value  compress MB/s (input bytes), input and output resident in HBM
  e2e    same through the host-buffer C-ABI call (H2D + kernels + D2H timed)
  roofline / cpu_baseline: see DESIGN.md "Measurement"

N > 1 (torchrun): every rank compresses its own shard of the stream (weak
scaling, no data-path collective) and the per-block encoded sizes are
all-gathered over NCCL inside the timed step.
"""
import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import tempfile
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "compress_MBps_input"
UNIT = "MB/s"

WORKLOADS = {
    # name: (generator, MiB per GPU, description)
    "english256": ("english", 256, "256 MiB + 4096 B English-like text (BASELINE configs[1])"),
    "dna1g": ("dna", 1024, "1 GiB + 4096 B DNA-like with planted repeats (configs[2])"),
    "single512": ("single", 512, "512 MiB + 4096 B single repeated byte (configs[4]a)"),
    "random512": ("random", 512, "512 MiB + 4096 B uniform random bytes (configs[4]b)"),
    "mixed": ("mixed", 1024, "1 GiB + 4096 B per GPU of the 8 GiB mixed stream (configs[3])"),
}


def make_input(kind, size, seed, block_len):
    """Seeded synthetic input.  SALZ_BENCH_CACHE=<dir> keeps generated inputs as .npy files
    (A/B scripts that run bench.py many times in one box visit; never set by the driver)."""
    cache = os.environ.get("SALZ_BENCH_CACHE")
    if cache:
        path = os.path.join(cache, "%s_%d_%d_%d.npy" % (kind, size, seed, block_len))
        if os.path.exists(path):
            return np.load(path)
        data = _make_input(kind, size, seed, block_len)
        os.makedirs(cache, exist_ok=True)
        np.save(path, data)
        return data
    return _make_input(kind, size, seed, block_len)


def _make_input(kind, size, seed, block_len):
    from salz_b200 import synth
    if kind == "english":
        return synth.english_like(size, seed=seed)
    if kind == "dna":
        return synth.dna_like(size, seed=seed + 1, block=block_len)
    if kind == "single":
        return synth.single_byte(size)
    if kind == "random":
        return synth.uniform_random(size, seed=seed + 3)
    if kind == "mixed":
        return synth.mixed_stream(size, seed=seed + 2)
    raise ValueError(kind)


def host_cores():
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


class ClockSampler:
    """nvidia-smi clocks/throttle reasons during the timed region (B200_PROFILING.md)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
         "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index = index
        self.proc = None
        self.path = None

    def start(self):
        try:
            fd, self.path = tempfile.mkstemp(suffix=".csv")
            os.close(fd)
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits",
                 "-lms", "100"], stdout=open(self.path, "w"), stderr=subprocess.DEVNULL)
        except OSError:
            self.proc = None

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        sm, mx, reasons, power = [], [], set(), []
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in open(self.path):
            f = [x.strip() for x in line.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1])); mx.append(float(f[2])); power.append(float(f[3]))
            except ValueError:
                continue
            for nm, v in zip(names, f[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        os.unlink(self.path)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"]}
        return {"sm_mhz": float(np.median(sm)), "sm_max_mhz": float(max(mx)), "reasons": sorted(reasons),
                "samples": len(sm), "power_w_max": max(power)}


def run_cpu_reference(data, block_len, nblocks_sample, threads):
    """The reference's own encoder (oracle/_ref, compiled in place from the upstream
    sources) or, when that prebuilt library is absent, the oracle restatement, on
    `nblocks_sample` blocks with `threads` host threads.  Returns (seconds, bytes, kind)."""
    from concurrent.futures import ThreadPoolExecutor
    from oracle.binding import Port, Ref
    if Ref.available():
        orc, kind = Ref(), "reference"
    else:
        orc, kind = Port(), "port"
    nblocks = (len(data) + block_len - 1) // block_len
    nb = min(nblocks, nblocks_sample)
    blocks = [data[b * block_len: (b + 1) * block_len] for b in range(nb)]
    nbytes = sum(len(b) for b in blocks)
    t0 = time.perf_counter()
    if threads <= 1:
        for b in blocks:
            orc.encode(b)
    else:
        with ThreadPoolExecutor(threads) as ex:  # ctypes releases the GIL during the call
            list(ex.map(orc.encode, blocks))
    return time.perf_counter() - t0, nbytes, kind


def run_cpu_libsais(data, block_len, nblocks_sample, threads):
    """SA only, beside the device SA stage (CPU baseline iii): upstream
    `libsais()` from oracle/_ref on the same blocks -- one thread, one block per thread
    on all host cores, and `libsais_omp(threads)` block after block.  GB/s of text, or
    None without the compiled reference."""
    from concurrent.futures import ThreadPoolExecutor
    from oracle.binding import Ref
    if not Ref.available():
        return None
    ref = Ref()
    nblocks = (len(data) + block_len - 1) // block_len
    blocks = [data[b * block_len: (b + 1) * block_len] for b in range(min(nblocks, nblocks_sample))]
    blocks = [b[: len(b) - 8] for b in blocks if len(b) > 8]   # the reference sorts T[0..N-8)
    out = {"blocks": len(blocks), "threads": threads}
    one = blocks[: max(1, min(4, len(blocks)))]
    t0 = time.perf_counter()
    for b in one:
        ref.suffix_array(b)
    out["libsais_1thread_GBps"] = sum(len(b) for b in one) / (time.perf_counter() - t0) / 1e9
    t0 = time.perf_counter()
    with ThreadPoolExecutor(threads) as ex:
        list(ex.map(ref.suffix_array, blocks))
    out["libsais_block_per_thread_GBps"] = sum(len(b) for b in blocks) / (time.perf_counter() - t0) / 1e9
    omp = blocks[: max(1, min(16, len(blocks)))]
    t0 = time.perf_counter()
    for b in omp:
        ref.suffix_array(b, threads=threads)
    out["libsais_omp_GBps"] = sum(len(b) for b in omp) / (time.perf_counter() - t0) / 1e9
    return out


def reference_arm(args, kind, mib, block_len, desc):
    """bench.py --impl reference: the reference CPU implementation on the host cores."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    cores = host_cores()
    sample_blocks = max(cores, min(128, max(1, (64 << 20) // block_len)))
    size = min(mib << 20, sample_blocks * block_len) + 4096
    data = make_input(kind, size, args.seed, block_len)
    nb = (len(data) + block_len - 1) // block_len
    for _ in range(args.warmup):
        run_cpu_reference(data, block_len, min(nb, cores), cores)
    tot_s, tot_b = 0.0, 0
    k = "reference"
    for _ in range(args.steps):
        s, b, k = run_cpu_reference(data, block_len, nb, cores)
        tot_s += s
        tot_b += b
    v = tot_b / tot_s / 1e6
    sample = "%d blocks of %d B (%s) per step, %d threads, one block per reference call" % (
        nb, block_len, kind, cores)
    line = {
        "impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": tot_s / args.steps * 1e3, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "u8/i32", "data": "synthetic",
        "config": {"workload": desc, "level": args.level, "block_len": block_len},
        "cpu_baseline": {"value": v, "unit": UNIT, "cores": cores, "kind": k, "sample": sample},
        "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line))
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="english256", choices=sorted(WORKLOADS))
    ap.add_argument("--level", type=int, default=5)
    ap.add_argument("--mib", type=int, default=0, help="override MiB per GPU")
    ap.add_argument("--seed", type=int, default=1)
    ap.add_argument("--batch-blocks", type=int, default=0, help="blocks per device batch (0 = all)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    args = ap.parse_args()
    kind, mib, desc = WORKLOADS[args.workload]
    if args.mib:
        mib = args.mib
        desc += " [size overridden: %d MiB]" % mib
    block_len = 1 << (15 + args.level)

    if args.impl == "reference":
        return reference_arm(args, kind, mib, block_len, desc)

    import torch
    import salz_b200

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    dist = None
    if world > 1:
        import torch.distributed as dist
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        torch.cuda.set_device(local_rank)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (the product path has no CPU fallback)")
    dev = torch.device("cuda", local_rank)
    torch.cuda.set_device(dev)

    size = (mib << 20) + 4096
    data = make_input(kind, size, args.seed + rank, block_len)
    nblocks = (size + block_len - 1) // block_len
    batch = args.batch_blocks or nblocks
    batch = min(batch, (1 << 30) // block_len)
    enc = salz_b200.SalzCuda(device=local_rank, max_block_len=block_len, max_batch_blocks=batch)
    stream = torch.cuda.Stream(device=dev)
    enc.set_stream(stream.cuda_stream)

    h_src = torch.from_numpy(data).pin_memory()
    cap = salz_b200.api.stream_len_max(size, block_len)
    h_dst = torch.empty(cap, dtype=torch.uint8).pin_memory()
    d_src = h_src.to(dev)
    d_dst = torch.empty(cap, dtype=torch.uint8, device=dev)
    sizes = np.zeros(nblocks, np.uint32)
    d_sizes = torch.zeros(nblocks, dtype=torch.int32, device=dev)
    gathered = torch.zeros(world * nblocks, dtype=torch.int32, device=dev) if world > 1 else None

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    out_len = [0]

    def step_device():
        with torch.cuda.stream(stream):
            out_len[0] = enc.encode_stream_ptr(d_src.data_ptr(), size, block_len, d_dst.data_ptr(), cap, sizes,
                                               device=True)
            if dist is not None:
                # the only exchange of the path: per-block encoded sizes -> stream offsets
                d_sizes.copy_(torch.from_numpy(sizes.astype(np.int32)), non_blocking=True)
                dist.all_gather_into_tensor(gathered, d_sizes)

    def step_host():
        with torch.cuda.stream(stream):
            out_len[0] = enc.encode_stream_ptr(h_src.data_ptr(), size, block_len, h_dst.data_ptr(), cap, sizes,
                                               device=False)
            if dist is not None:
                d_sizes.copy_(torch.from_numpy(sizes.astype(np.int32)), non_blocking=True)
                dist.all_gather_into_tensor(gathered, d_sizes)
                gathered.cpu()

    def timed(fn, steps):
        ev0 = torch.cuda.Event(enable_timing=True)
        ev1 = torch.cuda.Event(enable_timing=True)
        barrier()
        t0 = time.perf_counter()
        with torch.cuda.stream(stream):
            ev0.record(stream)
        for _ in range(steps):
            fn()
        with torch.cuda.stream(stream):
            ev1.record(stream)
        barrier()
        wall = (time.perf_counter() - t0) * 1e3
        ms = ev0.elapsed_time(ev1)
        t = torch.tensor([ms, wall], dtype=torch.float64, device=dev)
        if dist is not None:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t[0]), float(t[1])

    # ---- device-resident arm ---------------------------------------------------
    for _ in range(args.warmup):
        step_device()
    enc.reset_stats()
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    ms_dev, wall_dev = timed(step_device, args.steps)
    clocks = sampler.stop() if rank == 0 else None
    st = enc.stats()
    out_bytes = out_len[0]

    # ---- end-to-end arm (host buffers through the C ABI) ---------------------------
    e2e = None
    if not args.no_e2e:
        step_host()
        ms_e2e, wall_e2e = timed(step_host, args.steps)
        e2e = {"value": world * size * args.steps / (ms_e2e * 1e-3) / 1e6, "unit": UNIT,
               "h2d_bytes_per_step": size, "d2h_bytes_per_step": int(out_len[0]),
               "ms_per_step": ms_e2e / args.steps}

    if rank != 0:
        if dist is not None:
            dist.barrier()
            dist.destroy_process_group()
        return 0

    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except (OSError, ValueError):
        pass
    peak = float(peaks.get("hbm_gbs", 6650.0))
    peak_src = "MEASURED_PEAKS.json hbm_gbs (measured copy)" if "hbm_gbs" in peaks else "fallback 6650 GB/s"

    # Per-kernel rooflines, algorithmic bytes over CUDA-event time on the library's
    # stream.  The dominant kernel (largest share of the step) is reported as "roofline"; since the
    # serial-chain parse took that place from the radix pass, the sort pass follows in "roofline_others".
    traffic = {}
    try:
        traffic = json.load(open(os.path.join(ROOT, "profiles", "r01_v15_traffic.json")))
    except (OSError, ValueError):
        pass

    def tr(name):
        t = traffic.get(name)
        return t["dram_bytes_per_launch"] if t else None

    positions = float(st["input_bytes"]) - 8.0 * st["blocks"]
    sort_bytes = float(st["sort_bytes"])
    sort_launches = max(1, st["sort_launches"])
    parse_launches = max(1, args.steps * ((nblocks + batch - 1) // batch))
    cands = [
        {"kernel": "k_parse", "bound": "hbm", "alg_bytes": 32.0 * positions, "ms": st["ms_parse"],
         "launches": parse_launches, "traffic": (tr("k_parse_warm") + tr("k_parse") + tr("k_choose"))
         if tr("k_parse_warm") and tr("k_parse") and tr("k_choose") else None,
         "note": "whole parse stage per batch (k_parse: shortened first pass + full pass, k_choose incl. the jump "
                 "pointers; its CUDA-event time): serial (min,+) chains, 4 per consumer warp, 16 per SM, "
                 "latency-bound by construction; 32 B/position = candidates 16 + cost 4 + choice 8 + jump 4; "
                 "traffic = sum over the stage's launches"},
        {"kernel": "k_scatter", "bound": "hbm", "alg_bytes": sort_bytes, "ms": st["ms_sort"],
         "launches": sort_launches, "traffic": tr("k_scatter<1, 0>"),
         "note": "one LSD radix pass: 16 B per list element (read + write of a packed element)"},
    ]
    rl = []
    for c in cands:
        ach = c["alg_bytes"] / (c["ms"] * 1e-3) / 1e9 if c["ms"] > 0 else 0.0
        rl.append({"bound": c["bound"], "kernel": c["kernel"], "achieved": ach, "peak": peak, "unit": "GB/s",
                   "frac": ach / peak, "traffic": c["traffic"], "peak_source": peak_src,
                   "alg_bytes_per_launch": c["alg_bytes"] / c["launches"],
                   "avg_launch_ms": c["ms"] / c["launches"], "launches": c["launches"],
                   "share_of_step": c["ms"] / st["ms_total"] if st["ms_total"] else None, "note": c["note"]})
    rl.sort(key=lambda r: -(r["share_of_step"] or 0))
    roofline = rl[0]
    roofline["pipeline_alg_bytes_per_input_byte"] = (st["alg_bytes_sa"] + st["alg_bytes_rest"]) / max(1, st["input_bytes"])
    roofline["pipeline_achieved"] = (st["alg_bytes_sa"] + st["alg_bytes_rest"]) / (st["ms_total"] * 1e-3) / 1e9
    roofline_others = rl[1:]

    cpu = None
    if not args.no_cpu_baseline:
        cores = host_cores()
        nb = min(nblocks, max(cores, min(128, (128 << 20) // block_len)))
        s, b, k = run_cpu_reference(data, block_len, nb, cores)
        cpu = {"value": b / s / 1e6, "unit": UNIT, "cores": cores, "kind": k,
               "sample": "first %d of %d blocks (%d B each), %d threads, one block per reference call, %.1f s wall"
                         % (nb, nblocks, block_len, cores, s)}
        nb1 = min(nblocks, 4)
        s1, b1, _ = run_cpu_reference(data, block_len, nb1, 1)
        cpu["single_thread_value"] = b1 / s1 / 1e6
        cpu["sa"] = run_cpu_libsais(data, block_len, nb, cores)

    n_in = st["input_bytes"]
    value = world * size * args.steps / (ms_dev * 1e-3) / 1e6
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms_dev / args.steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "u8/i32", "data": "synthetic",
        "config": {"workload": desc, "level": args.level, "block_len": block_len, "blocks_per_gpu": nblocks,
                   "bytes_per_gpu": size, "batch_blocks": batch, "seed": args.seed,
                   "l2": "inputs and workspace larger than L2 (%.0f MiB text, ~%.0f GiB workspace per pass)"
                         % (size / 2**20, 96.0 * size / 2**30),
                   "parallelism": "blocks sharded over %d GPU(s); NCCL all-gather of sizes" % world},
        "e2e": e2e, "gpu_launches": int(st["launches"]), "clocks": clocks,
        "roofline": roofline, "roofline_others": roofline_others, "cpu_baseline": cpu,
        "extra": {
            "compressed_bytes_per_gpu": int(out_bytes), "ratio": size / max(1, out_bytes),
            "wall_ms_per_step": wall_dev / args.steps,
            "sa_GBps": (n_in / 1e9) / (st["ms_sa"] * 1e-3) if st["ms_sa"] else None,
            "doubling_rounds_per_step": st["doubling_rounds"] / args.steps,
            "stage_ms_per_step": {k: st[k] / args.steps for k in
                                  ("ms_sa", "ms_sort", "ms_lcp", "ms_ansv", "ms_parse", "ms_emit", "ms_total")},
        },
    }
    print(json.dumps(line))
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
