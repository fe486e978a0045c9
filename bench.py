#!/usr/bin/env python
"""bench.py — rzip pre-process throughput (input GB/s) on B200, next to the reference's CPU path.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--size BYTES]

A step = one pass of the hot path (hash_search, rzip.c:507-667) over one chunk of synthetic input.
N=1 workload: BASELINE.json configs[1] — 4 GiB tar-like input, duplicate bodies at distances up to
4 GiB, one chunk, rzip level 7.  N>1 (weak scaling, no data-path collective): every rank processes its own
4 GiB input of the SAME kind (seed 2 + rank), one chunk per GPU, each starting with victim_round = 0 exactly
as N reference processes on N files would; `--workload vm` switches every rank (also N=1) to the VM-image-like
data of configs[2].  `--one-file` instead treats the N chunks as consecutive
chunks of ONE input and resolves the victim_round chain exactly (modern-rzip_b200/shard.py): chunks whose
speculated carry was wrong are re-run inside the timed region.
Prints ONE JSON line (rank 0).  `value` = device-resident throughput (CUDA events); `e2e` = the same
through mrz_rzip_chunk() with pinned host buffers (H2D + kernels + D2H inside the timed region).
"""
import argparse
import json
import os
import subprocess
import sys
import tempfile
import threading
import time

import numpy as np

if os.environ.get("NCCL_DEBUG", "VERSION").upper() == "VERSION":
    os.environ["NCCL_DEBUG"] = "WARN"     # keep NCCL's version banner off stdout: rank 0 prints ONE JSON line

os.environ.setdefault("CUDA_DEVICE_MAX_CONNECTIONS", "32")   # --variants: one hardware queue per automaton stream (before torch creates the context)

_REAL_STDOUT = 1

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(ROOT, "modern-rzip_b200"))
sys.path.insert(0, os.path.join(ROOT, "tests"))

GIB = 1 << 30
LEVEL = 7


def measured_traffic():
    """dram bytes per k_automaton launch from the committed ncu --set full capture (profiles/), or None."""
    try:
        with open(os.path.join(ROOT, "profiles", "r1_final_traffic.json")) as f:
            return json.load(f)
    except Exception:
        return None


def hbm_peak():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured"
    except Exception:
        return 6650.0, "fallback"


# ------------------------------------------------------------------ workloads
def gen_tar_like(n, seed=2, out=None):
    """configs[1]: 512-byte headers + bodies (log-uniform 1 KiB..64 MiB); 35 % verbatim copies of an
    earlier body, the rest half random bytes, half 64-symbol text."""
    rng = np.random.default_rng(seed)
    buf = out if out is not None else np.zeros(n, dtype=np.uint8)
    buf[:n] = 0
    alphabet = np.frombuffer(b"etaoinshrdlucmfwypvbgkqjxz ETAOINSHRDLUCMFWYPVBGKQJXZ.,;\n0123456", dtype=np.uint8)
    bodies = []
    pos = idx = 0
    max_body = min(64 << 20, max(4096, n // 8))
    while pos + 1024 <= n:
        size = int(np.exp(rng.uniform(np.log(1024), np.log(max_body))))
        size = min(size, n - pos - 512)
        name = ("member%08d.dat" % idx).encode()
        buf[pos:pos + len(name)] = np.frombuffer(name, dtype=np.uint8)
        sz = ("%011o" % size).encode()
        buf[pos + 124:pos + 124 + len(sz)] = np.frombuffer(sz, dtype=np.uint8)
        pos += 512
        r = rng.random()
        if r < 0.35 and bodies:
            s, l = bodies[int(rng.integers(0, len(bodies)))]
            size = min(l, size)
            buf[pos:pos + size] = buf[s:s + size]
        elif r < 0.675:
            buf[pos:pos + size] = rng.integers(0, 256, size, dtype=np.uint8)
        else:
            buf[pos:pos + size] = alphabet[rng.integers(0, 64, size, dtype=np.uint8)]
        bodies.append((pos, size))
        pos += (size + 511) // 512 * 512
        idx += 1
    return buf


def gen_vm_like(n, seed=3, out=None):
    """configs[2]: 4 KiB blocks — 35 % zero, 25 % copy of an earlier block, 20 % random, 20 % text."""
    rng = np.random.default_rng(seed)
    blk = 4096
    nb = n // blk
    buf = out if out is not None else np.zeros(n, dtype=np.uint8)
    buf[:n] = 0
    v = buf[:nb * blk].reshape(nb, blk)
    kinds = rng.random(nb)
    rnd = np.nonzero((kinds >= 0.60) & (kinds < 0.80))[0]
    txt = np.nonzero(kinds >= 0.80)[0]
    alphabet = np.frombuffer(b"abcdefghijklmnopqrstuvwxyzABCDEFGHIJKLMNOPQRSTUVWXYZ0123456789 \n", dtype=np.uint8)
    step = 1 << 14
    for i in range(0, rnd.size, step):
        ix = rnd[i:i + step]
        v[ix] = rng.integers(0, 256, (ix.size, blk), dtype=np.uint8)
    for i in range(0, txt.size, step):
        ix = txt[i:i + step]
        v[ix] = alphabet[rng.integers(0, 64, (ix.size, blk), dtype=np.uint8)]
    cp = np.nonzero((kinds >= 0.35) & (kinds < 0.60))[0]
    cp = cp[cp > 0]
    src = (rng.random(cp.size) * cp).astype(np.int64)  # uniformly chosen earlier block
    for d, s in zip(cp.tolist(), src.tolist()):        # sequential: a copy may copy a copy
        v[d] = v[s]
    return buf


def gen_dense(n, seed=5, out=None, base=1 << 20, mutations=100):
    """configs[4]: one random 1 MiB base block written over and over with 100 random point mutations per copy."""
    rng = np.random.default_rng(seed)
    buf = out if out is not None else np.zeros(n, dtype=np.uint8)
    blk = rng.integers(0, 256, base, dtype=np.uint8)
    for pos in range(0, n, base):
        m = min(base, n - pos)
        buf[pos:pos + m] = blk[:m]
        idx = rng.integers(0, m, mutations)
        buf[pos + idx] = rng.integers(0, 256, mutations, dtype=np.uint8)
    return buf


def gen_random(n, seed=4, out=None):
    """configs[3]: uniform random bytes (all literals, worst-case table pressure); drawn in 64 MiB pieces."""
    rng = np.random.default_rng(seed)
    buf = out if out is not None else np.empty(n, dtype=np.uint8)
    step = 64 << 20
    for pos in range(0, n, step):
        m = min(step, n - pos)
        buf[pos:pos + m] = rng.integers(0, 256, m, dtype=np.uint8)
    return buf


GENS = {"tar": gen_tar_like, "vm": gen_vm_like, "dense": gen_dense, "random": gen_random}
KINDS = {"tar": "configs[1]: tar-like, duplicate bodies at distances up to the chunk size",
         "vm": "configs[2]: VM-image-like (4 KiB blocks: zero / copies / random / text)",
         "dense": "configs[4]: dense repeats (1 MiB base block, 100 point mutations per copy)",
         "random": "configs[3]: incompressible uniform random bytes (all literals)"}


# ------------------------------------------------------------------ clocks
class ClockSampler:
    Q = "clocks.sm,clocks.max.sm,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown," \
        "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown," \
        "clocks_event_reasons.sw_power_cap"

    def __init__(self, gpu_index):
        self.idx = gpu_index
        self.samples = []
        self.proc = None

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.idx), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits",
                 "-lms", "200"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.samples.append(line.strip())

    def stop(self):
        if self.proc:
            self.proc.terminate()
        sm, mx, reasons = [], 0, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for s in self.samples:
            f = [x.strip() for x in s.split(",")]
            try:
                sm.append(float(f[0]))
                mx = max(mx, float(f[1]))
                for nm, val in zip(names, f[3:7]):
                    if val.lower().startswith("active"):
                        reasons.add(nm)
            except Exception:
                pass
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": mx or None,
                "reasons": sorted(reasons), "samples": len(sm)}


# ------------------------------------------------------------------ reference CPU arm
def run_reference_sample(sample, steps, warmup, level=LEVEL):
    """Times the unmodified reference binary (oracle/_ref/mrzip -n) on a bounded sample."""
    import helpers as H
    if not H.have_ref():
        raise RuntimeError("oracle/_ref/mrzip missing (make -C oracle ref)")
    tmpdir = "/dev/shm" if os.path.isdir("/dev/shm") else None
    times = []
    with tempfile.TemporaryDirectory(dir=tmpdir) as d:
        fin, fout = os.path.join(d, "in.bin"), os.path.join(d, "out.mar")
        sample.tofile(fin)
        mflag = str(max(1, (sample.size * 3 // 2) // (100 * 1048576) + 2))  # ramsize so that one chunk holds it
        for i in range(warmup + steps):
            t0 = time.perf_counter()
            H.run_reference(fin, fout, ("-m", mflag, "-p", "1", "-L", str(level)))
            dt = time.perf_counter() - t0
            if i >= warmup:
                times.append(dt)
    return times


def emit_line(line):
    """The ONE JSON line goes to the real stdout; anything libraries print to fd 1 (NCCL's banner) was sent to stderr."""
    os.write(_REAL_STDOUT, (json.dumps(line) + "\n").encode())


CHAIN = [1, 2, 2, 2, 3, 4, 6, 16, 32, 128]     # levels[].max_chain_len, rzip.c:65-73


def variants_bench(args, rank, world, local_rank, config, barrier):
    """One input of world x chunks_per_gpu chunks (chunk k -> rank k mod world).  Every chunk is resident on its GPU and
    its automaton runs for all max_chain_len values of victim_round_in concurrently (one CTA + one table each); the
    chunks of a rank run concurrently too (one host thread each).  The chain is resolved with one all-gather
    (shard.run_sharded_variants) and only the chosen variant of each chunk is emitted.  Timed by the wall clock between
    barriers (the critical path includes the host threads and the all-gather); max over ranks."""
    import torch
    import mrz_py
    import helpers as H
    import shard
    multi = world > 1
    if multi:
        import torch.distributed as dist
    os.environ.setdefault("MRZ_VARIANTS_CONCURRENT", "1")   # this mode measures the side-by-side form (include/mrz_gpu.h)
    config["MRZ_VARIANTS_CONCURRENT"] = os.environ["MRZ_VARIANTS_CONCURRENT"]
    n, C_, V = args.size, args.chunks_per_gpu, CHAIN[args.level]
    n_chunks = world * C_
    hidx, cb = H.hash_index(), H.chunk_bytes(n)
    prm = mrz_py.make_params(args.level, cb, hidx, 0)
    sets, pins = [], []
    for c in range(C_):
        k = c * world + rank
        vs = mrz_py.VariantSet(local_rank, n, V)
        pin = mrz_py.PinnedBuffer(n)
        GENS[args.workload](n, seed=2 + k, out=pin.array)
        sets.append(vs)
        pins.append(pin)
    pin_s0 = mrz_py.PinnedBuffer(n // 2 + 64)
    pin_s1 = mrz_py.PinnedBuffer(n + 16)
    chunks = [None] * n_chunks
    launches = 0
    last = {}

    def trace(msg):
        if os.environ.get("MRZ_BENCH_TRACE"):
            print("[variants rank %d] %s" % (rank, msg), file=sys.stderr, flush=True)

    def step(with_copies):
        nonlocal launches
        outs = [None] * C_

        def work(c):
            if with_copies:
                sets[c].upload(pins[c].array)
            outs[c] = sets[c].run(n, prm)
        th = [threading.Thread(target=work, args=(c,)) for c in range(1, C_)]
        for t in th:
            t.start()
        work(0)
        for t in th:
            t.join()
        trace("runs done")
        it = iter(range(C_))
        chosen, vr = shard.run_sharded_variants(
            chunks, lambda _c: [r.victim_round_out for r in outs[next(it)]], V, rank, world, device="cuda")
        b_alg = 0
        trace("chain done: %r" % (chosen,))
        for c in range(C_):
            v = chosen[c * world + rank]
            launches += sum(int(r.gpu_launches) for r in outs[c])
            if with_copies:
                r, s0, s1 = sets[c].emit(v, pin_s0.array, pin_s1.array)
            else:
                r = sets[c].ctxs[v].run_emit()
            launches += int(r.gpu_launches) - int(outs[c][v].gpu_launches)
            trace("emit chunk %d variant %d done" % (c, v))
            b_alg += n + r.s0_len + r.s1_len
            last[c] = (r, v, max(float(x.ms_automaton) for x in outs[c]), float(outs[c][v].ms_automaton))
        return b_alg, vr

    for c in range(C_):
        sets[c].upload(pins[c].array)
    for _ in range(args.warmup):
        step(False)
    clocks = ClockSampler(local_rank)
    barrier()
    clocks.start()
    launches = 0
    t0 = time.perf_counter()
    b_alg = 0
    for _ in range(args.steps):
        b, vr_final = step(False)
        b_alg += b
    barrier()
    wall_k = time.perf_counter() - t0
    clk = clocks.stop()
    barrier()
    t0 = time.perf_counter()
    d2h = 0
    for _ in range(args.steps):
        step(True)
    barrier()
    wall_e = time.perf_counter() - t0
    tk = torch.tensor([wall_k, wall_e, float(b_alg)], dtype=torch.float64, device="cuda")
    if multi:
        tmax = tk.clone()
        dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
        dist.all_reduce(tk, op=dist.ReduceOp.SUM)
        t_k, t_e, b_all = float(tmax[0]), float(tmax[1]), float(tk[2])
    else:
        t_k, t_e, b_all = [float(x) for x in tk.tolist()]
    total_bytes = n * n_chunks * args.steps
    peak, peak_kind = hbm_peak()
    config["workload"] = "%s, ONE input of %d chunks x %.2f GiB (chunk k on GPU k mod %d, %d chunk(s) resident per GPU), " \
        "rzip level %d, all %d victim_round_in variants of every chunk run side by side, chain resolved by one all-gather" % (
            KINDS[args.workload], n_chunks, n / GIB, world, C_, args.level, V)
    config["timing"] = "wall clock between barriers, max over ranks"
    config["automata_per_gpu"] = C_ * V
    line = {"metric": "rzip pre-process input GB/s", "value": total_bytes / t_k / 1e9, "unit": "GB/s", "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * t_k / args.steps,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u8/int64",
            "data": "synthetic", "config": config,
            "e2e": {"value": total_bytes / t_e / 1e9, "unit": "GB/s", "h2d_bytes_per_step": n * C_,
                    "d2h_bytes_per_step": int(sum(int(last[c][0].s0_len + last[c][0].s1_len) for c in range(C_)))},
            "gpu_launches": int(launches), "clocks": clk,
            "roofline": {"bound": "hbm", "kernel": "k_automaton x %d concurrent" % (C_ * V),
                         "achieved": b_all / t_k / 1e9 / world, "peak": peak, "peak_kind": peak_kind, "unit": "GB/s",
                         "frac": b_all / t_k / 1e9 / world / peak, "traffic": None,
                         "note": "algorithmic bytes of the CHOSEN variants of one GPU's chunks / wall time of the step"},
            "variants": {"per_chunk": V, "final_victim_round": int(vr_final),
                         "chosen": {str(c * world + rank): int(last[c][1]) for c in range(C_)},
                         "ms_automaton_slowest_variant": {str(c * world + rank): last[c][2] for c in range(C_)},
                         "ms_automaton_chosen_variant": {str(c * world + rank): last[c][3] for c in range(C_)}}}
    if rank == 0:
        emit_line(line)
    for vs in sets:
        vs.close()
    if multi:
        dist.destroy_process_group()
    return 0


def main():
    global _REAL_STDOUT
    sys.stdout.flush()
    _REAL_STDOUT = os.dup(1)
    os.dup2(2, 1)                         # library chatter on fd 1 -> stderr; emit_line() writes to the saved stdout
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=2)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200")
    ap.add_argument("--size", type=int, default=4 * GIB, help="chunk bytes (default 4 GiB = configs[1])")
    ap.add_argument("--cpu-sample", type=int, default=1024 << 20, help="bytes of the workload given to the CPU baseline")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--workload", default="tar", choices=["tar", "vm", "dense"], help="tar: configs[1] data; vm: configs[2] data; dense: configs[4] data")
    ap.add_argument("--level", type=int, default=LEVEL, help="rzip level (index into levels[], rzip.c:65-73)")
    ap.add_argument("--one-file", action="store_true", help="N>1: the ranks' chunks are consecutive chunks of one input (exact carry)")
    ap.add_argument("--variants", action="store_true",
                    help="chunks of ONE input, chunk k on rank k mod N, every chunk run for all max_chain_len values of "
                         "victim_round_in side by side (mrz_run_variants); one all-gather resolves the chain, no re-runs")
    ap.add_argument("--chunks-per-gpu", type=int, default=1, help="--variants: chunks resident and running concurrently per GPU")
    args = ap.parse_args()

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    n = args.size
    multi = world > 1
    kind = KINDS[args.workload]
    workload = "%s, %.2f GiB single chunk per GPU, rzip level %d" % (kind, n / GIB, args.level)
    if multi:
        workload += ", consecutive chunks of one input with the exact victim_round carry" if args.one_file else \
            ", one independent input per GPU (seed 2 + rank)"
    config = {"workload": workload, "chunk_bytes_per_gpu": n, "rzip_level": args.level,
              "l2": "input (%.1f GiB) and 64 MiB table re-initialised every step; input larger than L2" % (n / GIB)}

    # ---------------- reference arm: the reference's own CPU implementation on the host cores
    if args.impl == "reference":
        if rank != 0:
            return 0
        gen = GENS[args.workload]
        sample_n = min(args.cpu_sample, n)
        sample = gen(sample_n, seed=2)
        w = min(args.warmup, 1)
        times = run_reference_sample(sample, args.steps, w, args.level)
        gbs = sample_n / (sum(times) / len(times)) / 1e9
        line = {"impl": "reference", "metric": "rzip pre-process input GB/s", "value": gbs, "unit": "GB/s",
                "n_gpus": args.gpus, "steps": args.steps, "warmup": w, "ms_per_step": 1e3 * sum(times) / len(times),
                "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u8/int64",
                "data": "synthetic", "config": config,
                "cpu_baseline": {"value": gbs, "unit": "GB/s", "cores": 1, "kind": "reference",
                                 "sample": "first %d MiB of the workload generator as one chunk (the reference does most of its table "
                                           "work early in a chunk, so a prefix understates its whole-chunk rate), "
                                           "oracle/_ref/mrzip -n -L7 file->file "
                                           "(1 hash_search thread + checksum/writer helpers), host has %d cores"
                                           % (sample_n >> 20, os.cpu_count())},
                "e2e": {"value": gbs, "unit": "GB/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
        emit_line(line)
        return 0

    # ---------------- B200 arm
    import torch
    import mrz_py
    import helpers as H
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the product has no CPU fallback")
    torch.cuda.set_device(local_rank)
    if multi:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    def barrier():
        if multi:
            dist.barrier()
        torch.cuda.synchronize()

    if args.variants:
        return variants_bench(args, rank, world, local_rank, config, barrier)

    hidx = H.hash_index()
    cb = H.chunk_bytes(n)
    ctx = mrz_py.Context(local_rank, n)
    pin_in = mrz_py.PinnedBuffer(n)
    pin_s0 = mrz_py.PinnedBuffer(n // 2 + 64)
    pin_s1 = mrz_py.PinnedBuffer(n + 16)
    data = pin_in.array
    GENS[args.workload](n, seed=2 + rank, out=data)
    prm = mrz_py.make_params(args.level, cb, hidx, 0)

    clocks = ClockSampler(local_rank)
    reruns_total = 0

    def resident_step():
        """One pass over this rank's chunk; with --one-file the victim_round chain across ranks is resolved exactly
        (shard.py) and a chunk whose speculated carry was wrong runs again.  Returns (last result, summed timings)."""
        nonlocal reruns_total
        if not (multi and args.one_file):
            r = ctx.run_resident(n, prm)
            return r, (r.ms_kernels, r.ms_automaton, r.ms_tagscan, r.ms_emit, r.gpu_launches)
        import shard
        acc = [0.0, 0.0, 0.0, 0.0, 0]

        def run(_chunk, vr_in):
            r = ctx.run_resident(n, mrz_py.make_params(args.level, cb, hidx, vr_in))
            for i, v in enumerate((r.ms_kernels, r.ms_automaton, r.ms_tagscan, r.ms_emit, r.gpu_launches)):
                acc[i] += v
            return int(r.victim_events), int(r.victim_round_out), r
        results, _vr, reruns = shard.run_sharded([None] * world, run, rank, world, device="cuda")
        reruns_total += reruns
        return results[rank], tuple(acc)

    # device-resident timing (K)
    ctx.upload(data)
    for _ in range(args.warmup):
        resident_step()
    barrier()
    clocks.start()
    t0 = time.perf_counter()
    ms_k = ms_auto = ms_scan = ms_emit = 0.0
    launches = 0
    res = None
    for _ in range(args.steps):
        res, tms = resident_step()
        ms_k += tms[0]
        ms_auto += tms[1]
        ms_scan += tms[2]
        ms_emit += tms[3]
        launches += tms[4]
    barrier()
    wall_k = time.perf_counter() - t0
    clk = clocks.stop()
    # e2e timing (host buffers through the C-ABI call)
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        r2, s0, s1 = ctx.rzip_chunk(data, prm, pin_s0.array, pin_s1.array)
    barrier()
    wall_e = time.perf_counter() - t0

    tk = torch.tensor([ms_k / 1e3, wall_k, wall_e], dtype=torch.float64, device="cuda")
    if multi:
        dist.all_reduce(tk, op=dist.ReduceOp.MAX)
    t_dev, t_wall, t_e2e = [float(x) for x in tk.tolist()]
    if multi and args.one_file:
        t_dev = t_wall   # the chain serialises re-runs across ranks: the critical path is the wall clock between the barriers
        config["timing"] = "wall clock between barriers (victim_round chain across ranks); re-runs this rank: %d" % reruns_total
    total_bytes = n * world * args.steps
    value = total_bytes / t_dev / 1e9
    e2e = total_bytes / t_e2e / 1e9
    b_alg = n + res.s0_len + res.s1_len
    peak, peak_kind = hbm_peak()
    achieved = b_alg * args.steps / (ms_auto / 1e3) / 1e9
    line = {"metric": "rzip pre-process input GB/s", "value": value, "unit": "GB/s", "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * t_dev / args.steps,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u8/int64",
            "data": "synthetic", "config": config,
            "e2e": {"value": e2e, "unit": "GB/s", "h2d_bytes_per_step": n,
                    "d2h_bytes_per_step": int(res.s0_len + res.s1_len)},
            "gpu_launches": int(launches),
            "clocks": clk,
            "roofline": {"bound": "hbm", "kernel": "k_automaton", "achieved": achieved, "peak": peak,
                         "peak_kind": peak_kind, "unit": "GB/s", "frac": achieved / peak,
                         "traffic": (measured_traffic() or {}).get("dram_bytes_per_launch"),
                         "traffic_source": (measured_traffic() or {}).get("source"),
                         "launches_per_step": int(launches) // max(1, args.steps),
                         "algorithmic_bytes_per_step": int(b_alg),
                         "share_of_step": ms_auto / ms_k if ms_k else None,
                         "ms_per_step": {"tag_scan": ms_scan / args.steps, "automaton": ms_auto / args.steps,
                                         "emit+crc": ms_emit / args.steps}},
            "wall_s_per_step_resident": t_wall / args.steps,
            "result": {"matches": int(res.matches), "match_bytes": int(res.match_bytes),
                       "literal_bytes": int(res.literal_bytes), "inserts": int(res.inserts),
                       "n_ops": int(res.n_ops), "final_min_mask": int(res.final_min_mask),
                       "victim_events": int(res.victim_events)}}
    if rank == 0 and not multi and not args.no_cpu_baseline:
        sample_n = min(args.cpu_sample, n)
        times = run_reference_sample(np.array(data[:sample_n]), 1, 0, args.level)
        line["cpu_baseline"] = {"value": sample_n / times[0] / 1e9, "unit": "GB/s", "cores": 1, "kind": "reference",
                                "sample": "first %d MiB of the same input as one chunk (a prefix understates the reference's "
                                          "whole-chunk rate: its table work is front-loaded), oracle/_ref/mrzip -n -L7 "
                                          "file->file in /dev/shm (1 hash_search thread + helpers); host has %d cores"
                                          % (sample_n >> 20, os.cpu_count())}
    if rank == 0:
        emit_line(line)
    ctx.close()
    if multi:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
