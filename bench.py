#!/usr/bin/env python
"""bench.py — rzip pre-process throughput (input GB/s) on B200, next to the reference's CPU path.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--workload ...]

A step = one pass of the hot path (hash_search, rzip.c:507-667) over the step's chunks of synthetic input.
Default workload (`--workload tar`):
  N=1   BASELINE.json configs[1]: 4 GiB tar-like input, duplicate bodies at distances up to 4 GiB, ONE chunk, rzip
        level 7, one automaton (a thread-block cluster).  Golden vector: tests/golden/fullsize_vectors.json "tar4g".
  N>1   weak scaling of the same unit of work inside ONE input: the first N chunks of the 8-chunk tar-like input
        "tar32g_w41" (`-w 41`: chunks of 4,299,161,600 B), chunk k on rank k.  Chunks of one input are linked by the
        static victim_round of insert_hash (rzip.c:259): chunk 0 starts from 0, every later chunk is searched for
        ALL 16 values of victim_round_in in one kernel launch (mrz_run_variants), ONE all-gather carries the
        (chunk, vr_in) -> vr_out table, every rank walks the chain and emits the variant it selects.  No chunk waits
        for another one's search and none is searched twice.
Other workloads (lines kept under profiles/):
  --workload vm32 [--window 41|10]   configs[2]: the 32 GiB VM-image-like input through the product's C host
                                     (`mrzip -n -g 0,..,N-1`: ONE process drives the N GPUs, strong scaling), wall
                                     clock of the process, archive compared with the reference's (SHA-256)
  --case NAME                        a golden single-chunk case of tests/fullsize_cases.py (random8g, vm4g, dense4g_R9, ...)
  --workload vm|random|dense --size BYTES [--level L]   one chunk of another data kind (configs[3]: random 8 GiB;
                                     configs[4]: dense 16 GiB --level 9)
  --independent --chunks-per-gpu C   C independent inputs resident per GPU, all their automata in ONE launch
Prints ONE JSON line (rank 0).  `value` = device-resident throughput (tag map, CRC, automaton and emission every step);
`e2e` = the same with the host<->device copies of the step's inputs and streams inside the timed region (pinned host
buffers, through the C-ABI); at N=1 also `e2e_cli` (wall clock of `mrzip -n`, file -> archive) and `cpu_baseline` (the
unmodified reference on a bounded sample).  `--impl reference`: the unmodified reference binary on the WHOLE chunk, plus
the labelled k-processes-on-k-cores figure.  Once per run, outside the timed region, the streams are compared with the
reference's (SHA-256 per chunk): "parity_checked".
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

_REAL_STDOUT = 1

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(ROOT, "modern-rzip_b200"))
sys.path.insert(0, os.path.join(ROOT, "tests"))

GIB = 1 << 30
LEVEL = 7


def measured_traffic(workload):
    """dram bytes of the k_automaton launch from the committed ncu --set full capture at the benched size (profiles/), or None."""
    try:
        with open(os.path.join(ROOT, "profiles", "r2_traffic.json")) as f:
            return json.load(f).get(workload)
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


W41_CHUNK = 41 * 100 * 1048576     # `-w 41`: chunks of 4,299,161,600 bytes (rzip.c:46 CHUNK_MULTIPLE, :884)


def gen_tar_chunks(n, seed=2, out=None, chunk=W41_CHUNK, only=None):
    """A multi-chunk tar-like input: piece k (chunk bytes, the last one shorter) is gen_tar_like(seed + k).  Matches never
    cross rzip chunks, so independent pieces lose nothing, and a rank can generate its own chunk alone (only=k)."""
    if only is not None:
        lo = only * chunk
        m = min(chunk, n - lo)
        return gen_tar_like(m, seed=seed + only, out=out)
    buf = out if out is not None else np.zeros(n, dtype=np.uint8)
    for k, lo in enumerate(range(0, n, chunk)):
        m = min(chunk, n - lo)
        gen_tar_like(m, seed=seed + k, out=buf[lo:lo + m])
    return buf


GENS = {"tar": gen_tar_like, "vm": gen_vm_like, "dense": gen_dense, "random": gen_random, "tarchunks": gen_tar_chunks}
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
def ref_flags(n, level):
    mflag = str(max(1, (n * 3 + 1) // (100 * 1048576) + 2))   # ramsize/3 >= n: the chunk is mapped whole, one block per stream
    return ("-m", mflag, "-p", "1", "-L", str(level))


def time_reference(fin, fout, n, level, runs):
    """Wall clock of the unmodified reference binary (oracle/_ref/mrzip -n) file -> file, per run."""
    import helpers as H
    if not H.have_ref():
        raise RuntimeError("oracle/_ref/mrzip missing (make -C oracle ref)")
    times = []
    for _ in range(runs):
        t0 = time.perf_counter()
        H.run_reference(fin, fout, ref_flags(n, level))
        times.append(time.perf_counter() - t0)
    return times


def time_reference_many(files, n, level):
    """k reference processes on k files at once (the whole-box figure SURVEY §8d allows, labelled as such)."""
    import helpers as H
    env = dict(os.environ, MRZIP="NOCONFIG")
    t0 = time.perf_counter()
    ps = [subprocess.Popen([H.REF_BIN, "-n", "-q", "-f", *ref_flags(n, level), "-o", f + ".mar", f], env=env,
                           stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL) for f in files]
    for p in ps:
        if p.wait():
            raise RuntimeError("reference process failed")
    return time.perf_counter() - t0


def shm_dir():
    return tempfile.TemporaryDirectory(dir="/dev/shm" if os.path.isdir("/dev/shm") else None)


def emit_line(line):
    """The ONE JSON line goes to the real stdout; anything libraries print to fd 1 (NCCL's banner) was sent to stderr."""
    os.write(_REAL_STDOUT, (json.dumps(line) + "\n").encode())


def sha(a):
    import hashlib
    return hashlib.sha256(memoryview(np.ascontiguousarray(a))).hexdigest()


def golden(name):
    try:
        with open(os.path.join(ROOT, "tests", "golden", "fullsize_vectors.json")) as f:
            return json.load(f).get(name)
    except Exception:
        return None


def reference_arm(args, config, n, kind, seed):
    """`--impl reference`: the reference's own CPU implementation of the path on this box's host cores, on the SAME
    workload: the whole configured chunk, file -> file in /dev/shm.  One process = one core for hash_search plus its
    checksum/writer helper threads; the k-process figure next to it is k processes on k 512 MiB samples."""
    cores = os.cpu_count() or 1
    with shm_dir() as d:
        fin, fout = os.path.join(d, "in.bin"), os.path.join(d, "out.mar")
        mm = np.memmap(fin, dtype=np.uint8, mode="w+", shape=(n,))
        GENS[kind](n, seed=seed, out=mm)
        mm.flush()
        budget_runs = max(1, min(args.steps, int(240 // max(1.0, 22.0 * n / GIB))))   # ~22 s per GiB here: stay within minutes
        times = time_reference(fin, fout, n, args.level, budget_runs)
        sample = min(n, 512 << 20)
        files = []
        for i in range(cores):
            f = os.path.join(d, "s%d.bin" % i)
            np.asarray(mm[:sample]).tofile(f)
            files.append(f)
        del mm
        os.unlink(fin)
        t_many = time_reference_many(files, sample, args.level)
    t = sum(times) / len(times)
    gbs = n / t / 1e9
    config["steps_actually_run"] = len(times)
    line = {"impl": "reference", "metric": "rzip pre-process input GB/s", "value": gbs, "unit": "GB/s",
            "n_gpus": args.gpus, "steps": len(times), "warmup": 0, "ms_per_step": 1e3 * t,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u8/int64",
            "data": "synthetic", "config": config,
            "cpu_baseline": {"value": gbs, "unit": "GB/s", "cores": 1, "kind": "reference",
                             "sample": "the WHOLE configured chunk (%d bytes), oracle/_ref/mrzip -n -L%d file->file in /dev/shm, %d run(s): "
                                       "1 hash_search thread + checksum/writer helper threads; host has %d cores" % (n, args.level, len(times), cores),
                             "k_processes": {"k": cores, "value": cores * sample / t_many / 1e9, "unit": "GB/s",
                                             "sample": "%d reference processes at once, each on its own copy of the first %d MiB" % (cores, sample >> 20)}},
            "e2e": {"value": gbs, "unit": "GB/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    emit_line(line)
    return 0


CHAIN = [1, 2, 2, 2, 3, 4, 6, 16, 32, 128]     # levels[].max_chain_len, rzip.c:65-73


def job_for(args, world):
    """-> dict(kind, seed, n_total, chunk, golden, label): the input of this run and the golden vector that pins it."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from fullsize_cases import CASES
    if args.case:
        c = CASES[args.case]
        return dict(kind=c["kind"], seed=c["seed"], n_total=c["n"], chunk=c["n"], golden=args.case, level=9 if "R9" in args.case else args.level,
                    label="%s (tests/fullsize_cases.py), one chunk" % args.case)
    if args.workload != "tar" or args.size:
        n = args.size or 4 * GIB
        seed = {"tar": 2, "vm": 2, "random": 4, "dense": 5}[args.workload]
        return dict(kind=args.workload, seed=seed, n_total=n * (world if args.independent else 1), chunk=n, golden=None, level=args.level,
                    label="%s, %.2f GiB chunk, rzip level %d" % (KINDS[args.workload], n / GIB, args.level))
    if world == 1 and not args.independent:
        return dict(kind="tar", seed=2, n_total=4 * GIB, chunk=4 * GIB, golden="tar4g", level=args.level,
                    label="%s, 4 GiB single chunk, rzip level %d" % (KINDS["tar"], args.level))
    n_total = min(world * W41_CHUNK, 32 * GIB)
    return dict(kind="tarchunks", seed=2, n_total=n_total, chunk=W41_CHUNK, golden="tar32g_w41", level=args.level,
                label="%s: the first %d chunk(s) of ONE 8-chunk input (`-w 41`: 4,299,161,600 B per chunk, piece k = seed 2+k), chunk k on GPU k, "
                      "rzip level %d; exact victim_round carry: chunk 0 from 0, later chunks searched for all %d values at once, one all-gather"
                      % (KINDS["tar"], world, args.level, CHAIN[args.level]))


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
    ap.add_argument("--workload", default="tar", choices=["tar", "vm", "dense", "random", "vm32"])
    ap.add_argument("--case", default=None, help="a single-chunk case of tests/fullsize_cases.py (random8g, dense16g_R9, vm4g, ...): golden-checked")
    ap.add_argument("--size", type=int, default=0, help="chunk bytes for --workload vm|dense|random|tar (no golden vector)")
    ap.add_argument("--level", type=int, default=LEVEL, help="rzip level (index into levels[], rzip.c:65-73)")
    ap.add_argument("--window", type=int, default=41, help="--workload vm32: `-w` (41 -> 8 chunks, 10 -> 33 chunks)")
    ap.add_argument("--slots", type=int, default=0, help="--workload vm32: chunks in flight per device (mrzip -G)")
    ap.add_argument("--independent", action="store_true", help="independent inputs (victim_round 0 each), --chunks-per-gpu of them per GPU in ONE launch")
    ap.add_argument("--chunks-per-gpu", type=int, default=1)
    ap.add_argument("--cluster", type=int, default=0, help="CTAs per automaton (0: as many as fit)")
    ap.add_argument("--threads", type=int, default=0, help="threads per CTA (0: library default)")
    ap.add_argument("--profile", action="store_true", help="collect the automaton's phase counters (slower)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-cli", action="store_true", help="skip the mrzip -n file->archive wall-clock comparison (e2e_cli)")
    args = ap.parse_args()

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    multi = world > 1
    if args.workload == "vm32":
        return vm32_cli(args, rank, world)
    job = job_for(args, world)
    args.level = job["level"]
    n = job["chunk"]
    config = {"workload": job["label"], "chunk_bytes_per_gpu": n, "rzip_level": args.level,
              "l2": "input (%.1f GiB per GPU) and the 64 MiB table are re-initialised every step; input larger than L2" % (n / GIB)}

    # ---------------- reference arm: the reference's own CPU implementation on the host cores
    if args.impl == "reference":
        if rank != 0:
            return 0
        if job["kind"] == "tarchunks":      # the unit of the weak-scaling run: one chunk (a reference process walks chunks one by one)
            config["workload"] += " | reference arm: chunk 0 alone (4,299,161,600 B), one process"
            return reference_arm(args, config, W41_CHUNK, "tar", 2)
        return reference_arm(args, config, n, job["kind"], job["seed"])

    # ---------------- B200 arm
    import torch
    import mrz_py
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

    V = CHAIN[args.level]
    C = args.chunks_per_gpu if args.independent else 1
    one_file = multi and not args.independent
    cfg = {}
    if args.cluster:
        cfg["cluster_size"] = args.cluster
    if args.threads:
        cfg["threads"] = args.threads
    if args.profile:
        cfg["profile"] = 1
    hidx = mrz_py.hash_index()
    # this rank's chunks: one-file mode: chunk `rank` of the input; independent mode: C inputs of its own
    my = []
    for c in range(C):
        k = rank if one_file else rank * C + c
        lo = k * n
        m = min(n, job["n_total"] - lo) if one_file else n
        pin = mrz_py.PinnedBuffer(m)
        if job["kind"] == "tarchunks":
            gen_tar_chunks(job["n_total"], seed=job["seed"], out=pin.array, only=k)
        else:
            GENS[job["kind"]](m, seed=job["seed"] + (k if args.independent else 0), out=pin.array)
        my.append(dict(k=k, n=m, pin=pin, cb=mrz_py.chunk_bytes(m)))
    nvar = V if (one_file and rank > 0) else 1      # chunk 0 of an input starts from victim_round 0: nothing to speculate on
    sets = [mrz_py.VariantSet(local_rank, ch["n"], nvar, **cfg) for ch in my]
    pin_s0 = mrz_py.PinnedBuffer(n // 2 + 64)
    pin_s1 = mrz_py.PinnedBuffer(n + 16)
    state = {"launches": 0, "ms_auto": 0.0, "ms_map": 0.0, "ms_emit": 0.0, "ms_kernels": 0.0, "b_alg": 0, "chosen": {}, "res": {}}

    def step(with_copies, keep=None):
        """One pass over this rank's chunks.  with_copies: H2D of the inputs and D2H of the streams inside the step (e2e)."""
        if with_copies:
            for vs, ch in zip(sets, my):
                vs.upload(ch["pin"].array)
        else:       # resident inputs: the tag map and the CRC are part of the step, not leftovers of the upload
            for vs, ch in zip(sets, my):
                vs.parent.invalidate_map(ch["n"])
        if C == 1:
            vs, ch = sets[0], my[0]
            outs = [vs.run(ch["n"], mrz_py.make_params(args.level, ch["cb"], hidx, 0), nvar)]
        else:       # all resident chunks of this GPU in ONE launch
            outs = [[r] for r in mrz_py.run_many([vs.parent for vs in sets], [ch["n"] for ch in my],
                                                 [mrz_py.make_params(args.level, ch["cb"], hidx, 0) for ch in my])]
        vr_in = 0
        if one_file:    # the chain: one all-gather of every chunk's (vr_in -> vr_out) row; chunk 0's row is constant
            row = torch.tensor([int(outs[0][v].victim_round_out) if nvar > 1 else int(outs[0][0].victim_round_out) for v in range(V)],
                               dtype=torch.int64, device="cuda")
            table = [torch.empty_like(row) for _ in range(world)]
            dist.all_gather(table, row)
            table = torch.stack(table).tolist()
            vr = 0
            for k in range(world):
                if k == rank:
                    vr_in = vr
                vr = table[k][vr]
        for c, (vs, ch) in enumerate(zip(sets, my)):
            v = vr_in if nvar > 1 else 0
            r = vs.ctxs[v].run_emit()
            if with_copies:
                s0, s1 = vs.ctxs[v].download(r, pin_s0.array, pin_s1.array)
                if keep is not None:
                    keep[ch["k"]] = (sha(s0), sha(s1), int(r.s0_len), int(r.s1_len))
            a = outs[c][v]
            # the chunks of one run_many call share ONE launch (and its tag-map pass): every chunk reports that launch's
            # time, which counts once; the emissions are launches of their own
            first = C == 1 or c == 0
            state["launches"] += int(r.gpu_launches) if first else max(0, int(r.gpu_launches) - int(a.gpu_launches))
            state["ms_auto"] += float(a.ms_automaton) if first else 0.0
            state["ms_map"] += float(a.ms_tagscan) if first else 0.0
            state["ms_emit"] += float(r.ms_emit)
            state["ms_kernels"] += float(r.ms_kernels) if first else float(r.ms_emit)
            state["b_alg"] += ch["n"] + int(r.s0_len) + int(r.s1_len)
            state["chosen"][ch["k"]] = v
            state["res"][ch["k"]] = r

    def reset():
        for key in ("launches", "b_alg"):
            state[key] = 0
        for key in ("ms_auto", "ms_map", "ms_emit", "ms_kernels"):
            state[key] = 0.0

    # parity, once, outside the timed region: the streams of this rank's chunks against the reference's (golden SHA-256)
    got = {}
    step(True, keep=got)
    g = golden(job["golden"]) if job["golden"] else None
    parity = None
    if g is not None:
        parity = True
        for ch in my:
            want = g["chunks"][ch["k"]]
            parity = parity and got[ch["k"]] == (want["s0_sha256"], want["s1_sha256"], want["s0_len"], want["s1_len"])
    if multi:
        flag = torch.tensor([1 if parity in (True, None) else 0], device="cuda")
        dist.all_reduce(flag, op=dist.ReduceOp.MIN)
        if parity is not None or g is not None:
            parity = bool(flag.item())
    if parity is False:
        raise SystemExit("PARITY FAILURE: the streams differ from the reference's golden vector %s" % job["golden"])

    # device-resident timing (K): inputs already in HBM
    for _ in range(args.warmup):
        step(False)
    clocks = ClockSampler(local_rank)
    barrier()
    clocks.start()
    reset()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        step(False)
    barrier()
    wall_k = time.perf_counter() - t0
    clk = clocks.stop()
    k_state = dict(state)
    prof = sets[0].ctxs[state["chosen"][my[0]["k"]]].profile() if args.profile else None
    # e2e timing: pinned host buffers through the C-ABI, H2D and D2H inside
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        step(True)
    barrier()
    wall_e = time.perf_counter() - t0

    tk = torch.tensor([k_state["ms_kernels"] / 1e3, wall_k, wall_e], dtype=torch.float64, device="cuda")
    tsum = torch.tensor([float(k_state["b_alg"]), float(sum(ch["n"] for ch in my))], dtype=torch.float64, device="cuda")
    if multi:
        dist.all_reduce(tk, op=dist.ReduceOp.MAX)
        dist.all_reduce(tsum, op=dist.ReduceOp.SUM)
    t_dev, t_wall, t_e2e = [float(x) for x in tk.tolist()]
    b_alg_all, bytes_per_step = float(tsum[0]), float(tsum[1])
    if one_file:
        t_val = t_wall      # the chain crosses ranks (host + all-gather): the wall clock between the barriers is the critical path
        config["timing"] = "wall clock between barriers, max over ranks (CUDA-event time of the slowest rank: %.1f ms per step)" % (1e3 * t_dev / args.steps)
    else:
        t_val = t_dev
        config["timing"] = "CUDA events around all kernels of a step (tag map + CRC, table memsets, automaton, emission), summed over the steps, max over ranks"
    total_bytes = bytes_per_step * args.steps
    res0 = state["res"][my[0]["k"]]
    peak, peak_kind = hbm_peak()
    # roofline of the dominant kernel on rank 0: algorithmic bytes of this rank's chunks / its k_automaton time
    achieved = k_state["b_alg"] / (k_state["ms_auto"] / 1e3) / 1e9 if k_state["ms_auto"] else 0.0
    geom = sets[0].parent.profile()[1]
    config["automata_per_gpu"] = C * nvar
    config["geometry"] = "%d CTA(s) per cluster x %d threads, %d automaton/automata per launch" % tuple(geom)
    traffic = measured_traffic(job["golden"] or args.workload) or {}
    line = {"metric": "rzip pre-process input GB/s", "value": total_bytes / t_val / 1e9, "unit": "GB/s", "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * t_val / args.steps,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u8/int64",
            "data": "synthetic", "config": config, "parity_checked": parity,
            "e2e": {"value": total_bytes / t_e2e / 1e9, "unit": "GB/s", "h2d_bytes_per_step": int(sum(ch["n"] for ch in my)),
                    "d2h_bytes_per_step": int(sum(int(state["res"][ch["k"]].s0_len + state["res"][ch["k"]].s1_len) for ch in my))},
            "gpu_launches": int(k_state["launches"]),
            "clocks": clk,
            "roofline": {"bound": "hbm", "kernel": "k_automaton", "achieved": achieved, "peak": peak,
                         "peak_kind": peak_kind, "unit": "GB/s", "frac": achieved / peak,
                         "traffic": traffic.get("dram_bytes_per_launch"), "traffic_source": traffic.get("source"),
                         "launches_per_step": int(k_state["launches"]) // max(1, args.steps),
                         "algorithmic_bytes_per_step": int(k_state["b_alg"]) // max(1, args.steps),
                         "share_of_step": k_state["ms_auto"] / k_state["ms_kernels"] if k_state["ms_kernels"] else None,
                         "ms_per_step": {"tagmap+crc": k_state["ms_map"] / args.steps, "automaton": k_state["ms_auto"] / args.steps,
                                         "emit": k_state["ms_emit"] / args.steps},
                         "note": "one launch of k_automaton per step and GPU; a latency-bound exact automaton, see DESIGN.md §5"},
            "wall_s_per_step_resident": t_wall / args.steps,
            "result": {"matches": int(res0.matches), "match_bytes": int(res0.match_bytes), "literal_bytes": int(res0.literal_bytes),
                       "inserts": int(res0.inserts), "n_ops": int(res0.n_ops), "final_min_mask": int(res0.final_min_mask),
                       "victim_events": int(res0.victim_events), "victim_round_in_chosen": {str(k): int(v) for k, v in state["chosen"].items()}}}
    if prof:
        line["profile"] = {"counters": prof[0], "geometry": list(prof[1])}
    if rank == 0 and not multi and not args.independent:
        extra = host_side_lines(args, job, my[0]["pin"].array, cli=not args.no_cli, cpu=not args.no_cpu_baseline)
        line.update(extra)
    if rank == 0:
        emit_line(line)
    for vs in sets:
        vs.close()
    if multi:
        dist.destroy_process_group()
    return 0


def host_side_lines(args, job, data, cli=True, cpu=True):
    """N=1 extras, outside the timed regions: `cpu_baseline` (the reference binary on a bounded sample of the same input)
    and `e2e_cli` (wall clock of the product's `mrzip -n` file -> archive, next to nothing but the file system)."""
    import helpers as H
    from fullsize_cases import CASES
    out = {}
    n = data.size
    cores = os.cpu_count() or 1
    with shm_dir() as d:
        fin = os.path.join(d, "in.bin")
        if cli:
            np.asarray(data).tofile(fin)
            flags = list(CASES[job["golden"]]["flags"]) if job["golden"] else list(ref_flags(n, args.level)[:2]) + ["-R", str(args.level)]
            fout = os.path.join(d, "gpu.mar")
            best, err = None, ""
            for _ in range(2):
                t0 = time.perf_counter()
                p = subprocess.run([os.path.join(ROOT, "modern-rzip_b200", "mrzip"), "-n", "-vv", "-f", *flags, "-o", fout, fin],
                                   env=dict(os.environ, MRZIP="NOCONFIG"), capture_output=True, text=True)
                dt = time.perf_counter() - t0
                if p.returncode:
                    raise SystemExit("mrzip failed: " + p.stderr[-2000:])
                if best is None or dt < best:
                    best, err = dt, p.stderr
            g = golden(job["golden"]) if job["golden"] else None
            ident = (H.sha256_file(fout) == g["archive_sha256"]) if g else None
            import re
            m = re.search(r"chunks written after ([0-9.]+) s", err)
            out["e2e_cli"] = {"value": n / best / 1e9, "unit": "GB/s", "wall_s": best,
                              "what": "modern-rzip_b200/mrzip -n file -> archive in /dev/shm, process start to exit (CUDA context creation, "
                                      "pageable reads through pinned slabs, search, D2H, write, whole-file MD5 in one thread), best of 2",
                              "search_and_write_s": float(m.group(1)) if m else None,
                              "archive_identical_to_reference": ident, "flags": flags}
            os.unlink(fout)
        if cpu:
            sample_n = min(n, 1 << 30)
            fs = os.path.join(d, "sample.bin")
            np.asarray(data[:sample_n]).tofile(fs)
            t = time_reference(fs, os.path.join(d, "ref.mar"), sample_n, args.level, 1)[0]
            out["cpu_baseline"] = {"value": sample_n / t / 1e9, "unit": "GB/s", "cores": 1, "kind": "reference",
                                   "sample": "first %d MiB of the same input as one chunk, oracle/_ref/mrzip -n file->file in /dev/shm "
                                             "(1 hash_search thread + checksum/writer helper threads; the whole-chunk figure and the "
                                             "%d-process figure are in the `--impl reference` line); host has %d cores"
                                             % (sample_n >> 20, cores, cores)}
    return out


def vm32_cli(args, rank, world):
    """configs[2] at its real shape through the product's C host: ONE 32 GiB VM-image-like file, `-w` chunks dealt to the
    N GPUs of this box by modern-rzip_b200/mrzip (one process, worker threads per chunk slot, exact victim_round chain),
    archive compared with the reference's.  Scope E: process wall clock, whole-file MD5 (one thread) included."""
    if rank != 0:
        return 0
    world = max(world, args.gpus)      # ONE process drives all the GPUs (`-g 0,1,..`): no torchrun needed for this mode
    import helpers as H
    name = "vm32g_w%d" % args.window
    g = golden(name)
    n = 32 * GIB
    fin = "/dev/shm/mrz_bench_vm32.bin"
    if not (os.path.exists(fin) and os.path.getsize(fin) == n):
        mm = np.memmap(fin, dtype=np.uint8, mode="w+", shape=(n,))
        gen_vm_like(n, seed=3, out=mm)
        mm.flush()
        del mm
    fout = "/dev/shm/mrz_bench_vm32.mar"
    flags = ["-w", str(args.window), "-m", "140", "-p", "1", "-g", ",".join(str(i) for i in range(world))]
    if args.slots:
        flags += ["-G", str(args.slots)]
    walls, err = [], ""
    for _ in range(max(1, args.steps)):
        t0 = time.perf_counter()
        p = subprocess.run([os.path.join(ROOT, "modern-rzip_b200", "mrzip"), "-n", "-vv", "-f", *flags, "-o", fout, fin],
                           env=dict(os.environ, MRZIP="NOCONFIG"), capture_output=True, text=True)
        walls.append(time.perf_counter() - t0)
        err = p.stderr
        if p.returncode:
            raise SystemExit("mrzip failed: " + err[-2000:])
    ident = (H.sha256_file(fout) == g["archive_sha256"]) if g else None
    os.unlink(fout)
    import re
    m = re.search(r"chunks written after ([0-9.]+) s", err)
    best = min(walls)
    line = {"metric": "rzip pre-process input GB/s", "value": n / best / 1e9, "unit": "GB/s", "n_gpus": world, "steps": len(walls), "warmup": 0,
            "ms_per_step": 1e3 * best, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "u8/int64", "data": "synthetic",
            "config": {"workload": "configs[2]: ONE 32 GiB VM-image-like input (seed 3), -w %d (%d chunks), chunks dealt to %d GPU(s) by the C host "
                                   "(modern-rzip_b200/mrzip %s); scope E: process wall clock file -> archive in /dev/shm, whole-file MD5 included"
                                   % (args.window, len(g["chunks"]) if g else -1, world, " ".join(flags)),
                       "timing": "wall clock of the process, best of %d" % len(walls)},
            "parity_checked": ident, "search_and_write_s": float(m.group(1)) if m else None,
            "value_without_md5_wait": n / float(m.group(1)) / 1e9 if m else None,
            "reference_wall_s_build_container": g.get("reference_wall_s_build_container") if g else None,
            "mrzip_log_tail": err[-600:]}
    emit_line(line)
    return 0


if __name__ == "__main__":
    sys.exit(main())
