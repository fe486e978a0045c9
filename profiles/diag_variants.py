"""Diagnostic (not a bench): how the wall time of mrz_run_variants scales with the number of concurrent automata."""
import os
os.environ.setdefault("MRZ_VARIANTS_CONCURRENT", "1")
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "modern-rzip_b200"))
sys.path.insert(0, os.path.join(ROOT, "tests"))
sys.path.insert(0, ROOT)
import bench
import helpers as H
import mrz_py

n = int(sys.argv[1]) if len(sys.argv) > 1 else 128 << 20
kind = sys.argv[2] if len(sys.argv) > 2 else "vm"
data = bench.GENS[kind](n, seed=2)
hidx, cb = H.hash_index(), H.chunk_bytes(n)
vs = mrz_py.VariantSet(0, n, 16)
vs.upload(data)
prm = mrz_py.make_params(7, cb, hidx, 0)
vs.run(n, prm, 1)
for nvar in (1, 2, 4, 8, 16):
    t0 = time.perf_counter()
    res = vs.run(n, prm, nvar)
    dt = time.perf_counter() - t0
    print("n=%d MiB %s nvar=%2d wall=%.3f s | automaton ms min/max %.1f/%.1f | tag_scan ms max %.1f | launches %d | MAX_CONN=%s" % (
        n >> 20, kind, nvar, dt, min(r.ms_automaton for r in res), max(r.ms_automaton for r in res),
        max(r.ms_tagscan for r in res), res[0].gpu_launches, os.environ.get("CUDA_DEVICE_MAX_CONNECTIONS")), flush=True)
vs.close()
