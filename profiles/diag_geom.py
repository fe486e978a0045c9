#!/usr/bin/env python
"""Compare two launch geometries on growing prefixes of one input; print the first differing token.
usage: python profiles/diag_geom.py <kind> <level> <geomA> <geomB> [MiB ...]"""
import os
import sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "modern-rzip_b200")]
import numpy as np
import bench
import mrz_py

kind, level, ga, gb = sys.argv[1], int(sys.argv[2]), sys.argv[3], sys.argv[4]
sizes = [int(x) for x in sys.argv[5:]] or [64, 128, 256, 512, 1024]
nmax = max(sizes) << 20
pin = mrz_py.PinnedBuffer(nmax)
bench.GENS[kind](nmax, seed=2, out=pin.array)
hidx = mrz_py.hash_index()
for mib in sizes:
    n = mib << 20
    out = []
    for g in (ga, gb):
        cl, t = [int(x) for x in g.split("x")]
        ctx = mrz_py.Context(0, n, cluster_size=cl, threads=t)
        ctx.upload(pin.array[:n])
        r = ctx.run_resident(n, mrz_py.make_params(level, 5, hidx, 0))
        out.append((r, ctx.debug_tokens().copy()))
        ctx.close()
    (ra, ta), (rb, tb) = out
    same = ta.shape == tb.shape and np.array_equal(ta, tb)
    print(mib, "MiB:", "same" if same else "DIFFERENT", "n_ops", ra.n_ops, rb.n_ops, "inserts", ra.inserts, rb.inserts, "misses", ra.tag_misses, rb.tag_misses,
          "hits", ra.tag_hits, rb.tag_hits, "mask", ra.final_min_mask, rb.final_min_mask, "vev", ra.victim_events, rb.victim_events, flush=True)
    if not same:
        m = min(len(ta), len(tb))
        d = np.nonzero((ta[:m] != tb[:m]).any(axis=1))[0]
        i = int(d[0]) if d.size else m
        print("  first differing token", i, "of", len(ta), len(tb), "\n  A:", ta[max(0, i - 2):i + 3].tolist(), "\n  B:", tb[max(0, i - 2):i + 3].tolist(), flush=True)
        break
