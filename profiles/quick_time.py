#!/usr/bin/env python
"""Device-resident timing of one chunk under several launch geometries, with the automaton's phase counters.
usage: python profiles/quick_time.py <tar|vm|dense|random> <MiB> <level> [CLxT ...]"""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "modern-rzip_b200")]
import numpy as np
import bench
import mrz_py

NAMES = {56: "x:pre_op", 57: "x:sync_queue", 58: "x:push", 50: "lk:walk", 51: "lk:warp_cmp", 52: "lk:cta_ext", 53: "x:load_op", 54: "x:op()", 55: "x:after_op", 18: "cta_lookup_cycles", 19: "cta_lookups", 0: "ring", 1: "gather", 2: "plan", 3: "conflict", 4: "cut+commit", 5: "executor", 6: "replan", 7: "epilogue", 10: "exact_ops", 11: "rounds",
         16: "ops", 17: "batches", 21: "cut:filter+", 22: "cut:twin", 23: "cut:insert-complex", 24: "cut:conflict", 25: "cut:exact-other",
         30: "replans", 31: "forwarded", 32: "zone", 33: "giveup:follower", 34: "giveup:dep", 35: "giveup:w1owner", 36: "giveup:equal",
         37: "giveup:lesser", 38: "giveup:limits", 39: "giveup:w1claimed", 42: "x:lookup", 43: "x:lookup_n", 44: "x:insert_fast", 45: "x:insert_fast_n",
         46: "x:insert_slow", 47: "x:insert_slow_n", 48: "x:clean", 49: "x:clean_n"}


def main():
    kind, mib, level = sys.argv[1], int(sys.argv[2]), int(sys.argv[3])
    geoms = sys.argv[4:] or ["0x256"]
    n = mib << 20
    pin = mrz_py.PinnedBuffer(n)
    t0 = time.time()
    bench.GENS[kind](n, seed=2, out=pin.array)
    print("generated %s %d MiB in %.1fs" % (kind, mib, time.time() - t0), flush=True)
    hidx, cb = mrz_py.hash_index(), mrz_py.chunk_bytes(n)
    prm = mrz_py.make_params(level, cb, hidx, 0)
    first = None
    for g in geoms:
        cl, t = [int(x) for x in g.split("x")[:2]]
        extra = {}
        for kv in g.split("x")[2:]:
            k, v = kv.split("=")
            extra[k] = int(v)
        ctx = mrz_py.Context(0, n, cluster_size=cl, threads=t, profile=1, **extra)
        ctx.upload(pin.array)
        best = None
        for rep in range(2):
            r = ctx.run_resident(n, prm)
            if best is None or r.ms_kernels < best.ms_kernels:
                best = r
        cnt, geom = ctx.profile()
        sig = (best.s0_len, best.s1_len, best.matches, best.inserts, best.tag_hits, best.tag_misses, best.victim_round_out, best.crc32)
        if first is None:
            first = sig
        out = {"geom": g, "used": geom, "ms_kernels": round(best.ms_kernels, 2), "ms_map+crc": round(best.ms_tagscan, 2),
               "ms_automaton": round(best.ms_automaton, 2), "ms_emit": round(best.ms_emit, 2), "GB/s": round(n / best.ms_kernels / 1e6, 3),
               "same_result": sig == first, "n_ops": best.n_ops, "final_min_mask": best.final_min_mask, "victim_events": best.victim_events}
        prof = {NAMES[i]: cnt[i] for i in sorted(NAMES) if cnt[i]}
        print(json.dumps(out), flush=True)
        print("   ", json.dumps(prof), flush=True)
        ctx.close()


if __name__ == "__main__":
    main()
