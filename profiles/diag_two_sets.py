"""Diagnostic: two resident chunks on one GPU, their variant sets driven from two host threads."""
import os
os.environ.setdefault("MRZ_VARIANTS_CONCURRENT", "1")
import sys
import threading

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "modern-rzip_b200"))
sys.path.insert(0, os.path.join(ROOT, "tests"))
sys.path.insert(0, ROOT)
import bench
import helpers as H
import mrz_py

n = int(sys.argv[1]) if len(sys.argv) > 1 else 128 << 20
nvar = int(sys.argv[2]) if len(sys.argv) > 2 else 16
mode = sys.argv[3] if len(sys.argv) > 3 else "all"
hidx, cb = H.hash_index(), H.chunk_bytes(n)
prm = mrz_py.make_params(7, cb, hidx, 0)
sets = []
for k in range(2):
    vs = mrz_py.VariantSet(0, n, nvar)
    vs.upload(bench.GENS[os.environ.get("KIND", "vm")](n, seed=2 + k))
    sets.append(vs)
if mode in ("all", "seq"):
    for k in range(2):
        r = sets[k].run(n, prm)
        print("sequential set", k, "ok", [int(x.victim_round_out) for x in r], flush=True)
if mode in ("all", "par"):
    out = [None, None]
    def work(k):
        out[k] = sets[k].run(n, prm)
    th = [threading.Thread(target=work, args=(k,)) for k in range(2)]
    [t.start() for t in th]
    [t.join() for t in th]
    print("parallel ok", [[int(x.victim_round_out) for x in r] for r in out], flush=True)
if mode == "emit":
    for rep in range(3):
        out = [sets[k].run(n, prm) for k in range(2)]
        print("rep", rep, "run ok", flush=True)
        for k, v in ((0, 0), (1, 9 % nvar)):
            r = sets[k].ctxs[v].run_emit()
            print("rep", rep, "set", k, "variant", v, "emit ok", int(r.s0_len), int(r.s1_len), flush=True)
