"""TEST TOOLING: conflict / serial-op statistics of the speculation rules in spec_model.c per batch size."""
import sys, os, time
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests"))
import helpers as H
kind, mib = sys.argv[1], int(sys.argv[2])
batches = [int(x) for x in sys.argv[3:]] or [1024]
gen = {"tar": H.gen_tar_like, "rand": H.gen_random, "vm": H.gen_vm_like, "dense": H.gen_dense}[kind]
data = gen(mib << 20)
t = time.time(); ref = H.oracle_chunk(data); print("oracle %.1fs" % (time.time() - t), ref["stats"])
for b in batches:
    t = time.time(); r = H.spec_model_chunk(data, batch=b)
    ok = r["s0"] == ref["s0"] and r["s1"] == ref["s1"] and all(r["stats"][k] == ref["stats"][k] for k in ("inserts", "tag_misses", "tag_hits", "final_min_mask"))
    print("B=%d ok=%s %.1fs" % (b, ok, time.time() - t), r["spec"])
