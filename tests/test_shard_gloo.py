"""N>1 host path on CPU: two gloo ranks shard the chunks of one input (chunk k -> rank k mod 2), run them through
the CPU-emulator build of the product kernels, and resolve the victim_round chain (modern-rzip_b200/shard.py).
The result must equal the oracle run sequentially over the chunks with the carried static (rzip.c:259)."""
import os
import pickle
import socket
import subprocess
import sys
import tempfile

import numpy as np

import helpers as H

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LEVEL, NCHUNK, CHUNK = 1, 5, 640 << 10

WORKER = r'''
import os, pickle, sys
ROOT = sys.argv[1]; out = sys.argv[2]
sys.path.insert(0, os.path.join(ROOT, "tests")); sys.path.insert(0, os.path.join(ROOT, "modern-rzip_b200"))
import numpy as np, torch, torch.distributed as dist
import helpers as H, mrz_py, shard
dist.init_process_group("gloo")
rank, world = dist.get_rank(), dist.get_world_size()
LEVEL, NCHUNK, CHUNK = [int(x) for x in sys.argv[3:6]]
data = H.as_u8(H.gen_motif(NCHUNK * CHUNK, 5000))
chunks = [data[k * CHUNK:(k + 1) * CHUNK] for k in range(NCHUNK)]
ctx = mrz_py.Context(0, CHUNK, library=mrz_py.bind(os.path.join(ROOT, "tests", "cusim", "libmrz_sim.so")))
hidx = H.hash_index(); cb = H.chunk_bytes(CHUNK)
def run(chunk, vr_in):
    res, s0, s1 = ctx.rzip_chunk(chunk, mrz_py.make_params(LEVEL, cb, hidx, vr_in))
    return res.victim_events, res.victim_round_out, (s0.tobytes(), s1.tobytes(), vr_in)
if sys.argv[6] == "variants":
    # variant-parallel: all victim_round_in values of a chunk at once, one all-gather, no re-runs
    chain = [1, 2, 2, 2, 3, 4, 6, 16, 32, 128][LEVEL]
    sets = {}
    def run_variants(chunk):
        vs = mrz_py.VariantSet(0, CHUNK, chain, library=ctx.L)
        vs.upload(chunk)
        sets[id(chunk)] = vs
        return [r.victim_round_out for r in vs.run(chunk.size, mrz_py.make_params(LEVEL, cb, hidx, 0))]
    chosen, vr = shard.run_sharded_variants(chunks, run_variants, chain, rank, world)
    results = {}
    for k, v in chosen.items():
        r, s0, s1 = sets[id(chunks[k])].emit(v)
        results[k] = (s0.tobytes(), s1.tobytes(), v)
    reruns = 0
else:
    results, vr, reruns = shard.run_sharded(chunks, run, rank, world)
pickle.dump({"results": results, "vr": vr, "reruns": reruns}, open(os.path.join(out, "r%d.pkl" % rank), "wb"))
ctx.close(); dist.destroy_process_group()
'''


import pytest


@pytest.mark.parametrize("mode", ["rerun", "variants"])
def test_two_ranks_chain_equals_sequential_oracle(mode):
    subprocess.check_call(["make", "-C", os.path.join(ROOT, "tests", "cusim"), "libmrz_sim.so"], stdout=subprocess.DEVNULL)
    data = H.as_u8(H.gen_motif(NCHUNK * CHUNK, 5000))
    hidx, cb = H.hash_index(), H.chunk_bytes(CHUNK)
    vr, ref, events = 0, [], []
    for k in range(NCHUNK):
        r = H.oracle_chunk(data[k * CHUNK:(k + 1) * CHUNK], level=LEVEL, cb=cb, vr_in=vr, hidx=hidx)
        ref.append((r["s0"], r["s1"], vr))
        events.append(r["stats"]["victim_events"])
        vr = r["vr_out"]
    assert sum(e > 0 for e in events) >= 2, "the case must exercise the carry"
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    with tempfile.TemporaryDirectory() as d:
        w = os.path.join(d, "worker.py")
        open(w, "w").write(WORKER)
        procs = []
        for rank in range(2):
            env = dict(os.environ, RANK=str(rank), WORLD_SIZE="2", MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
            procs.append(subprocess.Popen([sys.executable, w, ROOT, d, str(LEVEL), str(NCHUNK), str(CHUNK), mode], env=env))
        for p in procs:
            assert p.wait(timeout=600) == 0
        got, reruns = {}, 0
        for rank in range(2):
            r = pickle.load(open(os.path.join(d, "r%d.pkl" % rank), "rb"))
            assert r["vr"] == vr
            assert set(r["results"]) == {k for k in range(NCHUNK) if k % 2 == rank}
            got.update(r["results"])
            reruns += r["reruns"]
    for k in range(NCHUNK):
        assert got[k][2] == ref[k][2], "chunk %d ran with the wrong victim_round_in" % k
        assert got[k][0] == ref[k][0] and got[k][1] == ref[k][1], "chunk %d differs from the sequential oracle" % k
    if mode == "rerun":
        assert reruns >= 1 or all(x[2] == 0 for x in ref)
    else:
        assert reruns == 0
