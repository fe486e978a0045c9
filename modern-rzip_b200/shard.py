"""Chunk sharding across ranks (one process per GPU) with the exact victim_round carry.

rzip chunks are independent except for one scalar: the function-local `static victim_round` of insert_hash
(/root/reference/src/rzip.c:259), which survives from chunk to chunk and is output-relevant as soon as a chunk
performs a victim eviction.  Protocol (no data-path collective; three tiny control all-gathers):

  1. chunk k belongs to rank k mod W; every rank runs its chunks speculatively with victim_round_in = 0;
  2. all ranks all-gather (victim_events, victim_round_out) per chunk;
  3. every rank walks the chain 0 -> vr_out(chunk 0) -> ... identically: a chunk without victim events passes
     the value through and its result is valid for any input; a chunk whose speculated input equals the
     chained value is valid; otherwise its owner re-runs it with the right input and the new output is
     all-gathered before the walk continues.

The engine is any callable run(chunk_bytes, vr_in) -> (victim_events, vr_out, payload); bench.py passes the
CUDA context, tests/test_shard_gloo.py the CPU-emulator build of the same kernels.
"""
import torch
import torch.distributed as dist


def owner(k, world):
    return k % world


def my_chunks(n_chunks, rank, world):
    return [k for k in range(n_chunks) if owner(k, world) == rank]


def _gather_pairs(local, n_chunks, device):
    """local: {chunk: (events, vr_out)} -> full list for all chunks (MAX-reduce of a -1-filled table)."""
    t = torch.full((n_chunks, 2), -1, dtype=torch.int64, device=device)
    for k, (ev, vo) in local.items():
        t[k, 0] = ev
        t[k, 1] = vo
    if dist.is_initialized() and dist.get_world_size() > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return [(int(a), int(b)) for a, b in t.tolist()]


def run_sharded(chunks, run, rank=0, world=1, device="cpu", vr_start=0):
    """chunks: list of (k, bytes) for ALL k this rank owns (others may be None placeholders by index).
    Returns ({k: payload} for owned chunks, final victim_round, number of re-runs this rank did)."""
    n_chunks = len(chunks)
    mine = [k for k in range(n_chunks) if owner(k, world) == rank]
    results, local = {}, {}
    for k in mine:
        ev, vo, payload = run(chunks[k], 0)
        results[k] = payload
        local[k] = (ev, vo)
    table = _gather_pairs(local, n_chunks, device)
    spec_in = [0] * n_chunks
    vr, reruns = vr_start, 0
    for k in range(n_chunks):
        ev, vo = table[k]
        if ev == 0:
            continue                      # independent of its input; value passes through
        if spec_in[k] != vr:
            upd = {}
            if owner(k, world) == rank:
                ev2, vo2, payload = run(chunks[k], vr)
                results[k] = payload
                upd[k] = (ev2, vo2)
                reruns += 1
            fresh = _gather_pairs(upd, n_chunks, device)
            ev, vo = fresh[k]
            table[k] = (ev, vo)
            spec_in[k] = vr
        vr = vo
    return results, vr, reruns
