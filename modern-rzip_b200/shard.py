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

`run_sharded_variants` is the variant-parallel form of the same chain (SURVEY.md §8e option 1): every rank runs each
of its chunks for ALL max_chain_len values of victim_round_in side by side on its GPU (mrz_run_variants: one CTA and
one table per variant), ONE all-gather carries the (chunk, vr_in) -> vr_out table, every rank walks the chain and
emits only the chosen variant of its own chunks.  No chunk ever waits for another and none is re-run.

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


def run_sharded_variants(chunks, run_variants, n_variants, rank=0, world=1, device="cpu", vr_start=0):
    """run_variants(chunk) -> [vr_out for vr_in in range(n_variants)] (the automata of all variants, no emission).
    Returns ({k: chosen vr_in} for the chunks this rank owns, final victim_round)."""
    n_chunks = len(chunks)
    t = torch.full((n_chunks, n_variants), -1, dtype=torch.int64, device=device)
    for k in range(n_chunks):
        if owner(k, world) == rank:
            outs = run_variants(chunks[k])
            assert len(outs) == n_variants
            t[k] = torch.tensor([int(x) for x in outs], dtype=torch.int64, device=device)
    if dist.is_initialized() and dist.get_world_size() > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)      # rows are disjoint: MAX over a -1-filled table is an all-gather
    table = t.tolist()
    chosen, vr = {}, vr_start
    for k in range(n_chunks):
        if owner(k, world) == rank:
            chosen[k] = vr
        vr = int(table[k][vr])
        assert 0 <= vr < n_variants
    return chosen, vr
