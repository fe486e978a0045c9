// automaton.h — the descriptor of one automaton (a chunk, or one victim_round_in variant of a chunk) as the host
// hands it to k_automaton (automaton.cu); an array of these lives in device memory, one entry per cluster.
#pragma once
#include "mrz_common.cuh"

struct Shared;   // cluster-shared state in global memory (automaton.cu)

struct AutoDesc {
    const u8 *buf;        // chunk bytes (MRZ_PAD zero bytes either side)
    const u8 *bmap;       // trailing-ones map of the chunk (tagmap.cu), shared by the variants of a chunk
    const u8 *bsum;       // its 64:1 summary
    const i64 *hidx;      // hash_index[256] (rzip.c:669-673)
    i64 *ttag, *toff;     // the hash table: tags[] and offsets[]
    u64 *filt;            // tag filter
    u32 *gsum;            // group summary
    unsigned *claim;      // per-slot claims of the current batch: two arrays of claim_stride entries, used alternately
    i64 claim_stride;
    Token *toks;          // un-split match tokens (output)
    Shared *sh;
    AutoState *st;        // in: constants + initial state; out: final state
    i64 *dbg;             // MRZ_NDBG profile counters, or NULL
};
