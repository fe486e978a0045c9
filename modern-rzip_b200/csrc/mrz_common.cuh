// Shared device-side definitions for the B200 rzip pre-processor.
// Semantics follow /root/reference/src/rzip.c (cited per function); the code is a new design:
// a per-position "bitness" map instead of a per-byte loop, a thread-block cluster per chunk,
// token triples instead of byte streams.
#pragma once
#ifndef MRZ_CUSIM   // tests/cusim compiles these sources for the CPU emulator with -include cusim.h
#include <cuda_runtime.h>
#define MRZ_LAUNCH(kernel, grid, block, stream, ...) kernel<<<(grid), (block), 0, (stream)>>>(__VA_ARGS__)
#endif
#include <stdint.h>

typedef long long i64;
typedef unsigned long long u64;
typedef unsigned char u8;
typedef unsigned int u32;

#define MRZ_MIN_MATCH 31      // rzip.c:49 MINIMUM_MATCH
#define MRZ_GREAT_MATCH 1024  // rzip.c:48 GREAT_MATCH
#define MRZ_PAD 256           // zero padding before and after the chunk in HBM

// ---- tag map (tagmap.cu) ------------------------------------------------------------------------
// bmap[p] = number of trailing one bits of tag[p] (rzip.c:348-358) for p in [0, n-30), 0 beyond; a position passes
// the lookup gate `(t & mask) == mask` (rzip.c:573) for mask = 2^k-1 iff bmap[p] >= k.  bsum[b] = max of bmap over
// the 64 positions [64b, 64b+64).
#define MAP_RUN 32                             // positions per thread
#define MAP_THREADS 256
#define MAP_TILE (MAP_THREADS * MAP_RUN)       // positions per CTA step (8192)
#define MAP_BLK 64                             // positions per summary byte
#define MAP_BLK_SHIFT 6

// ---- automaton (automaton.cu) -------------------------------------------------------------------
#define MRZ_MAXCL 16          // CTAs per cluster (16 needs the non-portable cluster size attribute)
#define MRZ_MAXT 512
#define MRZ_MAXB (MRZ_MAXCL * MRZ_MAXT)   // ops per batch (8192); claim keys hold 14 bits of op index
#define MRZ_EQ_CAP (1 << 18)  // ring of sweep-eligible slots (power of two)
#define MRZ_WLOG_CAP 64
#define MRZ_NDBG 64

// One hash-table slot (rzip.c:59-62 struct hash_entry): all-zero == empty (rzip.c:230).  Stored as two arrays.
// Un-split match token; literals are the gaps between matches.
struct Token {
    i64 pos, len, ofs;
};

// Automaton state exchanged with the host (rzip_state, mrzip_private.h:388-416): the host fills the constants and
// the initial running state, the kernel writes the final one.
struct AutoState {
    // chunk constants
    i64 n, end;               // chunk size, n - 31
    i64 hsize, hmask, limit;  // table geometry (rzip.c:521-529)
    int chain;                // levels[].max_chain_len
    int done;                 // 1 when p reached end
    int error;                // non-zero: 2 = token buffer full, other = internal inconsistency (site id)
    int filt_bits;            // log2 of the tag filter's 64-bit words
    // running state
    i64 p;                    // last position examined (loop variable of rzip.c:548)
    i64 last;                 // st->last_match
    i64 cur_p, cur_len, cur_ofs;   // "current" best (rzip.c:511-515)
    i64 min_mask, ins_mask;   // st->minimum_tag_mask, local tag_mask (rzip.c:509)
    i64 clean_ptr, count;     // st->tag_clean_ptr, st->hash_count
    i64 victim_round;         // static in insert_hash (rzip.c:259)
    // counters
    i64 tag_hits, tag_misses, inserts, victim_events, n_ops;
    i64 ntok, tok_cap;
    i64 err_a, err_b;         // MRZ_CHECK: offending index / bound of a failed range check
};

struct EmitTotals {
    i64 s0_len, s1_len, matches, match_bytes, literals, literal_bytes;
};

__device__ __forceinline__ int tz1(i64 t) {  // trailing one bits (rzip.c:247-252 lesser_bitness via ffsll(~x))
    u64 x = ~(u64)t;
    return x ? (__ffsll((long long)x) - 1) : -1;
}

// Eight consecutive i64 (64-byte aligned) with two 256-bit loads (LDG.E.256 on sm_100a), L2-coherent.
__device__ __forceinline__ void ld_i64x8_cg(const i64 *p, i64 t[8]) {
#ifdef MRZ_CUSIM
    for (int q = 0; q < 8; q++) t[q] = ((const volatile i64 *)p)[q];
#else
    asm volatile("ld.global.cg.v4.u64 {%0,%1,%2,%3}, [%4];" : "=l"(t[0]), "=l"(t[1]), "=l"(t[2]), "=l"(t[3]) : "l"(p));
    asm volatile("ld.global.cg.v4.u64 {%0,%1,%2,%3}, [%4];" : "=l"(t[4]), "=l"(t[5]), "=l"(t[6]), "=l"(t[7]) : "l"(p + 4));
#endif
}

// 8 bytes starting at byte index idx (any alignment); buf has MRZ_PAD bytes of slack on both sides.
__device__ __forceinline__ u64 ld_u64_unaligned(const u8 *buf, i64 idx) {
    const u8 *a = buf + idx;
    uintptr_t ai = (uintptr_t)a;
    const u64 *q = (const u64 *)(ai & ~(uintptr_t)7);
    unsigned sh = (unsigned)(ai & 7) * 8;
    u64 lo = __ldg(q);
    if (sh == 0) return lo;
    u64 hi = __ldg(q + 1);
    return (lo >> sh) | (hi << (64 - sh));
}
