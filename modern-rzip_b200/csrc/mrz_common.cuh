// Shared device-side definitions for the B200 rzip pre-processor.
// Semantics follow /root/reference/src/rzip.c (cited per function); the code is a new design:
// candidate lists instead of a per-byte loop, a warp-wide probe of 32 slots, token triples.
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
#define MRZ_FILT_WORDS (1 << 24)   // tag filter: 2^24 64-bit words = 128 MiB (automaton.cu FILT_BITS)

#define SCAN_THREADS 256
#define SCAN_RUN 32                            // positions per thread
#define SCAN_TILE (SCAN_THREADS * SCAN_RUN)    // positions per CTA tile (8192)
#define SEG_MAX_TILES 8192                     // <= 64 Mi positions per segment
#define SEG_CAND_CAP (4 << 20)                 // flat candidate capacity per segment
#define SEG_CAND_TARGET (1 << 20)              // aim for this many candidates per segment

// One hash-table slot (rzip.c:59-62 struct hash_entry): all-zero == empty (rzip.c:230).
struct __align__(16) Slot {
    i64 off;
    i64 tag;
};

// Un-split match token; literals are the gaps between matches.
struct Token {
    i64 pos, len, ofs;
};

// Automaton state, resident in HBM between kernels (rzip_state, mrzip_private.h:388-416).
struct AutoState {
    // chunk constants
    i64 n, end;               // chunk size, n - 31
    i64 hsize, hmask, limit;  // table geometry (rzip.c:521-529)
    int chain;                // levels[].max_chain_len
    int done;                 // 1 when p reached end
    int error;                // non-zero: internal inconsistency / token overflow
    int pad0;
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
    // current segment (written by the tag-scan kernels)
    i64 seg_lo, seg_hi;       // candidate positions cover [seg_lo, seg_hi)
    i64 seg_mask;             // lookup mask the list was filtered with
    i64 seg_ncand;            // candidates in the flat list
    int seg_tiles;            // tiles covered
    int pad1;
};

struct EmitTotals {
    i64 s0_len, s1_len, matches, match_bytes, literals, literal_bytes;
};

__device__ __forceinline__ int tz1(i64 t) {  // trailing one bits (rzip.c:247-252 lesser_bitness via ffsll(~x))
    u64 x = ~(u64)t;
    return x ? (__ffsll((long long)x) - 1) : -1;
}

__device__ __forceinline__ Slot ld_slot(const Slot *p) {
    longlong2 v = __ldcg(reinterpret_cast<const longlong2 *>(p));
    Slot s;
    s.off = v.x;
    s.tag = v.y;
    return s;
}
__device__ __forceinline__ void st_slot(Slot *p, i64 off, i64 tag) {
    longlong2 v;
    v.x = off;
    v.y = tag;
    __stcg(reinterpret_cast<longlong2 *>(p), v);
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
