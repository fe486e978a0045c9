// automaton.cu — the table automaton of hash_search, exact, as ONE persistent launch of thread-block clusters.
//
// Reproduces, operation for operation, the table semantics of src/rzip.c:
//   find_best_match (rzip.c:426-462), insert_hash (rzip.c:256-301), clean_one_from_hash
//   (rzip.c:305-328), single_match_len (rzip.c:372-397) and the greedy/lazy emission rule of
//   hash_search (rzip.c:571-599).
//
// B200 design (not the reference's loop):
//   * ONE LAUNCH, MANY AUTOMATA, NO HOST IN THE LOOP.  The grid is (automata) x (CTAs per cluster); automaton a —
//     a chunk, or one victim_round_in variant of a chunk — is cluster a and walks its whole chunk before it
//     returns.  A cluster is CL = 1..16 CTAs of T threads; a batch is B = CL*T candidate ops, one per thread.
//     One SM can only keep ~1 uncoalesced request per clock in flight, which is what bounded the one-CTA design;
//     a cluster spreads every phase of a batch over CL SMs and pays ~400 clk per cluster barrier
//     (barrier.cluster, hardware) instead of ~4000 for a grid barrier through L2.  The state the CTAs share
//     (running state, batch arrays, sweep queue) lives in a per-automaton block of global memory that stays in L2;
//     every field has ONE writer between two barriers.
//   * the per-byte loop is gone: candidates come from the bitness map of tagmap.cu.  Each batch the cluster reads
//     the 64:1 summary of the next stretch of the chunk, compacts the non-empty blocks, reads those 64 map bytes
//     into a bit mask and compacts the positions (two cluster-wide prefix sums); the 64-bit tag of a candidate is
//     recomputed from the bytes (31 look-ups);
//   * the table is stored as two arrays (tags[], offs[]): walks only ever look at tags;
//   * BATCHED SPECULATION: the B ops are *planned* in parallel against the live table read-only (insert walk,
//     one level of displacement) and committed in order; the first op that cannot commit from its plan cuts the
//     batch.  If all it suffers from is a conflict, everything behind the cut that lost its plan is simply planned
//     AGAIN against the updated table and the batch goes on (a round costs a few barriers); only ops that need a
//     lookup, a deep displacement, an emission or a serial sweep step run on the exact executor (below);
//   * EXACT CONFLICT RULE: an insert only ever RAISES the trailing-one count of the slot it writes
//     (empty / due-for-sweep -> >= need, lesser -> greater), so a walk that passed a slot still passes it: only a
//     write to a walk's own STOP slot can change its outcome, and an entry due for sweeping stops every walk, so a
//     slot of the sweep queue can only be a stop slot too.  Plans CLAIM their stop slots with
//     atomicMin((epoch, op index)) in a 16 MiB array (no hash, no clearing); a plan is invalid iff an earlier op
//     claims one of its stop slots;
//   * FORWARDING: when the earlier op's tag is no stop for the later walk (and not equal to its tag), the later
//     plan walks on from the next slot, re-claims and remembers the dependency; iterated to a fixed point;
//   * TAG FILTER: a blocked Bloom filter over every tag inserted in this chunk (no false negatives).  A negative
//     answer proves that no entry with an equal tag exists, so find_best_match's walk to the end of the probe
//     cluster is skipped altogether, and lookup-only ops touch no table slot;
//   * GROUP SUMMARY: per group of 8 slots a 16-bit thermometer of the minimum trailing-one count (L2 resident);
//     an insert walk skips 64 slots per 128-bit summary load; lowering is one atomicAnd, repair one atomicOr;
//   * TWINS: positions p and p+1 with equal tags (buf[p] == buf[p+31]) are planned as a pair;
//   * the sweep of clean_one_from_hash is pre-resolved: a parallel scan keeps a ring of the next eligible slots, so
//     the n-th overflow of a batch zeroes the n-th queued slot.  A plan that overwrites a queued slot conflicts
//     only when the sweep can reach that slot within this batch; otherwise it commits and the ring is cut there;
//   * THE EXACT EXECUTOR (rank 0 of the cluster): the probe sequence of find_best_match is walked 4*T slots per
//     step, equal-tag candidates are compared by the warps in parallel (__ballot_sync / __ffs memcmp), long
//     matches are extended by all threads; the insert walks are warp-cooperative.
//   oracle/spec_model.c is the CPU model of the speculation rules; tests/test_sim_parity.py runs this very file on
//   the CPU emulator (clusters included), tests/test_gpu_parity.py on the B200.
#ifndef MRZ_CUSIM
#include <mutex>
#endif
#include <string.h>
#include <stddef.h>
#include "mrz_common.cuh"
#include "automaton.h"

#define FULL 0xffffffffu
#define LK_XCAP 4096       // a candidate still matching after this many bytes is extended by the whole CTA
#define LK_XCAP2 (1 << 20)  // ... unless more than two candidates are that long: then the warps go on in parallel up to here
#define LK_CAP 256         // equal-tag entries the CTA-wide lookup collects; more: the warp walk
#define FWD_ITERS 6        // forwarding rounds per conflict phase; later collisions are planned again after the cut
#define MAXWALK 16384      // longer insert walks are left to the exact executor
#define GRP 8              // slots per summary group (= one 64-byte read of tags)
#define GRP_SHIFT 3
#define DUP_SALT 0x5bd1e9955bd1e995LL

enum { F_COMPLEX = 1, F_INSERT = 2, F_W1 = 4, F_CONF = 8, F_FOLLOW = 16, F_W2ELIG = 32, F_W1ELIG = 64,
       F_ZONE = 128 };   // F_ZONE: the conflict is with the sweep, not with another plan: planning again cannot resolve it
#define MAXDEP 3            // claimed slots a plan may walk past (forwarding)
#define IDX_BITS 14         // MRZ_MAXB = 2^14 ops per batch
#ifdef MRZ_CUSIM
#define CLAIM_EPOCH_MAX 700u       // the CPU-emulator tests exercise the wrap-around (wipe + restart) path
#else
#define CLAIM_EPOCH_MAX 0x1FFF0u
#endif
#define EPOCH_TOP 0x1FFFFu
__device__ __forceinline__ unsigned claim_key(unsigned epoch, int j, int is_w1) {
    return ((EPOCH_TOP - epoch) << (IDX_BITS + 1)) | ((unsigned)j << 1) | (unsigned)is_w1;
}
#define KEY_IDX(o) ((int)(((o) >> 1) & (unsigned)(MRZ_MAXB - 1)))
#define KEY_EPOCH(o) ((o) >> (IDX_BITS + 1))

// ---- range checks of the -DMRZ_CHECK build: a violation records (site, index, bound) and clamps the index ----
#ifdef MRZ_CHECK
#define CHK(sh, x, site, idx, bound)                                                                 \
    do {                                                                                             \
        if ((i64)(idx) < 0 || (i64)(idx) >= (i64)(bound)) {                                          \
            (sh)->err_a = (i64)(idx); (sh)->err_b = (i64)(bound); __stcg(&(sh)->err_site, 100 + (site)); \
            (idx) = 0;                                                                               \
        }                                                                                            \
    } while (0)
#else
#define CHK(sh, x, site, idx, bound) do { } while (0)
#endif

// Blocked Bloom filter over inserted tags: word index and 3-bit mask (oracle/spec_model.c filt_hash).
struct Filt {
    u64 *w;
    int shift;   // 64 - log2(words)
};
__device__ __forceinline__ void filt_hash(const Filt &F, i64 tag, u64 &word, u64 &bits) {
    u64 x = (u64)tag;
    x ^= x >> 33; x *= 0xff51afd7ed558ccdULL; x ^= x >> 33; x *= 0xc4ceb9fe1a85ec53ULL; x ^= x >> 33;
    word = x >> F.shift;
    bits = ((u64)1 << (x & 63)) | ((u64)1 << ((x >> 6) & 63)) | ((u64)1 << ((x >> 12) & 63));
}
__device__ __forceinline__ bool filt_query(const Filt &F, i64 tag) {
    u64 w, b;
    filt_hash(F, tag, w, b);
    return (__ldcg(F.w + w) & b) == b;
}
__device__ __forceinline__ void filt_set(const Filt &F, i64 tag) {
    u64 w, b;
    filt_hash(F, tag, w, b);
    atomicOr((unsigned long long *)(F.w + w), (unsigned long long)b);
}

__device__ __forceinline__ int tz1c(i64 t) {   // trailing one bits, capped at 15 (summary thermometer)
    int z = tz1(t);
    return z > 15 ? 15 : (z < 0 ? 15 : z);
}
// GROUP SUMMARY, in global memory (L2 resident, 1 MiB for the 64 MiB table): per group of 8 slots a 16-bit
// thermometer - bit i set <=> every tag of the group has more than i trailing one bits (as far as is known: the
// code is a lower bound, all-zero = nothing known).  Lowering is one atomicAnd, raising one atomicOr: both fire
// and forget, so plans and commits on any SM maintain it without read-modify-write loops.
__device__ __forceinline__ void sum_raise(u32 *gs, i64 g, int v) {      // plan phases: the whole group was seen, min = v
    if (v > 0) atomicOr(gs + (g >> 1), (((1u << v) - 1u) & 0x7fffu) << ((g & 1) * 16));
}
__device__ __forceinline__ void sum_lower(u32 *gs, i64 g, int v) {      // commit / serial phases: a tag with v trailing ones (0: zeroed)
    atomicAnd(gs + (g >> 1), ~(0xffffu << ((g & 1) * 16)) | ((((1u << v) - 1u) & 0x7fffu) << ((g & 1) * 16)));
}

// ---- cluster plumbing ----------------------------------------------------------------------------
// The CTAs of a cluster share nothing through global memory that is hot: every CTA keeps a REPLICA of the
// automaton's scalar state in its own shared memory.  A replica changes in one of two ways only: (1) every CTA
// applies the same deterministic update from the same inputs (commit, gather, ring refill), or (2) the one CTA that
// ran the exact executor PUSHES the result into every replica through distributed shared memory before the next
// cluster barrier (remote stores are fire-and-forget; the readers then read their own shared memory, ~30 clk,
// instead of L2, ~375 clk, or a volatile global load, ~745 clk: profiles/r2_micro_lat.txt).  Per-rank values that
// everybody needs (scan totals, local minima) are pushed by each rank into slot [rank] of every replica.  A slot
// that is rewritten in CONSECUTIVE barrier intervals is double-buffered: a CTA that has passed barrier k may write
// before a slower CTA has read what it received before barrier k.
#ifdef MRZ_CUSIM
__device__ __forceinline__ int cl_rank() { return cusim::cluster_rank(); }
__device__ __forceinline__ int cl_size() { return cusim::cluster_size(); }
__device__ __forceinline__ void csync(int) {
    __syncthreads();
    if (threadIdx.x == 0) cusim::sync_cluster();
    __syncthreads();
}
template <class T> __device__ __forceinline__ T *remote_ptr(T *p, int r) { return reinterpret_cast<T *>(cusim::map_shared_rank((void *)p, r)); }
__device__ __forceinline__ void prefetch_l2(const void *) {}
#else
__device__ __forceinline__ int cl_rank() { unsigned r; asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r)); return (int)r; }
__device__ __forceinline__ int cl_size() { unsigned r; asm volatile("mov.u32 %0, %%cluster_nctarank;" : "=r"(r)); return (int)r; }
// Barrier over all threads of the cluster; orders their global and (distributed) shared memory accesses.
__device__ __forceinline__ void csync(int CL) {
    if (CL == 1) { __syncthreads(); return; }
    asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// generic address of the same shared-memory object in the CTA of rank r
template <class T> __device__ __forceinline__ T *remote_ptr(T *p, int r) {
    u64 out;
    asm volatile("mapa.u64 %0, %1, %2;" : "=l"(out) : "l"((u64)(uintptr_t)p), "r"(r));
    return reinterpret_cast<T *>((uintptr_t)out);
}
__device__ __forceinline__ void prefetch_l2(const void *p) { asm volatile("prefetch.global.L2 [%0];" ::"l"(p)); }
#endif
// store v into slot *p of the replica of rank r
template <class T> __device__ __forceinline__ void push_to(T *p, int r, T v) { *(volatile T *)remote_ptr(p, r) = v; }

// Running state of the automaton (rzip_state, mrzip_private.h:388-416).
struct Run {
    i64 p, last, cur_p, cur_len, cur_ofs, min_mask, ins_mask, clean_ptr, count, vr;
    i64 hits, misses, inserts, vevents, nops, ntok;
    int error, pad;
};

#define NSITE 4   // cluster-wide prefix sums per batch, each with its own exchange slots
// What the exact executor (rank 0) changes; pushed as one block of words into every replica.
struct XRes {
    Run run;
    i64 gp;                  // gather cursor: every position < gp has been gathered or skipped
    i64 e_scan;              // slots in [run.clean_ptr, e_scan) have been examined for the ring
    unsigned e_idx, e_cnt;   // ring counters (monotonic)
    int ended, jumped, nlog, pos_after;
    int wlog[MRZ_WLOG_CAP];  // (slot << 2) | (overwrote an entry that was due for sweeping) << 1 | zeroed
};
// Replica of the automaton's scalars, one per CTA, in shared memory.
struct Rep {
    XRes x;
    int exhausted, pad;
    unsigned long long tot[NSITE][MRZ_MAXCL];   // scan totals of rank r
    unsigned long long lcut[2][MRZ_MAXCL];      // first op of rank r that cannot commit from its plan: idx << 1 | exact (0x7fffffff: none)
    unsigned long long lcutpre[MRZ_MAXCL];      // ... and the prefix sum over the ops of rank r before it
    i64 xmin[MRZ_MAXCL];                        // gather: first candidate of rank r that did not fit the batch
    int qmin[MRZ_MAXCL];                        // commit: smallest ring slot a plan of rank r overwrote
    unsigned chg[2][MRZ_MAXCL];                 // conflict phase: epoch for which rank r reported a change (double-buffered by iteration)
};

// Per-op arrays and the sweep ring, in global memory (L2); read with ld.cg, one writer per element between barriers.
struct Shared {
    i64 err_a, err_b;
    int err_site, pad;       // MRZ_CHECK build: site of a failed range check (polled by every CTA)
    int E[MRZ_EQ_CAP];       // ring of eligible slots, ascending
    i64 b_pos[MRZ_MAXB], b_tag[MRZ_MAXB];   // batch: position / tag of op j
    int pl_fl[MRZ_MAXB], pl_w2[MRZ_MAXB];   // plan flags / insert slot of op j (twins, commit)
    unsigned long long pre[MRZ_MAXB];       // inclusive prefix over ops holding a valid plan
};
#define LDG_(x) __ldcg(&(x))
#define STG_(x, v) __stcg(&(x), (v))

#define PK_D(x) ((unsigned)((x) & 0xffffu))            // hash_count delta
#define PK_I(x) ((unsigned)(((x) >> 16) & 0xffffu))    // inserts
#define PK_M(x) ((unsigned)(((x) >> 32) & 0xffffu))    // tag misses (twins)
#define PK_V(x) ((unsigned)(((x) >> 48) & 0xffffu))    // ops

extern "C" size_t mrz_shared_bytes() { return sizeof(Shared); }

// ---- shared memory of a CTA ----------------------------------------------------------------------
struct Smem {
    Rep rep;
    i64 hx[256];
    i64 stk[3 * 64];
    unsigned long long wsum[40];
    unsigned long long cbase, ctotal;
    i64 lgp;                // CTA-local minimum for xmin
    int lk_fe, lk_fe2, lk_n, lmin, lq, lchg;
    int lk_pos[LK_CAP];     // CTA-wide lookup: positions (relative to the primary slot) of equal tags, ascending
    i64 lk_len[LK_CAP], lk_rev[LK_CAP], lk_off[LK_CAP];   // match_len results per candidate
    i64 lk_f[LK_CAP], lk_r[LK_CAP];                       // capped forward / backward lengths (first pass)
};
extern "C" size_t mrz_automaton_smem_bytes() { return sizeof(Smem); }

#ifdef MRZ_CUSIM
// Emulator build only (tests/cusim): self-checks of the automaton's bookkeeping against the table itself, once per batch.
// tests/test_sim_parity.py turns them on and asserts that none fired.
static bool g_selfcheck_on = false;
static int g_selfcheck_failures = 0;
static unsigned g_selfcheck_tick = 0;
static unsigned g_selfcheck_mask = 63;      // mrz_sim_selfcheck(1): every 64th batch; (2): every batch
extern "C" void mrz_sim_selfcheck(int on) { g_selfcheck_on = on != 0; g_selfcheck_mask = on >= 2 ? 0 : 63; g_selfcheck_failures = 0; }
extern "C" int mrz_sim_selfcheck_failures() { return g_selfcheck_failures; }
#define MRZ_SMEM_PTR(T, name) T *name = reinterpret_cast<T *>(cusim::dyn_smem())
#else
#define MRZ_SMEM_PTR(T, name) extern __shared__ __align__(16) unsigned char name##_raw[]; T *name = reinterpret_cast<T *>(name##_raw)
#endif

// exclusive prefix sum of v over the CTA (thread order); total in `total`.  Three block barriers.
__device__ __forceinline__ unsigned long long block_excl_scan(unsigned long long v, unsigned long long *wsum, unsigned long long &total) {
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
    unsigned long long inc = v;
    for (int o = 1; o < 32; o <<= 1) {
        unsigned long long u = __shfl_up_sync(FULL, inc, o);
        if (lane >= o) inc += u;
    }
    if (lane == 31) wsum[wid] = inc;
    __syncthreads();
    if (wid == 0) {
        unsigned long long w = lane < nw ? wsum[lane] : 0, wi = w;
        for (int o = 1; o < 32; o <<= 1) {
            unsigned long long u = __shfl_up_sync(FULL, wi, o);
            if (lane >= o) wi += u;
        }
        wsum[lane] = wi - w;
        if (lane == 31) wsum[32] = wi;
    }
    __syncthreads();
    total = wsum[32];
    const unsigned long long r = wsum[wid] + inc - v;
    __syncthreads();   // wsum may be reused right away
    return r;
}

// exclusive prefix sum of v over the CLUSTER (rank-major, then thread order); total in `total`.  ONE cluster barrier:
// every rank pushes its total into slot [rank] of every replica, then sums the slots below its own.
// `extra(ex)` is called by all threads between the push and the barrier, with the thread's exclusive prefix inside
// its CTA: whatever else has to cross the cluster in the same hop.
template <class F>
__device__ __forceinline__ unsigned long long cluster_excl_scan(unsigned long long v, Smem *S, int site, int CL, int rank,
                                                                unsigned long long &total, F extra) {
    unsigned long long btot;
    const unsigned long long ex = block_excl_scan(v, S->wsum, btot);
    if ((int)threadIdx.x < CL) push_to(&S->rep.tot[site][rank], (int)threadIdx.x, btot);
    extra(ex);
    csync(CL);
    if (threadIdx.x < 32) {
        const int lane = threadIdx.x;
        const unsigned long long t = lane < CL ? S->rep.tot[site][lane] : 0ULL;
        unsigned long long below = lane < rank ? t : 0ULL, all = t;
        for (int o = 16; o; o >>= 1) { below += __shfl_xor_sync(FULL, below, o); all += __shfl_xor_sync(FULL, all, o); }
        if (lane == 0) { S->cbase = below; S->ctotal = all; }
    }
    __syncthreads();
    total = S->ctotal;
    const unsigned long long r = S->cbase + ex;
    __syncthreads();
    return r;
}
struct NoExtra { __device__ __forceinline__ void operator()(unsigned long long) const {} };

// ---- byte compares with aligned 128-bit loads ---------------------------------------------------------------------
// match_len (rzip.c:372-397) compares two byte streams of the chunk from an arbitrary pair of positions.  Stream A is
// read on its own 16-byte grid (the grid starts up to 15 bytes before the first byte that counts; those bytes are
// masked), stream B wherever that puts it: two aligned vectors and a funnel shift per unit.  A thread that handles G
// consecutive units issues 2G+1 LDG.128 for 16G bytes (the 8-byte unaligned compare it replaces: 4 LDG.64 per 8 bytes).
__device__ __forceinline__ void funnel16(const uint4 v0, const uint4 v1, int sb, u64 &lo, u64 &hi) {
    u64 w0 = (u64)v0.x | ((u64)v0.y << 32), w1 = (u64)v0.z | ((u64)v0.w << 32);
    u64 w2 = (u64)v1.x | ((u64)v1.y << 32), w3 = (u64)v1.z | ((u64)v1.w << 32);
    if (sb >= 8) { w0 = w1; w1 = w2; w2 = w3; sb -= 8; }
    const int s = sb * 8;
    if (s == 0) { lo = w0; hi = w1; }
    else { lo = (w0 >> s) | (w1 << (64 - s)); hi = (w1 >> s) | (w2 << (64 - s)); }
}
// XOR of units [u, u+G) of A (Abase 16-aligned) and B (Bbase any alignment): x[2c] = low 8 bytes, x[2c+1] = high 8 bytes
template <int G>
__device__ __forceinline__ void xor_units(const u8 *Abase, const u8 *Bbase, i64 u, u64 (&x)[2 * G]) {
    const uint4 *pa = reinterpret_cast<const uint4 *>(Abase + 16 * u);
    const uintptr_t bb = (uintptr_t)(Bbase + 16 * u);
    const uint4 *pb = reinterpret_cast<const uint4 *>(bb & ~(uintptr_t)15);
    const int sb = (int)(bb & 15);
    uint4 va[G], vb[G + 1];
#pragma unroll
    for (int c = 0; c < G; c++) va[c] = __ldg(pa + c);
#pragma unroll
    for (int c = 0; c <= G; c++) vb[c] = __ldg(pb + c);
#pragma unroll
    for (int c = 0; c < G; c++) {
        u64 bl, bh;
        funnel16(vb[c], vb[c + 1], sb, bl, bh);
        x[2 * c] = ((u64)va[c].x | ((u64)va[c].y << 32)) ^ bl;
        x[2 * c + 1] = ((u64)va[c].z | ((u64)va[c].w << 32)) ^ bh;
    }
}
// One thread's share of a compare step.  A step covers S units (16 S bytes); this thread handles units
// [step * S + tu * G, + G).  Forward: the streams start at a0 / b0; returns the byte index (relative to the step's first
// unit) of the thread's first mismatch at or behind a0, or 0x7fffffff.  `skip` = a0 & 15 bytes before a0 on the grid.
#define CMP_G 2
#define CMP_LONG 4          // steps per round trip of a long compare (cta_cmp2) ...
#define CMP_LONG_AFTER 16   // ... once this many single steps (64 KiB at 256 threads) have matched
template <bool FWD>
__device__ __forceinline__ int cmp_thread(const u8 *buf, i64 a0, i64 b0, i64 maxlen, i64 step, int S, int tu) {
    const i64 u = step * S + (i64)tu * CMP_G;             // first unit of this thread, counted from the grid origin
    u64 x[2 * CMP_G];
    if (FWD) {
        const int skip = (int)((uintptr_t)(buf + a0) & 15);
        if (16 * u - skip >= maxlen) return 0x7fffffff;   // nothing of these units counts
        xor_units<CMP_G>(buf + a0 - skip, buf + b0 - skip, u, x);
        if (u == 0 && skip) {                             // bytes before a0
            if (skip >= 8) { x[0] = 0; x[1] &= ~(((u64)1 << ((skip - 8) * 8)) - 1); }
            else x[0] &= ~(((u64)1 << (skip * 8)) - 1);
        }
        int first = 0x7fffffff;
#pragma unroll
        for (int c = 2 * CMP_G - 1; c >= 0; c--)
            if (x[c]) first = (tu * CMP_G * 16) + c * 8 + ((__ffsll((long long)x[c]) - 1) >> 3);
        return first;
    } else {
        // backwards: the grid origin is the first 16-byte boundary at or behind a0; unit u covers the 16 bytes that end
        // 16 u bytes before it; `skip` bytes between a0 and the origin do not count
        const int skip = (int)((16 - ((uintptr_t)(buf + a0) & 15)) & 15);
        if (16 * u - skip >= maxlen) return 0x7fffffff;
        // units are stored descending in memory: the thread's G units are [origin - 16 (u + G), origin - 16 u)
        xor_units<CMP_G>(buf + a0 + skip - 16 * (u + CMP_G), buf + b0 + skip - 16 * (u + CMP_G), 0, x);
        // x[2 * CMP_G - 1] holds the bytes nearest to the origin
        if (u == 0 && skip) {                             // bytes at or behind a0
            if (skip > 8) { x[2 * CMP_G - 1] = 0; x[2 * CMP_G - 2] &= (((u64)1 << ((16 - skip) * 8)) - 1); }
            else if (skip == 8) x[2 * CMP_G - 1] = 0;
            else x[2 * CMP_G - 1] &= (((u64)1 << ((8 - skip) * 8)) - 1);
        }
        int first = 0x7fffffff;                           // distance from the step's origin of the nearest mismatch
#pragma unroll
        for (int c = 0; c < 2 * CMP_G; c++)
            if (x[c]) first = (tu * CMP_G * 16) + (2 * CMP_G - 1 - c) * 8 + (__clzll((long long)x[c]) >> 3);
        return first;
    }
}

// The same compare by ONE warp (many candidates: a warp each): 1 KiB per step and direction.
// Longest common prefix of buf[a0..] / buf[b0..] (FWD) or longest common suffix of buf[..a0) / buf[..b0), at most maxlen.
template <bool FWD>
__device__ __forceinline__ i64 warp_extend(const u8 *buf, i64 a0, i64 b0, i64 maxlen, int lane) {
    if (maxlen <= 0) return 0;
    const int skip = FWD ? (int)((uintptr_t)(buf + a0) & 15) : (int)((16 - ((uintptr_t)(buf + a0) & 15)) & 15);
    const int S = 32 * CMP_G;
    for (i64 step = 0;; step++) {
        int first = cmp_thread<FWD>(buf, a0, b0, maxlen, step, S, lane);
        for (int o = 16; o; o >>= 1) {
            const int u = __shfl_xor_sync(FULL, first, o);
            first = u < first ? u : first;
        }
        if (first != 0x7fffffff) {
            const i64 k = 16 * step * S + first - skip;
            return k > maxlen ? maxlen : k;
        }
        if (16 * (step + 1) * S - skip >= maxlen) return maxlen;
    }
}
// Both directions, the first step of each in ONE round trip.
__device__ __forceinline__ void warp_cmp2(const u8 *buf, i64 a0, i64 b0, i64 maxf, i64 maxb, int lane, i64 &f, i64 &r) {
    const int S = 32 * CMP_G;
    const int skf = (int)((uintptr_t)(buf + a0) & 15), skb = (int)((16 - ((uintptr_t)(buf + a0) & 15)) & 15);
    int ff = maxf > 0 ? cmp_thread<true>(buf, a0, b0, maxf, 0, S, lane) : 0x7fffffff;
    int fr = maxb > 0 ? cmp_thread<false>(buf, a0, b0, maxb, 0, S, lane) : 0x7fffffff;
    for (int o = 16; o; o >>= 1) {
        const int u = __shfl_xor_sync(FULL, ff, o), v = __shfl_xor_sync(FULL, fr, o);
        ff = u < ff ? u : ff;
        fr = v < fr ? v : fr;
    }
    if (maxf <= 0) f = 0;
    else if (ff != 0x7fffffff) { f = ff - skf; if (f > maxf) f = maxf; }
    else if (16 * S - skf >= maxf) f = maxf;
    else { const i64 done = 16 * S - skf; f = done + warp_extend<true>(buf, a0 + done, b0 + done, maxf - done, lane); }
    if (maxb <= 0) r = 0;
    else if (fr != 0x7fffffff) { r = fr - skb; if (r > maxb) r = maxb; }
    else if (16 * S - skb >= maxb) r = maxb;
    else { const i64 done = 16 * S - skb; r = done + warp_extend<false>(buf, a0 - done, b0 - done, maxb - done, lane); }
}

// ---- the exact executor: one warp runs one op on the live table with the reference semantics -----
struct Warp {
    // constants
    const u8 *buf;
    i64 *ttag, *toff;   // the table: tags and offsets (struct hash_entry, rzip.c:59-62, split in two arrays)
    Filt filt;
    u32 *gsum;          // group summary (global memory)
    Token *toks;
    i64 n, end, hmask, hsize, limit, tok_cap;
    int chain, lane;
    // running state (identical in all lanes)
    i64 p, last, cur_p, cur_len, cur_ofs, min_mask, ins_mask, clean_ptr, count, vr;
    i64 hits, misses, inserts, vevents, nops, ntok;
    int error;
    // write log of the current op: (slot << 1) | zeroed; nlog > MRZ_WLOG_CAP means overflow
    int *wlog;
    int nlog;
    i64 clean_hint;   // no eligible slot exists in [clean_ptr, clean_hint)
    i64 *prof;        // profile build: cycles / calls of the parts of a serial op

    __device__ __forceinline__ void load(const Run &r) {
        p = r.p; last = r.last; cur_p = r.cur_p; cur_len = r.cur_len; cur_ofs = r.cur_ofs;
        min_mask = r.min_mask; ins_mask = r.ins_mask; clean_ptr = r.clean_ptr; count = r.count; vr = r.vr;
        hits = r.hits; misses = r.misses; inserts = r.inserts; vevents = r.vevents; nops = r.nops; ntok = r.ntok;
        error = r.error;
        __syncwarp();   // every lane holds the state before lane 0 may store the next one
    }
    __device__ __forceinline__ void store(Run &r) const {
        if (lane == 0) {
            r.p = p; r.last = last; r.cur_p = cur_p; r.cur_len = cur_len; r.cur_ofs = cur_ofs;
            r.min_mask = min_mask; r.ins_mask = ins_mask; r.clean_ptr = clean_ptr; r.count = count; r.vr = vr;
            r.hits = hits; r.misses = misses; r.inserts = inserts; r.vevents = vevents; r.nops = nops; r.ntok = ntok;
            if (error) r.error = error;
        }
        __syncwarp();
    }
    // table writes of the serial path: lane 0 stores, the write is logged, the summary follows
    // elig: the slot held an entry due for sweeping (rzip.c:267-269) - the only kind of slot the sweep ring can hold
    __device__ __forceinline__ void write_slot(i64 slot, i64 off, i64 tag, bool elig = false) {
        if (lane == 0) {
            __stcg(toff + slot, off);
            __stcg(ttag + slot, tag);
            sum_lower(gsum, slot >> GRP_SHIFT, tag ? tz1c(tag) : 0);
            if (nlog < MRZ_WLOG_CAP) wlog[nlog] = (int)(slot << 2) | (elig ? 2 : 0) | (tag ? 0 : 1);
        }
        nlog++;
        __syncwarp();
    }

    // rzip.c:372-397.  Longest common prefix of buf[a0..] / buf[b0..] (fwd) or longest common suffix of buf[..a0) /
    // buf[..b0), at most maxlen bytes; 1 KiB per step, aligned 128-bit loads (cmp_thread).
    __device__ __forceinline__ i64 extend(i64 a0, i64 b0, i64 maxlen, bool fwd) {
        return fwd ? warp_extend<true>(buf, a0, b0, maxlen, lane) : warp_extend<false>(buf, a0, b0, maxlen, lane);
    }
    __device__ __forceinline__ i64 match_len(i64 p0, i64 op, i64 &rev) {
        rev = 0;
        if (op >= p0) return 0;
        const i64 lo = last > 0 ? last : 0;
        i64 maxb = p0 - lo;
        if (op < maxb) maxb = op;
        const i64 maxf = end - p0;
        i64 f, r;
        warp_cmp2(buf, p0, op, maxf, maxb, lane, f, r);
        rev = r;
        const i64 tot = f + r;
        return tot < MRZ_MIN_MATCH ? 0 : tot;
    }

    // rzip.c:426-462.  128 slots per step: every lane examines four consecutive slots.
    __device__ __forceinline__ i64 lookup(i64 t, i64 pp, i64 &bofs, i64 &brev) {
        i64 best = 0;
        i64 h = t & hmask;
        for (;;) {
            i64 et[4];
#pragma unroll
            for (int q = 0; q < 4; q++) et[q] = __ldcg(ttag + ((h + lane * 4 + q) & hmask));
            unsigned em4 = 0, eq4 = 0;
#pragma unroll
            for (int q = 0; q < 4; q++) {
                if (et[q] == 0) em4 |= 1u << q;
                if (et[q] == t) eq4 |= 1u << q;
            }
            const unsigned eml = __ballot_sync(FULL, em4 != 0);
            int fe = 128;                       // position of the first empty slot in this window
            if (eml) {
                const int l = __ffs(eml) - 1;
                const unsigned m = __shfl_sync(FULL, em4, l);
                fe = l * 4 + (__ffs(m) - 1);
            }
#pragma unroll
            for (int q = 0; q < 4; q++)
                if (lane * 4 + q >= fe) eq4 &= ~(1u << q);
            for (;;) {                          // equal tags before the first empty slot, in slot order
                const unsigned ql = __ballot_sync(FULL, eq4 != 0);
                if (!ql) break;
                const int l = __ffs(ql) - 1;
                const unsigned m = __shfl_sync(FULL, eq4, l);
                const int q = __ffs(m) - 1;
                if (lane == l) eq4 &= eq4 - 1;
                i64 off = __ldcg(toff + ((h + l * 4 + q) & hmask));
                i64 rev;
                i64 mlen = match_len(pp, off, rev);
                if (mlen) {
                    hits++;
                    if (mlen > best) { best = mlen; bofs = off - rev; brev = rev; }
                } else
                    misses++;
            }
            if (fe < 128) break;
            h = (h + 128) & hmask;
        }
        return best;
    }

    // The selection half of find_best_match (rzip.c:445-456) over candidates the CTA has already compared
    // (every warp ran match_len for a share of the equal-tag slots; results in slot order).
    __device__ __forceinline__ i64 lookup_pre(i64 &bofs, i64 &brev, int n, const i64 *len, const i64 *rv, const i64 *off) {
        i64 best = 0;
        for (int q = 0; q < n; q++) {
            const i64 m = len[q];
            if (m) {
                hits++;
                if (m > best) { best = m; bofs = off[q] - rv[q]; brev = rv[q]; }
            } else
                misses++;
        }
        return best;
    }

    // rzip.c:256-301; recursion unrolled onto stk (pending outer writes).
    __device__ __forceinline__ void insert(i64 t, i64 off, i64 *stk) {
        int sp = 0;
        i64 ct = t, coff = off;
        for (;;) {
            i64 h = ct & hmask, round = 0, victim_h = 0, target = 0;
            i64 btm = (min_mask << 1) | 1;
            int ctz = tz1(ct);
            bool displaced = false, over_elig = false;
            i64 otag = 0;
            for (;;) {
                i64 et = __ldcg(ttag + ((h + lane) & hmask));
                bool empty = et == 0;
                bool minbit = !empty && (et & btm) != btm;
                bool lesser = !empty && !minbit && tz1(et) < ctz;
                bool equal = !empty && !minbit && !lesser && et == ct;
                unsigned emm = __ballot_sync(FULL, empty);
                unsigned mbm = __ballot_sync(FULL, minbit);
                unsigned lsm = __ballot_sync(FULL, lesser);
                unsigned stopm = emm | mbm | lsm;
                int s = stopm ? (__ffs(stopm) - 1) : 32;
                unsigned eqm = __ballot_sync(FULL, equal);
                if (s < 32) eqm &= (1u << s) - 1u;
                int ne = __popc(eqm);
                if (vr >= round && vr - round < ne) {
                    int want = (int)(vr - round);  // index among the equal entries of this window
                    unsigned m = eqm;
                    for (int q = 0; q < want; q++) m &= m - 1;
                    victim_h = (h + (__ffs(m) - 1)) & hmask;
                }
                if (round + ne >= chain) {  // rzip.c:284-290
                    target = victim_h;
                    count--;
                    vevents++;
                    vr = (vr + 1 == chain) ? 0 : vr + 1;
                    break;
                }
                round += ne;
                if (s < 32) {
                    target = (h + s) & hmask;
                    if ((mbm >> s) & 1u) { count--; over_elig = true; }      // rzip.c:267-269: due for cleaning, overwrite
                    else if ((lsm >> s) & 1u) {        // rzip.c:275-277: re-insert the occupant first
                        displaced = true;
                        otag = __shfl_sync(FULL, et, s);
                    }
                    break;
                }
                h = (h + 32) & hmask;
            }
            if (displaced) {
                if (sp >= 60) { error = 3; return; }
                if (lane == 0) { stk[3 * sp] = target; stk[3 * sp + 1] = ct; stk[3 * sp + 2] = coff; }
                __syncwarp();
                sp++;
                coff = __ldcg(toff + target);
                ct = otag;
                continue;
            }
            write_slot(target, coff, ct, over_elig);
            break;
        }
        while (sp > 0) {
            sp--;
            write_slot(stk[3 * sp], stk[3 * sp + 2], stk[3 * sp + 1]);
        }
    }

    // find_stop for the serial path, by one warp: the first probe and 32 summary quad-words (2048 slots of summary) are
    // fetched in ONE round trip, the candidate group in a second (8 lanes, one tag each); a stale candidate is repaired
    // and the next one taken from the summary already in registers.  Same result as find_stop.
    __device__ __forceinline__ i64 find_stop_warp(i64 h, int need, i64 &stop_tag) {
        if (h >= hsize) return -1;
        const i64 ngroups_all = hsize >> GRP_SHIFT;
        i64 ngroups = ((h + MAXWALK) >> GRP_SHIFT) + 1;
        if (ngroups > ngroups_all) ngroups = ngroups_all;
        const u32 pass_bit = 1u << ((need > 15 ? 15 : need) - 1);
        const i64 g0 = h >> GRP_SHIFT;
        // round trip 1: the rest of the first group (lanes 0..7) and the summary of the groups behind it (all lanes)
        i64 wbase = (g0 + 1) >> 3;                       // first summary quad-word (8 groups each)
        uint4 w4 = __ldcg(reinterpret_cast<const uint4 *>(gsum + ((wbase + lane) << 2)));
        {
            const i64 sl = (g0 << GRP_SHIFT) + (lane & (GRP - 1));
            const i64 tg = __ldcg(ttag + sl);
            int z = tz1(tg);
            if (z < 0) z = 64;
            const unsigned m = __ballot_sync(FULL, lane < GRP && sl >= h && z < need);
            if (m) {
                const int l = __ffs(m) - 1;
                stop_tag = __shfl_sync(FULL, tg, l);
                return (g0 << GRP_SHIFT) + l;
            }
            if ((h & (GRP - 1)) == 0) {                  // the whole group was seen: repair its summary
                int zc = lane < GRP ? (z > 15 ? 15 : z) : 15;
                for (int o = GRP / 2; o; o >>= 1) { const int u = __shfl_xor_sync(FULL, zc, o); zc = u < zc ? u : zc; }
                if (lane == 0) sum_raise(gsum, g0, zc);
            }
        }
        i64 g = g0 + 1;
        for (;;) {
            if (g >= ngroups) return -1;
            // candidate groups of my quad-word: bit q set <=> group (wbase + lane) * 8 + q may hold a stop
            const u32 wv[4] = {w4.x, w4.y, w4.z, w4.w};
            unsigned cand = 0;
#pragma unroll
            for (int q = 0; q < 8; q++) {
                const i64 gg = ((wbase + lane) << 3) + q;
                if (gg >= g && gg < ngroups && !((wv[q >> 1] >> ((q & 1) * 16)) & pass_bit)) cand |= 1u << q;
            }
            const unsigned lm = __ballot_sync(FULL, cand != 0);
            if (!lm) {                                   // nothing in these 2048 slots: next window of summary
                g = (wbase + 32) << 3;
                wbase += 32;
                if (g >= ngroups) return -1;
                w4 = __ldcg(reinterpret_cast<const uint4 *>(gsum + ((wbase + lane) << 2)));
                continue;
            }
            const int ll = __ffs(lm) - 1;
            const unsigned cm = __shfl_sync(FULL, cand, ll);
            const i64 gc = ((wbase + ll) << 3) + (__ffs(cm) - 1);
            // round trip 2: the candidate group, one tag per lane
            const i64 sl = (gc << GRP_SHIFT) + (lane & (GRP - 1));
            const i64 tg = __ldcg(ttag + sl);
            int z = tz1(tg);
            if (z < 0) z = 64;
            const unsigned m = __ballot_sync(FULL, lane < GRP && z < need);
            if (m) {
                const int l = __ffs(m) - 1;
                stop_tag = __shfl_sync(FULL, tg, l);
                return (gc << GRP_SHIFT) + l;
            }
            int zc = lane < GRP ? (z > 15 ? 15 : z) : 15;
            for (int o = GRP / 2; o; o >>= 1) { const int u = __shfl_xor_sync(FULL, zc, o); zc = u < zc ? u : zc; }
            zc = __shfl_sync(FULL, zc, 0);
            if (lane == 0) sum_raise(gsum, gc, zc);      // stale summary: repair, go on behind this group
            g = gc + 1;
        }
    }

    // rzip.c:256-301 for a tag whose equal entries (if any) are known: every level of the displacement chain is a
    // summary-accelerated find_stop walk (all walks see the table before any write, as the reference's recursion does).
    // Returns false, having written nothing, when some level needs the general walk of insert().
    // n_eq / eq_pos: the slots (relative to the primary slot, ascending) that hold a tag equal to t, as collected by the
    // CTA-wide lookup; with them the equal-tag round counting and the victim choice of rzip.c:278-291 need no walk.
    __device__ __forceinline__ bool insert_fast(i64 t, i64 off, i64 *stk, int n_eq, const int *eq_pos) {
        const i64 btm = (min_mask << 1) | 1;
        const int kb = __popcll((u64)btm);
        int sp = 0;
        i64 ct = t, coff = off;
        for (;;) {
            const int z = tz1(ct);
            i64 stg;
            const i64 sl = find_stop_warp(ct & hmask, z > kb ? z : kb, stg);
            if (sl < 0) return false;
            if (sp == 0 && n_eq > 0) {
                const i64 rel = sl - (t & hmask);
                int ne = 0;                              // equal tags the walk meets before its stop slot
                for (int q = lane; q < n_eq; q += 32) ne += eq_pos[q] < rel;
                for (int o = 16; o; o >>= 1) ne += __shfl_xor_sync(FULL, ne, o);
                if (ne >= chain) {                       // rzip.c:284-290: the chain is full, replace the victim_round-th entry
                    const i64 target = ((t & hmask) + eq_pos[vr]) & hmask;
                    __syncwarp();
                    count--;
                    vevents++;
                    vr = (vr + 1 == chain) ? 0 : vr + 1;
                    write_slot(target, off, t);
                    return true;
                }
            }
            if (stg == 0 || (stg & btm) != btm) {
                __syncwarp();                          // every lane has read the table before lane 0 writes it
                if (stg != 0) count--;                 // rzip.c:267-269
                write_slot(sl, coff, ct, stg != 0);
                while (sp > 0) {
                    sp--;
                    write_slot(stk[3 * sp], stk[3 * sp + 2], stk[3 * sp + 1]);
                }
                return true;
            }
            // rzip.c:275-277: a lesser occupant is re-inserted first; it must be unique (it meets only itself on its walk)
            if (chain == 1 || sp >= 60) return false;
            const i64 ooff = __ldcg(toff + sl);          // independent of the duplicate-filter word: one round trip for both
            if (filt_query(filt, stg ^ DUP_SALT)) return false;
            __syncwarp();
            if (lane == 0) { stk[3 * sp] = sl; stk[3 * sp + 1] = ct; stk[3 * sp + 2] = coff; }
            __syncwarp();
            sp++;
            coff = ooff;
            ct = stg;
        }
    }

    // rzip.c:305-328
    __device__ __forceinline__ void clean_one() {
        if (clean_ptr < clean_hint) clean_ptr = clean_hint;   // nothing eligible before the hint
        clean_hint = 0;
        for (;;) {
            i64 btm = (min_mask << 1) | 1;
            while (clean_ptr < hsize) {
                i64 idx = clean_ptr + lane;
                bool elig = false;
                if (idx < hsize) {
                    i64 et = __ldcg(ttag + idx);
                    elig = et != 0 && (et & btm) != btm;
                }
                unsigned m = __ballot_sync(FULL, elig);
                if (m) {
                    clean_ptr += __ffs(m) - 1;
                    write_slot(clean_ptr, 0, 0);
                    count--;
                    ins_mask = btm;
                    return;
                }
                clean_ptr += 32;
            }
            min_mask = btm;
            clean_ptr = 0;
        }
    }

    // One position that passed the lookup gate (rzip.c:573-599).  Returns true when a match was
    // emitted (p jumped to last_match).
    // pre_n >= 0: the CTA has already walked the probe sequence and compared the pre_n equal-tag candidates;
    // pre_n == -2: the filter says no equal tag exists; pre_n == -1: ask the filter / walk on this warp.
    __device__ __noinline__ bool op(i64 pp, i64 t, i64 *stk, int pre_n, const i64 *pre_len, const i64 *pre_rev,
                                    const i64 *pre_off, const int *pre_pos) {
#define PROF(slot, t0) do { if (prof && lane == 0) { prof[slot] += clock64() - (t0); prof[(slot) + 1] += 1; } } while (0)
        nops++;
        i64 bofs = 0, brev = 0;
        i64 c0 = clock64();
        const bool maybe = pre_n >= 0 ? true : pre_n == -2 ? false : filt_query(filt, t);   // negative: nothing to find
        i64 best = 0;
        if (maybe) best = pre_n >= 0 ? lookup_pre(bofs, brev, pre_n, pre_len, pre_rev, pre_off) : lookup(t, pp, bofs, brev);
        if (maybe) PROF(2, c0);
        if ((t & ins_mask) == ins_mask) {
            inserts++;
            count++;
            bool placed = false;
            c0 = clock64();
            if (!maybe) { placed = insert_fast(t, pp, stk, 0, nullptr); PROF(4, c0); }   // no equal tag exists
            else if (pre_n >= 0) { placed = insert_fast(t, pp, stk, pre_n, pre_pos); PROF(4, c0); }   // all equal tags are known
            c0 = clock64();
            if (!placed) { insert(t, pp, stk); PROF(6, c0); }
            if (lane == 0) {
                filt_set(filt, t);
                if (maybe) filt_set(filt, t ^ DUP_SALT);              // this tag may now be in the table twice
            }
            __syncwarp();
            c0 = clock64();
            if (count > limit) { clean_one(); PROF(8, c0); }
        }
        if (best > cur_len) { cur_p = pp - brev; cur_len = best; cur_ofs = bofs; }
        if ((cur_len >= MRZ_GREAT_MATCH || pp >= cur_p + MRZ_MIN_MATCH) && cur_len >= MRZ_MIN_MATCH) {
            if (ntok >= tok_cap) { error = 2; return true; }
            if (lane == 0) { toks[ntok].pos = cur_p; toks[ntok].len = cur_len; toks[ntok].ofs = cur_ofs; }
            ntok++;
            last = cur_p + cur_len;
            cur_p = p = last;
            cur_len = 0;
            return true;
        }
        return false;
#undef PROF
    }
};

struct Plan {
    int flags, dcount;
    int w1_slot, w2_slot;               // displaced occupant's new home; where (tag, pos) goes
    i64 w1_off, w1_tag;
    int ndep;                           // forwarding: slots this plan walked past because an EARLIER op of the batch
    int dep_slot[MAXDEP], dep_own[MAXDEP];   // claims them with a tag that is no stop for this walk
};

// First slot s >= h with tz1(tags[s]) < need (an empty slot has tag 0, i.e. zero trailing ones), or -1 when
// the walk reaches the end of the table / gets too long.  Whole groups whose summary says "every tag has >= need
// trailing ones" are skipped unseen (one 128-bit load covers 8 groups = 64 slots); a group that is scanned
// completely without a hit has its summary repaired.  Tags are read eight at a time (two 256-bit loads).
__device__ __forceinline__ i64 find_stop(const i64 *ttag, u32 *gsum, i64 h, int need, i64 hsize, i64 &stop_tag) {
    if (h >= hsize) return -1;
    const i64 h0 = h;
    i64 g = h >> GRP_SHIFT;
    i64 ngroups = hsize >> GRP_SHIFT;
    {   // never look further than MAXWALK slots
        i64 lim = ((h0 + MAXWALK) >> GRP_SHIFT) + 1;
        if (lim < ngroups) ngroups = lim;
    }
    const u32 pass_bit = 1u << ((need > 15 ? 15 : need) - 1);   // need >= 1: set <=> the group holds no stop for this walk
    bool whole = (h & (GRP - 1)) == 0;
    for (;;) {
        int mn = 15;
        if ((h & (GRP - 1)) == GRP - 1) {   // the usual case at mask levels >= 3 (primary slots end in ...111): one tag, one load
            const i64 t = __ldcg(ttag + h);
            int z = tz1(t);
            if (z < 0) z = 64;
            if (z < need) { stop_tag = t; return h; }
        } else {
            const i64 o = h & ~(i64)7;
            i64 v[8];
            ld_i64x8_cg(ttag + o, v);
#pragma unroll
            for (int q = 0; q < 8; q++) {
                const i64 t = v[q];
                if (o + q >= h) {
                    int z = tz1(t);
                    if (z < 0) z = 64;
                    if (z < need) { stop_tag = t; return o + q; }
                    int zc = z > 15 ? 15 : z;
                    mn = zc < mn ? zc : mn;
                }
            }
        }
        if (whole) sum_raise(gsum, g, mn);
        // skip groups that cannot contain a stop
        g++;
        bool found = false;
        while (g < ngroups && !found) {
            const uint4 w4 = __ldcg(reinterpret_cast<const uint4 *>(gsum + ((g >> 3) << 2)));
            const u32 wv[4] = {w4.x, w4.y, w4.z, w4.w};
            for (;;) {
                if (!((wv[(g & 7) >> 1] >> ((g & 1) * 16)) & pass_bit)) { found = true; break; }
                g++;
                if ((g & 7) == 0 || g >= ngroups) break;
            }
        }
        if (!found) return -1;
        h = g << GRP_SHIFT;
        whole = true;
    }
}

__device__ __forceinline__ void plan_init(Plan &r) {
    r.flags = 0; r.dcount = 0; r.ndep = 0;
    r.w1_slot = -1; r.w2_slot = -1; r.w1_off = 0; r.w1_tag = 0;
}

// Plan the insert of (tag) against the table without modifying anything (oracle/spec_model.c plan_op).  The caller
// has established through the tag filter that no equal tag exists.  The walk starts at `start` (the tag's primary
// slot, or the slot after a claimed one when the plan is forwarded).
__device__ __forceinline__ void plan_op(Plan &r, const i64 *ttag, const i64 *toff, u32 *gsum, const Filt &filt, i64 tag,
                                        i64 start, i64 hmask, i64 hsize, i64 min_mask, int chain) {
    const i64 btm = (min_mask << 1) | 1;
    const int kb = __popcll((u64)btm);                 // entries with fewer trailing ones are due for sweeping
    const int myz = tz1(tag);
    r.flags = (r.flags & ~(F_W1 | F_W2ELIG | F_W1ELIG)) | F_INSERT;
    i64 st;
    i64 s = find_stop(ttag, gsum, start, myz > kb ? myz : kb, hsize, st);
    if (s < 0) { r.flags |= F_COMPLEX; return; }
    r.w2_slot = (int)s;
    if (st == 0) r.dcount = 1;                         // empty slot
    else if ((st & btm) != btm) { r.dcount = 0; r.flags |= F_W2ELIG; }   // rzip.c:267-269: overwrite an entry due for sweeping
    else {
        // rzip.c:275-277: the occupant is lesser; where does it go when re-inserted?  Only planned when the
        // occupant's tag is unique in the table (no equal-tag chain to count) and nothing has to move further.
        if (chain == 1 || filt_query(filt, st ^ DUP_SALT)) { r.flags |= F_COMPLEX; return; }
        const int oz = tz1(st);
        i64 st2;
        i64 g = find_stop(ttag, gsum, st & hmask, oz > kb ? oz : kb, hsize, st2);
        if (g < 0) { r.flags |= F_COMPLEX; return; }
        int d2 = 0;
        if (st2 != 0) {
            if ((st2 & btm) != btm) { d2 = -1; r.flags |= F_W1ELIG; }
            else { r.flags |= F_COMPLEX; return; }     // a second displacement: exact executor
        }
        r.flags |= F_W1;
        r.w1_slot = (int)g; r.w1_off = __ldcg(toff + s); r.w1_tag = st;
        r.dcount = 1 + d2;
    }
}

// "Twin" (oracle/spec_model.c plan_follower): position p+1 has the same tag as p and both insert.  B's
// lookup meets exactly the entry A has just inserted and match_len(p+1, p) fails (tag_misses++) unless
// the bytes form a genuine run; B's insert walk passes A's entry and goes on from the next slot - or,
// before the first overflow, overwrites A's entry when the tag itself is due for sweeping.
__device__ __forceinline__ void plan_follower(Plan &r, const i64 *ttag, u32 *gsum, const u8 *buf, i64 pos, i64 tag,
                                              int a_flags, int a_w2, i64 hsize, i64 min_mask, int chain, i64 end, i64 last) {
    if (a_flags & F_COMPLEX) { r.flags |= F_COMPLEX; return; }
    if (!(a_flags & F_INSERT)) return;                       // both lookup-only, nothing to find
    if ((a_flags & F_W1) || chain == 1) { r.flags |= F_COMPLEX; return; }
    {   // match_len(pos, pos - 1) must be < 31 (rzip.c:372-397); each direction capped at 31
        int f = 0, b = 0;
        while (f < MRZ_MIN_MATCH && pos + f < end && buf[pos + f] == buf[pos - 1 + f]) f++;
        const i64 lo = last > 0 ? last : 0;
        while (b < MRZ_MIN_MATCH && pos - b > lo && pos - 1 - b > 0 && buf[pos - 2 - b] == buf[pos - 1 - b]) b++;
        if (f + b >= MRZ_MIN_MATCH) { r.flags |= F_COMPLEX; return; }
    }
    const i64 btm = (min_mask << 1) | 1;
    const int kb = __popcll((u64)btm);
    const int myz = tz1(tag);
    r.flags |= F_INSERT | F_FOLLOW;
    if ((tag & btm) != btm) {      // the leader's own entry is due for sweeping (before the first overflow): overwrite it
        r.w2_slot = a_w2; r.dcount = 0;
    } else {
        i64 st;
        i64 s = find_stop(ttag, gsum, (i64)a_w2 + 1, myz > kb ? myz : kb, hsize, st);
        if (s < 0) { r.flags |= F_COMPLEX; return; }
        if (st != 0 && (st & btm) == btm) { r.flags |= F_COMPLEX; return; }   // lesser occupant: displacement
        r.w2_slot = (int)s; r.dcount = st == 0 ? 1 : 0;
        if (st != 0) r.flags |= F_W2ELIG;
    }
}

// Refill the ring of eligible slots (those clean_one_from_hash will zero next, in order) by a parallel scan of the
// table from e_scan, eight slots (one 64-byte line) per thread of the cluster; only meaningful once
// ins_mask == 2*min_mask+1.  The scan never goes past scan_end.  Every thread of the cluster calls it; every CTA
// updates its own replica of the ring counters identically.  One cluster barrier per iteration.
__device__ __forceinline__ void refill_queue(Shared *sh, Smem *S, const i64 *ttag, i64 btm, i64 scan_end, int want, int CL, int rank, int j, int B) {
    for (int iter = 0;; iter++) {
        const unsigned e_cnt = S->rep.x.e_cnt;
        const int avail = (int)(e_cnt - S->rep.x.e_idx);
        const i64 scan = S->rep.x.e_scan;
        if (avail >= want || scan >= scan_end || avail + 8 * B > MRZ_EQ_CAP) break;
        const i64 base = scan & ~(i64)7;
        const i64 s0 = base + 8 * (i64)j;
        unsigned m = 0;
        if (s0 < scan_end) {
            i64 v[8];
            ld_i64x8_cg(ttag + s0, v);
#pragma unroll
            for (int q = 0; q < 8; q++) {
                const i64 s = s0 + q, t = v[q];
                if (s >= scan && s < scan_end && t != 0 && (t & btm) != btm) m |= 1u << q;
            }
        }
        unsigned long long total;
        // (sites 0 and 3 alternate: a CTA that is already in the next iteration must not overwrite totals a slower one still reads)
        const unsigned long long ex = cluster_excl_scan((unsigned long long)__popc(m), S, (iter & 1) ? 3 : 0, CL, rank, total, NoExtra());
        unsigned at = e_cnt + (unsigned)ex;
        for (int q = 0; q < 8; q++)
            if ((m >> q) & 1u) STG_(sh->E[at++ & (MRZ_EQ_CAP - 1)], (int)(s0 + q));
        if (threadIdx.x == 0) {
            S->rep.x.e_cnt = e_cnt + (unsigned)total;
            const i64 ns = base + 8 * (i64)B;
            S->rep.x.e_scan = ns < scan_end ? ns : scan_end;
        }
        __syncthreads();
    }
}

// First ring index in [a, b) whose slot is >= key (b if none); 32-ary search by one warp over the ring in global memory.
__device__ __forceinline__ unsigned ring_lower_bound(const Shared *sh, unsigned a, unsigned b, int key, int lane) {
    while (b - a > 32) {
        const unsigned len = b - a;
        const unsigned idx = a + (unsigned)(((unsigned long long)len * lane) / 32);
        const int v = LDG_(sh->E[idx & (MRZ_EQ_CAP - 1)]);
        const unsigned ge = __ballot_sync(FULL, v >= key);
        if (!ge) a = a + (unsigned)(((unsigned long long)len * 31) / 32) + 1;
        else {
            const int l = __ffs(ge) - 1;
            if (l == 0) return a;
            const unsigned na = a + (unsigned)(((unsigned long long)len * (l - 1)) / 32) + 1;
            b = a + (unsigned)(((unsigned long long)len * l) / 32) + 1;
            a = na;
        }
    }
    const bool in = a + lane < b;
    const int v = in ? LDG_(sh->E[(a + lane) & (MRZ_EQ_CAP - 1)]) : 0x7fffffff;
    const unsigned ge = __ballot_sync(FULL, in && v >= key);
    return ge ? a + (__ffs(ge) - 1) : b;
}

// After a serial op (warp 0 of rank 0, all lanes): bring the ring back in step with what the op wrote.  Only a
// sweep step (zeroed) or an insert over an entry that was due for sweeping can touch a slot of the ring.
// head0 / head1: the first two slots of the ring as read BEFORE the op (-1: not read): the usual sweep step zeroes head0
// and the ring advances by one without another look at it.
__device__ __forceinline__ void sync_queue(Shared *sh, XRes *x, const int *wlog, int nlog, int lane, int head0, int head1) {
    bool fresh = head1 >= 0;                     // the ring is still as it was when head0 / head1 were read
    for (int w = 0; w < nlog; w++) {
        const int e = wlog[w], sl = e >> 2;
        unsigned e_idx = x->e_idx, e_cnt = x->e_cnt;
        const i64 e_scan = x->e_scan;
        __syncwarp();                            // every lane has read the counters before lane 0 updates them
        if (e & 1) {                             // zeroed by clean_one: consume the ring up to it (usually just its head)
            if (e_idx != e_cnt) {
                if (fresh && head0 == sl && head1 > sl) e_idx++;
                else if (e_cnt - e_idx > 1 && LDG_(sh->E[e_idx & (MRZ_EQ_CAP - 1)]) == sl && LDG_(sh->E[(e_idx + 1) & (MRZ_EQ_CAP - 1)]) > sl) e_idx++;
                else e_idx = ring_lower_bound(sh, e_idx, e_cnt, sl + 1, lane);
            }
            fresh = false;
            if (lane == 0) x->e_idx = e_idx;
        } else if ((e & 2) && sl < e_scan && e_idx != e_cnt) {   // overwritten by an insert: if queued, cut the ring there
            const unsigned a = ring_lower_bound(sh, e_idx, e_cnt, sl, lane);
            const bool hit = a != e_cnt && LDG_(sh->E[a & (MRZ_EQ_CAP - 1)]) == sl;
            if (hit && lane == 0) { x->e_cnt = a; x->e_scan = sl; }
            fresh = false;
        }
        __syncwarp();
    }
}

// First op index in [a, b) whose position is > p (b if none); 32-ary search by one warp over b_pos in global memory.
__device__ __forceinline__ int pos_upper_bound(const Shared *sh, int a, int b, i64 p, int lane) {
    while (b - a > 32) {
        const int len = b - a;
        const int idx = a + (len * lane) / 32;
        const i64 v = LDG_(sh->b_pos[idx]);
        const unsigned gt = __ballot_sync(FULL, v > p);
        if (!gt) a = a + (len * 31) / 32 + 1;
        else {
            const int l = __ffs(gt) - 1;
            if (l == 0) return a;
            const int na = a + (len * (l - 1)) / 32 + 1;
            b = a + (len * l) / 32 + 1;
            a = na;
        }
    }
    const bool in = a + lane < b;
    const i64 v = in ? LDG_(sh->b_pos[a + lane]) : 0;
    const unsigned gt = __ballot_sync(FULL, in && v > p);
    return gt ? a + (__ffs(gt) - 1) : b;
}

// Match extension by the whole CTA: longest common prefix of buf[a0..] / buf[b0..] (fwd) or longest common suffix of
// buf[..a0) / buf[..b0), at most maxlen bytes; 32*T bytes per step, four independent 8-byte compares per thread in flight.
__device__ __forceinline__ i64 cta_extend(Smem *S, const u8 *buf, i64 a0, i64 b0, i64 maxlen, bool fwd) {
    const int tid = threadIdx.x, T = blockDim.x;
    i64 f = 0;
    while (f < maxlen) {
        if (tid == 0) S->lk_fe = 0x7fffffff;
        __syncthreads();
        int first = 0x7fffffff;
        for (int c = 3; c >= 0; c--) {
            const int ci = c * T + tid;
            const i64 k = f + (i64)ci * 8;
            u64 x = 0;
            if (k < maxlen) x = fwd ? ld_u64_unaligned(buf, a0 + k) ^ ld_u64_unaligned(buf, b0 + k)
                                    : ld_u64_unaligned(buf, a0 - 8 - k) ^ ld_u64_unaligned(buf, b0 - 8 - k);
            if (x) first = (ci << 3) | (fwd ? ((__ffsll((long long)x) - 1) >> 3) : (__clzll((long long)x) >> 3));
        }
        if (first != 0x7fffffff) atomicMin(&S->lk_fe, first);
        __syncthreads();
        const int r = S->lk_fe;
        __syncthreads();
        if (r != 0x7fffffff) {
            f += (i64)(r >> 3) * 8 + (r & 7);
            break;
        }
        f += (i64)4 * T * 8;
    }
    return f > maxlen ? maxlen : f;
}

// single_match_len's two scans (rzip.c:372-397) for ONE candidate by the whole CTA, both directions at once: the lower
// half of the threads compares forwards from (a0, b0), the upper half backwards, 16 * T bytes per direction and step
// (aligned 128-bit loads, cmp_thread); a step is one memory round trip and three block barriers.
__device__ __forceinline__ void cta_cmp2(Smem *S, const u8 *buf, i64 a0, i64 b0, i64 maxf, i64 maxb, i64 &f_out, i64 &r_out) {
    const int tid = threadIdx.x, T2 = blockDim.x >> 1;
    const bool back = tid >= T2;
    const int t = back ? tid - T2 : tid;
    const int SU = T2 * CMP_G;                            // units per step and direction
    const int skf = (int)((uintptr_t)(buf + a0) & 15), skb = (int)((16 - ((uintptr_t)(buf + a0) & 15)) & 15);
    i64 f = 0, r = 0;
    bool fdone = maxf <= 0, rdone = maxb <= 0;
    // one step per round trip at first (most candidates differ within it; VM-like and dense matches end within tens of
    // KiB, and a wider trip costs them more than it saves: measured); a match that has survived CMP_LONG_AFTER steps is a
    // long one (tar-like bodies: MiB) and goes on CMP_LONG steps per round trip
    for (i64 step = 0, ks = 1; !(fdone && rdone); step += ks, ks = step >= CMP_LONG_AFTER ? CMP_LONG : 1) {
        if (tid == 0) { S->lk_fe = 0x7fffffff; S->lk_fe2 = 0x7fffffff; }
        __syncthreads();
        int first = 0x7fffffff;
        if (ks == 1) {
            if (back) { if (!rdone) first = cmp_thread<false>(buf, a0, b0, maxb, step, SU, t); }
            else if (!fdone) first = cmp_thread<true>(buf, a0, b0, maxf, step, SU, t);
        } else {
#pragma unroll
            for (int q = CMP_LONG - 1; q >= 0; q--) {
                int v = 0x7fffffff;
                if (back) { if (!rdone) v = cmp_thread<false>(buf, a0, b0, maxb, step + q, SU, t); }
                else if (!fdone) v = cmp_thread<true>(buf, a0, b0, maxf, step + q, SU, t);
                if (v != 0x7fffffff) first = q * 16 * SU + v;
            }
        }
        if (first != 0x7fffffff) atomicMin(back ? &S->lk_fe2 : &S->lk_fe, first);
        __syncthreads();
        const int rf = S->lk_fe, rr = S->lk_fe2;
        __syncthreads();
        if (!fdone) {
            if (rf != 0x7fffffff) { f = 16 * step * SU + rf - skf; fdone = true; }
            else if (16 * (step + ks) * SU - skf >= maxf) { f = maxf; fdone = true; }
        }
        if (!rdone) {
            if (rr != 0x7fffffff) { r = 16 * step * SU + rr - skb; rdone = true; }
            else if (16 * (step + ks) * SU - skb >= maxb) { r = maxb; rdone = true; }
        }
    }
    f_out = f > maxf ? maxf : f;
    r_out = r > maxb ? maxb : r;
}

// The 64-bit tag of position p (rzip.c:348-358) from the bytes: two 256-bit loads, 31 look-ups in shared memory.
__device__ __forceinline__ i64 tag_at(const u8 *buf, const i64 *hx, i64 p) {
#ifdef MRZ_CUSIM
    i64 t = 0;
    for (int i = 0; i < MRZ_MIN_MATCH; i++) t ^= hx[buf[p + i]];
    return t;
#else
    const u8 *a = buf + (p & ~(i64)31);
    u64 v[8];
    asm volatile("ld.global.nc.v4.u64 {%0,%1,%2,%3}, [%4];" : "=l"(v[0]), "=l"(v[1]), "=l"(v[2]), "=l"(v[3]) : "l"(a));
    asm volatile("ld.global.nc.v4.u64 {%0,%1,%2,%3}, [%4];" : "=l"(v[4]), "=l"(v[5]), "=l"(v[6]), "=l"(v[7]) : "l"(a + 32));
    const int ws = (int)(p & 31) >> 3, sh = (int)(p & 7) * 8;
    u64 a0, a1, a2, a3, a4;
    switch (ws) {
        case 0: a0 = v[0]; a1 = v[1]; a2 = v[2]; a3 = v[3]; a4 = v[4]; break;
        case 1: a0 = v[1]; a1 = v[2]; a2 = v[3]; a3 = v[4]; a4 = v[5]; break;
        case 2: a0 = v[2]; a1 = v[3]; a2 = v[4]; a3 = v[5]; a4 = v[6]; break;
        default: a0 = v[3]; a1 = v[4]; a2 = v[5]; a3 = v[6]; a4 = v[7]; break;
    }
    u64 w[4];
    if (sh) {
        w[0] = (a0 >> sh) | (a1 << (64 - sh)); w[1] = (a1 >> sh) | (a2 << (64 - sh));
        w[2] = (a2 >> sh) | (a3 << (64 - sh)); w[3] = (a3 >> sh) | (a4 << (64 - sh));
    } else { w[0] = a0; w[1] = a1; w[2] = a2; w[3] = a3; }
    i64 t = 0;
#pragma unroll
    for (int i = 0; i < MRZ_MIN_MATCH; i++) t ^= hx[(w[i >> 3] >> ((i & 7) * 8)) & 255u];
    return t;
#endif
}

// high bit of every byte of w that is >= k (all bytes < 128, k < 128), gathered into a nibble: bit i <=> byte i
__device__ __forceinline__ unsigned ge_nibble(u32 w, u32 k4) {
    const u32 x = ((w | 0x80808080u) - k4) & 0x80808080u;
    return (((x >> 7) * 0x01020408u) >> 24) & 0xfu;
}

// Candidates (map byte >= kgate) of the span [a, a + s) of positions, s a power of two, in ascending order, restricted
// to [lo, end]; f(pos) returns false to stop.  s <= 64: the map bytes are read directly; s > 64: the 64:1 summary
// first, then the map bytes of the non-empty blocks only.
template <class F>
__device__ __forceinline__ void span_candidates(const u8 *__restrict__ bmap, const u8 *__restrict__ bsum, i64 a, int s, int kgate, i64 lo,
                                                i64 end, i64 nblocks, F f) {
    if (a > end || a + s <= lo) return;
    const u32 k4 = (u32)kgate * 0x01010101u;
    if (s <= 64) {
        const i64 v0 = a & ~(i64)15;
        const int nvec = s <= 16 ? 1 : s >> 4;
        u64 m = 0;
        for (int q = 0; q < nvec; q++) {
            const uint4 v = __ldg(reinterpret_cast<const uint4 *>(bmap + v0) + q);
            const u64 m16 = (u64)(ge_nibble(v.x, k4) | (ge_nibble(v.y, k4) << 4) | (ge_nibble(v.z, k4) << 8) | (ge_nibble(v.w, k4) << 12));
            m |= m16 << (16 * q);
        }
        if (s < 16) m &= (((u64)1 << s) - 1) << (int)(a - v0);
        if (v0 < lo) m &= ~(((u64)1 << (int)(lo - v0)) - 1);            // lo - v0 < 64 here
        if (v0 + 63 > end) m &= (((u64)1 << (int)(end - v0)) << 1) - 1;   // end - v0 in 0..62
        while (m) {
            const int b = __ffsll((long long)m) - 1;
            m &= m - 1;
            if (!f(v0 + b)) return;
        }
        return;
    }
    const i64 blk0 = a >> MAP_BLK_SHIFT, blk_lo = lo >> MAP_BLK_SHIFT;
    const int nb = s >> MAP_BLK_SHIFT;                 // blocks in the span: 2, 4, ... 1024
    const i64 vb0 = blk0 & ~(i64)15;
    const int nvec = nb <= 16 ? 1 : nb >> 4;
    for (int q = 0; q < nvec; q++) {
        const i64 bq = vb0 + 16 * q;
        if (bq >= nblocks) return;
        const uint4 v = __ldg(reinterpret_cast<const uint4 *>(bsum + bq));
        unsigned m = ge_nibble(v.x, k4) | (ge_nibble(v.y, k4) << 4) | (ge_nibble(v.z, k4) << 8) | (ge_nibble(v.w, k4) << 12);
        if (nb < 16) m &= ((1u << nb) - 1u) << (int)(blk0 - vb0);
        if (bq < blk_lo) m &= (blk_lo - bq >= 16) ? 0u : ~((1u << (int)(blk_lo - bq)) - 1u);
        if (bq + 16 > nblocks) m &= (1u << (int)(nblocks - bq)) - 1u;
        while (m) {
            const int bb = __ffs(m) - 1;
            m &= m - 1;
            const i64 p0 = (bq + bb) << MAP_BLK_SHIFT;
            const uint4 *src = reinterpret_cast<const uint4 *>(bmap + p0);
            u64 cm = 0;
#pragma unroll
            for (int w = 0; w < 4; w++) {
                const uint4 x = __ldg(src + w);
                const u64 m16 = (u64)(ge_nibble(x.x, k4) | (ge_nibble(x.y, k4) << 4) | (ge_nibble(x.z, k4) << 8) | (ge_nibble(x.w, k4) << 12));
                cm |= m16 << (16 * w);
            }
            if (p0 < lo) cm &= ~(((u64)1 << (int)(lo - p0)) - 1);
            if (p0 + 63 > end) cm &= (((u64)1 << (int)(end - p0)) << 1) - 1;
            while (cm) {
                const int b = __ffsll((long long)cm) - 1;
                cm &= cm - 1;
                if (!f(p0 + b)) return;
            }
        }
    }
}

// ---- the kernel -----------------------------------------------------------------------------------
// Cluster c of the grid runs automaton c: descs[c].  Persistent: returns when the chunk is done (or on error).
__global__ void __launch_bounds__(MRZ_MAXT, 1)
k_automaton(const AutoDesc *__restrict__ descs) {
    const int CL = cl_size(), rank = cl_rank();
    const AutoDesc D = descs[blockIdx.x / CL];
    AutoState *st = D.st;
    if (st->done || st->error) return;
    MRZ_SMEM_PTR(Smem, S);
    Shared *sh = D.sh;
    XRes &x = S->rep.x;
    const u8 *__restrict__ buf = D.buf;
    const u8 *__restrict__ bmap = D.bmap;
    const u8 *__restrict__ bsum = D.bsum;
    i64 *ttag = D.ttag, *toff = D.toff;
    u32 *gsum = D.gsum;
    // two claim arrays, used alternately (epoch parity): the claims of the NEXT epoch are made while the ones of the current
    // epoch are still being read, and must not disturb them
    unsigned *const claim0 = D.claim, *const claim1 = D.claim + D.claim_stride;
    i64 *dbg = D.dbg;
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5, T = blockDim.x;
    const int B = CL * T, j = rank * T + tid;
    Filt filt;
    filt.w = D.filt; filt.shift = 64 - st->filt_bits;
    for (int q = tid; q < 256; q += T) S->hx[q] = D.hidx[q];
    const i64 n = st->n, end = st->end, hmask = st->hmask, hsize = st->hsize, limit = st->limit;
    const int chain = st->chain;
    const i64 nblocks = (end >> MAP_BLK_SHIFT) + 1;          // 64-position blocks that hold a position <= end
    if (tid == 0) {                                          // every CTA builds the same replica
        x.run.p = st->p; x.run.last = st->last; x.run.cur_p = st->cur_p; x.run.cur_len = st->cur_len; x.run.cur_ofs = st->cur_ofs;
        x.run.min_mask = st->min_mask; x.run.ins_mask = st->ins_mask; x.run.clean_ptr = st->clean_ptr; x.run.count = st->count;
        x.run.vr = st->victim_round;
        x.run.hits = st->tag_hits; x.run.misses = st->tag_misses; x.run.inserts = st->inserts; x.run.vevents = st->victim_events;
        x.run.nops = st->n_ops; x.run.ntok = st->ntok; x.run.error = 0; x.run.pad = 0;
        x.gp = st->p + 1; x.e_idx = x.e_cnt = 0; x.e_scan = st->clean_ptr;
        x.ended = x.jumped = x.nlog = x.pos_after = 0;
        S->rep.exhausted = 0;
        for (int r = 0; r < MRZ_MAXCL; r++) S->rep.chg[0][r] = S->rep.chg[1][r] = 0xffffffffu;
    }
    Warp w;   // used by warp 0 of rank 0 only
    w.buf = buf; w.ttag = ttag; w.toff = toff; w.filt = filt; w.gsum = gsum; w.toks = D.toks; w.lane = lane;
    w.n = n; w.end = end; w.hmask = hmask; w.hsize = hsize; w.limit = limit; w.tok_cap = st->tok_cap; w.chain = chain;
    w.wlog = x.wlog; w.nlog = 0; w.clean_hint = 0; w.prof = dbg ? dbg + 40 : nullptr;
    if (j == 0) { sh->err_a = sh->err_b = 0; STG_(sh->err_site, 0); }
    unsigned epoch = 0;      // claim epoch: uniform over the cluster (every thread advances it identically)
    i64 tm[12];
#pragma unroll
    for (int q = 0; q < 12; q++) tm[q] = 0;
    i64 t_prev = clock64();
#define TICK(slot) do { if (dbg) { const i64 t_now = clock64(); tm[slot] += t_now - t_prev; t_prev = t_now; } } while (0)
    csync(CL);   // replicas are built; the host has filled claim[] with 0xff: every key compares greater than any live one

    for (;;) {
        // ================================================================ batch set-up (every CTA reads its own replica)
#ifdef MRZ_CHECK
        if (LDG_(sh->err_site)) { if (tid == 0) x.run.error = LDG_(sh->err_site); __syncthreads(); }
#endif
        if (x.run.error || S->rep.exhausted) break;
        const i64 min_mask = x.run.min_mask, ins_mask = x.run.ins_mask;
        const i64 btm = (min_mask << 1) | 1;
        const bool steady = ins_mask == btm;   // inserts can no longer create sweep-eligible entries
        const i64 p_now = x.run.p, last_now = x.run.last;
        const int kgate = __popcll((u64)min_mask);
        const i64 gp0 = x.gp;
        const i64 lo = gp0 > p_now + 1 ? gp0 : p_now + 1;
        if (lo > end) {                        // every position up to `end` has been examined or jumped over (rzip.c:548)
            __syncthreads();
            if (tid == 0) S->rep.exhausted = 1;
            __syncthreads();
            continue;
        }
        // ================================================================ sweep ring
        if (steady) refill_queue(sh, S, ttag, btm, hsize, B, CL, rank, j, B);
#ifdef MRZ_CUSIM
        // (the ring entries of the refill just made are read by other CTAs only behind the next cluster barrier - the gather's
        // prefix sum; the check reads them right here and needs that barrier itself)
        if (g_selfcheck_on) csync(CL);
        if (j == 0 && g_selfcheck_on && (g_selfcheck_tick++ & g_selfcheck_mask) == 0) {   // emulator self-check (every 64th batch): hash_count and the ring against the table itself
            i64 used = 0;
            for (i64 sl = 0; sl < hsize; sl++) used += ttag[sl] != 0 || toff[sl] != 0;
            if (used != x.run.count) {
                g_selfcheck_failures++;
                fprintf(stderr, "SELFCHECK count: batch after p=%lld: the table holds %lld entries, hash_count says %lld\n", (long long)p_now, (long long)used, (long long)x.run.count);
            }
            unsigned q = x.e_idx;
            i64 bad = -1;
            for (i64 sl = x.run.clean_ptr; steady && sl < x.e_scan; sl++) {   // the ring lists exactly the eligible slots of [clean_ptr, e_scan)
                const i64 t = ttag[sl];
                if (t != 0 && (t & btm) != btm) {
                    if (q == x.e_cnt || sh->E[q & (MRZ_EQ_CAP - 1)] != (int)sl) { bad = sl; break; }
                    q++;
                }
            }
            if (steady && bad < 0 && q != x.e_cnt) bad = -2;
            if (bad != -1) {
                g_selfcheck_failures++;
                fprintf(stderr, "SELFCHECK ring: batch after p=%lld: slot %lld (clean_ptr %lld e_scan %lld e_idx %u e_cnt %u)\n", (long long)p_now, (long long)bad,
                        (long long)x.run.clean_ptr, (long long)x.e_scan, x.e_idx, x.e_cnt);
            }
        }
#endif
        TICK(0);
        // ================================================================ gather the next candidates from the map
        // thread j looks at the s = 2^kgate positions [a, a + s): one candidate expected.  One cluster-wide prefix sum
        // orders them; the cluster exchanges them through b_pos[] (thread j then takes candidate j).
        const int ks = kgate < 16 ? kgate : 16, s = 1 << ks;
        const i64 gbase = lo & ~((i64)s - 1);
        const i64 a = gbase + (i64)j * s;
        int nr;
        i64 n_ins_batch = 0;         // candidates of this batch that pass the insert gate: an upper bound on its sweep steps
        {
            // (the same sum also counts the candidates that pass the INSERT gate - the map byte tells: only those can make
            // the sweep move, which bounds how far it gets within this batch, see s_safe below)
            int cnt = 0, icnt = 0;
            const int kins = __popcll((u64)ins_mask);
            span_candidates(bmap, bsum, a, s, kgate, lo, end, nblocks, [&](i64 pos) { cnt++; icnt += (int)__ldg(bmap + pos) >= kins; return true; });
            if (tid == 0) S->lgp = 0x7fffffffffffffffLL;
            unsigned long long total;
            unsigned long long at = cluster_excl_scan((unsigned long long)cnt | ((unsigned long long)icnt << 32), S, 1, CL, rank, total, NoExtra());
            n_ins_batch = (i64)(total >> 32);
            total &= 0xffffffffULL;
            at &= 0xffffffffULL;
            if (cnt) {
                i64 drop = -1;
                span_candidates(bmap, bsum, a, s, kgate, lo, end, nblocks, [&](i64 pos) {
                    if (at < (unsigned long long)B) {
                        STG_(sh->b_pos[at], pos);
                        prefetch_l2(buf + (pos & ~(i64)31));
                        prefetch_l2(buf + (pos & ~(i64)31) + 63);
                        at++;
                        return true;
                    }
                    drop = pos;                  // first candidate left for the next batch
                    return false;
                });
                if (drop >= 0) atomicMin((long long *)&S->lgp, (long long)drop);
            }
            __syncthreads();
            if (tid < CL) push_to(&S->rep.xmin[rank], tid, S->lgp);
            csync(CL);
            nr = total < (unsigned long long)B ? (int)total : B;
            if (tid == 0) {
                i64 g = gbase + (i64)B * s;
                for (int r = 0; r < CL; r++) g = S->rep.xmin[r] < g ? S->rep.xmin[r] : g;
                x.gp = g;
            }
        }
        if (nr == 0) { __syncthreads(); TICK(1); continue; }
        i64 my_pos = 0, my_tag = 0, prev1 = -1, prev2 = -1;
        if (j < nr) {
            my_pos = LDG_(sh->b_pos[j]);
            if (j > 0) prev1 = LDG_(sh->b_pos[j - 1]);
            if (j > 1) prev2 = LDG_(sh->b_pos[j - 2]);
            my_tag = tag_at(buf, S->hx, my_pos);
            STG_(sh->b_tag[j], my_tag);
            // the plan's first reads (filter word, first group of tags, its summary) miss L2 when the table is larger
            // than L2: start the last two now, they arrive while the filter word is on its way
            const i64 h0 = my_tag & hmask;
            prefetch_l2(ttag + h0);
            prefetch_l2(gsum + ((((h0 >> GRP_SHIFT) + 1) >> 3) << 2));
        }
        TICK(1);
        if (dbg && j == 0) { dbg[16] += nr; dbg[17] += 1; }
        // ================================================================ phase A: plan (read-only against the table)
        Plan pl;
        plan_init(pl);
        bool follower = false;
        int lead_w2 = -1;
        if (j < nr) {
            // twin: same tag as the position just before, which is not itself the second of a pair
            bool same = false, same_prev = false;
            if (prev1 == my_pos - 1) {
                same = S->hx[buf[my_pos - 1]] == S->hx[buf[my_pos + 30]];                                       // tag[p-1] == tag[p]
                if (same && prev2 == my_pos - 2) same_prev = S->hx[buf[my_pos - 2]] == S->hx[buf[my_pos + 29]];   // tag[p-2] == tag[p-1]
            }
            follower = same && !same_prev;
        }
        {
            // own plan; a follower in lane 0 plans its leader's op itself (same tag => same plan), the others get it by shuffle
            if (j < nr && (!follower || lane == 0)) {
                if (filt_query(filt, my_tag)) pl.flags |= F_COMPLEX;          // an equal tag may exist
                else if ((my_tag & ins_mask) == ins_mask)
                    plan_op(pl, ttag, toff, gsum, filt, my_tag, my_tag & hmask, hmask, hsize, min_mask, chain);
            }
            int a_flags = __shfl_up_sync(FULL, pl.flags, 1), a_w2 = __shfl_up_sync(FULL, pl.w2_slot, 1);
            if (follower) {
                if (lane == 0) { a_flags = pl.flags; a_w2 = pl.w2_slot; }
                lead_w2 = a_w2;
                plan_init(pl);
                plan_follower(pl, ttag, gsum, buf, my_pos, my_tag, a_flags, a_w2, hsize, min_mask, chain, end, last_now);
            }
        }
        if (j < nr) { STG_(sh->pl_fl[j], pl.flags); STG_(sh->pl_w2[j], pl.w2_slot); }
        TICK(2);
        const int my_need = tz1(my_tag) > __popcll((u64)btm) ? tz1(my_tag) : __popcll((u64)btm);
        int pos = 0;                 // ops before pos are done (committed, executed or skipped)
        int pre0 = 0;                // the prefix sums in pre[] cover the ops of [pre0, nr)
        bool need_conflict = true;   // plans were made since the last round of claims
        bool batch_ended = false;
#ifdef MRZ_CUSIM
        const i64 dbg_nops0 = x.run.nops;   // emulator build: every op of the batch is accounted for exactly once
        int dbg_done = 0, dbg_skip = 0;
#endif
        const i64 ins_at_start = x.run.inserts;
        int qacc = 0x7fffffff;       // smallest ring slot a committed plan overwrote (the ring is cut there at the end of the batch)
        i64 w_lo = x.run.clean_ptr, w_hi = steady ? x.e_scan : w_lo, s_safe = w_lo - 1;
        unsigned long long incl = 0;

        for (;;) {   // ====================================================== rounds: [conflicts ->] cut -> commit -> (executor | plan again)
            // the sweep can reach the slots of the ring up to s_safe within this batch (at most one sweep per op)
            const bool ovf_possible = steady ? (int)(x.e_cnt - x.e_idx) < nr - pos : x.run.count + (nr - pos) > limit;
            // zone: a stop slot the sweep may reach in this batch: the order of sweeps and inserts matters
#define ZONE_CHECK()                                                                                                   \
    if (steady && !(pl.flags & F_CONF) && (pl.flags & (F_COMPLEX | F_INSERT)) == F_INSERT &&                           \
        (((pl.flags & F_W2ELIG) && pl.w2_slot >= w_lo && pl.w2_slot <= s_safe) ||                                       \
         ((pl.flags & F_W1ELIG) && pl.w1_slot >= w_lo && pl.w1_slot <= s_safe))) {                                      \
        pl.flags |= F_CONF | F_ZONE;                                                                                   \
        if (dbg) atomicAdd((unsigned long long *)&dbg[32], 1ULL);                                                      \
    }
#define CLAIM(ep_)                                                                                                      \
    if (j >= pos && j < nr && (pl.flags & (F_COMPLEX | F_INSERT | F_CONF)) == F_INSERT) {                               \
        unsigned *const cw_ = ((ep_) & 1u) ? claim1 : claim0;                                                           \
        int s2 = pl.w2_slot; CHK(sh, x, 1, s2, hsize + 8);                                                              \
        atomicMin(cw_ + s2, claim_key(ep_, j, 0));                                                                      \
        if (pl.flags & F_W1) { int s1 = pl.w1_slot; CHK(sh, x, 2, s1, hsize + 8); atomicMin(cw_ + s1, claim_key(ep_, j, 1)); } \
    }
            if (tid == 0) { S->lmin = 0x7fffffff; S->lq = 0x7fffffff; }
            __syncthreads();
            bool heavy = false;                    // this round computes the prefix sums afresh
            unsigned long long heavy_total = 0;
            if (need_conflict) {
                // ------------------------------------------------------------ conflicts: exact stop-slot claims with forwarding
                // Every live plan claims its stop slot(s) in claim[] (atomicMin of (epoch, op index)).  A plan whose stop slot
                // is claimed by an EARLIER op either walks on from the next slot - when that op's tag is no stop for this walk,
                // the claimed slot is remembered as a dependency - or gives up (F_CONF).  Iterated to a fixed point; the claims
                // of the next epoch are made BEFORE the barrier that tells whether anybody changed: one hop per iteration.
                {   // a dependency on an op that has been committed since is satisfied for good
                    int nd = 0;
                    for (int d = 0; d < pl.ndep; d++)
                        if (pl.dep_own[d] >= pos) { pl.dep_slot[nd] = pl.dep_slot[d]; pl.dep_own[nd] = pl.dep_own[d]; nd++; }
                    pl.ndep = nd;
                }
                w_lo = x.run.clean_ptr; w_hi = steady ? x.e_scan : w_lo; s_safe = w_lo - 1;
                if (steady && w_hi > w_lo) {
                    // at most one sweep step per op that is still to run - and only an op that inserts can cause one
                    i64 ins_left = n_ins_batch - (x.run.inserts - ins_at_start);
                    if (ins_left < 0) ins_left = 0;
                    const unsigned ei = x.e_idx, ec = x.e_cnt, reach = (unsigned)(ins_left < (i64)(nr - pos) ? ins_left : (i64)(nr - pos));
                    s_safe = (ec - ei > reach) ? (i64)LDG_(sh->E[(ei + reach) & (MRZ_EQ_CAP - 1)]) : w_hi - 1;
                }
                if (epoch + 4 * FWD_ITERS + 16 >= CLAIM_EPOCH_MAX) {   // claim keys nearly exhausted: wipe the arrays, restart the epochs
                    csync(CL);
                    for (i64 q = j; q < hsize + 8; q += B) claim0[q] = claim1[q] = 0xffffffffu;
                    if (tid == 0) for (int r = 0; r < MRZ_MAXCL; r++) S->rep.chg[0][r] = S->rep.chg[1][r] = 0xffffffffu;
                    epoch = 0;
                    csync(CL);
                }
                unsigned ep = epoch++;
                ZONE_CHECK();
                CLAIM(ep);
                csync(CL);
                for (int it = 0;; it++) {
                    if (tid == 0) S->lchg = 0;
                    __syncthreads();
                    const unsigned ep_hi = KEY_EPOCH(claim_key(ep, 0, 0));
                    const unsigned *const claim = (ep & 1u) ? claim1 : claim0;     // the claims of epoch ep (complete since the last barrier)
                    const bool live = j >= pos && j < nr && (pl.flags & (F_COMPLEX | F_INSERT | F_CONF)) == F_INSERT;
                    bool chg = false;
                    if (live) {
                        bool bad = false;
                        // a follower's plan builds on its leader's write: only older claims conflict, and the leader must not have moved
                        const int before = (pl.flags & F_FOLLOW) ? j - 1 : j;
                        int why = 0;
                        // (a leader that was planned again may keep its slot and yet displace the occupant now: its second write
                        // can land on the slot this plan found free)
                        if ((pl.flags & F_FOLLOW) && j - 1 >= pos &&
                            (LDG_(sh->pl_w2[j - 1]) != lead_w2 || (LDG_(sh->pl_fl[j - 1]) & (F_W1 | F_COMPLEX | F_INSERT)) != F_INSERT)) { bad = true; why = 33; }
                        for (int d = 0; d < pl.ndep && !bad; d++) {   // slots I walked past on the strength of an earlier op's claim
                            const unsigned o = __ldcg(claim + pl.dep_slot[d]);
                            const int oi = KEY_IDX(o);
                            if (KEY_EPOCH(o) != ep_hi || (o & 1u) || oi >= j) { bad = true; why = 34; break; }
                            const i64 ot = LDG_(sh->b_tag[oi]);
                            if (tz1(ot) < my_need || (ot & btm) != btm || ot == my_tag) { bad = true; why = 34; break; }
                            pl.dep_own[d] = oi;
                        }
                        if (!bad) {
                            const unsigned o = __ldcg(claim + pl.w2_slot);
                            const int oi = KEY_IDX(o);
                            if (oi < before) {   // my stop slot goes to an earlier op
                                const i64 ot = LDG_(sh->b_tag[oi]);
                                if ((pl.flags & F_FOLLOW) || (o & 1u) || pl.ndep == MAXDEP || it >= FWD_ITERS || tz1(ot) < my_need ||
                                    (ot & btm) != btm || ot == my_tag) {   // an equal tag: my lookup must meet its entry
                                    bad = true;
                                    why = (pl.flags & F_FOLLOW) ? 33 : (o & 1u) ? 35 : ot == my_tag ? 36 : (tz1(ot) < my_need || (ot & btm) != btm) ? 37 : 38;
                                } else {         // its tag is no stop for my walk: go on from the next slot
                                    pl.dep_slot[pl.ndep] = pl.w2_slot; pl.dep_own[pl.ndep] = oi; pl.ndep++;
                                    plan_op(pl, ttag, toff, gsum, filt, my_tag, (i64)pl.w2_slot + 1, hmask, hsize, min_mask, chain);
                                    STG_(sh->pl_w2[j], pl.w2_slot);
                                    chg = true;
                                    if (dbg) atomicAdd((unsigned long long *)&dbg[31], 1ULL);
                                }
                            } else if ((pl.flags & F_W1) && KEY_IDX(__ldcg(claim + pl.w1_slot)) < j) {
                                bad = true;
                                why = 39;
                            }
                        }
                        // an earlier op of the batch inserts an EQUAL tag (or is this plan's moved leader): planned again, the filter
                        // would find that op's entry and send this op to the exact executor anyway - send it there right away and
                        // save the round of planning and claims in between
                        if (bad) { pl.flags |= (why == 36 || why == 33) ? F_COMPLEX : F_CONF; chg = true; if (dbg) atomicAdd((unsigned long long *)&dbg[why], 1ULL); }
                        if (chg) { STG_(sh->pl_fl[j], pl.flags); S->lchg = 1; }
                    }
                    const unsigned ep2 = epoch++;
                    ZONE_CHECK();
                    CLAIM(ep2);          // in case anybody changed: every live plan is validated again under the next epoch
                    __syncthreads();
                    // (slot [it & 1]: a CTA that is already in the next iteration must not overwrite what a slower one still reads)
                    if (tid < CL && S->lchg) push_to(&S->rep.chg[it & 1][rank], tid, ep2);
                    csync(CL);
                    bool any = false;
                    for (int r = 0; r < CL; r++) any = any || S->rep.chg[it & 1][r] == ep2;
                    if (!any) break;
                    ep = ep2;
                }
                TICK(3);
            }
            // ---------------------------------------------------------------- [prefix over the valid plans of [pos, nr)] and the cut
            // (what the commit will ask about the next op's plan is fetched now: the latency hides behind the cut's hop)
            int nx_fl = 0, nx_w2 = -1;
            if (j >= pos && j + 1 < nr && (pl.flags & (F_INSERT | F_COMPLEX | F_CONF)) == F_INSERT) { nx_fl = LDG_(sh->pl_fl[j + 1]); nx_w2 = LDG_(sh->pl_w2[j + 1]); }
            const i64 cnt0 = x.run.count;
            const unsigned e_idx0 = x.e_idx;
            const int avail = (int)(x.e_cnt - e_idx0);
            bool c_free = false, exact = false;     // the part of the cut condition that does not depend on the prefix sums
            if (j >= pos && j < nr) {
                exact = (pl.flags & (F_COMPLEX | F_ZONE)) != 0 || (x.run.cur_len >= MRZ_MIN_MATCH && my_pos >= x.run.cur_p + MRZ_MIN_MATCH);   // an emission is due
                c_free = exact || (pl.flags & F_CONF) != 0;
            }
            {
                const unsigned bal = __ballot_sync(FULL, c_free), exb = __ballot_sync(FULL, exact);
                if (bal && lane == 0) {
                    const int l = __ffs(bal) - 1;
                    atomicMin(&S->lmin, ((j + l) << 1) | (int)((exb >> l) & 1u));
                }
            }
            if (need_conflict) {
                unsigned long long v = 0;
                if (j >= pos && j < nr && !(pl.flags & (F_COMPLEX | F_CONF))) {
                    v = 1ULL << 48;
                    if (pl.flags & F_INSERT) v |= (unsigned long long)pl.dcount | (1ULL << 16) | ((pl.flags & F_FOLLOW) ? (1ULL << 32) : 0ULL);
                }
                unsigned long long total;
                incl = cluster_excl_scan(v, S, 2, CL, rank, total, [&](unsigned long long ex) {
                    // the local minimum is complete (the atomics came before block_excl_scan's barriers): it crosses the cluster in
                    // the same hop as the scan totals, together with the prefix sum at the cut (what commits before it)
                    const int lm = S->lmin;
                    if (tid < CL) push_to(&S->rep.lcut[0][rank], tid, (unsigned long long)(unsigned)lm);
                    if (lm != 0x7fffffff && j == (lm >> 1))
                        for (int r = 0; r < CL; r++) push_to(&S->rep.lcutpre[rank], r, ex);
                }) + v;
                if (j < nr) STG_(sh->pre[j], incl);      // for the later rounds of this batch (visible behind the commit barrier)
                heavy_total = total;
                pre0 = pos;
                heavy = true;
                need_conflict = false;
            } else {
                __syncthreads();
                if (tid < CL) push_to(&S->rep.lcut[0][rank], tid, (unsigned long long)(unsigned)S->lmin);
                csync(CL);
            }
            int k = nr;          // first op of [pos, nr) that cannot be committed from its plan
            bool k_exact = false;
            {
                unsigned m = 0x7fffffffu;
                for (int r = 0; r < CL; r++) { const unsigned vv = (unsigned)S->rep.lcut[0][r]; m = vv < m ? vv : m; }
                if (m != 0x7fffffffu) { k = (int)(m >> 1); k_exact = (m & 1u) != 0; }
            }
            const unsigned long long basev = pos > pre0 ? LDG_(sh->pre[pos - 1]) : 0ULL;
            // d: the sums over the ops of [pos, k).  Right after the scan nobody can read pre[] of another CTA yet: the sums
            // come from the scan totals of the ranks below the cut's and the prefix its owner pushed along with it.
            unsigned long long d = 0;
            if (heavy) {
                if (k >= nr) d = heavy_total;
                else {
                    const int rk = k / T;
                    for (int r = 0; r < rk; r++) d += S->rep.tot[2][r];
                    d += S->rep.lcutpre[rk];
                }
            } else if (k > pos) d = LDG_(sh->pre[k - 1]) - basev;
            if (ovf_possible) {   // rare: an insert may need a sweep step the ring cannot serve (first overflow / ring at the end of the table)
                if (tid == 0) S->lmin = 0x7fffffff;
                __syncthreads();
                bool c = false;
                if (j >= pos && j < k && (pl.flags & F_INSERT) && !(pl.flags & (F_COMPLEX | F_CONF)) && pl.dcount == 1) {
                    const i64 Tc = cnt0 + (i64)PK_D(incl - basev);
                    if (Tc > limit && (!steady || Tc - limit > avail)) c = true;   // needs a serial sweep step
                }
                const unsigned bal = __ballot_sync(FULL, c);
                if (bal && lane == 0) atomicMin(&S->lmin, ((j + __ffs(bal) - 1) << 1) | 1);
                __syncthreads();
                if (tid < CL) push_to(&S->rep.lcut[1][rank], tid, (unsigned long long)(unsigned)S->lmin);
                csync(CL);
                unsigned m = 0x7fffffffu;
                for (int r = 0; r < CL; r++) { const unsigned vv = (unsigned)S->rep.lcut[1][r]; m = vv < m ? vv : m; }
                if (m != 0x7fffffffu && (int)(m >> 1) < k) {
                    k = (int)(m >> 1); k_exact = true;
                    d = k > pos ? LDG_(sh->pre[k - 1]) - basev : 0ULL;      // pre[] was stored before the barrier above
                }
            }
            // ---------------------------------------------------------------- commit [pos, k)
            int cleans = 0;
            i64 Tk = cnt0;
            if (k > pos) {
                Tk = cnt0 + (i64)PK_D(d);
                cleans = Tk > limit ? (int)(Tk - limit) : 0;
            }
            if (j >= pos && j < k && (pl.flags & F_INSERT)) {
                if (pl.flags & F_W1) {
                    int s1 = pl.w1_slot; CHK(sh, x, 3, s1, hsize);
                    __stcg(toff + s1, pl.w1_off);
                    __stcg(ttag + s1, pl.w1_tag);
                    sum_lower(gsum, s1 >> GRP_SHIFT, tz1c(pl.w1_tag));
                    if (steady && (pl.flags & F_W1ELIG) && s1 >= w_lo && s1 < w_hi) atomicMin(&S->lq, s1);   // a ring slot the sweep does not reach in this batch
                }
                // a twin that overwrites its leader's entry in the same commit must win the slot
                const bool overwritten = j + 1 < k && (nx_fl & F_FOLLOW) && nx_w2 == pl.w2_slot;
                int s2 = pl.w2_slot; CHK(sh, x, 4, s2, hsize);
                if (!overwritten) {
                    __stcg(toff + s2, my_pos);
                    __stcg(ttag + s2, my_tag);
                    sum_lower(gsum, s2 >> GRP_SHIFT, tz1c(my_tag));
                }
                if (steady && (pl.flags & F_W2ELIG) && s2 >= w_lo && s2 < w_hi) atomicMin(&S->lq, s2);
                filt_set(filt, my_tag);
                if (pl.flags & F_FOLLOW) filt_set(filt, my_tag ^ DUP_SALT);   // the tag is in the table twice now
            }
            for (int q = j; q < cleans; q += B) {
                int sl = LDG_(sh->E[(e_idx0 + (unsigned)q) & (MRZ_EQ_CAP - 1)]); CHK(sh, x, 5, sl, hsize);
                __stcg(ttag + sl, (i64)0);
                __stcg(toff + sl, (i64)0);
                sum_lower(gsum, sl >> GRP_SHIFT, 0);
            }
#ifdef MRZ_CUSIM
            dbg_done += k - pos;
            if (j == 0 && k > pos && (int)PK_V(d) != k - pos)
                g_selfcheck_failures++, fprintf(stderr, "ACCOUNTING: commit [%d,%d) of nr %d counts %u ops (heavy %d pre0 %d)\n", pos, k, nr, PK_V(d), (int)heavy, pre0);
#endif
            if (tid == 0 && k > pos) {           // every CTA applies the same update to its replica
                x.run.count = Tk - cleans;
                x.run.inserts += PK_I(d);
                x.run.misses += PK_M(d);
                x.run.nops += PK_V(d);
                if (cleans) {
                    x.e_idx = e_idx0 + (unsigned)cleans;
                    x.run.clean_ptr = LDG_(sh->E[(e_idx0 + (unsigned)cleans - 1) & (MRZ_EQ_CAP - 1)]);   // pointer stays on the last zeroed slot
                }
            }
            if (k > pos) {       // (nothing committed - an exact op right behind another one: no writes to publish, no hop)
                __syncthreads();
                if (tid < CL) push_to(&S->rep.qmin[rank], tid, S->lq);
                csync(CL);
                for (int r = 0; r < CL; r++) qacc = S->rep.qmin[r] < qacc ? S->rep.qmin[r] : qacc;
            }
            TICK(4);
            if (k >= nr) break;
            if (dbg && j == k) {   // why does this op cut the batch?
                const int why = (pl.flags & F_COMPLEX) ? (follower ? 22 : ((pl.flags & F_INSERT) ? 23 : 21)) : k_exact ? 25 : 24;
                atomicAdd((unsigned long long *)&dbg[why], 1ULL);
            }
            if (dbg && j == 0) tm[11] += 1;
            if (k_exact) {
                if (dbg && j == 0) tm[10] += 1;
                // ---- the lookup of op k (find_best_match, rzip.c:426-462) by the whole CTA (rank 0): the probe sequence is
                // walked 4*T slots per step, equal tags are collected in slot order, and the warps share the match_len calls
                // At the levels with long chains (8, 9: up to 32 / 128 equal tags per lookup, dense repeats) EVERY CTA of the
                // cluster walks the probe sequence itself - same table, same result - and, when there are more candidates than
                // one CTA has warps, takes its share of the compares; their results reach rank 0 in one more hop.
                const bool lead = rank == 0;
                const bool help = CL > 1 && chain > 2 * (T >> 5);
                bool many = false;          // uniform over the cluster: the candidates of this lookup are shared out
                if (lead || help) {
                    int pre_n = -2;   // -2: the filter says no equal tag; -1: ask / walk on the warp; >= 0: candidates compared
                    const i64 c_x0 = clock64();
                    const i64 tk = LDG_(sh->b_tag[k]), pk = LDG_(sh->b_pos[k]);
                    const i64 c_lk0 = clock64();
                    i64 c_lk1 = c_lk0, c_lk2 = c_lk0;
                    if (dbg && lead && tid == 0) dbg[53] += c_lk0 - c_x0;
                    {
                        // ONE round trip: the filter word and the first 4*T slots of the probe sequence, tags and offsets alike
                        const i64 h = tk & hmask;
                        u64 fw, fb;
                        filt_hash(filt, tk, fw, fb);
                        const u64 fword = __ldcg(filt.w + fw);
                        i64 et[4], eo[4];
#pragma unroll
                        for (int q = 0; q < 4; q++) et[q] = __ldcg(ttag + ((h + tid * 4 + q) & hmask));
#pragma unroll
                        for (int q = 0; q < 4; q++) eo[q] = __ldcg(toff + ((h + tid * 4 + q) & hmask));
                        if ((fword & fb) == fb) {
                            pre_n = -1;
                            if (tid == 0) S->lk_fe = 0x7fffffff;
                            __syncthreads();
                            int n_tot = 0;
                            bool found = false;
                            for (int base = 0;; base += T * 4) {
                                if (base) {
#pragma unroll
                                    for (int q = 0; q < 4; q++) et[q] = __ldcg(ttag + ((h + base + tid * 4 + q) & hmask));
#pragma unroll
                                    for (int q = 0; q < 4; q++) eo[q] = __ldcg(toff + ((h + base + tid * 4 + q) & hmask));
                                }
                                unsigned eq4 = 0;
#pragma unroll
                                for (int q = 0; q < 4; q++) {
                                    if (et[q] == 0) atomicMin(&S->lk_fe, base + tid * 4 + q);
                                    else if (et[q] == tk) eq4 |= 1u << q;
                                }
                                __syncthreads();
                                const int fe = S->lk_fe;
#pragma unroll
                                for (int q = 0; q < 4; q++)
                                    if (base + tid * 4 + q >= fe) eq4 &= ~(1u << q);   // beyond the first empty slot: not on the walk
                                unsigned long long total;
                                int ex = (int)block_excl_scan((unsigned long long)__popc(eq4), S->wsum, total);
#pragma unroll
                                for (int q = 0; q < 4; q++)
                                    if ((eq4 >> q) & 1u) {
                                        if (n_tot + ex < LK_CAP) { S->lk_pos[n_tot + ex] = base + tid * 4 + q; S->lk_off[n_tot + ex] = eo[q]; }
                                        ex++;
                                    }
                                n_tot += (int)total;
                                found = fe != 0x7fffffff;
                                if (found || n_tot > LK_CAP || base + T * 4 >= hsize) break;
                            }
                            if (tid == 0) S->lk_n = 0;      // candidates still matching at the cap
                            __syncthreads();
                            c_lk1 = c_lk2 = clock64();
                            if (found && n_tot <= LK_CAP) {
                                // single_match_len (rzip.c:372-397) per candidate: the warps share the candidates; the first 2 KiB of
                                // both directions come in one round trip, LK_XCAP bytes either side by the warp; whatever still matches
                                // there is extended by the whole CTA
                                w.last = x.run.last;
                                const i64 lo_m = w.last > 0 ? w.last : 0;
                                const i64 maxf = end - pk;
                                const int NW = T >> 5;
                                many = help && n_tot > NW;
                                if (many) {
                                    // candidate q belongs to rank q mod CL: its warps compare (up to LK_XCAP2 either side), the
                                    // whole CTA extends what still matches there, and the result goes to rank 0
                                    const int mine = rank < n_tot ? (n_tot - rank + CL - 1) / CL : 0;
                                    for (int i = wid; i < mine; i += NW) {
                                        const int q = rank + CL * i;
                                        const i64 off = S->lk_off[q];
                                        i64 f = 0, r = 0;
                                        if (off < pk) {
                                            i64 maxb = pk - lo_m;
                                            if (off < maxb) maxb = off;
                                            const i64 mf = maxf < LK_XCAP ? maxf : LK_XCAP, mb = maxb < LK_XCAP ? maxb : LK_XCAP;
                                            warp_cmp2(buf, pk, off, mf, mb, lane, f, r);
                                            if (f == LK_XCAP && maxf > LK_XCAP) f += w.extend(pk + LK_XCAP, off + LK_XCAP, (maxf < LK_XCAP2 ? maxf : (i64)LK_XCAP2) - LK_XCAP, true);
                                            if (r == LK_XCAP && maxb > LK_XCAP) r += w.extend(pk - LK_XCAP, off - LK_XCAP, (maxb < LK_XCAP2 ? maxb : (i64)LK_XCAP2) - LK_XCAP, false);
                                            __syncwarp();
                                        }
                                        if (lane == 0) { S->lk_f[q] = f; S->lk_r[q] = r; }
                                    }
                                    __syncthreads();
                                    for (int i = 0; i < mine; i++) {
                                        const int q = rank + CL * i;
                                        const i64 off = S->lk_off[q];
                                        i64 f = S->lk_f[q], r = S->lk_r[q], m = 0;
                                        if (off < pk) {
                                            i64 maxb = pk - lo_m;
                                            if (off < maxb) maxb = off;
                                            if (f == LK_XCAP2 && maxf > LK_XCAP2) f += cta_extend(S, buf, pk + LK_XCAP2, off + LK_XCAP2, maxf - LK_XCAP2, true);
                                            if (r == LK_XCAP2 && maxb > LK_XCAP2) r += cta_extend(S, buf, pk - LK_XCAP2, off - LK_XCAP2, maxb - LK_XCAP2, false);
                                            m = f + r < MRZ_MIN_MATCH ? 0 : f + r;
                                        }
                                        if (tid == 0) { push_to(&S->lk_len[q], 0, m); push_to(&S->lk_rev[q], 0, r); }
                                    }
                                } else if (!lead) {
                                    // (nothing to help with)
                                } else if (n_tot <= NW) {
                                    // few candidates (the usual case): one after the other, each by the whole CTA, complete
                                    for (int q = 0; q < n_tot; q++) {
                                        const i64 off = S->lk_off[q];
                                        i64 f = 0, r = 0;
                                        if (off < pk) {
                                            i64 maxb = pk - lo_m;
                                            if (off < maxb) maxb = off;
                                            cta_cmp2(S, buf, pk, off, maxf, maxb, f, r);
                                        }
                                        if (tid == 0) { S->lk_len[q] = (off < pk && f + r >= MRZ_MIN_MATCH) ? f + r : 0; S->lk_rev[q] = r; }
                                    }
                                    __syncthreads();
                                    c_lk2 = clock64();
                                } else {
                                for (int q = wid; q < n_tot; q += NW) {
                                    const i64 off = S->lk_off[q];
                                    i64 f = 0, r = 0;
                                    if (off < pk) {
                                        i64 maxb = pk - lo_m;
                                        if (off < maxb) maxb = off;
                                        const i64 mf = maxf < LK_XCAP ? maxf : LK_XCAP, mb = maxb < LK_XCAP ? maxb : LK_XCAP;
                                        warp_cmp2(buf, pk, off, mf, mb, lane, f, r);
                                        if (lane == 0 && ((f == LK_XCAP && maxf > LK_XCAP) || (r == LK_XCAP && maxb > LK_XCAP))) atomicAdd(&S->lk_n, 1);
                                    }
                                    if (lane == 0) { S->lk_f[q] = f; S->lk_r[q] = r; }
                                }
                                __syncthreads();
                                c_lk2 = clock64();
                                // many long candidates (dense repeats, deep chains): the warps carry on in parallel, one candidate
                                // each, up to LK_XCAP2; the whole CTA then only extends what is still matching there
                                const i64 cap = S->lk_n > 2 ? (i64)LK_XCAP2 : (i64)LK_XCAP;
                                if (cap != LK_XCAP) {
                                    for (int q = wid; q < n_tot; q += NW) {
                                        const i64 off = S->lk_off[q];
                                        if (off >= pk) continue;
                                        i64 f = S->lk_f[q], r = S->lk_r[q];
                                        i64 maxb = pk - lo_m;
                                        if (off < maxb) maxb = off;
                                        if (f == LK_XCAP && maxf > LK_XCAP) f += w.extend(pk + LK_XCAP, off + LK_XCAP, (maxf < cap ? maxf : cap) - LK_XCAP, true);
                                        if (r == LK_XCAP && maxb > LK_XCAP) r += w.extend(pk - LK_XCAP, off - LK_XCAP, (maxb < cap ? maxb : cap) - LK_XCAP, false);
                                        __syncwarp();
                                        if (lane == 0) { S->lk_f[q] = f; S->lk_r[q] = r; }
                                    }
                                    __syncthreads();
                                }
                                if (S->lk_n == 0) {          // the usual case: nothing reaches the cap, no CTA-wide extension
                                    for (int q = tid; q < n_tot; q += T) {
                                        const i64 f = S->lk_f[q], r = S->lk_r[q];
                                        S->lk_len[q] = (S->lk_off[q] < pk && f + r >= MRZ_MIN_MATCH) ? f + r : 0;
                                        S->lk_rev[q] = r;
                                    }
                                } else
                                for (int q = 0; q < n_tot; q++) {
                                    const i64 off = S->lk_off[q];
                                    i64 f = S->lk_f[q], r = S->lk_r[q], m = 0;
                                    if (off < pk) {
                                        i64 maxb = pk - lo_m;
                                        if (off < maxb) maxb = off;
                                        if (f == cap && maxf > cap) f += cta_extend(S, buf, pk + cap, off + cap, maxf - cap, true);
                                        if (r == cap && maxb > cap) r += cta_extend(S, buf, pk - cap, off - cap, maxb - cap, false);
                                        m = f + r < MRZ_MIN_MATCH ? 0 : f + r;
                                    }
                                    if (tid == 0) { S->lk_len[q] = m; S->lk_rev[q] = r; }
                                }
                                }
                                __syncthreads();
                                pre_n = n_tot;
                            }
                        }
                    }
                    // Every CTA has finished its walk before rank 0's op may change the table (otherwise a slow CTA could see
                    // the op's own insert and count one candidate more); the shared-out compares arrive by the same barrier.
                    if (help) csync(CL);
                    if (dbg && lead && tid == 0) {
                        const i64 c_now = clock64();
                        dbg[18] += c_now - c_lk0; dbg[19] += pre_n != -2;
                        dbg[50] += c_lk1 - c_lk0; dbg[51] += c_lk2 - c_lk1; dbg[52] += c_now - c_lk2;
                    }
                    // ---- op k on the live table, reference semantics (warp 0), on the local replica
                    if (lead && wid == 0) {
                        const i64 c_pre0 = clock64();
                        w.load(x.run);
                        w.nlog = 0;
                        i64 hint = w.clean_ptr;
                        int head0 = -1, head1 = -1;              // the ring's first two slots, read in one round trip
                        if (steady && x.e_cnt != x.e_idx) {
                            head0 = LDG_(sh->E[x.e_idx & (MRZ_EQ_CAP - 1)]);
                            if (x.e_cnt - x.e_idx > 1) head1 = LDG_(sh->E[(x.e_idx + 1) & (MRZ_EQ_CAP - 1)]);
                            hint = head0;
                        } else if (steady)
                            hint = x.e_scan > hint ? x.e_scan : hint;
                        w.clean_hint = hint;
                        w.p = pk;
                        const i64 c_op0 = clock64();
                        const bool jumped = w.op(pk, tk, S->stk, pre_n, S->lk_len, S->lk_rev, S->lk_off, S->lk_pos);
                        const i64 c_op1 = clock64();
                        w.clean_hint = 0;
                        w.store(x.run);
                        const bool mask_changed = w.min_mask != min_mask || w.ins_mask != ins_mask;
                        const bool overflow = w.nlog > MRZ_WLOG_CAP;
                        int kp = nr;
                        // an emission skips the ops up to the end of the match; the plans beyond it stay good (only a write to a
                        // plan's stop slot can spoil it): carry on with the first op past the jump
                        // (a short match can end BEFORE op k: p goes back, rzip.c:597: the batch ends)
                        if (jumped && !mask_changed && !overflow && !w.error && w.p >= pk) kp = pos_upper_bound(sh, k + 1, nr, w.p, lane);
                        const bool ends = mask_changed || overflow || w.error || (jumped && kp >= nr);
                        if (lane == 0) {
                            x.jumped = jumped;
                            x.nlog = overflow ? 0 : w.nlog;
                            x.ended = ends;
                            if (ends) {
                                x.e_idx = x.e_cnt = 0; x.e_scan = w.clean_ptr;    // rebuild the ring
                                // what this batch had gathered behind p is gathered again (a jump past the batch keeps the cursor)
                                if (w.p + 1 < x.gp) x.gp = w.p + 1;
                            } else
                                x.pos_after = jumped ? kp : k + 1;
                        }
                        __syncwarp();
                        const i64 c_sq0 = clock64();
                        if (!ends && steady) {
                            sync_queue(sh, &x, x.wlog, w.nlog, lane, head0, head1);
                            if (lane == 0 && x.e_cnt == x.e_idx && x.e_scan < w.clean_ptr) x.e_scan = w.clean_ptr;
                            __syncwarp();
                        }
                        if (dbg && lane == 0) {
                            const i64 c_end = clock64();
                            dbg[54] += c_op1 - c_op0; dbg[55] += c_end - c_op1;
                            dbg[56] += c_op0 - c_pre0; dbg[57] += c_end - c_sq0;
                        }
                    }
                    // the result goes into every other replica: one destination CTA per warp (a remote store costs the issuing
                    // warp ~200 clk; one warp pushing to 7 CTAs in turn was 1.5k clk of every exact op).  The write log is the
                    // tail of XRes: only its used part travels - usually one store per lane.
                    if (lead && CL > 1) {
                        __syncthreads();
                        const i64 c_pu0 = clock64();
                        const int nwords = (int)((offsetof(XRes, wlog) + 4 * (size_t)(x.nlog > 0 ? x.nlog : 0) + 7) / 8);
                        const unsigned long long *src = reinterpret_cast<const unsigned long long *>(&x);
                        for (int r = 1 + wid; r < CL; r += T >> 5) {
                            volatile unsigned long long *dst = reinterpret_cast<volatile unsigned long long *>(remote_ptr(&x, r));
                            for (int q = lane; q < nwords; q += 32) dst[q] = src[q];
                        }
                        if (dbg && tid == 0) dbg[58] += clock64() - c_pu0;
                    }
                }
                csync(CL);
                TICK(5);
                if (x.ended) { batch_ended = true; break; }
                // the op overwrote a queued slot and the ring was cut there: rebuild its tail, but only inside the
                // window this batch was planned with
                if (steady && x.e_scan < w_hi) refill_queue(sh, S, ttag, btm, w_hi, MRZ_EQ_CAP, CL, rank, j, B);
                const int npos = x.pos_after;   // k + 1, or the first op past an emission's jump (the ops in between never run)
#ifdef MRZ_CUSIM
                dbg_done += 1; dbg_skip += npos - k - 1;
#endif
                // later plans that read a slot this op wrote are no longer valid; nor is its twin's plan
                if (j == npos && (pl.flags & F_FOLLOW)) pl.flags |= F_CONF;
                if (j >= npos && j < nr)   // a plan forwarded over the claim of op k / of a skipped op assumed that op's planned write
                    for (int dd = 0; dd < pl.ndep; dd++)
                        if (pl.dep_own[dd] >= k && pl.dep_own[dd] < npos) pl.flags |= F_CONF;
                if (j >= npos && j < nr && (pl.flags & F_FOLLOW) && j - 1 >= k && j - 1 < npos) pl.flags |= F_CONF;   // its leader was skipped
                if (j >= npos && j < nr && !(pl.flags & (F_COMPLEX | F_CONF))) {
                    const int nl = x.nlog;
                    for (int q = 0; q < nl; q++) {
                        const int sl = x.wlog[q] >> 2;
                        bool hit = sl == pl.w2_slot || ((pl.flags & F_W1) && sl == pl.w1_slot);
                        for (int dd = 0; dd < pl.ndep; dd++) hit = hit || sl == pl.dep_slot[dd];
                        if (hit) { pl.flags |= F_CONF; break; }
                    }
                }
                pos = npos;
            } else {
                // ------------------------------------------------------------ plan again what lost its plan
                // All ops before `pos` are in the table now, so the first of these plans is exact; the others are as
                // speculative as a first plan and are validated by the next round of claims.  A twin becomes an ordinary
                // op: if its leader has been committed, the filter finds the leader's entry.
                pos = k;
                if (j >= pos && j < nr && (pl.flags & F_CONF) && !(pl.flags & F_COMPLEX)) {
                    plan_init(pl);
                    follower = false;
                    lead_w2 = -1;
                    if (filt_query(filt, my_tag)) pl.flags |= F_COMPLEX;
                    else if ((my_tag & ins_mask) == ins_mask)
                        plan_op(pl, ttag, toff, gsum, filt, my_tag, my_tag & hmask, hmask, hsize, min_mask, chain);
                    STG_(sh->pl_fl[j], pl.flags); STG_(sh->pl_w2[j], pl.w2_slot);
                    if (dbg) atomicAdd((unsigned long long *)&dbg[30], 1ULL);
                }
                need_conflict = true;
            }
            TICK(6);
        }
#undef ZONE_CHECK
#undef CLAIM
#ifdef MRZ_CUSIM
        if (j == 0 && !batch_ended && (dbg_done + dbg_skip != nr || x.run.nops - dbg_nops0 != dbg_done))
            g_selfcheck_failures++, fprintf(stderr, "ACCOUNTING: batch of %d ops: %d done + %d skipped, n_ops grew by %lld\n", nr, dbg_done, dbg_skip, (long long)(x.run.nops - dbg_nops0));
        if (j == 0 && batch_ended && x.run.nops - dbg_nops0 != dbg_done + 1)
            g_selfcheck_failures++, fprintf(stderr, "ACCOUNTING: ended batch of %d ops: %d done + the cut op, n_ops grew by %lld\n", nr, dbg_done, (long long)(x.run.nops - dbg_nops0));
#endif
        // ================================================================ batch epilogue
        // a committed plan overwrote a slot of the ring that the sweep had not reached: cut the ring there (every CTA, its replica)
        if (!batch_ended && steady && qacc != 0x7fffffff && wid == 0) {
            const unsigned a = ring_lower_bound(sh, x.e_idx, x.e_cnt, qacc, lane);
            if (lane == 0 && a != x.e_cnt) { x.e_cnt = a; x.e_scan = qacc; }
        }
        __syncthreads();
        TICK(7);
    }
    csync(CL);
    if (j == 0 && dbg) {
        for (int q = 0; q < 12; q++) dbg[q] += tm[q];
    }
    if (j == 0) {
        i64 p = x.run.p;
        const int err = x.run.error;
        if (!err && p < end) p = end;
        st->p = p; st->last = x.run.last; st->cur_p = x.run.cur_p; st->cur_len = x.run.cur_len; st->cur_ofs = x.run.cur_ofs;
        st->min_mask = x.run.min_mask; st->ins_mask = x.run.ins_mask; st->clean_ptr = x.run.clean_ptr; st->count = x.run.count;
        st->victim_round = x.run.vr;
        st->tag_hits = x.run.hits; st->tag_misses = x.run.misses; st->inserts = x.run.inserts; st->victim_events = x.run.vevents;
        st->n_ops = x.run.nops; st->ntok = x.run.ntok;
        st->err_a = sh->err_a; st->err_b = sh->err_b;
        if (err) st->error = err;
        else st->done = 1;
    }
#undef TICK
}

#ifndef MRZ_CUSIM
// function attributes: once per device, never rewritten under a concurrent launch from another host thread
static int set_attributes_once() {
    static std::mutex mu;
    static bool set_for[64];
    int dev = 0;
    cudaGetDevice(&dev);
    std::lock_guard<std::mutex> g(mu);
    if (set_for[dev & 63]) return 0;
    cudaError_t e = cudaFuncSetAttribute(k_automaton, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(Smem));
    if (e != cudaSuccess) return (int)e;
    e = cudaFuncSetAttribute(k_automaton, cudaFuncAttributeNonPortableClusterSizeAllowed, 1);
    if (e != cudaSuccess) return (int)e;
    // room for a persisting window over one automaton's hash table (mrz_launch_automata); failure is not an error
    {
        int maxp = 0;
        if (cudaDeviceGetAttribute(&maxp, cudaDevAttrMaxPersistingL2CacheSize, dev) == cudaSuccess && maxp > 0)
            cudaDeviceSetLimit(cudaLimitPersistingL2CacheSize, (size_t)maxp);
        cudaGetLastError();
    }
    set_for[dev & 63] = true;
    return 0;
}
static void cluster_config(cudaLaunchConfig_t &cfg, cudaLaunchAttribute *at, int nclusters, int CL, int T, cudaStream_t s) {
    memset(&cfg, 0, sizeof cfg);
    cfg.gridDim = dim3((unsigned)(nclusters * CL));
    cfg.blockDim = dim3((unsigned)T);
    cfg.dynamicSmemBytes = sizeof(Smem);
    cfg.stream = s;
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = (unsigned)CL; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
    cfg.attrs = at; cfg.numAttrs = 1;
}
#endif

// How many clusters of CL CTAs x T threads the device holds at once (0 on error).
extern "C" int mrz_max_active_clusters(int CL, int T) {
#ifndef MRZ_CUSIM
    if (set_attributes_once()) return 0;
    cudaLaunchConfig_t cfg;
    cudaLaunchAttribute at[1];
    cluster_config(cfg, at, 1, CL, T, 0);
    int n = 0;
    if (cudaOccupancyMaxActiveClusters(&n, k_automaton, &cfg) != cudaSuccess) { cudaGetLastError(); return 0; }
    return n;
#else
    (void)CL; (void)T;
    return 1 << 20;
#endif
}

// nclusters automata (descs[0..nclusters)), CL CTAs of T threads each, one launch.
// hot / hot_bytes: a range the launch asks L2 to keep (the hash table of a lone automaton: every walk of the plan phase and
// of the exact executor is a chain of dependent reads of it, 375 clk from L2 against 1334 clk from HBM); 0 bytes = none.
extern "C" int mrz_launch_automata(const AutoDesc *descs, int nclusters, int CL, int T, cudaStream_t s, const void *hot, size_t hot_bytes) {
#ifndef MRZ_CUSIM
    cudaError_t e;
    {
        int rc = set_attributes_once();
        if (rc) return rc;
    }
    cudaLaunchConfig_t cfg;
    cudaLaunchAttribute at[2];
    cluster_config(cfg, at, nclusters, CL, T, s);
    if (hot && hot_bytes) {
        int dev = 0, maxw = 0, maxp = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&maxw, cudaDevAttrMaxAccessPolicyWindowSize, dev);
        cudaDeviceGetAttribute(&maxp, cudaDevAttrMaxPersistingL2CacheSize, dev);
        if (maxw > 0 && maxp > 0) {
            const size_t nb = hot_bytes < (size_t)maxw ? hot_bytes : (size_t)maxw;
            at[1].id = cudaLaunchAttributeAccessPolicyWindow;
            at[1].val.accessPolicyWindow.base_ptr = const_cast<void *>(hot);
            at[1].val.accessPolicyWindow.num_bytes = nb;
            at[1].val.accessPolicyWindow.hitRatio = (size_t)maxp >= nb ? 1.0f : (float)((double)maxp / (double)nb);
            at[1].val.accessPolicyWindow.hitProp = cudaAccessPropertyPersisting;
            at[1].val.accessPolicyWindow.missProp = cudaAccessPropertyStreaming;
            cfg.numAttrs = 2;
        }
        cudaGetLastError();
    }
    e = cudaLaunchKernelEx(&cfg, k_automaton, descs);
    return (int)e;
#else
    (void)hot; (void)hot_bytes;
    MRZ_LAUNCH_CLUSTER(k_automaton, nclusters * CL, CL, T, sizeof(Smem), s, descs);
    return 0;
#endif
}

