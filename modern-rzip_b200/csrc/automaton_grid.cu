// automaton_grid.cu — the table automaton of hash_search, exact, spread over many SMs (v5).
//
// Same semantics as automaton.cu (find_best_match rzip.c:426-462, insert_hash rzip.c:256-301,
// clean_one_from_hash rzip.c:305-328, single_match_len rzip.c:372-397, the greedy rule rzip.c:571-599);
// what changes is where the speculation runs.  The one-CTA automaton is bound by what a single SM can
// keep in flight (its L1 wavefronts and its ~64 B/clk port to L2): a batch of 1024 plans took ~80 us.
// Here one cooperative grid of G CTAs x T threads plans a batch of (G-1)*T candidate ops, one op per
// thread, each SM holding only T ops; the phases of a batch are separated by grid barriers and all
// shared state (running state, sweep queue, prefix sums, claims) lives in L2-resident global memory:
//
//   phase A  gather + plan (read-only against the table) + CLAIM the stop slots: claim[slot] =
//            atomicMin((epoch, op index)) in a 4 Mi-entry array - no hash, no clearing;
//            meanwhile the last CTA (the "service" CTA) refills the sweep queue of clean_one_from_hash;
//   phase B  conflict = some earlier op of the batch claimed one of MY stop slots.  That is exact, not
//            conservative, because an insert only ever RAISES the trailing-one count of the slot it
//            writes (empty / due-for-sweep -> >= need, lesser -> greater): a walk that passed a slot
//            still passes it, so only a write to a walk's stop slot can change its outcome; and an
//            entry due for sweeping stops every walk, so a slot of the sweep queue can only be a stop
//            slot too (oracle/spec_model.c, proved equivalent by tests/test_spec_model.py);
//            then the prefix sums (count delta, inserts, misses, ops) across the grid;
//   phase C  first op that cannot commit from its plan (grid-wide atomicMin);
//   phase D  commit of the prefix in parallel, sweeps from the queue; then the cut op runs on warp 0
//            of CTA 0 with the reference semantics and later plans that claimed a slot it wrote are
//            invalidated.
#include "mrz_common.cuh"

namespace g5 {

#define FULL 0xffffffffu
#define G5_EQ_CAP 16384       // eligible-slot ring (power of two)
#define G5_MAXWALK 16384      // longer insert walks are left to the serial path
#define G5_WLOG_CAP 64
#define G5_MAXB 8192          // ops per batch (13 bits of a claim key)
#define G5_MAXG 160
#define GRP 16                // slots per summary group
#define GRP_SHIFT 4
#define DUP_SALT 0x5bd1e9955bd1e995LL

enum { F_COMPLEX = 1, F_INSERT = 2, F_W1 = 4, F_CONF = 8, F_FOLLOW = 16, F_W2ELIG = 32, F_W1ELIG = 64 };

// Blocked Bloom filter over inserted tags (see automaton.cu / oracle/spec_model.c filt_hash).
#define FILT_BITS 24
__device__ __forceinline__ void filt_hash(i64 tag, u64 &word, u64 &bits) {
    u64 x = (u64)tag;
    x ^= x >> 33; x *= 0xff51afd7ed558ccdULL; x ^= x >> 33; x *= 0xc4ceb9fe1a85ec53ULL; x ^= x >> 33;
    word = x >> (64 - FILT_BITS);
    bits = ((u64)1 << (x & 63)) | ((u64)1 << ((x >> 6) & 63)) | ((u64)1 << ((x >> 12) & 63));
}
__device__ __forceinline__ bool filt_query(const u64 *F, i64 tag) {
    u64 w, b;
    filt_hash(tag, w, b);
    return (__ldcg(F + w) & b) == b;
}
__device__ __forceinline__ void filt_set(u64 *F, i64 tag) {
    u64 w, b;
    filt_hash(tag, w, b);
    atomicOr((unsigned long long *)(F + w), (unsigned long long)b);
}

__device__ __forceinline__ int tz1c(i64 t) {   // trailing one bits, capped at 15 (summary nibble)
    int z = tz1(t);
    return z > 15 ? 15 : (z < 0 ? 15 : z);
}
// Group summary in global memory (L2 resident, 128 KiB): one nibble per 16 slots, a lower bound of the
// minimum trailing-one count in the group.
__device__ __forceinline__ int nib_of(u32 w, i64 g) { return (int)((w >> ((g & 7) * 4)) & 15u); }
// plan phases only (nobody lowers concurrently): a lost update merely leaves a smaller, still valid, bound
__device__ __forceinline__ void nib_raise(u32 *gs, i64 g, int v) {
    u32 w = __ldcg(gs + (g >> 3));
    int sh = (int)(g & 7) * 4;
    if ((int)((w >> sh) & 15u) < v) __stcg(gs + (g >> 3), (w & ~(15u << sh)) | ((u32)v << sh));
}
// commit / serial phases: the group now holds a tag with v trailing ones (0 for a zeroed slot)
__device__ __forceinline__ void nib_lower(u32 *gs, i64 g, int v) {
    u32 *p = gs + (g >> 3);
    int sh = (int)(g & 7) * 4;
    u32 old = __ldcg(p);
    for (;;) {
        if ((int)((old >> sh) & 15u) <= v) return;
        u32 nw = (old & ~(15u << sh)) | ((u32)v << sh);
        u32 seen = atomicCAS(p, old, nw);
        if (seen == old) return;
        old = seen;
    }
}

// First slot s >= h with tz1(tags[s]) < need (an empty slot has tag 0), or -1 when the walk reaches the
// end of the table / gets too long.  Groups whose summary nibble is >= need are skipped unseen; a group
// scanned completely without a hit has its nibble repaired.
__device__ __forceinline__ i64 find_stop(const i64 *ttag, u32 *gsum, i64 h, int need, i64 hsize, i64 &stop_tag) {
    if (h >= hsize) return -1;
    const i64 h0 = h;
    i64 g = h >> GRP_SHIFT;
    i64 ngroups = hsize >> GRP_SHIFT;
    {
        i64 lim = ((h0 + G5_MAXWALK) >> GRP_SHIFT) + 1;
        if (lim < ngroups) ngroups = lim;
    }
    bool whole = (h & (GRP - 1)) == 0;
    for (;;) {
        const i64 gend = (g + 1) << GRP_SHIFT;
        int mn = 15;
        for (i64 o = h & ~(i64)7; o < gend; o += 8) {
            longlong2 v[4];
#pragma unroll
            for (int q = 0; q < 4; q++) v[q] = __ldcg(reinterpret_cast<const longlong2 *>(ttag + o) + q);
#pragma unroll
            for (int q = 0; q < 8; q++) {
                const i64 t = (q & 1) ? v[q >> 1].y : v[q >> 1].x;
                if (o + q >= h) {
                    int z = tz1(t);
                    if (z < 0) z = 64;
                    if (z < need) { stop_tag = t; return o + q; }
                    int zc = z > 15 ? 15 : z;
                    mn = zc < mn ? zc : mn;
                }
            }
        }
        if (whole) nib_raise(gsum, g, mn);
        g++;
        // skip groups that cannot contain a stop: one 32-bit word covers 8 groups (128 slots)
        while (g < ngroups) {
            u32 w = __ldcg(gsum + (g >> 3));
            bool found = false;
            for (; g < ngroups; g++) {
                if (nib_of(w, g) < need) { found = true; break; }
                if ((g & 7) == 7) { g++; break; }
            }
            if (found) break;
        }
        if (g >= ngroups) return -1;
        h = g << GRP_SHIFT;
        whole = true;
    }
}

// Running state of the automaton (rzip_state, mrzip_private.h:388-416).
struct Run {
    i64 p, last, cur_p, cur_len, cur_ofs, min_mask, ins_mask, clean_ptr, count, vr;
    i64 hits, misses, inserts, vevents, nops, ntok;
    int error, pad;
};

// State shared by the whole grid, in global memory (L2).
struct GState {
    Run run;
    i64 e_scan;              // slots in [run.clean_ptr, e_scan) have been examined for the queue
    i64 li;                  // next candidate-list index
    unsigned long long kcut; // cut search: min over (~round << 32 | op index)
    unsigned e_idx, e_cnt;   // ring counters (monotonic)
    unsigned bar;            // grid barrier counter (zeroed by the host before every launch)
    unsigned epoch;          // batch number of this chunk: claims of older batches compare greater
    int pos, ended, jumped, bmax, nlog, pad;
    int wlog[G5_WLOG_CAP];
    int E[G5_EQ_CAP];        // ring of eligible slots, ascending
    unsigned long long cta_tot[G5_MAXG];
    unsigned long long pre[G5_MAXB];   // inclusive prefix over ops holding a valid plan
};

#define PK_D(x) ((unsigned)((x) & 0xffffu))            // hash_count delta
#define PK_I(x) ((unsigned)(((x) >> 16) & 0xffffu))    // inserts
#define PK_M(x) ((unsigned)(((x) >> 32) & 0xffffu))    // tag misses (twins)
#define PK_V(x) ((unsigned)(((x) >> 48) & 0xffffu))    // ops

// The serial executor: one warp runs one op on the live table with the reference semantics.
struct Warp {
    const u8 *buf;
    i64 *ttag, *toff;
    u64 *filt;
    u32 *gsum;
    Token *toks;
    i64 n, end, hmask, hsize, limit, tok_cap;
    int chain, lane;
    i64 p, last, cur_p, cur_len, cur_ofs, min_mask, ins_mask, clean_ptr, count, vr;
    i64 hits, misses, inserts, vevents, nops, ntok;
    int error;
    int *wlog;        // (slot << 1) | zeroed; nlog > G5_WLOG_CAP means overflow
    int nlog;
    i64 clean_hint;   // no eligible slot exists in [clean_ptr, clean_hint)

    __device__ __forceinline__ void load(const Run *r) {
        p = r->p; last = r->last; cur_p = r->cur_p; cur_len = r->cur_len; cur_ofs = r->cur_ofs;
        min_mask = r->min_mask; ins_mask = r->ins_mask; clean_ptr = r->clean_ptr; count = r->count; vr = r->vr;
        hits = r->hits; misses = r->misses; inserts = r->inserts; vevents = r->vevents; nops = r->nops; ntok = r->ntok;
        error = r->error;
    }
    __device__ __forceinline__ void store(Run *r) const {
        if (lane == 0) {
            r->p = p; r->last = last; r->cur_p = cur_p; r->cur_len = cur_len; r->cur_ofs = cur_ofs;
            r->min_mask = min_mask; r->ins_mask = ins_mask; r->clean_ptr = clean_ptr; r->count = count; r->vr = vr;
            r->hits = hits; r->misses = misses; r->inserts = inserts; r->vevents = vevents; r->nops = nops; r->ntok = ntok;
            r->error = error;
        }
    }
    __device__ __forceinline__ void write_slot(i64 slot, i64 off, i64 tag) {
        if (lane == 0) {
            __stcg(toff + slot, off);
            __stcg(ttag + slot, tag);
            nib_lower(gsum, slot >> GRP_SHIFT, tag ? tz1c(tag) : 0);
            if (nlog < G5_WLOG_CAP) wlog[nlog] = (int)(slot << 1) | (tag ? 0 : 1);
        }
        nlog++;
        __syncwarp();
    }

    // rzip.c:372-397.  Returns total length (0 if < 31) and the backward part in rev.
    __device__ __forceinline__ i64 match_len(i64 p0, i64 op, i64 &rev) {
        rev = 0;
        if (op >= p0) return 0;
        i64 maxf = end - p0, f = 0;
        while (f < maxf) {
            i64 k = f + lane * 8;
            u64 x = 0;
            if (k < maxf) x = ld_u64_unaligned(buf, p0 + k) ^ ld_u64_unaligned(buf, op + k);
            unsigned mm = __ballot_sync(FULL, x != 0);
            if (mm) {
                int l = __ffs(mm) - 1;
                int nb = x ? ((__ffsll((long long)x) - 1) >> 3) : 8;
                nb = __shfl_sync(FULL, nb, l);
                f += l * 8 + nb;
                break;
            }
            f += 256;
        }
        if (f > maxf) f = maxf;
        i64 lo = last > 0 ? last : 0;
        i64 maxb = p0 - lo;
        if (op < maxb) maxb = op;
        i64 r = 0;
        while (r < maxb) {
            i64 k = r + lane * 8;
            u64 x = 0;
            if (k < maxb) x = ld_u64_unaligned(buf, p0 - 8 - k) ^ ld_u64_unaligned(buf, op - 8 - k);
            unsigned mm = __ballot_sync(FULL, x != 0);
            if (mm) {
                int l = __ffs(mm) - 1;
                int nb = x ? (__clzll((long long)x) >> 3) : 8;
                nb = __shfl_sync(FULL, nb, l);
                r += l * 8 + nb;
                break;
            }
            r += 256;
        }
        if (r > maxb) r = maxb;
        rev = r;
        i64 tot = f + r;
        return tot < MRZ_MIN_MATCH ? 0 : tot;
    }

    // rzip.c:426-462
    __device__ __forceinline__ i64 lookup(i64 t, i64 pp, i64 &bofs, i64 &brev) {
        i64 best = 0;
        i64 h = t & hmask;
        for (;;) {
            i64 et = __ldcg(ttag + ((h + lane) & hmask));
            unsigned em = __ballot_sync(FULL, et == 0);
            int fe = em ? (__ffs(em) - 1) : 32;
            unsigned eq = __ballot_sync(FULL, et == t);
            if (fe < 32) eq &= (1u << fe) - 1u;
            while (eq) {
                int l = __ffs(eq) - 1;
                eq &= eq - 1;
                i64 off = __ldcg(toff + ((h + l) & hmask));
                i64 rev;
                i64 m = match_len(pp, off, rev);
                if (m) {
                    hits++;
                    if (m > best) { best = m; bofs = off - rev; brev = rev; }
                } else
                    misses++;
            }
            if (fe < 32) break;
            h = (h + 32) & hmask;
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
            bool displaced = false;
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
                    int want = (int)(vr - round);
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
                    if ((mbm >> s) & 1u) count--;      // rzip.c:267-269
                    else if ((lsm >> s) & 1u) {        // rzip.c:275-277
                        displaced = true;
                        otag = __shfl_sync(FULL, et, s);
                    }
                    break;
                }
                h = (h + 32) & hmask;
            }
            if (displaced) {
                if (sp >= 60) { error = 3; return; }
                stk[3 * sp] = target; stk[3 * sp + 1] = ct; stk[3 * sp + 2] = coff;
                sp++;
                coff = __ldcg(toff + target);
                ct = otag;
                continue;
            }
            write_slot(target, coff, ct);
            break;
        }
        while (sp > 0) {
            sp--;
            write_slot(stk[3 * sp], stk[3 * sp + 2], stk[3 * sp + 1]);
        }
    }

    // rzip.c:305-328
    __device__ __forceinline__ void clean_one() {
        if (clean_ptr < clean_hint) clean_ptr = clean_hint;
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

    // One position that passed the lookup gate (rzip.c:573-599).  True when a match was emitted.
    __device__ __forceinline__ bool op(i64 pp, i64 t, i64 *stk) {
        nops++;
        i64 bofs = 0, brev = 0;
        const bool maybe = filt_query(filt, t);
        i64 best = maybe ? lookup(t, pp, bofs, brev) : 0;
        if ((t & ins_mask) == ins_mask) {
            inserts++;
            count++;
            bool placed = false;
            if (!maybe) {   // no equal tag exists: the insert slot is the first empty / due-for-sweep / lesser entry
                const i64 btm = (min_mask << 1) | 1;
                const int kb = __popcll((u64)btm), myz = tz1(t);
                i64 stg;
                i64 sl = find_stop(ttag, gsum, t & hmask, myz > kb ? myz : kb, hsize, stg);
                __syncwarp();
                if (sl >= 0 && (stg == 0 || (stg & btm) != btm)) {
                    if (stg != 0) count--;             // rzip.c:267-269
                    write_slot(sl, pp, t);
                    placed = true;
                }
            }
            if (!placed) insert(t, pp, stk);
            if (lane == 0) {
                filt_set(filt, t);
                if (maybe) filt_set(filt, t ^ DUP_SALT);
            }
            __syncwarp();
            if (count > limit) clean_one();
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
    }
};

__device__ __forceinline__ i64 xor_scan_incl(i64 v, int lane) {
    for (int o = 1; o < 32; o <<= 1) {
        i64 u = __shfl_up_sync(FULL, v, o);
        if (lane >= o) v ^= u;
    }
    return v;
}

// First index in [lo, hi) whose cand_pos >= rel (hi if none); 32-ary search by one warp.
__device__ __forceinline__ i64 list_seek(const u32 *cand_pos, i64 lo, i64 hi, u32 rel, int lane) {
    while (hi - lo > 32) {
        i64 len = hi - lo;
        i64 idx = lo + (len * lane) / 32;
        u32 v = __ldcg(cand_pos + idx);
        unsigned ge = __ballot_sync(FULL, v >= rel);
        if (!ge) {
            lo = lo + (len * 31) / 32 + 1;
        } else {
            int l = __ffs(ge) - 1;
            if (l == 0) return lo;
            i64 nlo = lo + (len * (l - 1)) / 32 + 1;
            hi = lo + (len * l) / 32 + 1;
            lo = nlo;
        }
    }
    bool in = lo + lane < hi;
    u32 v = in ? __ldcg(cand_pos + lo + lane) : 0xffffffffu;
    unsigned ge = __ballot_sync(FULL, in && v >= rel);
    return ge ? lo + (__ffs(ge) - 1) : hi;
}

struct Smem {
    i64 hx[256];
    i64 stk[3 * 64];
    unsigned long long wsum[40];
    unsigned long long cta_base;
    int red;
};

#ifdef MRZ_CUSIM
#define G5_SMEM_PTR(T, name) T *name = reinterpret_cast<T *>(cusim::dyn_smem())
#else
#define G5_SMEM_PTR(T, name) extern __shared__ __align__(16) unsigned char name##_raw[]; T *name = reinterpret_cast<T *>(name##_raw)
#endif

// Grid barrier: every thread of every CTA calls it.  One arrive per CTA on a monotonic counter in L2.
__device__ __forceinline__ void grid_sync(GState *gs, unsigned &nbar) {
#ifdef MRZ_CUSIM
    __syncthreads();
    if (threadIdx.x == 0) cusim::sync_grid();
    __syncthreads();
    (void)gs; (void)nbar;
#else
    __syncthreads();
    nbar++;
    if (threadIdx.x == 0) {
        const unsigned target = nbar * gridDim.x;
        __threadfence();
        atomicAdd(&gs->bar, 1u);
        while ((int)(*(volatile unsigned *)&gs->bar - target) < 0) { }
        __threadfence();
    }
    __syncthreads();
#endif
}

// exclusive prefix sum of v over the CTA (thread order); total in `total`.  Two block barriers.
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
    return wsum[wid] + inc - v;
}

struct Plan {
    int flags, dcount;
    int w1_slot, w2_slot;   // displaced occupant's new home; where (tag, pos) goes
    i64 w1_off, w1_tag;
};
__device__ __forceinline__ void plan_init(Plan &r) {
    r.flags = 0; r.dcount = 0; r.w1_slot = -1; r.w2_slot = -1; r.w1_off = 0; r.w1_tag = 0;
}

// Plan the insert of `tag` against the table without modifying anything (oracle/spec_model.c plan_op).
// The caller has established through the tag filter that no equal tag exists.
__device__ __forceinline__ void plan_op(Plan &r, const i64 *ttag, const i64 *toff, u32 *gsum, const u64 *filt, i64 tag,
                                        i64 hmask, i64 hsize, i64 min_mask, int chain) {
    const i64 btm = (min_mask << 1) | 1;
    const int kb = __popcll((u64)btm);                 // entries with fewer trailing ones are due for sweeping
    const int myz = tz1(tag);
    r.flags |= F_INSERT;
    i64 st;
    i64 s = find_stop(ttag, gsum, tag & hmask, myz > kb ? myz : kb, hsize, st);
    if (s < 0) { r.flags |= F_COMPLEX; return; }
    r.w2_slot = (int)s;
    if (st == 0) r.dcount = 1;                         // empty slot
    else if ((st & btm) != btm) { r.dcount = 0; r.flags |= F_W2ELIG; }   // rzip.c:267-269
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
            else { r.flags |= F_COMPLEX; return; }     // a second displacement: serial
        }
        r.flags |= F_W1;
        r.w1_slot = (int)g; r.w1_off = __ldcg(toff + s); r.w1_tag = st;
        r.dcount = 1 + d2;
    }
}

// "Twin" (oracle/spec_model.c plan_follower): position p+1 has the same tag as p and both insert.
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

// Service CTA: refill the ring of eligible slots (those clean_one_from_hash will zero next, in order) by a
// scan of the table from e_scan; only meaningful once ins_mask == 2*min_mask+1.
__device__ __forceinline__ void refill_queue(GState *gs, Smem *S, const i64 *ttag, i64 hsize, i64 btm, int want) {
    const int tid = threadIdx.x, T = blockDim.x;
    for (;;) {
        __syncthreads();
        const unsigned e_cnt = *(volatile unsigned *)&gs->e_cnt;
        const int avail = (int)(e_cnt - *(volatile unsigned *)&gs->e_idx);
        const i64 scan = *(volatile i64 *)&gs->e_scan;
        if (avail >= want || scan >= hsize || avail + 8 * T > G5_EQ_CAP) break;
        const i64 s0 = scan + 8 * (i64)tid;   // 8 consecutive slots per thread (scan is 8-aligned or clipped below)
        unsigned m = 0;
#pragma unroll
        for (int q = 0; q < 8; q++) {
            i64 s = s0 + q;
            if (s < hsize) {
                i64 t = __ldcg(ttag + s);
                if (t != 0 && (t & btm) != btm) m |= 1u << q;
            }
        }
        unsigned long long total;
        unsigned long long ex = block_excl_scan((unsigned long long)__popc(m), S->wsum, total);
        unsigned base = e_cnt + (unsigned)ex;
        for (int q = 0; q < 8; q++)
            if ((m >> q) & 1u) gs->E[base++ & (G5_EQ_CAP - 1)] = (int)(s0 + q);
        __syncthreads();
        if (tid == 0) {
            gs->e_cnt = e_cnt + (unsigned)total;
            i64 ns = scan + 8 * (i64)T;
            gs->e_scan = ns < hsize ? ns : hsize;
            __threadfence();
        }
    }
}

// After a serial op (one thread): bring the queue back in step with what the op wrote.
__device__ __forceinline__ void sync_queue(GState *gs, int nlog) {
    for (int w = 0; w < nlog; w++) {
        int sl = gs->wlog[w] >> 1;
        if (gs->wlog[w] & 1) {                   // zeroed by clean_one: consume the queue up to it
            while (gs->e_idx != gs->e_cnt && gs->E[gs->e_idx & (G5_EQ_CAP - 1)] <= sl) gs->e_idx++;
        } else if (sl < gs->e_scan) {            // overwritten by an insert: if queued, cut the queue there
            unsigned a = gs->e_idx, b = gs->e_cnt;
            while (a != b) {
                unsigned m = a + (b - a) / 2;
                if (gs->E[m & (G5_EQ_CAP - 1)] < sl) a = m + 1; else b = m;
            }
            if (a != gs->e_cnt && gs->E[a & (G5_EQ_CAP - 1)] == sl) { gs->e_cnt = a; gs->e_scan = sl; }
        }
    }
}

#ifdef MRZ_CUSIM
#define CLAIM_EPOCH_MAX 300u          // the CPU-emulator tests exercise the wrap-around path
#else
#define CLAIM_EPOCH_MAX 0x7FFF0u
#endif
__device__ __forceinline__ unsigned claim_key(unsigned epoch, int j) { return ((0x7FFFFu - epoch) << 13) | (unsigned)j; }

// Runs the automaton over the current segment's candidate list (or to the end of the chunk).
// Grid: G CTAs of T threads, cooperative.  CTAs 0..G-2 plan one op per thread, CTA G-1 serves the sweep queue.
__global__ void __launch_bounds__(256, 1)
k_automaton_grid(const u8 *__restrict__ buf, i64 *ttag, i64 *toff, u64 *filt, u32 *gsum, unsigned *claim, GState *gs_,
                 const i64 *__restrict__ hidx, AutoState *st, const i64 *__restrict__ tile_off,
                 const u32 *__restrict__ cand_pos, const i64 *__restrict__ cand_tag, Token *toks, i64 *dbg) {
    if (st->done || st->error) return;
    G5_SMEM_PTR(Smem, S);
    volatile GState *gs = gs_;
    unsigned nbar = 0;
    i64 tm[12];
#pragma unroll
    for (int q = 0; q < 12; q++) tm[q] = 0;
    i64 t_prev = clock64();
#define TICK(slot) do { i64 t_now = clock64(); tm[slot] += t_now - t_prev; t_prev = t_now; } while (0)
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5, T = blockDim.x;
    const int G = gridDim.x, cta = blockIdx.x;
    const bool service = cta == G - 1;
    const int j = cta * T + tid;             // my op index within a batch (planner CTAs)
    const int B = (G - 1) * T;               // ops per batch
    for (int q = tid; q < 256; q += T) S->hx[q] = hidx[q];
    const i64 n = st->n, end = st->end, hmask = st->hmask, hsize = st->hsize, limit = st->limit;
    const int chain = st->chain;
    const i64 seg_lo = st->seg_lo, seg_hi = st->seg_hi, ncand = st->seg_ncand;
    if (j == 0) {
        volatile Run *r = &gs->run;
        r->p = st->p; r->last = st->last; r->cur_p = st->cur_p; r->cur_len = st->cur_len; r->cur_ofs = st->cur_ofs;
        r->min_mask = st->min_mask; r->ins_mask = st->ins_mask; r->clean_ptr = st->clean_ptr; r->count = st->count;
        r->vr = st->victim_round;
        r->hits = st->tag_hits; r->misses = st->tag_misses; r->inserts = st->inserts; r->vevents = st->victim_events;
        r->nops = st->n_ops; r->ntok = st->ntok; r->error = 0;
        gs->e_idx = 0; gs->e_cnt = 0; gs->e_scan = r->clean_ptr;
        gs->li = 0; gs->bmax = B; gs->kcut = ~0ULL;
    }
    Warp w;   // used by warp 0 of CTA 0 only
    w.buf = buf; w.ttag = ttag; w.toff = toff; w.filt = filt; w.gsum = gsum; w.toks = toks; w.lane = lane;
    w.n = n; w.end = end; w.hmask = hmask; w.hsize = hsize; w.limit = limit; w.tok_cap = st->tok_cap; w.chain = chain;
    w.wlog = const_cast<int *>(gs->wlog); w.nlog = 0; w.clean_hint = 0;
    unsigned round = 0;                      // cut-search round (identical in every thread)
    grid_sync(gs_, nbar);

    for (;;) {
        // ---------------------------------------------------------------- batch set-up (every thread reads the same state)
        if (gs->run.error || gs->li >= ncand) break;
        const i64 li0 = gs->li;
        const i64 min_mask = gs->run.min_mask, ins_mask = gs->run.ins_mask;
        const i64 btm = (min_mask << 1) | 1;
        const bool steady = ins_mask == btm;   // inserts can no longer create sweep-eligible entries
        const int bmax = gs->bmax;
        const i64 p_now = gs->run.p, last_now = gs->run.last, clean0 = gs->run.clean_ptr;
        unsigned epoch = gs->epoch;
        int nr = bmax < B ? bmax : B;
        if (ncand - li0 < nr) nr = (int)(ncand - li0);
        if (epoch >= CLAIM_EPOCH_MAX) {      // claim keys exhausted: wipe the claim array, restart the epochs
            for (i64 q = (i64)blockIdx.x * T + tid; q < hsize + 8; q += (i64)G * T) claim[q] = 0xffffffffu;
            epoch = 0;
        }
        const unsigned kbase = claim_key(epoch, 0);
        // ---------------------------------------------------------------- phase A: gather + plan + claim | queue refill
        Plan pl;
        plan_init(pl);
        i64 my_pos = 0, my_tag = 0;
        bool valid = false, follower = false, next_follows = false;
        if (service) {
            if (steady) {
                if (tid == 0 && gs->e_scan < clean0) gs->e_scan = clean0;
                refill_queue(gs_, S, ttag, hsize, btm, nr + 64);
            } else if (tid == 0) {
                gs->e_idx = 0; gs->e_cnt = 0; gs->e_scan = clean0;
            }
        } else {
            if (j < nr) {
                const u32 rp = __ldcg(cand_pos + li0 + j);
                my_tag = __ldcg(cand_tag + li0 + j);
                my_pos = seg_lo + (i64)rp;
                valid = my_pos > p_now && (my_tag & min_mask) == min_mask;
                // twin: same tag as the position just before (itself a live op of this batch, not the second of a pair)
                bool same = false, same_prev = false;
                if (j > 0 && my_pos - 1 > p_now) same = __ldcg(cand_pos + li0 + j - 1) + 1 == rp && __ldcg(cand_tag + li0 + j - 1) == my_tag;
                if (j > 1 && my_pos - 2 > p_now) same_prev = __ldcg(cand_pos + li0 + j - 2) + 2 == rp && __ldcg(cand_tag + li0 + j - 2) == my_tag;
                follower = valid && same && !same_prev;
                if (valid && j + 1 < nr && !same) {   // am I the leader of a pair?
                    next_follows = __ldcg(cand_pos + li0 + j + 1) == rp + 1 && __ldcg(cand_tag + li0 + j + 1) == my_tag;
                }
            }
            // own plan; a follower in lane 0 plans its leader's op itself (same tag => same plan), the others get it by shuffle
            if (valid && (!follower || lane == 0)) {
                if (filt_query(filt, my_tag)) pl.flags |= F_COMPLEX;          // an equal tag may exist
                else if ((my_tag & ins_mask) == ins_mask)
                    plan_op(pl, ttag, toff, gsum, filt, my_tag, hmask, hsize, min_mask, chain);
            }
            int a_flags = __shfl_up_sync(FULL, pl.flags, 1), a_w2 = __shfl_up_sync(FULL, pl.w2_slot, 1);
            if (follower) {
                if (lane == 0) { a_flags = pl.flags; a_w2 = pl.w2_slot; }
                plan_init(pl);
                plan_follower(pl, ttag, gsum, buf, my_pos, my_tag, a_flags, a_w2, hsize, min_mask, chain, end, last_now);
            }
            if (valid && (pl.flags & (F_COMPLEX | F_INSERT)) == F_INSERT) {
                atomicMin(claim + pl.w2_slot, kbase | (unsigned)j);
                if (pl.flags & F_W1) atomicMin(claim + pl.w1_slot, kbase | (unsigned)j);
            }
        }
        if (j == 0) gs->epoch = epoch + 1;
        TICK(0);
        grid_sync(gs_, nbar);
        TICK(1);
        // ---------------------------------------------------------------- phase B: conflicts + CTA prefix
        const i64 w_lo = clean0, w_hi = steady ? gs->e_scan : clean0;   // the sweep queue covers [w_lo, w_hi)
        unsigned long long v = 0, incl = 0;
        if (!service) {
            if (valid && (pl.flags & (F_COMPLEX | F_INSERT)) == F_INSERT) {
                // a follower's plan builds on its leader's write: only older claims conflict
                const unsigned before = kbase | (unsigned)((pl.flags & F_FOLLOW) ? j - 1 : j);
                bool cf = __ldcg(claim + pl.w2_slot) < before;
                if (!cf && (pl.flags & F_W1)) cf = __ldcg(claim + pl.w1_slot) < (kbase | (unsigned)j);
                // a stop slot that sits in the sweep queue: the order of sweeps and inserts matters
                if ((pl.flags & F_W2ELIG) && pl.w2_slot >= w_lo && pl.w2_slot < w_hi) cf = true;
                if ((pl.flags & F_W1ELIG) && pl.w1_slot >= w_lo && pl.w1_slot < w_hi) cf = true;
                if (cf) pl.flags |= F_CONF;
            }
            if (valid && !(pl.flags & (F_COMPLEX | F_CONF))) {
                v = 1ULL << 48;
                if (pl.flags & F_INSERT) v |= (unsigned long long)pl.dcount | (1ULL << 16) | ((pl.flags & F_FOLLOW) ? (1ULL << 32) : 0ULL);
            }
            unsigned long long total;
            incl = block_excl_scan(v, S->wsum, total) + v;
            if (tid == 0) gs->cta_tot[cta] = total;
        }
        grid_sync(gs_, nbar);
        TICK(2);
        // ---------------------------------------------------------------- phase C: grid prefix, first cut search
        if (!service) {
            if (wid == 0) {
                unsigned long long a = 0;
                for (int c = lane; c < cta; c += 32) a += gs->cta_tot[c];
                for (int o = 16; o; o >>= 1) a += __shfl_xor_sync(FULL, a, o);
                if (lane == 0) S->cta_base = a;
            }
            __syncthreads();
            incl += S->cta_base;
            if (j < nr) gs->pre[j] = incl;
        }
        int pos = 0;
        for (;;) {   // ---- rounds: cut search -> commit -> serial op -> invalidation
            round++;
            const unsigned long long rkey = (unsigned long long)(~round) << 32;
            const i64 cnt0 = gs->run.count;
            const unsigned e_idx0 = gs->e_idx;
            const int avail = (int)(gs->e_cnt - e_idx0);
            const unsigned long long base = pos ? gs->pre[pos - 1] : 0ULL;
            if (!service) {
                const i64 cur_len = gs->run.cur_len, cur_p = gs->run.cur_p;
                bool c = false;
                if (j >= pos && j < nr && valid) {
                    c = (pl.flags & (F_COMPLEX | F_CONF)) != 0;
                    c = c || (cur_len >= MRZ_MIN_MATCH && my_pos >= cur_p + MRZ_MIN_MATCH);   // an emission is due
                    if (!c && (pl.flags & F_INSERT) && pl.dcount == 1) {
                        i64 Tc = cnt0 + (i64)PK_D(incl - base);
                        if (Tc > limit && (!steady || Tc - limit > avail)) c = true;           // needs a serial sweep step
                    }
                }
                const unsigned bal = __ballot_sync(FULL, c);
                if (lane == 0 && bal) atomicMin(const_cast<unsigned long long *>(&gs->kcut), rkey | (unsigned)(j + __ffs(bal) - 1));
            }
            grid_sync(gs_, nbar);
            TICK(3);
            // ---- phase D: commit [pos, k)
            int k = nr;
            {
                const unsigned long long kc = gs->kcut;
                if ((kc >> 32) == (rkey >> 32)) k = (int)(kc & 0xffffffffu);
            }
            unsigned long long d = 0;
            int cleans = 0;
            i64 Tk = cnt0;
            if (k > pos) {
                d = gs->pre[k - 1] - base;
                Tk = cnt0 + (i64)PK_D(d);
                cleans = Tk > limit ? (int)(Tk - limit) : 0;
            }
            if (!service && j >= pos && j < k && valid && (pl.flags & F_INSERT)) {
                if (pl.flags & F_W1) {
                    __stcg(toff + pl.w1_slot, pl.w1_off);
                    __stcg(ttag + pl.w1_slot, pl.w1_tag);
                    nib_lower(gsum, pl.w1_slot >> GRP_SHIFT, tz1c(pl.w1_tag));
                }
                // a twin that overwrites its leader's entry in the same commit must win the slot
                const bool overwritten = next_follows && j + 1 < k && (my_tag & btm) != btm;
                if (!overwritten) {
                    __stcg(toff + pl.w2_slot, my_pos);
                    __stcg(ttag + pl.w2_slot, my_tag);
                    nib_lower(gsum, pl.w2_slot >> GRP_SHIFT, tz1c(my_tag));
                }
                filt_set(filt, my_tag);
                if (pl.flags & F_FOLLOW) filt_set(filt, my_tag ^ DUP_SALT);   // the tag is in the table twice now
            }
            if (!service && j < cleans) {
                const int sl = gs->E[(e_idx0 + (unsigned)j) & (G5_EQ_CAP - 1)];
                __stcg(ttag + sl, (i64)0);
                __stcg(toff + sl, (i64)0);
                nib_lower(gsum, sl >> GRP_SHIFT, 0);
            }
            if (j == 0 && k > pos) {
                volatile Run *r = &gs->run;
                r->count = Tk - cleans;
                r->inserts += PK_I(d);
                r->misses += PK_M(d);
                r->nops += PK_V(d);
                const i64 pk1 = seg_lo + (i64)__ldcg(cand_pos + li0 + k - 1);
                if (pk1 > r->p) r->p = pk1;
                if (cleans) {
                    gs->e_idx = e_idx0 + (unsigned)cleans;
                    r->clean_ptr = gs->E[(e_idx0 + (unsigned)cleans - 1) & (G5_EQ_CAP - 1)];   // pointer stays on the last zeroed slot
                }
            }
            if (j == 0) { gs->ended = 0; gs->jumped = 0; }
            grid_sync(gs_, nbar);
            TICK(4);
            if (k >= nr) break;
            tm[10] += 1;
            if (dbg && !service && j == k) {   // why is this op serial?
                int why = (pl.flags & F_COMPLEX) ? (follower ? 12 : ((pl.flags & F_INSERT) ? 13 : 11)) : (pl.flags & F_CONF) ? 14 : 15;
                atomicAdd((unsigned long long *)&dbg[why], 1ULL);
            }
            // ---- op k on the live table, reference semantics (warp 0 of CTA 0)
            if (cta == 0 && wid == 0) {
                w.load(const_cast<const Run *>(&gs_->run));
                w.nlog = 0;
                i64 hint = w.clean_ptr;
                if (steady) {
                    const unsigned ei = gs->e_idx, ec = gs->e_cnt;
                    const i64 es = gs->e_scan;
                    hint = (ec != ei) ? (i64)gs->E[ei & (G5_EQ_CAP - 1)] : (es > hint ? es : hint);
                }
                w.clean_hint = hint;
                const i64 pk = seg_lo + (i64)__ldcg(cand_pos + li0 + k), tk = __ldcg(cand_tag + li0 + k);
                w.p = pk;
                bool jumped = w.op(pk, tk, S->stk);
                w.clean_hint = 0;
                if (jumped && !w.error) {
                    // dense-match regime: examine the next positions directly from the bytes until a
                    // window of 32 positions passes without an emission, then re-synchronise with the list
                    for (;;) {
                        if (w.error || w.p >= w.end) break;
                        i64 b0 = w.p + 1;                       // positions b0 + lane
                        i64 v0 = S->hx[__ldg(buf + b0 + lane)];
                        i64 v1 = S->hx[__ldg(buf + b0 + 32 + lane)];
                        i64 P0 = xor_scan_incl(v0, lane);
                        i64 tot0 = __shfl_sync(FULL, P0, 31);
                        i64 P1 = tot0 ^ xor_scan_incl(v1, lane);
                        int hi_i = lane + 30;
                        i64 a0 = __shfl_sync(FULL, P0, hi_i & 31);
                        i64 a1 = __shfl_sync(FULL, P1, hi_i & 31);
                        i64 lo_v = __shfl_up_sync(FULL, P0, 1);
                        if (lane == 0) lo_v = 0;
                        i64 tq = (hi_i < 32 ? a0 : a1) ^ lo_v;
                        bool again = false;
                        int lim = (w.end - w.p) < 32 ? (int)(w.end - w.p) : 32;   // positions b0..min(b0+31,end)
                        for (int q = 0; q < lim; q++) {
                            i64 tj = __shfl_sync(FULL, tq, q);
                            if ((tj & w.min_mask) != w.min_mask) continue;
                            w.p = b0 + q;
                            if (w.op(b0 + q, tj, S->stk)) { again = true; break; }
                        }
                        if (again) continue;
                        w.p = b0 + lim - 1;
                        break;
                    }
                    i64 nli = ncand;
                    if (!w.error && w.p < w.end && w.p + 1 < seg_hi) {
                        i64 rel = w.p + 1 - seg_lo;
                        i64 tile = rel / SCAN_TILE;
                        nli = list_seek(cand_pos, tile_off[tile], tile_off[tile + 1], (u32)rel, lane);
                    }
                    if (lane == 0) gs->li = nli;
                }
                w.store(const_cast<Run *>(&gs_->run));
                if (lane == 0) {
                    bool mask_changed = w.min_mask != min_mask || w.ins_mask != ins_mask;
                    bool overflow = w.nlog > G5_WLOG_CAP;
                    gs->jumped = jumped;
                    gs->nlog = overflow ? 0 : w.nlog;
                    if (jumped || mask_changed || overflow || w.error) {
                        gs->ended = 1;
                        gs->e_idx = 0; gs->e_cnt = 0; gs->e_scan = w.clean_ptr;    // rebuild the queue
                        if (!jumped) gs->li = li0 + k + 1;
                    } else {
                        if (steady) sync_queue(gs_, w.nlog);
                    }
                    __threadfence();
                }
            }
            grid_sync(gs_, nbar);
            TICK(5);
            if (gs->ended) break;
            pos = k + 1;
            // later plans that claimed a slot this op wrote are no longer valid; nor is its twin's plan
            if (!service && j > k && j < nr && valid && !(pl.flags & (F_COMPLEX | F_CONF))) {
                if (j == k + 1 && (pl.flags & F_FOLLOW)) pl.flags |= F_CONF;
                if (pl.flags & F_INSERT) {
                    const int nl = gs->nlog;
                    for (int q = 0; q < nl; q++) {
                        const int sl = gs->wlog[q] >> 1;
                        if (sl == pl.w2_slot || ((pl.flags & F_W1) && sl == pl.w1_slot)) {
                            pl.flags |= F_CONF;
                            if (dbg) atomicAdd((unsigned long long *)&dbg[16], 1ULL);
                            break;
                        }
                    }
                }
            }
            TICK(6);
        }
        // ---------------------------------------------------------------- batch epilogue
        if (j == 0) {
            if (!gs->ended) gs->li = li0 + nr;               // whole batch consumed
            // batch size: small after an emission (the plans behind a jump are wasted), growing back
            if (gs->jumped) gs->bmax = 64;
            else if (gs->bmax < B) gs->bmax = gs->bmax * 2;
            tm[8] += nr; tm[9] += 1;
        }
        grid_sync(gs_, nbar);
        TICK(7);
    }
    if (j == 0 && dbg)
        for (int q = 0; q < 11; q++) dbg[q] += tm[q];
    // segment exhausted (or chunk finished)
    if (j == 0) {
        volatile Run *r = &gs->run;
        i64 p = r->p;
        if (!r->error && p < seg_hi - 1) p = seg_hi - 1;
        st->p = p; st->last = r->last; st->cur_p = r->cur_p; st->cur_len = r->cur_len; st->cur_ofs = r->cur_ofs;
        st->min_mask = r->min_mask; st->ins_mask = r->ins_mask; st->clean_ptr = r->clean_ptr; st->count = r->count;
        st->victim_round = r->vr;
        st->tag_hits = r->hits; st->tag_misses = r->misses; st->inserts = r->inserts; st->victim_events = r->vevents;
        st->n_ops = r->nops; st->ntok = r->ntok;
        if (r->error) st->error = r->error;
        if (p >= end) st->done = 1;
    }
#undef TICK
}

}  // namespace g5

extern "C" size_t mrz_grid_state_bytes() { return sizeof(g5::GState); }

// G CTAs x T threads, cooperative launch (all CTAs resident: they wait on one another at the grid barriers).
extern "C" int mrz_launch_automaton_grid(int G, int T, const u8 *buf, i64 *ttag, i64 *toff, u64 *filt, u32 *gsum, unsigned *claim,
                                         void *gstate, const i64 *hidx, AutoState *st, const i64 *tile_off, const u32 *cand_pos,
                                         const i64 *cand_tag, Token *toks, i64 *dbg, cudaStream_t s) {
    g5::GState *gs = (g5::GState *)gstate;
#ifndef MRZ_CUSIM
    cudaError_t e = cudaMemsetAsync(&gs->bar, 0, sizeof(unsigned), s);
    if (e != cudaSuccess) return (int)e;
    void *args[] = {(void *)&buf, (void *)&ttag, (void *)&toff, (void *)&filt, (void *)&gsum, (void *)&claim, (void *)&gs, (void *)&hidx,
                    (void *)&st, (void *)&tile_off, (void *)&cand_pos, (void *)&cand_tag, (void *)&toks, (void *)&dbg};
    e = cudaLaunchCooperativeKernel((const void *)g5::k_automaton_grid, dim3(G), dim3(T), args, sizeof(g5::Smem), s);
    return (int)e;
#else
    MRZ_LAUNCH_COOP(g5::k_automaton_grid, G, T, sizeof(g5::Smem), s, buf, ttag, toff, filt, gsum, claim, gs, hidx, st, tile_off, cand_pos,
                    cand_tag, toks, dbg);
    return 0;
#endif
}
