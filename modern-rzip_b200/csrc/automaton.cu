// automaton.cu — the table automaton of hash_search, exact, as one 1024-thread CTA per chunk.
//
// Reproduces, operation for operation, the table semantics of src/rzip.c:
//   find_best_match (rzip.c:426-462), insert_hash (rzip.c:256-301), clean_one_from_hash
//   (rzip.c:305-328), single_match_len (rzip.c:372-397) and the greedy/lazy emission rule of
//   hash_search (rzip.c:571-599).
//
// B200 design (not the reference's loop):
//   * the per-byte loop is gone: candidates (positions whose tag passes the lookup gate) come from
//     tag_scan.cu as a position-ordered list;
//   * BATCHED SPECULATION: the next <=1024 candidate ops are *planned* in parallel, one thread
//     each, against the live table read-only (probe walk, insert target, one level of
//     displacement).  Plans are then committed in order: a plan is valid as long as no earlier op
//     of the batch wrote a slot inside its read ranges (checked exactly through a shared-memory
//     hash of the batch's write slots), it met no equal tag (=> match_len needed), and it stays
//     clear of the sweep window of clean_one_from_hash.  Everything else - typically <1 % of the
//     ops - is executed by warp 0 on the live table with the reference semantics (32-slot probe,
//     ballots for walk termination / equal-tag chains / the victim index, 256-byte-per-step
//     match extension), after which later plans that read a slot it wrote are invalidated.
//   * TAG FILTER: a blocked Bloom filter over every tag inserted in this chunk (one 64-bit word,
//     three bits per tag, no false negatives).  A negative answer proves that no entry with an
//     equal tag exists, so find_best_match's walk to the end of the probe cluster - hundreds of
//     slots at high mask levels - is skipped altogether, and lookup-only ops touch no table slot.
//   * TWINS: positions p and p+1 with equal tags (buf[p] == buf[p+31]) are planned as a pair.
//   * the sweep of clean_one_from_hash is pre-resolved per batch: a parallel scan keeps a queue of
//     the next eligible slots, so the n-th overflow of a batch zeroes the n-th queued slot.
//   oracle/spec_model.c is the CPU model of exactly this scheme; tests/test_spec_model.py proves
//   it equivalent to the sequential restatement, tests/test_sim_parity.py runs this very file on
//   the CPU emulator, tests/test_gpu_parity.py on the B200.
#include "mrz_common.cuh"

#define FULL 0xffffffffu
#define AT 1024            // threads per CTA = ops per batch
#define EQ_CAP 4096        // eligible-slot ring (power of two)
#define WH_CAP 4096        // write-slot hash (power of two, >= 2 * AT)
#define MAXWALK 4096       // longer probe walks are left to the serial path
#define WLOG_CAP 64

enum { F_COMPLEX = 1, F_INSERT = 2, F_W1 = 4, F_CONF = 8, F_FOLLOW = 16 };

// Blocked Bloom filter over inserted tags: word index and 3-bit mask (oracle/spec_model.c filt_hash).
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

// Running state of the automaton, shared by the whole CTA (rzip_state, mrzip_private.h:388-416).
struct Run {
    i64 p, last, cur_p, cur_len, cur_ofs, min_mask, ins_mask, clean_ptr, count, vr;
    i64 hits, misses, inserts, vevents, nops, ntok;
    int error;
};

struct Warp {
    // constants
    const u8 *buf;
    Slot *tab;
    u64 *filt;
    Token *toks;
    i64 n, end, hmask, hsize, limit, tok_cap;
    int chain, lane;
    // running state (identical in all lanes)
    i64 p, last, cur_p, cur_len, cur_ofs, min_mask, ins_mask, clean_ptr, count, vr;
    i64 hits, misses, inserts, vevents, nops, ntok;
    int error;
    // write log of the current op: (slot << 1) | zeroed; nlog > WLOG_CAP means overflow
    int *wlog;
    int nlog;
    i64 clean_hint;   // no eligible slot exists in [clean_ptr, clean_hint)

    __device__ __forceinline__ void load(const Run &r) {
        p = r.p; last = r.last; cur_p = r.cur_p; cur_len = r.cur_len; cur_ofs = r.cur_ofs;
        min_mask = r.min_mask; ins_mask = r.ins_mask; clean_ptr = r.clean_ptr; count = r.count; vr = r.vr;
        hits = r.hits; misses = r.misses; inserts = r.inserts; vevents = r.vevents; nops = r.nops; ntok = r.ntok;
        error = r.error;
    }
    __device__ __forceinline__ void store(Run &r) const {
        if (lane == 0) {
            r.p = p; r.last = last; r.cur_p = cur_p; r.cur_len = cur_len; r.cur_ofs = cur_ofs;
            r.min_mask = min_mask; r.ins_mask = ins_mask; r.clean_ptr = clean_ptr; r.count = count; r.vr = vr;
            r.hits = hits; r.misses = misses; r.inserts = inserts; r.vevents = vevents; r.nops = nops; r.ntok = ntok;
            r.error = error;
        }
    }
    __device__ __forceinline__ void log_write(i64 slot, int zeroed) {
        if (nlog < WLOG_CAP && lane == 0) wlog[nlog] = (int)(slot << 1) | zeroed;
        nlog++;
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
            Slot e = ld_slot(&tab[(h + lane) & hmask]);
            bool empty = (e.off | e.tag) == 0;
            unsigned em = __ballot_sync(FULL, empty);
            int fe = em ? (__ffs(em) - 1) : 32;
            unsigned eq = __ballot_sync(FULL, !empty && e.tag == t);
            if (fe < 32) eq &= (1u << fe) - 1u;
            while (eq) {
                int l = __ffs(eq) - 1;
                eq &= eq - 1;
                i64 off = __shfl_sync(FULL, e.off, l);
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
            i64 otag = 0, ooff = 0;
            for (;;) {
                Slot e = ld_slot(&tab[(h + lane) & hmask]);
                bool empty = (e.off | e.tag) == 0;
                bool minbit = !empty && (e.tag & btm) != btm;
                bool lesser = !empty && !minbit && tz1(e.tag) < ctz;
                bool equal = !empty && !minbit && !lesser && e.tag == ct;
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
                    if ((mbm >> s) & 1u) count--;      // rzip.c:267-269: due for cleaning, overwrite
                    else if ((lsm >> s) & 1u) {        // rzip.c:275-277: re-insert the occupant first
                        displaced = true;
                        otag = __shfl_sync(FULL, e.tag, s);
                        ooff = __shfl_sync(FULL, e.off, s);
                    }
                    break;
                }
                h = (h + 32) & hmask;
            }
            if (displaced) {
                if (sp >= 60) { error = 3; return; }
                stk[3 * sp] = target; stk[3 * sp + 1] = ct; stk[3 * sp + 2] = coff;
                sp++;
                ct = otag; coff = ooff;
                continue;
            }
            if (lane == 0) st_slot(&tab[target], coff, ct);
            log_write(target, 0);
            __syncwarp();
            break;
        }
        while (sp > 0) {
            sp--;
            if (lane == 0) st_slot(&tab[stk[3 * sp]], stk[3 * sp + 2], stk[3 * sp + 1]);
            log_write(stk[3 * sp], 0);
            __syncwarp();
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
                    Slot e = ld_slot(&tab[idx]);
                    elig = (e.off | e.tag) != 0 && (e.tag & btm) != btm;
                }
                unsigned m = __ballot_sync(FULL, elig);
                if (m) {
                    int l = __ffs(m) - 1;
                    clean_ptr += l;
                    if (lane == l) st_slot(&tab[idx], 0, 0);
                    log_write(clean_ptr, 1);
                    __syncwarp();
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
    __device__ __forceinline__ bool op(i64 pp, i64 t, i64 *stk) {
        nops++;
        i64 bofs = 0, brev = 0;
        i64 best = filt_query(filt, t) ? lookup(t, pp, bofs, brev) : 0;   // negative: nothing to find
        if ((t & ins_mask) == ins_mask) {
            inserts++;
            count++;
            insert(t, pp, stk);
            if (lane == 0) filt_set(filt, t);
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

// ---- shared memory of the CTA ------------------------------------------------------------------
struct Smem {
    i64 hx[256];
    i64 stk[3 * 64];
    i64 b_tag[AT];          // batch: tag of op j
    Run run;
    i64 e_scan;             // slots in [run.clean_ptr, e_scan) have been examined for the queue
    i64 li;                 // next candidate-list index
    u32 b_rel[AT];          // batch: position of op j relative to seg_lo
    int b_li[AT];           // batch: candidate-list index of op j
    int pre[AT];            // inclusive prefix over planned ops: misses << 22 | inserts << 11 | dcount sum
    int pl_fl[AT], pl_w2[AT];   // plan flags / insert slot of op j, for its twin
    int E[EQ_CAP];          // ring of eligible slots, ascending
    int wh_key[WH_CAP];     // write-slot hash: slot (or -1)
    int wh_val[WH_CAP];     //                  smallest op index writing it
    int wlog[WLOG_CAP];
    int wsum[40];
    unsigned e_idx, e_cnt;  // ring counters (monotonic)
    int nr, pos, k, nlog, jumped, ended, bmax, red;
};

#ifdef MRZ_CUSIM
#define MRZ_SMEM_PTR(T, name) static T name##_storage; T *name = &name##_storage
#else
#define MRZ_SMEM_PTR(T, name) extern __shared__ __align__(16) unsigned char name##_raw[]; T *name = reinterpret_cast<T *>(name##_raw)
#endif

// exclusive prefix sum of v over the CTA (thread order); total in `total`.  Two barriers.
__device__ __forceinline__ int block_excl_scan(int v, int *wsum, int &total) {
    int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    int inc = v;
    for (int o = 1; o < 32; o <<= 1) {
        int u = __shfl_up_sync(FULL, inc, o);
        if (lane >= o) inc += u;
    }
    if (lane == 31) wsum[wid] = inc;
    __syncthreads();
    if (wid == 0) {
        int w = wsum[lane], wi = w;
        for (int o = 1; o < 32; o <<= 1) {
            int u = __shfl_up_sync(FULL, wi, o);
            if (lane >= o) wi += u;
        }
        wsum[lane] = wi - w;
        if (lane == 31) wsum[32] = wi;
    }
    __syncthreads();
    total = wsum[32];
    int r = wsum[wid] + inc - v;
    return r;
}

// minimum of v over the CTA.  Two barriers; `red` must not be in use.
__device__ __forceinline__ int block_min(int v, int *red) {
    for (int o = 16; o; o >>= 1) {
        int u = __shfl_xor_sync(FULL, v, o);
        v = u < v ? u : v;
    }
    if (threadIdx.x == 0) *red = 0x7fffffff;
    __syncthreads();
    if ((threadIdx.x & 31) == 0) atomicMin(red, v);
    __syncthreads();
    return *red;
}

__device__ __forceinline__ unsigned wh_hash(int slot) { return ((unsigned)slot * 2654435761u) >> 20; }  // 12 bits

__device__ __forceinline__ void wh_insert(Smem *S, int slot, int j) {
    unsigned h = wh_hash(slot) & (WH_CAP - 1);
    for (;;) {
        int k = atomicCAS(&S->wh_key[h], -1, slot);
        if (k == -1 || k == slot) { atomicMin(&S->wh_val[h], j); return; }
        h = (h + 1) & (WH_CAP - 1);
    }
}
// true when some op with index < j writes a slot in [lo, hi]
__device__ __forceinline__ bool wh_range_hit(const Smem *S, int lo, int hi, int j) {
    for (int s = lo; s <= hi; s++) {
        unsigned h = wh_hash(s) & (WH_CAP - 1);
        for (;;) {
            int k = S->wh_key[h];
            if (k == -1) break;
            if (k == s) { if (S->wh_val[h] < j) return true; break; }
            h = (h + 1) & (WH_CAP - 1);
        }
    }
    return false;
}

struct Plan {
    int flags, dcount;
    int rd_lo, rd_hi, rd2_lo, rd2_hi;   // inclusive slot ranges read; rd2_lo < 0: none
    int w1_slot, w2_slot;               // displaced occupant's new home; where (tag, pos) goes
    i64 w1_off, w1_tag;
};

__device__ __forceinline__ bool slot_empty(const Slot &e) { return (e.off | e.tag) == 0; }

// Phase A: plan one op against the table without modifying anything (oracle/spec_model.c plan_op).
// Only the insert walk is needed: the caller has established (tag filter) that no equal tag exists.
__device__ __forceinline__ void plan_op(Plan &r, const Slot *tab, i64 tag, i64 hmask, i64 hsize, i64 min_mask,
                                        int chain, i64 w_lo, i64 w_hi) {
    const i64 btm = (min_mask << 1) | 1;
    const int myz = tz1(tag);
    const i64 h0 = tag & hmask;
    i64 h = h0;
    int kind = -1;           // 0: empty slot, 1: overwrite a due-for-sweep entry, 2: displace a lesser entry
    i64 otag = 0, ooff = 0;
    r.flags |= F_INSERT;
    r.rd_lo = (int)h0;
    for (;;) {
        Slot e[4];
#pragma unroll
        for (int q = 0; q < 4; q++) e[q] = ld_slot(&tab[h + q]);   // the table has slack slots past hsize
#pragma unroll
        for (int q = 0; q < 4; q++) {
            if (kind >= 0) continue;
            if (h + q >= hsize) { r.flags |= F_COMPLEX; return; }
            if (slot_empty(e[q])) { kind = 0; h += q; }
            else if (e[q].tag == tag) { r.flags |= F_COMPLEX; return; }          // cannot happen with a sound filter
            else if ((e[q].tag & btm) != btm) { kind = 1; h += q; }
            else if (tz1(e[q].tag) < myz) { kind = 2; h += q; otag = e[q].tag; ooff = e[q].off; }
        }
        if (kind >= 0) break;
        h += 4;
        if (h - h0 > MAXWALK) { r.flags |= F_COMPLEX; return; }
    }
    r.w2_slot = (int)h; r.rd_hi = (int)h;
    if (kind == 0) r.dcount = 1;
    else if (kind == 1) r.dcount = 0;
    else {
        // one level of displacement (rzip.c:275-277): where does the occupant go?
        const int oz = tz1(otag);
        const i64 g0 = otag & hmask;
        i64 g = g0;
        int round = 0, d2 = 0;
        bool done = false;
        for (;;) {
            Slot e[4];
#pragma unroll
            for (int q = 0; q < 4; q++) e[q] = ld_slot(&tab[g + q]);
#pragma unroll
            for (int q = 0; q < 4; q++) {
                if (done) continue;
                if (g + q >= hsize) { r.flags |= F_COMPLEX; return; }
                if (slot_empty(e[q])) { g += q; done = true; continue; }
                if ((e[q].tag & btm) != btm) { d2 = -1; g += q; done = true; continue; }
                if (tz1(e[q].tag) < oz) { r.flags |= F_COMPLEX; return; }
                if (e[q].tag == otag && ++round == chain) { r.flags |= F_COMPLEX; return; }
            }
            if (done) break;
            g += 4;
            if (g - g0 > MAXWALK) { r.flags |= F_COMPLEX; return; }
        }
        r.flags |= F_W1;
        r.rd2_lo = (int)g0; r.rd2_hi = (int)g;
        r.w1_slot = (int)g; r.w1_off = ooff; r.w1_tag = otag;
        r.dcount = 1 + d2;
    }
    // anything touching the sweep window is left to the serial path
    if (r.rd_hi >= w_lo && r.rd_lo < w_hi) r.flags |= F_COMPLEX;
    if (r.rd2_lo >= 0 && r.rd2_hi >= w_lo && r.rd2_lo < w_hi) r.flags |= F_COMPLEX;
}

// "Twin" (oracle/spec_model.c plan_follower): position p+1 has the same tag as p and both insert.  B's
// lookup meets exactly the entry A has just inserted and match_len(p+1, p) fails (tag_misses++) unless
// the bytes form a genuine run; B's insert walk passes A's entry and goes on from the next slot - or,
// before the first overflow, overwrites A's entry when the tag itself is due for sweeping.
__device__ __forceinline__ void plan_follower(Plan &r, const Slot *tab, const u8 *buf, i64 pos, i64 tag, int a_flags,
                                              int a_w2, i64 hmask, i64 hsize, i64 min_mask, int chain, i64 end, i64 last,
                                              i64 w_lo, i64 w_hi) {
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
    const int myz = tz1(tag);
    r.flags |= F_INSERT | F_FOLLOW;
    r.rd_lo = (int)(tag & hmask);
    if ((tag & btm) != btm) {
        r.w2_slot = a_w2; r.rd_hi = a_w2; r.dcount = 0;
    } else {
        i64 h = (i64)a_w2 + 1;
        const i64 hs = h;
        int kind = -1;
        for (;;) {
            Slot e[4];
#pragma unroll
            for (int q = 0; q < 4; q++) e[q] = ld_slot(&tab[h + q]);
#pragma unroll
            for (int q = 0; q < 4; q++) {
                if (kind >= 0) continue;
                if (h + q >= hsize) { r.flags |= F_COMPLEX; return; }
                if (slot_empty(e[q])) { kind = 0; h += q; }
                else if (e[q].tag == tag) { r.flags |= F_COMPLEX; return; }
                else if ((e[q].tag & btm) != btm) { kind = 1; h += q; }
                else if (tz1(e[q].tag) < myz) { r.flags |= F_COMPLEX; return; }
            }
            if (kind >= 0) break;
            h += 4;
            if (h - hs > MAXWALK) { r.flags |= F_COMPLEX; return; }
        }
        r.w2_slot = (int)h; r.rd_hi = (int)h; r.dcount = kind == 0 ? 1 : 0;
    }
    if (r.rd_hi >= w_lo && r.rd_lo < w_hi) r.flags |= F_COMPLEX;
}

// Refill the ring of eligible slots (those clean_one_from_hash will zero next, in order) by a
// parallel scan of the table from e_scan; only meaningful once ins_mask == 2*min_mask+1.
__device__ __forceinline__ void refill_queue(Smem *S, const Slot *tab, i64 hsize, int want) {
    const int tid = threadIdx.x;
    const i64 btm = (S->run.min_mask << 1) | 1;
    for (;;) {
        __syncthreads();
        int avail = (int)(S->e_cnt - S->e_idx);
        i64 scan = S->e_scan;
        if (avail >= want || scan >= hsize || avail + 2 * AT > EQ_CAP) break;
        // each thread examines two consecutive slots
        i64 s0 = scan + 2 * tid;
        bool g0 = false, g1 = false;
        if (s0 < hsize) { Slot e = ld_slot(&tab[s0]); g0 = !slot_empty(e) && (e.tag & btm) != btm; }
        if (s0 + 1 < hsize) { Slot e = ld_slot(&tab[s0 + 1]); g1 = !slot_empty(e) && (e.tag & btm) != btm; }
        int total;
        int ex = block_excl_scan((int)g0 + (int)g1, S->wsum, total);
        unsigned base = S->e_cnt + (unsigned)ex;
        if (g0) S->E[base++ & (EQ_CAP - 1)] = (int)s0;
        if (g1) S->E[base & (EQ_CAP - 1)] = (int)(s0 + 1);
        __syncthreads();
        if (tid == 0) {
            S->e_cnt += (unsigned)total;
            i64 ns = scan + 2 * AT;
            S->e_scan = ns < hsize ? ns : hsize;
        }
    }
}

// After a serial op (thread 0): bring the queue back in step with what the op wrote.
__device__ __forceinline__ void sync_queue(Smem *S, int nlog) {
    for (int w = 0; w < nlog; w++) {
        int sl = S->wlog[w] >> 1;
        if (S->wlog[w] & 1) {                    // zeroed by clean_one: consume the queue up to it
            while (S->e_idx != S->e_cnt && S->E[S->e_idx & (EQ_CAP - 1)] <= sl) S->e_idx++;
        } else if (sl < S->e_scan) {             // overwritten by an insert: if queued, cut the queue there
            unsigned a = S->e_idx, b = S->e_cnt;
            while (a != b) {
                unsigned m = a + (b - a) / 2;
                if (S->E[m & (EQ_CAP - 1)] < sl) a = m + 1; else b = m;
            }
            if (a != S->e_cnt && S->E[a & (EQ_CAP - 1)] == sl) { S->e_cnt = a; S->e_scan = sl; }
        }
    }
}

// Runs the automaton over the current segment's candidate list (or to the end of the chunk).
__global__ void __launch_bounds__(AT, 1)
k_automaton(const u8 *__restrict__ buf, Slot *tab, u64 *filt, const i64 *__restrict__ hidx, AutoState *st,
            const i64 *__restrict__ tile_off, const u32 *__restrict__ cand_pos,
            const i64 *__restrict__ cand_tag, Token *toks, i64 *dbg) {
    if (st->done || st->error) return;
    MRZ_SMEM_PTR(Smem, S);
    i64 tm[12];
#pragma unroll
    for (int q = 0; q < 12; q++) tm[q] = 0;
    i64 t_prev = clock64();
#define TICK(slot) do { i64 t_now = clock64(); tm[slot] += t_now - t_prev; t_prev = t_now; } while (0)
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    if (tid < 256) S->hx[tid] = hidx[tid];
    const i64 n = st->n, end = st->end, hmask = st->hmask, hsize = st->hsize, limit = st->limit;
    const int chain = st->chain;
    const i64 seg_lo = st->seg_lo, seg_hi = st->seg_hi, ncand = st->seg_ncand;
    if (tid == 0) {
        Run &r = S->run;
        r.p = st->p; r.last = st->last; r.cur_p = st->cur_p; r.cur_len = st->cur_len; r.cur_ofs = st->cur_ofs;
        r.min_mask = st->min_mask; r.ins_mask = st->ins_mask; r.clean_ptr = st->clean_ptr; r.count = st->count;
        r.vr = st->victim_round;
        r.hits = st->tag_hits; r.misses = st->tag_misses; r.inserts = st->inserts; r.vevents = st->victim_events;
        r.nops = st->n_ops; r.ntok = st->ntok; r.error = 0;
        S->e_idx = S->e_cnt = 0; S->e_scan = r.clean_ptr;
        S->li = 0; S->bmax = AT;
    }
    Warp w;   // used by warp 0 only
    w.buf = buf; w.tab = tab; w.filt = filt; w.toks = toks; w.lane = lane;
    w.n = n; w.end = end; w.hmask = hmask; w.hsize = hsize; w.limit = limit; w.tok_cap = st->tok_cap; w.chain = chain;
    w.wlog = S->wlog; w.nlog = 0; w.clean_hint = 0;
    __syncthreads();

    for (;;) {
        // ---------------------------------------------------------------- batch set-up
        if (S->run.error || S->li >= ncand) break;
        const i64 min_mask = S->run.min_mask, ins_mask = S->run.ins_mask;
        const i64 btm = (min_mask << 1) | 1;
        const bool steady = ins_mask == btm;   // inserts can no longer create sweep-eligible entries
        const int bmax = S->bmax;
        if (steady) {
            if (tid == 0 && S->e_scan < S->run.clean_ptr) S->e_scan = S->run.clean_ptr;
            refill_queue(S, tab, hsize, bmax);
        } else if (tid == 0) {
            S->e_idx = S->e_cnt = 0; S->e_scan = S->run.clean_ptr;
        }
        __syncthreads();
        TICK(0);   // queue refill
        const i64 w_lo = S->run.clean_ptr, w_hi = steady ? S->e_scan : S->run.clean_ptr;
        // ---------------------------------------------------------------- gather the next candidates
        int nr = 0;
        i64 li = S->li;
        const i64 p_now = S->run.p, last_now = S->run.last;
        for (;;) {
            int room = bmax - nr;
            if (room <= 0 || li >= ncand) break;
            int m = (ncand - li) < room ? (int)(ncand - li) : room;
            bool valid = false;
            u32 rp = 0; i64 tg = 0;
            if (tid < m) {
                rp = __ldcg(cand_pos + li + tid); tg = __ldcg(cand_tag + li + tid);
                valid = seg_lo + (i64)rp > p_now && (tg & min_mask) == min_mask;
            }
            int total;
            int ex = block_excl_scan((int)valid, S->wsum, total);
            if (valid) { S->b_rel[nr + ex] = rp; S->b_tag[nr + ex] = tg; S->b_li[nr + ex] = (int)(li + tid); }
            nr += total;
            li += m;
            if (nr * 2 >= bmax) break;
        }
        __syncthreads();
        TICK(1);   // gather
        if (nr == 0) { if (tid == 0) S->li = li; __syncthreads(); continue; }
        tm[8] += nr; tm[9] += 1;
        // ---------------------------------------------------------------- phase A: plan
        Plan pl;
        pl.flags = 0; pl.dcount = 0; pl.rd_lo = 1; pl.rd_hi = 0; pl.rd2_lo = -1; pl.rd2_hi = -1;
        pl.w1_slot = -1; pl.w2_slot = -1; pl.w1_off = 0; pl.w1_tag = 0;
        i64 my_pos = 0, my_tag = 0;
        bool follower = false;
        for (int q = tid; q < WH_CAP; q += AT) { S->wh_key[q] = -1; S->wh_val[q] = 0x7fffffff; }
        if (tid < nr) {
            const u32 rel = S->b_rel[tid];
            my_pos = seg_lo + (i64)rel; my_tag = S->b_tag[tid];
            // twin: same tag as the position just before, which is not itself the second of a pair
            bool same = tid > 0 && S->b_rel[tid - 1] + 1 == rel && S->b_tag[tid - 1] == my_tag;
            bool same_prev = tid > 1 && S->b_rel[tid - 2] + 2 == rel && S->b_tag[tid - 2] == my_tag;
            follower = same && !same_prev;
            if (!follower) {
                if (filt_query(filt, my_tag)) pl.flags |= F_COMPLEX;          // an equal tag may exist
                else if ((my_tag & ins_mask) == ins_mask)
                    plan_op(pl, tab, my_tag, hmask, hsize, min_mask, chain, w_lo, w_hi);
            }
            S->pl_fl[tid] = pl.flags; S->pl_w2[tid] = pl.w2_slot;
        }
        __syncthreads();
        if (follower) {
            plan_follower(pl, tab, buf, my_pos, my_tag, S->pl_fl[tid - 1], S->pl_w2[tid - 1], hmask, hsize, min_mask, chain,
                          end, last_now, w_lo, w_hi);
            S->pl_fl[tid] = pl.flags; S->pl_w2[tid] = pl.w2_slot;    // its leader looks at this when committing
        }
        TICK(2);   // plan
        const bool planned_ins = tid < nr && (pl.flags & (F_COMPLEX | F_INSERT)) == F_INSERT;
        if (planned_ins) {
            wh_insert(S, pl.w2_slot, tid);
            if (pl.flags & F_W1) wh_insert(S, pl.w1_slot, tid);
        }
        __syncthreads();
        if (planned_ins) {   // a follower's plan builds on its leader's write: only older writes conflict
            const int before = (pl.flags & F_FOLLOW) ? tid - 1 : tid;
            if (wh_range_hit(S, pl.rd_lo, pl.rd_hi, before) ||
                (pl.rd2_lo >= 0 && wh_range_hit(S, pl.rd2_lo, pl.rd2_hi, tid)))
                pl.flags |= F_CONF;
        }
        {   // inclusive prefix over ops that still hold a valid plan: misses << 22 | inserts << 11 | dcount sum
            int v = 0;
            if (tid < nr && (pl.flags & (F_COMPLEX | F_CONF | F_INSERT)) == F_INSERT)
                v = pl.dcount | (1 << 11) | ((pl.flags & F_FOLLOW) ? (1 << 22) : 0);
            int total;
            int ex = block_excl_scan(v, S->wsum, total);
            S->pre[tid] = ex + v;
        }
        if (tid == 0) { S->pos = 0; S->ended = 0; S->jumped = 0; }
        __syncthreads();
        TICK(3);   // conflicts + prefix
        // ---------------------------------------------------------------- phase B: in-order commit
        for (;;) {
            const int pos = S->pos;
            const int base = pos ? S->pre[pos - 1] : 0;
            const i64 cnt0 = S->run.count;
            const int avail = (int)(S->e_cnt - S->e_idx);
            const i64 cur_len = S->run.cur_len, cur_p = S->run.cur_p;
            int cut = nr;
            if (tid >= pos && tid < nr) {
                bool c = (pl.flags & (F_COMPLEX | F_CONF)) != 0;
                c = c || (cur_len >= MRZ_MIN_MATCH && my_pos >= cur_p + MRZ_MIN_MATCH);   // an emission is due
                if (!c && (pl.flags & F_INSERT) && pl.dcount == 1) {
                    i64 T = cnt0 + (((unsigned)S->pre[tid] - (unsigned)base) & 0x7ffu);
                    if (T > limit && (!steady || T - limit > avail)) c = true;             // needs a serial sweep step
                }
                if (c) cut = tid;
            }
            const int k = block_min(cut, &S->red);
            // commit [pos, k)
            if (tid >= pos && tid < k && (pl.flags & F_INSERT)) {
                if (pl.flags & F_W1) st_slot(&tab[pl.w1_slot], pl.w1_off, pl.w1_tag);
                // a twin that overwrites its leader's entry in the same commit must win the slot
                const bool overwritten = tid + 1 < k && (S->pl_fl[tid + 1] & F_FOLLOW) && S->pl_w2[tid + 1] == pl.w2_slot;
                if (!overwritten) st_slot(&tab[pl.w2_slot], my_pos, my_tag);
                filt_set(filt, my_tag);
            }
            int cleans = 0;
            if (k > pos) {
                const unsigned d = (unsigned)S->pre[k - 1] - (unsigned)base;
                i64 T = cnt0 + (d & 0x7ffu);
                cleans = T > limit ? (int)(T - limit) : 0;
                if (tid < cleans) st_slot(&tab[S->E[(S->e_idx + tid) & (EQ_CAP - 1)]], 0, 0);
                __syncthreads();   // everyone has read run.count / e_idx
                if (tid == 0) {
                    Run &r = S->run;
                    r.count = T - cleans;
                    r.inserts += (d >> 11) & 0x7ffu;
                    r.misses += d >> 22;
                    r.nops += k - pos;
                    r.p = seg_lo + (i64)S->b_rel[k - 1];
                    if (cleans) {
                        S->e_idx += (unsigned)cleans;
                        r.clean_ptr = S->E[(S->e_idx - 1) & (EQ_CAP - 1)];   // pointer stays on the last zeroed slot
                    }
                }
            }
            if (k >= nr) { __syncthreads(); TICK(4); break; }
            __syncthreads();
            TICK(4);   // cut search + commit
            tm[10] += 1;
            // ---- op k on the live table, reference semantics (warp 0)
            if (wid == 0) {
                w.load(S->run);
                w.nlog = 0;
                i64 hint = w.clean_ptr;
                if (steady) hint = (S->e_cnt != S->e_idx) ? (i64)S->E[S->e_idx & (EQ_CAP - 1)] : (S->e_scan > hint ? S->e_scan : hint);
                w.clean_hint = hint;
                const i64 pk = seg_lo + (i64)S->b_rel[k], tk = S->b_tag[k];
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
                        for (int j = 0; j < lim; j++) {
                            i64 tj = __shfl_sync(FULL, tq, j);
                            if ((tj & w.min_mask) != w.min_mask) continue;
                            w.p = b0 + j;
                            if (w.op(b0 + j, tj, S->stk)) { again = true; break; }
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
                    if (lane == 0) S->li = nli;
                }
                w.store(S->run);
                if (lane == 0) {
                    bool mask_changed = w.min_mask != min_mask || w.ins_mask != ins_mask;
                    bool overflow = w.nlog > WLOG_CAP;
                    S->jumped = jumped;
                    S->nlog = overflow ? 0 : w.nlog;
                    if (jumped || mask_changed || overflow || w.error) {
                        S->ended = 1;
                        S->e_idx = S->e_cnt = 0; S->e_scan = w.clean_ptr;    // rebuild the queue
                        if (!jumped) S->li = (i64)S->b_li[k] + 1;
                    } else {
                        if (steady) sync_queue(S, w.nlog);
                        S->pos = k + 1;
                    }
                }
            }
            __syncthreads();
            TICK(5);   // serial op
            if (S->ended) break;
            // later plans that read a slot this op wrote are no longer valid; nor is its twin's plan
            if (tid == k + 1 && (pl.flags & F_FOLLOW)) pl.flags |= F_CONF;
            if (tid > k && tid < nr && !(pl.flags & (F_COMPLEX | F_CONF))) {
                const int nl = S->nlog;
                for (int q = 0; q < nl; q++) {
                    int sl = S->wlog[q] >> 1;
                    if ((sl >= pl.rd_lo && sl <= pl.rd_hi) || (pl.rd2_lo >= 0 && sl >= pl.rd2_lo && sl <= pl.rd2_hi)) {
                        pl.flags |= F_CONF;
                        break;
                    }
                }
            }
            __syncthreads();
            TICK(6);   // invalidation
        }
        // ---------------------------------------------------------------- batch epilogue
        if (tid == 0) {
            if (!S->ended) S->li = li;                       // whole batch consumed
            // batch size: small after an emission (the plans behind a jump are wasted), growing back
            if (S->jumped) S->bmax = 64;
            else if (S->bmax < AT) S->bmax *= 2;
        }
        __syncthreads();
        TICK(7);
    }
    if (tid == 0 && dbg)
        for (int q = 0; q < 12; q++) dbg[q] += tm[q];
    // segment exhausted (or chunk finished)
    if (tid == 0) {
        const Run &r = S->run;
        i64 p = r.p;
        if (!r.error && p < seg_hi - 1) p = seg_hi - 1;
        st->p = p; st->last = r.last; st->cur_p = r.cur_p; st->cur_len = r.cur_len; st->cur_ofs = r.cur_ofs;
        st->min_mask = r.min_mask; st->ins_mask = r.ins_mask; st->clean_ptr = r.clean_ptr; st->count = r.count;
        st->victim_round = r.vr;
        st->tag_hits = r.hits; st->tag_misses = r.misses; st->inserts = r.inserts; st->victim_events = r.vevents;
        st->n_ops = r.nops; st->ntok = r.ntok;
        if (r.error) st->error = r.error;
        if (p >= end) st->done = 1;
    }
}

extern "C" void mrz_launch_automaton(const u8 *buf, Slot *tab, u64 *filt, const i64 *hidx, AutoState *st, const i64 *tile_off,
                                     const u32 *cand_pos, const i64 *cand_tag, Token *toks, i64 *dbg, cudaStream_t s) {
#ifndef MRZ_CUSIM
    cudaFuncSetAttribute(k_automaton, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(Smem));
    k_automaton<<<1, AT, sizeof(Smem), s>>>(buf, tab, filt, hidx, st, tile_off, cand_pos, cand_tag, toks, dbg);
#else
    MRZ_LAUNCH(k_automaton, 1, AT, s, buf, tab, filt, hidx, st, tile_off, cand_pos, cand_tag, toks, dbg);
#endif
}
