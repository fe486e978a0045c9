// automaton.cu — the exact sequential core of hash_search, one warp per chunk automaton.
//
// Reproduces, operation for operation, the table semantics of src/rzip.c:
//   find_best_match (rzip.c:426-462), insert_hash (rzip.c:256-301), clean_one_from_hash
//   (rzip.c:305-328), single_match_len (rzip.c:372-397) and the greedy/lazy emission rule of
//   hash_search (rzip.c:571-599).
// B200 design: the per-byte loop is gone (candidates come from tag_scan.cu); every probe looks at
// 32 consecutive 16-byte slots at once (one coalesced 512 B read from the L2-resident table) and
// resolves walk termination, equal-tag chains and the victim index with ballots; match extension
// compares 256 bytes per warp step; recursion in insert_hash is an explicit stack.
#include "mrz_common.cuh"

#define FULL 0xffffffffu

struct Warp {
    // constants
    const u8 *buf;
    Slot *tab;
    Token *toks;
    i64 n, end, hmask, hsize, limit, tok_cap;
    int chain, lane;
    // running state (identical in all lanes)
    i64 p, last, cur_p, cur_len, cur_ofs, min_mask, ins_mask, clean_ptr, count, vr;
    i64 hits, misses, inserts, vevents, nops, ntok;
    int error;

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
            __syncwarp();
            break;
        }
        while (sp > 0) {
            sp--;
            if (lane == 0) st_slot(&tab[stk[3 * sp]], stk[3 * sp + 2], stk[3 * sp + 1]);
            __syncwarp();
        }
    }

    // rzip.c:305-328
    __device__ __forceinline__ void clean_one() {
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
        i64 best = lookup(t, pp, bofs, brev);
        if ((t & ins_mask) == ins_mask) {
            inserts++;
            count++;
            insert(t, pp, stk);
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

// First index in [lo, hi) whose cand_pos >= rel (hi if none); 32-ary search.
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

// Runs the automaton over the current segment's candidate list (or to the end of the chunk).
__global__ void __launch_bounds__(32, 1)
k_automaton(const u8 *__restrict__ buf, Slot *tab, const i64 *__restrict__ hidx, AutoState *st,
            const i64 *__restrict__ tile_off, const u32 *__restrict__ cand_pos,
            const i64 *__restrict__ cand_tag, Token *toks) {
    if (st->done || st->error) return;
    __shared__ i64 hx[256];
    __shared__ i64 stk[3 * 64];
    const int lane = threadIdx.x;
    for (int i = lane; i < 256; i += 32) hx[i] = hidx[i];
    __syncwarp();

    Warp w;
    w.buf = buf; w.tab = tab; w.toks = toks; w.lane = lane;
    w.n = st->n; w.end = st->end; w.hmask = st->hmask; w.hsize = st->hsize; w.limit = st->limit;
    w.tok_cap = st->tok_cap; w.chain = st->chain;
    w.p = st->p; w.last = st->last; w.cur_p = st->cur_p; w.cur_len = st->cur_len; w.cur_ofs = st->cur_ofs;
    w.min_mask = st->min_mask; w.ins_mask = st->ins_mask; w.clean_ptr = st->clean_ptr; w.count = st->count;
    w.vr = st->victim_round;
    w.hits = st->tag_hits; w.misses = st->tag_misses; w.inserts = st->inserts; w.vevents = st->victim_events;
    w.nops = st->n_ops; w.ntok = st->ntok; w.error = 0;

    const i64 seg_lo = st->seg_lo, seg_hi = st->seg_hi, ncand = st->seg_ncand;
    i64 i = 0;
    while (i < ncand && !w.error) {
        i64 idx = i + lane;
        u32 rp = 0; i64 tg = 0;
        if (idx < ncand) { rp = __ldcg(cand_pos + idx); tg = __ldcg(cand_tag + idx); }
        int cnt = (ncand - i) < 32 ? (int)(ncand - i) : 32;
        bool jumped = false;
        for (int j = 0; j < cnt; j++) {
            i64 pj = seg_lo + (i64)__shfl_sync(FULL, rp, j);
            i64 tj = __shfl_sync(FULL, tg, j);
            if (pj <= w.p) continue;
            if ((tj & w.min_mask) != w.min_mask) continue;
            w.p = pj;
            if (w.op(pj, tj, stk)) { jumped = true; break; }
        }
        if (!jumped) { i += 32; continue; }
        // ---- after a match: p jumped to last_match.  Examine the next positions directly from the
        // bytes (dense-match regime), then re-synchronise with the list.
        for (;;) {
            if (w.error || w.p >= w.end) break;
            i64 base = w.p + 1;                       // positions base + lane
            i64 v0 = hx[__ldg(buf + base + lane)];
            i64 v1 = hx[__ldg(buf + base + 32 + lane)];
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
            int lim = (w.end - w.p) < 32 ? (int)(w.end - w.p) : 32;   // positions base..min(base+31,end)
            for (int j = 0; j < lim; j++) {
                i64 tj = __shfl_sync(FULL, tq, j);
                if ((tj & w.min_mask) != w.min_mask) continue;
                w.p = base + j;
                if (w.op(base + j, tj, stk)) { again = true; break; }
            }
            if (again) continue;
            w.p = base + lim - 1;
            break;
        }
        if (w.error || w.p >= w.end || w.p + 1 >= seg_hi) break;
        {
            i64 rel = w.p + 1 - seg_lo;
            i64 tile = rel / SCAN_TILE;
            i = list_seek(cand_pos, tile_off[tile], tile_off[tile + 1], (u32)rel, lane);
        }
    }
    // segment exhausted (or chunk finished)
    if (!w.error && w.p < seg_hi - 1) w.p = seg_hi - 1;
    if (lane == 0) {
        st->p = w.p; st->last = w.last; st->cur_p = w.cur_p; st->cur_len = w.cur_len; st->cur_ofs = w.cur_ofs;
        st->min_mask = w.min_mask; st->ins_mask = w.ins_mask; st->clean_ptr = w.clean_ptr; st->count = w.count;
        st->victim_round = w.vr;
        st->tag_hits = w.hits; st->tag_misses = w.misses; st->inserts = w.inserts; st->victim_events = w.vevents;
        st->n_ops = w.nops; st->ntok = w.ntok;
        if (w.error) st->error = w.error;
        if (w.p >= w.end) st->done = 1;
    }
}

extern "C" void mrz_launch_automaton(const u8 *buf, Slot *tab, const i64 *hidx, AutoState *st, const i64 *tile_off,
                                     const u32 *cand_pos, const i64 *cand_tag, Token *toks, cudaStream_t s) {
    MRZ_LAUNCH(k_automaton, 1, 32, s, buf, tab, hidx, st, tile_off, cand_pos, cand_tag, toks);
}
