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
//   * the table is stored as two arrays (tags[], offs[]) instead of the reference's 16-byte
//     entries: walks only ever look at tags, so a probe step moves half the bytes;
//   * BATCHED SPECULATION: the next <=1024 candidate ops are *planned* in parallel, one thread
//     each, against the live table read-only (insert walk, one level of displacement), then
//     committed in order; the first op that cannot commit from its plan cuts the batch, runs
//     exactly (below), and the batch carries on behind it - also behind an emitted match;
//   * EXACT CONFLICT RULE: an insert only ever RAISES the trailing-one count of the slot it writes
//     (empty / due-for-sweep -> >= need, lesser -> greater), so a walk that passed a slot still
//     passes it: only a write to a walk's own STOP slot can change its outcome, and an entry due for
//     sweeping stops every walk, so a slot of the sweep queue can only be a stop slot too.  Plans
//     CLAIM their stop slots with atomicMin((epoch, op index)) in a 16 MiB array (no hash, no
//     clearing); a plan is invalid iff an earlier op claims one of its stop slots;
//   * FORWARDING: when the earlier op's tag is no stop for the later walk (and not equal to its
//     tag), the later plan walks on from the next slot, re-claims and remembers the dependency;
//     iterated to a fixed point with __syncthreads rounds;
//   * TAG FILTER: a blocked Bloom filter over every tag inserted in this chunk (one 64-bit word,
//     three bits per tag, no false negatives).  A negative answer proves that no entry with an
//     equal tag exists, so find_best_match's walk to the end of the probe cluster - hundreds of
//     slots at high mask levels - is skipped altogether, and lookup-only ops touch no table slot;
//   * GROUP SUMMARY: the insert walk looks for the first slot whose tag has fewer trailing one bits
//     than `need` (empty, due for sweeping, or lesser than the new tag: rzip.c:264-277).  Global
//     memory (L2) holds, per group of 16 slots, a 16-bit thermometer of the minimum trailing-one
//     count; a walk skips 128 slots per 128-bit summary load and reads one group (256-bit loads);
//     lowering is one atomicAnd, repair one atomicOr;
//   * TWINS: positions p and p+1 with equal tags (buf[p] == buf[p+31]) are planned as a pair;
//   * the sweep of clean_one_from_hash is pre-resolved per batch: a parallel scan keeps a queue of
//     the next eligible slots, so the n-th overflow of a batch zeroes the n-th queued slot;
//   * SERIAL OPS USE THE WHOLE CTA: the probe sequence of find_best_match is walked 4096 slots per
//     step, equal-tag candidates are compared by the 32 warps in parallel (__ballot_sync / __ffs
//     memcmp), long matches are extended 32 KiB per step by all threads; the insert walks of the
//     cut op are warp-cooperative (first probe + 4096 slots of summary in one round trip).
//   oracle/spec_model.c is the CPU model of the speculation rules; tests/test_spec_model.py proves
//   it equivalent to the sequential restatement, tests/test_sim_parity.py runs this very file on
//   the CPU emulator, tests/test_gpu_parity.py on the B200.
#ifndef MRZ_CUSIM
#include <mutex>
#endif
#include "mrz_common.cuh"

#define FULL 0xffffffffu
#define AT 1024            // threads per CTA = ops per batch
#define EQ_CAP 4096        // eligible-slot ring (power of two)
#define LK_XCAP 2048       // a candidate still matching after this many bytes is extended by the whole CTA
#define LK_XCAP2 (1 << 20)  // ... unless more than two candidates are that long: then the warps go on in parallel up to here
#define LK_CAP 256         // equal-tag entries the CTA-wide lookup collects; more: the warp walk
#define DENSE_LOOP 0       // 1: after an emission warp 0 examines the next 32 positions from the bytes (measured slower than letting the batch carry on)
#define FWD_ITERS 6         // forwarding rounds per batch; later collisions go serial
#define MAXWALK 16384      // longer insert walks are left to the serial path
#define WLOG_CAP 64
#define GRP 8              // slots per summary group (= one 64-byte read of tags)
#define GRP_SHIFT 3
#define DUP_SALT 0x5bd1e9955bd1e995LL

enum { F_COMPLEX = 1, F_INSERT = 2, F_W1 = 4, F_CONF = 8, F_FOLLOW = 16, F_W2ELIG = 32, F_W1ELIG = 64 };
#define MAXDEP 3            // claimed slots a plan may walk past (forwarding)
#ifdef MRZ_CUSIM
#define CLAIM_EPOCH_MAX 700u       // the CPU-emulator tests exercise the wrap-around (wipe + restart) path
#else
#define CLAIM_EPOCH_MAX 0x1FFFF0u
#endif
__device__ __forceinline__ unsigned claim_key(unsigned epoch, int j, int is_w1) { return ((0x1FFFFFu - epoch) << 11) | ((unsigned)j << 1) | (unsigned)is_w1; }

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

__device__ __forceinline__ int tz1c(i64 t) {   // trailing one bits, capped at 15 (summary nibble)
    int z = tz1(t);
    return z > 15 ? 15 : (z < 0 ? 15 : z);
}
// GROUP SUMMARY, in global memory (L2 resident, 512 KiB for the 64 MiB table): per group of 16 slots a 16-bit
// thermometer - bit i set <=> every tag of the group has more than i trailing one bits (as far as is known: the
// code is a lower bound, all-zero = nothing known).  Lowering is one atomicAnd, raising one atomicOr: both fire
// and forget, so plans on any SM and commits on the main CTA maintain it without read-modify-write loops.
__device__ __forceinline__ void sum_raise(u32 *gs, i64 g, int v) {      // plan phases: the whole group was seen, min = v
    if (v > 0) atomicOr(gs + (g >> 1), (((1u << v) - 1u) & 0x7fffu) << ((g & 1) * 16));
}
__device__ __forceinline__ void sum_lower(u32 *gs, i64 g, int v) {      // commit / serial phases: a tag with v trailing ones (0: zeroed)
    atomicAnd(gs + (g >> 1), ~(0xffffu << ((g & 1) * 16)) | ((((1u << v) - 1u) & 0x7fffu) << ((g & 1) * 16)));
}

__device__ __forceinline__ i64 find_stop(const i64 *ttag, u32 *gsum, i64 h, int need, i64 hsize, i64 &stop_tag);

// Running state of the automaton, shared by the whole CTA (rzip_state, mrzip_private.h:388-416).
struct Run {
    i64 p, last, cur_p, cur_len, cur_ofs, min_mask, ins_mask, clean_ptr, count, vr;
    i64 hits, misses, inserts, vevents, nops, ntok;
    int error;
};

struct Warp {
    // constants
    const u8 *buf;
    i64 *ttag, *toff;   // the table: tags and offsets (struct hash_entry, rzip.c:59-62, split in two arrays)
    u64 *filt;
    u32 *gsum;          // group summary (global memory)
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
    i64 *prof;        // MRZ_PROFILE: cycles / calls of the parts of a serial op (dbg[20..])

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
    // table writes of the serial path: lane 0 stores, the write is logged, the summary follows
    __device__ __forceinline__ void write_slot(i64 slot, i64 off, i64 tag) {
        if (lane == 0) {
            __stcg(toff + slot, off);
            __stcg(ttag + slot, tag);
            sum_lower(gsum, slot >> GRP_SHIFT, tag ? tz1c(tag) : 0);
            if (nlog < WLOG_CAP) wlog[nlog] = (int)(slot << 1) | (tag ? 0 : 1);
        }
        nlog++;
        __syncwarp();
    }

    // rzip.c:372-397.  Returns total length (0 if < 31) and the backward part in rev.
    // The first step compares 256 bytes (most candidates fail there); a match that survives it is extended 4 KiB per
    // step - 16 independent 8-byte compares per lane in flight - so that long matches run at memory speed, not at
    // one dependent round trip per 256 bytes.
    __device__ __forceinline__ i64 extend(i64 a0, i64 b0, i64 maxlen, bool fwd) {
        // fwd: longest common prefix of buf[a0..] and buf[b0..]; else longest common suffix of buf[..a0) and buf[..b0)
        i64 f = 0;
        int K = 1;
        while (f < maxlen) {
            int first = 0x7fffffff;            // (chunk index << 3 | byte) of my first mismatch in this step
            for (int c = K - 1; c >= 0; c--) {
                const i64 k = f + (i64)(c * 32 + lane) * 8;
                u64 x = 0;
                if (k < maxlen) x = fwd ? ld_u64_unaligned(buf, a0 + k) ^ ld_u64_unaligned(buf, b0 + k)
                                        : ld_u64_unaligned(buf, a0 - 8 - k) ^ ld_u64_unaligned(buf, b0 - 8 - k);
                if (x) {
                    const int nb = fwd ? ((__ffsll((long long)x) - 1) >> 3) : (__clzll((long long)x) >> 3);
                    first = ((c * 32 + lane) << 3) | nb;
                }
            }
            for (int o = 16; o; o >>= 1) {
                const int u = __shfl_xor_sync(FULL, first, o);
                first = u < first ? u : first;
            }
            if (first != 0x7fffffff) {
                f += (i64)(first >> 3) * 8 + (first & 7);
                break;
            }
            f += (i64)K * 256;
            K = 16;
        }
        return f > maxlen ? maxlen : f;
    }
    __device__ __forceinline__ i64 match_len(i64 p0, i64 op, i64 &rev) {
        rev = 0;
        if (op >= p0) return 0;
        const i64 f = extend(p0, op, end - p0, true);
        const i64 lo = last > 0 ? last : 0;
        i64 maxb = p0 - lo;
        if (op < maxb) maxb = op;
        const i64 r = extend(p0, op, maxb, false);
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
    // (k_automaton: every warp ran match_len for a share of the equal-tag slots; results in slot order).
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

    // find_stop for the serial path, by one warp: the first probe and 32 summary quad-words (4096 slots of summary) are
    // fetched in ONE round trip, the candidate group in a second (16 lanes, one tag each); a stale candidate is repaired
    // and the next one taken from the summary already in registers.  Same result as find_stop.
    __device__ __forceinline__ i64 find_stop_warp(i64 h, int need, i64 &stop_tag) {
        if (h >= hsize) return -1;
        const i64 ngroups_all = hsize >> GRP_SHIFT;
        i64 ngroups = ((h + MAXWALK) >> GRP_SHIFT) + 1;
        if (ngroups > ngroups_all) ngroups = ngroups_all;
        const u32 pass_bit = 1u << ((need > 15 ? 15 : need) - 1);
        const i64 g0 = h >> GRP_SHIFT;
        // round trip 1: the rest of the first group (lanes 0..15) and the summary of the groups behind it (all lanes)
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
            if (!lm) {                                   // nothing in these 4096 slots: next window of summary
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

    // rzip.c:256-301 for a tag that has no equal entry in the table: every level of the displacement chain is a
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
                write_slot(sl, coff, ct);
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
    // pre_n >= 0: the CTA has already walked the probe sequence and compared the pre_n equal-tag candidates
    __device__ __noinline__ bool op(i64 pp, i64 t, i64 *stk, int pre_n = -1, const i64 *pre_len = nullptr, const i64 *pre_rev = nullptr,
                                    const i64 *pre_off = nullptr, const int *pre_pos = nullptr) {
#define PROF(slot, t0) do { if (prof && lane == 0) { prof[slot] += clock64() - (t0); prof[(slot) + 1] += 1; } } while (0)
        nops++;
        i64 bofs = 0, brev = 0;
        i64 c0 = clock64();
        const bool maybe = pre_n >= 0 ? true : pre_n == -2 ? false : filt_query(filt, t);   // negative: nothing to find
        PROF(0, c0); c0 = clock64();
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
    int lk_fe, lk_n;        // CTA-wide lookup: first empty position, equal-tag count
    int lk_pos[LK_CAP];     //                  positions (relative to the primary slot) of equal tags, ascending
    i64 lk_len[LK_CAP], lk_rev[LK_CAP], lk_off[LK_CAP];   // match_len results per candidate
    i64 lk_f[LK_CAP], lk_r[LK_CAP];                       // capped forward / backward lengths (first pass)
    unsigned epoch;         // claim epoch (monotonic over the chunk; AutoState.pad1 between launches)
    int changed[2];
    int wlog[WLOG_CAP];
    int wsum[40];
    unsigned e_idx, e_cnt;  // ring counters (monotonic)
    unsigned cutmask[AT / 32];
    int nr, pos, k, nlog, jumped, ended, bmax, red;
};

#ifdef MRZ_CUSIM
#define MRZ_SMEM_PTR(T, name) T *name = reinterpret_cast<T *>(cusim::dyn_smem())
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

struct Plan {
    int flags, dcount;
    int w1_slot, w2_slot;               // displaced occupant's new home; where (tag, pos) goes
    i64 w1_off, w1_tag;
    int ndep;                           // forwarding: slots this plan walked past because an EARLIER op of the batch
    int dep_slot[MAXDEP], dep_own[MAXDEP];   // claims them with a tag that is no stop for this walk
};

// First slot s >= h with tz1(tags[s]) < need (an empty slot has tag 0, i.e. zero trailing ones), or -1 when
// the walk reaches the end of the table / gets too long.  Whole groups whose summary says "every tag has >= need
// trailing ones" are skipped unseen (one 128-bit load covers 8 groups = 128 slots); a group that is scanned
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
        const i64 gend = (g + 1) << GRP_SHIFT;
        int mn = 15;
        if ((h & (GRP - 1)) == GRP - 1) {   // the usual case at mask levels >= 4 (primary slots end in ...1111): one tag, one load
            const i64 t = __ldcg(ttag + h);
            int z = tz1(t);
            if (z < 0) z = 64;
            if (z < need) { stop_tag = t; return h; }
        } else
        for (i64 o = h & ~(i64)7; o < gend; o += 8) {
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

// Phase A: plan the insert of (tag) against the table without modifying anything (oracle/spec_model.c
// plan_op).  The caller has established through the tag filter that no equal tag exists.  The walk starts at
// `start` (the tag's primary slot, or the slot after a claimed one when the plan is forwarded).
__device__ __forceinline__ void plan_op(Plan &r, const i64 *ttag, const i64 *toff, u32 *gsum, const u64 *filt, i64 tag,
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
            else { r.flags |= F_COMPLEX; return; }     // a second displacement: serial
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
                                              int a_flags, int a_w2, i64 hmask, i64 hsize, i64 min_mask, int chain, i64 end,
                                              i64 last) {
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
    if ((tag & btm) != btm) {
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

// Refill the ring of eligible slots (those clean_one_from_hash will zero next, in order) by a
// parallel scan of the table from e_scan; only meaningful once ins_mask == 2*min_mask+1.
// The scan never goes past scan_end (the table end, or the window the current batch was planned with).
__device__ __forceinline__ void refill_queue(Smem *S, const i64 *ttag, i64 scan_end, int want) {
    const int tid = threadIdx.x;
    const i64 btm = (S->run.min_mask << 1) | 1;
    const i64 hsize = scan_end;
    for (;;) {
        __syncthreads();
        int avail = (int)(S->e_cnt - S->e_idx);
        i64 scan = S->e_scan;
        if (avail >= want || scan >= hsize || avail + 2 * AT > EQ_CAP) break;
        // each thread examines two consecutive slots
        i64 s0 = scan + 2 * tid;
        bool g0 = false, g1 = false;
        if (s0 < hsize) { i64 t = __ldcg(ttag + s0); g0 = t != 0 && (t & btm) != btm; }
        if (s0 + 1 < hsize) { i64 t = __ldcg(ttag + s0 + 1); g1 = t != 0 && (t & btm) != btm; }
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

// PLAN OFFLOAD.  The plan phase of a batch is bound by what one SM can issue (about one uncoalesced request per
// cycle: ~20 k cycles for 1024 plans), so it is spread over H helper CTAs of the same cooperative grid: the main CTA
// publishes the batch, every helper thread plans op (tid * H + helper) against the table - read-only, through L2 -
// and writes the plan record; two flag hand-offs per batch, no grid barrier.  Commits, conflicts, sweeps and the
// serial ops stay on the main CTA.
struct PlanIO {
    unsigned seq, done;       // batch number published by the main CTA / helper threads finished (both zeroed per launch)
    int nr, chain;
    i64 min_mask, ins_mask, last_now, hmask, hsize, end, seg_lo;
    u32 rel[AT];
    i64 tag[AT];
    int flags[AT], dcount[AT], w1_slot[AT], w2_slot[AT];
    i64 w1_off[AT], w1_tag[AT];
};
__device__ __forceinline__ unsigned wait_change(volatile unsigned *a, unsigned old) {
#ifdef MRZ_CUSIM
    return cusim::wait_change(a, old);
#else
    unsigned v;
    while ((v = *a) == old) __nanosleep(40);
    return v;
#endif
}

// Match extension by the whole CTA: longest common prefix of buf[a0..] / buf[b0..] (fwd) or longest common suffix of
// buf[..a0) / buf[..b0), at most maxlen bytes; 32 KiB per step, four independent 8-byte compares per thread in flight.
__device__ __forceinline__ i64 cta_extend(Smem *S, const u8 *buf, i64 a0, i64 b0, i64 maxlen, bool fwd) {
    const int tid = threadIdx.x;
    i64 f = 0;
    while (f < maxlen) {
        if (tid == 0) S->lk_fe = 0x7fffffff;
        __syncthreads();
        int first = 0x7fffffff;
        for (int c = 3; c >= 0; c--) {
            const int ci = c * AT + tid;
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
        f += (i64)4 * AT * 8;
    }
    return f > maxlen ? maxlen : f;
}

// Runs the automaton over the current segment's candidate list (or to the end of the chunk).
__global__ void __launch_bounds__(AT, 1)
k_automaton(const u8 *__restrict__ buf, i64 *ttag, i64 *toff, u64 *filt, u32 *gsum_global, unsigned *claim, const i64 *__restrict__ hidx,
            AutoState *st, const i64 *__restrict__ tile_off, const u32 *__restrict__ cand_pos,
            const i64 *__restrict__ cand_tag, Token *toks, i64 *dbg, PlanIO *io) {
    if (st->done || st->error) return;
    const int H = (int)gridDim.x - 1;             // helper CTAs (0: the main CTA plans by itself)
    if (blockIdx.x > 0) {
        // ------------------------------------------------------------ helper CTA: plans on request, touches nothing else
        const int hid = (int)blockIdx.x - 1, per = AT / H;
        if ((int)threadIdx.x >= per) return;
        unsigned seen = 0;
        for (;;) {
            const unsigned sq = wait_change(&io->seq, seen);
            if (sq == 0xffffffffu) return;
            seen = sq;
            __threadfence();
            const int nr = *(volatile int *)&io->nr, chain = *(volatile int *)&io->chain;
            const i64 min_mask = *(volatile i64 *)&io->min_mask, ins_mask = *(volatile i64 *)&io->ins_mask;
            const i64 hmask = *(volatile i64 *)&io->hmask, hsize = *(volatile i64 *)&io->hsize;
            const int j = (int)threadIdx.x * H + hid;
            if (j < nr) {
                const u32 rel = __ldcg(io->rel + j);
                const i64 tag = __ldcg(io->tag + j);
                const bool same = j > 0 && __ldcg(io->rel + j - 1) + 1 == rel && __ldcg(io->tag + j - 1) == tag;
                const bool same_prev = j > 1 && __ldcg(io->rel + j - 2) + 2 == rel && __ldcg(io->tag + j - 2) == tag;
                Plan pl;
                plan_init(pl);
                // a twin plans its leader's op itself: same tag, same table, same plan
                if (filt_query(filt, tag)) pl.flags |= F_COMPLEX;
                else if ((tag & ins_mask) == ins_mask)
                    plan_op(pl, ttag, toff, gsum_global, filt, tag, tag & hmask, hmask, hsize, min_mask, chain);
                if (same && !same_prev) {
                    const int a_flags = pl.flags, a_w2 = pl.w2_slot;
                    plan_init(pl);
                    plan_follower(pl, ttag, gsum_global, buf, *(volatile i64 *)&io->seg_lo + (i64)rel, tag, a_flags, a_w2, hmask, hsize,
                                  min_mask, chain, *(volatile i64 *)&io->end, *(volatile i64 *)&io->last_now);
                }
                __stcg(io->flags + j, pl.flags); __stcg(io->dcount + j, pl.dcount);
                __stcg(io->w1_slot + j, pl.w1_slot); __stcg(io->w2_slot + j, pl.w2_slot);
                __stcg(io->w1_off + j, pl.w1_off); __stcg(io->w1_tag + j, pl.w1_tag);
            }
            __threadfence();
            __syncwarp();
            if ((threadIdx.x & 31) == 0) atomicAdd(&io->done, 32u);
        }
    }
    unsigned plan_seq = 0;
    MRZ_SMEM_PTR(Smem, S);
    i64 tm[20];
#pragma unroll
    for (int q = 0; q < 20; q++) tm[q] = 0;
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
        S->epoch = (unsigned)st->pad1; S->changed[0] = S->changed[1] = 0;
    }
    Warp w;   // used by warp 0 only
    w.buf = buf; w.ttag = ttag; w.toff = toff; w.filt = filt; w.gsum = gsum_global; w.toks = toks; w.lane = lane;
    w.n = n; w.end = end; w.hmask = hmask; w.hsize = hsize; w.limit = limit; w.tok_cap = st->tok_cap; w.chain = chain;
    w.wlog = S->wlog; w.nlog = 0; w.clean_hint = 0; w.prof = dbg ? dbg + 20 : nullptr;
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
            refill_queue(S, ttag, hsize, bmax);
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
        plan_init(pl);
        i64 my_pos = 0, my_tag = 0;
        bool follower = false;
        int lead_w2 = -1;
        const i64 t_plan0 = clock64();
        if (dbg && tid == 0) S->red = 0;
        if (H > 0) {
            // publish the batch, wait for the helpers' plan records
            if (tid < nr) { __stcg(io->rel + tid, S->b_rel[tid]); __stcg(io->tag + tid, S->b_tag[tid]); }
            if (tid == 0) {
                io->nr = nr; io->chain = chain; io->min_mask = min_mask; io->ins_mask = ins_mask; io->last_now = last_now;
                io->hmask = hmask; io->hsize = hsize; io->end = end; io->seg_lo = seg_lo;
            }
            plan_seq++;
            __syncthreads();
            if (tid == 0) {
                __threadfence();
                *(volatile unsigned *)&io->seq = plan_seq;
                const unsigned target = plan_seq * (unsigned)AT;
                unsigned v = *(volatile unsigned *)&io->done;
                while (v != target) v = wait_change(&io->done, v);
                __threadfence();
            }
            __syncthreads();
            if (tid < nr) {
                const u32 rel = S->b_rel[tid];
                my_pos = seg_lo + (i64)rel; my_tag = S->b_tag[tid];
                bool same = tid > 0 && S->b_rel[tid - 1] + 1 == rel && S->b_tag[tid - 1] == my_tag;
                bool same_prev = tid > 1 && S->b_rel[tid - 2] + 2 == rel && S->b_tag[tid - 2] == my_tag;
                follower = same && !same_prev;
                pl.flags = __ldcg(io->flags + tid); pl.dcount = __ldcg(io->dcount + tid);
                pl.w1_slot = __ldcg(io->w1_slot + tid); pl.w2_slot = __ldcg(io->w2_slot + tid);
                pl.w1_off = __ldcg(io->w1_off + tid); pl.w1_tag = __ldcg(io->w1_tag + tid);
                S->pl_fl[tid] = pl.flags; S->pl_w2[tid] = pl.w2_slot;
            }
            __syncthreads();
            if (follower) lead_w2 = S->pl_w2[tid - 1];
        } else {
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
                    plan_op(pl, ttag, toff, gsum_global, filt, my_tag, my_tag & hmask, hmask, hsize, min_mask, chain);
            }
            S->pl_fl[tid] = pl.flags; S->pl_w2[tid] = pl.w2_slot;
            if (dbg) {   // per-op plan latency: mean and per-batch maximum (the barrier waits for the slowest walk)
                const i64 dt = clock64() - t_plan0;
                atomicAdd((unsigned long long *)&dbg[17], (unsigned long long)dt);
                atomicMax(&S->red, (int)dt);
            }
        }
        __syncthreads();
        if (dbg && tid == 0) dbg[18] += S->red;
        if (follower) {
            lead_w2 = S->pl_w2[tid - 1];
            plan_follower(pl, ttag, gsum_global, buf, my_pos, my_tag, S->pl_fl[tid - 1], lead_w2, hmask, hsize, min_mask,
                          chain, end, last_now);
        }
        }
        TICK(2);   // plan
        // ---------------------------------------------------------------- conflicts: exact stop-slot claims with forwarding
        // Every planned insert claims its stop slot(s) in claim[] (atomicMin of (epoch, op index)).  A plan whose stop slot
        // is claimed by an EARLIER op either walks on from the next slot - when that op's tag is no stop for this walk,
        // the claimed slot is remembered as a dependency - or gives up (F_CONF: serial).  Iterated to a fixed point.
        const bool planned_ins = tid < nr && (pl.flags & (F_COMPLEX | F_INSERT)) == F_INSERT;
        const int my_need = tz1(my_tag) > __popcll((u64)btm) ? tz1(my_tag) : __popcll((u64)btm);
        for (int it = 0;; it++) {
            if (S->epoch >= CLAIM_EPOCH_MAX) {   // claim keys exhausted: wipe the array, restart the epochs
                __syncthreads();
                for (i64 q = tid; q < hsize + 8; q += AT) claim[q] = 0xffffffffu;
                __syncthreads();
                if (tid == 0) S->epoch = 0;
                __syncthreads();
            }
            const unsigned ep = S->epoch;
            const unsigned ep_hi = claim_key(ep, 0, 0) >> 11;
            bool live = planned_ins && !(pl.flags & (F_COMPLEX | F_CONF));
            if (live && steady &&   // a stop slot that sits in the sweep queue: the order of sweeps and inserts matters
                (((pl.flags & F_W2ELIG) && pl.w2_slot >= w_lo && pl.w2_slot < w_hi) ||
                 ((pl.flags & F_W1ELIG) && pl.w1_slot >= w_lo && pl.w1_slot < w_hi))) {
                pl.flags |= F_CONF;
                live = false;
                if (dbg) atomicAdd((unsigned long long *)&dbg[32], 1ULL);   // stop slot in the sweep queue
            }
            if (live) {
                atomicMin(claim + pl.w2_slot, claim_key(ep, tid, 0));
                if (pl.flags & F_W1) atomicMin(claim + pl.w1_slot, claim_key(ep, tid, 1));
            }
            __threadfence_block();
            __syncthreads();
            if (tid == 0) { S->epoch = ep + 1; S->changed[(it + 1) & 1] = 0; }
            bool chg = false;
            if (planned_ins && !live && it == 0) chg = false;
            if (live) {
                bool bad = false;
                // a follower's plan builds on its leader's write: only older claims conflict, and the leader must not have moved
                const int before = (pl.flags & F_FOLLOW) ? tid - 1 : tid;
                int why = 0;
                if ((pl.flags & F_FOLLOW) && S->pl_w2[tid - 1] != lead_w2) { bad = true; why = 33; }
                for (int d = 0; d < pl.ndep && !bad; d++) {   // slots I walked past on the strength of an earlier op's claim
                    const unsigned o = __ldcg(claim + pl.dep_slot[d]);
                    const int oi = (int)((o >> 1) & 1023u);
                    if ((o >> 11) != ep_hi || (o & 1u) || oi >= tid) { bad = true; why = 34; break; }
                    const i64 ot = S->b_tag[oi];
                    if (tz1(ot) < my_need || (ot & btm) != btm || ot == my_tag) { bad = true; why = 34; break; }
                    pl.dep_own[d] = oi;
                }
                if (!bad) {
                    const unsigned o = __ldcg(claim + pl.w2_slot);
                    const int oi = (int)((o >> 1) & 1023u);
                    if (oi < before) {   // my stop slot goes to an earlier op
                        const i64 ot = S->b_tag[oi];
                        if ((pl.flags & F_FOLLOW) || (o & 1u) || pl.ndep == MAXDEP || it >= FWD_ITERS || tz1(ot) < my_need ||
                            (ot & btm) != btm || ot == my_tag) {   // an equal tag: my lookup must meet its entry (serial)
                            bad = true;
                            why = (pl.flags & F_FOLLOW) ? 33 : (o & 1u) ? 35 : ot == my_tag ? 36 : (tz1(ot) < my_need || (ot & btm) != btm) ? 37 : 38;
                        }
                        else {           // its tag is no stop for my walk: go on from the next slot
                            pl.dep_slot[pl.ndep] = pl.w2_slot; pl.dep_own[pl.ndep] = oi; pl.ndep++;
                            plan_op(pl, ttag, toff, gsum_global, filt, my_tag, (i64)pl.w2_slot + 1, hmask, hsize, min_mask, chain);
                            S->pl_w2[tid] = pl.w2_slot;
                            chg = true;
                            if (dbg) atomicAdd((unsigned long long *)&dbg[16], 1ULL);
                        }
                    } else if ((pl.flags & F_W1) && (int)((__ldcg(claim + pl.w1_slot) >> 1) & 1023u) < tid) {
                        bad = true;
                        why = 39;
                    }
                }
                if (bad) { pl.flags |= F_CONF; chg = true; if (dbg) atomicAdd((unsigned long long *)&dbg[why], 1ULL); }
                if (chg) S->pl_fl[tid] = pl.flags;
            }
            if (chg) S->changed[it & 1] = 1;
            __syncthreads();
            if (!S->changed[it & 1]) break;
        }
        if (tid < nr) { S->pl_fl[tid] = pl.flags; S->pl_w2[tid] = pl.w2_slot; }
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
            const unsigned e_idx0 = S->e_idx;
            bool c = false;
            if (tid >= pos && tid < nr) {
                c = (pl.flags & (F_COMPLEX | F_CONF)) != 0;
                c = c || (cur_len >= MRZ_MIN_MATCH && my_pos >= cur_p + MRZ_MIN_MATCH);   // an emission is due
                if (!c && (pl.flags & F_INSERT) && pl.dcount == 1) {
                    i64 T = cnt0 + (((unsigned)S->pre[tid] - (unsigned)base) & 0x7ffu);
                    if (T > limit && (!steady || T - limit > avail)) c = true;             // needs a serial sweep step
                }
            }
            {
                const unsigned bal = __ballot_sync(FULL, c);
                if (lane == 0) S->cutmask[wid] = bal;
            }
            __syncthreads();   // every thread has read pos / count / queue state; the mask is complete
            int k = nr;        // first op of [pos, nr) that cannot be committed from its plan
            for (int wd = pos >> 5; wd < ((nr + 31) >> 5); wd++) {
                const unsigned m = S->cutmask[wd];
                if (m) { k = wd * 32 + __ffs(m) - 1; break; }
            }
            // commit [pos, k)
            if (tid >= pos && tid < k && (pl.flags & F_INSERT)) {
                if (pl.flags & F_W1) {
                    __stcg(toff + pl.w1_slot, pl.w1_off);
                    __stcg(ttag + pl.w1_slot, pl.w1_tag);
                    sum_lower(gsum_global, pl.w1_slot >> GRP_SHIFT, tz1c(pl.w1_tag));
                }
                // a twin that overwrites its leader's entry in the same commit must win the slot
                const bool overwritten = tid + 1 < k && (S->pl_fl[tid + 1] & F_FOLLOW) && S->pl_w2[tid + 1] == pl.w2_slot;
                if (!overwritten) {
                    __stcg(toff + pl.w2_slot, my_pos);
                    __stcg(ttag + pl.w2_slot, my_tag);
                    sum_lower(gsum_global, pl.w2_slot >> GRP_SHIFT, tz1c(my_tag));
                }
                filt_set(filt, my_tag);
                if (pl.flags & F_FOLLOW) filt_set(filt, my_tag ^ DUP_SALT);   // the tag is in the table twice now
            }
            int cleans = 0;
            if (k > pos) {
                const unsigned d = (unsigned)S->pre[k - 1] - (unsigned)base;
                i64 T = cnt0 + (d & 0x7ffu);
                cleans = T > limit ? (int)(T - limit) : 0;
                if (tid < cleans) {
                    const int sl = S->E[(e_idx0 + tid) & (EQ_CAP - 1)];
                    __stcg(ttag + sl, (i64)0);
                    __stcg(toff + sl, (i64)0);
                    sum_lower(gsum_global, sl >> GRP_SHIFT, 0);
                }
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
            if (tid == k) {   // why is this op serial?
                int why = (pl.flags & F_COMPLEX) ? (follower ? 12 : ((pl.flags & F_INSERT) ? 13 : 11)) : (pl.flags & F_CONF) ? 14 : 15;
                if (dbg) atomicAdd((unsigned long long *)&dbg[why], 1ULL);
            }
            // ---- the lookup of op k (find_best_match, rzip.c:426-462) by the whole CTA: the probe sequence is walked
            // 4096 slots per step, equal tags are collected in slot order, and the 32 warps share the match_len calls
            int pre_n = -2;   // -2: the filter says no equal tag; -1: ask / walk on the warp; >= 0: candidates compared
            const i64 c_lk0 = clock64();
            {
                const i64 tk = S->b_tag[k], pk = seg_lo + (i64)S->b_rel[k];
                if (filt_query(filt, tk)) {
                    pre_n = -1;
                    const i64 h = tk & hmask;
                    if (tid == 0) S->lk_fe = 0x7fffffff;
                    __syncthreads();
                    int n_tot = 0;
                    bool found = false;
                    for (int base = 0;; base += AT * 4) {
                        i64 et[4];
#pragma unroll
                        for (int q = 0; q < 4; q++) et[q] = __ldcg(ttag + ((h + base + tid * 4 + q) & hmask));
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
                        int total;
                        int ex = block_excl_scan(__popc(eq4), S->wsum, total);
                        for (int q = 0; q < 4; q++)
                            if ((eq4 >> q) & 1u) {
                                if (n_tot + ex < LK_CAP) S->lk_pos[n_tot + ex] = base + tid * 4 + q;
                                ex++;
                            }
                        n_tot += total;
                        found = fe != 0x7fffffff;
                        if (found || n_tot > LK_CAP || base + AT * 4 >= hsize) break;
                    }
                    __syncthreads();
                    if (found && n_tot <= LK_CAP) {
                        // single_match_len (rzip.c:372-397) per candidate: the warps share the candidates for the first
                        // LK_XCAP bytes either side; whatever still matches there is extended by the whole CTA, 32 KiB per step
                        w.last = S->run.last;
                        const i64 lo = w.last > 0 ? w.last : 0;
                        const i64 maxf = end - pk;
                        if (tid == 0) S->lk_n = 0;      // candidates still matching at the cap
                        __syncthreads();
                        for (int q = wid; q < n_tot; q += AT / 32) {
                            const i64 off = __ldcg(toff + ((h + S->lk_pos[q]) & hmask));
                            i64 f = 0, r = 0;
                            if (off < pk) {
                                i64 maxb = pk - lo;
                                if (off < maxb) maxb = off;
                                const i64 mf = maxf < LK_XCAP ? maxf : LK_XCAP, mb = maxb < LK_XCAP ? maxb : LK_XCAP;
                                {   // the first 256 bytes of both directions in one round trip
                                    const i64 kk = lane * 8;
                                    u64 xf = 0, xb = 0;
                                    if (kk < mf) xf = ld_u64_unaligned(buf, pk + kk) ^ ld_u64_unaligned(buf, off + kk);
                                    if (kk < mb) xb = ld_u64_unaligned(buf, pk - 8 - kk) ^ ld_u64_unaligned(buf, off - 8 - kk);
                                    unsigned mm = __ballot_sync(FULL, xf != 0);
                                    if (mm) {
                                        const int l = __ffs(mm) - 1;
                                        int nb = xf ? ((__ffsll((long long)xf) - 1) >> 3) : 8;
                                        nb = __shfl_sync(FULL, nb, l);
                                        f = l * 8 + nb;
                                    } else
                                        f = mf > 256 ? 256 + w.extend(pk + 256, off + 256, mf - 256, true) : mf;
                                    if (f > mf) f = mf;
                                    mm = __ballot_sync(FULL, xb != 0);
                                    if (mm) {
                                        const int l = __ffs(mm) - 1;
                                        int nb = xb ? (__clzll((long long)xb) >> 3) : 8;
                                        nb = __shfl_sync(FULL, nb, l);
                                        r = l * 8 + nb;
                                    } else
                                        r = mb > 256 ? 256 + w.extend(pk - 256, off - 256, mb - 256, false) : mb;
                                    if (r > mb) r = mb;
                                }
                                if (lane == 0 && ((f == LK_XCAP && maxf > LK_XCAP) || (r == LK_XCAP && maxb > LK_XCAP))) atomicAdd(&S->lk_n, 1);
                            }
                            if (lane == 0) { S->lk_f[q] = f; S->lk_r[q] = r; S->lk_off[q] = off; }
                        }
                        __syncthreads();
                        // many long candidates (dense repeats, deep chains): the warps carry on in parallel, one candidate
                        // each, up to LK_XCAP2; the whole CTA then only extends what is still matching there
                        const i64 cap = S->lk_n > 2 ? (i64)LK_XCAP2 : (i64)LK_XCAP;
                        if (cap != LK_XCAP) {
                            for (int q = wid; q < n_tot; q += AT / 32) {
                                const i64 off = S->lk_off[q];
                                if (off >= pk) continue;
                                i64 f = S->lk_f[q], r = S->lk_r[q];
                                i64 maxb = pk - lo;
                                if (off < maxb) maxb = off;
                                if (f == LK_XCAP && maxf > LK_XCAP) f += w.extend(pk + LK_XCAP, off + LK_XCAP, (maxf < cap ? maxf : cap) - LK_XCAP, true);
                                if (r == LK_XCAP && maxb > LK_XCAP) r += w.extend(pk - LK_XCAP, off - LK_XCAP, (maxb < cap ? maxb : cap) - LK_XCAP, false);
                                __syncwarp();
                                if (lane == 0) { S->lk_f[q] = f; S->lk_r[q] = r; }
                            }
                            __syncthreads();
                        }
                        for (int q = 0; q < n_tot; q++) {
                            const i64 off = S->lk_off[q];
                            i64 f = S->lk_f[q], r = S->lk_r[q], m = 0;
                            if (off < pk) {
                                i64 maxb = pk - lo;
                                if (off < maxb) maxb = off;
                                if (f == cap && maxf > cap) f += cta_extend(S, buf, pk + cap, off + cap, maxf - cap, true);
                                if (r == cap && maxb > cap) r += cta_extend(S, buf, pk - cap, off - cap, maxb - cap, false);
                                m = f + r < MRZ_MIN_MATCH ? 0 : f + r;
                            }
                            if (tid == 0) { S->lk_len[q] = m; S->lk_rev[q] = r; }
                        }
                        __syncthreads();
                        pre_n = n_tot;
                    }
                }
            }
            if (dbg && tid == 0) dbg[19] += clock64() - c_lk0;
            // ---- op k on the live table, reference semantics (warp 0)
            if (wid == 0) {
                w.load(S->run);
                w.nlog = 0;
                i64 hint = w.clean_ptr;
                if (steady) hint = (S->e_cnt != S->e_idx) ? (i64)S->E[S->e_idx & (EQ_CAP - 1)] : (S->e_scan > hint ? S->e_scan : hint);
                w.clean_hint = hint;
                const i64 pk = seg_lo + (i64)S->b_rel[k], tk = S->b_tag[k];
                w.p = pk;
                bool jumped = w.op(pk, tk, S->stk, pre_n, S->lk_len, S->lk_rev, S->lk_off, S->lk_pos);
                w.clean_hint = 0;
                const i64 c_dense = clock64();
                if (jumped && !w.error) {
                    // dense-match regime: examine the next positions directly from the bytes until a
                    // window of 32 positions passes without an emission, then re-synchronise with the list
                    for (; DENSE_LOOP;) {
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
                        bool again = false, bail = false;
                        int lim = (w.end - w.p) < 32 ? (int)(w.end - w.p) : 32;   // positions b0..min(b0+31,end)
                        // one round trip for the 32 filter words (op() asks again: a tag inserted inside this window is seen there)
                        const unsigned fmask = __ballot_sync(FULL, filt_query(filt, tq));
                        for (int j = 0; j < lim; j++) {
                            i64 tj = __shfl_sync(FULL, tq, j);
                            if ((tj & w.min_mask) != w.min_mask) continue;
                            if ((fmask >> j) & 1u) {   // an equal tag may exist: its lookup belongs to the whole CTA (next cut op)
                                w.p = b0 + j - 1;
                                bail = true;
                                break;
                            }
                            w.p = b0 + j;
                            if (w.op(b0 + j, tj, S->stk)) { again = true; break; }
                        }
                        if (again) continue;
                        if (!bail) w.p = b0 + lim - 1;
                        break;
                    }
                    // re-synchronise with the candidate list - unless the batch simply carries on past the jump (below)
                    bool stay = !w.error && w.nlog <= WLOG_CAP && w.min_mask == min_mask && w.ins_mask == ins_mask && w.p >= pk &&
                                (i64)S->b_rel[nr - 1] > w.p - seg_lo;
                    if (!stay) {
                        i64 nli = ncand;
                        if (!w.error && w.p < w.end && w.p + 1 < seg_hi) {
                            i64 rel = w.p + 1 - seg_lo;
                            i64 tile = rel / SCAN_TILE;
                            nli = list_seek(cand_pos, tile_off[tile], tile_off[tile + 1], (u32)rel, lane);
                        }
                        if (lane == 0) S->li = nli;
                    }
                    if (dbg && lane == 0) { dbg[30] += clock64() - c_dense; dbg[31] += 1; }
                }
                w.store(S->run);
                if (lane == 0) {
                    bool mask_changed = w.min_mask != min_mask || w.ins_mask != ins_mask;
                    bool overflow = w.nlog > WLOG_CAP;
                    S->jumped = jumped;
                    S->nlog = overflow ? 0 : w.nlog;
                    int kp = nr;
                    if (jumped && !mask_changed && !overflow && !w.error && w.p >= pk) {   // (a short match can end BEFORE op k: p goes back, rzip.c:597)
                        // an emission skips the ops up to the end of the match; the plans beyond it stay good (only a
                        // write to a plan's stop slot can spoil it): carry on with the first op past the jump
                        int a = k + 1, b = nr;
                        const i64 relp = w.p - seg_lo;
                        while (a < b) {
                            const int m = (a + b) >> 1;
                            if ((i64)S->b_rel[m] <= relp) a = m + 1; else b = m;
                        }
                        kp = a;
                    }
                    if (mask_changed || overflow || w.error || (jumped && kp >= nr)) {
                        S->ended = 1;
                        S->e_idx = S->e_cnt = 0; S->e_scan = w.clean_ptr;    // rebuild the queue
                        if (!jumped) S->li = (i64)S->b_li[k] + 1;
                    } else {
                        if (steady) sync_queue(S, w.nlog);
                        S->pos = jumped ? kp : k + 1;
                    }
                }
            }
            __syncthreads();
            TICK(5);   // serial op
            if (S->ended) break;
            // the op overwrote a queued slot and the queue was cut there: rebuild its tail, but only inside the
            // window this batch was planned with
            if (steady && S->e_scan < w_hi) refill_queue(S, ttag, w_hi, EQ_CAP);
            // later plans that read a slot this op wrote are no longer valid; nor is its twin's plan
            const int npos = S->pos;   // k + 1, or the first op past an emission's jump (the ops in between never run)
            if (tid == npos && (pl.flags & F_FOLLOW)) pl.flags |= F_CONF;
            if (tid >= npos && tid < nr)   // a plan forwarded over the claim of op k / of a skipped op assumed that op's planned write
                for (int d = 0; d < pl.ndep; d++)
                    if (pl.dep_own[d] >= k && pl.dep_own[d] < npos) pl.flags |= F_CONF;
            if (tid >= npos && tid < nr && !(pl.flags & (F_COMPLEX | F_CONF))) {
                const int nl = S->nlog;
                for (int q = 0; q < nl; q++) {
                    int sl = S->wlog[q] >> 1;
                    bool hit = sl == pl.w2_slot || ((pl.flags & F_W1) && sl == pl.w1_slot);
                    for (int d = 0; d < pl.ndep; d++) hit = hit || sl == pl.dep_slot[d];
                    if (hit) {
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
            if (S->jumped) S->bmax = st->pad0 > 0 ? st->pad0 : AT;   // measured: restarting small after an emission only adds batches
            else if (S->bmax < AT) S->bmax *= 2;
        }
        __syncthreads();
        TICK(7);
    }
    if (tid == 0 && dbg)
        for (int q = 0; q < 11; q++) dbg[q] += tm[q];
    if (tid == 0 && H > 0) {   // release the helper CTAs
        __threadfence();
        *(volatile unsigned *)&io->seq = 0xffffffffu;
    }
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
        st->pad1 = (int)S->epoch;
    }
}

extern "C" size_t mrz_planio_bytes() { return sizeof(PlanIO); }

// helpers = number of plan-offload CTAs (0, or a power of two <= 32); > 0 needs a cooperative launch: the CTAs wait on one another.
extern "C" int mrz_launch_automaton(const u8 *buf, i64 *ttag, i64 *toff, u64 *filt, u32 *gsum, unsigned *claim, const i64 *hidx, AutoState *st,
                                    const i64 *tile_off, const u32 *cand_pos, const i64 *cand_tag, Token *toks, i64 *dbg, int helpers,
                                    void *planio, cudaStream_t s) {
    PlanIO *io = (PlanIO *)planio;
#ifndef MRZ_CUSIM
    // once per device: mrz_run_variants launches this kernel from many host threads at once, and the attribute must not
    // be rewritten under a concurrent launch of the same function
    cudaError_t e = cudaSuccess;
    {
        static std::mutex mu;
        static bool set_for[64];
        int dev = 0;
        cudaGetDevice(&dev);
        std::lock_guard<std::mutex> g(mu);
        if (!set_for[dev & 63]) {
            e = cudaFuncSetAttribute(k_automaton, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(Smem));
            if (e != cudaSuccess) return (int)e;
            set_for[dev & 63] = true;
        }
    }
    if (helpers > 0) {
        e = cudaMemsetAsync(io, 0, 8, s);   // seq, done
        if (e != cudaSuccess) return (int)e;
        void *args[] = {(void *)&buf, (void *)&ttag, (void *)&toff, (void *)&filt, (void *)&gsum, (void *)&claim, (void *)&hidx, (void *)&st,
                        (void *)&tile_off, (void *)&cand_pos, (void *)&cand_tag, (void *)&toks, (void *)&dbg, (void *)&io};
        return (int)cudaLaunchCooperativeKernel((const void *)k_automaton, dim3(1 + helpers), dim3(AT), args, sizeof(Smem), s);
    }
    k_automaton<<<1, AT, sizeof(Smem), s>>>(buf, ttag, toff, filt, gsum, claim, hidx, st, tile_off, cand_pos, cand_tag, toks, dbg, io);
    return (int)cudaGetLastError();
#else
    if (helpers > 0) {
        memset(io, 0, 8);
        MRZ_LAUNCH_COOP(k_automaton, 1 + helpers, AT, sizeof(Smem), s, buf, ttag, toff, filt, gsum, claim, hidx, st, tile_off, cand_pos, cand_tag,
                        toks, dbg, io);
    } else
        MRZ_LAUNCH_COOP(k_automaton, 1, AT, sizeof(Smem), s, buf, ttag, toff, filt, gsum, claim, hidx, st, tile_off, cand_pos, cand_tag, toks, dbg, io);
    return 0;
#endif
}
