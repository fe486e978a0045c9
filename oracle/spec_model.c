/* TEST INFRASTRUCTURE — CPU model of the B200 automaton's batched speculation (automaton.cu v2).
 *
 * Same inputs/outputs as orc_rzip_chunk, but the table automaton is driven the way the GPU drives it:
 *   phase A  every candidate of a batch is *planned* against a snapshot of the table (read-only):
 *            lookup walk, insert target/kind, one level of displacement;
 *   phase B  plans are committed in order; an op whose read ranges contain a slot written by an
 *            earlier op of the batch, or that needs state a plan cannot know (equal tags -> match_len,
 *            victim eviction, deep displacement, the clean window, a pending emission, a mask
 *            change), is executed serially on the live table with the reference semantics.
 * tests/test_spec_model.py checks that this produces byte-identical streams to the plain
 * restatement for every batch size — i.e. that the conflict rules are sufficient.  Nothing in the
 * product links this file.
 */
#include "rzip_oracle.c"
#include <stdio.h>

enum { F_SKIP = 1, F_COMPLEX = 2, F_INSERT = 4, F_W1 = 8, F_FOLLOW = 16 };
#define MAXWALK 16384

typedef struct {
    int flags, dcount;
    i64 pos, tag;
    i64 rd_lo, rd_hi, rd2_lo, rd2_hi; /* inclusive slot ranges read; rd2_lo < 0: none */
    i64 w1_slot, w1_off, w1_tag;      /* displaced occupant's new home */
    i64 w2_slot;                      /* where (tag,pos) goes */
} rec;

typedef struct {
    i64 batches, planned_commits, serial_ops, static_conflicts, extra_conflicts, complex_ops, emit_cuts, count_cuts;
} spec_stats;

static int nonempty(const slot *s) { return (s->off | s->tag) != 0; }
static int g_range_rule = 0;   /* SPEC_RANGE=1: the older, conservative read-range rules */
static int in_win(i64 s, i64 w_lo, i64 w_hi) { return s >= w_lo && s < w_hi; }

/* Tag filter: one 64-bit word per hash bucket, 3 bits per tag (blocked Bloom filter).  No false
 * negatives: every tag ever inserted into the table during this chunk has its bits set, so a
 * negative answer proves that no entry with this tag exists => the lookup walk can be skipped. */
#define FILT_BITS 24
typedef uint64_t u64;
static inline void filt_hash(i64 tag, u64 *word, u64 *bits) {
    u64 x = (u64)tag;
    x ^= x >> 33; x *= 0xff51afd7ed558ccdULL; x ^= x >> 33; x *= 0xc4ceb9fe1a85ec53ULL; x ^= x >> 33;
    *word = x >> (64 - FILT_BITS);
    *bits = ((u64)1 << (x & 63)) | ((u64)1 << ((x >> 6) & 63)) | ((u64)1 << ((x >> 12) & 63));
}
static inline int filt_query(const u64 *F, i64 tag) { u64 w, b; filt_hash(tag, &w, &b); return (F[w] & b) == b; }
static inline void filt_set(u64 *F, i64 tag) { u64 w, b; filt_hash(tag, &w, &b); F[w] |= b; }

/* phase A: plan one op against the current table without modifying anything */
static void plan_op(const ctx *c, const u64 *F, rec *r, i64 pos, i64 tag, i64 ins_mask, i64 w_lo, i64 w_hi) {
    memset(r, 0, sizeof *r);
    r->pos = pos; r->tag = tag; r->rd2_lo = -1;
    r->rd_lo = 1; r->rd_hi = 0;                       /* empty read range */
    if (filt_query(F, tag)) { r->flags |= F_COMPLEX; return; }   /* an equal tag may exist: match_len, chains */
    if ((tag & ins_mask) != ins_mask) return;         /* lookup only, and nothing to find */
    i64 h0 = tag & c->hmask, h, steps = 0;
    i64 btm = (c->min_mask << 1) | 1;
    int w2_elig = 0, w1_elig = 0;   /* the stop slot holds an entry due for sweeping (it may sit in the sweep queue) */
    r->flags |= F_INSERT;
    r->rd_lo = h0;
    for (h = h0;; h++) {
        if (h == c->hsize || ++steps > MAXWALK) { r->flags |= F_COMPLEX; return; }
        const slot *o = &c->tab[h];
        if (!nonempty(o)) { r->w2_slot = h; r->dcount = 1; break; }
        if (o->tag == tag) { r->flags |= F_COMPLEX; return; }   /* cannot happen with a sound filter */
        if ((o->tag & btm) != btm) { r->w2_slot = h; r->dcount = 0; w2_elig = 1; break; }
        if (tz1(o->tag) < tz1(tag)) {
            /* one level of displacement (rzip.c:275-277) */
            i64 ot = o->tag, g = ot & c->hmask, round = 0, st2 = 0;
            int d2 = 0;
            for (;; g++) {
                if (g == c->hsize || ++st2 > MAXWALK) { r->flags |= F_COMPLEX; return; }
                const slot *q = &c->tab[g];
                if (!nonempty(q)) break;
                if ((q->tag & btm) != btm) { d2 = -1; w1_elig = 1; break; }
                if (tz1(q->tag) < tz1(ot)) { r->flags |= F_COMPLEX; return; }
                if (q->tag == ot && ++round == c->chain) { r->flags |= F_COMPLEX; return; }
            }
            r->flags |= F_W1;
            r->rd2_lo = ot & c->hmask; r->rd2_hi = g;
            r->w1_slot = g; r->w1_off = o->off; r->w1_tag = ot;
            r->w2_slot = h; r->dcount = 1 + d2;
            break;
        }
    }
    r->rd_hi = r->w2_slot;
    if (g_range_rule) {
        /* anything touching the clean window is left to the serial path */
        if (r->rd_hi >= w_lo && r->rd_lo < w_hi) r->flags |= F_COMPLEX;
        if (r->rd2_lo >= 0 && r->rd2_hi >= w_lo && r->rd2_lo < w_hi) r->flags |= F_COMPLEX;
    } else {
        /* An entry due for sweeping stops every walk (its trailing-one count is below every op's need), so a slot of
         * the sweep queue can only ever be a walk's STOP slot: only then does the order of sweeps and inserts matter. */
        if (w2_elig && in_win(r->w2_slot, w_lo, w_hi)) r->flags |= F_COMPLEX;
        if (w1_elig && in_win(r->w1_slot, w_lo, w_hi)) r->flags |= F_COMPLEX;
    }
}

/* "Twin": position p+1 has the same tag as p (buf[p] == buf[p+31]) and both pass the gates.  The second op
 * (B) sees the entry the first (A) has just inserted: its lookup meets exactly that one equal tag (the filter
 * was negative, so no other exists) and match_len(p+1, p) almost always fails (tag_misses++); its insert walk
 * passes A's entry (round = 1 < max_chain_len) and goes on from the next slot.  Planned when A's plan is the
 * simple kind (no displacement); everything else is left to the serial path. */
static void plan_follower(const ctx *c, const rec *A, rec *B, i64 pos, i64 tag, i64 w_lo, i64 w_hi) {
    memset(B, 0, sizeof *B);
    B->pos = pos; B->tag = tag; B->rd2_lo = -1; B->rd_lo = 1; B->rd_hi = 0;
    if (A->flags & F_COMPLEX) { B->flags |= F_COMPLEX; return; }
    if (!(A->flags & F_INSERT)) return;               /* both lookup-only, nothing to find */
    if ((A->flags & F_W1) || c->chain == 1) { B->flags |= F_COMPLEX; return; }
    i64 rev = 0;
    if (match_len(c, pos, A->pos, &rev)) { B->flags |= F_COMPLEX; return; }   /* a genuine run: real match */
    i64 btm = (c->min_mask << 1) | 1, steps = 0, h;
    B->flags |= F_INSERT | F_FOLLOW;
    B->rd_lo = A->rd_lo;
    if ((tag & btm) != btm) {   /* A's own entry is due for sweeping (only before the first overflow): B overwrites it */
        B->w2_slot = A->w2_slot; B->rd_hi = A->w2_slot; B->dcount = 0;
        if (B->rd_hi >= w_lo && B->rd_lo < w_hi) B->flags |= F_COMPLEX;
        return;
    }
    for (h = A->w2_slot + 1;; h++) {
        if (h == c->hsize || ++steps > MAXWALK) { B->flags |= F_COMPLEX; return; }
        const slot *o = &c->tab[h];
        if (!nonempty(o)) { B->dcount = 1; break; }
        if (o->tag == tag) { B->flags |= F_COMPLEX; return; }
        if ((o->tag & btm) != btm) { B->dcount = 0; break; }
        if (tz1(o->tag) < tz1(tag)) { B->flags |= F_COMPLEX; return; }
    }
    B->w2_slot = h; B->rd_hi = h;
    if (g_range_rule) { if (B->rd_hi >= w_lo && B->rd_lo < w_hi) B->flags |= F_COMPLEX; }
    else if (B->dcount == 0 && in_win(h, w_lo, w_hi)) B->flags |= F_COMPLEX;
}

static int in_ranges(const rec *r, i64 s) {
    if (!g_range_rule) return ((r->flags & F_INSERT) && s == r->w2_slot) || ((r->flags & F_W1) && s == r->w1_slot);
    return (s >= r->rd_lo && s <= r->rd_hi) || (r->rd2_lo >= 0 && s >= r->rd2_lo && s <= r->rd2_hi);
}

typedef struct { i64 slot; int idx; } wkey;
static int wcmp(const void *a, const void *b) {
    const wkey *x = a, *y = b;
    if (x->slot != y->slot) return x->slot < y->slot ? -1 : 1;
    return x->idx - y->idx;
}
/* exact-slot rule: some op with index < j writes exactly slot s */
static int slot_hit(const wkey *w, int nw, i64 s, int j) {
    int a = 0, b = nw;
    while (a < b) { int m = (a + b) / 2; if (w[m].slot < s) a = m + 1; else b = m; }
    for (; a < nw && w[a].slot == s; a++)
        if (w[a].idx < j) return 1;
    return 0;
}
static int range_hit(const wkey *w, int nw, i64 lo, i64 hi, int j) {
    int a = 0, b = nw;
    while (a < b) { int m = (a + b) / 2; if (w[m].slot < lo) a = m + 1; else b = m; }
    for (; a < nw && w[a].slot <= hi; a++)
        if (w[a].idx < j) return 1;
    return 0;
}

int orc_rzip_chunk_spec(const u8 *buf, i64 n, const i64 hash_index[256], int level, int chunk_bytes,
                        i64 *victim_round, u8 *s0, i64 s0_cap, i64 *s0_len, u8 *s1, i64 s1_cap, i64 *s1_len,
                        orc_stats *stats, int B, spec_stats *ss) {
    ctx c;
    memset(&c, 0, sizeof c);
    memset(stats, 0, sizeof *stats);
    memset(ss, 0, sizeof *ss);
    c.buf = buf; c.n = n; c.end = n - MIN_MATCH; c.hidx = hash_index;
    i64 hashsize = LEVELS[level].mb * (1048576 / 16);
    int bits = 0;
    while (((i64)1 << bits) < hashsize) bits++;
    c.hsize = (i64)1 << bits; c.hmask = c.hsize - 1;
    c.limit = c.hsize / 3 * 2;
    c.chain = LEVELS[level].chain;
    c.tab = calloc((size_t)c.hsize, sizeof(slot));
    c.victim_round = victim_round; c.st = stats;
    c.s0 = s0; c.s1 = s1; c.s0_cap = s0_cap; c.s1_cap = s1_cap; c.cb = chunk_bytes;
    i64 wlog[256];
    c.wlog = wlog;
    const int exact_rule = getenv("SPEC_RANGE") == NULL;
    g_range_rule = !exact_rule;

    i64 ins_mask = ((i64)1 << LEVELS[level].freq) - 1;
    c.min_mask = ins_mask;
    i64 p = 0;
    struct { i64 p, ofs, len; } cur = {0, 0, 0};
    /* all tags up front (the GPU gets them from tag_scan.cu) */
    i64 npos = c.end > 0 ? c.end + 1 : 0;
    i64 *tags = malloc((size_t)(npos + 1) * sizeof(i64));
    if (npos) {
        tags[0] = full_tag(&c, 0);
        for (i64 q = 1; q < npos; q++) tags[q] = tags[q - 1] ^ hash_index[buf[q - 1]] ^ hash_index[buf[q + MIN_MATCH - 1]];
    }
    u64 *F = calloc((size_t)1 << FILT_BITS, sizeof(u64));
    rec *R = malloc((size_t)B * sizeof(rec));
    wkey *W = malloc((size_t)B * 2 * sizeof(wkey));
    char *conf = malloc((size_t)B);
    /* eligible-slot queue for clean_one (slots in [clean_ptr, e_scan) already examined) */
    i64 ecap = 2 * (i64)B + 8, *E = malloc((size_t)ecap * sizeof(i64)), e_idx = 0, e_cnt = 0, e_scan = 0;

    while (p < c.end) {
        /* ---- refill the eligible queue, fix the clean window */
        i64 btm = (c.min_mask << 1) | 1;
        /* The queue is only meaningful once inserts can no longer create eligible entries, i.e. after
         * the first overflow made ins_mask == btm (rzip.c:583); before that no clean can be planned. */
        i64 w_lo = 0, w_hi = 0; /* clean window [w_lo, w_hi): empty while filling */
        if (ins_mask == btm) {
            if (e_cnt - e_idx < B) {
                memmove(E, E + e_idx, (size_t)(e_cnt - e_idx) * sizeof(i64));
                e_cnt -= e_idx; e_idx = 0;
                if (e_scan < c.clean_ptr) e_scan = c.clean_ptr;
                for (; e_scan < c.hsize && e_cnt < ecap; e_scan++) {
                    const slot *o = &c.tab[e_scan];
                    if (nonempty(o) && (o->tag & btm) != btm) E[e_cnt++] = e_scan;
                }
            }
            w_lo = c.clean_ptr; w_hi = e_scan;
        } else {
            e_idx = e_cnt = 0; e_scan = c.clean_ptr;
        }
        /* ---- batch of the next B candidates */
        int nr = 0;
        i64 q = p + 1;
        for (; q <= c.end && nr < B; q++)
            if ((tags[q] & c.min_mask) == c.min_mask) {
                if (nr > 0 && R[nr - 1].pos == q - 1 && R[nr - 1].tag == tags[q] && !(R[nr - 1].flags & F_FOLLOW) && !getenv("SPEC_NOTWIN"))
                    plan_follower(&c, &R[nr - 1], &R[nr], q, tags[q], w_lo, w_hi);
                else
                    plan_op(&c, F, &R[nr], q, tags[q], ins_mask, w_lo, w_hi);
                nr++;
            }
        i64 scanned_to = q - 1; /* positions <= scanned_to were examined for this batch */
        if (!nr) { p = c.end; break; }
        ss->batches++;
        /* ---- static conflicts: a planned write of an earlier op inside my read ranges */
        int nw = 0;
        for (int j = 0; j < nr; j++) {
            if ((R[j].flags & (F_COMPLEX | F_INSERT)) != F_INSERT) continue;
            W[nw].slot = R[j].w2_slot; W[nw++].idx = j;
            if (R[j].flags & F_W1) { W[nw].slot = R[j].w1_slot; W[nw++].idx = j; }
        }
        qsort(W, (size_t)nw, sizeof(wkey), wcmp);
        for (int j = 0; j < nr; j++) {
            conf[j] = 0;
            if (R[j].flags & F_COMPLEX) { ss->complex_ops++; continue; }
            int before = (R[j].flags & F_FOLLOW) ? j - 1 : j;
            if (exact_rule) {
                /* steady state: an insert only ever raises the trailing-one count of the slot it writes, so a walk that
                 * passed a slot still passes it; only a write to one of my own stop slots changes my plan */
                if (!(R[j].flags & F_INSERT)) continue;
                if (slot_hit(W, nw, R[j].w2_slot, before) || ((R[j].flags & F_W1) && slot_hit(W, nw, R[j].w1_slot, j))) { conf[j] = 1; ss->static_conflicts++; }
                continue;
            }
            if (range_hit(W, nw, R[j].rd_lo, R[j].rd_hi, before) ||
                (R[j].rd2_lo >= 0 && range_hit(W, nw, R[j].rd2_lo, R[j].rd2_hi, j))) { conf[j] = 1; ss->static_conflicts++; }
        }
        /* ---- phase B */
        int pos = 0, ended = 0;
        while (pos < nr && !ended) {
            int k = nr;
            for (int j = pos; j < nr; j++) {
                int emit = cur.len >= MIN_MATCH && R[j].pos >= cur.p + MIN_MATCH;
                if ((R[j].flags & F_COMPLEX) || conf[j] || emit) { if (emit && !(R[j].flags & F_COMPLEX) && !conf[j]) ss->emit_cuts++; k = j; break; }
            }
            /* running hash_count over [pos, k): a clean that changes ins_mask, or an empty queue, cuts */
            i64 cnt = c.count, cleans = 0, avail = e_cnt - e_idx;
            for (int j = pos; j < k; j++) {
                if (!(R[j].flags & F_INSERT)) continue;
                if (cnt + R[j].dcount > c.limit) {
                    if (ins_mask != btm || cleans == avail) { k = j; ss->count_cuts++; break; }
                    cleans++;
                    cnt += R[j].dcount - 1;
                } else
                    cnt += R[j].dcount;
            }
            /* commit [pos, k) */
            for (int j = pos; j < k; j++) {
                stats->lookups++;
                if (!(R[j].flags & F_INSERT)) continue;
                stats->inserts++;
                if (R[j].flags & F_W1) { c.tab[R[j].w1_slot].off = R[j].w1_off; c.tab[R[j].w1_slot].tag = R[j].w1_tag; stats->displaced++; }
                c.tab[R[j].w2_slot].off = R[j].pos; c.tab[R[j].w2_slot].tag = R[j].tag;
                if (R[j].flags & F_FOLLOW) { stats->tag_misses++;
                    }
                filt_set(F, R[j].tag);
                ss->planned_commits++;
            }
            for (i64 e = 0; e < cleans; e++) { c.tab[E[e_idx + e]].off = 0; c.tab[E[e_idx + e]].tag = 0; }
            if (cleans) { e_idx += cleans; c.clean_ptr = E[e_idx - 1]; }
            c.count = cnt;
            if (k > pos) p = R[k - 1].pos;
            if (k == nr) break;
            /* ---- serial op k on the live table (reference semantics) */
            ss->serial_ops++;
            rec *r = &R[k];
            i64 t = r->tag, old_min = c.min_mask, old_ins = ins_mask;
            p = r->pos;
            c.nlog = 0;
            i64 best = 0, bofs = 0, brev = 0;
            stats->lookups++;
            if (filt_query(F, t))   /* a negative filter proves the walk would find nothing */
            for (i64 h = t & c.hmask; nonempty(&c.tab[h]); h = (h + 1) & c.hmask)
                if (c.tab[h].tag == t) {
                    i64 rev = 0, m = match_len(&c, p, c.tab[h].off, &rev);
                    if (m) { if (m > best) { best = m; bofs = c.tab[h].off - rev; brev = rev; } stats->tag_hits++; }
                    else stats->tag_misses++;
                }
            if ((t & ins_mask) == ins_mask) {
                stats->inserts++;
                c.count++;
                insert(&c, t, p);
                filt_set(F, t);
                if (c.count > c.limit) ins_mask = clean_one(&c);
            }
            if (best > cur.len) { cur.p = p - brev; cur.len = best; cur.ofs = bofs; }
            int jumped = 0;
            if ((cur.len >= GREAT || p >= cur.p + MIN_MATCH) && cur.len >= MIN_MATCH) {
                if (c.last < cur.p) put_literal(&c, c.last, cur.p);
                put_match(&c, cur.p, cur.ofs, cur.len);
                c.last = cur.p + cur.len;
                cur.p = p = c.last;
                cur.len = 0;
                jumped = 1;
            }
            /* bring the eligible queue back in step with what the serial op wrote inside the window */
            for (int w = 0; w < c.nlog; w++) {
                i64 sl = wlog[w];
                if (sl < w_lo || sl >= w_hi) continue;
                if (!nonempty(&c.tab[sl])) { /* zeroed by its clean_one: consume the queue up to it */
                    while (e_idx < e_cnt && E[e_idx] <= sl) e_idx++;
                } else { /* an insert overwrote it: no longer eligible (inserted tags satisfy the mask) */
                    for (i64 e = e_idx; e < e_cnt; e++)
                        if (E[e] == sl) { memmove(E + e, E + e + 1, (size_t)(e_cnt - e - 1) * sizeof(i64)); e_cnt--; break; }
                }
            }
            if (c.min_mask != old_min || ins_mask != old_ins) { e_idx = e_cnt = 0; e_scan = c.clean_ptr; } /* new mask: rebuild */
            if (jumped || c.min_mask != old_min || ins_mask != old_ins) { ended = 1; break; }
            if (k + 1 < nr && (R[k + 1].flags & F_FOLLOW)) conf[k + 1] = 1;   /* its plan assumed the leader's planned write */
            for (int j = k + 1; j < nr; j++)
                if (!conf[j] && !(R[j].flags & F_COMPLEX))
                    for (int w = 0; w < c.nlog; w++)
                        if (in_ranges(&R[j], wlog[w])) { conf[j] = 1; ss->extra_conflicts++; break; }
            pos = k + 1;
        }
        if (!ended && p < scanned_to) p = scanned_to; /* whole batch consumed: skip the non-candidates too */
    }
    if (c.last < n) put_literal(&c, c.last, n);
    put_literal(&c, 0, 0);
    uint32_t crc = orc_crc32(0, buf, n);
    u8 cb4[4] = {(u8)(crc >> 24), (u8)(crc >> 16), (u8)(crc >> 8), (u8)crc};
    s0_put(&c, cb4, 4);
    stats->final_min_mask = c.min_mask;
    *s0_len = c.s0_len; *s1_len = c.s1_len;
    free(c.tab); free(tags); free(F); free(R); free(W); free(conf); free(E);
    return c.overflow ? -1 : 0;
}
