/* TEST INFRASTRUCTURE — see rzip_oracle.h.  CPU restatement of src/rzip.c's hash_search. */
#include "rzip_oracle.h"
#include <stdlib.h>
#include <string.h>
#include <stdio.h>
#include <stdlib.h>
static FILE *g_oplog;

typedef int64_t i64;
typedef uint8_t u8;

#define MIN_MATCH 31   /* rzip.c:49 */
#define GREAT 1024     /* rzip.c:48 */

static const struct { i64 mb; int freq; int chain; } LEVELS[10] = { /* rzip.c:65-73 */
    {1, 4, 1}, {2, 4, 2}, {4, 4, 2}, {8, 4, 2}, {16, 4, 3},
    {32, 4, 4}, {32, 2, 6}, {64, 1, 16}, {64, 1, 32}, {64, 1, 128}};

void orc_level(int level, i64 *mb_used, int *initial_freq, int *max_chain_len) {
    *mb_used = LEVELS[level].mb;
    *initial_freq = LEVELS[level].freq;
    *max_chain_len = LEVELS[level].chain;
}

/* glibc random_r TYPE_3 (degree 31, separation 3), srandom(1) state; rzip.c:669-673 */
void orc_hash_index(i64 hash_index[256], i64 skip) {
    enum { DEG = 31, SEP = 3 };
    int32_t r[DEG];
    int i;
    i64 word = 1;
    r[0] = 1;
    for (i = 1; i < DEG; i++) {
        i64 hi = word / 127773, lo = word % 127773;
        word = 16807 * lo - 2836 * hi;
        if (word < 0) word += 2147483647;
        r[i] = (int32_t)word;
    }
    int f = SEP, b = 0;
    i64 total = 310 + skip + 512, k;
    i64 draws[512];
    for (k = 0; k < total; k++) {
        uint32_t v = (uint32_t)r[f] + (uint32_t)r[b];
        r[f] = (int32_t)v;
        if (k >= 310 + skip) draws[k - 310 - skip] = (i64)(v >> 1);
        if (++f >= DEG) f = 0;
        if (++b >= DEG) b = 0;
    }
    for (i = 0; i < 256; i++) hash_index[i] = (draws[2 * i] << 16) ^ draws[2 * i + 1];
}

int orc_chunk_bytes(i64 chunk_size) { /* rzip.c:1006-1008 */
    int bits = 8;
    while (chunk_size >> bits > 0) bits++;
    return bits / 8 + (bits % 8 ? 1 : 0);
}

static uint32_t crc_tab[8][256];
static int crc_ready;
static void crc_init(void) {
    for (uint32_t i = 0; i < 256; i++) {
        uint32_t c = i;
        for (int k = 0; k < 8; k++) c = (c >> 1) ^ (0xEDB88320u & (0u - (c & 1)));
        crc_tab[0][i] = c;
    }
    for (uint32_t i = 0; i < 256; i++)
        for (int t = 1; t < 8; t++) crc_tab[t][i] = (crc_tab[t - 1][i] >> 8) ^ crc_tab[0][crc_tab[t - 1][i] & 255];
    crc_ready = 1;
}
uint32_t orc_crc32(uint32_t crc, const u8 *p, i64 n) {
    if (!crc_ready) crc_init();
    crc = ~crc;
    while (n >= 8) {
        uint32_t a, b;
        memcpy(&a, p, 4);
        memcpy(&b, p + 4, 4);
        a ^= crc;
        crc = crc_tab[7][a & 255] ^ crc_tab[6][(a >> 8) & 255] ^ crc_tab[5][(a >> 16) & 255] ^ crc_tab[4][a >> 24] ^
              crc_tab[3][b & 255] ^ crc_tab[2][(b >> 8) & 255] ^ crc_tab[1][(b >> 16) & 255] ^ crc_tab[0][b >> 24];
        p += 8;
        n -= 8;
    }
    while (n--) crc = (crc >> 8) ^ crc_tab[0][(crc ^ *p++) & 255];
    return ~crc;
}

typedef struct { i64 off; i64 tag; } slot;

typedef struct {
    const u8 *buf;
    i64 n, end;
    const i64 *hidx;
    slot *tab;
    i64 hmask;       /* 2^bits - 1 */
    i64 hsize;
    i64 limit;
    int chain;
    i64 min_mask, clean_ptr, count, last;
    i64 *victim_round;
    orc_stats *st;
    u8 *s0, *s1;
    i64 s0_cap, s1_cap, s0_len, s1_len;
    int cb;
    int overflow;
    orc_token *tok;
    i64 tok_cap, ntok;
    i64 *wlog;       /* optional log of table slots written (used by spec_model.c only) */
    int nlog;
} ctx;

static inline int tz1(i64 t) { /* trailing one bits; ffsll(~t)-1 with ffsll(0)=0 -> -1 */
    uint64_t x = ~(uint64_t)t;
    return x ? __builtin_ctzll(x) : -1;
}

static inline i64 full_tag(const ctx *c, i64 p) { /* rzip.c:348-358 */
    i64 t = 0;
    for (int i = 0; i < MIN_MATCH; i++) t ^= c->hidx[c->buf[p + i]];
    return t;
}

/* rzip.c:372-397 */
static i64 match_len(const ctx *c, i64 p0, i64 op, i64 *rev) {
    if (op >= p0) return 0;
    const u8 *b = c->buf;
    i64 p = p0;
    while (p < c->end && b[p] == b[op]) { p++; op++; }
    i64 len = p - p0;
    p = p0;
    op -= len;
    i64 lo = c->last > 0 ? c->last : 0;
    while (p > lo && op > 0 && b[op - 1] == b[p - 1]) { op--; p--; }
    *rev = p0 - p;
    len += *rev;
    return len < MIN_MATCH ? 0 : len;
}

/* rzip.c:256-301 */
static void insert(ctx *c, i64 t, i64 off) {
    i64 h = t & c->hmask, victim_h = 0, round = 0;
    i64 btm = (c->min_mask << 1) | 1;
    while (c->tab[h].off | c->tab[h].tag) {
        slot *o = &c->tab[h];
        if ((o->tag & btm) != btm) { c->count--; c->st->minbit_overwrites++; break; }
        if (tz1(o->tag) < tz1(t)) { c->st->displaced++; insert(c, o->tag, o->off); break; }
        if (o->tag == t) {
            if (round == *c->victim_round) victim_h = h;
            if (++round == c->chain) {
                h = victim_h;
                c->count--;
                c->st->victim_events++;
                if (++*c->victim_round == c->chain) *c->victim_round = 0;
                break;
            }
        }
        h = (h + 1) & c->hmask;
    }
    c->tab[h].tag = t;
    c->tab[h].off = off;
    if (c->wlog) c->wlog[c->nlog++] = h;
}

/* rzip.c:305-328 */
static i64 clean_one(ctx *c) {
    for (;;) {
        i64 btm = (c->min_mask << 1) | 1;
        for (; c->clean_ptr < c->hsize; c->clean_ptr++) {
            slot *o = &c->tab[c->clean_ptr];
            if (!(o->off | o->tag)) continue;
            if ((o->tag & btm) != btm) {
                o->off = 0; o->tag = 0;
                c->count--;
                if (c->wlog) c->wlog[c->nlog++] = c->clean_ptr;
                return btm;
            }
        }
        c->min_mask = btm;
        c->clean_ptr = 0;
        c->st->sweeps_wrapped++;
    }
}

static void s0_put(ctx *c, const u8 *p, i64 n) {
    if (c->s0_len + n > c->s0_cap) { c->overflow = 1; return; }
    memcpy(c->s0 + c->s0_len, p, n);
    c->s0_len += n;
}
static void s0_header(ctx *c, u8 head, i64 len) { /* rzip.c:176-179 */
    u8 h[3] = {head, (u8)(len & 255), (u8)((len >> 8) & 255)};
    s0_put(c, h, 3);
}
static void record(ctx *c, i64 pos, i64 len, i64 ofs) {
    if (!c->tok) return;
    if (c->ntok >= c->tok_cap) { c->overflow = 1; return; }
    c->tok[c->ntok].pos = pos; c->tok[c->ntok].len = len; c->tok[c->ntok].ofs = ofs;
    c->ntok++;
}
/* rzip.c:212-227 */
static void put_literal(ctx *c, i64 last, i64 p) {
    do {
        i64 len = p - last;
        if (len > 0xFFFF) len = 0xFFFF;
        c->st->literals++;
        c->st->literal_bytes += len;
        s0_header(c, 0, len);
        if (len) {
            if (c->s1_len + len > c->s1_cap) c->overflow = 1;
            else { memcpy(c->s1 + c->s1_len, c->buf + last, len); c->s1_len += len; }
        }
        last += len;
    } while (p > last);
}
/* rzip.c:181-195 */
static void put_match(ctx *c, i64 p, i64 offset, i64 len) {
    do {
        i64 n = len > 0xFFFF ? 0xFFFF : len;
        i64 ofs = p - offset;
        u8 d[8];
        for (int i = 0; i < 8; i++) d[i] = (u8)((uint64_t)ofs >> (8 * i));
        s0_header(c, 1, n);
        s0_put(c, d, c->cb);
        c->st->matches++;
        c->st->match_bytes += n;
        len -= n; p += n; offset += n;
    } while (len);
}

int orc_rzip_chunk(const u8 *buf, i64 n, const i64 hash_index[256], int level, int chunk_bytes,
                   i64 *victim_round, u8 *s0, i64 s0_cap, i64 *s0_len, u8 *s1, i64 s1_cap, i64 *s1_len,
                   orc_token *tokens, i64 tok_cap, i64 *ntok, orc_stats *stats) {
    ctx c;
    memset(&c, 0, sizeof c);
    memset(stats, 0, sizeof *stats);
    c.buf = buf; c.n = n; c.end = n - MIN_MATCH; c.hidx = hash_index;
    i64 hashsize = LEVELS[level].mb * (1048576 / 16);
    int bits = 0;
    while (((i64)1 << bits) < hashsize) bits++;        /* rzip.c:521-523 */
    c.hsize = (i64)1 << bits; c.hmask = c.hsize - 1;
    c.limit = c.hsize / 3 * 2;                         /* rzip.c:529 */
    c.chain = LEVELS[level].chain;
    c.tab = calloc((size_t)c.hsize, sizeof(slot));
    if (!c.tab) return -1;
    c.victim_round = victim_round; c.st = stats;
    c.s0 = s0; c.s1 = s1; c.s0_cap = s0_cap; c.s1_cap = s1_cap; c.cb = chunk_bytes;
    c.tok = tokens; c.tok_cap = tok_cap;

    i64 ins_mask = ((i64)1 << LEVELS[level].freq) - 1;
    c.min_mask = ins_mask;
    i64 p = 0, t = 0;
    struct { i64 p, ofs, len; } cur = {0, 0, 0};
    if (c.end > 0) t = full_tag(&c, 0);
    g_oplog = getenv("ORC_OPLOG") ? fopen(getenv("ORC_OPLOG"), "wb") : NULL;

    while (p < c.end) {                                /* rzip.c:548-615 */
        ++p;
        t ^= hash_index[buf[p - 1]] ^ hash_index[buf[p + MIN_MATCH - 1]];
        if ((t & c.min_mask) != c.min_mask) continue;

        i64 best = 0, bofs = 0, brev = 0;              /* find_best_match rzip.c:426-462 */
        stats->lookups++;
        if (g_oplog) fwrite(&p, 8, 1, g_oplog);      /* debugging aid: ORC_OPLOG=<file> records the position of every lookup */
        for (i64 h = t & c.hmask; c.tab[h].off | c.tab[h].tag; h = (h + 1) & c.hmask) {
            stats->slots_scanned++;
            if (c.tab[h].tag == t) {
                i64 rev = 0, m = match_len(&c, p, c.tab[h].off, &rev);
                if (m) {
                    if (m > best) { best = m; bofs = c.tab[h].off - rev; brev = rev; }
                    stats->tag_hits++;
                } else stats->tag_misses++;
            }
        }
        if ((t & ins_mask) == ins_mask) {
            stats->inserts++;
            c.count++;
            insert(&c, t, p);
            if (c.count > c.limit) ins_mask = clean_one(&c);
        }
        if (best > cur.len) { cur.p = p - brev; cur.len = best; cur.ofs = bofs; }
        if ((cur.len >= GREAT || p >= cur.p + MIN_MATCH) && cur.len >= MIN_MATCH) {
            if (c.last < cur.p) { record(&c, c.last, cur.p - c.last, -1); put_literal(&c, c.last, cur.p); }
            record(&c, cur.p, cur.len, cur.ofs);
            put_match(&c, cur.p, cur.ofs, cur.len);
            c.last = cur.p + cur.len;
            cur.p = p = c.last;
            cur.len = 0;
            t = full_tag(&c, p);
        }
    }
    if (c.last < n) { record(&c, c.last, n - c.last, -1); put_literal(&c, c.last, n); }
    put_literal(&c, 0, 0);                             /* terminator rzip.c:664; counts as a literal */
    uint32_t crc = orc_crc32(0, buf, n);
    u8 cb4[4] = {(u8)(crc >> 24), (u8)(crc >> 16), (u8)(crc >> 8), (u8)crc};  /* rzip.c:662-665 */
    s0_put(&c, cb4, 4);
    stats->final_min_mask = c.min_mask;
    *s0_len = c.s0_len; *s1_len = c.s1_len;
    if (ntok) *ntok = c.ntok;
    free(c.tab);
    return c.overflow ? -1 : 0;
}
