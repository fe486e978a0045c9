/* md5.c — MD5 (RFC 1321) for the whole-input hash trailer.  The reference takes it from libgcrypt
 * (gcry_md_open/write/read with GCRY_MD_MD5: src/rzip.c:495, :829, :1069-1091); this is an
 * independent implementation of the public algorithm, checked against hashlib in the tests. */
#include <string.h>

#include "mrzip_host.h"

static const uint32_t K[64] = {
    0xd76aa478, 0xe8c7b756, 0x242070db, 0xc1bdceee, 0xf57c0faf, 0x4787c62a, 0xa8304613, 0xfd469501,
    0x698098d8, 0x8b44f7af, 0xffff5bb1, 0x895cd7be, 0x6b901122, 0xfd987193, 0xa679438e, 0x49b40821,
    0xf61e2562, 0xc040b340, 0x265e5a51, 0xe9b6c7aa, 0xd62f105d, 0x02441453, 0xd8a1e681, 0xe7d3fbc8,
    0x21e1cde6, 0xc33707d6, 0xf4d50d87, 0x455a14ed, 0xa9e3e905, 0xfcefa3f8, 0x676f02d9, 0x8d2a4c8a,
    0xfffa3942, 0x8771f681, 0x6d9d6122, 0xfde5380c, 0xa4beea44, 0x4bdecfa9, 0xf6bb4b60, 0xbebfbc70,
    0x289b7ec6, 0xeaa127fa, 0xd4ef3085, 0x04881d05, 0xd9d4d039, 0xe6db99e5, 0x1fa27cf8, 0xc4ac5665,
    0xf4292244, 0x432aff97, 0xab9423a7, 0xfc93a039, 0x655b59c3, 0x8f0ccc92, 0xffeff47d, 0x85845dd1,
    0x6fa87e4f, 0xfe2ce6e0, 0xa3014314, 0x4e0811a1, 0xf7537e82, 0xbd3af235, 0x2ad7d2bb, 0xeb86d391};
static const unsigned char S[64] = {7, 12, 17, 22, 7, 12, 17, 22, 7, 12, 17, 22, 7, 12, 17, 22,
                                    5, 9,  14, 20, 5, 9,  14, 20, 5, 9,  14, 20, 5, 9,  14, 20,
                                    4, 11, 16, 23, 4, 11, 16, 23, 4, 11, 16, 23, 4, 11, 16, 23,
                                    6, 10, 15, 21, 6, 10, 15, 21, 6, 10, 15, 21, 6, 10, 15, 21};

#define ROL(x, s) (((x) << (s)) | ((x) >> (32 - (s))))

static void md5_block(uint32_t h[4], const uchar *p) {
    uint32_t m[16];
    memcpy(m, p, 64); /* little-endian host */
    uint32_t a = h[0], b = h[1], c = h[2], d = h[3];
#define STEP(f, g, i)                                   \
    do {                                                \
        uint32_t t_ = a + (f) + K[i] + m[g];            \
        a = d; d = c; c = b;                            \
        b = b + ROL(t_, S[i]);                          \
    } while (0)
    for (int i = 0; i < 16; i++) STEP((b & c) | (~b & d), i, i);
    for (int i = 16; i < 32; i++) STEP((d & b) | (~d & c), (5 * i + 1) & 15, i);
    for (int i = 32; i < 48; i++) STEP(b ^ c ^ d, (3 * i + 5) & 15, i);
    for (int i = 48; i < 64; i++) STEP(c ^ (b | ~d), (7 * i) & 15, i);
#undef STEP
    h[0] += a; h[1] += b; h[2] += c; h[3] += d;
}

void md5_init(md5_ctx *c) {
    c->h[0] = 0x67452301; c->h[1] = 0xefcdab89; c->h[2] = 0x98badcfe; c->h[3] = 0x10325476;
    c->len = 0;
    c->fill = 0;
}

void md5_update(md5_ctx *c, const void *data, size_t len) {
    const uchar *p = data;
    c->len += len;
    if (c->fill) {
        size_t n = 64 - (size_t)c->fill;
        if (n > len) n = len;
        memcpy(c->blk + c->fill, p, n);
        c->fill += (int)n; p += n; len -= n;
        if (c->fill == 64) { md5_block(c->h, c->blk); c->fill = 0; }
    }
    while (len >= 64) { md5_block(c->h, p); p += 64; len -= 64; }
    if (len) { memcpy(c->blk, p, len); c->fill = (int)len; }
}

void md5_final(md5_ctx *c, uchar out[16]) {
    uint64_t bits = c->len * 8;
    uchar pad[72] = {0x80};
    size_t padlen = (c->fill < 56) ? (size_t)(56 - c->fill) : (size_t)(120 - c->fill);
    md5_update(c, pad, padlen);
    uchar lenb[8];
    for (int i = 0; i < 8; i++) lenb[i] = (uchar)(bits >> (8 * i));
    md5_update(c, lenb, 8);
    memcpy(out, c->h, 16);
}
