/* md5.c — MD5 (RFC 1321) for the whole-input hash trailer.  The reference takes it from libgcrypt
 * (gcry_md_open/write/read with GCRY_MD_MD5: src/rzip.c:495, :829, :1069-1091); this is an
 * independent implementation of the public algorithm, checked against hashlib in the tests. */
#include <string.h>

#include "mrzip_host.h"

#define ROL(x, s) (((x) << (s)) | ((x) >> (32 - (s))))

/* One 64-byte block, fully unrolled: constant rotations and no register shuffling (the rolled form with table-driven
 * rotations ran at 0.29 GB/s; this one at more than twice that, and the whole-file MD5 thread bounds `mrzip -n`). */
static void md5_block(uint32_t h[4], const uchar *p) {
    uint32_t m[16];
    memcpy(m, p, 64); /* little-endian host */
    uint32_t a = h[0], b = h[1], c = h[2], d = h[3];
    a += (d ^ (b & (c ^ d))) + 0xd76aa478u + m[0]; a = b + ROL(a, 7);
    d += (c ^ (a & (b ^ c))) + 0xe8c7b756u + m[1]; d = a + ROL(d, 12);
    c += (b ^ (d & (a ^ b))) + 0x242070dbu + m[2]; c = d + ROL(c, 17);
    b += (a ^ (c & (d ^ a))) + 0xc1bdceeeu + m[3]; b = c + ROL(b, 22);
    a += (d ^ (b & (c ^ d))) + 0xf57c0fafu + m[4]; a = b + ROL(a, 7);
    d += (c ^ (a & (b ^ c))) + 0x4787c62au + m[5]; d = a + ROL(d, 12);
    c += (b ^ (d & (a ^ b))) + 0xa8304613u + m[6]; c = d + ROL(c, 17);
    b += (a ^ (c & (d ^ a))) + 0xfd469501u + m[7]; b = c + ROL(b, 22);
    a += (d ^ (b & (c ^ d))) + 0x698098d8u + m[8]; a = b + ROL(a, 7);
    d += (c ^ (a & (b ^ c))) + 0x8b44f7afu + m[9]; d = a + ROL(d, 12);
    c += (b ^ (d & (a ^ b))) + 0xffff5bb1u + m[10]; c = d + ROL(c, 17);
    b += (a ^ (c & (d ^ a))) + 0x895cd7beu + m[11]; b = c + ROL(b, 22);
    a += (d ^ (b & (c ^ d))) + 0x6b901122u + m[12]; a = b + ROL(a, 7);
    d += (c ^ (a & (b ^ c))) + 0xfd987193u + m[13]; d = a + ROL(d, 12);
    c += (b ^ (d & (a ^ b))) + 0xa679438eu + m[14]; c = d + ROL(c, 17);
    b += (a ^ (c & (d ^ a))) + 0x49b40821u + m[15]; b = c + ROL(b, 22);
    a += (c ^ (d & (b ^ c))) + 0xf61e2562u + m[1]; a = b + ROL(a, 5);
    d += (b ^ (c & (a ^ b))) + 0xc040b340u + m[6]; d = a + ROL(d, 9);
    c += (a ^ (b & (d ^ a))) + 0x265e5a51u + m[11]; c = d + ROL(c, 14);
    b += (d ^ (a & (c ^ d))) + 0xe9b6c7aau + m[0]; b = c + ROL(b, 20);
    a += (c ^ (d & (b ^ c))) + 0xd62f105du + m[5]; a = b + ROL(a, 5);
    d += (b ^ (c & (a ^ b))) + 0x02441453u + m[10]; d = a + ROL(d, 9);
    c += (a ^ (b & (d ^ a))) + 0xd8a1e681u + m[15]; c = d + ROL(c, 14);
    b += (d ^ (a & (c ^ d))) + 0xe7d3fbc8u + m[4]; b = c + ROL(b, 20);
    a += (c ^ (d & (b ^ c))) + 0x21e1cde6u + m[9]; a = b + ROL(a, 5);
    d += (b ^ (c & (a ^ b))) + 0xc33707d6u + m[14]; d = a + ROL(d, 9);
    c += (a ^ (b & (d ^ a))) + 0xf4d50d87u + m[3]; c = d + ROL(c, 14);
    b += (d ^ (a & (c ^ d))) + 0x455a14edu + m[8]; b = c + ROL(b, 20);
    a += (c ^ (d & (b ^ c))) + 0xa9e3e905u + m[13]; a = b + ROL(a, 5);
    d += (b ^ (c & (a ^ b))) + 0xfcefa3f8u + m[2]; d = a + ROL(d, 9);
    c += (a ^ (b & (d ^ a))) + 0x676f02d9u + m[7]; c = d + ROL(c, 14);
    b += (d ^ (a & (c ^ d))) + 0x8d2a4c8au + m[12]; b = c + ROL(b, 20);
    a += (b ^ c ^ d) + 0xfffa3942u + m[5]; a = b + ROL(a, 4);
    d += (a ^ b ^ c) + 0x8771f681u + m[8]; d = a + ROL(d, 11);
    c += (d ^ a ^ b) + 0x6d9d6122u + m[11]; c = d + ROL(c, 16);
    b += (c ^ d ^ a) + 0xfde5380cu + m[14]; b = c + ROL(b, 23);
    a += (b ^ c ^ d) + 0xa4beea44u + m[1]; a = b + ROL(a, 4);
    d += (a ^ b ^ c) + 0x4bdecfa9u + m[4]; d = a + ROL(d, 11);
    c += (d ^ a ^ b) + 0xf6bb4b60u + m[7]; c = d + ROL(c, 16);
    b += (c ^ d ^ a) + 0xbebfbc70u + m[10]; b = c + ROL(b, 23);
    a += (b ^ c ^ d) + 0x289b7ec6u + m[13]; a = b + ROL(a, 4);
    d += (a ^ b ^ c) + 0xeaa127fau + m[0]; d = a + ROL(d, 11);
    c += (d ^ a ^ b) + 0xd4ef3085u + m[3]; c = d + ROL(c, 16);
    b += (c ^ d ^ a) + 0x04881d05u + m[6]; b = c + ROL(b, 23);
    a += (b ^ c ^ d) + 0xd9d4d039u + m[9]; a = b + ROL(a, 4);
    d += (a ^ b ^ c) + 0xe6db99e5u + m[12]; d = a + ROL(d, 11);
    c += (d ^ a ^ b) + 0x1fa27cf8u + m[15]; c = d + ROL(c, 16);
    b += (c ^ d ^ a) + 0xc4ac5665u + m[2]; b = c + ROL(b, 23);
    a += (c ^ (b | ~d)) + 0xf4292244u + m[0]; a = b + ROL(a, 6);
    d += (b ^ (a | ~c)) + 0x432aff97u + m[7]; d = a + ROL(d, 10);
    c += (a ^ (d | ~b)) + 0xab9423a7u + m[14]; c = d + ROL(c, 15);
    b += (d ^ (c | ~a)) + 0xfc93a039u + m[5]; b = c + ROL(b, 21);
    a += (c ^ (b | ~d)) + 0x655b59c3u + m[12]; a = b + ROL(a, 6);
    d += (b ^ (a | ~c)) + 0x8f0ccc92u + m[3]; d = a + ROL(d, 10);
    c += (a ^ (d | ~b)) + 0xffeff47du + m[10]; c = d + ROL(c, 15);
    b += (d ^ (c | ~a)) + 0x85845dd1u + m[1]; b = c + ROL(b, 21);
    a += (c ^ (b | ~d)) + 0x6fa87e4fu + m[8]; a = b + ROL(a, 6);
    d += (b ^ (a | ~c)) + 0xfe2ce6e0u + m[15]; d = a + ROL(d, 10);
    c += (a ^ (d | ~b)) + 0xa3014314u + m[6]; c = d + ROL(c, 15);
    b += (d ^ (c | ~a)) + 0x4e0811a1u + m[13]; b = c + ROL(b, 21);
    a += (c ^ (b | ~d)) + 0xf7537e82u + m[4]; a = b + ROL(a, 6);
    d += (b ^ (a | ~c)) + 0xbd3af235u + m[11]; d = a + ROL(d, 10);
    c += (a ^ (d | ~b)) + 0x2ad7d2bbu + m[2]; c = d + ROL(c, 15);
    b += (d ^ (c | ~a)) + 0xeb86d391u + m[9]; b = c + ROL(b, 21);
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
