// emit.cu — parallel emission of the two rzip streams and the chunk CRC32.
//
// Replaces put_literal / put_match / put_header / put_vchars / write_sbstream (rzip.c:161-227),
// the tail of hash_search (rzip.c:617-666) and the per-page checksum feed (rzip.c:488-505,
// 601-661).  The automaton leaves un-split match triples (pos, len, ofs); literals are the gaps.
// Item i (0 <= i <= m) = [literal before match i][match i]; item m = final literal + terminator
// + CRC.  Sizes -> exclusive scans -> every item writes its stream-0 bytes at its own offset and
// stream 1 is gathered output-centrically (16 output bytes per thread) at HBM speed.
#include <stdio.h>
#include "mrz_common.cuh"

#define PIECE 0xFFFF  // rzip.c:185, :216

__device__ __forceinline__ void item_geometry(const Token *toks, i64 m, i64 n, i64 i, i64 &lit_start, i64 &litlen) {
    lit_start = i == 0 ? 0 : toks[i - 1].pos + toks[i - 1].len;
    i64 lit_end = i < m ? toks[i].pos : n;
    litlen = lit_end - lit_start;
}

__global__ void k_item_sizes(const Token *__restrict__ toks, i64 m, i64 n, int cb, i64 *__restrict__ a0,
                             i64 *__restrict__ a1, EmitTotals *tot) {
    i64 i = (i64)blockIdx.x * blockDim.x + threadIdx.x;
    i64 lp = 0, mp = 0, mb = 0;
    if (i <= m) {
        i64 ls, ll;
        item_geometry(toks, m, n, i, ls, ll);
        lp = (ll + PIECE - 1) / PIECE;
        i64 s0 = 3 * lp;
        if (i < m) {
            mb = toks[i].len;
            mp = (mb + PIECE - 1) / PIECE;
            s0 += mp * (3 + cb);
        } else {
            s0 += 3 + 4;
            lp += 1;  // the terminator is counted by put_literal (rzip.c:219, :664)
        }
        a0[i] = s0;
        a1[i] = ll;
    }
    // block-reduce the three counters
    for (int o = 16; o; o >>= 1) {
        lp += __shfl_xor_sync(0xffffffffu, lp, o);
        mp += __shfl_xor_sync(0xffffffffu, mp, o);
        mb += __shfl_xor_sync(0xffffffffu, mb, o);
    }
    if ((threadIdx.x & 31) == 0) {
        if (lp) atomicAdd((u64 *)&tot->literals, (u64)lp);
        if (mp) atomicAdd((u64 *)&tot->matches, (u64)mp);
        if (mb) atomicAdd((u64 *)&tot->match_bytes, (u64)mb);
    }
}

// ---- exclusive scan of i64, three phases -------------------------------------------------------
#define SCAN_B 1024
#define SCAN_PER 4
#define SCAN_BLK (SCAN_B * SCAN_PER)

__device__ __forceinline__ i64 block_excl_scan(i64 v, i64 *wsum, i64 &total) {
    int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    i64 inc = v;
    for (int o = 1; o < 32; o <<= 1) {
        i64 u = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += u;
    }
    if (lane == 31) wsum[wid] = inc;
    __syncthreads();
    if (wid == 0) {
        i64 w = wsum[lane], wi = w;
        for (int o = 1; o < 32; o <<= 1) {
            i64 u = __shfl_up_sync(0xffffffffu, wi, o);
            if (lane >= o) wi += u;
        }
        wsum[lane] = wi - w;
        if (lane == 31) wsum[32] = wi;
    }
    __syncthreads();
    total = wsum[32];
    i64 r = wsum[wid] + inc - v;
    __syncthreads();
    return r;
}

__global__ void __launch_bounds__(SCAN_B) k_scan_local(const i64 *__restrict__ in, i64 *__restrict__ out, i64 cnt,
                                                       i64 *__restrict__ bsum) {
    __shared__ i64 wsum[33];
    i64 base = (i64)blockIdx.x * SCAN_BLK + (i64)threadIdx.x * SCAN_PER;
    i64 v[SCAN_PER], s = 0;
#pragma unroll
    for (int k = 0; k < SCAN_PER; k++) {
        v[k] = (base + k < cnt) ? in[base + k] : 0;
        s += v[k];
    }
    i64 total;
    i64 ex = block_excl_scan(s, wsum, total);
#pragma unroll
    for (int k = 0; k < SCAN_PER; k++) {
        if (base + k < cnt) out[base + k] = ex;
        ex += v[k];
    }
    if (threadIdx.x == 0) bsum[blockIdx.x] = total;
}

// single CTA: exclusive scan of the block sums in place; grand total -> *total_out
__global__ void __launch_bounds__(SCAN_B) k_scan_bsums(i64 *bsum, i64 nb, i64 *total_out) {
    __shared__ i64 wsum[33];
    i64 carry = 0;
    for (i64 base = 0; base < nb; base += SCAN_B) {
        i64 idx = base + threadIdx.x;
        i64 v = idx < nb ? bsum[idx] : 0;
        i64 total;
        i64 ex = block_excl_scan(v, wsum, total);
        if (idx < nb) bsum[idx] = carry + ex;
        carry += total;
    }
    if (threadIdx.x == 0) *total_out = carry;
}

__global__ void __launch_bounds__(SCAN_B) k_scan_add(i64 *__restrict__ out, i64 cnt, const i64 *__restrict__ bsum) {
    i64 base = (i64)blockIdx.x * SCAN_BLK + (i64)threadIdx.x * SCAN_PER;
    i64 add = bsum[blockIdx.x];
#pragma unroll
    for (int k = 0; k < SCAN_PER; k++)
        if (base + k < cnt) out[base + k] += add;
}

// ---- stream 0: one warp per item ---------------------------------------------------------------
__global__ void k_emit_s0(const Token *__restrict__ toks, i64 m, i64 n, int cb, const i64 *__restrict__ off0,
                          u8 *__restrict__ s0, const u32 *__restrict__ crc) {
    i64 i = ((i64)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    int lane = threadIdx.x & 31;
    if (i > m) return;
    i64 ls, ll;
    item_geometry(toks, m, n, i, ls, ll);
    u8 *o = s0 + off0[i];
    i64 lp = (ll + PIECE - 1) / PIECE;
    for (i64 q = lane; q < lp; q += 32) {  // literal headers: 00 len16 (rzip.c:176-179, :221)
        i64 len = (q == lp - 1) ? ll - q * PIECE : PIECE;
        u8 *w = o + 3 * q;
        w[0] = 0; w[1] = (u8)len; w[2] = (u8)(len >> 8);
    }
    o += 3 * lp;
    if (i < m) {  // match pieces: 01 len16 dist[cb], same dist in every piece (rzip.c:181-195)
        i64 len = toks[i].len, dist = toks[i].pos - toks[i].ofs;
        i64 mp = (len + PIECE - 1) / PIECE;
        int w8 = 3 + cb;
        for (i64 q = lane; q < mp; q += 32) {
            i64 l = (q == mp - 1) ? len - q * PIECE : PIECE;
            u8 *w = o + w8 * q;
            w[0] = 1; w[1] = (u8)l; w[2] = (u8)(l >> 8);
            for (int b = 0; b < cb; b++) w[3 + b] = (u8)((u64)dist >> (8 * b));
        }
    } else if (lane == 0) {  // terminator + CRC, big-endian (rzip.c:662-665)
        u32 c = *crc;
        o[0] = 0; o[1] = 0; o[2] = 0;
        o[3] = (u8)(c >> 24); o[4] = (u8)(c >> 16); o[5] = (u8)(c >> 8); o[6] = (u8)c;
    }
}

// ---- stream 1: 16 output bytes per thread ------------------------------------------------------
__global__ void __launch_bounds__(256) k_emit_s1(const u8 *__restrict__ buf, const Token *__restrict__ toks, i64 m,
                                                 i64 n, const i64 *__restrict__ off1, i64 s1_len,
                                                 u8 *__restrict__ s1) {
    i64 o = ((i64)blockIdx.x * blockDim.x + threadIdx.x) * 16;
    if (o >= s1_len) return;
    // item i with off1[i] <= o < off1[i+1]  (off1 has m+2 entries; off1[m+1] = s1_len)
    i64 lo = 0, hi = m + 1;
    while (hi - lo > 1) {
        i64 mid = (lo + hi) >> 1;
        if (off1[mid] <= o) lo = mid; else hi = mid;
    }
    i64 i = lo;
    i64 ls, ll;
    item_geometry(toks, m, n, i, ls, ll);
    i64 within = o - off1[i];
    i64 left = s1_len - o;
    if (ll - within >= 16 && left >= 16) {
        uint4 v;
        u64 a = ld_u64_unaligned(buf, ls + within), b = ld_u64_unaligned(buf, ls + within + 8);
        v.x = (u32)a; v.y = (u32)(a >> 32); v.z = (u32)b; v.w = (u32)(b >> 32);
        *reinterpret_cast<uint4 *>(s1 + o) = v;
        return;
    }
    int todo = left < 16 ? (int)left : 16;
    for (int k = 0; k < todo; k++) {
        while (within >= ll) {  // advance to the next non-empty literal
            i++;
            item_geometry(toks, m, n, i, ls, ll);
            within = 0;
        }
        s1[o + k] = buf[ls + within];
        within++;
    }
}

// ---- CRC32 (IEEE, reflected) -------------------------------------------------------------------
// level 1: one warp per 16 KiB block, 512 B per lane with slicing-by-4, combined in a 5-step tree
// using shift matrices; raw CRCs (init 0, no final xor) so that blocks combine linearly.
#define CRC_SUB 512
#define CRC_BLK (32 * CRC_SUB)

struct CrcConst {
    u32 tab[4][256];
    u32 shift_sub[5][32];   // multiply by x^(8*512*2^j)
    u32 shift_blk[5][32];   // multiply by x^(8*16384*2^j)
};
__constant__ CrcConst c_crc;

__device__ __forceinline__ u32 gf2_apply(const u32 *mat, u32 v) {
    u32 r = 0;
#pragma unroll
    for (int b = 0; b < 32; b++) r ^= mat[b] & (0u - ((v >> b) & 1u));
    return r;
}

__global__ void __launch_bounds__(128) k_crc_blocks(const u8 *__restrict__ buf, i64 nblk, u32 *__restrict__ part) {
    __shared__ u32 tab[4][256];
    for (int i = threadIdx.x; i < 1024; i += blockDim.x) tab[i >> 8][i & 255] = c_crc.tab[i >> 8][i & 255];
    __syncthreads();
    i64 blk = ((i64)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    int lane = threadIdx.x & 31;
    if (blk >= nblk) return;
    const uint4 *src = reinterpret_cast<const uint4 *>(buf + blk * CRC_BLK + lane * CRC_SUB);
    u32 crc = 0;
#pragma unroll 2
    for (int v = 0; v < CRC_SUB / 16; v++) {
        uint4 x = __ldg(src + v);
        u32 wds[4] = {x.x, x.y, x.z, x.w};
#pragma unroll
        for (int k = 0; k < 4; k++) {
            u32 y = crc ^ wds[k];
            crc = tab[3][y & 255] ^ tab[2][(y >> 8) & 255] ^ tab[1][(y >> 16) & 255] ^ tab[0][y >> 24];
        }
    }
    // tree: at step j, lanes whose bit j is 0 absorb the partner's CRC: A||B = shift_len(B)(A) ^ B
#pragma unroll
    for (int j = 0; j < 5; j++) {
        u32 other = __shfl_down_sync(0xffffffffu, crc, 1 << j);
        if ((lane & ((2 << j) - 1)) == 0) crc = gf2_apply(c_crc.shift_sub[j], crc) ^ other;
    }
    if (lane == 0) part[blk] = crc;
}

// level 2: one warp combines 32 consecutive level-1 partials (512 KiB)
__global__ void __launch_bounds__(128) k_crc_level2(const u32 *__restrict__ part, i64 ngrp, u32 *__restrict__ part2) {
    i64 g = ((i64)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    int lane = threadIdx.x & 31;
    if (g >= ngrp) return;
    u32 crc = part[g * 32 + lane];
#pragma unroll
    for (int j = 0; j < 5; j++) {
        u32 other = __shfl_down_sync(0xffffffffu, crc, 1 << j);
        if ((lane & ((2 << j) - 1)) == 0) crc = gf2_apply(c_crc.shift_blk[j], crc) ^ other;
    }
    if (lane == 0) part2[g] = crc;
}

// tail (< 16 KiB): one thread, bytewise
__global__ void k_crc_tail(const u8 *__restrict__ buf, i64 start, i64 len, u32 *out) {
    u32 crc = 0;
    for (i64 i = 0; i < len; i++) crc = (crc >> 8) ^ c_crc.tab[0][(crc ^ buf[start + i]) & 255];
    *out = crc;
}

// ---- launchers ---------------------------------------------------------------------------------
extern "C" void mrz_crc_upload_consts(const void *host_consts, cudaStream_t s) {
#ifdef MRZ_CUSIM
    (void)s;
    cudaMemcpyToSymbol(c_crc, host_consts, sizeof(CrcConst));
#else
    cudaMemcpyToSymbolAsync(c_crc, host_consts, sizeof(CrcConst), 0, cudaMemcpyHostToDevice, s);
#endif
}

extern "C" void mrz_launch_crc(const u8 *buf, i64 n, u32 *part, u32 *part2, u32 *tail, cudaStream_t s, int *launches) {
    i64 nblk = n / CRC_BLK, ngrp = nblk / 32;
    if (nblk) { MRZ_LAUNCH(k_crc_blocks, (unsigned)((nblk + 3) / 4), 128, s, buf, nblk, part); (*launches)++; }
    if (ngrp) { MRZ_LAUNCH(k_crc_level2, (unsigned)((ngrp + 3) / 4), 128, s, part, ngrp, part2); (*launches)++; }
    MRZ_LAUNCH(k_crc_tail, 1, 1, s, buf, nblk * CRC_BLK, n - nblk * CRC_BLK, tail);
    (*launches)++;
}

extern "C" void mrz_launch_exclusive_scan(const i64 *in, i64 *out, i64 cnt, i64 *bsum, i64 *total, cudaStream_t s,
                                          int *launches) {
    i64 nb = (cnt + SCAN_BLK - 1) / SCAN_BLK;
    MRZ_LAUNCH(k_scan_local, (unsigned)nb, SCAN_B, s, in, out, cnt, bsum);
    MRZ_LAUNCH(k_scan_bsums, 1, SCAN_B, s, bsum, nb, total);
    MRZ_LAUNCH(k_scan_add, (unsigned)nb, SCAN_B, s, out, cnt, bsum);
    *launches += 3;
}

extern "C" void mrz_launch_item_sizes(const Token *toks, i64 m, i64 n, int cb, i64 *a0, i64 *a1, EmitTotals *tot,
                                      cudaStream_t s) {
    i64 items = m + 1;
    MRZ_LAUNCH(k_item_sizes, (unsigned)((items + 255) / 256), 256, s, toks, m, n, cb, a0, a1, tot);
}

extern "C" void mrz_launch_emit(const u8 *buf, const Token *toks, i64 m, i64 n, int cb, const i64 *off0,
                                const i64 *off1, i64 s1_len, u8 *s0, u8 *s1, const u32 *crc, cudaStream_t s,
                                int *launches) {
    i64 items = m + 1;
    MRZ_LAUNCH(k_emit_s0, (unsigned)((items * 32 + 255) / 256), 256, s, toks, m, n, cb, off0, s0, crc);
    (*launches)++;
    if (s1_len > 0) {
        i64 thr = (s1_len + 15) / 16;
        MRZ_LAUNCH(k_emit_s1, (unsigned)((thr + 255) / 256), 256, s, buf, toks, m, n, off1, s1_len, s1);
        (*launches)++;
    }
}

// ---- CRC32 without any allocation ----------------------------------------------------------------------------------
// mrz_crc32_noalloc (include/mrz_gpu.h): the chunk CRC32 for callers that cannot call cudaMalloc (CUDA graph capture,
// tight serving loops, harnesses that forbid it).  Same kernels as the chunk path (k_crc_blocks / k_crc_level2 /
// k_crc_tail); their scratch - the per-block partial CRCs - lives in ONE file-scope device array of fixed size instead
// of a context's buffers, and the final combine, which the chunk path does on the host, is a kernel here.
// Every GPU that loads the library holds the array until exit.  All calls on a device share it: calls that overlap on
// the GPU (different streams) race.  Nothing in it has to start at zero: a call overwrites every entry it reads.
#define NOALLOC_MAX_BYTES ((i64)1 << 30)
#define NOALLOC_L1 (NOALLOC_MAX_BYTES / CRC_BLK)          // 65536 level-1 partials
#define NOALLOC_L2 (NOALLOC_L1 / 32)                      // 2048 level-2 partials
static __device__ __align__(256) u32 g_noalloc_scratch[NOALLOC_L1 + NOALLOC_L2 + 64];   // [L1 | L2 | tail]

__device__ __forceinline__ u32 gf2_apply_rt(const u32 *mat, u32 v) {
    u32 r = 0;
    for (int b = 0; v; v >>= 1, b++)
        if (v & 1u) r ^= mat[b];
    return r;
}
__device__ void gf2_square(u32 *dst, const u32 *m) {
    for (int b = 0; b < 32; b++) dst[b] = gf2_apply_rt(m, m[b]);
}
// register -> register after nbytes zero bytes (the host's shift_matrix / crc_shift, mrz_gpu.cu), applied to v
__device__ u32 gf2_shift(u32 v, u64 nbytes) {
    u32 cur[32], tmp[32];
    cur[0] = 0xEDB88320u;                                  // one zero bit
    for (int b = 1; b < 32; b++) cur[b] = 1u << (b - 1);
    for (int q = 0; q < 3; q++) { gf2_square(tmp, cur); for (int b = 0; b < 32; b++) cur[b] = tmp[b]; }   // one zero byte
    while (nbytes) {
        if (nbytes & 1) v = gf2_apply_rt(cur, v);
        nbytes >>= 1;
        if (nbytes) { gf2_square(tmp, cur); for (int b = 0; b < 32; b++) cur[b] = tmp[b]; }
    }
    return v;
}
// the combine of mrz_gpu.cu run_phase_b, on the device: one thread
__global__ void k_crc_final(const u32 *__restrict__ part, const u32 *__restrict__ part2, const u32 *__restrict__ tail, i64 n, u32 *out) {
    const i64 nblk = n / CRC_BLK, ngrp = nblk / 32, nrest = nblk - ngrp * 32;
    u32 m512k[32];
    gf2_square(m512k, c_crc.shift_blk[4]);                 // x^(8 * 256 KiB) squared
    u32 raw = 0;
    for (i64 g = 0; g < ngrp; g++) raw = gf2_apply_rt(m512k, raw) ^ part2[g];
    for (i64 r = 0; r < nrest; r++) raw = gf2_apply_rt(c_crc.shift_blk[0], raw) ^ part[ngrp * 32 + r];
    raw = gf2_shift(raw, (u64)(n - nblk * CRC_BLK)) ^ *tail;
    *out = raw ^ gf2_shift(0xFFFFFFFFu, (u64)n) ^ 0xFFFFFFFFu;
}

extern "C" int mrz_crc_ensure_device(int device, cudaStream_t s);     // mrz_gpu.cu: the CRC tables of this device (__constant__), once

#ifndef MRZ_CUSIM
#include <mutex>
static std::mutex g_noalloc_mu;
#endif
static u32 *g_noalloc_ptr[64];        // device address of g_noalloc_scratch per device (the symbol's name is NOT an address in host code)
static char g_noalloc_err[256];

extern "C" const char *mrz_noalloc_last_error(void) { return g_noalloc_err; }

extern "C" int mrz_crc32_noalloc(int device, const uint8_t *d_data, int64_t n, uint32_t *d_crc_out, void *stream) {
    cudaStream_t s = (cudaStream_t)stream;
    g_noalloc_err[0] = 0;
    if (device < 0 || device >= 64 || n < 0 || !d_crc_out || (n && !d_data)) { snprintf(g_noalloc_err, sizeof g_noalloc_err, "bad argument"); return -3; }
    if (n > NOALLOC_MAX_BYTES) {
        snprintf(g_noalloc_err, sizeof g_noalloc_err, "mrz_crc32_noalloc: %lld bytes exceed the fixed scratch (maximum %lld bytes per call)",
                 (long long)n, (long long)NOALLOC_MAX_BYTES);
        return -5;
    }
    if (((uintptr_t)d_data & 15) != 0) { snprintf(g_noalloc_err, sizeof g_noalloc_err, "mrz_crc32_noalloc: the data pointer must be 16-byte aligned"); return -3; }
    cudaError_t e = cudaSetDevice(device);
    if (e != cudaSuccess) { snprintf(g_noalloc_err, sizeof g_noalloc_err, "cudaSetDevice: %s", cudaGetErrorString(e)); return -2; }
    {
#ifndef MRZ_CUSIM
        std::lock_guard<std::mutex> g(g_noalloc_mu);
#endif
        if (!g_noalloc_ptr[device]) {            // once per device, with that device current
            void *p = nullptr;
            e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
            if (e != cudaSuccess || !p) { snprintf(g_noalloc_err, sizeof g_noalloc_err, "cudaGetSymbolAddress: %s", cudaGetErrorString(e)); return -2; }
            g_noalloc_ptr[device] = (u32 *)p;
        }
    }
    int rc = mrz_crc_ensure_device(device, s);
    if (rc) { snprintf(g_noalloc_err, sizeof g_noalloc_err, "cannot upload the CRC tables"); return rc; }
    u32 *part = g_noalloc_ptr[device], *part2 = part + NOALLOC_L1, *tail = part2 + NOALLOC_L2;
    int launches = 0;
    mrz_launch_crc(d_data, n, part, part2, tail, s, &launches);
    MRZ_LAUNCH(k_crc_final, 1, 1, s, part, part2, tail, n, d_crc_out);
    e = cudaGetLastError();
    if (e != cudaSuccess) { snprintf(g_noalloc_err, sizeof g_noalloc_err, "launch: %s", cudaGetErrorString(e)); return -2; }
    return 0;
}
