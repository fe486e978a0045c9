// tagmap.cu — the rolling tag of hash_search for every position of a chunk, in one streaming pass.
//
// Replaces the per-byte loop of hash_search (rzip.c:548-573): single_next_tag (rzip.c:330-337),
// single_full_tag (rzip.c:348-358) and the lookup / insert gates (rzip.c:573, :579).  The gates only ask how
// many trailing one bits a tag has (`(t & mask) == mask` with mask = 2^k - 1), so the product of this pass is one
// byte per position, bmap[p] = trailing ones of tag[p], plus a 64:1 summary (bsum[b] = max over 64 positions) that
// lets the automaton find its next candidates at any mask level without touching the bytes in between.  The map is
// computed ONCE per chunk and shared by every victim_round variant of that chunk; the 64-bit tag of a candidate is
// recomputed on demand by the automaton (31 table look-ups).
//
// Layout of the work: a thread owns 32 consecutive positions.  It loads the 63 bytes they cover with four 128-bit
// loads, looks every byte up ONCE in a 32-bit copy of hash_index[] that is replicated 32 times in shared memory
// (lane l only ever touches bank l: conflict-free), forms the first tag from 31 values and rolls it over its run
// from registers.  Only the low 32 bits of a tag are needed to count trailing ones; the one-in-2^32 tag whose low
// word is all ones is redone in 64 bits.  HBM traffic: 1 byte read + 1 byte written per position (+ 1/64).
#include "mrz_common.cuh"

__device__ __forceinline__ i64 full_tag64(const u8 *buf, const i64 *hidx, i64 p) {   // rzip.c:348-358
    i64 t = 0;
    for (int i = 0; i < MRZ_MIN_MATCH; i++) t ^= __ldg(hidx + buf[p + i]);
    return t;
}

__global__ void __launch_bounds__(MAP_THREADS, 3)
k_tagmap(const u8 *__restrict__ buf, const i64 *__restrict__ hidx, i64 n, u8 *__restrict__ bmap, u8 *__restrict__ bsum,
         i64 ntiles) {
    __shared__ u32 lut[256 * 32];
    for (int i = threadIdx.x; i < 256 * 32; i += MAP_THREADS) lut[i] = (u32)(u64)hidx[i >> 5];
    __syncthreads();
    const int lane = threadIdx.x & 31;
    const u32 *L = lut + lane;
    const i64 end = n - MRZ_MIN_MATCH;          // positions 0..end carry a tag (rzip.c:541, :548)
    for (i64 tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
        const i64 a = tile * MAP_TILE + (i64)threadIdx.x * MAP_RUN;
        u32 w[16];
        if (a <= n + MRZ_PAD - 64) {            // the zero padding behind the chunk keeps these loads in bounds
            const uint4 *src = reinterpret_cast<const uint4 *>(buf + a);
#pragma unroll
            for (int q = 0; q < 4; q++) {
                const uint4 v = __ldg(src + q);
                w[4 * q] = v.x; w[4 * q + 1] = v.y; w[4 * q + 2] = v.z; w[4 * q + 3] = v.w;
            }
        } else {
#pragma unroll
            for (int q = 0; q < 16; q++) w[q] = 0;
        }
        u32 h[63];
#pragma unroll
        for (int i = 0; i < 63; i++) h[i] = L[((w[i >> 2] >> ((i & 3) * 8)) & 255u) << 5];
        u32 t = 0;
#pragma unroll
        for (int i = 0; i < MRZ_MIN_MATCH; i++) t ^= h[i];
        u32 o[8];
        int mx = 0;
        if (a + MAP_RUN - 1 <= end) {           // the whole run carries tags (all but the last thread of a chunk)
            bool wide = false;                  // some tag has >= 32 trailing ones (one in 2^32): redone below
#pragma unroll
            for (int i = 0; i < MAP_RUN; i++) {
                const int z = __ffs((int)~t) - 1;          // -1 when the low word is all ones
                wide = wide || z < 0;
                mx = z > mx ? z : mx;
                const u32 zb = (u32)z & 255u;              // 255 marks a wide tag
                if ((i & 3) == 0) o[i >> 2] = zb; else o[i >> 2] |= zb << ((i & 3) * 8);
                if (i + 1 < MAP_RUN) t ^= h[i] ^ h[i + MRZ_MIN_MATCH];
            }
            if (wide) {
                for (int i = 0; i < MAP_RUN; i++)
                    if (((o[i >> 2] >> ((i & 3) * 8)) & 255u) == 255u) {
                        const int z = tz1(full_tag64(buf, hidx, a + i));   // count in 64 bits
                        o[i >> 2] = (o[i >> 2] & ~(255u << ((i & 3) * 8))) | ((u32)z << ((i & 3) * 8));
                        mx = z > mx ? z : mx;
                    }
            }
        } else {
#pragma unroll
            for (int i = 0; i < MAP_RUN; i++) {
                const u32 x = ~t;
                int z = x ? (__ffs((int)x) - 1) : 32;
                if (a + i > end) z = 0;
                else if (z == 32) z = tz1(full_tag64(buf, hidx, a + i));   // >= 32 trailing ones: count in 64 bits
                mx = z > mx ? z : mx;
                if ((i & 3) == 0) o[i >> 2] = (u32)z; else o[i >> 2] |= (u32)z << ((i & 3) * 8);
                if (i + 1 < MAP_RUN) t ^= h[i] ^ h[i + MRZ_MIN_MATCH];
            }
        }
        uint4 *dst = reinterpret_cast<uint4 *>(bmap + a);
        dst[0] = make_uint4(o[0], o[1], o[2], o[3]);
        dst[1] = make_uint4(o[4], o[5], o[6], o[7]);
        const int other = __shfl_xor_sync(0xffffffffu, mx, 1);
        if (!(lane & 1)) bsum[a >> MAP_BLK_SHIFT] = (u8)(other > mx ? other : mx);
    }
}

// Debug / parity: the full 64-bit tag of every position p in [0, count).
__global__ void __launch_bounds__(256)
k_debug_tags(const u8 *__restrict__ buf, const i64 *__restrict__ hidx, i64 count, i64 *__restrict__ tags) {
    const i64 p = (i64)blockIdx.x * blockDim.x + threadIdx.x;
    if (p < count) tags[p] = full_tag64(buf, hidx, p);
}

extern "C" i64 mrz_tagmap_tiles(i64 n) { return (n + MAP_TILE - 1) / MAP_TILE + 1; }   // map bytes = tiles * MAP_TILE

extern "C" void mrz_launch_tagmap(const u8 *buf, const i64 *hidx, i64 n, u8 *bmap, u8 *bsum, int sms, cudaStream_t s) {
    const i64 ntiles = mrz_tagmap_tiles(n);
    i64 grid = (i64)sms * 3;    // 80 registers, 32 KiB of shared memory: three CTAs per SM
    if (grid > ntiles) grid = ntiles;
    MRZ_LAUNCH(k_tagmap, (unsigned)grid, MAP_THREADS, s, buf, hidx, n, bmap, bsum, ntiles);
}

extern "C" void mrz_launch_debug_tags(const u8 *buf, const i64 *hidx, i64 count, i64 *tags, cudaStream_t s) {
    if (count > 0) MRZ_LAUNCH(k_debug_tags, (unsigned)((count + 255) / 256), 256, s, buf, hidx, count, tags);
}
