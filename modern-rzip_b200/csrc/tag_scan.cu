// tag_scan.cu — rolling-tag computation and candidate compaction.
//
// Replaces the per-byte loop of hash_search (rzip.c:548-573): single_next_tag (rzip.c:330-337),
// single_full_tag (rzip.c:348-358) and the lookup gate (rzip.c:573).  Instead of rolling one tag
// through the chunk on one core, every CTA stages an 8 KiB tile (+30 byte halo) in shared memory
// with 128-bit loads, each thread rolls a private tag over 32 consecutive positions, and positions
// whose tag passes the current lookup mask are compacted, in position order, into a flat
// (pos:u32, tag:i64) list that the automaton consumes.
//
// Three launches per segment: count -> offsets -> emit.  All segment geometry is derived on the
// device from AutoState so the host never has to know where the automaton stands.
#include "mrz_common.cuh"

// padded shared-memory layout: one extra word every 8 words => a stride of 9 words between the
// 32-byte runs of neighbouring threads (conflict-free byte reads).
#define SM_WORDS ((SCAN_TILE + 64) / 4)
#define SM_PADDED (SM_WORDS + (SM_WORDS >> 3) + 8)

__device__ __forceinline__ unsigned sm_byte(const u32 *sm, int b) {
    int w = b >> 2;
    return (sm[w + (w >> 3)] >> ((b & 3) * 8)) & 255u;
}

// Geometry of the segment that starts at st->p + 1 (identical in all three kernels).
__device__ __forceinline__ void seg_geometry(const AutoState *st, i64 &seg_lo, i64 &seg_hi_max, int &tiles) {
    seg_lo = st->p + 1;
    int k = __popcll((u64)st->min_mask);
    i64 len = (i64)SEG_CAND_TARGET << (k < 20 ? k : 20);
    i64 maxlen = (i64)SEG_MAX_TILES * SCAN_TILE;
    if (len > maxlen) len = maxlen;
    seg_hi_max = seg_lo + len;
    if (seg_hi_max > st->end + 1) seg_hi_max = st->end + 1;  // positions examined are 1..end
    i64 span = seg_hi_max - seg_lo;
    tiles = span > 0 ? (int)((span + SCAN_TILE - 1) / SCAN_TILE) : 0;
}

// Stage bytes [lo, lo + SCAN_TILE + 30] of the chunk; returns d = lo - aligned base.
__device__ __forceinline__ int stage_tile(const u8 *buf, i64 lo, i64 n, u32 *sm) {
    uintptr_t a = (uintptr_t)(buf + lo);
    const uint4 *src = (const uint4 *)(a & ~(uintptr_t)15);
    int d = (int)(a & 15);
    const int nvec = (SCAN_TILE + 64) / 16;  // covers d + TILE + 31 <= TILE + 46
    i64 vmax = (n + MRZ_PAD - (lo - d)) / 16;  // vectors that stay inside the padded allocation
    for (int v = threadIdx.x; v < nvec; v += SCAN_THREADS) {
        uint4 x = make_uint4(0, 0, 0, 0);
        if (v < vmax) x = __ldg(src + v);
        int w = v * 4;
        int pw = w + (w >> 3);
        sm[pw] = x.x; sm[pw + 1] = x.y; sm[pw + 2] = x.z; sm[pw + 3] = x.w;
    }
    return d;
}

// Roll the tag over this thread's run; calls f(i, tag) for each position lo + tid*RUN + i < hi.
template <class F>
__device__ __forceinline__ void roll_run(const u32 *sm, const i64 *hx, int d, i64 lo, i64 hi, F f) {
    int b0 = d + threadIdx.x * SCAN_RUN;
    i64 p0 = lo + (i64)threadIdx.x * SCAN_RUN;
    if (p0 >= hi) return;
    i64 t = 0;
#pragma unroll
    for (int i = 0; i < MRZ_MIN_MATCH; i++) t ^= hx[sm_byte(sm, b0 + i)];
    int cnt = (hi - p0) < SCAN_RUN ? (int)(hi - p0) : SCAN_RUN;
    f(0, t);
    for (int i = 1; i < cnt; i++) {
        t ^= hx[sm_byte(sm, b0 + i - 1)] ^ hx[sm_byte(sm, b0 + i + MRZ_MIN_MATCH - 1)];
        f(i, t);
    }
}

__global__ void __launch_bounds__(SCAN_THREADS)
k_scan_count(const u8 *__restrict__ buf, const i64 *__restrict__ hidx, const AutoState *__restrict__ st,
             int *__restrict__ tile_cnt) {
    if (st->done) return;
    i64 seg_lo, seg_hi; int tiles;
    seg_geometry(st, seg_lo, seg_hi, tiles);
    if ((int)blockIdx.x >= tiles) return;
    __shared__ u32 sm[SM_PADDED];
    __shared__ i64 hx[256];
    __shared__ int total;
    hx[threadIdx.x] = hidx[threadIdx.x];
    if (threadIdx.x == 0) total = 0;
    i64 lo = seg_lo + (i64)blockIdx.x * SCAN_TILE;
    int d = stage_tile(buf, lo, st->n, sm);
    __syncthreads();
    i64 mask = st->min_mask;
    int c = 0;
    roll_run(sm, hx, d, lo, seg_hi, [&](int, i64 t) { c += ((t & mask) == mask); });
    for (int o = 16; o; o >>= 1) c += __shfl_xor_sync(0xffffffffu, c, o);
    if ((threadIdx.x & 31) == 0 && c) atomicAdd(&total, c);
    __syncthreads();
    if (threadIdx.x == 0) tile_cnt[blockIdx.x] = total;
}

// Exclusive scan of tile counts (single CTA, 1024 threads x 8 tiles) and segment bookkeeping.
__global__ void __launch_bounds__(1024)
k_scan_offsets(AutoState *st, const int *__restrict__ tile_cnt, i64 *__restrict__ tile_off) {
    if (st->done) return;
    i64 seg_lo, seg_hi; int tiles;
    seg_geometry(st, seg_lo, seg_hi, tiles);
    __shared__ i64 wsum[32];
    __shared__ int nfit_s;
    const int PER = SEG_MAX_TILES / 1024;
    int base = threadIdx.x * PER;
    i64 loc[PER];
    i64 s = 0;
#pragma unroll
    for (int i = 0; i < PER; i++) {
        int tl = base + i;
        loc[i] = s;
        s += (tl < tiles) ? tile_cnt[tl] : 0;
    }
    i64 inc = s;
    for (int o = 1; o < 32; o <<= 1) {
        i64 v = __shfl_up_sync(0xffffffffu, inc, o);
        if ((threadIdx.x & 31) >= o) inc += v;
    }
    if ((threadIdx.x & 31) == 31) wsum[threadIdx.x >> 5] = inc;
    if (threadIdx.x == 0) nfit_s = tiles;
    __syncthreads();
    if (threadIdx.x < 32) {
        i64 w = wsum[threadIdx.x], wi = w;
        for (int o = 1; o < 32; o <<= 1) {
            i64 v = __shfl_up_sync(0xffffffffu, wi, o);
            if (threadIdx.x >= o) wi += v;
        }
        wsum[threadIdx.x] = wi - w;
    }
    __syncthreads();
    i64 excl = wsum[threadIdx.x >> 5] + inc - s;
#pragma unroll
    for (int i = 0; i < PER; i++) {
        int tl = base + i;
        if (tl <= tiles) tile_off[tl] = excl + loc[i];
        // first tile whose end would overflow the flat list
        if (tl < tiles && excl + loc[i] + tile_cnt[tl] > SEG_CAND_CAP) atomicMin(&nfit_s, tl);
    }
    if (threadIdx.x == 1023 && tiles == SEG_MAX_TILES) tile_off[SEG_MAX_TILES] = excl + s;
    __syncthreads();
    if (threadIdx.x == 0) {
        int nfit = nfit_s;
        i64 hi = seg_lo + (i64)nfit * SCAN_TILE;
        if (hi > seg_hi) hi = seg_hi;
        st->seg_lo = seg_lo;
        st->seg_hi = hi;
        st->seg_mask = st->min_mask;
        st->seg_tiles = nfit;
        st->seg_ncand = tile_off[nfit];
    }
}

__global__ void __launch_bounds__(SCAN_THREADS)
k_scan_emit(const u8 *__restrict__ buf, const i64 *__restrict__ hidx, const AutoState *__restrict__ st,
            const i64 *__restrict__ tile_off, u32 *__restrict__ cand_pos, i64 *__restrict__ cand_tag) {
    if (st->done) return;
    if ((int)blockIdx.x >= st->seg_tiles) return;
    __shared__ u32 sm[SM_PADDED];
    __shared__ i64 hx[256];
    __shared__ int wsum[SCAN_THREADS / 32];
    hx[threadIdx.x] = hidx[threadIdx.x];
    i64 seg_lo = st->seg_lo, seg_hi = st->seg_hi;
    i64 lo = seg_lo + (i64)blockIdx.x * SCAN_TILE;
    int d = stage_tile(buf, lo, st->n, sm);
    __syncthreads();
    i64 mask = st->seg_mask;
    u32 bits = 0;  // which of my 32 positions pass
    roll_run(sm, hx, d, lo, seg_hi, [&](int i, i64 t) { bits |= (u32)((t & mask) == mask) << i; });
    int c = __popc(bits);
    int inc = c;
    for (int o = 1; o < 32; o <<= 1) {
        int v = __shfl_up_sync(0xffffffffu, inc, o);
        if ((threadIdx.x & 31) >= o) inc += v;
    }
    if ((threadIdx.x & 31) == 31) wsum[threadIdx.x >> 5] = inc;
    __syncthreads();
    int woff = 0;
    for (int w = 0; w < (int)(threadIdx.x >> 5); w++) woff += wsum[w];
    i64 out = tile_off[blockIdx.x] + woff + inc - c;
    if (bits) {
        // second roll to write tags in order (cheaper than keeping 32 tags in registers)
        u32 rel0 = (u32)(lo - seg_lo) + threadIdx.x * SCAN_RUN;
        roll_run(sm, hx, d, lo, seg_hi, [&](int i, i64 t) {
            if ((bits >> i) & 1u) {
                cand_pos[out] = rel0 + i;
                cand_tag[out] = t;
                out++;
            }
        });
    }
}

// Debug / parity: tags of all positions p in [0, count) with the same staging + rolling code.
__global__ void __launch_bounds__(SCAN_THREADS)
k_debug_tags(const u8 *__restrict__ buf, const i64 *__restrict__ hidx, i64 n, i64 count, i64 *__restrict__ tags) {
    __shared__ u32 sm[SM_PADDED];
    __shared__ i64 hx[256];
    hx[threadIdx.x] = hidx[threadIdx.x];
    i64 lo = (i64)blockIdx.x * SCAN_TILE;
    int d = stage_tile(buf, lo, n, sm);
    __syncthreads();
    i64 p0 = lo + (i64)threadIdx.x * SCAN_RUN;
    roll_run(sm, hx, d, lo, count, [&](int i, i64 t) { tags[p0 + i] = t; });
}

extern "C" void mrz_launch_scan(const u8 *buf, const i64 *hidx, AutoState *st, int *tile_cnt, i64 *tile_off,
                                u32 *cand_pos, i64 *cand_tag, int tiles, cudaStream_t s) {
    if (tiles < 1) tiles = 1;
    MRZ_LAUNCH(k_scan_count, tiles, SCAN_THREADS, s, buf, hidx, st, tile_cnt);
    MRZ_LAUNCH(k_scan_offsets, 1, 1024, s, st, tile_cnt, tile_off);
    MRZ_LAUNCH(k_scan_emit, tiles, SCAN_THREADS, s, buf, hidx, st, tile_off, cand_pos, cand_tag);
}

extern "C" void mrz_launch_debug_tags(const u8 *buf, const i64 *hidx, i64 n, i64 count, i64 *tags, cudaStream_t s) {
    int tiles = (int)((count + SCAN_TILE - 1) / SCAN_TILE);
    if (tiles > 0) MRZ_LAUNCH(k_debug_tags, tiles, SCAN_THREADS, s, buf, hidx, n, count, tags);
}
