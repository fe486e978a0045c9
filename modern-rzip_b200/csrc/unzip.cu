// unzip.cu — token replay of one chunk on the device: the inverse of emit.cu.
//
// Replaces unzip_literal / unzip_match and the loop of runzip_chunk (runzip.c:120-207, :226-327): stream 0 is a list of
// headers (00 len16: `len` literal bytes follow in stream 1; 01 len16 ofs[chunk_bytes]: copy `len` bytes from `ofs`
// bytes back in the output, runzip.c:162-201), ended by 00 00 00 and the chunk's CRC32 (rzip.c:662-665).
//
// The header list is parsed on the host (it is a few bytes per token and strictly sequential: a header's size depends
// on its first byte); the bytes move on the device:
//   k_unzip_literals: every literal run goes to its place in the chunk - all runs at once, the output position of a
//                     run is a prefix sum over the token lengths.  Bound: HBM, 1 B read + 1 B written per literal byte.
//   k_unzip_matches:  the matches in stream order (a match may copy what an earlier match wrote, and its own first
//                     bytes when ofs < len: `n = MIN(len, offset)` repeated, runzip.c:175-201, i.e. a byte-wise forward
//                     copy).  Runs of consecutive matches that do not read anything written inside the run are found on
//                     the host ("waves") and copied by the whole grid at once; a wave boundary is a kernel boundary.
#include "mrz_common.cuh"

struct UnzipItem {
    i64 dst;    // position in the chunk
    i64 src;    // literal: offset in stream 1; match: dst - ofs
    i64 len;
};

// One CTA per item (grid-stride): byte copy with 16-byte vectors where both sides allow it.
__global__ void __launch_bounds__(256)
k_unzip_literals(const UnzipItem *__restrict__ items, i64 count, const u8 *__restrict__ s1, u8 *__restrict__ out) {
    for (i64 it = blockIdx.x; it < count; it += gridDim.x) {
        const UnzipItem e = items[it];
        const u8 *s = s1 + e.src;
        u8 *d = out + e.dst;
        i64 head = (16 - ((uintptr_t)d & 15)) & 15;      // bytes before the destination is 16-byte aligned
        if (head > e.len) head = e.len;
        for (i64 i = threadIdx.x; i < head; i += blockDim.x) d[i] = s[i];
        const i64 body = (e.len - head) >> 4;
        const u8 *sb = s + head;
        uint4 *db = reinterpret_cast<uint4 *>(d + head);
        if (((uintptr_t)sb & 15) == 0) {
            const uint4 *sv = reinterpret_cast<const uint4 *>(sb);
            for (i64 i = threadIdx.x; i < body; i += blockDim.x) db[i] = __ldg(sv + i);
        } else {
            for (i64 i = threadIdx.x; i < body; i += blockDim.x) {
                const u64 lo = ld_u64_unaligned(sb, 16 * i), hi = ld_u64_unaligned(sb, 16 * i + 8);
                db[i] = make_uint4((u32)lo, (u32)(lo >> 32), (u32)hi, (u32)(hi >> 32));
            }
        }
        const i64 done = head + (body << 4);
        for (i64 i = done + threadIdx.x; i < e.len; i += blockDim.x) d[i] = s[i];
    }
}

// The matches of one wave [first, first + count): none of them reads a byte that another match of the wave writes, so
// they are independent; a match whose offset is shorter than its length repeats its first `ofs` bytes (all of them
// final before the wave starts).  One CTA per match (grid-stride).
__global__ void __launch_bounds__(256)
k_unzip_matches(const UnzipItem *__restrict__ items, i64 first, i64 count, u8 *__restrict__ out) {
    for (i64 it = blockIdx.x; it < count; it += gridDim.x) {
        const UnzipItem e = items[first + it];
        const i64 ofs = e.dst - e.src;
        u8 *d = out + e.dst;
        const u8 *s = out + e.src;
        if (ofs >= e.len) {
            for (i64 i = threadIdx.x; i < e.len; i += blockDim.x) d[i] = s[i];
        } else {
            for (i64 i = threadIdx.x; i < e.len; i += blockDim.x) d[i] = s[i % ofs];
        }
    }
}

extern "C" void mrz_launch_unzip_literals(const UnzipItem *items, i64 count, const u8 *s1, u8 *out, int sms, cudaStream_t s) {
    if (count <= 0) return;
    i64 grid = (i64)sms * 8;
    if (grid > count) grid = count;
    MRZ_LAUNCH(k_unzip_literals, (unsigned)grid, 256, s, items, count, s1, out);
}

extern "C" void mrz_launch_unzip_matches(const UnzipItem *items, i64 first, i64 count, u8 *out, int sms, cudaStream_t s) {
    if (count <= 0) return;
    i64 grid = (i64)sms * 8;
    if (grid > count) grid = count;
    MRZ_LAUNCH(k_unzip_matches, (unsigned)grid, 256, s, items, first, count, out);
}
