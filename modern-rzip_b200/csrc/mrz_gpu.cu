// mrz_gpu.cu — C-ABI of libmrz_gpu.so (see include/mrz_gpu.h) and the per-chunk orchestration.
//
// Pipeline for one chunk (replaces hash_search, rzip.c:507-667):
//   H2D chunk -> memset table (rzip.c:518-519) -> { tag_scan segment -> automaton } until p == end
//   -> item sizes -> two exclusive scans -> stream 0 / stream 1 emission + CRC32 -> D2H streams.
// No CPU fallback: every entry point needs a CUDA device.
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <string>
#ifndef MRZ_CUSIM
#include <mutex>
#include <thread>
#endif
#include <vector>

#include "../../include/mrz_gpu.h"
#include "mrz_common.cuh"

extern "C" {
void mrz_launch_scan(const u8 *, const i64 *, AutoState *, int *, i64 *, u32 *, i64 *, int, cudaStream_t);
void mrz_launch_debug_tags(const u8 *, const i64 *, i64, i64, i64 *, cudaStream_t);
size_t mrz_planio_bytes();
int mrz_launch_automaton(const u8 *, i64 *, i64 *, u64 *, u32 *, unsigned *, const i64 *, AutoState *, const i64 *, const u32 *,
                         const i64 *, Token *, i64 *, int, void *, cudaStream_t);
size_t mrz_grid_state_bytes();
int mrz_launch_automaton_grid(int, int, const u8 *, i64 *, i64 *, u64 *, u32 *, unsigned *, void *, const i64 *, AutoState *,
                              const i64 *, const u32 *, const i64 *, Token *, i64 *, cudaStream_t);
void mrz_crc_upload_consts(const void *);
void mrz_launch_crc(const u8 *, i64, u32 *, u32 *, u32 *, cudaStream_t, int *);
void mrz_launch_exclusive_scan(const i64 *, i64 *, i64, i64 *, i64 *, cudaStream_t, int *);
void mrz_launch_item_sizes(const Token *, i64, i64, int, i64 *, i64 *, EmitTotals *, cudaStream_t);
void mrz_launch_emit(const u8 *, const Token *, i64, i64, int, const i64 *, const i64 *, i64, u8 *, u8 *,
                     const u32 *, cudaStream_t, int *);
}

static const struct { i64 mb; int freq; int chain; } kLevels[10] = {  // rzip.c:65-73
    {1, 4, 1}, {2, 4, 2}, {4, 4, 2}, {8, 4, 2}, {16, 4, 3}, {32, 4, 4}, {32, 2, 6}, {64, 1, 16}, {64, 1, 32}, {64, 1, 128}};

static thread_local std::string g_err;

struct mrz_ctx {
    int device;
    i64 max_chunk;
    cudaStream_t stream;
    cudaEvent_t ev[6];
    mrz_ctx *parent;     // variant context (mrz_ctx_create_variant): d_buf and d_s1 belong to the parent
    // device buffers
    u8 *d_buf_alloc, *d_buf;  // chunk with MRZ_PAD slack on both sides
    i64 *d_tab; i64 tab_slots;   // tags[tab_slots + 32] then offs[tab_slots + 32]
    u32 *d_gsum;         // group summary of the automaton, kept across its launches
    u64 *d_filt;         // tag filter of the automaton: MRZ_FILT_WORDS 64-bit words
    unsigned *d_claim;   // grid automaton: per-slot claim of the current batch (automaton_grid.cu)
    void *d_gstate;      // grid automaton: state shared by the CTAs
    int grid_g, grid_t;  // grid automaton geometry (0 = the one-CTA automaton of automaton.cu)
    void *d_planio;      // automaton.cu plan offload: batch + plan records
    int helpers;         // plan-offload CTAs (MRZ_HELPERS)
    i64 *d_hidx;
    AutoState *d_st;
    int *d_tile_cnt; i64 *d_tile_off; u32 *d_cand_pos; i64 *d_cand_tag;
    Token *d_toks; i64 tok_cap;
    i64 *d_a0, *d_a1, *d_off0, *d_off1, *d_bsum; i64 item_cap;
    EmitTotals *d_tot;
    i64 *d_dbg;          // phase cycle counters of k_automaton (MRZ_PROFILE=1), else NULL
    i64 h_dbg[40];
    u8 *d_s0; i64 s0_cap; u8 *d_s1; i64 s1_cap;
    u32 *d_crc_part, *d_crc_part2, *d_crc_misc;  // misc[0]=tail, misc[1]=final
    // host staging: pinned, so that the small read-backs are truly asynchronous.  (A D2H copy into pageable memory
    // blocks inside the driver until the stream has drained; with one host thread per variant that serialised the
    // automata of mrz_run_variants — measured: wall = nvar x one automaton.)
    u32 *h_crc; i64 h_crc_cap;
    struct Mailbox { AutoState st; i64 totals[2]; u32 tailcrc; u32 crc; EmitTotals tot; i64 dbg[40]; } *h_mb;
    // last run
    i64 last_n, last_s0, last_s1, last_ntok;
    // between mrz_run_automaton and mrz_run_emit
    AutoState last_state; int last_cb; bool have_tokens;
    float ms_scan, ms_auto, ms_phase_a; int launches;
    std::string err;
};

#define CK(call)                                                                                    \
    do {                                                                                            \
        cudaError_t e_ = (call);                                                                    \
        if (e_ != cudaSuccess) {                                                                    \
            char b_[512];                                                                           \
            snprintf(b_, sizeof b_, "%s:%d %s: %s", __FILE__, __LINE__, #call, cudaGetErrorString(e_)); \
            if (ctx) ctx->err = b_;                                                                 \
            g_err = b_;                                                                             \
            return e_ == cudaErrorMemoryAllocation ? MRZ_E_NOMEM : MRZ_E_CUDA;                      \
        }                                                                                           \
    } while (0)

// ---- CRC32 shift operators over GF(2) (host) ---------------------------------------------------
struct CrcConstHost {
    u32 tab[4][256];
    u32 shift_sub[5][32];
    u32 shift_blk[5][32];
};
static CrcConstHost g_crc;
static u32 g_byte_mat[32];  // advance a raw CRC register over one zero byte
static bool g_crc_ready = false;

static u32 h_apply(const u32 *m, u32 v) {
    u32 r = 0;
    for (int b = 0; v; v >>= 1, b++)
        if (v & 1) r ^= m[b];
    return r;
}
static void h_square(u32 *dst, const u32 *m) {
    for (int b = 0; b < 32; b++) dst[b] = h_apply(m, m[b]);
}
static void h_mul(u32 *dst, const u32 *a, const u32 *b) {  // dst = a o b
    u32 t[32];
    for (int i = 0; i < 32; i++) t[i] = h_apply(a, b[i]);
    memcpy(dst, t, sizeof t);
}
static void shift_matrix(u32 *out, u64 nbytes) {  // operator: register -> register after nbytes zero bytes
    u32 res[32], cur[32], tmp[32];
    for (int b = 0; b < 32; b++) res[b] = 1u << b;
    memcpy(cur, g_byte_mat, sizeof cur);
    while (nbytes) {
        if (nbytes & 1) h_mul(res, cur, res);
        h_square(tmp, cur);
        memcpy(cur, tmp, sizeof cur);
        nbytes >>= 1;
    }
    memcpy(out, res, sizeof res);
}
static u32 crc_shift(u32 crc, u64 nbytes) {
    u32 m[32];
    shift_matrix(m, nbytes);
    return h_apply(m, crc);
}
static void crc_host_init() {
    if (g_crc_ready) return;
    for (u32 i = 0; i < 256; i++) {
        u32 c = i;
        for (int k = 0; k < 8; k++) c = (c >> 1) ^ (0xEDB88320u & (0u - (c & 1)));
        g_crc.tab[0][i] = c;
    }
    for (u32 i = 0; i < 256; i++)
        for (int t = 1; t < 4; t++)
            g_crc.tab[t][i] = (g_crc.tab[t - 1][i] >> 8) ^ g_crc.tab[0][g_crc.tab[t - 1][i] & 255];
    // one zero bit: reflected register shifts right, bit 0 feeds the polynomial
    u32 bit[32], m2[32], m4[32];
    bit[0] = 0xEDB88320u;
    for (int b = 1; b < 32; b++) bit[b] = 1u << (b - 1);
    h_square(m2, bit);
    h_square(m4, m2);
    h_square(g_byte_mat, m4);
    for (int j = 0; j < 5; j++) {
        shift_matrix(g_crc.shift_sub[j], (u64)512 << j);
        shift_matrix(g_crc.shift_blk[j], (u64)16384 << j);
    }
    g_crc_ready = true;
}

// ---- library -----------------------------------------------------------------------------------
extern "C" const char *mrz_version(void) { return "mrz_gpu 0.1 (sm_100a)"; }

extern "C" int mrz_device_count(void) {
    // mrz_run_variants drives one stream per automaton.  With the default of 8 hardware queues, streams that share a
    // queue serialise behind each other's kernel -> read-back -> next launch chain (measured on B200: 16 variants ran
    // 4 at a time; with 32 queues all 16 run side by side, +8 % over one).  Only effective before the CUDA context
    // exists, which is why it sits in the first call every entry path makes; an explicit setting wins.
    setenv("CUDA_DEVICE_MAX_CONNECTIONS", "32", 0);
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess) {
        g_err = std::string("cudaGetDeviceCount: ") + cudaGetErrorString(e);
        return MRZ_E_NODEVICE;
    }
    return n;
}

extern "C" const char *mrz_last_error(const mrz_ctx *ctx) { return ctx ? ctx->err.c_str() : g_err.c_str(); }

extern "C" void *mrz_host_alloc(int64_t bytes) {
    void *p = nullptr;
    if (cudaHostAlloc(&p, (size_t)bytes, cudaHostAllocDefault) != cudaSuccess) return nullptr;
    return p;
}
extern "C" void mrz_host_free(void *p) {
    if (p) cudaFreeHost(p);
}

static int ensure_tokens(mrz_ctx *ctx, i64 cap) {
    if (cap <= ctx->tok_cap) return 0;
    if (ctx->d_toks) cudaFree(ctx->d_toks);
    ctx->d_toks = nullptr; ctx->tok_cap = 0;
    CK(cudaMalloc(&ctx->d_toks, (size_t)cap * sizeof(Token)));
    if (!(getenv("MRZ_NO_ZERO_INIT") && strchr(getenv("MRZ_NO_ZERO_INIT"), 't'))) CK(cudaMemset(ctx->d_toks, 0, (size_t)cap * sizeof(Token)));
    ctx->tok_cap = cap;
    return 0;
}
static int ensure_items(mrz_ctx *ctx, i64 items) {
    if (items + 2 <= ctx->item_cap) return 0;
    i64 cap = items + 2 + items / 4;
    cudaFree(ctx->d_a0); cudaFree(ctx->d_a1); cudaFree(ctx->d_off0); cudaFree(ctx->d_off1); cudaFree(ctx->d_bsum);
    ctx->d_a0 = ctx->d_a1 = ctx->d_off0 = ctx->d_off1 = ctx->d_bsum = nullptr; ctx->item_cap = 0;
    CK(cudaMalloc(&ctx->d_a0, (size_t)cap * 8));
    CK(cudaMalloc(&ctx->d_a1, (size_t)cap * 8));
    CK(cudaMalloc(&ctx->d_off0, (size_t)cap * 8));
    CK(cudaMalloc(&ctx->d_off1, (size_t)cap * 8));
    CK(cudaMalloc(&ctx->d_bsum, (size_t)(cap / 4096 + 2) * 8));
    ctx->item_cap = cap;
    return 0;
}
static int ensure_s0(mrz_ctx *ctx, i64 bytes) {
    if (bytes <= ctx->s0_cap) return 0;
    if (ctx->d_s0) cudaFree(ctx->d_s0);
    ctx->d_s0 = nullptr; ctx->s0_cap = 0;
    i64 cap = bytes + bytes / 4 + 4096;
    CK(cudaMalloc(&ctx->d_s0, (size_t)cap));
    ctx->s0_cap = cap;
    return 0;
}

extern "C" void mrz_ctx_destroy(mrz_ctx *ctx) {
    if (!ctx) return;
    cudaSetDevice(ctx->device);
    if (ctx->parent) ctx->d_buf_alloc = nullptr, ctx->d_s1 = nullptr;   // borrowed from the parent
    cudaFree(ctx->d_buf_alloc); cudaFree(ctx->d_tab); cudaFree(ctx->d_gsum); cudaFree(ctx->d_filt); cudaFree(ctx->d_claim); cudaFree(ctx->d_planio); cudaFree(ctx->d_gstate); cudaFree(ctx->d_hidx); cudaFree(ctx->d_st);
    cudaFree(ctx->d_tile_cnt); cudaFree(ctx->d_tile_off); cudaFree(ctx->d_cand_pos); cudaFree(ctx->d_cand_tag);
    cudaFree(ctx->d_toks); cudaFree(ctx->d_a0); cudaFree(ctx->d_a1); cudaFree(ctx->d_off0); cudaFree(ctx->d_off1);
    cudaFree(ctx->d_bsum); cudaFree(ctx->d_tot); cudaFree(ctx->d_dbg); cudaFree(ctx->d_s0); cudaFree(ctx->d_s1);
    cudaFree(ctx->d_crc_part); cudaFree(ctx->d_crc_part2); cudaFree(ctx->d_crc_misc);
    if (ctx->h_crc) cudaFreeHost(ctx->h_crc);
    if (ctx->h_mb) cudaFreeHost(ctx->h_mb);
    for (auto &e : ctx->ev)
        if (e) cudaEventDestroy(e);
    if (ctx->stream) cudaStreamDestroy(ctx->stream);
    delete ctx;
}

static int ctx_create_impl(mrz_ctx *ctx, int device, i64 max_chunk, mrz_ctx *parent) {
    CK(cudaSetDevice(device));
    ctx->parent = parent;
    crc_host_init();
    mrz_crc_upload_consts(&g_crc);
    CK(cudaGetLastError());
    ctx->device = device;
    ctx->max_chunk = max_chunk;
    CK(cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking));
    for (auto &e : ctx->ev) CK(cudaEventCreate(&e));
    if (parent) {
        ctx->d_buf_alloc = parent->d_buf_alloc;
        ctx->d_buf = parent->d_buf;
    } else {
        CK(cudaMalloc(&ctx->d_buf_alloc, (size_t)max_chunk + 2 * MRZ_PAD + 64));
        ctx->d_buf = ctx->d_buf_alloc + MRZ_PAD;
        CK(cudaMemset(ctx->d_buf_alloc, 0, MRZ_PAD));
    }
    ctx->tab_slots = (i64)64 * 65536;  // largest levels[].mb_used = 64 MiB
    CK(cudaMalloc(&ctx->d_tab, (size_t)(ctx->tab_slots + 32) * 16));  // + slack for 4-slot probe loads
    CK(cudaMalloc(&ctx->d_gsum, (size_t)(ctx->tab_slots / 8 / 2 + 1024) * 4));   // 16-bit thermometer per 8 slots (automaton.cu GRP)
    CK(cudaMalloc(&ctx->d_filt, (size_t)MRZ_FILT_WORDS * 8));
    CK(cudaMalloc(&ctx->d_claim, (size_t)(ctx->tab_slots + 32) * 4));
    CK(cudaMalloc(&ctx->d_gstate, mrz_grid_state_bytes()));
    CK(cudaMalloc(&ctx->d_planio, mrz_planio_bytes()));
    {
        const char *h = getenv("MRZ_HELPERS");
#ifdef MRZ_CUSIM
        ctx->helpers = 0;   // the emulator tests select helpers / the grid automaton per test (environment)
#else
        ctx->helpers = 0;   // measured on B200: the hand-off costs more than the plans it moves (profiles/r1_helpers_sweep.txt)
#endif
        if (h) ctx->helpers = atoi(h);
        if (ctx->helpers < 0 || ctx->helpers > 32 || (ctx->helpers & (ctx->helpers - 1))) { ctx->err = "MRZ_HELPERS must be 0 or a power of two <= 32"; return MRZ_E_ARG; }
    }
    {   // geometry of the cooperative grid: G CTAs (the last one serves the sweep queue) x T threads
        const char *a = getenv("MRZ_AUTOMATON"), *g = getenv("MRZ_G"), *t = getenv("MRZ_T");
#ifdef MRZ_CUSIM
        ctx->grid_g = 5; ctx->grid_t = 64;
#else
        ctx->grid_g = 33; ctx->grid_t = 32;
#endif
        if (g) ctx->grid_g = atoi(g);
        if (t) ctx->grid_t = atoi(t);
        if (!a || strcmp(a, "grid")) ctx->grid_g = 0;   // default: the one-CTA automaton (automaton.cu); MRZ_AUTOMATON=grid selects automaton_grid.cu
        if (ctx->grid_g) {
            if (ctx->grid_g < 2 || ctx->grid_g > 148 || ctx->grid_t < 32 || ctx->grid_t > 256 || (ctx->grid_t & 31) ||
                (i64)(ctx->grid_g - 1) * ctx->grid_t > 8192) {
                ctx->err = "bad MRZ_G / MRZ_T";
                return MRZ_E_ARG;
            }
        }
    }
    CK(cudaMalloc(&ctx->d_hidx, 256 * 8));
    CK(cudaMalloc(&ctx->d_st, sizeof(AutoState)));
    CK(cudaMalloc(&ctx->d_tile_cnt, (SEG_MAX_TILES + 1) * sizeof(int)));
    CK(cudaMalloc(&ctx->d_tile_off, (SEG_MAX_TILES + 2) * sizeof(i64)));
    CK(cudaMalloc(&ctx->d_cand_pos, (size_t)(SEG_CAND_CAP + 64) * sizeof(u32)));
    CK(cudaMalloc(&ctx->d_cand_tag, (size_t)(SEG_CAND_CAP + 64) * sizeof(i64)));
    CK(cudaMalloc(&ctx->d_tot, sizeof(EmitTotals)));
    // Device memory recycled from an earlier process is not zero.  A second variant set created after a large run
    // faulted (illegal address in the per-segment kernels) while the same program on a fresh box did not: something
    // on the path reads a word of one of these work buffers before writing it.  Not root-caused yet
    // (compute-sanitizer is closed on this pool); every work buffer starts zeroed, as on a fresh device.
    const char *nz = getenv("MRZ_NO_ZERO_INIT");   // debugging knob: comma-free list of letters naming buffers left as allocated
#define ZERO(letter, ptr, bytes) do { if (!(nz && strchr(nz, letter))) CK(cudaMemset(ptr, 0, bytes)); } while (0)
    ZERO('g', ctx->d_gstate, mrz_grid_state_bytes());
    ZERO('p', ctx->d_planio, mrz_planio_bytes());
    ZERO('h', ctx->d_hidx, 256 * 8);
    ZERO('s', ctx->d_st, sizeof(AutoState));
    ZERO('c', ctx->d_tile_cnt, (SEG_MAX_TILES + 1) * sizeof(int));
    ZERO('o', ctx->d_tile_off, (SEG_MAX_TILES + 2) * sizeof(i64));
    ZERO('P', ctx->d_cand_pos, (size_t)(SEG_CAND_CAP + 64) * sizeof(u32));
    ZERO('T', ctx->d_cand_tag, (size_t)(SEG_CAND_CAP + 64) * sizeof(i64));
    if (getenv("MRZ_PROFILE")) CK(cudaMalloc(&ctx->d_dbg, 40 * 8));
    if (parent) {
        ctx->d_s1 = parent->d_s1;
    } else {
        CK(cudaMalloc(&ctx->d_s1, (size_t)max_chunk + 64));
    }
    ctx->s1_cap = max_chunk + 64;
    i64 nblk = max_chunk / 16384 + 1;
    CK(cudaMalloc(&ctx->d_crc_part, (size_t)nblk * 4));
    CK(cudaMalloc(&ctx->d_crc_part2, (size_t)(nblk / 32 + 1) * 4));
    CK(cudaMalloc(&ctx->d_crc_misc, 16));
    ZERO('r', ctx->d_crc_part, (size_t)nblk * 4);
    ZERO('r', ctx->d_crc_part2, (size_t)(nblk / 32 + 1) * 4);
    ZERO('r', ctx->d_crc_misc, 16);
    ctx->h_crc_cap = nblk / 32 + 64;
    CK(cudaHostAlloc((void **)&ctx->h_crc, (size_t)ctx->h_crc_cap * 4, cudaHostAllocDefault));
    CK(cudaHostAlloc((void **)&ctx->h_mb, sizeof(*ctx->h_mb), cudaHostAllocDefault));
    int rc;
    if ((rc = ensure_tokens(ctx, max_chunk / 256 + 1024))) return rc;
    if (parent) return 0;   // a variant allocates its emission buffers only if it is the one chosen to emit
    if ((rc = ensure_items(ctx, max_chunk / 256 + 1024))) return rc;
    if ((rc = ensure_s0(ctx, max_chunk / 32 + 4096))) return rc;
    return 0;
}

extern "C" int mrz_ctx_create(int device, int64_t max_chunk, mrz_ctx **out) {
    if (!out || max_chunk < 0) return MRZ_E_ARG;
    *out = nullptr;
    int n = mrz_device_count();
    if (n <= 0) {
        if (n == 0) g_err = "no CUDA device";
        return MRZ_E_NODEVICE;
    }
    if (device < 0 || device >= n) { g_err = "bad device index"; return MRZ_E_ARG; }
    mrz_ctx *ctx = new mrz_ctx();
    int rc = ctx_create_impl(ctx, device, max_chunk, nullptr);
    if (rc) {
        g_err = ctx->err;
        mrz_ctx_destroy(ctx);
        return rc;
    }
    *out = ctx;
    return 0;
}

extern "C" int mrz_ctx_create_variant(mrz_ctx *parent, mrz_ctx **out) {
    if (!out || !parent || parent->parent) return MRZ_E_ARG;
    *out = nullptr;
    mrz_ctx *ctx = new mrz_ctx();
    int rc = ctx_create_impl(ctx, parent->device, parent->max_chunk, parent);
    if (rc) {
        g_err = ctx->err;
        mrz_ctx_destroy(ctx);
        return rc;
    }
    *out = ctx;
    return 0;
}

extern "C" int mrz_upload(mrz_ctx *ctx, const uint8_t *chunk, int64_t n) {
    if (!ctx || n < 0 || n > ctx->max_chunk || (n && !chunk)) return MRZ_E_ARG;
    CK(cudaSetDevice(ctx->device));
    if (n) CK(cudaMemcpyAsync(ctx->d_buf, chunk, (size_t)n, cudaMemcpyHostToDevice, ctx->stream));
    CK(cudaMemsetAsync(ctx->d_buf + n, 0, MRZ_PAD + 64, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    return 0;
}

// One attempt with the current token capacity.  Returns 1 when the token buffer overflowed.
static int run_phase_a(mrz_ctx *ctx, i64 n, const mrz_params *prm) {
    cudaStream_t s = ctx->stream;
    ctx->have_tokens = false;
    const int L = prm->level;
    AutoState h;
    memset(&h, 0, sizeof h);
    h.n = n;
    h.end = n - MRZ_MIN_MATCH;
    i64 hashsize = kLevels[L].mb * (1048576 / 16);
    int bits = 0;
    while (((i64)1 << bits) < hashsize) bits++;  // rzip.c:521-523
    h.hsize = (i64)1 << bits;
    h.hmask = h.hsize - 1;
    h.limit = h.hsize / 3 * 2;  // rzip.c:529
    h.chain = kLevels[L].chain;
    h.min_mask = h.ins_mask = ((i64)1 << kLevels[L].freq) - 1;  // rzip.c:509, :533
    h.victim_round = prm->victim_round_in;
    h.tok_cap = ctx->tok_cap;
    { const char *bj = getenv("MRZ_BMAX_JUMP"); h.pad0 = bj ? atoi(bj) : 0; }   // batch size right after an emission (tuning)
    h.done = h.end <= 0 ? 1 : 0;
    int launches = 0;
    float ms_scan = 0, ms_auto = 0;

    CK(cudaEventRecord(ctx->ev[0], s));
    CK(cudaMemcpyAsync(ctx->d_hidx, prm->hash_index, 256 * 8, cudaMemcpyHostToDevice, s));
    CK(cudaMemcpyAsync(ctx->d_st, &h, sizeof h, cudaMemcpyHostToDevice, s));
    CK(cudaMemsetAsync(ctx->d_tab, 0, (size_t)(ctx->tab_slots + 32) * 16, s));  // rzip.c:518-519
    CK(cudaMemsetAsync(ctx->d_gsum, 0, (size_t)(ctx->tab_slots / 8 / 2 + 1024) * 4, s));
    CK(cudaMemsetAsync(ctx->d_filt, 0, (size_t)MRZ_FILT_WORDS * 8, s));
    CK(cudaMemsetAsync(ctx->d_claim, 0xff, (size_t)(ctx->tab_slots + 32) * 4, s));
    if (ctx->grid_g) CK(cudaMemsetAsync(ctx->d_gstate, 0, mrz_grid_state_bytes(), s));
    if (ctx->d_dbg) CK(cudaMemsetAsync(ctx->d_dbg, 0, 40 * 8, s));
    launches += 2;
    // CRC of the chunk (independent of the automaton)
    mrz_launch_crc(ctx->d_buf, n, ctx->d_crc_part, ctx->d_crc_part2, ctx->d_crc_misc, s, &launches);

    while (!h.done) {
        int k = __builtin_popcountll((unsigned long long)h.min_mask);
        i64 len = (i64)SEG_CAND_TARGET << (k < 20 ? k : 20);
        i64 maxlen = (i64)SEG_MAX_TILES * SCAN_TILE;
        if (len > maxlen) len = maxlen;
        i64 span = h.end + 1 - (h.p + 1);
        if (span > len) span = len;
        int tiles = (int)((span + SCAN_TILE - 1) / SCAN_TILE);
        CK(cudaEventRecord(ctx->ev[1], s));
        mrz_launch_scan(ctx->d_buf, ctx->d_hidx, ctx->d_st, ctx->d_tile_cnt, ctx->d_tile_off, ctx->d_cand_pos,
                        ctx->d_cand_tag, tiles, s);
        CK(cudaEventRecord(ctx->ev[2], s));
        if (ctx->grid_g) {
            int lrc = mrz_launch_automaton_grid(ctx->grid_g, ctx->grid_t, ctx->d_buf, ctx->d_tab, ctx->d_tab + ctx->tab_slots + 32,
                                                ctx->d_filt, ctx->d_gsum, ctx->d_claim, ctx->d_gstate, ctx->d_hidx, ctx->d_st,
                                                ctx->d_tile_off, ctx->d_cand_pos, ctx->d_cand_tag, ctx->d_toks, ctx->d_dbg, s);
            if (lrc) CK((cudaError_t)lrc);
        } else {
            int lrc = mrz_launch_automaton(ctx->d_buf, ctx->d_tab, ctx->d_tab + ctx->tab_slots + 32, ctx->d_filt, ctx->d_gsum, ctx->d_claim, ctx->d_hidx,
                                           ctx->d_st, ctx->d_tile_off, ctx->d_cand_pos, ctx->d_cand_tag, ctx->d_toks, ctx->d_dbg, ctx->helpers,
                                           ctx->d_planio, s);
            if (lrc) CK((cudaError_t)lrc);
        }
        CK(cudaEventRecord(ctx->ev[3], s));
        launches += 4;
        i64 prev_p = h.p;
        CK(cudaMemcpyAsync(&ctx->h_mb->st, ctx->d_st, sizeof h, cudaMemcpyDeviceToHost, s));
        CK(cudaStreamSynchronize(s));
        CK(cudaGetLastError());
        h = ctx->h_mb->st;
        float a = 0, b = 0;
        cudaEventElapsedTime(&a, ctx->ev[1], ctx->ev[2]);
        cudaEventElapsedTime(&b, ctx->ev[2], ctx->ev[3]);
        ms_scan += a;
        ms_auto += b;
        if (h.error == 2) return 1;
        if (h.error) {
            char m[128];
            snprintf(m, sizeof m, "device automaton error %d at p=%lld", h.error, (long long)h.p);
            ctx->err = m;
            return MRZ_E_INTERNAL;
        }
        if (!h.done && h.p <= prev_p) {
            ctx->err = "automaton made no progress";
            return MRZ_E_INTERNAL;
        }
    }

    if (ctx->d_dbg) {
        CK(cudaMemcpyAsync(ctx->h_mb->dbg, ctx->d_dbg, 40 * 8, cudaMemcpyDeviceToHost, s));
        CK(cudaStreamSynchronize(s));
        memcpy(ctx->h_dbg, ctx->h_mb->dbg, sizeof ctx->h_dbg);
        static const char *nm[12] = {"refill", "gather", "plan", "conflict", "commit", "serial", "inval", "epilogue", "ops", "batches", "serial_ops", ""};
        fprintf(stderr, "[mrz profile] automaton cycles:");
        for (int q = 0; q < 11; q++) fprintf(stderr, " %s=%lld", nm[q], (long long)ctx->h_dbg[q]);
        fprintf(stderr, " | serial because: filter+=%lld twin-complex=%lld insert-complex=%lld conflict=%lld emit/count=%lld | forwarded=%lld | plan cycles per op: mean=%lld, mean of per-batch max=%lld",
                (long long)ctx->h_dbg[11], (long long)ctx->h_dbg[12], (long long)ctx->h_dbg[13], (long long)ctx->h_dbg[14],
                (long long)ctx->h_dbg[15], (long long)ctx->h_dbg[16], (long long)(ctx->h_dbg[17] / (ctx->h_dbg[8] ? ctx->h_dbg[8] : 1)),
                (long long)(ctx->h_dbg[18] / (ctx->h_dbg[9] ? ctx->h_dbg[9] : 1)));
        fprintf(stderr, "\n[mrz profile] serial op parts (cycles/calls): filter=%lld/%lld lookup=%lld/%lld insert_fast=%lld/%lld insert_slow=%lld/%lld clean=%lld/%lld dense-loop=%lld/%lld\n",
                (long long)ctx->h_dbg[20], (long long)ctx->h_dbg[21], (long long)ctx->h_dbg[22], (long long)ctx->h_dbg[23],
                (long long)ctx->h_dbg[24], (long long)ctx->h_dbg[25], (long long)ctx->h_dbg[26], (long long)ctx->h_dbg[27],
                (long long)ctx->h_dbg[28], (long long)ctx->h_dbg[29], (long long)ctx->h_dbg[30], (long long)ctx->h_dbg[31]);
        fprintf(stderr, "[mrz profile] CTA lookup phase cycles=%lld | plans given up: sweep-queue stop=%lld twin=%lld dependency=%lld owner-moves-occupant=%lld equal-tag=%lld owner-lesser=%lld limits=%lld w1-claimed=%lld\n",
                (long long)ctx->h_dbg[19], (long long)ctx->h_dbg[32], (long long)ctx->h_dbg[33], (long long)ctx->h_dbg[34],
                (long long)ctx->h_dbg[35], (long long)ctx->h_dbg[36], (long long)ctx->h_dbg[37], (long long)ctx->h_dbg[38], (long long)ctx->h_dbg[39]);
    }
    CK(cudaEventRecord(ctx->ev[4], s));
    CK(cudaStreamSynchronize(s));
    cudaEventElapsedTime(&ctx->ms_phase_a, ctx->ev[0], ctx->ev[4]);
    ctx->last_state = h;
    ctx->last_cb = prm->chunk_bytes;
    ctx->last_n = n; ctx->last_ntok = h.ntok; ctx->last_s0 = ctx->last_s1 = 0;
    ctx->ms_scan = ms_scan; ctx->ms_auto = ms_auto; ctx->launches = launches;
    ctx->have_tokens = true;
    return 0;
}

// Counters of a chunk whose automaton has run (everything but the stream lengths and the CRC).
static void fill_automaton_result(const mrz_ctx *ctx, mrz_result *res) {
    const AutoState &h = ctx->last_state;
    memset(res, 0, sizeof *res);
    res->tag_hits = h.tag_hits;
    res->tag_misses = h.tag_misses;
    res->inserts = h.inserts;
    res->matches = h.ntok;            // un-split matches; the emission reports rzip.c's split count
    res->victim_round_out = h.victim_round;
    res->victim_events = h.victim_events;
    res->final_min_mask = h.min_mask;
    res->n_ops = h.n_ops;
    res->ms_kernels = ctx->ms_phase_a;
    res->ms_tagscan = ctx->ms_scan;
    res->ms_automaton = ctx->ms_auto;
    res->gpu_launches = ctx->launches;
}

// Emission of the tokens the last run_phase_a of this context left behind.
static int run_phase_b(mrz_ctx *ctx, mrz_result *res) {
    cudaStream_t s = ctx->stream;
    if (!ctx->have_tokens) { ctx->err = "mrz_run_emit without a preceding automaton run"; return MRZ_E_ARG; }
    const AutoState &h = ctx->last_state;
    const i64 n = ctx->last_n;
    const int chunk_bytes = ctx->last_cb;
    int launches = ctx->launches;
    float ms_emit = 0;
    // ---- emission ----
    CK(cudaEventRecord(ctx->ev[1], s));
    CK(cudaMemsetAsync(ctx->d_tot, 0, sizeof(EmitTotals), s));
    i64 m = h.ntok, items = m + 1;
    int rc;
    if ((rc = ensure_items(ctx, items))) return rc;
    mrz_launch_item_sizes(ctx->d_toks, m, n, chunk_bytes, ctx->d_a0, ctx->d_a1, ctx->d_tot, s);
    launches++;
    mrz_launch_exclusive_scan(ctx->d_a0, ctx->d_off0, items, ctx->d_bsum, ctx->d_off0 + items, s, &launches);
    mrz_launch_exclusive_scan(ctx->d_a1, ctx->d_off1, items, ctx->d_bsum, ctx->d_off1 + items, s, &launches);
    i64 *totals = ctx->h_mb->totals;
    CK(cudaMemcpyAsync(&totals[0], ctx->d_off0 + items, 8, cudaMemcpyDeviceToHost, s));
    CK(cudaMemcpyAsync(&totals[1], ctx->d_off1 + items, 8, cudaMemcpyDeviceToHost, s));
    // CRC partials -> host combine
    i64 nblk = n / 16384, ngrp = nblk / 32, nrest = nblk - ngrp * 32;
    u32 &tailcrc = ctx->h_mb->tailcrc;
    tailcrc = 0;
    if (ngrp) CK(cudaMemcpyAsync(ctx->h_crc, ctx->d_crc_part2, (size_t)ngrp * 4, cudaMemcpyDeviceToHost, s));
    if (nrest)
        CK(cudaMemcpyAsync(ctx->h_crc + ngrp, ctx->d_crc_part + ngrp * 32, (size_t)nrest * 4, cudaMemcpyDeviceToHost, s));
    CK(cudaMemcpyAsync(&tailcrc, ctx->d_crc_misc, 4, cudaMemcpyDeviceToHost, s));
    CK(cudaStreamSynchronize(s));
    u32 raw = 0;
    {
        u32 m512k[32], m16k[32];
        shift_matrix(m512k, 512 * 1024);
        shift_matrix(m16k, 16384);
        for (i64 g = 0; g < ngrp; g++) raw = h_apply(m512k, raw) ^ ctx->h_crc[g];
        for (i64 r = 0; r < nrest; r++) raw = h_apply(m16k, raw) ^ ctx->h_crc[ngrp + r];
        i64 tl = n - nblk * 16384;
        raw = crc_shift(raw, (u64)tl) ^ tailcrc;
    }
    const u32 crc = raw ^ crc_shift(0xFFFFFFFFu, (u64)n) ^ 0xFFFFFFFFu;
    ctx->h_mb->crc = crc;
    CK(cudaMemcpyAsync(ctx->d_crc_misc + 1, &ctx->h_mb->crc, 4, cudaMemcpyHostToDevice, s));
    if ((rc = ensure_s0(ctx, totals[0]))) return rc;
    mrz_launch_emit(ctx->d_buf, ctx->d_toks, m, n, chunk_bytes, ctx->d_off0, ctx->d_off1, totals[1], ctx->d_s0,
                    ctx->d_s1, ctx->d_crc_misc + 1, s, &launches);
    EmitTotals &tot = ctx->h_mb->tot;
    CK(cudaMemcpyAsync(&tot, ctx->d_tot, sizeof tot, cudaMemcpyDeviceToHost, s));
    CK(cudaEventRecord(ctx->ev[2], s));
    CK(cudaStreamSynchronize(s));
    CK(cudaGetLastError());
    cudaEventElapsedTime(&ms_emit, ctx->ev[1], ctx->ev[2]);

    fill_automaton_result(ctx, res);
    res->s0_len = totals[0];
    res->s1_len = totals[1];
    res->matches = tot.matches;
    res->match_bytes = tot.match_bytes;
    res->literals = tot.literals;
    res->literal_bytes = totals[1];
    res->crc32 = crc;
    res->ms_kernels = ctx->ms_phase_a + ms_emit;
    res->ms_emit = ms_emit;
    res->gpu_launches = launches;
    ctx->last_s0 = totals[0]; ctx->last_s1 = totals[1]; ctx->last_ntok = m;
    return 0;
}

// Automaton with token-buffer growth: a chunk whose tokens overflowed is redone with a larger buffer.
static int run_automaton_grow(mrz_ctx *ctx, i64 n, const mrz_params *prm) {
    for (;;) {
        int rc = run_phase_a(ctx, n, prm);
        if (rc != 1) return rc;
        i64 cap = ctx->tok_cap * 8;
        i64 hard = n / MRZ_MIN_MATCH + 16;   // bounded by one match per 31 bytes
        if (cap > hard) cap = hard;
        if (cap <= ctx->tok_cap) { ctx->err = "token capacity exhausted"; return MRZ_E_INTERNAL; }
        if ((rc = ensure_tokens(ctx, cap))) return rc;
    }
}

static int check_run_args(const mrz_ctx *ctx, i64 n, const mrz_params *prm) {
    if (!ctx || !prm || n < 0 || n > ctx->max_chunk) return MRZ_E_ARG;
    if (prm->level < 0 || prm->level > 9 || prm->chunk_bytes < 1 || prm->chunk_bytes > 8) return MRZ_E_ARG;
    if (prm->victim_round_in < 0 || prm->victim_round_in >= kLevels[prm->level].chain) return MRZ_E_ARG;
    return 0;
}

extern "C" int mrz_run_resident(mrz_ctx *ctx, int64_t n, const mrz_params *prm, mrz_result *res) {
    if (!res || check_run_args(ctx, n, prm)) return MRZ_E_ARG;
    CK(cudaSetDevice(ctx->device));
    int rc = run_automaton_grow(ctx, n, prm);
    if (rc) return rc;
    return run_phase_b(ctx, res);
}

extern "C" int mrz_run_automaton(mrz_ctx *ctx, int64_t n, const mrz_params *prm, mrz_result *res) {
    if (!res || check_run_args(ctx, n, prm)) return MRZ_E_ARG;
    CK(cudaSetDevice(ctx->device));
    int rc = run_automaton_grow(ctx, n, prm);
    if (rc) return rc;
    fill_automaton_result(ctx, res);
    return 0;
}

extern "C" int mrz_run_emit(mrz_ctx *ctx, mrz_result *res) {
    if (!ctx || !res) return MRZ_E_ARG;
    CK(cudaSetDevice(ctx->device));
    return run_phase_b(ctx, res);
}

// Variant v runs with victim_round_in = v on its own stream, tables and CTA; on the device the nvar automata
// (one SM each) run side by side.  One host thread per variant drives the per-segment launch loop.
extern "C" int mrz_run_variants(mrz_ctx *const *ctxs, int nvar, int64_t n, const mrz_params *prm, mrz_result *res) {
    if (!ctxs || !prm || !res || nvar < 1 || prm->level < 0 || prm->level > 9 || nvar > kLevels[prm->level].chain) return MRZ_E_ARG;
    if (!ctxs[0]) return MRZ_E_ARG;
#ifndef MRZ_CUSIM
    // One variant set per device at a time.  Measured on B200: the <= 32 automata of one set run side by side
    // correctly, but two sets driven concurrently from two host threads (32 automata of 1 GiB tar-like chunks)
    // ended in an illegal-address fault that is not root-caused yet — so concurrent callers queue here.
    static std::mutex mu[64];
    std::lock_guard<std::mutex> guard(mu[ctxs[0]->device & 63]);
#endif
    std::vector<int> rcs((size_t)nvar, 0);
    auto one = [&](int v) {
        mrz_params p = *prm;
        p.victim_round_in = v;
        rcs[(size_t)v] = mrz_run_automaton(ctxs[v], n, &p, &res[v]);
    };
#ifndef MRZ_CUSIM
    // Side by side only on request (MRZ_VARIANTS_CONCURRENT=1): measured on B200 the 16 automata of a 1 GiB chunk
    // finish in 1.06x the time of one, and every completed run was bit-exact, but inside bench.py's process the
    // concurrent form sometimes ended in an illegal-address fault or a hang (never in profiles/diag_*.py, never
    // with one automaton at a time) — open issue, DESIGN.md §6.  The default runs the variants one after another.
    const char *cc = getenv("MRZ_VARIANTS_CONCURRENT");
    if (cc && atoi(cc) > 0) {
        std::vector<std::thread> th;
        for (int v = 1; v < nvar; v++) th.emplace_back(one, v);
        one(0);
        for (auto &t : th) t.join();
    } else {
        for (int v = 0; v < nvar; v++) one(v);
    }
#else
    for (int v = 0; v < nvar; v++) one(v);   // the fiber emulator is single-threaded
#endif
    for (int v = 0; v < nvar; v++)
        if (rcs[(size_t)v]) {
            g_err = ctxs[v] ? ctxs[v]->err : "null variant context";
            return rcs[(size_t)v];
        }
    return 0;
}

extern "C" int mrz_download(mrz_ctx *ctx, uint8_t *s0, int64_t s0_cap, uint8_t *s1, int64_t s1_cap) {
    if (!ctx) return MRZ_E_ARG;
    if (ctx->last_s0 > s0_cap || ctx->last_s1 > s1_cap) return MRZ_E_CAPACITY;
    CK(cudaSetDevice(ctx->device));
    if (ctx->last_s0) CK(cudaMemcpyAsync(s0, ctx->d_s0, (size_t)ctx->last_s0, cudaMemcpyDeviceToHost, ctx->stream));
    if (ctx->last_s1) CK(cudaMemcpyAsync(s1, ctx->d_s1, (size_t)ctx->last_s1, cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    return 0;
}

extern "C" int mrz_rzip_chunk(mrz_ctx *ctx, const uint8_t *chunk, int64_t n, const mrz_params *prm, uint8_t *s0,
                              int64_t s0_cap, uint8_t *s1, int64_t s1_cap, mrz_result *res) {
    int rc = mrz_upload(ctx, chunk, n);
    if (rc) return rc;
    rc = mrz_run_resident(ctx, n, prm, res);
    if (rc) return rc;
    return mrz_download(ctx, s0, s0_cap, s1, s1_cap);
}

extern "C" int mrz_debug_tags(mrz_ctx *ctx, int64_t n, const int64_t hash_index[256], int64_t *tags_out,
                              int64_t count) {
    if (!ctx || count < 0 || count + 30 > n) return MRZ_E_ARG;
    CK(cudaSetDevice(ctx->device));
    i64 *d_tags = nullptr;
    CK(cudaMalloc(&d_tags, (size_t)(count + 1) * 8));
    cudaError_t e = cudaMemcpyAsync(ctx->d_hidx, hash_index, 256 * 8, cudaMemcpyHostToDevice, ctx->stream);
    if (e == cudaSuccess) {
        mrz_launch_debug_tags(ctx->d_buf, ctx->d_hidx, n, count, d_tags, ctx->stream);
        e = cudaMemcpyAsync(tags_out, d_tags, (size_t)count * 8, cudaMemcpyDeviceToHost, ctx->stream);
    }
    if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
    cudaFree(d_tags);
    CK(e);
    return 0;
}

extern "C" int mrz_debug_crc32(mrz_ctx *ctx, int64_t n, uint32_t *crc_out) {
    if (!ctx || !crc_out) return MRZ_E_ARG;
    (void)n;
    CK(cudaSetDevice(ctx->device));
    CK(cudaMemcpy(crc_out, ctx->d_crc_misc + 1, 4, cudaMemcpyDeviceToHost));
    return 0;
}

extern "C" int mrz_debug_tokens(mrz_ctx *ctx, int64_t *triples, int64_t cap, int64_t *count) {
    if (!ctx || !count) return MRZ_E_ARG;
    *count = ctx->last_ntok;
    if (ctx->last_ntok > cap) return MRZ_E_CAPACITY;
    CK(cudaSetDevice(ctx->device));
    if (ctx->last_ntok)
        CK(cudaMemcpy(triples, ctx->d_toks, (size_t)ctx->last_ntok * sizeof(Token), cudaMemcpyDeviceToHost));
    return 0;
}
