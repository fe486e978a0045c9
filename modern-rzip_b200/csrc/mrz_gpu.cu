// mrz_gpu.cu — C-ABI of libmrz_gpu.so (see include/mrz_gpu.h) and the per-chunk orchestration.
//
// Pipeline for one chunk (replaces hash_search, rzip.c:507-667):
//   H2D chunk -> tag map + CRC32 (once per chunk) -> memset table (rzip.c:518-519) -> ONE persistent launch of the
//   automaton (a thread-block cluster per chunk / per victim_round variant) -> item sizes -> two exclusive scans
//   -> stream 0 / stream 1 emission -> D2H streams.
// No CPU fallback: every entry point needs a CUDA device.  No environment variables: see mrz_config.
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <string>
#ifndef MRZ_CUSIM
#include <mutex>
#endif
#include <vector>

#include "../../include/mrz_gpu.h"
#include "automaton.h"

extern "C" {
i64 mrz_tagmap_tiles(i64 n);
void mrz_launch_tagmap(const u8 *, const i64 *, i64, u8 *, u8 *, int, cudaStream_t);
void mrz_launch_debug_tags(const u8 *, const i64 *, i64, i64 *, cudaStream_t);
size_t mrz_shared_bytes();
size_t mrz_automaton_smem_bytes();
int mrz_launch_automata(const AutoDesc *, int, int, int, cudaStream_t, const void *, size_t);
struct UnzipItem { i64 dst, src, len; };   // unzip.cu
void mrz_launch_unzip_literals(const UnzipItem *, i64, const u8 *, u8 *, int, cudaStream_t);
void mrz_launch_unzip_matches(const UnzipItem *, i64, i64, u8 *, int, cudaStream_t);
int mrz_max_active_clusters(int, int);
void mrz_crc_upload_consts(const void *, cudaStream_t);
int mrz_crc_ensure_device(int, cudaStream_t);
void mrz_launch_crc(const u8 *, i64, u32 *, u32 *, u32 *, cudaStream_t, int *);
void mrz_launch_exclusive_scan(const i64 *, i64 *, i64, i64 *, i64 *, cudaStream_t, int *);
void mrz_launch_item_sizes(const Token *, i64, i64, int, i64 *, i64 *, EmitTotals *, cudaStream_t);
void mrz_launch_emit(const u8 *, const Token *, i64, i64, int, const i64 *, const i64 *, i64, u8 *, u8 *,
                     const u32 *, cudaStream_t, int *);
}

static const struct { i64 mb; int freq; int chain; } kLevels[10] = {  // rzip.c:65-73
    {1, 4, 1}, {2, 4, 2}, {4, 4, 2}, {8, 4, 2}, {16, 4, 3}, {32, 4, 4}, {32, 2, 6}, {64, 1, 16}, {64, 1, 32}, {64, 1, 128}};

static thread_local std::string g_err;
#define MAX_LAUNCH 256   // automata per launch

struct mrz_ctx {
    int device, sms;
    i64 max_chunk;
    mrz_config cfg;
    cudaStream_t stream;
    cudaEvent_t ev[6];
    mrz_ctx *parent;     // variant context (mrz_ctx_create_variant): chunk, tag map and d_s1 belong to the parent
    // chunk-level device buffers (owned by the root context)
    u8 *d_buf_alloc, *d_buf;  // chunk with MRZ_PAD slack on both sides
    u8 *d_bmap, *d_bsum;      // tag map and its summary
    i64 *d_hidx;
    u32 *d_crc_part, *d_crc_part2, *d_crc_misc;  // misc[0]=tail, misc[1]=final
    i64 upload_gen, map_gen, map_n;   // the map is valid for (upload_gen, map_n, map_hidx)
    i64 map_hidx[256];
    // automaton slot (every context)
    i64 *d_tab; i64 tab_slots;   // tags[tab_slots + 32] then offs[tab_slots + 32]
    u32 *d_gsum;         // group summary
    u64 *d_filt;         // tag filter: 2^filt_bits 64-bit words
    int filt_bits;
    unsigned *d_claim;   // per-slot claim of the current batch
    Shared *d_sh;        // cluster-shared state
    AutoState *d_st;
    Token *d_toks; i64 tok_cap;
    i64 *d_dbg;          // phase counters (cfg.profile), else NULL
    AutoDesc *d_desc;    // launch descriptors (root contexts)
    // emission
    i64 *d_a0, *d_a1, *d_off0, *d_off1, *d_bsumscan; i64 item_cap;
    EmitTotals *d_tot;
    u8 *d_s0; i64 s0_cap; u8 *d_s1; i64 s1_cap;
    // host staging: pinned, so that the small read-backs are truly asynchronous
    u32 *h_crc; i64 h_crc_cap;
    struct Mailbox { AutoState st[MAX_LAUNCH]; AutoDesc desc[MAX_LAUNCH]; i64 totals[2]; u32 tailcrc; u32 crc; EmitTotals tot; i64 dbg[MRZ_NDBG]; } *h_mb;
    // last run
    i64 last_n, last_s0, last_s1, last_ntok;
    AutoState last_state; int last_cb; bool have_tokens;
    float ms_map, ms_auto, ms_phase_a; int launches;
    int geom[3];
    i64 h_dbg[MRZ_NDBG];
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
#ifndef MRZ_CUSIM
static std::once_flag g_crc_once;
static std::mutex g_dev_mu;
static bool g_crc_on_device[64];
#else
static bool g_crc_ready = false;
#endif

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
static void crc_host_build() {
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
}
static void crc_host_init() {
#ifndef MRZ_CUSIM
    std::call_once(g_crc_once, crc_host_build);
#else
    if (!g_crc_ready) { crc_host_build(); g_crc_ready = true; }
#endif
}

// The CRC constants live in __constant__ memory: uploaded once per device, on a stream, complete before anyone launches.
extern "C" int mrz_crc_ensure_device(int device, cudaStream_t s) {
    crc_host_init();
#ifndef MRZ_CUSIM
    std::lock_guard<std::mutex> g(g_dev_mu);
    if (g_crc_on_device[device & 63]) return 0;
    mrz_crc_upload_consts(&g_crc, s);
    if (cudaStreamSynchronize(s) != cudaSuccess) return MRZ_E_CUDA;
    g_crc_on_device[device & 63] = true;
#else
    (void)device;
    mrz_crc_upload_consts(&g_crc, s);
#endif
    return 0;
}

// ---- library -----------------------------------------------------------------------------------
extern "C" const char *mrz_version(void) { return "mrz_gpu 0.2 (sm_100a, cluster automaton)"; }

extern "C" int mrz_device_count(void) {
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess) {
        g_err = std::string("cudaGetDeviceCount: ") + cudaGetErrorString(e);
        return MRZ_E_NODEVICE;
    }
    return n;
}

extern "C" const char *mrz_last_error(const mrz_ctx *ctx) { return ctx ? ctx->err.c_str() : g_err.c_str(); }

extern "C" void mrz_config_default(mrz_config *cfg) {
    if (!cfg) return;
    memset(cfg, 0, sizeof *cfg);
}

extern "C" void *mrz_host_alloc(int64_t bytes) {
    void *p = nullptr;
    if (cudaHostAlloc(&p, (size_t)(bytes > 0 ? bytes : 1), cudaHostAllocDefault) != cudaSuccess) return nullptr;
    return p;
}
extern "C" void mrz_host_free(void *p) {
    if (p) cudaFreeHost(p);
}

// device allocation; cfg.poison fills it with 0xA5 so that a read-before-write cannot hide behind a fresh (zero) page
static int dev_alloc(mrz_ctx *ctx, void **p, size_t bytes) {
    *p = nullptr;
    CK(cudaMalloc(p, bytes ? bytes : 16));
    if (ctx->cfg.poison) CK(cudaMemsetAsync(*p, 0xA5, bytes ? bytes : 16, ctx->stream));
    return 0;
}
#define DA(ptr, bytes) do { int rc_ = dev_alloc(ctx, (void **)&(ptr), (size_t)(bytes)); if (rc_) return rc_; } while (0)

static int ensure_tokens(mrz_ctx *ctx, i64 cap) {
    if (cap <= ctx->tok_cap) return 0;
    CK(cudaStreamSynchronize(ctx->stream));
    if (ctx->d_toks) cudaFree(ctx->d_toks);
    ctx->d_toks = nullptr; ctx->tok_cap = 0;
    DA(ctx->d_toks, (size_t)cap * sizeof(Token));
    ctx->tok_cap = cap;
    return 0;
}
static int ensure_items(mrz_ctx *ctx, i64 items) {
    if (items + 2 <= ctx->item_cap) return 0;
    i64 cap = items + 2 + items / 4;
    CK(cudaStreamSynchronize(ctx->stream));
    cudaFree(ctx->d_a0); cudaFree(ctx->d_a1); cudaFree(ctx->d_off0); cudaFree(ctx->d_off1); cudaFree(ctx->d_bsumscan);
    ctx->d_a0 = ctx->d_a1 = ctx->d_off0 = ctx->d_off1 = ctx->d_bsumscan = nullptr; ctx->item_cap = 0;
    DA(ctx->d_a0, (size_t)cap * 8);
    DA(ctx->d_a1, (size_t)cap * 8);
    DA(ctx->d_off0, (size_t)cap * 8);
    DA(ctx->d_off1, (size_t)cap * 8);
    DA(ctx->d_bsumscan, (size_t)(cap / 4096 + 2) * 8);
    ctx->item_cap = cap;
    return 0;
}
static int ensure_s0(mrz_ctx *ctx, i64 bytes) {
    if (bytes <= ctx->s0_cap) return 0;
    CK(cudaStreamSynchronize(ctx->stream));
    if (ctx->d_s0) cudaFree(ctx->d_s0);
    ctx->d_s0 = nullptr; ctx->s0_cap = 0;
    i64 cap = bytes + bytes / 4 + 4096;
    DA(ctx->d_s0, (size_t)cap);
    ctx->s0_cap = cap;
    return 0;
}

extern "C" void mrz_ctx_destroy(mrz_ctx *ctx) {
    if (!ctx) return;
    cudaSetDevice(ctx->device);
    if (ctx->stream) cudaStreamSynchronize(ctx->stream);
    if (!ctx->parent) {
        cudaFree(ctx->d_buf_alloc); cudaFree(ctx->d_bmap); cudaFree(ctx->d_bsum); cudaFree(ctx->d_hidx); cudaFree(ctx->d_s1);
        cudaFree(ctx->d_crc_part); cudaFree(ctx->d_crc_part2); cudaFree(ctx->d_crc_misc); cudaFree(ctx->d_desc);
    }
    cudaFree(ctx->d_tab); cudaFree(ctx->d_gsum); cudaFree(ctx->d_filt); cudaFree(ctx->d_claim); cudaFree(ctx->d_sh); cudaFree(ctx->d_st);
    cudaFree(ctx->d_toks); cudaFree(ctx->d_a0); cudaFree(ctx->d_a1); cudaFree(ctx->d_off0); cudaFree(ctx->d_off1);
    cudaFree(ctx->d_bsumscan); cudaFree(ctx->d_tot); cudaFree(ctx->d_dbg); cudaFree(ctx->d_s0);
    if (ctx->h_crc) cudaFreeHost(ctx->h_crc);
    if (ctx->h_mb) cudaFreeHost(ctx->h_mb);
    for (auto &e : ctx->ev)
        if (e) cudaEventDestroy(e);
    if (ctx->stream) cudaStreamDestroy(ctx->stream);
    delete ctx;
}

static int ctx_create_impl(mrz_ctx *ctx, int device, i64 max_chunk, const mrz_config *cfg, mrz_ctx *parent) {
    CK(cudaSetDevice(device));
    ctx->parent = parent;
    ctx->device = device;
    ctx->max_chunk = max_chunk;
    if (cfg) ctx->cfg = *cfg; else mrz_config_default(&ctx->cfg);
    mrz_config &c = ctx->cfg;
    if (c.cluster_size != 0 && c.cluster_size != 1 && c.cluster_size != 2 && c.cluster_size != 4 && c.cluster_size != 8 && c.cluster_size != 16) {
        ctx->err = "mrz_config.cluster_size must be 0, 1, 2, 4, 8 or 16"; return MRZ_E_ARG;
    }
    if (c.threads == 0) c.threads = 256;
    if (c.threads < 32 || c.threads > MRZ_MAXT || (c.threads & 31)) { ctx->err = "mrz_config.threads must be a multiple of 32 in 32..512"; return MRZ_E_ARG; }
    if (c.filter_log2_words == 0) c.filter_log2_words = 24;
    if (c.filter_log2_words < 10 || c.filter_log2_words > 26) { ctx->err = "mrz_config.filter_log2_words must be in 10..26"; return MRZ_E_ARG; }
    if (c.token_divisor == 0) c.token_divisor = 1024;
    if (c.token_divisor < 31) { ctx->err = "mrz_config.token_divisor must be >= 31"; return MRZ_E_ARG; }
    {
        int sms = 0;
        CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device));
        ctx->sms = sms > 0 ? sms : 1;
    }
    crc_host_init();
    CK(cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking));
    for (auto &e : ctx->ev) CK(cudaEventCreate(&e));
    if (mrz_crc_ensure_device(device, ctx->stream)) { ctx->err = "cannot upload the CRC tables"; return MRZ_E_CUDA; }
    if (parent) {
        ctx->d_buf_alloc = parent->d_buf_alloc; ctx->d_buf = parent->d_buf;
        ctx->d_bmap = parent->d_bmap; ctx->d_bsum = parent->d_bsum; ctx->d_hidx = parent->d_hidx;
        ctx->d_crc_part = parent->d_crc_part; ctx->d_crc_part2 = parent->d_crc_part2; ctx->d_crc_misc = parent->d_crc_misc;
        ctx->d_s1 = parent->d_s1;
        ctx->d_desc = nullptr;
    } else {
        DA(ctx->d_buf_alloc, (size_t)max_chunk + 2 * MRZ_PAD + 64);
        ctx->d_buf = ctx->d_buf_alloc + MRZ_PAD;
        CK(cudaMemsetAsync(ctx->d_buf_alloc, 0, MRZ_PAD, ctx->stream));
        const i64 tiles = mrz_tagmap_tiles(max_chunk);
        DA(ctx->d_bmap, (size_t)tiles * MAP_TILE + 64);
        DA(ctx->d_bsum, (size_t)tiles * (MAP_TILE / MAP_BLK) + 64);
        DA(ctx->d_hidx, 256 * 8);
        DA(ctx->d_s1, (size_t)max_chunk + 64);
        i64 nblk = max_chunk / 16384 + 1;
        DA(ctx->d_crc_part, (size_t)nblk * 4);
        DA(ctx->d_crc_part2, (size_t)(nblk / 32 + 1) * 4);
        DA(ctx->d_crc_misc, 16);
        DA(ctx->d_desc, sizeof(AutoDesc) * MAX_LAUNCH);
    }
    ctx->s1_cap = max_chunk + 64;
    ctx->upload_gen = 1; ctx->map_gen = 0; ctx->map_n = -1;
    ctx->tab_slots = (i64)64 * 65536;  // largest levels[].mb_used = 64 MiB
    DA(ctx->d_tab, (size_t)(ctx->tab_slots + 32) * 16);  // + slack for 8-slot probe loads
    DA(ctx->d_gsum, (size_t)(ctx->tab_slots / 8 / 2 + 1024) * 4);   // 16-bit thermometer per 8 slots (automaton.cu GRP)
    ctx->filt_bits = c.filter_log2_words;
    DA(ctx->d_filt, (size_t)8 << ctx->filt_bits);
    DA(ctx->d_claim, (size_t)(ctx->tab_slots + 32) * 4 * 2);    // two arrays (automaton.cu: alternating epochs)
    DA(ctx->d_sh, mrz_shared_bytes());
    DA(ctx->d_st, sizeof(AutoState));
    DA(ctx->d_tot, sizeof(EmitTotals));
    if (c.profile) DA(ctx->d_dbg, MRZ_NDBG * 8);
    i64 nblk = max_chunk / 16384 + 1;
    ctx->h_crc_cap = nblk / 32 + 64;
    CK(cudaHostAlloc((void **)&ctx->h_crc, (size_t)ctx->h_crc_cap * 4, cudaHostAllocDefault));
    CK(cudaHostAlloc((void **)&ctx->h_mb, sizeof(*ctx->h_mb), cudaHostAllocDefault));
    int rc;
    if ((rc = ensure_tokens(ctx, max_chunk / c.token_divisor + 1024))) return rc;
    if (!parent) {   // a variant allocates its emission buffers only if it is the one chosen to emit
        if ((rc = ensure_items(ctx, max_chunk / 256 + 1024))) return rc;
        if ((rc = ensure_s0(ctx, max_chunk / 32 + 4096))) return rc;
    }
    CK(cudaStreamSynchronize(ctx->stream));   // every initialisation above has executed before the context is handed out
    return 0;
}

extern "C" int mrz_ctx_create_ex(int device, int64_t max_chunk, const mrz_config *cfg, mrz_ctx **out) {
    if (!out || max_chunk < 0) return MRZ_E_ARG;
    *out = nullptr;
    int n = mrz_device_count();
    if (n <= 0) {
        if (n == 0) g_err = "no CUDA device";
        return MRZ_E_NODEVICE;
    }
    if (device < 0 || device >= n) { g_err = "bad device index"; return MRZ_E_ARG; }
    mrz_ctx *ctx = new mrz_ctx();
    int rc = ctx_create_impl(ctx, device, max_chunk, cfg, nullptr);
    if (rc) {
        g_err = ctx->err;
        mrz_ctx_destroy(ctx);
        return rc;
    }
    *out = ctx;
    return 0;
}
extern "C" int mrz_ctx_set_cluster(mrz_ctx *ctx, int cluster_size) {
    if (!ctx) return MRZ_E_ARG;
    if (cluster_size != 0 && cluster_size != 1 && cluster_size != 2 && cluster_size != 4 && cluster_size != 8 && cluster_size != 16) return MRZ_E_ARG;
    ctx->cfg.cluster_size = cluster_size;
    return 0;
}

extern "C" int mrz_ctx_create(int device, int64_t max_chunk, mrz_ctx **out) { return mrz_ctx_create_ex(device, max_chunk, nullptr, out); }

extern "C" int mrz_ctx_create_variant(mrz_ctx *parent, mrz_ctx **out) {
    if (!out || !parent || parent->parent) return MRZ_E_ARG;
    *out = nullptr;
    mrz_ctx *ctx = new mrz_ctx();
    int rc = ctx_create_impl(ctx, parent->device, parent->max_chunk, &parent->cfg, parent);
    if (rc) {
        g_err = ctx->err;
        mrz_ctx_destroy(ctx);
        return rc;
    }
    *out = ctx;
    return 0;
}

// ---- upload / download ---------------------------------------------------------------------------
extern "C" int mrz_upload_part(mrz_ctx *ctx, int64_t offset, const uint8_t *src, int64_t len) {
    if (!ctx || ctx->parent || offset < 0 || len < 0 || offset + len > ctx->max_chunk || (len && !src)) return MRZ_E_ARG;
    CK(cudaSetDevice(ctx->device));
    if (len) CK(cudaMemcpyAsync(ctx->d_buf + offset, src, (size_t)len, cudaMemcpyHostToDevice, ctx->stream));
    return 0;
}
extern "C" int mrz_upload_wait(mrz_ctx *ctx) {
    if (!ctx) return MRZ_E_ARG;
    CK(cudaSetDevice(ctx->device));
    CK(cudaStreamSynchronize(ctx->stream));
    return 0;
}
extern "C" int mrz_upload_finish(mrz_ctx *ctx, int64_t n) {
    if (!ctx || ctx->parent || n < 0 || n > ctx->max_chunk) return MRZ_E_ARG;
    CK(cudaSetDevice(ctx->device));
    CK(cudaMemsetAsync(ctx->d_buf + n, 0, MRZ_PAD + 64, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    ctx->upload_gen++;
    return 0;
}
extern "C" int mrz_upload(mrz_ctx *ctx, const uint8_t *chunk, int64_t n) {
    if (!ctx || n < 0 || n > ctx->max_chunk || (n && !chunk)) return MRZ_E_ARG;
    int rc = mrz_upload_part(ctx, 0, chunk, n);
    if (rc) return rc;
    return mrz_upload_finish(ctx, n);
}

// The chunk CRC32 from the partial CRCs the CRC kernels left in the root's buffers (mrz_launch_crc): drains the stream.
static int combine_crc(mrz_ctx *ctx, i64 n, cudaStream_t s, u32 *crc_out) {
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
    *crc_out = raw ^ crc_shift(0xFFFFFFFFu, (u64)n) ^ 0xFFFFFFFFu;
    return 0;
}

static mrz_ctx *root_of(mrz_ctx *c) { return c->parent ? c->parent : c; }

// tag map + CRC of the resident chunk of `root`, unless they are already valid for (upload, n, hash_index)
static int ensure_map(mrz_ctx *ctx, mrz_ctx *root, i64 n, const mrz_params *prm, cudaStream_t s, int *launches) {
    if (root->map_gen == root->upload_gen && root->map_n == n && !memcmp(root->map_hidx, prm->hash_index, sizeof root->map_hidx)) return 0;
    CK(cudaMemcpyAsync(root->d_hidx, prm->hash_index, 256 * 8, cudaMemcpyHostToDevice, s));
    mrz_launch_tagmap(root->d_buf, root->d_hidx, n, root->d_bmap, root->d_bsum, root->sms, s);
    (*launches)++;
    mrz_launch_crc(root->d_buf, n, root->d_crc_part, root->d_crc_part2, root->d_crc_misc, s, launches);   // independent of the automaton
    CK(cudaGetLastError());
    root->map_gen = root->upload_gen; root->map_n = n;
    memcpy(root->map_hidx, prm->hash_index, sizeof root->map_hidx);
    return 0;
}

static void init_state(AutoState &h, const mrz_ctx *c, i64 n, const mrz_params *prm) {
    memset(&h, 0, sizeof h);
    const int L = prm->level;
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
    h.tok_cap = c->tok_cap;
    h.filt_bits = c->filt_bits;
    h.done = h.end <= 0 ? 1 : 0;
}

static int check_run_args(const mrz_ctx *ctx, i64 n, const mrz_params *prm) {
    if (!ctx || !prm || n < 0 || n > ctx->max_chunk) return MRZ_E_ARG;
    if (prm->level < 0 || prm->level > 9 || prm->chunk_bytes < 1 || prm->chunk_bytes > 8) return MRZ_E_ARG;
    if (prm->victim_round_in < 0 || prm->victim_round_in >= kLevels[prm->level].chain) return MRZ_E_ARG;
    return 0;
}

// CTAs per cluster for `count` automata in one launch: the largest size of which `count` clusters are resident at once
static int pick_cluster(const mrz_ctx *ctx, int count) {
    if (ctx->cfg.cluster_size) return ctx->cfg.cluster_size;
#ifdef MRZ_CUSIM
    return count > 2 ? 1 : 2;
#else
    // measured on B200 (profiles/r2_geometry_*.txt): 8 CTAs x 256 threads beats 16 x 128 / 16 x 256 / 8 x 512 for one automaton
    for (int cl = 8; cl > 1; cl >>= 1) {
        int m = mrz_max_active_clusters(cl, ctx->cfg.threads);
        if (m >= count) return cl;
    }
    return 1;
#endif
}

// The automata of `count` contexts (same device) in ONE launch.  An automaton whose token buffer overflowed is redone
// with a larger buffer.  On return every context holds its final AutoState and token list.
static int run_phase_a_many(mrz_ctx *const *ctxs, int count, const i64 *ns, const mrz_params *prms) {
    mrz_ctx *ctx = ctxs[0];
    cudaStream_t s = ctx->stream;
    mrz_ctx *droot = root_of(ctx);           // its descriptor array and mailbox serve the launch
    if (count < 1 || count > MAX_LAUNCH) { ctx->err = "too many automata in one launch"; return MRZ_E_ARG; }
    std::vector<int> todo((size_t)count);
    for (int i = 0; i < count; i++) { todo[(size_t)i] = i; ctxs[i]->have_tokens = false; }
    int launches = 0;
    float ms_map = 0, ms_auto = 0;
    CK(cudaEventRecord(ctx->ev[0], s));
    for (int i = 0; i < count; i++) {         // tag map + CRC once per distinct chunk
        int rc = ensure_map(ctx, root_of(ctxs[i]), ns[i], &prms[i], s, &launches);
        if (rc) return rc;
    }
    CK(cudaEventRecord(ctx->ev[1], s));
    const int T = ctx->cfg.threads;
    int CL = 1;
    while (!todo.empty()) {
        const int m = (int)todo.size();
        CL = pick_cluster(ctx, m);
        for (int q = 0; q < m; q++) {
            mrz_ctx *c = ctxs[todo[(size_t)q]];
            mrz_ctx *r = root_of(c);
            AutoState &h = droot->h_mb->st[q];
            init_state(h, c, ns[todo[(size_t)q]], &prms[todo[(size_t)q]]);
            AutoDesc &d = droot->h_mb->desc[q];
            d.buf = r->d_buf; d.bmap = r->d_bmap; d.bsum = r->d_bsum; d.hidx = r->d_hidx;
            d.ttag = c->d_tab; d.toff = c->d_tab + c->tab_slots + 32; d.filt = c->d_filt; d.gsum = c->d_gsum; d.claim = c->d_claim; d.claim_stride = c->tab_slots + 32;
            d.toks = c->d_toks; d.sh = c->d_sh; d.st = c->d_st; d.dbg = c->d_dbg;
            CK(cudaMemcpyAsync(c->d_st, &h, sizeof h, cudaMemcpyHostToDevice, s));
            CK(cudaMemsetAsync(c->d_tab, 0, (size_t)(c->tab_slots + 32) * 16, s));  // rzip.c:518-519
            CK(cudaMemsetAsync(c->d_gsum, 0, (size_t)(c->tab_slots / 8 / 2 + 1024) * 4, s));
            CK(cudaMemsetAsync(c->d_filt, 0, (size_t)8 << c->filt_bits, s));
            CK(cudaMemsetAsync(c->d_claim, 0xff, (size_t)(c->tab_slots + 32) * 4 * 2, s));
            if (c->d_dbg) CK(cudaMemsetAsync(c->d_dbg, 0, MRZ_NDBG * 8, s));
        }
        CK(cudaMemcpyAsync(droot->d_desc, droot->h_mb->desc, sizeof(AutoDesc) * (size_t)m, cudaMemcpyHostToDevice, s));
        CK(cudaEventRecord(ctx->ev[2], s));
        bool any = false;
        for (int q = 0; q < m; q++) any = any || !droot->h_mb->st[q].done;
        if (any) {
            // a lone automaton: its table (tags, then offsets) should stay in L2
            mrz_ctx *c0 = ctxs[todo[0]];
            const bool pin = m == 1 && ctx->cfg.l2_policy == 0;
            int lrc = mrz_launch_automata(droot->d_desc, m, CL, T, s, pin ? c0->d_tab : nullptr, pin ? (size_t)(c0->tab_slots + 32) * 16 : 0);
            if (lrc) CK((cudaError_t)lrc);
            launches++;
        }
        CK(cudaEventRecord(ctx->ev[3], s));
        for (int q = 0; q < m; q++)
            CK(cudaMemcpyAsync(&droot->h_mb->st[q], ctxs[todo[(size_t)q]]->d_st, sizeof(AutoState), cudaMemcpyDeviceToHost, s));
        CK(cudaStreamSynchronize(s));
        CK(cudaGetLastError());
        float b = 0;
        cudaEventElapsedTime(&b, ctx->ev[2], ctx->ev[3]);
        ms_auto += b;
        std::vector<int> again;
        for (int q = 0; q < m; q++) {
            mrz_ctx *c = ctxs[todo[(size_t)q]];
            const AutoState &h = droot->h_mb->st[q];
            if (h.error == 2) {              // token buffer full: redo this automaton with a larger one
                i64 n = ns[todo[(size_t)q]];
                i64 cap = c->tok_cap * 8, hard = n / MRZ_MIN_MATCH + 16;   // bounded by one match per 31 bytes
                if (cap > hard) cap = hard;
                if (cap <= c->tok_cap) { ctx->err = c->err = "token capacity exhausted"; return MRZ_E_INTERNAL; }
                int rc = ensure_tokens(c, cap);
                if (rc) { ctx->err = c->err; return rc; }
                again.push_back(todo[(size_t)q]);
                continue;
            }
            if (h.error || !h.done) {
                char msg[192];
                snprintf(msg, sizeof msg, "device automaton error %d at p=%lld (index %lld, bound %lld)", h.error, (long long)h.p,
                         (long long)h.err_a, (long long)h.err_b);
                ctx->err = c->err = msg;
                return MRZ_E_INTERNAL;
            }
            c->last_state = h;
            c->last_cb = prms[todo[(size_t)q]].chunk_bytes;
            c->last_n = ns[todo[(size_t)q]]; c->last_ntok = h.ntok; c->last_s0 = c->last_s1 = 0;
            c->have_tokens = true;
            c->geom[0] = CL; c->geom[1] = T; c->geom[2] = m;
        }
        todo.swap(again);
    }
    for (int i = 0; i < count; i++) {
        mrz_ctx *c = ctxs[i];
        if (c->d_dbg) {
            CK(cudaMemcpyAsync(droot->h_mb->dbg, c->d_dbg, MRZ_NDBG * 8, cudaMemcpyDeviceToHost, s));
            CK(cudaStreamSynchronize(s));
            memcpy(c->h_dbg, droot->h_mb->dbg, sizeof c->h_dbg);
        }
    }
    CK(cudaEventRecord(ctx->ev[4], s));
    CK(cudaStreamSynchronize(s));
    float a = 0, tot = 0;
    cudaEventElapsedTime(&a, ctx->ev[0], ctx->ev[1]);
    cudaEventElapsedTime(&tot, ctx->ev[0], ctx->ev[4]);
    ms_map = a;
    for (int i = 0; i < count; i++) {
        mrz_ctx *c = ctxs[i];
        c->ms_map = ms_map; c->ms_auto = ms_auto; c->ms_phase_a = tot; c->launches = launches;
    }
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
    res->ms_tagscan = ctx->ms_map;
    res->ms_automaton = ctx->ms_auto;
    res->gpu_launches = ctx->launches;
}

// Emission of the tokens the last automaton run of this context left behind.
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
    mrz_launch_exclusive_scan(ctx->d_a0, ctx->d_off0, items, ctx->d_bsumscan, ctx->d_off0 + items, s, &launches);
    mrz_launch_exclusive_scan(ctx->d_a1, ctx->d_off1, items, ctx->d_bsumscan, ctx->d_off1 + items, s, &launches);
    i64 *totals = ctx->h_mb->totals;
    CK(cudaMemcpyAsync(&totals[0], ctx->d_off0 + items, 8, cudaMemcpyDeviceToHost, s));
    CK(cudaMemcpyAsync(&totals[1], ctx->d_off1 + items, 8, cudaMemcpyDeviceToHost, s));
    // CRC partials -> host combine
    u32 crc = 0;
    if ((rc = combine_crc(ctx, n, s, &crc))) return rc;
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

extern "C" int mrz_run_many(mrz_ctx *const *ctxs, int count, const int64_t *n, const mrz_params *prms, mrz_result *res) {
    if (!ctxs || !n || !prms || !res || count < 1 || count > MAX_LAUNCH) return MRZ_E_ARG;
    for (int i = 0; i < count; i++) {
        if (check_run_args(ctxs[i], n[i], &prms[i])) return MRZ_E_ARG;
        if (ctxs[i]->device != ctxs[0]->device) return MRZ_E_ARG;
        for (int k = 0; k < i; k++)
            if (ctxs[k] == ctxs[i]) return MRZ_E_ARG;
    }
    mrz_ctx *ctx = ctxs[0];
    CK(cudaSetDevice(ctx->device));
    int rc = run_phase_a_many(ctxs, count, (const i64 *)n, prms);
    if (rc) { g_err = ctx->err; return rc; }
    for (int i = 0; i < count; i++) fill_automaton_result(ctxs[i], &res[i]);
    return 0;
}

extern "C" int mrz_run_automaton(mrz_ctx *ctx, int64_t n, const mrz_params *prm, mrz_result *res) {
    if (!res || check_run_args(ctx, n, prm)) return MRZ_E_ARG;
    return mrz_run_many(&ctx, 1, &n, prm, res);
}

extern "C" int mrz_run_resident(mrz_ctx *ctx, int64_t n, const mrz_params *prm, mrz_result *res) {
    if (!res || check_run_args(ctx, n, prm)) return MRZ_E_ARG;
    int rc = mrz_run_many(&ctx, 1, &n, prm, res);
    if (rc) return rc;
    return run_phase_b(ctx, res);
}

extern "C" int mrz_run_emit(mrz_ctx *ctx, mrz_result *res) {
    if (!ctx || !res) return MRZ_E_ARG;
    CK(cudaSetDevice(ctx->device));
    return run_phase_b(ctx, res);
}

// Variant v runs with victim_round_in = v on its own table and cluster: one launch for all of them.
extern "C" int mrz_run_variants(mrz_ctx *const *ctxs, int nvar, int64_t n, const mrz_params *prm, mrz_result *res) {
    if (!ctxs || !prm || !res || nvar < 1 || nvar > MAX_LAUNCH || prm->level < 0 || prm->level > 9 || nvar > kLevels[prm->level].chain) return MRZ_E_ARG;
    std::vector<mrz_params> p((size_t)nvar, *prm);
    std::vector<int64_t> ns((size_t)nvar, n);
    for (int v = 0; v < nvar; v++) p[(size_t)v].victim_round_in = v;
    return mrz_run_many(ctxs, nvar, ns.data(), p.data(), res);
}

extern "C" int mrz_download_part(mrz_ctx *ctx, int streamno, int64_t offset, uint8_t *dst, int64_t len) {
    if (!ctx || (streamno != 0 && streamno != 1) || offset < 0 || len < 0) return MRZ_E_ARG;
    const i64 have = streamno ? ctx->last_s1 : ctx->last_s0;
    if (offset + len > have) return MRZ_E_CAPACITY;
    CK(cudaSetDevice(ctx->device));
    if (len) CK(cudaMemcpyAsync(dst, (streamno ? ctx->d_s1 : ctx->d_s0) + offset, (size_t)len, cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
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

extern "C" int mrz_get_profile(mrz_ctx *ctx, int64_t counters[64], int32_t geom[3]) {
    if (!ctx) return MRZ_E_ARG;
    if (counters) memcpy(counters, ctx->h_dbg, sizeof ctx->h_dbg);
    if (geom) { geom[0] = ctx->geom[0]; geom[1] = ctx->geom[1]; geom[2] = ctx->geom[2]; }
    return 0;
}

// ---- runzip_chunk on the device (runzip.c:226-327): the two streams in, the chunk out ------------------------------
// The header list (stream 0) is walked on the host - a header's size depends on its first byte - into literal runs and
// matches with their places in the chunk; the literal runs are scattered by one kernel, the matches copied in stream
// order, wave by wave (a wave: consecutive matches none of which reads a byte another match of the wave writes).
extern "C" int mrz_runzip_chunk(mrz_ctx *ctx, const uint8_t *s0, int64_t s0_len, const uint8_t *s1, int64_t s1_len, int chunk_bytes,
                                uint8_t *out, int64_t out_cap, mrz_unzip_result *res) {
    if (!ctx || ctx->parent || !res || !s0 || s0_len < 7 || s1_len < 0 || (s1_len && !s1) || chunk_bytes < 1 || chunk_bytes > 8)
        return MRZ_E_ARG;
    memset(res, 0, sizeof *res);
    CK(cudaSetDevice(ctx->device));
    std::vector<UnzipItem> lits, mats;
    std::vector<i64> wave_first;        // index of the first match of every wave
    i64 i = 0, j = 0, pos = 0;
    bool ended = false;
    while (i + 3 <= s0_len) {
        const int head = s0[i];
        const i64 len = (i64)s0[i + 1] | ((i64)s0[i + 2] << 8);
        if (head == 0) {                // runzip.c:120-147 unzip_literal
            i += 3;
            if (len == 0) { ended = true; break; }      // the loop condition of runzip.c:276
            if (j + len > s1_len) { ctx->err = "stream 1 is shorter than the literal headers say"; return MRZ_E_DATA; }
            lits.push_back(UnzipItem{pos, j, len});
            j += len;
        } else {                        // runzip.c:159-207 unzip_match
            if (i + 3 + chunk_bytes > s0_len) break;
            i64 ofs = 0;
            for (int b = 0; b < chunk_bytes; b++) ofs |= (i64)s0[i + 3 + b] << (8 * b);
            i += 3 + chunk_bytes;
            if (len < 1 || ofs < 1 || ofs > pos) { ctx->err = "a match reaches in front of the chunk (corrupt archive)"; return MRZ_E_DATA; }
            // does it read what a match of the current wave writes?  (their destinations ascend)
            const i64 rd_lo = pos - ofs, rd_hi = rd_lo + (len < ofs ? len : ofs);
            const i64 w0 = wave_first.empty() ? 0 : wave_first.back();
            bool clash = wave_first.empty();
            if (!clash) {
                i64 lo = w0, hi = (i64)mats.size();     // first match of the wave whose end is behind rd_lo
                while (lo < hi) {
                    const i64 mid = (lo + hi) >> 1;
                    if (mats[(size_t)mid].dst + mats[(size_t)mid].len > rd_lo) hi = mid; else lo = mid + 1;
                }
                clash = lo < (i64)mats.size() && mats[(size_t)lo].dst < rd_hi;
            }
            if (clash) wave_first.push_back((i64)mats.size());
            mats.push_back(UnzipItem{pos, rd_lo, len});
        }
        pos += len;
        if (pos > ctx->max_chunk) { ctx->err = "the streams describe a chunk larger than this context"; return MRZ_E_CAPACITY; }
    }
#ifdef MRZ_CUSIM
    // emulator build (its CTAs run one after the other, which would hide a wrong wave): no match of a wave may read a byte
    // that another match of the same wave writes - checked against every pair
    for (size_t w = 0; w < wave_first.size(); w++) {
        const i64 a = wave_first[w], b = w + 1 < wave_first.size() ? wave_first[w + 1] : (i64)mats.size();
        for (i64 p = a; p < b; p++)
            for (i64 q = a; q < b; q++) {
                if (p == q) continue;
                const UnzipItem &rd = mats[(size_t)p], &wr = mats[(size_t)q];
                const i64 ofs = rd.dst - rd.src, rd_hi = rd.src + (rd.len < ofs ? rd.len : ofs);
                if (rd.src < wr.dst + wr.len && wr.dst < rd_hi) { fprintf(stderr, "unzip: wave %zu holds dependent matches %lld and %lld\n", w, (long long)p, (long long)q); abort(); }
            }
    }
#endif
    if (!ended || i + 4 > s0_len) { ctx->err = "stream 0 ends without terminator and checksum"; return MRZ_E_DATA; }
    if (j != s1_len) { ctx->err = "stream 1 is longer than the literal headers say"; return MRZ_E_DATA; }
    const u32 stored = ((u32)s0[i] << 24) | ((u32)s0[i + 1] << 16) | ((u32)s0[i + 2] << 8) | (u32)s0[i + 3];   // rzip.c:662-665
    const i64 n = pos;
    if (n > out_cap) return MRZ_E_CAPACITY;
    cudaStream_t s = ctx->stream;
    UnzipItem *d_items = nullptr;
    const size_t nl = lits.size(), nm = mats.size();
    if (nl + nm) {
        if (cudaMalloc((void **)&d_items, (nl + nm) * sizeof(UnzipItem)) != cudaSuccess) { cudaGetLastError(); ctx->err = "cudaMalloc of the item table failed"; return MRZ_E_NOMEM; }
    }
    int launches = 0;
    cudaError_t e = cudaSuccess;
    auto step = [&](cudaError_t r) { if (e == cudaSuccess) e = r; };
    if (nl) step(cudaMemcpyAsync(d_items, lits.data(), nl * sizeof(UnzipItem), cudaMemcpyHostToDevice, s));
    if (nm) step(cudaMemcpyAsync(d_items + nl, mats.data(), nm * sizeof(UnzipItem), cudaMemcpyHostToDevice, s));
    if (s1_len) step(cudaMemcpyAsync(ctx->d_s1, s1, (size_t)s1_len, cudaMemcpyHostToDevice, s));
    step(cudaEventRecord(ctx->ev[0], s));
    if (e == cudaSuccess) {
        if (nl) { mrz_launch_unzip_literals(d_items, (i64)nl, ctx->d_s1, ctx->d_buf, ctx->sms, s); launches++; }
        for (size_t w = 0; w < wave_first.size(); w++) {
            const i64 a = wave_first[w], b = w + 1 < wave_first.size() ? wave_first[w + 1] : (i64)nm;
            mrz_launch_unzip_matches(d_items + nl, a, b - a, ctx->d_buf, ctx->sms, s);
            launches++;
        }
        step(cudaMemsetAsync(ctx->d_buf + n, 0, MRZ_PAD + 64, s));
        mrz_launch_crc(ctx->d_buf, n, ctx->d_crc_part, ctx->d_crc_part2, ctx->d_crc_misc, s, &launches);
        step(cudaEventRecord(ctx->ev[1], s));
        step(cudaGetLastError());
    }
    u32 crc = 0;
    int rc = 0;
    if (e == cudaSuccess) rc = combine_crc(ctx, n, s, &crc);
    if (e == cudaSuccess && !rc && n) step(cudaMemcpyAsync(out, ctx->d_buf, (size_t)n, cudaMemcpyDeviceToHost, s));
    step(cudaStreamSynchronize(s));
    if (d_items) cudaFree(d_items);
    ctx->upload_gen++;                  // the resident chunk changed: its tag map is stale
    ctx->have_tokens = false;
    CK(e);
    if (rc) return rc;
    float ms = 0;
    cudaEventElapsedTime(&ms, ctx->ev[0], ctx->ev[1]);
    res->out_len = n;
    res->literals = (i64)nl;
    res->matches = (i64)nm;
    res->waves = (i64)wave_first.size();
    res->crc32 = crc;
    res->crc32_stored = stored;
    res->ms_kernels = ms;
    res->gpu_launches = launches;
    if (crc != stored) {
        char b[128];
        snprintf(b, sizeof b, "Bad checksum: 0x%08x - expected: 0x%08x", crc, stored);   // runzip.c:319
        ctx->err = b;
        return MRZ_E_DATA;
    }
    return 0;
}

extern "C" int mrz_debug_tags(mrz_ctx *ctx, int64_t n, const int64_t hash_index[256], int64_t *tags_out,
                              int64_t count) {
    if (!ctx || count < 0 || count + 30 > n) return MRZ_E_ARG;
    CK(cudaSetDevice(ctx->device));
    i64 *d_tags = nullptr;
    CK(cudaMalloc(&d_tags, (size_t)(count + 1) * 8));
    mrz_ctx *root = root_of(ctx);
    cudaError_t e = cudaMemcpyAsync(root->d_hidx, hash_index, 256 * 8, cudaMemcpyHostToDevice, ctx->stream);
    root->map_gen = 0;   // d_hidx no longer matches the map
    if (e == cudaSuccess) {
        mrz_launch_debug_tags(ctx->d_buf, root->d_hidx, count, d_tags, ctx->stream);
        e = cudaMemcpyAsync(tags_out, d_tags, (size_t)count * 8, cudaMemcpyDeviceToHost, ctx->stream);
    }
    if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
    cudaFree(d_tags);
    CK(e);
    return 0;
}

extern "C" int mrz_debug_tagmap(mrz_ctx *ctx, int64_t n, const int64_t hash_index[256], uint8_t *map_out, uint8_t *summary_out) {
    if (!ctx || n < 0 || n > ctx->max_chunk || !hash_index) return MRZ_E_ARG;
    CK(cudaSetDevice(ctx->device));
    mrz_ctx *root = root_of(ctx);
    mrz_params prm;
    memset(&prm, 0, sizeof prm);
    memcpy(prm.hash_index, hash_index, sizeof prm.hash_index);
    int launches = 0;
    root->map_gen = 0;
    int rc = ensure_map(ctx, root, n, &prm, ctx->stream, &launches);
    if (rc) return rc;
    if (map_out && n) CK(cudaMemcpyAsync(map_out, root->d_bmap, (size_t)n, cudaMemcpyDeviceToHost, ctx->stream));
    if (summary_out && n) CK(cudaMemcpyAsync(summary_out, root->d_bsum, (size_t)((n + 63) / 64), cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    return 0;
}

extern "C" int mrz_debug_crc32(mrz_ctx *ctx, int64_t n, uint32_t *crc_out) {
    if (!ctx || !crc_out) return MRZ_E_ARG;
    (void)n;
    CK(cudaSetDevice(ctx->device));
    CK(cudaMemcpyAsync(&ctx->h_mb->crc, ctx->d_crc_misc + 1, 4, cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    *crc_out = ctx->h_mb->crc;
    return 0;
}

extern "C" int mrz_debug_tokens(mrz_ctx *ctx, int64_t *triples, int64_t cap, int64_t *count) {
    if (!ctx || !count) return MRZ_E_ARG;
    *count = ctx->last_ntok;
    if (ctx->last_ntok > cap) return MRZ_E_CAPACITY;
    CK(cudaSetDevice(ctx->device));
    if (ctx->last_ntok) {
        CK(cudaMemcpyAsync(triples, ctx->d_toks, (size_t)ctx->last_ntok * sizeof(Token), cudaMemcpyDeviceToHost, ctx->stream));
        CK(cudaStreamSynchronize(ctx->stream));
    }
    return 0;
}
