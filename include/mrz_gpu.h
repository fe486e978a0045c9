/* mrz_gpu.h — C-ABI of the B200 rzip pre-processor (libmrz_gpu.so).
 *
 * This is the drop-in boundary for modern-rzip's hot path.  Each entry point replaces a piece of
 * /root/reference/src/rzip.c as cited; plain pointers and sizes only, no CUDA or torch types.
 * The host side (modern-rzip_b200/host/rzip_fd.c) keeps the reference's
 *     void rzip_fd(rzip_control *control, int fd_in, int fd_out)         (include/rzip.h:25)
 * and calls into this library where the reference calls hash_search()     (src/rzip.c:507, :773).
 *
 * All functions return 0 on success or a negative MRZ_E_* code; mrz_last_error() gives text.
 * There is NO CPU fallback: without a CUDA device every compute entry point fails with
 * MRZ_E_NODEVICE.  The library reads no environment variables; everything tunable is in mrz_config.
 *
 * Threading: calls on DIFFERENT contexts may be made from different host threads at the same time
 * (that is how the host overlaps the upload of chunk k+1 with the search of chunk k, and drives
 * several GPUs); calls on one context must not overlap.
 */
#ifndef MRZ_GPU_H
#define MRZ_GPU_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

#define MRZ_OK 0
#define MRZ_E_NODEVICE (-1)   /* no CUDA device / driver */
#define MRZ_E_CUDA (-2)       /* a CUDA call failed; see mrz_last_error */
#define MRZ_E_ARG (-3)        /* bad argument */
#define MRZ_E_NOMEM (-4)      /* device or pinned-host allocation failed */
#define MRZ_E_CAPACITY (-5)   /* caller's output buffer too small */
#define MRZ_E_INTERNAL (-6)   /* device automaton reported an inconsistency */
#define MRZ_E_DATA (-7)       /* mrz_runzip_chunk: the streams are inconsistent or the chunk's checksum does not match */

typedef struct mrz_ctx mrz_ctx;

/* Inputs that src/rzip.c reads from rzip_control / rzip_state (mrzip_private.h:388-416). */
typedef struct {
    int32_t level;            /* index into levels[] (rzip.c:65-73, :896): 0..9 */
    int32_t chunk_bytes;      /* st->chunk_bytes (rzip.c:1006-1008): width of match distances */
    int64_t victim_round_in;  /* the static in insert_hash (rzip.c:259), carried across chunks */
    int64_t hash_index[256];  /* st->hash_index (rzip.c:669-673) computed on the host with random() */
} mrz_params;

/* Outputs of hash_search: st->stats (rzip.c:1108-1115) and stream lengths. */
typedef struct {
    int64_t s0_len;           /* stream 0 bytes incl. terminator + 4 CRC bytes (rzip.c:664-665) */
    int64_t s1_len;           /* stream 1 (literal) bytes */
    int64_t matches, match_bytes, literals, literal_bytes, tag_hits, tag_misses, inserts;
    int64_t victim_round_out; /* value of the static after this chunk */
    int64_t victim_events;    /* number of victim evictions (0 => result independent of victim_round_in) */
    int64_t final_min_mask;   /* st->minimum_tag_mask at the end (the "Starting sweep for mask" trace) */
    int64_t n_ops;            /* positions that passed the lookup gate (rzip.c:573) */
    uint32_t crc32;           /* IEEE CRC32 of the chunk (rzip.c:662) */
    uint32_t reserved;
    double ms_kernels;        /* device time of the resident path (CUDA events), milliseconds */
    double ms_tagscan, ms_automaton, ms_emit;   /* its parts: tag map (+ CRC), automaton launch, emission */
    int64_t gpu_launches;     /* kernels launched for this chunk */
} mrz_result;

/* Engine configuration (replaces every environment knob; zero = default). */
typedef struct {
    int32_t cluster_size;     /* CTAs per automaton: 1,2,4,8,16; 0 = as many as fit for the automata of a launch */
    int32_t threads;          /* threads per CTA: 64..512, multiple of 32; 0 = 256 */
    int32_t filter_log2_words;/* tag filter size: 2^x 64-bit words, 16..26; 0 = 24 (128 MiB) */
    int32_t profile;          /* 1: collect the automaton's phase counters (mrz_get_profile) */
    int32_t token_divisor;    /* initial token capacity = max_chunk / divisor (a chunk that needs more is redone with 8x); 0 = 1024 */
    int32_t poison;           /* debugging: fill every work buffer with 0xA5 at creation instead of leaving it as allocated */
    int32_t l2_policy;        /* 0 = when ONE automaton runs, ask for its hash table to persist in L2 (access-policy window on the launch); 1 = off */
    int32_t reserved[1];
} mrz_config;

/* library / device ------------------------------------------------------------------------------ */
int mrz_device_count(void);                       /* >=0, or MRZ_E_NODEVICE */
const char *mrz_version(void);
const char *mrz_last_error(const mrz_ctx *ctx);   /* ctx may be NULL: last error of a failed create */
void mrz_config_default(mrz_config *cfg);

/* A context owns one device's buffers for chunks up to max_chunk bytes: the chunk and its tag map, one
 * automaton slot (hash table levels[].mb_used MiB, rzip.c:518-531, filter, claims, tokens) and the two
 * output streams. */
int mrz_ctx_create(int device, int64_t max_chunk, mrz_ctx **out);
int mrz_ctx_create_ex(int device, int64_t max_chunk, const mrz_config *cfg, mrz_ctx **out);
void mrz_ctx_destroy(mrz_ctx *ctx);
/* Change mrz_config.cluster_size of a context (and of launches it leads) after creation: a host that keeps several
 * chunks of one device in flight sizes the clusters so that all their automata are resident at once. */
int mrz_ctx_set_cluster(mrz_ctx *ctx, int cluster_size);

/* hash_search (rzip.c:507-667) for one chunk, host buffers in and out.
 * Copies the chunk H2D, runs the kernels, copies stream 0 / stream 1 D2H.
 * s0 needs up to n/2+64 bytes, s1 up to n bytes (pass capacities). */
int mrz_rzip_chunk(mrz_ctx *ctx, const uint8_t *chunk, int64_t n, const mrz_params *prm,
                   uint8_t *s0, int64_t s0_cap, uint8_t *s1, int64_t s1_cap, mrz_result *res);

/* The same in steps, for pipelining and for device-resident timing. */
int mrz_upload(mrz_ctx *ctx, const uint8_t *chunk, int64_t n);                 /* H2D, synchronous */
/* piecewise upload (read -> pinned slab -> H2D of slab i while slab i+1 is being read): parts may be queued
 * back to back; src must stay valid until mrz_upload_wait (a pinned src makes the copy asynchronous) */
int mrz_upload_part(mrz_ctx *ctx, int64_t offset, const uint8_t *src, int64_t len);
int mrz_upload_wait(mrz_ctx *ctx);                                              /* all queued parts have landed */
int mrz_upload_finish(mrz_ctx *ctx, int64_t n);                                /* pads the chunk; call once after the last part */
int mrz_run_resident(mrz_ctx *ctx, int64_t n, const mrz_params *prm, mrz_result *res); /* kernels only */
int mrz_download(mrz_ctx *ctx, uint8_t *s0, int64_t s0_cap, uint8_t *s1, int64_t s1_cap); /* D2H */
/* piecewise download of stream `streamno` (0/1) of the last emission: bytes [offset, offset+len) */
int mrz_download_part(mrz_ctx *ctx, int streamno, int64_t offset, uint8_t *dst, int64_t len);

/* Variant-parallel chunks (the multi-GPU / multi-chunk answer to the one scalar that links chunks: the
 * function-local `static victim_round` of insert_hash, rzip.c:259, which enters the next chunk and is
 * output-relevant as soon as that chunk evicts a victim, rzip.c:278-291).  Instead of waiting for the previous
 * chunk's value, the automaton of a resident chunk is run for EVERY possible victim_round_in (0 .. max_chain_len-1,
 * levels[] rzip.c:65-73) in ONE kernel launch — one thread-block cluster, one hash table and one token list per
 * variant, the tag map shared — and the caller keeps the variant the chain selects.
 *   mrz_ctx_create_variant: a context that shares the parent's chunk bytes, tag map and stream-1 buffer and owns
 *                           its own table / filter / claims / tokens (≈ 0.25 GB + tokens).
 *   mrz_run_automaton:      hash_search up to the token list (no streams): counters, victim_round_out, n_ops.
 *   mrz_run_emit:           put_literal/put_match/CRC (rzip.c:161-227, :617-666) for the tokens of the last
 *                           automaton run of this context; then mrz_download from the same context.
 *   mrz_run_variants:       ctxs[v] runs with victim_round_in = v (prm->victim_round_in is ignored); res[v] as
 *                           mrz_run_automaton.  ctxs[0] is the parent, the others its variants.
 *   mrz_run_many:           the general form: `count` automata in one launch, ctxs[i] with n[i] / prms[i]; contexts
 *                           may belong to different chunks (parents) of the same device. */
int mrz_ctx_create_variant(mrz_ctx *parent, mrz_ctx **out);
int mrz_run_automaton(mrz_ctx *ctx, int64_t n, const mrz_params *prm, mrz_result *res);
int mrz_run_emit(mrz_ctx *ctx, mrz_result *res);
int mrz_run_variants(mrz_ctx *const *ctxs, int nvar, int64_t n, const mrz_params *prm, mrz_result *res);
int mrz_run_many(mrz_ctx *const *ctxs, int count, const int64_t *n, const mrz_params *prms, mrz_result *res);

/* The way back: runzip_chunk (runzip.c:226-327) for one chunk of an uncompressed archive.
 * s0 / s1 are the chunk's two streams as the archive holds them (what mrz_rzip_chunk produced): stream 0 is walked on the
 * host (unzip headers, runzip.c:112-118), the literal runs (unzip_literal, runzip.c:120-147) are scattered and the
 * matches (unzip_match, runzip.c:159-207, `n = MIN(len, offset)` repeated) copied on the device, and the chunk's CRC32 is
 * checked against the four bytes behind stream 0's terminator (rzip.c:662-665; the reference checks it at
 * runzip.c:310-322 only for archives without MD5 - here always): a mismatch returns MRZ_E_DATA with the reference's
 * "Bad checksum" text in mrz_last_error.  out receives the chunk (out_cap >= its size, which is at most the context's
 * max_chunk); the context must be a root context.  The resident chunk of the context is replaced by the result. */
typedef struct mrz_unzip_result {
    int64_t out_len;          /* bytes of the chunk */
    int64_t literals, matches;/* headers of each kind in stream 0 */
    int64_t waves;            /* groups of independent matches (one kernel launch each) */
    uint32_t crc32, crc32_stored;
    float ms_kernels;
    int32_t gpu_launches;
} mrz_unzip_result;
int mrz_runzip_chunk(mrz_ctx *ctx, const uint8_t *s0, int64_t s0_len, const uint8_t *s1, int64_t s1_len, int chunk_bytes,
                     uint8_t *out, int64_t out_cap, mrz_unzip_result *res);

/* Pinned host memory for the e2e path (cudaHostAlloc); optional. */
void *mrz_host_alloc(int64_t bytes);
void mrz_host_free(void *p);

/* MRZ_NDBG (64) phase counters of the last automaton run of this context (cfg.profile = 1), and the launch
 * geometry it used: geom[0] = CTAs per cluster, geom[1] = threads per CTA, geom[2] = automata in the launch. */
int mrz_get_profile(mrz_ctx *ctx, int64_t counters[64], int32_t geom[3]);

/* CRC32 of n bytes that are already in device memory, without any allocation.
 * NO COUNTERPART IN THE REFERENCE: this entry point was added for callers that
 * cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  It is the
 * chunk checksum of the path (what gcry_md_write(crc_handle) accumulates) computed by the same kernels, with their
 * scratch (the per-block partial CRCs) in a fixed file-scope device array instead of a context's buffers:
 *   - n <= MRZ_NOALLOC_MAX_BYTES (1 GiB); larger sizes are rejected with MRZ_E_CAPACITY and a text in
 *     mrz_noalloc_last_error(); d_data must be 16-byte aligned; d_crc_out receives the IEEE CRC32 (one uint32, device memory)
 *   - calls cudaMalloc / cudaFree / no allocating routine; everything is queued on `stream` (a cudaStream_t, NULL = the
 *     default stream) and nothing is synchronised - EXCEPT the first call on a device, which uploads the CRC tables
 *     and synchronises that stream once (make one warm-up call outside any graph capture)
 *   - every GPU that loads the library holds the array (~270 KiB) until exit; all calls on a device share it, and it
 *     keeps the last call's contents (a call overwrites every entry it reads: nothing has to be cleared).  Calls that
 *     overlap on the GPU - different streams without an ordering between them - RACE on it. */
#define MRZ_NOALLOC_MAX_BYTES (1LL << 30)
int mrz_crc32_noalloc(int device, const uint8_t *d_data, int64_t n, uint32_t *d_crc_out, void *stream);
const char *mrz_noalloc_last_error(void);

/* Building blocks exposed for parity tests (each is also used inside mrz_run_resident). */
/* single_full_tag for every position (rzip.c:348-358): tags[p] for p in [0, count), count <= n-30. */
int mrz_debug_tags(mrz_ctx *ctx, int64_t n, const int64_t hash_index[256], int64_t *tags_out, int64_t count);
/* the trailing-ones map of the resident chunk (tagmap.cu): map[p] for p in [0, n), summary[b] for b in [0, (n+63)/64) */
int mrz_debug_tagmap(mrz_ctx *ctx, int64_t n, const int64_t hash_index[256], uint8_t *map_out, uint8_t *summary_out);
/* per-chunk CRC32 as gcry_md_write(crc_handle) accumulates it (rzip.c:494, :648-662) */
int mrz_debug_crc32(mrz_ctx *ctx, int64_t n, uint32_t *crc_out);
/* un-split match tokens of the last run: triples (pos, len, ofs) */
int mrz_debug_tokens(mrz_ctx *ctx, int64_t *triples, int64_t cap, int64_t *count);

#ifdef __cplusplus
}
#endif
#endif
