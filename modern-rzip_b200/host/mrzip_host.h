/* mrzip_host.h — host-side (C) mirror of the reference's interfaces around the rzip hot path.
 *
 * Same names and argument meaning as the reference so that a maintainer can diff call sites:
 *   rzip_fd            include/rzip.h:25, src/rzip.c:807
 *   open_stream_out / write_stream / flush_buffer / close_stream_out
 *                      include/stream.h:39-44, src/stream.c:771,1574,1347,1623
 *   write_magic / compress_file   src/mrzip.c:127,1053
 *   setup_ram / round_to_page     src/util.c:156-169
 * Only the fields of rzip_control (include/mrzip_private.h:419-524) that the `-n` path reads exist.
 * The hash search itself is NOT here: rzip_chunk() calls mrz_rzip_chunk() from libmrz_gpu.so.
 */
#ifndef MRZIP_HOST_H
#define MRZIP_HOST_H
#include <stdbool.h>
#include <stdint.h>
#include <stdio.h>

typedef int64_t i64;
typedef unsigned char uchar;

#define ONE_MB 1048576
#define STREAM_BUFSIZE (ONE_MB * 10)      /* mrzip_private.h:27 */
#ifndef CHUNK_MULTIPLE                  /* (tests/cusim builds a binary with tiny chunks to exercise the chunk scheduler on the CPU) */
#define CHUNK_MULTIPLE (100 * (i64)ONE_MB) /* rzip.c:46 */
#endif
#define NUM_STREAMS 2
#define MAGIC_LEN 20
#define CTYPE_NONE 3
#define MRZIP_MAJOR 0
#define MRZIP_MINOR 9

#define FLAG_SHOW_PROGRESS (1 << 0)
#define FLAG_KEEP_FILES (1 << 1)
#define FLAG_FORCE_REPLACE (1 << 2)
#define FLAG_NO_COMPRESS (1 << 3)
#define FLAG_UNLIMITED (1 << 4)
#define FLAG_VERBOSE (1 << 5)
#define FLAG_MAX_VERBOSE (1 << 6)
#define FLAG_STDIN (1 << 7)
#define FLAG_STDOUT (1 << 8)
#define FLAG_KEEP_BROKEN (1 << 9)

struct mrz_ctx;

typedef struct rzip_control {
    char *infile, *outname, *outfile, *outdir, *suffix;
    char *comment;
    int comment_length;
    int flags;
    int compression_level;       /* -L, default 7 (mrzip.c:1374) */
    int rzip_compression_level;  /* -R, defaults to -L (main.c:575) */
    i64 ramsize, maxram, usable_ram, max_chunk, max_mmap;
    i64 window;                  /* -w, units of 100 MiB */
    i64 st_size;
    i64 page_size;
    int threads;                 /* -p; forced to 1 for -n (stream.c:737) */
    int eof;
    int hash_code;               /* 1 = MD5 (main.c:583) */
    uchar hash_resblock[16];
    int fd_out;
    int devices[16];             /* -g: CUDA devices the chunks of a file are spread over (chunk k -> slot k mod slots) */
    int n_devices;
    int slots_per_device;        /* -G: chunks kept in flight per device (0 = chosen from the chunk count) */
    void *engine;                /* the GPU engine of rzip_fd.c (slots, contexts, staging), kept across files */
    /* statistics as printed by rzip.c:1108-1115 */
    i64 matches, match_bytes, literals, literal_bytes, tag_hits, tag_misses, inserts;
    double gpu_ms, gpu_scan_ms, gpu_auto_ms, gpu_emit_ms;
    i64 gpu_launches, variant_launches;
} rzip_control;

#define NO_COMPRESS (control->flags & FLAG_NO_COMPRESS)
#define UNLIMITED (control->flags & FLAG_UNLIMITED)
#define VERBOSE (control->flags & FLAG_VERBOSE)
#define MAX_VERBOSE (control->flags & FLAG_MAX_VERBOSE)
#define SHOW_PROGRESS (control->flags & FLAG_SHOW_PROGRESS)
#define FORCE_REPLACE (control->flags & FLAG_FORCE_REPLACE)
#define KEEP_FILES (control->flags & FLAG_KEEP_FILES)
#define STDIN (control->flags & FLAG_STDIN)
#define STDOUT (control->flags & FLAG_STDOUT)

/* util */
void fatal(rzip_control *control, const char *fmt, ...);
void setup_ram(rzip_control *control);
void round_to_page(rzip_control *control, i64 *size);
i64 round_up_page(rzip_control *control, i64 len);
i64 get_ram(void);

/* stream sink (write side, CTYPE_NONE) */
struct stream {
    uchar *buf;
    i64 buflen;
    i64 last_head;
};
struct stream_info {
    struct stream s[NUM_STREAMS];
    int num_streams;
    int fd;
    i64 bufsize, size, cur_pos, initial_pos;
    int chunks;
    char chunk_bytes;
    /* positional mode (open_stream_out_at): every write is a pwrite at fpos, the shared file offset is never used, so
     * several chunks of one archive can be written at the same time */
    int positional, eof;
    i64 fpos, expect_end;   /* expect_end != 0: close_stream_out checks the chunk ended there */
};
void *open_stream_out(rzip_control *control, int f, unsigned int n, i64 chunk_limit, char cbytes);
/* the same sink for a chunk whose first byte goes to file offset `start` (eof: this is the archive's last chunk) */
void *open_stream_out_at(rzip_control *control, int f, unsigned int n, i64 chunk_limit, char cbytes, i64 start, int eof);
/* bytes the chunk will occupy on disk (headers and blocks), known from the stream lengths alone */
i64 stream_out_chunk_size(const struct stream_info *sinfo, i64 s0_len, i64 s1_len);
void write_stream(rzip_control *control, void *ss, int streamno, const uchar *p, i64 len);
void flush_buffer(rzip_control *control, struct stream_info *sinfo, int streamno);
int close_stream_out(rzip_control *control, void *ss);
/* replay device-produced streams into the sink in the reference's write order */
void replay_streams(rzip_control *control, void *ss, const uchar *s0, i64 s0_len, const uchar *s1, i64 s1_len,
                    int chunk_bytes);

/* container */
bool write_magic(rzip_control *control);
bool compress_file(rzip_control *control);

/* the hot path's outer entry (include/rzip.h:25) */
void rzip_fd(rzip_control *control, int fd_in, int fd_out);
/* the way back for rzip-only archives (runzip_fd.c): include/runzip.h runzip_fd, mrzip.c decompress_file */
i64 runzip_fd(rzip_control *control, int fd_in, int fd_out, i64 first_chunk, i64 expected_size, int has_md5);
bool decompress_file(rzip_control *control);
void rzip_release_gpu(rzip_control *control);
i64 rzip_victim_round(void);     /* the process-lifetime static of insert_hash (rzip.c:259), for tests */
bool initialise_control(rzip_control *control); /* mrzip.c:1362 */

/* MD5 (RFC 1321) — the reference gets it from libgcrypt (rzip.c:495, :829) */
typedef struct {
    uint32_t h[4];
    uint64_t len;
    uchar blk[64];
    int fill;
} md5_ctx;
void md5_init(md5_ctx *c);
void md5_update(md5_ctx *c, const void *data, size_t len);
void md5_final(md5_ctx *c, uchar out[16]);

#endif
