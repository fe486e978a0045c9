/* runzip_fd.c — the way back for rzip-only (`-n`) archives: read side of the stream multiplexer + token replay.
 *
 * Mirrors, for CTYPE_NONE blocks, the reference's runzip_fd / runzip_chunk (src/runzip.c:226-327, :332-460), the read
 * side of the block multiplexer (open_stream_in src/stream.c:941-1080, fill_buffer :1412-1560, read_stream :1593-1620,
 * close_stream_in :1652-1680) and read_magic (src/mrzip.c:322-360).  New design: a chunk's two streams are gathered
 * whole (following each stream's chain of block headers), handed to ONE device call - mrz_runzip_chunk: header walk on
 * the host, literal scatter and match copies on the GPU, CRC32 of the result checked against the stored value - and
 * the chunk is written out; the reference replays token by token through a history file descriptor.  The stored MD5 is
 * checked over the output (runzip.c:393-424), messages as the reference prints them.
 */
#include <errno.h>
#include <fcntl.h>
#include <pthread.h>
#include <stdlib.h>
#include <string.h>
#include <sys/stat.h>
#include <sys/time.h>
#include <unistd.h>

#include "../../include/mrz_gpu.h"
#include "mrzip_host.h"

static void pread_all(rzip_control *control, int fd, void *buf, i64 len, i64 pos, const char *what) {
    uchar *b = buf;
    while (len > 0) {
        i64 n = len > (1 << 30) ? (1 << 30) : len;
        ssize_t r = pread(fd, b, (size_t)n, (off_t)pos);
        if (r < 0 && errno == EINTR) continue;
        if (r <= 0) fatal(control, "Failed to read %s: unexpected end of the archive\n", what);
        b += r;
        pos += r;
        len -= r;
    }
}

static i64 get_val(const uchar *p, int len) { /* read_val, stream.c:556-564: little-endian, `len` bytes */
    i64 v = 0;
    for (int i = 0; i < len; i++) v |= (i64)p[i] << (8 * i);
    return v;
}

struct md5_span {
    md5_ctx *ctx;
    const uchar *p;
    i64 len;
};
static void *md5_span_thread(void *arg) {
    struct md5_span *s = arg;
    md5_update(s->ctx, s->p, (size_t)s->len);
    return NULL;
}

struct strm {
    uchar *buf;
    i64 len, cap;
};

static void strm_reserve(rzip_control *control, struct strm *s, i64 need) {
    if (need <= s->cap) return;
    i64 cap = s->cap ? s->cap * 2 : (1 << 20);
    if (cap < need) cap = need;
    uchar *p = realloc(s->buf, (size_t)cap + 64);
    if (!p) fatal(control, "Unable to allocate a stream buffer of %lld bytes\n", (long long)cap);
    s->buf = p;
    s->cap = cap;
}

/* One chunk starting at file offset `pos`: gathers its streams; returns the offset behind the chunk. */
static i64 read_chunk_streams(rzip_control *control, int fd, i64 pos, i64 file_end, struct strm st[NUM_STREAMS], int *chunk_bytes, int *eof,
                              i64 *chunk_size) {
    uchar b2[2];
    pread_all(control, fd, b2, 2, pos, "the chunk header");
    const int wl = b2[0];
    if (wl < 1 || wl > 8) fatal(control, "chunk_bytes %d is invalid in runzip_chunk\n", wl); /* runzip.c:264 */
    *chunk_bytes = wl;
    *eof = b2[1];
    uchar v[8];
    pread_all(control, fd, v, wl, pos + 2, "the chunk size");
    *chunk_size = get_val(v, wl);
    const i64 base = pos + 2 + wl; /* initial_pos: block positions are relative to it (stream.c:1010-1030) */
    const i64 hdr = 1 + 3 * (i64)wl;
    i64 end = base + NUM_STREAMS * hdr;
    for (int s = 0; s < NUM_STREAMS; s++) {
        st[s].len = 0;
        uchar h[1 + 3 * 8];
        pread_all(control, fd, h, hdr, base + s * hdr, "a stream header");
        if (h[0] != CTYPE_NONE || get_val(h + 1, wl) || get_val(h + 1 + wl, wl))
            fatal(control, "Unexpected initial tag %d in streams: this build reads rzip-only (-n) archives\n", h[0]); /* stream.c:1046 */
        i64 next = get_val(h + 1 + 2 * wl, wl), last_at = base;
        while (next) {
            const i64 at = base + next;
            if (at + hdr > file_end) fatal(control, "Invalid data: a block header lies beyond the end of the archive\n");
            if (at <= last_at) fatal(control, "Invalid data: the block chain of a stream goes backwards\n"); /* blocks are appended: no cycles */
            last_at = at;
            pread_all(control, fd, h, hdr, at, "a block header");
            const i64 c_len = get_val(h + 1, wl), u_len = get_val(h + 1 + wl, wl);
            if (h[0] != CTYPE_NONE) fatal(control, "Block with compression type %d: this build reads rzip-only (-n) archives\n", h[0]);
            if (c_len != u_len || c_len < 0 || at + hdr + c_len > file_end)
                fatal(control, "Invalid data compressed len %lld uncompressed %lld\n", (long long)c_len, (long long)u_len); /* stream.c:1465 */
            strm_reserve(control, &st[s], st[s].len + c_len);
            pread_all(control, fd, st[s].buf + st[s].len, c_len, at + hdr, "a block");
            st[s].len += c_len;
            if (at + hdr + c_len > end) end = at + hdr + c_len;
            next = get_val(h + 1 + 2 * wl, wl);
        }
    }
    return end;
}

/* Decode the archive on fd_in (positioned anywhere) into fd_out; returns the bytes written.  expected_size and the
 * hash flag come from read_magic. */
i64 runzip_fd(rzip_control *control, int fd_in, int fd_out, i64 first_chunk, i64 expected_size, int has_md5) {
    struct stat sb;
    if (fstat(fd_in, &sb)) fatal(control, "Failed to stat the archive\n");
    const i64 data_end = sb.st_size - (has_md5 ? 16 : 0);
    struct timeval t0, t1;
    gettimeofday(&t0, NULL);
    md5_ctx md5;
    md5_init(&md5);
    struct strm st[NUM_STREAMS];
    memset(st, 0, sizeof st);
    mrz_ctx *ctx = NULL;
    i64 ctx_cap = 0, out_cap = 0, total = 0, pos = first_chunk;
    uchar *out = NULL;
    double gpu_ms = 0;
    i64 launches = 0, waves = 0;
    int eof = 0, chunks = 0;
    const int device = control->n_devices ? control->devices[0] : 0;
    while (!eof && pos < data_end) {
        int wl;
        i64 csize;
        pos = read_chunk_streams(control, fd_in, pos, data_end, st, &wl, &eof, &csize);
        if (MAX_VERBOSE) fprintf(stderr, "Chunk byte width: %d\n", wl); /* runzip.c:268 */
        /* the chunk is at most what its header says (the encoder's chunk size); literals alone bound it from below */
        i64 need = csize > st[1].len ? csize : st[1].len;
        if (need < 1) need = 1;
        if (!ctx || need > ctx_cap) {
            if (ctx) mrz_ctx_destroy(ctx);
            int rc = mrz_ctx_create(device, need, &ctx);
            if (rc) fatal(control, "Cannot set up the GPU rzip engine on device %d (%d): %s\n", device, rc, mrz_last_error(NULL));
            ctx_cap = need;
        }
        if (need > out_cap) {
            if (out) mrz_host_free(out);
            out = mrz_host_alloc(need + 64);
            if (!out) fatal(control, "Failed to allocate the chunk buffer of %lld bytes\n", (long long)need);
            out_cap = need;
        }
        mrz_unzip_result r;
        int rc = mrz_runzip_chunk(ctx, st[0].buf, st[0].len, st[1].buf, st[1].len, wl, out, out_cap, &r);
        if (rc) fatal(control, "%s\n", mrz_last_error(ctx)); /* "Bad checksum: 0x.. - expected: 0x.." (runzip.c:319) / corrupt archive */
        if (MAX_VERBOSE) fprintf(stderr, "Checksum for block: 0x%08x\n", r.crc32); /* runzip.c:321 */
        /* the MD5 chain of this chunk runs beside its write(2) (both only read `out`) */
        struct md5_span span = {&md5, out, r.out_len};
        pthread_t th;
        const int hashing = has_md5 && r.out_len > (1 << 20) && !pthread_create(&th, NULL, md5_span_thread, &span);
        for (i64 o = 0; o < r.out_len;) { /* write_1g */
            i64 n = r.out_len - o > (1 << 30) ? (1 << 30) : r.out_len - o;
            ssize_t w = write(fd_out, out + o, (size_t)n);
            if (w < 0 && errno == EINTR) continue;
            if (w <= 0) fatal(control, "Failed to write %lld bytes of the chunk\n", (long long)n);
            o += w;
        }
        if (hashing)
            pthread_join(th, NULL);
        else if (has_md5)
            md5_update(&md5, out, (size_t)r.out_len);
        total += r.out_len;
        gpu_ms += r.ms_kernels;
        launches += r.gpu_launches;
        waves += r.waves;
        chunks++;
    }
    if (expected_size && total != expected_size)
        fatal(control, "Failed to runzip_chunk in runzip_fd: %lld bytes, the archive says %lld\n", (long long)total, (long long)expected_size);
    gettimeofday(&t1, NULL);
    double tdiff = (t1.tv_sec - t0.tv_sec) + 1e-6 * (t1.tv_usec - t0.tv_usec);
    if (tdiff <= 0) tdiff = 1e-6;
    if (SHOW_PROGRESS) fprintf(stderr, "\nAverage DeCompression Speed: %6.3fMB/s\n", (total / (double)ONE_MB) / tdiff);
    if (has_md5) { /* runzip.c:393-424 */
        uchar stored[16], got[16];
        pread_all(control, fd_in, stored, 16, sb.st_size - 16, "MD5 data");
        md5_final(&md5, got);
        if (memcmp(stored, got, 16)) {
            fprintf(stderr, "MD5 CHECK FAILED.\nStored:");
            for (int j = 0; j < 16; j++) fprintf(stderr, "%02x", stored[j]);
            fprintf(stderr, "\nOutput file:");
            for (int j = 0; j < 16; j++) fprintf(stderr, "%02x", got[j]);
            fatal(control, "\n");
        }
        if (SHOW_PROGRESS) {
            fprintf(stderr, "MD5:");
            for (int j = 0; j < 16; j++) fprintf(stderr, "%02x", got[j]);
            fprintf(stderr, "\n");
        }
    } else if (SHOW_PROGRESS)
        fprintf(stderr, "Note this mrzip archive did not have a stored hash value.\n"
                        "The archive decompression was validated with crc32 and the was calculated on decompression\n");
    if (MAX_VERBOSE)
        fprintf(stderr, "gpu: %d chunk(s), %.1f ms in kernels, %lld launches (%lld waves of independent matches)\n", chunks, gpu_ms,
                (long long)launches, (long long)waves);
    if (out) mrz_host_free(out);
    if (ctx) mrz_ctx_destroy(ctx);
    for (int s = 0; s < NUM_STREAMS; s++) free(st[s].buf);
    return total;
}

/* decompress_file (mrzip.c:700-1050) for a regular file holding an unencrypted rzip-only archive. */
bool decompress_file(rzip_control *control) {
    const char *suffix = control->suffix ? control->suffix : ".lrz";
    int fd_in = open(control->infile, O_RDONLY);
    if (fd_in == -1) fatal(control, "Failed to open %s\n", control->infile);
    uchar magic[MAGIC_LEN];
    if (read(fd_in, magic, MAGIC_LEN) != MAGIC_LEN) fatal(control, "Failed to read magic header\n");
    if (memcmp(magic, "MRZI", 4)) fatal(control, "Not an mrzip file\n"); /* mrzip.c:338 */
    if (magic[4] != MRZIP_MAJOR || magic[5] != MRZIP_MINOR)
        fatal(control, "mrzip version %d.%d archive is not supported. Aborting\n", magic[4], magic[5]); /* mrzip.c:313 */
    if (magic[15]) fatal(control, "Encrypted archives are read by the reference build\n");
    if (magic[16]) fatal(control, "This build reads rzip-only (-n) archives; the back-end filters stay with the reference build\n");
    const i64 expected_size = get_val(magic + 6, 8);
    const int has_md5 = magic[14] != 0; /* hash_code: 1 = MD5 (mrzip.c get_hash_from_magic); others stay with the reference */
    if (magic[14] > 1) fatal(control, "Hash code %d: this build checks MD5 only\n", magic[14]);
    const i64 first_chunk = MAGIC_LEN + magic[19]; /* a comment follows the magic (mrzip.c:286-296) */

    char *outfile;
    if (control->outname)
        outfile = strdup(control->outname);
    else { /* mrzip.c:726-760: strip the suffix, or append ".out" */
        const char *tmp, *in = control->infile;
        if (control->outdir && (tmp = strrchr(in, '/'))) in = tmp + 1;
        size_t l = strlen(in), sl = strlen(suffix);
        outfile = malloc((control->outdir ? strlen(control->outdir) : 0) + l + 8);
        if (!outfile) fatal(control, "Failed to allocate outfile name\n");
        outfile[0] = 0;
        if (control->outdir) strcpy(outfile, control->outdir);
        if (l > sl && !strcmp(in + l - sl, suffix))
            strncat(outfile, in, l - sl);
        else {
            strcat(outfile, in);
            strcat(outfile, ".out");
        }
    }
    if (!strcmp(control->infile, outfile)) fatal(control, "Input and Output files are the same. %s. Exiting\n", control->infile);
    int fd_out = open(outfile, O_WRONLY | O_CREAT | O_EXCL, 0666);
    if (FORCE_REPLACE && fd_out == -1 && errno == EEXIST) {
        if (unlink(outfile)) fatal(control, "Failed to unlink an existing file: %s\n", outfile);
        fd_out = open(outfile, O_WRONLY | O_CREAT | O_EXCL, 0666);
    }
    if (fd_out == -1) fatal(control, "Failed to create %s\n", outfile);
    struct stat st;
    if (!fstat(fd_in, &st)) (void)!fchmod(fd_out, st.st_mode & 0777);
    if (SHOW_PROGRESS) fprintf(stderr, "Decompressing...\n");
    runzip_fd(control, fd_in, fd_out, first_chunk, expected_size, has_md5);
    if (close(fd_out)) fatal(control, "Failed to close fd_out\n");
    close(fd_in);
    if (SHOW_PROGRESS) fprintf(stderr, "Output filename is: %s: [OK] - %lld bytes\n", outfile, (long long)expected_size);
    if (!KEEP_FILES && unlink(control->infile)) fatal(control, "Failed to unlink %s\n", control->infile);
    free(outfile);
    return true;
}
