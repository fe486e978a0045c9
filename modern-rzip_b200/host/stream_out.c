/* stream_out.c — write side of the two-stream block multiplexer for uncompressed (CTYPE_NONE)
 * blocks, byte-compatible with the reference's archive layout.
 *
 * Interface and on-disk format follow src/stream.c: open_stream_out (:771-938, buffer sizing
 * :803-921), write_stream (:1574-1590), flush_buffer/clear_buffer (:1307-1349), the writer half of
 * compthread (:1199-1299) and close_stream_out (:1623-1648).  New design: with no back-end there is
 * nothing to run in parallel, so blocks are written synchronously in flush order instead of through
 * detached threads serialised by a condition variable — the resulting bytes are the same.
 */
#include <errno.h>
#include <stdlib.h>
#include <string.h>
#include <unistd.h>

#include "mrzip_host.h"

/* stream.c:68-70: decided once per process */
static i64 stream_bufsize = 0;

static void write_all(rzip_control *control, struct stream_info *sinfo, const void *p, i64 len) {
    const uchar *b = p;
    while (len > 0) {
        i64 n = len > (1 << 30) ? (1 << 30) : len; /* write_1g, stream.c:519 */
        ssize_t r = sinfo->positional ? pwrite(sinfo->fd, b, (size_t)n, (off_t)sinfo->fpos) : write(sinfo->fd, b, (size_t)n);
        if (r < 0) {
            if (errno == EINTR) continue;
            fatal(control, "Failed to write %lld bytes: %s\n", (long long)n, strerror(errno));
        }
        b += r;
        len -= r;
        sinfo->fpos += r;
    }
}

static void write_val(rzip_control *control, struct stream_info *sinfo, i64 v, int len) {
    uchar b[8];
    for (int i = 0; i < 8; i++) b[i] = (uchar)((uint64_t)v >> (8 * i)); /* htole64, low `len` bytes */
    write_all(control, sinfo, b, len);
}

static void seekto(rzip_control *control, struct stream_info *sinfo, i64 pos) {
    if (sinfo->positional) {
        sinfo->fpos = sinfo->initial_pos + pos;
        return;
    }
    if (lseek(sinfo->fd, sinfo->initial_pos + pos, SEEK_SET) == -1) fatal(control, "Failed to seek in stream writer\n");
}

void *open_stream_out(rzip_control *control, int f, unsigned int n, i64 chunk_limit, char cbytes) {
    struct stream_info *sinfo = calloc(1, sizeof *sinfo);
    if (!sinfo) return NULL;
    if (chunk_limit < control->page_size) chunk_limit = control->page_size; /* stream.c:779 */
    sinfo->bufsize = sinfo->size = chunk_limit;
    sinfo->chunk_bytes = cbytes;
    sinfo->num_streams = (int)n;
    sinfo->fd = f;

    if (!stream_bufsize) {
        /* stream.c:803-914 with NO_COMPRESS: one test buffer, one thread, no back-end overhead */
        i64 limit = control->usable_ram;
        if (control->st_size > 0 && control->st_size < limit)
            limit = control->st_size > STREAM_BUFSIZE ? control->st_size : STREAM_BUFSIZE;
        else if (limit > chunk_limit)
            limit = chunk_limit;
        for (;;) { /* retest_malloc, stream.c:882-896 */
            void *t = malloc((size_t)limit);
            if (t) {
                free(t);
                break;
            }
            limit = limit / 10 * 9;
            if (limit < 100000000) {
                free(sinfo);
                return NULL;
            }
        }
        i64 per = limit / 1 > STREAM_BUFSIZE ? limit : STREAM_BUFSIZE; /* MAX(limit/threads, STREAM_BUFSIZE) */
        stream_bufsize = round_up_page(control, limit < per ? limit : per);
    }
    sinfo->bufsize = stream_bufsize;
    for (unsigned i = 0; i < n; i++) {
        /* the reference callocs bufsize up front; we grow on demand (same bytes on disk) */
        sinfo->s[i].buf = NULL;
        sinfo->s[i].buflen = 0;
    }
    return sinfo;
}

/* Writer half of compthread (stream.c:1199-1299) for one finished block. */
static void write_block(rzip_control *control, struct stream_info *sinfo, int streamno, const uchar *data, i64 len) {
    int wl = sinfo->chunk_bytes;
    if (!sinfo->chunks++) {
        uchar b2[2] = {(uchar)sinfo->chunk_bytes, (uchar)(sinfo->positional ? sinfo->eof : control->eof)};
        write_all(control, sinfo, b2, 2);
        write_val(control, sinfo, sinfo->size, wl);
        if (sinfo->positional)
            sinfo->initial_pos = sinfo->fpos;
        else {
            sinfo->initial_pos = lseek(sinfo->fd, 0, SEEK_CUR);
            if (sinfo->initial_pos == -1) fatal(control, "Failed to get position of chunk header\n");
        }
        for (int j = 0; j < sinfo->num_streams; j++) {
            uchar t = CTYPE_NONE;
            sinfo->s[j].last_head = sinfo->cur_pos + 1 + (wl * 2);
            write_all(control, sinfo, &t, 1);
            write_val(control, sinfo, 0, wl);
            write_val(control, sinfo, 0, wl);
            write_val(control, sinfo, 0, wl);
            sinfo->cur_pos += 1 + (wl * 3);
        }
    }
    /* link the previous header of this stream to the new block */
    seekto(control, sinfo, sinfo->s[streamno].last_head);
    write_val(control, sinfo, sinfo->cur_pos, wl);
    sinfo->s[streamno].last_head = sinfo->cur_pos + 1 + (wl * 2);
    seekto(control, sinfo, sinfo->cur_pos);
    uchar t = CTYPE_NONE;
    write_all(control, sinfo, &t, 1);
    write_val(control, sinfo, len, wl); /* c_len */
    write_val(control, sinfo, len, wl); /* s_len */
    write_val(control, sinfo, 0, wl);   /* next_head, patched later */
    sinfo->cur_pos += 1 + (wl * 3);
    write_all(control, sinfo, data, len);
    sinfo->cur_pos += len;
}

/* What write_block will have written for a chunk, from the stream lengths alone: the chunk header (2 + wl bytes), one
 * initial header per stream, then per stream floor(len / bufsize) full blocks and the final flush (close_stream_out writes
 * a block for every stream, even an empty one), each with a 1 + 3 wl header. */
i64 stream_out_chunk_size(const struct stream_info *sinfo, i64 s0_len, i64 s1_len) {
    const i64 wl = sinfo->chunk_bytes, hdr = 1 + 3 * wl;
    const i64 lens[2] = {s0_len, s1_len};
    i64 total = 2 + wl + (i64)sinfo->num_streams * hdr;
    for (int i = 0; i < sinfo->num_streams && i < 2; i++) total += (lens[i] / sinfo->bufsize + 1) * hdr + lens[i];
    return total;
}

void *open_stream_out_at(rzip_control *control, int f, unsigned int n, i64 chunk_limit, char cbytes, i64 start, int eof) {
    struct stream_info *sinfo = open_stream_out(control, f, n, chunk_limit, cbytes);
    if (!sinfo) return NULL;
    sinfo->positional = 1;
    sinfo->eof = eof;
    sinfo->fpos = start;
    return sinfo;
}

void flush_buffer(rzip_control *control, struct stream_info *sinfo, int streamno) {
    write_block(control, sinfo, streamno, sinfo->s[streamno].buf, sinfo->s[streamno].buflen);
    sinfo->s[streamno].buflen = 0;
}

static void reserve(rzip_control *control, struct stream *s, i64 need, i64 bufsize) {
    static const i64 step = 1 << 20;
    i64 have = s->buf ? *(i64 *)(s->buf - 16) : 0;
    if (need <= have) return;
    i64 cap = have ? have * 2 : step;
    if (cap < need) cap = need;
    if (cap > bufsize) cap = bufsize;
    uchar *raw = s->buf ? s->buf - 16 : NULL;
    raw = realloc(raw, (size_t)cap + 16);
    if (!raw) fatal(control, "Unable to allocate stream buffer of %lld bytes\n", (long long)cap);
    *(i64 *)raw = cap;
    s->buf = raw + 16;
}

void write_stream(rzip_control *control, void *ss, int streamno, const uchar *p, i64 len) {
    struct stream_info *sinfo = ss;
    struct stream *s = &sinfo->s[streamno];
    while (len) {
        i64 n = sinfo->bufsize - s->buflen;
        if (n > len) n = len;
        if (n == sinfo->bufsize && s->buflen == 0) {
            /* a whole block straight from the caller's memory: skip the staging copy */
            write_block(control, sinfo, streamno, p, n);
        } else {
            reserve(control, s, s->buflen + n, sinfo->bufsize);
            memcpy(s->buf + s->buflen, p, (size_t)n);
            s->buflen += n;
            if (s->buflen == sinfo->bufsize) flush_buffer(control, sinfo, streamno);
        }
        p += n;
        len -= n;
    }
}

int close_stream_out(rzip_control *control, void *ss) {
    struct stream_info *sinfo = ss;
    /* stream.c:1627: every stream is flushed, stream 0 first, even when empty */
    for (int i = 0; i < sinfo->num_streams; i++) {
        write_block(control, sinfo, i, sinfo->s[i].buf, sinfo->s[i].buflen);
        if (sinfo->s[i].buf) free(sinfo->s[i].buf - 16);
        sinfo->s[i].buf = NULL;
    }
    /* leave the fd at the end of the chunk for the next chunk / the hash trailer (positional mode: the caller seeks) */
    seekto(control, sinfo, sinfo->cur_pos);
    if (sinfo->positional && sinfo->expect_end && sinfo->fpos != sinfo->expect_end)
        fatal(control, "stream writer: the chunk ends at %lld, its place was computed as ending at %lld\n", (long long)sinfo->fpos,
              (long long)sinfo->expect_end);
    free(sinfo);
    return 0;
}

/* Feed device-produced streams to the sink in the order hash_search would have written them:
 * per token, header bytes to stream 0 then literal bytes to stream 1 (rzip.c:181-227).  When
 * neither stream can fill a buffer the interleaving is irrelevant and both go in one call each. */
void replay_streams(rzip_control *control, void *ss, const uchar *s0, i64 s0_len, const uchar *s1, i64 s1_len,
                    int chunk_bytes) {
    struct stream_info *sinfo = ss;
    if (s0_len < sinfo->bufsize && s1_len < sinfo->bufsize) {
        write_stream(control, ss, 0, s0, s0_len);
        write_stream(control, ss, 1, s1, s1_len);
        return;
    }
    i64 i = 0, j = 0;
    while (i < s0_len) {
        uchar head = s0[i];
        i64 len = s0[i + 1] | ((i64)s0[i + 2] << 8);
        if (head == 0) {
            write_stream(control, ss, 0, s0 + i, 3);
            i += 3;
            if (len == 0) { /* terminator: the CRC follows */
                write_stream(control, ss, 0, s0 + i, s0_len - i);
                i = s0_len;
                break;
            }
            write_stream(control, ss, 1, s1 + j, len);
            j += len;
        } else {
            write_stream(control, ss, 0, s0 + i, 3 + chunk_bytes);
            i += 3 + chunk_bytes;
        }
    }
    if (j != s1_len) fatal(control, "stream replay out of sync (%lld of %lld literal bytes)\n", (long long)j, (long long)s1_len);
}
