/* stream_writer_test.c - TEST INFRASTRUCTURE.  The positional stream writer of modern-rzip_b200/host/stream_out.c
 * (open_stream_out_at: pwrite at tracked offsets, chunk size known from the stream lengths) against the sequential one
 * (open_stream_out: the shared file offset, as the reference's stream.c uses it): same bytes, and the size predicted by
 * stream_out_chunk_size is the size written - with streams several blocks long (bufsize is 10 MiB here).
 * usage: stream_writer_test <dir> <literal runs per chunk> [<literal runs> ...]   prints "ok <bytes>" per case */
#include <fcntl.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <unistd.h>

#include "../../modern-rzip_b200/host/mrzip_host.h"

/* util.c (fatal, compress_file) refers to the hot path's entry points; never called here */
void rzip_fd(rzip_control *control, int fd_in, int fd_out) { (void)control; (void)fd_in; (void)fd_out; abort(); }
void rzip_release_gpu(rzip_control *control) { (void)control; }

static void make_streams(int runs, uchar **s0, i64 *s0_len, uchar **s1, i64 *s1_len, unsigned seed) {
    /* `runs` literal runs of 65535 bytes with a match token (01 len16 ofs[4]) between them, terminator, CRC */
    const int cb = 4;
    *s0 = malloc((size_t)runs * (3 + 3 + cb) + 16);
    *s1 = malloc((size_t)runs * 65535 + 16);
    i64 a = 0, b = 0;
    for (int r = 0; r < runs; r++) {
        (*s0)[a++] = 0; (*s0)[a++] = 0xff; (*s0)[a++] = 0xff;
        for (int i = 0; i < 65535; i++) (*s1)[b++] = (uchar)(seed = seed * 1103515245u + 12345u) >> 3;
        (*s0)[a++] = 1; (*s0)[a++] = 0x40; (*s0)[a++] = 0;
        for (int i = 0; i < cb; i++) (*s0)[a++] = (uchar)(r + i);
    }
    (*s0)[a++] = 0; (*s0)[a++] = 0; (*s0)[a++] = 0;
    for (int i = 0; i < 4; i++) (*s0)[a++] = (uchar)(0xc0 + i);
    *s0_len = a; *s1_len = b;
}

int main(int argc, char **argv) {
    if (argc < 3) return 2;
    rzip_control control;
    memset(&control, 0, sizeof control);
    control.page_size = 4096;
    control.usable_ram = 1LL << 33;
    control.st_size = 1 << 20;       /* bufsize = max(st_size, STREAM_BUFSIZE) = 10 MiB */
    char pa[512], pb[512];
    snprintf(pa, sizeof pa, "%s/seq.bin", argv[1]);
    snprintf(pb, sizeof pb, "%s/pos.bin", argv[1]);
    int fa = open(pa, O_RDWR | O_CREAT | O_TRUNC, 0600), fb = open(pb, O_RDWR | O_CREAT | O_TRUNC, 0600);
    if (fa < 0 || fb < 0) return 3;
    const char magic[20] = "container header...";
    if (write(fa, magic, 20) != 20 || write(fb, magic, 20) != 20) return 3;
    const int ncase = argc - 2;
    /* sequential: chunk after chunk through the shared file offset */
    i64 place = 20;
    for (int c = 0; c < ncase; c++) {
        uchar *s0, *s1; i64 l0, l1;
        make_streams(atoi(argv[2 + c]), &s0, &l0, &s1, &l1, 77u + (unsigned)c);
        control.eof = c == ncase - 1;
        void *ss = open_stream_out(&control, fa, NUM_STREAMS, l1 + 100, 4);
        replay_streams(&control, ss, s0, l0, s1, l1, 4);
        close_stream_out(&control, ss);
        free(s0); free(s1);
    }
    /* positional: places from the lengths first, then the chunks written in REVERSE order */
    i64 start[64], endp[64];
    for (int c = 0; c < ncase; c++) {
        uchar *s0, *s1; i64 l0, l1;
        make_streams(atoi(argv[2 + c]), &s0, &l0, &s1, &l1, 77u + (unsigned)c);
        struct stream_info *ss = open_stream_out_at(&control, fb, NUM_STREAMS, l1 + 100, 4, place, 0);
        start[c] = place;
        place += stream_out_chunk_size(ss, l0, l1);
        endp[c] = place;
        free(ss);
        free(s0); free(s1);
    }
    control.eof = 0;
    for (int c = ncase - 1; c >= 0; c--) {
        uchar *s0, *s1; i64 l0, l1;
        make_streams(atoi(argv[2 + c]), &s0, &l0, &s1, &l1, 77u + (unsigned)c);
        struct stream_info *ss = open_stream_out_at(&control, fb, NUM_STREAMS, l1 + 100, 4, start[c], c == ncase - 1);
        replay_streams(&control, ss, s0, l0, s1, l1, 4);
        ss->expect_end = endp[c];
        close_stream_out(&control, ss);   /* fatal on a size mismatch */
        free(s0); free(s1);
    }
    const i64 ea = lseek(fa, 0, SEEK_CUR), eb = lseek(fb, 0, SEEK_END);
    if (ea != place || eb != place) { fprintf(stderr, "sizes: sequential %lld positional %lld predicted %lld\n", (long long)ea, (long long)eb, (long long)place); return 1; }
    printf("ok %lld\n", (long long)place);
    return 0;
}
