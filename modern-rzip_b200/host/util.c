/* util.c — the few helpers of src/util.c and src/mrzip.c the `-n` path needs:
 * fatal (util.c:64-83, util.h:50-62), setup_ram / round_to_page / round_up_page (util.c:156-176),
 * get_ram (mrzip.c:92-112), initialise_control (mrzip.c:1362), write_magic (mrzip.c:127-188),
 * compress_file (mrzip.c:1053-1165). */
#define _GNU_SOURCE
#include <errno.h>
#include <fcntl.h>
#include <stdarg.h>
#include <stdlib.h>
#include <string.h>
#include <sys/stat.h>
#include <unistd.h>

#include "mrzip_host.h"

void fatal(rzip_control *control, const char *fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vfprintf(stderr, fmt, ap);
    va_end(ap);
    /* util.c:64-83: remove the partial output unless it pre-existed or -K */
    if (control && control->outfile && control->fd_out > 0 && !(control->flags & FLAG_KEEP_BROKEN)) unlink(control->outfile);
    fprintf(stderr, "Fatal error - exiting\n");
    exit(1);
}

i64 get_ram(void) {
    i64 ramsize = (i64)sysconf(_SC_PHYS_PAGES) * (i64)sysconf(_SC_PAGE_SIZE);
    if (ramsize <= 0) {
        FILE *f = fopen("/proc/meminfo", "r");
        long long kb = 0;
        char line[256];
        while (f && fgets(line, sizeof line, f))
            if (sscanf(line, "MemTotal: %lld kB", &kb) == 1) break;
        if (f) fclose(f);
        ramsize = (i64)kb * 1024;
    }
    if (ramsize <= 0) fatal(NULL, "No memory or can't determine ram? Can't continue.\n");
    return ramsize;
}

void round_to_page(rzip_control *control, i64 *size) {
    *size -= *size % control->page_size;
    if (!*size) *size = control->page_size;
}

i64 round_up_page(rzip_control *control, i64 len) {
    i64 rem = len % control->page_size;
    if (rem) len += control->page_size - rem;
    return len;
}

void setup_ram(rzip_control *control) {
    /* util.c:156-164 (no stdout mode here: 1/3 of ram) */
    control->maxram = control->ramsize / 3;
    control->usable_ram = control->maxram;
    round_to_page(control, &control->maxram);
}

bool initialise_control(rzip_control *control) {
    memset(control, 0, sizeof *control);
    control->flags = FLAG_SHOW_PROGRESS | FLAG_KEEP_FILES;
    control->compression_level = 7;
    control->rzip_compression_level = 0;
    control->ramsize = get_ram();
    control->threads = (int)sysconf(_SC_NPROCESSORS_ONLN);
    control->page_size = sysconf(_SC_PAGE_SIZE);
    control->hash_code = 1; /* MD5, main.c:583 */
    return true;
}

bool write_magic(rzip_control *control) {
    uchar magic[MAGIC_LEN] = {'M', 'R', 'Z', 'I', MRZIP_MAJOR, MRZIP_MINOR};
    if (control->eof) {
        uint64_t sz = (uint64_t)control->st_size;
        for (int i = 0; i < 8; i++) magic[6 + i] = (uchar)(sz >> (8 * i));
    }
    magic[14] = (uchar)control->hash_code;
    magic[16] = 0;
    magic[18] = (uchar)((control->rzip_compression_level << 4) + control->compression_level);
    magic[19] = (uchar)control->comment_length;
    if (lseek(control->fd_out, 0, SEEK_SET) == -1) fatal(control, "Failed to seek to BOF to write Magic Header\n");
    if (write(control->fd_out, magic, MAGIC_LEN) != MAGIC_LEN) fatal(control, "Failed to write magic header\n");
    if (magic[19] && write(control->fd_out, control->comment, control->comment_length) != control->comment_length)
        fatal(control, "Failed to write comment after magic header\n");
    return true;
}

bool compress_file(rzip_control *control) {
    int fd_in, fd_out, len = MAGIC_LEN + control->comment_length;
    struct stat st;
    char *header = calloc((size_t)len, 1);

    fd_in = open(control->infile, O_RDONLY);
    if (fd_in == -1) fatal(control, "Failed to open %s\n", control->infile);

    if (control->outname)
        control->outfile = strdup(control->outname);
    else {
        const char *tmp, *tmpinfile = control->infile;
        if (control->outdir && (tmp = strrchr(control->infile, '/'))) tmpinfile = tmp + 1;
        control->outfile = malloc((control->outdir ? strlen(control->outdir) : 0) + strlen(tmpinfile) + strlen(control->suffix) + 1);
        if (!control->outfile) fatal(control, "Failed to allocate outfile name\n");
        control->outfile[0] = 0;
        if (control->outdir) strcpy(control->outfile, control->outdir);
        strcat(control->outfile, tmpinfile);
        strcat(control->outfile, control->suffix);
    }
    if (!strcmp(control->infile, control->outfile)) fatal(control, "Input and Output files are the same. %s. Exiting\n", control->infile);

    fd_out = open(control->outfile, O_RDWR | O_CREAT | O_EXCL, 0666);
    if (FORCE_REPLACE && fd_out == -1 && errno == EEXIST) {
        if (unlink(control->outfile)) fatal(control, "Failed to unlink an existing file: %s\n", control->outfile);
        fd_out = open(control->outfile, O_RDWR | O_CREAT | O_EXCL, 0666);
    }
    if (fd_out == -1) {
        control->flags |= FLAG_KEEP_BROKEN;
        fatal(control, "Failed to create %s\n", control->outfile);
    }
    control->fd_out = fd_out;
    if (!fstat(fd_in, &st)) (void)!fchmod(fd_out, st.st_mode & 0777); /* preserve_perms */

    if (write(fd_out, header, (size_t)len) != len) fatal(control, "Cannot write file header\n");
    control->eof = 0;
    rzip_fd(control, fd_in, fd_out);
    if (!write_magic(control)) return false;
    close(fd_in);
    if (close(fd_out)) fatal(control, "Failed to close fd_out\n");
    if (!KEEP_FILES && unlink(control->infile)) fatal(control, "Failed to unlink %s\n", control->infile);
    free(control->outfile);
    control->outfile = NULL;
    free(header);
    return true;
}
