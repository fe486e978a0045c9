/* rzip_fd.c — host side of the rzip pre-processor: same entry point and chunking rules as the
 * reference's src/rzip.c, with the hot loop (hash_search, rzip.c:507-667) replaced by one call into
 * libmrz_gpu.so per chunk.
 *
 *   rzip_fd        rzip.c:807-1132   chunk sizing (:875-894), per-chunk loop (:915-1061), hash trailer
 *   rzip_chunk     rzip.c:763-792    open_stream_out -> [device] -> close_stream_out
 *   init_hash_indexes rzip.c:669-673 (the literal random() calls stay on the host: process-global state)
 *
 * What is different by design: the chunk is handed to the GPU whole (no sliding mmap, rzip.c:75-158:
 * a chunk fits HBM), the per-page checksum threads (rzip.c:488-505, :601-661) are gone — CRC32 comes
 * back from the device inside stream 0 and MD5 runs in ONE helper thread over the mapped chunk while
 * the GPU works — and the static victim_round of insert_hash (rzip.c:259) is carried explicitly from
 * chunk to chunk (and file to file) through mrz_params/mrz_result.
 * There is no CPU implementation of the search here: without a CUDA device rzip_fd() is fatal.
 */
#define _GNU_SOURCE
#include <errno.h>
#include <pthread.h>
#include <stdlib.h>
#include <string.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <sys/statvfs.h>
#include <sys/time.h>
#include <unistd.h>

#include "../../include/mrz_gpu.h"
#include "mrzip_host.h"

/* process-lifetime state, as in the reference: the static in insert_hash (rzip.c:259) */
static i64 victim_round = 0;

static void init_hash_indexes(i64 hash_index[256]) { /* rzip.c:669-673; gcc evaluates left to right */
    for (int i = 0; i < 256; i++) {
        i64 a = random();
        i64 b = random();
        hash_index[i] = (a << 16) ^ b;
    }
}

struct md5_job {
    md5_ctx *ctx;
    const uchar *p;
    i64 len;
};
static void *md5_thread(void *arg) {
    struct md5_job *j = arg;
    const i64 step = 64 << 20;
    for (i64 o = 0; o < j->len; o += step) md5_update(j->ctx, j->p + o, (size_t)(j->len - o < step ? j->len - o : step));
    return NULL;
}

/* device + pinned staging, grown to the largest chunk seen */
static void ensure_gpu(rzip_control *control, i64 chunk) {
    if (control->gpu && control->gpu_cap >= chunk) return;
    if (control->gpu) {
        mrz_ctx_destroy(control->gpu);
        mrz_host_free(control->pin_s0);
        mrz_host_free(control->pin_s1);
        control->gpu = NULL;
    }
    int rc = mrz_ctx_create(control->device, chunk, &control->gpu);
    if (rc) fatal(control, "Cannot set up the GPU rzip engine (%d): %s\n", rc, mrz_last_error(NULL));
    control->gpu_cap = chunk;
    control->s0_cap = chunk / 2 + 4096;
    control->s1_cap = chunk + 64;
    control->pin_s0 = mrz_host_alloc(control->s0_cap);
    control->pin_s1 = mrz_host_alloc(control->s1_cap);
    if (!control->pin_s0 || !control->pin_s1) fatal(control, "Cannot allocate pinned stream buffers\n");
}

/* rzip.c:763-792 with hash_search on the device */
static void rzip_chunk(rzip_control *control, const uchar *buf, i64 chunk_size, int chunk_bytes, const i64 *hash_index,
                       md5_ctx *md5, int fd_out) {
    void *ss = open_stream_out(control, fd_out, NUM_STREAMS, chunk_size, (char)chunk_bytes);
    if (!ss) fatal(control, "Failed to open streams in rzip_chunk\n");

    pthread_t th;
    struct md5_job job = {md5, buf, chunk_size};
    if (pthread_create(&th, NULL, md5_thread, &job)) fatal(control, "Failed to start the checksum thread\n");

    mrz_params prm;
    mrz_result res;
    memset(&prm, 0, sizeof prm);
    prm.level = control->rzip_compression_level;
    prm.chunk_bytes = chunk_bytes;
    prm.victim_round_in = victim_round;
    memcpy(prm.hash_index, hash_index, sizeof prm.hash_index);
    ensure_gpu(control, chunk_size);
    int rc = mrz_rzip_chunk(control->gpu, buf, chunk_size, &prm, control->pin_s0, control->s0_cap, control->pin_s1,
                            control->s1_cap, &res);
    if (rc) fatal(control, "GPU rzip failed (%d): %s\n", rc, mrz_last_error(control->gpu));
    victim_round = res.victim_round_out;

    control->matches += res.matches;
    control->match_bytes += res.match_bytes;
    control->literals += res.literals;
    control->literal_bytes += res.literal_bytes;
    control->tag_hits += res.tag_hits;
    control->tag_misses += res.tag_misses;
    control->inserts += res.inserts;
    control->gpu_ms += res.ms_kernels;
    control->gpu_scan_ms += res.ms_tagscan;
    control->gpu_auto_ms += res.ms_automaton;
    control->gpu_emit_ms += res.ms_emit;
    control->gpu_launches += res.gpu_launches;

    replay_streams(control, ss, control->pin_s0, res.s0_len, control->pin_s1, res.s1_len, chunk_bytes);
    pthread_join(th, NULL);
    if (close_stream_out(control, ss)) fatal(control, "Failed to flush/close streams in rzip_chunk\n");
}

void rzip_fd(rzip_control *control, int fd_in, int fd_out) {
    struct stat s, s2;
    struct statvfs fbuf;
    struct timeval start, current;
    i64 len = 0, hash_index[256];
    int pass = 0;
    md5_ctx md5;

    if (mrz_device_count() <= 0) fatal(control, "No CUDA device: the rzip pre-processor of this build runs on the GPU only\n");
    if (STDIN) fatal(control, "stdin input is not supported by this build yet; give a file name\n");
    md5_init(&md5);
    if (fstat(fd_in, &s)) fatal(control, "Failed to stat fd_in in rzip_fd\n");
    len = control->st_size = s.st_size;
    if (VERBOSE) fprintf(stderr, "File size: %lld\n", (long long)len);

    if (fstatvfs(fd_out, &fbuf)) fatal(control, "Failed to fstatvfs in compress_file\n");
    if ((i64)fbuf.f_bsize * (i64)fbuf.f_bavail < control->st_size) {
        if (FORCE_REPLACE)
            fprintf(stderr, "Warning, possibly inadequate free space detected, but attempting to compress due to -f option being used.\n");
        else
            fatal(control, "Possibly inadequate free space to compress file, use -f to override.\n");
    }

    /* rzip.c:875-894 */
    control->max_mmap = control->maxram;
    round_to_page(control, &control->max_mmap);
    if (UNLIMITED)
        control->max_chunk = control->st_size;
    else if (control->window)
        control->max_chunk = control->window * CHUNK_MULTIPLE;
    else
        control->max_chunk = control->ramsize / 3 * 2;
    if (control->max_mmap > control->max_chunk) control->max_mmap = control->max_chunk;
    if (control->max_chunk < control->st_size) round_to_page(control, &control->max_chunk);

    init_hash_indexes(hash_index);
    gettimeofday(&start, NULL);
    control->threads = 1; /* prepare_streamout_threads, stream.c:737: no back end => one writer */

    while (!pass || len > 0) {
        i64 offset = s.st_size - len;
        i64 chunk_size = control->max_chunk < len ? control->max_chunk : len;
        int bits = 8, chunk_bytes;
        const uchar *buf = NULL;

        if (chunk_size) {
            /* the reference maps the chunk (or slides a smaller window over it); the GPU needs the bytes once */
            buf = mmap(NULL, (size_t)chunk_size, PROT_READ, MAP_SHARED, fd_in, offset);
            if (buf == MAP_FAILED) fatal(control, "Failed to mmap %s: %s\n", control->infile, strerror(errno));
            madvise((void *)buf, (size_t)chunk_size, MADV_SEQUENTIAL);
        }
        if (VERBOSE && !pass && chunk_size) {
            i64 passes = s.st_size / chunk_size + !!(s.st_size % chunk_size);
            fprintf(stderr, passes == 1 ? "Will take 1 pass\n" : "Will take %lld passes\n", (long long)passes);
        }
        while (chunk_size >> bits > 0) bits++; /* rzip.c:1006-1008 */
        chunk_bytes = bits / 8 + (bits % 8 ? 1 : 0);
        if (MAX_VERBOSE) fprintf(stderr, "Chunk size: %lld\nByte width: %d\n", (long long)chunk_size, chunk_bytes);
        pass++;
        if (chunk_size == len) control->eof = 1; /* rzip.c:1049 */
        rzip_chunk(control, buf, chunk_size, chunk_bytes, hash_index, &md5, fd_out);
        if (buf) munmap((void *)buf, (size_t)chunk_size);
        len -= chunk_size;
    }

    /* hash trailer, rzip.c:1069-1091 */
    md5_final(&md5, control->hash_resblock);
    if (MAX_VERBOSE) {
        fprintf(stderr, "MD5: ");
        for (int j = 0; j < 16; j++) fprintf(stderr, "%02x", control->hash_resblock[j]);
        fprintf(stderr, "\n");
    }
    if (write(fd_out, control->hash_resblock, 16) != 16) fatal(control, "Failed to write md5 in rzip_fd\n");

    gettimeofday(&current, NULL);
    double tdiff = (current.tv_sec - start.tv_sec) + 1e-6 * (current.tv_usec - start.tv_usec);
    if (tdiff <= 0) tdiff = 1e-6;
    fstat(fd_out, &s2);
    if (MAX_VERBOSE) { /* rzip.c:1108-1115, same casts */
        fprintf(stderr, "matches=%u match_bytes=%u\n", (unsigned)control->matches, (unsigned)control->match_bytes);
        fprintf(stderr, "literals=%u literal_bytes=%u\n", (unsigned)control->literals, (unsigned)control->literal_bytes);
        fprintf(stderr, "true_tag_positives=%u false_tag_positives=%u\n", (unsigned)control->tag_hits, (unsigned)control->tag_misses);
        fprintf(stderr, "inserts=%u match %.3f\n", (unsigned)control->inserts, (1.0 + control->match_bytes) / control->literal_bytes);
        fprintf(stderr, "gpu: %.1f ms in kernels (tag scan %.1f, automaton %.1f, emit+crc %.1f), %lld launches\n", control->gpu_ms,
                control->gpu_scan_ms, control->gpu_auto_ms, control->gpu_emit_ms, (long long)control->gpu_launches);
    }
    if (SHOW_PROGRESS)
        fprintf(stderr, "%s - Compression Ratio: %.3f. bpb: %.3f. Average Compression Speed: %6.3fMB/s.\n", control->infile,
                1.0 * s.st_size / s2.st_size, 8.0 * s2.st_size / (s.st_size ? s.st_size : 1), (s.st_size / (double)ONE_MB) / tdiff);
}

void rzip_release_gpu(rzip_control *control) {
    if (!control->gpu) return;
    mrz_ctx_destroy(control->gpu);
    mrz_host_free(control->pin_s0);
    mrz_host_free(control->pin_s1);
    control->gpu = NULL;
    control->pin_s0 = control->pin_s1 = NULL;
}
