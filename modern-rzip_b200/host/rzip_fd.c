/* rzip_fd.c — host side of the rzip pre-processor: same entry point and chunking rules as the
 * reference's src/rzip.c, with the hot loop (hash_search, rzip.c:507-667) replaced by calls into
 * libmrz_gpu.so.
 *
 *   rzip_fd        rzip.c:807-1132   chunk sizing (:875-894), per-chunk loop (:915-1061), hash trailer
 *   rzip_chunk     rzip.c:763-792    open_stream_out -> [device] -> close_stream_out
 *   init_hash_indexes rzip.c:669-673 (the literal random() calls stay on the host: process-global state)
 *
 * What is different by design:
 *  * the reference walks the chunks of a file one after another on one core (rzip.c:915-1061).  Here the chunks
 *    are handed, in order, to SLOTS: `-G` chunks in flight on each of the `-g` devices, one worker thread per
 *    slot.  A worker reads its chunk through two pinned slabs (read of slab i+1 overlaps the H2D copy of slab i),
 *    runs the device search, brings the two streams back and hands them to the stream writer when it is the
 *    chunk's turn: archives are written strictly in chunk order (rzip.c:1049 eof on the last one).
 *  * the one scalar that links chunks - the static victim_round of insert_hash (rzip.c:259) - is carried exactly:
 *    a chunk whose predecessor has finished starts with its victim_round_out; a chunk that is ready earlier runs
 *    ALL max_chain_len values of victim_round_in at once (mrz_run_variants: one kernel launch, one thread-block
 *    cluster and one hash table per value) and keeps the variant the chain selects once the predecessor's value is
 *    published.  No chunk waits for another one's search, and none is searched twice.  The chain is 16 integers
 *    per chunk inside one process: no collective is needed (the multi-process form of bench.py uses one all-gather).
 *  * the chunk is handed to the GPU whole (no sliding mmap, rzip.c:75-158: a chunk fits HBM), the per-page checksum
 *    threads (rzip.c:488-505, :601-661) are gone - CRC32 comes back from the device inside stream 0 and MD5 runs
 *    in ONE helper thread over the file while the GPUs work.
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

#define SLAB ((i64)32 << 20)      /* pinned staging slab */
#define MAX_VARIANTS 32           /* run all victim_round_in values side by side up to this chain length (levels <= 8) */
static const int kChain[10] = {1, 2, 2, 2, 3, 4, 6, 16, 32, 128}; /* levels[].max_chain_len, rzip.c:65-73 */

/* process-lifetime state, as in the reference: the static in insert_hash (rzip.c:259) */
static i64 victim_round = 0;
i64 rzip_victim_round(void) { return victim_round; }

static void init_hash_indexes(i64 hash_index[256]) { /* rzip.c:669-673; gcc evaluates left to right */
    for (int i = 0; i < 256; i++) {
        i64 a = random();
        i64 b = random();
        hash_index[i] = (a << 16) ^ b;
    }
}

struct slot {
    int device, index;
    mrz_ctx *ctx[MAX_VARIANTS]; /* ctx[0] owns the chunk; the others are its variant contexts (created on demand) */
    int nctx;
    i64 cap;                    /* chunk capacity of ctx[0] */
    uchar *pin[2];              /* pinned slabs */
    uchar *s0, *s1;             /* the chunk's two streams on the host */
    i64 s0_cap, s1_cap;
};
struct engine {
    int nslots, slots_per_device;
    struct slot *slots;
};

/* one rzip_fd call */
struct job {
    rzip_control *control;
    int fd_in, fd_out, level, nvar;
    i64 len, max_chunk;
    int n_chunks;
    const i64 *hash_index;
    pthread_mutex_t mu;
    pthread_cond_t cv;
    int next_chunk;   /* next chunk to hand out */
    int place_turn;   /* next chunk to be given its place in the archive */
    i64 next_pos;     /* file offset of that chunk */
    i64 *vr_in;       /* vr_in[k]: victim_round entering chunk k, valid when vr_known[k] */
    char *vr_known;
    struct engine *eng;
};
struct worker_arg {
    struct job *job;
    struct slot *slot;
};

struct md5_job {
    rzip_control *control;
    md5_ctx *ctx;
    int fd;
    i64 len;
};
static void *md5_file_thread(void *arg) {
    struct md5_job *j = arg;
    /* MD5 is one serial chain over the whole input (~0.7 GB/s on one core): this thread bounds `mrzip -n` on large files,
     * so it hashes straight from the page cache through 256 MiB mappings (no copy); pread is the fallback */
    const i64 win = 256 << 20;
    i64 done = 0;
    while (done < j->len) {
        i64 want = j->len - done < win ? j->len - done : win;
        void *m = mmap(NULL, (size_t)want, PROT_READ, MAP_SHARED, j->fd, (off_t)done);
        if (m == MAP_FAILED) break;
        madvise(m, (size_t)want, MADV_SEQUENTIAL);
        madvise(m, (size_t)want, MADV_WILLNEED);
        md5_update(j->ctx, m, (size_t)want);
        munmap(m, (size_t)want);
        done += want;
    }
    if (done >= j->len) return NULL;
    const i64 step = 8 << 20;
    uchar *buf = malloc((size_t)step);
    if (!buf) fatal(j->control, "Failed to allocate the checksum buffer\n");
    for (i64 o = done; o < j->len;) {
        i64 want = j->len - o < step ? j->len - o : step;
        ssize_t r = pread(j->fd, buf, (size_t)want, o);
        if (r < 0 && errno == EINTR) continue;
        if (r <= 0) fatal(j->control, "Failed to read %s for the checksum\n", j->control->infile);
        md5_update(j->ctx, buf, (size_t)r);
        o += r;
    }
    free(buf);
    return NULL;
}

static int pow2_floor(int x) {
    int p = 1;
    while (p * 2 <= x) p *= 2;
    return p;
}

static void release_slot(struct slot *s) {
    for (int v = s->nctx - 1; v >= 0; v--) mrz_ctx_destroy(s->ctx[v]);
    s->nctx = 0;
    s->cap = 0;
}

static void ensure_slot(rzip_control *control, struct slot *s, i64 chunk) {
    if (s->nctx && s->cap >= chunk) return;
    release_slot(s);
    int rc = mrz_ctx_create(s->device, chunk, &s->ctx[0]);
    if (rc) fatal(control, "Cannot set up the GPU rzip engine on device %d (%d): %s\n", s->device, rc, mrz_last_error(NULL));
    s->nctx = 1;
    s->cap = chunk;
    if (!s->pin[0]) {
        s->pin[0] = mrz_host_alloc(SLAB);
        s->pin[1] = mrz_host_alloc(SLAB);
        if (!s->pin[0] || !s->pin[1]) fatal(control, "Cannot allocate pinned staging slabs\n");
    }
}
static void ensure_variants(rzip_control *control, struct slot *s, int nvar) {
    while (s->nctx < nvar) {
        int rc = mrz_ctx_create_variant(s->ctx[0], &s->ctx[s->nctx]);
        if (rc) fatal(control, "Cannot set up a variant context on device %d (%d): %s\n", s->device, rc, mrz_last_error(NULL));
        s->nctx++;
    }
}
static void ensure_host(rzip_control *control, uchar **p, i64 *cap, i64 need) {
    if (*cap >= need) return;
    free(*p);
    *p = malloc((size_t)need);
    if (!*p) fatal(control, "Cannot allocate %lld bytes for a stream\n", (long long)need);
    *cap = need;
}

/* read chunk [off, off+n) of the file into the slot's device context: pread into one pinned slab while the other is
 * in flight to the device */
static void upload_chunk(rzip_control *control, struct slot *s, int fd_in, i64 off, i64 n) {
    int which = 0;
    for (i64 o = 0; o < n;) {
        i64 want = n - o < SLAB ? n - o : SLAB, got = 0;
        while (got < want) {
            ssize_t r = pread(fd_in, s->pin[which] + got, (size_t)(want - got), off + o + got);
            if (r < 0 && errno == EINTR) continue;
            if (r <= 0) fatal(control, "Failed to read %s\n", control->infile);
            got += r;
        }
        /* the copy that used the OTHER slab two parts ago has to be finished before we overwrite it next round */
        if (o) {
            int rc = mrz_upload_wait(s->ctx[0]);
            if (rc) fatal(control, "GPU upload failed (%d): %s\n", rc, mrz_last_error(s->ctx[0]));
        }
        int rc = mrz_upload_part(s->ctx[0], o, s->pin[which], want);
        if (rc) fatal(control, "GPU upload failed (%d): %s\n", rc, mrz_last_error(s->ctx[0]));
        which ^= 1;
        o += want;
    }
    int rc = mrz_upload_finish(s->ctx[0], n);
    if (rc) fatal(control, "GPU upload failed (%d): %s\n", rc, mrz_last_error(s->ctx[0]));
}

/* stream `streamno` of the chosen context to pageable host memory through the pinned slabs */
static void download_stream(rzip_control *control, struct slot *s, mrz_ctx *ctx, int streamno, uchar *dst, i64 len) {
    for (i64 o = 0; o < len; o += SLAB) {
        i64 want = len - o < SLAB ? len - o : SLAB;
        int rc = mrz_download_part(ctx, streamno, o, s->pin[0], want);
        if (rc) fatal(control, "GPU download failed (%d): %s\n", rc, mrz_last_error(ctx));
        memcpy(dst + o, s->pin[0], (size_t)want);
    }
}

static void *worker(void *argp) {
    struct worker_arg *wa = argp;
    struct job *job = wa->job;
    struct slot *s = wa->slot;
    rzip_control *control = job->control;
    const int spd = job->eng->slots_per_device;
    mrz_result res[MAX_VARIANTS];
    for (;;) {
        pthread_mutex_lock(&job->mu);
        const int k = job->next_chunk++;
        pthread_mutex_unlock(&job->mu);
        if (k >= job->n_chunks) break;
        const i64 off = (i64)k * job->max_chunk;
        const i64 n = job->len - off < job->max_chunk ? job->len - off : job->max_chunk;
        int bits = 8, chunk_bytes;
        while (n >> bits > 0) bits++; /* rzip.c:1006-1008 */
        chunk_bytes = bits / 8 + (bits % 8 ? 1 : 0);

        upload_chunk(control, s, job->fd_in, off, n);

        mrz_params prm;
        memset(&prm, 0, sizeof prm);
        prm.level = job->level;
        prm.chunk_bytes = chunk_bytes;
        memcpy(prm.hash_index, job->hash_index, sizeof prm.hash_index);

        pthread_mutex_lock(&job->mu);
        const int known = job->vr_known[k];
        i64 vr = job->vr_in[k];
        pthread_mutex_unlock(&job->mu);
        mrz_ctx *chosen = s->ctx[0];
        int rc;
        if (known || job->nvar == 1 || job->nvar > MAX_VARIANTS) {
            if (!known) { /* chain too long to run side by side: wait for the predecessor (sequential carry) */
                pthread_mutex_lock(&job->mu);
                while (!job->vr_known[k]) pthread_cond_wait(&job->cv, &job->mu);
                vr = job->vr_in[k];
                pthread_mutex_unlock(&job->mu);
            }
            prm.victim_round_in = vr;
            mrz_ctx_set_cluster(s->ctx[0], spd == 1 ? 0 : pow2_floor(148 / spd > 8 ? 8 : 148 / spd));
            rc = mrz_run_automaton(s->ctx[0], n, &prm, &res[0]);
            if (rc) fatal(control, "GPU rzip failed (%d): %s\n", rc, mrz_last_error(s->ctx[0]));
            vr = res[0].victim_round_out;
        } else {
            /* the predecessor is still being searched: every victim_round_in at once, one launch */
            ensure_variants(control, s, job->nvar);
            int cl = 148 / (spd * job->nvar);
            mrz_ctx_set_cluster(s->ctx[0], cl < 1 ? 1 : pow2_floor(cl > 8 ? 8 : cl));
            rc = mrz_run_variants(s->ctx, job->nvar, n, &prm, res);
            if (rc) fatal(control, "GPU rzip (variants) failed (%d): %s\n", rc, mrz_last_error(NULL));
            pthread_mutex_lock(&job->mu);
            while (!job->vr_known[k]) pthread_cond_wait(&job->cv, &job->mu);
            const i64 vin = job->vr_in[k];
            pthread_mutex_unlock(&job->mu);
            chosen = s->ctx[vin];
            vr = res[vin].victim_round_out;
            control->variant_launches++;
        }
        /* the next chunk can pick its variant now */
        pthread_mutex_lock(&job->mu);
        job->vr_in[k + 1] = vr;
        job->vr_known[k + 1] = 1;
        pthread_cond_broadcast(&job->cv);
        pthread_mutex_unlock(&job->mu);

        mrz_result r;
        rc = mrz_run_emit(chosen, &r);
        if (rc) fatal(control, "GPU emission failed (%d): %s\n", rc, mrz_last_error(chosen));
        /* A chunk's place in the archive follows from the stream LENGTHS of the chunks before it (rzip.c:763-792 writes
         * chunk after chunk): places are handed out strictly in chunk order, then every worker downloads and writes its
         * own chunk with positional writes - several chunks are on their way to the file at the same time. */
        pthread_mutex_lock(&job->mu);
        while (job->place_turn != k) pthread_cond_wait(&job->cv, &job->mu);
        const int last = k == job->n_chunks - 1;
        if (last) control->eof = 1; /* rzip.c:1049 */
        if (MAX_VERBOSE) fprintf(stderr, "Chunk size: %lld\nByte width: %d\n", (long long)n, chunk_bytes);
        struct stream_info *ss = open_stream_out_at(control, job->fd_out, NUM_STREAMS, n, (char)chunk_bytes, job->next_pos, last);
        if (!ss) fatal(control, "Failed to open streams in rzip_chunk\n");
        const i64 chunk_end = job->next_pos + stream_out_chunk_size(ss, r.s0_len, r.s1_len);
        job->next_pos = chunk_end;
        job->place_turn = k + 1;
        pthread_cond_broadcast(&job->cv);
        pthread_mutex_unlock(&job->mu);

        ensure_host(control, &s->s0, &s->s0_cap, r.s0_len + 64);
        ensure_host(control, &s->s1, &s->s1_cap, r.s1_len + 64);
        download_stream(control, s, chosen, 0, s->s0, r.s0_len);
        download_stream(control, s, chosen, 1, s->s1, r.s1_len);
        replay_streams(control, ss, s->s0, r.s0_len, s->s1, r.s1_len, chunk_bytes);
        ss->expect_end = chunk_end;
        if (close_stream_out(control, ss)) fatal(control, "Failed to flush/close streams in rzip_chunk\n");
        pthread_mutex_lock(&job->mu);
        control->matches += r.matches;
        control->match_bytes += r.match_bytes;
        control->literals += r.literals;
        control->literal_bytes += r.literal_bytes;
        control->tag_hits += r.tag_hits;
        control->tag_misses += r.tag_misses;
        control->inserts += r.inserts;
        control->gpu_ms += r.ms_kernels;
        control->gpu_scan_ms += r.ms_tagscan;
        control->gpu_auto_ms += r.ms_automaton;
        control->gpu_emit_ms += r.ms_emit;
        control->gpu_launches += r.gpu_launches;
        pthread_mutex_unlock(&job->mu);
    }
    return NULL;
}

void rzip_release_gpu(rzip_control *control);
static struct engine *get_engine(rzip_control *control, int n_chunks) {
    struct engine *eng = control->engine;
    if (!control->n_devices) { control->n_devices = 1; control->devices[0] = 0; }
    int spd = control->slots_per_device;
    if (!spd) { /* chunks in flight per device: as many as the file has to offer, up to 4 */
        spd = (n_chunks + control->n_devices - 1) / control->n_devices;
        if (spd > 4) spd = 4;
        if (spd < 1) spd = 1;
    }
    if (eng && eng->slots_per_device != spd) { /* a later file of the same command line wants another geometry */
        rzip_release_gpu(control);
        eng = NULL;
    }
    if (!eng) {
        int ndev = mrz_device_count();
        for (int d = 0; d < control->n_devices; d++)
            if (control->devices[d] < 0 || control->devices[d] >= ndev) fatal(control, "No CUDA device %d (the box has %d)\n", control->devices[d], ndev);
        eng = calloc(1, sizeof *eng);
        eng->slots_per_device = spd;
        eng->nslots = spd * control->n_devices;
        eng->slots = calloc((size_t)eng->nslots, sizeof *eng->slots);
        for (int i = 0; i < eng->nslots; i++) { /* slot i -> device i mod G: consecutive chunks go to different devices */
            eng->slots[i].index = i;
            eng->slots[i].device = control->devices[i % control->n_devices];
        }
        control->engine = eng;
    }
    return eng;
}

void rzip_fd(rzip_control *control, int fd_in, int fd_out) {
    struct stat s, s2;
    struct statvfs fbuf;
    struct timeval start, current;
    i64 len = 0, hash_index[256];
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

    /* the chunks of the file (rzip.c:915-924: every chunk is max_chunk bytes but the last); an empty file is one empty chunk */
    struct job job;
    memset(&job, 0, sizeof job);
    job.control = control;
    job.fd_in = fd_in;
    job.fd_out = fd_out;
    job.level = control->rzip_compression_level;
    job.nvar = kChain[job.level];
    job.len = len;
    job.max_chunk = control->max_chunk > 0 ? control->max_chunk : 1;
    job.n_chunks = len ? (int)((len + job.max_chunk - 1) / job.max_chunk) : 1;
    job.hash_index = hash_index;
    job.vr_in = calloc((size_t)job.n_chunks + 2, sizeof(i64));
    job.vr_known = calloc((size_t)job.n_chunks + 2, 1);
    job.vr_in[0] = victim_round; /* carried over from the previous file of this process */
    job.vr_known[0] = 1;
    job.next_pos = lseek(fd_out, 0, SEEK_CUR); /* the first chunk starts where the container header ended */
    if (job.next_pos == -1) fatal(control, "Failed to get the position of the first chunk\n");
    pthread_mutex_init(&job.mu, NULL);
    pthread_cond_init(&job.cv, NULL);
    if (VERBOSE) fprintf(stderr, job.n_chunks == 1 ? "Will take 1 pass\n" : "Will take %d passes\n", job.n_chunks);

    struct engine *eng = get_engine(control, job.n_chunks);
    job.eng = eng;
    int nworkers = eng->nslots < job.n_chunks ? eng->nslots : job.n_chunks;
    const i64 cap = len < job.max_chunk ? len : job.max_chunk;
    for (int i = 0; i < nworkers; i++) ensure_slot(control, &eng->slots[i], cap);
    if (MAX_VERBOSE)
        fprintf(stderr, "GPU engine: %d device(s), %d chunk(s) in flight per device, %d victim_round variants per early chunk\n",
                control->n_devices, eng->slots_per_device, job.nvar <= MAX_VARIANTS ? job.nvar : 1);

    pthread_t md5_th;
    struct md5_job mj = {control, &md5, fd_in, len};
    if (pthread_create(&md5_th, NULL, md5_file_thread, &mj)) fatal(control, "Failed to start the checksum thread\n");
    pthread_t th[256];
    struct worker_arg wa[256];
    for (int i = 0; i < nworkers; i++) {
        wa[i].job = &job;
        wa[i].slot = &eng->slots[i];
        if (pthread_create(&th[i], NULL, worker, &wa[i])) fatal(control, "Failed to start a chunk worker\n");
    }
    for (int i = 0; i < nworkers; i++) pthread_join(th[i], NULL);
    victim_round = job.vr_in[job.n_chunks];
    gettimeofday(&current, NULL);
    const double t_chunks = (current.tv_sec - start.tv_sec) + 1e-6 * (current.tv_usec - start.tv_usec);
    pthread_join(md5_th, NULL);
    free(job.vr_in);
    free(job.vr_known);

    /* hash trailer, rzip.c:1069-1091: behind the last chunk (the chunks were written with positional writes) */
    if (lseek(fd_out, job.next_pos, SEEK_SET) == -1) fatal(control, "Failed to seek to the end of the chunks\n");
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
        fprintf(stderr, "gpu: %.1f ms in kernels (tag map + crc %.1f, automaton %.1f, emit %.1f), %lld launches, %lld chunk(s) searched for every victim_round\n",
                control->gpu_ms, control->gpu_scan_ms, control->gpu_auto_ms, control->gpu_emit_ms, (long long)control->gpu_launches,
                (long long)control->variant_launches);
        fprintf(stderr, "chunks written after %.3f s; rzip_fd (with the whole-file MD5 thread) %.3f s\n", t_chunks, tdiff);
    }
    if (SHOW_PROGRESS)
        fprintf(stderr, "%s - Compression Ratio: %.3f. bpb: %.3f. Average Compression Speed: %6.3fMB/s.\n", control->infile,
                1.0 * s.st_size / s2.st_size, 8.0 * s2.st_size / (s.st_size ? s.st_size : 1), (s.st_size / (double)ONE_MB) / tdiff);
}

void rzip_release_gpu(rzip_control *control) {
    struct engine *eng = control->engine;
    if (!eng) return;
    for (int i = 0; i < eng->nslots; i++) {
        struct slot *s = &eng->slots[i];
        release_slot(s);
        mrz_host_free(s->pin[0]);
        mrz_host_free(s->pin[1]);
        free(s->s0);
        free(s->s1);
    }
    free(eng->slots);
    free(eng);
    control->engine = NULL;
}
