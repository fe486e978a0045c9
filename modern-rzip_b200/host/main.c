/* main.c — `mrzip` command line for the B200 build: the rzip-only (`-n`) path of the reference's
 * src/main.c (option table :252-303, parsing :356-569, defaults :575-583, per-file loop :651-747).
 * Same flags and meanings for the subset that shapes an `-n` archive; the back-end compressors,
 * decompression and info stay with the reference's host code and are refused here with a message.
 *   -n            rzip pre-processing only (required in this build to compress)
 *   -d            decompress an rzip-only archive (the device token replay of runzip_fd.c)
 *   -L/-R <1-9>   compression level / rzip level (levels[] row, rzip.c:65-73, :896)
 *   -w <n>        window, units of 100 MiB     -m <n>  ram size, units of 100 MiB    -U  unlimited window
 *   -p <n>        threads (accepted; forced to 1 for -n as in stream.c:737)
 *   -o <file> -O <dir> -S <suffix> -f -D -K -q -v -vv -V -h
 *   -g <d,d,...>  CUDA devices to spread the chunks of a file over (new; default 0)
 *   -G <n>        chunks kept in flight per device (new; default chosen from the chunk count)
 */
#define _GNU_SOURCE
#include <getopt.h>
#include <stdlib.h>
#include <string.h>
#include <sys/stat.h>
#include <sys/time.h>

#include "mrzip_host.h"

static void usage(void) {
    fprintf(stderr,
            "mrzip (B200 rzip pre-processor build) 0.9.0\n"
            "usage: mrzip -n [options] <file...>   |   mrzip -d [options] <archive...>\n"
            "  -n  rzip pre-processing only (the only mode of this build; back-ends stay with the reference)\n"
            "  -L level  -R rzip-level  -w window(x100MB)  -m ram(x100MB)  -U  -p threads\n"
            "  -o outfile  -O outdir  -S suffix  -f  -D  -K  -q  -v  -V  -g cuda-devices(0,1,..)  -G chunks-in-flight-per-device\n");
}

int main(int argc, char *argv[]) {
    rzip_control control_s, *control = &control_s;
    int c;
    char *endptr;
    initialise_control(control);
    bool no_compress = false, decompress = false;

    static const struct option long_options[] = {
        {"no-compress", no_argument, 0, 'n'}, {"level", required_argument, 0, 'L'}, {"rzip-level", required_argument, 0, 'R'},
        {"window", required_argument, 0, 'w'}, {"maxram", required_argument, 0, 'm'}, {"unlimited", no_argument, 0, 'U'},
        {"threads", required_argument, 0, 'p'}, {"outfile", required_argument, 0, 'o'}, {"outdir", required_argument, 0, 'O'},
        {"suffix", required_argument, 0, 'S'}, {"force", no_argument, 0, 'f'}, {"delete", no_argument, 0, 'D'},
        {"keep-broken", no_argument, 0, 'K'}, {"quiet", no_argument, 0, 'q'}, {"verbose", no_argument, 0, 'v'},
        {"version", no_argument, 0, 'V'}, {"help", no_argument, 0, 'h'}, {"decompress", no_argument, 0, 'd'},
        {"test", no_argument, 0, 't'}, {"info", no_argument, 0, 'i'}, {0, 0, 0, 0}};

    while ((c = getopt_long(argc, argv, "nL:R:w:m:Up:o:O:S:fDKqvVhdtig:G:", long_options, NULL)) != -1) {
        switch (c) {
            case 'n': no_compress = true; control->flags |= FLAG_NO_COMPRESS; break;
            case 'L':
                control->compression_level = (int)strtol(optarg, &endptr, 10);
                if (*endptr) fatal(control, "Extra characters after compression level: '%s'\n", endptr);
                if (control->compression_level < 1 || control->compression_level > 9) fatal(control, "Invalid compression level (must be 1-9)\n");
                break;
            case 'R':
                control->rzip_compression_level = (int)strtol(optarg, &endptr, 10);
                if (*endptr) fatal(control, "Extra characters after rzip compression level: '%s'\n", endptr);
                if (control->rzip_compression_level < 1 || control->rzip_compression_level > 9) fatal(control, "Invalid rzip compression level (must be 1-9)\n");
                break;
            case 'w':
                control->window = strtol(optarg, &endptr, 10);
                if (*endptr) fatal(control, "Extra characters after window size: '%s'\n", endptr);
                if (control->window < 1) fatal(control, "Window must be positive\n");
                break;
            case 'm':
                control->ramsize = strtol(optarg, &endptr, 10) * ONE_MB * 100;
                if (*endptr) fatal(control, "Extra characters after ramsize: '%s'\n", endptr);
                break;
            case 'U': control->flags |= FLAG_UNLIMITED; break;
            case 'p':
                control->threads = (int)strtol(optarg, &endptr, 10);
                if (*endptr) fatal(control, "Extra characters after number of threads: '%s'\n", endptr);
                if (control->threads < 1) fatal(control, "Must have at least one thread\n");
                break;
            case 'o': control->outname = optarg; break;
            case 'O': {
                size_t l = strlen(optarg);
                control->outdir = malloc(l + 2);
                strcpy(control->outdir, optarg);
                if (l && optarg[l - 1] != '/') strcat(control->outdir, "/");
                break;
            }
            case 'S': control->suffix = optarg; break;
            case 'f': control->flags |= FLAG_FORCE_REPLACE; break;
            case 'D': control->flags &= ~FLAG_KEEP_FILES; break;
            case 'K': control->flags |= FLAG_KEEP_BROKEN; break;
            case 'q': control->flags &= ~FLAG_SHOW_PROGRESS; break;
            case 'v':
                control->flags |= FLAG_SHOW_PROGRESS;
                if (control->flags & FLAG_VERBOSE) control->flags |= FLAG_MAX_VERBOSE;
                control->flags |= FLAG_VERBOSE;
                break;
            case 'g': { /* device list: -g 0,1,2,3 */
                char *q = optarg;
                control->n_devices = 0;
                while (*q && control->n_devices < 16) {
                    control->devices[control->n_devices++] = (int)strtol(q, &endptr, 10);
                    if (endptr == q) fatal(control, "Bad device list '%s'\n", optarg);
                    q = *endptr == ',' ? endptr + 1 : endptr;
                    if (*endptr && *endptr != ',') fatal(control, "Bad device list '%s'\n", optarg);
                }
                break;
            }
            case 'G':
                control->slots_per_device = (int)strtol(optarg, &endptr, 10);
                if (*endptr || control->slots_per_device < 1 || control->slots_per_device > 16) fatal(control, "-G wants 1..16 chunks in flight per device\n");
                break;
            case 'V': printf("modern-rzip (B200 rzip pre-processor) version 0.9.0\n"); return 0;
            case 'd': decompress = true; break;
            case 't': case 'i':
                fatal(control, "This build implements the rzip pre-processor (-n) and its inverse (-d); use the reference mrzip for -%c\n", c);
                break;
            case 'h': default: usage(); return c == 'h' ? 0 : 1;
        }
    }
    argc -= optind;
    argv += optind;
    if (!no_compress && !decompress)
        fatal(control, "This build accelerates the rzip pre-processor only: pass -n (the back-end compressors stay on the host in the reference build)\n");
    if (!control->rzip_compression_level) control->rzip_compression_level = control->compression_level; /* main.c:575 */
    if (control->outname && argc > 1) fatal(control, "Cannot specify output filename with more than 1 file\n");
    if (!control->suffix && !control->outname) control->suffix = ".lrz"; /* main.c:581 */
    if (UNLIMITED && control->window) {
        fprintf(stderr, "If -U used, cannot specify a window size with -w.\n");
        control->window = 0;
    }
    if (argc < 1) fatal(control, "stdin input is not supported by this build yet; give a file name\n");

    rzip_control base = *control;
    for (int i = 0; i < argc; i++) { /* main.c:651: files share random() state and victim_round */
        struct stat istat;
        struct timeval t0, t1;
        rzip_control local = base;
        local.engine = control->engine;
        local.infile = argv[i];
        if (stat(local.infile, &istat)) fatal(&local, "Failed to stat %s\n", local.infile);
        if (!S_ISREG(istat.st_mode)) fatal(&local, "mrzip only works directly on regular FILES.\n");
        setup_ram(&local);
        gettimeofday(&t0, NULL);
        if (decompress) {
            if (!decompress_file(&local)) return -1;
        } else if (!compress_file(&local))
            return -1;
        gettimeofday(&t1, NULL);
        control->engine = local.engine;
        double tt = (t1.tv_sec - t0.tv_sec) + 1e-6 * (t1.tv_usec - t0.tv_usec);
        if (local.flags & FLAG_SHOW_PROGRESS)
            fprintf(stderr, "Total time: %02d:%02d:%05.2f\n", (int)tt / 3600, (int)(tt / 60) % 60, tt - (int)tt / 3600 * 3600 - (int)(tt / 60) % 60 * 60);
    }
    rzip_release_gpu(control);
    return 0;
}
