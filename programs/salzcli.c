/*
 * salzcli.c - `salz` command line front end on top of the B200-native library.
 *
 * Same options, file naming, framing and log lines as the reference front end
 * (reference programs/salzcli.c:272-496), written afresh.  What differs is
 * the compress loop: instead of one salz_encode_safe call per block
 * (reference programs/salzcli.c:143-179) whole batches of blocks are handed
 * to salz_cuda_encode_stream, which returns the framed records
 * (u32 encoded_len + block) ready to be written.
 *
 * Container: u32 magic 0x53414C5A, u32 block size, then the records
 * (reference programs/salzcli.c:115-121,162-169).
 *
 * Not in the reference: --gpus N (first N devices, 0 = all visible).
 * Environment: SALZ_CUDA_DEVICE (default 0), SALZ_CUDA_DEVICES (count or a,b,c
 * list), SALZ_BATCH_MIB (per-GPU batch, default 256).
 */
#include <errno.h>
#include <getopt.h>
#include <limits.h>
#include <pthread.h>
#include <stdbool.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <sys/stat.h>
#include <unistd.h>

#include "common.h"
#include "salz.h"
#include "salz_cuda.h"

#define SALZ_MAGIC 0x53414C5Au
#define SUFFIX ".salz"

enum { MODE_COMPRESS, MODE_DECOMPRESS };
enum { QUIET_ALL = 0, QUIET_CRIT, QUIET_ERR, VERBOSE_INFO };

static int mode = MODE_COMPRESS;
static int verbosity = VERBOSE_INFO;
static bool force_overwrite;
static bool keep_source;
static int level = 5;
static int gpus_opt = -1;   /* --gpus N: first N devices, 0 = all; -1 = not given */

#define say(lvl, ...)                      \
    do {                                   \
        if ((lvl) <= verbosity) {          \
            fprintf(stderr, __VA_ARGS__);  \
            fputc('\n', stderr);           \
        }                                  \
    } while (0)

static const char *base_name(const char *path)
{
    const char *slash = strrchr(path, '/');
    return slash ? slash + 1 : path;
}

/* Output path: strip ".salz" when decompressing, append it (truncating over-long
 * names so that the result still fits NAME_MAX / PATH_MAX) when compressing. */
static void derive_outpath(const char *path, char out[PATH_MAX])
{
    size_t plen = strlen(path), slen = strlen(SUFFIX);
    if (mode == MODE_DECOMPRESS) {
        size_t keep = plen - slen;
        memcpy(out, path, keep);
        out[keep] = '\0';
        return;
    }
    size_t nlen = strlen(base_name(path));
    size_t keep = plen;
    if (nlen + slen + 1 > NAME_MAX)
        keep = plen - nlen + (NAME_MAX - (slen + 1));
    if (plen + slen + 1 > PATH_MAX)
        keep = PATH_MAX - (slen + 1);
    memcpy(out, path, keep);
    memcpy(out + keep, SUFFIX, slen);
    out[keep + slen] = '\0';
}

/* ---- compress: reader -> GPU(s) -> writer -------------------------------------
 * The reference reads a block, encodes it and writes it, one after the other
 * (reference programs/salzcli.c:143-179).  Here three threads share a ring of
 * NSLOT pinned buffer pairs: the reader fills slot i+1 and the writer drains
 * slot i-1 while the encoder (this thread) has slot i on the GPUs, so file I/O
 * runs under the kernels.  A slot holds `gpus x per-GPU batch` blocks; the
 * multi-GPU entry cuts it into one contiguous block range per device and
 * assembles the records after the NCCL all-gather of their sizes.  The encoder
 * context (device workspaces, NCCL ranks) is created once per process. */
#define NSLOT 3

enum { SLOT_FREE, SLOT_FILLED, SLOT_ENCODED };

struct slot {
    uint8_t *in, *out;
    size_t got, produced;
    bool last;
    int state;
};

struct pipe_ctx {
    FILE *in, *out;
    struct slot slot[NSLOT];
    size_t batch_bytes, block_len;
    pthread_mutex_t mu;
    pthread_cond_t cv;
    bool failed;          /* any stage failed: everybody drains and leaves */
    const char *why;      /* message of the first failure */
};

static void pipe_fail(struct pipe_ctx *p, const char *why)
{
    pthread_mutex_lock(&p->mu);
    if (!p->failed) {
        p->failed = true;
        p->why = why;
    }
    pthread_cond_broadcast(&p->cv);
    pthread_mutex_unlock(&p->mu);
}

static void *reader_main(void *arg)
{
    struct pipe_ctx *p = arg;
    for (int i = 0;; i = (i + 1) % NSLOT) {
        struct slot *s = &p->slot[i];
        pthread_mutex_lock(&p->mu);
        while (s->state != SLOT_FREE && !p->failed)
            pthread_cond_wait(&p->cv, &p->mu);
        const bool stop = p->failed;
        pthread_mutex_unlock(&p->mu);
        if (stop)
            break;
        const size_t got = fread(s->in, 1, p->batch_bytes, p->in);
        if (got != p->batch_bytes && ferror(p->in)) {
            pipe_fail(p, "Couldn't read from input stream");
            break;
        }
        /* The reference encodes one more (short or empty) block after every
         * full-size read, so a file whose size is a multiple of the block size,
         * or whose tail is <= 8 bytes, fails there; keep that behaviour. */
        const bool last = got != p->batch_bytes;
        if (got == 0 || (last && got % p->block_len == 0)) {
            pipe_fail(p, "Couldn't encode segment");
            break;
        }
        pthread_mutex_lock(&p->mu);
        s->got = got;
        s->last = last;
        s->state = SLOT_FILLED;
        pthread_cond_broadcast(&p->cv);
        pthread_mutex_unlock(&p->mu);
        if (last)
            break;
    }
    return NULL;
}

static void *writer_main(void *arg)
{
    struct pipe_ctx *p = arg;
    for (int i = 0;; i = (i + 1) % NSLOT) {
        struct slot *s = &p->slot[i];
        pthread_mutex_lock(&p->mu);
        while (s->state != SLOT_ENCODED && !p->failed)
            pthread_cond_wait(&p->cv, &p->mu);
        const bool stop = p->failed;
        pthread_mutex_unlock(&p->mu);
        if (stop)
            break;
        if (fwrite(s->out, 1, s->produced, p->out) != s->produced) {
            pipe_fail(p, "Couldn't write encoded segment to output stream");
            break;
        }
        const bool last = s->last;
        pthread_mutex_lock(&p->mu);
        s->state = SLOT_FREE;
        pthread_cond_broadcast(&p->cv);
        pthread_mutex_unlock(&p->mu);
        if (last)
            break;
    }
    return NULL;
}

/* Encoder context of the process: created by the first file, reused by the rest. */
static salz_cuda_multi *g_enc;
static uint8_t *g_in[NSLOT], *g_out[NSLOT];
static bool g_pinned[NSLOT];
static size_t g_batch_bytes, g_out_cap;

/* Slot buffers.  A stream of several slots reuses them many times: page-locked memory, so
 * that the copies run at full PCIe speed under the kernels.  A file that fits ONE slot is
 * copied once: pinning its buffers (~1 GB/s) would cost more than it saves, plain memory. */
static bool slot_alloc(int i, bool pinned)
{
    g_pinned[i] = pinned;
    g_in[i] = pinned ? salz_cuda_host_alloc(g_batch_bytes) : malloc(g_batch_bytes);
    g_out[i] = pinned ? salz_cuda_host_alloc(g_out_cap) : malloc(g_out_cap);
    return g_in[i] != NULL && g_out[i] != NULL;
}

static void encoder_release(void)
{
    for (int i = 0; i < NSLOT; i++) {
        if (g_pinned[i]) {
            salz_cuda_host_free(g_in[i]);
            salz_cuda_host_free(g_out[i]);
        } else {
            free(g_in[i]);
            free(g_out[i]);
        }
        g_in[i] = g_out[i] = NULL;
    }
    salz_cuda_multi_destroy(g_enc);
    g_enc = NULL;
}

/* Devices: --gpus N / SALZ_CUDA_DEVICES=N (first N devices, 0 = all) or
 * SALZ_CUDA_DEVICES=a,b,c (that list); default: SALZ_CUDA_DEVICE or device 0. */
static int pick_devices(int devs[64])
{
    const char *list = getenv("SALZ_CUDA_DEVICES");
    const int have = salz_cuda_device_count();
    int n = 0;
    if (gpus_opt >= 0) {
        n = gpus_opt == 0 ? have : gpus_opt;
        for (int i = 0; i < n && i < 64; i++)
            devs[i] = i;
    } else if (list && strchr(list, ',')) {
        char tmp[256];
        snprintf(tmp, sizeof(tmp), "%s", list);
        for (char *t = strtok(tmp, ","); t && n < 64; t = strtok(NULL, ","))
            devs[n++] = atoi(t);
    } else if (list) {
        n = atoi(list) == 0 ? have : atoi(list);
        for (int i = 0; i < n && i < 64; i++)
            devs[i] = i;
    } else {
        const char *one = getenv("SALZ_CUDA_DEVICE");
        devs[0] = one ? atoi(one) : 0;
        n = 1;
    }
    return n > 64 ? 64 : n;
}

/* SALZ_CLI_TIMING=1: where the wall time of a compress run goes (stderr). */
static void phase(const char *what, uint64_t *t)
{
    static int on = -1;
    if (on < 0)
        on = getenv("SALZ_CLI_TIMING") != NULL;
    uint64_t now = 0;
    get_time_ns(&now);
    if (on && *t)
        fprintf(stderr, "salz: %-28s %8.1f ms\n", what, (double)(now - *t) / 1e6);
    *t = now;
}

static int encoder_acquire(size_t block_len, size_t file_size)
{
    if (g_enc != NULL)
        return 0;
    uint64_t t = 0;
    phase("", &t);
    int devs[64];
    const int ndev = pick_devices(devs);
    if (ndev <= 0) {
        say(QUIET_ERR, "Couldn't create CUDA encoder context (no GPU?)");
        return -1;
    }
    /* Per-GPU batch: large enough to fill the device (default 256 MiB: ~23 GB of
     * workspace), no larger than this file's share of blocks needs. */
    const char *env = getenv("SALZ_BATCH_MIB");
    size_t per_gpu = (size_t)(env ? atoi(env) : 256) << 20;
    if (per_gpu < block_len)
        per_gpu = block_len;
    size_t blocks = per_gpu / block_len;
    const size_t file_blocks = file_size / block_len + 1;
    const size_t share = divup(file_blocks, (size_t)ndev);
    if (blocks > share || share <= blocks + blocks / 4)
        blocks = share;   /* the whole file in one slot (a little over the default batch still is) */
    if (blocks > 65535)
        blocks = 65535;
    g_enc = salz_cuda_multi_create(devs, ndev, block_len, blocks);
    if (g_enc == NULL) {
        say(QUIET_ERR, "Couldn't create CUDA encoder context (no GPU?)");
        return -1;
    }
    phase("contexts (CUDA init, HBM)", &t);
    g_batch_bytes = blocks * block_len * (size_t)ndev;
    g_out_cap = blocks * (size_t)ndev * ((size_t)salz_encoded_len_max(block_len) + 4);
    /* a file that fits one slot needs one buffer pair only */
    const int nslot = file_size < g_batch_bytes ? 1 : NSLOT;
    for (int i = 0; i < nslot; i++) {
        if (!slot_alloc(i, nslot > 1)) {
            say(QUIET_ERR, "Couldn't allocate memory: (%zu bytes)", g_batch_bytes + g_out_cap);
            encoder_release();
            return -1;
        }
    }
    phase("host buffers", &t);
    return 0;
}

static int do_compress(FILE *in, FILE *out, size_t file_size)
{
    const uint32_t block_len = 1u << (15 + level);
    const uint32_t magic = SALZ_MAGIC;
    if (encoder_acquire(block_len, file_size) != 0)
        return -1;
    const int want = file_size < g_batch_bytes ? 1 : NSLOT;   /* a later, larger file: all slots */
    for (int i = 0; i < want; i++)
        if (g_in[i] == NULL || g_out[i] == NULL) {
            if (!slot_alloc(i, want > 1)) {
                say(QUIET_ERR, "Couldn't allocate memory: (%zu bytes)", g_batch_bytes + g_out_cap);
                return -1;
            }
        }
    if (fwrite(&magic, 1, 4, out) != 4 || fwrite(&block_len, 1, 4, out) != 4) {
        say(QUIET_ERR, "Couldn't write SALZ header to output");
        return -1;
    }
    struct pipe_ctx p;
    memset(&p, 0, sizeof(p));
    p.in = in;
    p.out = out;
    p.batch_bytes = g_batch_bytes;
    p.block_len = block_len;
    pthread_mutex_init(&p.mu, NULL);
    pthread_cond_init(&p.cv, NULL);
    int nslot = 0;
    for (int i = 0; i < NSLOT && g_in[i] && g_out[i]; i++) {
        p.slot[i].in = g_in[i];
        p.slot[i].out = g_out[i];
        p.slot[i].state = SLOT_FREE;
        nslot++;
    }
    int rc = -1;
    if (nslot < NSLOT) {
        /* one slot: the file fits it; nothing to overlap */
        struct slot *s = &p.slot[0];
        uint64_t t = 0;
        phase("", &t);
        s->got = fread(s->in, 1, p.batch_bytes, in);
        phase("read", &t);
        if (s->got != p.batch_bytes && ferror(in)) {
            say(QUIET_ERR, "Couldn't read from input stream");
        } else if (s->got == 0 || s->got % block_len == 0 || s->got == p.batch_bytes) {
            say(QUIET_ERR, "Couldn't encode segment");
        } else if (salz_cuda_multi_encode_stream(g_enc, s->in, s->got, block_len, s->out, g_out_cap, &s->produced, NULL,
                                                 NULL) != 0) {
            say(QUIET_ERR, "Couldn't encode segment");
        } else if (phase("encode (H2D, kernels, D2H)", &t), fwrite(s->out, 1, s->produced, out) != s->produced) {
            say(QUIET_ERR, "Couldn't write encoded segment to output stream");
        } else {
            phase("write", &t);
            rc = 0;
        }
        goto done;
    }
    pthread_t rd, wr;
    pthread_create(&rd, NULL, reader_main, &p);
    pthread_create(&wr, NULL, writer_main, &p);
    for (int i = 0;; i = (i + 1) % NSLOT) {
        struct slot *s = &p.slot[i];
        pthread_mutex_lock(&p.mu);
        while (s->state != SLOT_FILLED && !p.failed)
            pthread_cond_wait(&p.cv, &p.mu);
        const bool stop = p.failed;
        pthread_mutex_unlock(&p.mu);
        if (stop)
            break;
        if (salz_cuda_multi_encode_stream(g_enc, s->in, s->got, block_len, s->out, g_out_cap, &s->produced, NULL,
                                          NULL) != 0) {
            pipe_fail(&p, "Couldn't encode segment");
            break;
        }
        const bool last = s->last;
        pthread_mutex_lock(&p.mu);
        s->state = SLOT_ENCODED;
        pthread_cond_broadcast(&p.cv);
        pthread_mutex_unlock(&p.mu);
        if (last)
            break;
    }
    pthread_join(rd, NULL);
    pthread_join(wr, NULL);
    if (p.failed)
        say(QUIET_ERR, "%s", p.why);
    else
        rc = 0;
done:
    pthread_mutex_destroy(&p.mu);
    pthread_cond_destroy(&p.cv);
    return rc;
}

static int do_decompress(FILE *in, FILE *out)
{
    uint8_t hdr[8];
    uint32_t magic, block_len;
    if (fread(hdr, 1, 8, in) != 8) {
        say(QUIET_ERR, "Couldn't read SALZ header from input");
        return -1;
    }
    memcpy(&magic, hdr, 4);
    memcpy(&block_len, hdr + 4, 4);
    if (magic != SALZ_MAGIC) {
        say(QUIET_ERR, "Not a SALZ header, unexpected magic number");
        return -1;
    }
    const size_t in_cap = (size_t)salz_encoded_len_max(block_len);
    /* 8 spare bytes: unlike the reference decoder ours never writes past the
     * end of a factor, the slack is only defensive */
    uint8_t *inbuf = malloc(in_cap), *outbuf = malloc((size_t)block_len + 8);
    int rc = -1;
    if (inbuf == NULL || outbuf == NULL) {
        say(QUIET_ERR, "Couldn't allocate memory (%zu bytes)", in_cap + block_len);
        goto done;
    }
    for (;;) {
        uint32_t rec_len;
        size_t got = fread(&rec_len, 1, 4, in);
        if (got != 4) {
            if (ferror(in)) {
                say(QUIET_ERR, "Couldn't read encoded segments length from input stream");
                goto done;
            }
            if (got == 0 && feof(in))
                break;
            say(QUIET_ERR, "Couldn't read encoded segments length from input stream");
            goto done;
        }
        if (rec_len > in_cap) {
            say(QUIET_ERR, "Encoded segment too large to fit into input buffer");
            goto done;
        }
        if (fread(inbuf, 1, rec_len, in) != rec_len) {
            say(QUIET_ERR, "Couldn't read encoded segment from input stream");
            goto done;
        }
        size_t plain = block_len;
        if (salz_decode_safe(inbuf, rec_len, outbuf, &plain) != 0) {
            say(QUIET_ERR, "Couldn't decode segment");
            goto done;
        }
        if (fwrite(outbuf, 1, plain, out) != plain) {
            say(QUIET_ERR, "Couldn't write decoded segment to output stream");
            goto done;
        }
    }
    rc = 0;
done:
    free(inbuf);
    free(outbuf);
    return rc;
}

static int handle_file(const char *path)
{
    char outpath[PATH_MAX];
    struct stat st;
    const bool suffixed = strstr(path, SUFFIX) != NULL;

    if (mode == MODE_COMPRESS && suffixed) {
        say(QUIET_ERR, "\"%s\" path already has \".salz\" suffix", path);
        return -1;
    }
    if (mode == MODE_DECOMPRESS && !suffixed) {
        say(QUIET_ERR, "\"%s\" path has unknown suffix", path);
        return -1;
    }
    if (stat(path, &st) != 0) {
        say(QUIET_ERR, "Couldn't stat \"%s\" path (err: %d)", path, errno);
        return -1;
    }
    if (!S_ISREG(st.st_mode)) {
        say(QUIET_ERR, "\"%s\" path is not a regular file", path);
        return -1;
    }
    const long in_size = (long)st.st_size;
    FILE *in = fopen(path, "r");
    if (in == NULL) {
        say(QUIET_ERR, "Couldn't open \"%s\" path (err: %d)", path, errno);
        return -1;
    }
    derive_outpath(path, outpath);
    if (!force_overwrite && stat(outpath, &st) == 0) {
        say(QUIET_ERR, "\"%s\" path already exists", outpath);
        fclose(in);
        return -1;
    }
    FILE *out = fopen(outpath, "w");
    if (out == NULL) {
        say(QUIET_ERR, "Couldn't open \"%s\" path (err: %d)", outpath, errno);
        fclose(in);
        return -1;
    }

    uint64_t t0 = 0, t1 = 0;
    get_time_ns(&t0);
    int rc = (mode == MODE_COMPRESS) ? do_compress(in, out, (size_t)in_size) : do_decompress(in, out);
    get_time_ns(&t1);
    fclose(out);
    fclose(in);

    if (rc != 0) {
        say(QUIET_ERR, "Operation failed");
        unlink(outpath);
        return -1;
    }
    if (!keep_source)
        unlink(path);
    if (stat(outpath, &st) != 0) {
        say(QUIET_ERR, "Couldn't stat \"%s\" path (err: %d)", outpath, errno);
        return -1;
    }
    const double secs = (double)(t1 - t0) / NS_IN_SEC;
    if (mode == MODE_COMPRESS)
        say(VERBOSE_INFO, "%s: compressed %ld bytes to %ld bytes (ratio: %.3f) in %.3f seconds", path, in_size,
            (long)st.st_size, 1.0 * in_size / (double)st.st_size, secs);
    else
        say(VERBOSE_INFO, "%s: decompressed %ld bytes in %.3f seconds", path, in_size, secs);
    return 0;
}

static void usage(const char *name)
{
    printf("salz, a Suffix Array-based Lempel-Ziv data compressor\n\n"
           "  usage: %s [options] input_file ...\n\n"
           "  -c --stdout        write to standard output, keep input file\n"
           "  -d --decompress    force decompression mode\n"
           "  -f --force         force overwrite of output file\n"
           "  -h --help          print this message\n"
           "  -k --keep          keep input file\n"
           "  -l --list          print information about salz-compressed file\n"
           "  -q --quiet         suppress output\n"
           "                     (specify twice to all but non-critical errors)\n"
           "  -0 ... -9          compression level [default: 5]\n"
           "                     (note that memory usage grows exponentially)\n"
           "  --fast             alias of \"-1\"\n"
           "  --best             alias of \"-9\"\n"
           "  --gpus N           compress on the first N GPUs (0 = all) [default: 1]\n\n"
           "  Default action is to compress.\n"
           "  If invoked as \"unsalz\", default action is to decompress.\n"
           "                \"salzcat\", default action is to decompress to stdout.\n\n"
           "  If no input file is given, or - is provided instead, salz compresses\n"
           "  or decompresses from standard input to standard output.\n",
           name);
}

int main(int argc, char *argv[])
{
    static const struct option longopts[] = {
        { "stdout", no_argument, NULL, 'c' },     { "decompress", no_argument, NULL, 'd' },
        { "force", no_argument, NULL, 'f' },      { "help", no_argument, NULL, 'h' },
        { "keep", no_argument, NULL, 'k' },       { "list", no_argument, NULL, 'l' },
        { "quiet", no_argument, NULL, 'q' },      { "fast", no_argument, NULL, '1' },
        { "best", no_argument, NULL, '9' },       { "gpus", required_argument, NULL, 'G' },
        { NULL, 0, NULL, 0 },
    };
    const char *self = base_name(argv[0]);
    int opt;
    while ((opt = getopt_long(argc, argv, "cdfhklq0123456789", longopts, NULL)) != -1) {
        if (opt >= '0' && opt <= '9') {
            level = opt - '0';
            continue;
        }
        switch (opt) {
        case 'c':
            fprintf(stderr, "writing to stdout not supported\n");
            return -1;
        case 'G':
            gpus_opt = atoi(optarg);
            break;
        case 'd':
            mode = MODE_DECOMPRESS;
            break;
        case 'f':
            force_overwrite = true;
            break;
        case 'h':
            usage(self);
            return 0;
        case 'k':
            keep_source = true;
            break;
        case 'l':
            fprintf(stderr, "listing info not supported\n");
            return -1;
        case 'q':
            if (verbosity > 0)
                verbosity--;
            break;
        default:
            fprintf(stderr, "See \"%s --help\" for more information.\n", self);
            return -1;
        }
    }
    if (strncmp(self, "unsalz", 6) == 0)
        mode = MODE_DECOMPRESS;
    if (strncmp(self, "salzcat", 7) == 0) {
        fprintf(stderr, "writing to stdout not supported\n");
        return -1;
    }
    if (optind >= argc || argv[optind][0] == '-') {
        fprintf(stderr, "compressing from stdin not supported\n");
        return -1;
    }
    int worst = 0;
    for (int i = optind; i < argc; i++) {
        int rc = handle_file(argv[i]);
        worst = min(worst, rc);
    }
    /* Every output file is closed.  Leave without tearing the encoder down: unmapping the
     * device workspace and the pinned buffers takes longer than compressing a small file,
     * and the driver reclaims them with the process. */
    fflush(NULL);
    if (g_enc != NULL)
        _exit(worst == 0 ? 0 : 255);
    return worst;
}
