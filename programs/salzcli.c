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
 * Environment: SALZ_CUDA_DEVICE (default 0), SALZ_BATCH_MIB (default 64).
 */
#include <errno.h>
#include <getopt.h>
#include <limits.h>
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

static int do_compress(FILE *in, FILE *out)
{
    const uint32_t block_len = 1u << (15 + level);
    const uint32_t magic = SALZ_MAGIC;
    const char *env = getenv("SALZ_BATCH_MIB");
    size_t batch_bytes = (size_t)(env ? atoi(env) : 64) << 20;
    if (batch_bytes < block_len)
        batch_bytes = block_len;
    const size_t batch_blocks = batch_bytes / block_len;
    batch_bytes = batch_blocks * block_len;
    const size_t out_cap = batch_blocks * ((size_t)salz_encoded_len_max(block_len) + 4);
    env = getenv("SALZ_CUDA_DEVICE");

    int rc = -1;
    uint8_t *inbuf = NULL, *outbuf = NULL;
    salz_cuda_ctx *ctx = salz_cuda_ctx_create(env ? atoi(env) : 0, block_len, batch_blocks);
    if (ctx == NULL) {
        say(QUIET_ERR, "Couldn't create CUDA encoder context (no GPU?)");
        return -1;
    }
    inbuf = salz_cuda_host_alloc(batch_bytes);
    outbuf = salz_cuda_host_alloc(out_cap);
    if (inbuf == NULL || outbuf == NULL) {
        say(QUIET_ERR, "Couldn't allocate memory: (%zu bytes)", batch_bytes + out_cap);
        goto done;
    }
    if (fwrite(&magic, 1, 4, out) != 4 || fwrite(&block_len, 1, 4, out) != 4) {
        say(QUIET_ERR, "Couldn't write SALZ header to output");
        goto done;
    }
    for (;;) {
        size_t got = fread(inbuf, 1, batch_bytes, in);
        if (got != batch_bytes && ferror(in)) {
            say(QUIET_ERR, "Couldn't read from input stream");
            goto done;
        }
        /* The reference encodes one more (short or empty) block after every
         * full-size read, so a file whose size is a multiple of the block size,
         * or whose tail is <= 8 bytes, fails there; keep that behaviour. */
        const bool last = got != batch_bytes;
        if (got == 0 || (last && got % block_len == 0)) {
            say(QUIET_ERR, "Couldn't encode segment");
            goto done;
        }
        size_t produced = 0;
        if (salz_cuda_encode_stream(ctx, inbuf, got, block_len, outbuf, out_cap, &produced, NULL, NULL) != 0) {
            say(QUIET_ERR, "Couldn't encode segment");
            goto done;
        }
        if (fwrite(outbuf, 1, produced, out) != produced) {
            say(QUIET_ERR, "Couldn't write encoded segment to output stream");
            goto done;
        }
        if (last)
            break;
    }
    rc = 0;
done:
    salz_cuda_host_free(inbuf);
    salz_cuda_host_free(outbuf);
    salz_cuda_ctx_destroy(ctx);
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
    int rc = (mode == MODE_COMPRESS) ? do_compress(in, out) : do_decompress(in, out);
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
           "  --best             alias of \"-9\"\n\n"
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
        { "best", no_argument, NULL, '9' },       { NULL, 0, NULL, 0 },
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
    return worst;
}
