// rnghunt -- drop-in for software/rnghunt/src/bin/rnghunt.rs on the B200 library.
//
// With no arguments it behaves like the reference binary: prints
//   "Starting search for n=192"                                   (rnghunt.rs:15)
// searches random sparse 192 x 192 GF(2) matrices until one has a primitive characteristic
// polynomial of degree n, prints
//   "Found suitable recurrence matrix for n=192, writing out"     (rnghunt.rs:45)
// writes ./out (n lines of n '0'/'1' characters, line r = row r; rnghunt.rs:46-53) and exits.
// The candidates come from the library's reproducible candidate stream; like the reference
// (OS-seeded thread_rng, binary_matrix.rs:84) the default seed differs from run to run.
//
// New, optional surface:  --n N  --seed S  --first C  --batch B  --k K  --all COUNT  --gpu D  --stats
//   --all COUNT   sweep exactly COUNT candidates and list the K smallest accepted indices (best-k)
//                 instead of stopping at the first hit; `out` then holds the smallest one.
#include <cerrno>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <random>
#include <string>
#include <vector>

#include "../include/bbb_b200.h"

static void die(bbb_ctx *c, const char *what, int rc)
{
    fprintf(stderr, "rnghunt: %s failed (%d): %s\n", what, rc, c ? bbb_last_error(c) : "");
    exit(2);
}

static void write_out(bbb_ctx *c, uint32_t n, uint64_t seed, uint64_t cand)
{
    const uint32_t wpc = (n + 63) / 64;
    std::vector<uint64_t> m((size_t)n * wpc);
    int rc = bbb_rnghunt_candidate(c, n, seed, cand, m.data());
    if (rc) die(c, "bbb_rnghunt_candidate", rc);
    FILE *f = fopen("out", "w");
    if (!f) {
        // the reference panics with this text (rnghunt.rs:48)
        fprintf(stderr, "thread 'main' panicked at 'Couldn't create out: %s'\n", strerror(errno));
        exit(101);
    }
    std::string line(n + 1, '\n');
    for (uint32_t r = 0; r < n; r++) {
        for (uint32_t col = 0; col < n; col++)
            line[col] = ((m[(size_t)col * wpc + r / 64] >> (63 - r % 64)) & 1) ? '1' : '0';   // a.row(r)
        fwrite(line.data(), 1, n + 1, f);
    }
    fclose(f);
}

int main(int argc, char **argv)
{
    uint32_t n = 192, k = 16;                       // rnghunt.rs:14
    int gpu = 0;
    bool stats = false;
    uint64_t seed = 0, first = 0, batch = 1u << 16, all = 0;
    bool have_seed = false;
    for (int i = 1; i < argc; i++) {
        auto val = [&](const char *name) -> uint64_t {
            if (i + 1 >= argc) { fprintf(stderr, "rnghunt: %s needs a value\n", name); exit(2); }
            return strtoull(argv[++i], nullptr, 0);
        };
        if (!strcmp(argv[i], "--n")) n = (uint32_t)val("--n");
        else if (!strcmp(argv[i], "--seed")) { seed = val("--seed"); have_seed = true; }
        else if (!strcmp(argv[i], "--first")) first = val("--first");
        else if (!strcmp(argv[i], "--batch")) batch = val("--batch");
        else if (!strcmp(argv[i], "--k")) k = (uint32_t)val("--k");
        else if (!strcmp(argv[i], "--all")) all = val("--all");
        else if (!strcmp(argv[i], "--gpu")) gpu = (int)val("--gpu");
        else if (!strcmp(argv[i], "--stats")) stats = true;
        else { fprintf(stderr, "rnghunt: unknown argument %s\n", argv[i]); return 2; }
    }
    if (!have_seed) {
        std::random_device rd;
        seed = ((uint64_t)rd() << 32) | rd();
    }
    if (!batch) batch = 1;
    printf("Starting search for n=%u\n", n);
    fflush(stdout);
    bbb_ctx *c = nullptr;
    int rc = bbb_create(gpu, &c);
    if (rc) { fprintf(stderr, "rnghunt: no usable B200 (bbb_create: %d); there is no CPU fallback\n", rc); return 2; }
    uint64_t counts[BBB_NSTAGES] = {0, 0, 0, 0, 0};
    std::vector<uint64_t> best(k ? k : 1);
    uint32_t nfound = 0;
    uint64_t done = 0;
    if (all) {
        rc = bbb_rnghunt_sweep(c, n, seed, first, 1, all, k, best.data(), &nfound, counts, nullptr);
        if (rc) die(c, "bbb_rnghunt_sweep", rc);
        done = all;
    } else {
        // first hit wins (rnghunt.rs:54-55): sweep batch after batch, keep the smallest index
        while (!nfound) {
            rc = bbb_rnghunt_sweep(c, n, seed, first + done, 1, batch, 1, best.data(), &nfound, counts, nullptr);
            if (rc) die(c, "bbb_rnghunt_sweep", rc);
            done += batch;
            if (batch < (1u << 20)) batch *= 2;
        }
    }
    if (nfound) {
        printf("Found suitable recurrence matrix for n=%u, writing out\n", n);
        fflush(stdout);
        write_out(c, n, seed, best[0]);
    }
    if (stats || all) {
        fprintf(stderr, "seed=%llu candidates=%llu degree=%llu filter=%llu order=%llu factor=%llu primitive=%llu\n",
                (unsigned long long)seed, (unsigned long long)done, (unsigned long long)counts[0],
                (unsigned long long)counts[1], (unsigned long long)counts[2], (unsigned long long)counts[3],
                (unsigned long long)counts[4]);
        for (uint32_t i = 0; i < nfound; i++) fprintf(stderr, "best[%u]=%llu\n", i, (unsigned long long)best[i]);
    }
    bbb_destroy(c);
    return 0;
}
