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
// The reference starts one worker per CPU (rnghunt.rs:16-18, num_cpus::get()); this binary starts one
// host thread with its own bbb_ctx per visible B200 (--gpus G to limit, --gpu D for one given device).
// Device g takes the candidates c = g (mod G) of every round, so the accepted index reported -- the
// smallest one -- does not depend on how many devices took part.
//
// New, optional surface:  --n N  --seed S  --first C  --batch B  --k K  --all COUNT  --gpus G  --gpu D
//                         --devices LIST  --rank index|weight  --stats
//   --all COUNT    sweep exactly COUNT candidates and list the K best accepted ones (best-k) instead of
//                  stopping at the first hit; `out` then holds the best one.
//   --rank weight  (with --all) order the accepted candidates by the number of terms of their
//                  characteristic polynomial (the figure util/verify.py:19-28 looks at; 103 for the shipped
//                  n = 192 matrix), ties by index; default order is by candidate index.
#include <algorithm>
#include <cerrno>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <random>
#include <string>
#include <thread>
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

struct Hit {
    uint64_t cand;
    uint32_t weight;
};

struct Worker {
    bbb_ctx *ctx = nullptr;
    int rc = 0;
    uint64_t counts[BBB_NSTAGES] = {0, 0, 0, 0, 0};
    std::vector<Hit> hits;
};

// candidates first + g, first + g + G, ... (count of them) on worker w; up to kmax hits, with weights if asked
static void sweep_shard(Worker &w, uint32_t n, uint64_t seed, uint64_t first, uint64_t stride, uint64_t count,
                        uint32_t kmax, bool want_weight)
{
    w.hits.clear();
    if (!count) return;
    std::vector<uint64_t> best(kmax ? kmax : 1);
    std::vector<uint32_t> wt(kmax ? kmax : 1, 0);
    uint32_t nf = 0;
    if (want_weight)
        w.rc = bbb_rnghunt_sweep_polys(w.ctx, n, seed, first, stride, count, kmax, best.data(), &nf, w.counts, nullptr,
                                       wt.data());
    else
        w.rc = bbb_rnghunt_sweep(w.ctx, n, seed, first, stride, count, kmax, best.data(), &nf, w.counts, nullptr);
    for (uint32_t i = 0; i < nf; i++) w.hits.push_back(Hit{best[i], wt[i]});
}

int main(int argc, char **argv)
{
    uint32_t n = 192, k = 16;                       // rnghunt.rs:14
    int gpu = -1, gpus = 0;
    bool stats = false, by_weight = false;
    uint64_t seed = 0, first = 0, batch = 1u << 16, all = 0;
    bool have_seed = false;
    std::vector<int> dev_list;
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
        else if (!strcmp(argv[i], "--gpus")) gpus = (int)val("--gpus");
        else if (!strcmp(argv[i], "--devices")) {          // explicit ordinals, repeats allowed: "0,1" or "0,0,0"
            if (i + 1 >= argc) { fprintf(stderr, "rnghunt: --devices needs a list\n"); return 2; }
            for (char *tok = strtok(argv[++i], ","); tok; tok = strtok(nullptr, ",")) dev_list.push_back(atoi(tok));
        }
        else if (!strcmp(argv[i], "--stats")) stats = true;
        else if (!strcmp(argv[i], "--rank")) {
            if (i + 1 >= argc) { fprintf(stderr, "rnghunt: --rank needs index|weight\n"); return 2; }
            const char *v = argv[++i];
            if (!strcmp(v, "weight")) by_weight = true;
            else if (strcmp(v, "index")) { fprintf(stderr, "rnghunt: --rank index|weight\n"); return 2; }
        } else { fprintf(stderr, "rnghunt: unknown argument %s\n", argv[i]); return 2; }
    }
    if (!have_seed) {
        std::random_device rd;
        seed = ((uint64_t)rd() << 32) | rd();
    }
    if (!batch) batch = 1;
    printf("Starting search for n=%u\n", n);
    fflush(stdout);
    // one worker per device, the analogue of ThreadPool::new(num_cpus::get()) (rnghunt.rs:16-18)
    const int visible = bbb_device_count();
    if (visible <= 0) { fprintf(stderr, "rnghunt: no usable B200 (there is no CPU fallback)\n"); return 2; }
    std::vector<int> devs;
    if (!dev_list.empty()) devs = dev_list;
    else if (gpu >= 0) devs.push_back(gpu);
    else for (int d = 0; d < visible && (gpus <= 0 || d < gpus); d++) devs.push_back(d);
    const uint64_t G = devs.size();
    std::vector<Worker> ws(G);
    for (uint64_t g = 0; g < G; g++) {
        int rc = bbb_create(devs[g], &ws[g].ctx);
        if (rc) { fprintf(stderr, "rnghunt: no usable B200 at device %d (bbb_create: %d); there is no CPU fallback\n", devs[g], rc); return 2; }
    }
    auto run_round = [&](uint64_t lo, uint64_t count, uint32_t kmax, bool want_weight) {
        // candidates [lo, lo + count), device g taking lo + g, lo + g + G, ...
        std::vector<std::thread> th;
        for (uint64_t g = 0; g < G; g++) {
            const uint64_t mine = count > g ? (count - g + G - 1) / G : 0;
            th.emplace_back(sweep_shard, std::ref(ws[g]), n, seed, lo + g, G, mine, kmax, want_weight);
        }
        for (auto &t : th) t.join();
        for (uint64_t g = 0; g < G; g++) if (ws[g].rc) die(ws[g].ctx, "bbb_rnghunt_sweep", ws[g].rc);
    };
    std::vector<Hit> found;
    uint64_t done = 0;
    if (all) {
        // every hit is needed to rank by weight; by index the k smallest of each device suffice
        const uint32_t kmax = by_weight ? (uint32_t)std::min<uint64_t>(all, 1u << 20) : k;
        run_round(first, all, kmax, by_weight);
        for (auto &w : ws) found.insert(found.end(), w.hits.begin(), w.hits.end());
        done = all;
    } else {
        // first hit wins (rnghunt.rs:54-55): round after round of G * batch consecutive candidates; the smallest
        // accepted index of the first round that has one is the first hit of the whole stream
        while (found.empty()) {
            run_round(first + done, G * batch, 1, false);
            for (auto &w : ws) found.insert(found.end(), w.hits.begin(), w.hits.end());
            done += G * batch;
            if (batch < (1u << 20)) batch *= 2;
        }
    }
    std::sort(found.begin(), found.end(), [&](const Hit &a, const Hit &b) {
        if (by_weight && a.weight != b.weight) return a.weight < b.weight;
        return a.cand < b.cand;
    });
    if (!all && found.size() > 1) found.resize(1);
    if (found.size() > k && all) found.resize(k);
    if (!found.empty()) {
        printf("Found suitable recurrence matrix for n=%u, writing out\n", n);
        fflush(stdout);
        write_out(ws[0].ctx, n, seed, found[0].cand);
    }
    if (stats || all) {
        uint64_t counts[BBB_NSTAGES] = {0, 0, 0, 0, 0};
        for (auto &w : ws) for (int i = 0; i < BBB_NSTAGES; i++) counts[i] += w.counts[i];
        fprintf(stderr, "seed=%llu candidates=%llu degree=%llu filter=%llu order=%llu factor=%llu primitive=%llu gpus=%llu\n",
                (unsigned long long)seed, (unsigned long long)done, (unsigned long long)counts[0],
                (unsigned long long)counts[1], (unsigned long long)counts[2], (unsigned long long)counts[3],
                (unsigned long long)counts[4], (unsigned long long)G);
        for (size_t i = 0; i < found.size(); i++) {
            if (by_weight) fprintf(stderr, "best[%zu]=%llu weight=%u\n", i, (unsigned long long)found[i].cand, found[i].weight);
            else fprintf(stderr, "best[%zu]=%llu\n", i, (unsigned long long)found[i].cand);
        }
    }
    for (auto &w : ws) bbb_destroy(w.ctx);
    return 0;
}
