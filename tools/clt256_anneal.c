// Simulated annealing of the load order for the one-pass clt256 kernel; same cost model as
// tools/clt256_sched.py:live_profile (kept in step with it; the Python version re-checks the result).
// usage: clt256_anneal <iters> <seed> <lead> <resident_budget> < rows.txt > pos.txt
//   rows.txt: 256 lines "k t0 t1 .. t(k-1)"   (taps of each row, reference rng_recurrences.py:172-259)
//   pos.txt : 256 integers, pos[bit] = load time
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#define N 256
static int ntap[N], tap[N][4];
static int lead_g = 0;
static double w_elig = 0.0;   // reward per plane that can stay in a register across the loop edge (short wraps first)
static int done_t[N];

static uint64_t rng_s = 88172645463325252ull;
static inline uint64_t rnd64(void)
{
    rng_s ^= rng_s << 13;
    rng_s ^= rng_s >> 7;
    rng_s ^= rng_s << 17;
    return rng_s;
}
static inline double rndu(void) { return (rnd64() >> 11) * (1.0 / 9007199254740992.0); }

static void sort4(int *a, int n)
{
    for (int i = 1; i < n; i++) {
        int v = a[i], j = i - 1;
        while (j >= 0 && a[j] > v) {
            a[j + 1] = a[j];
            j--;
        }
        a[j + 1] = v;
    }
}

// cost = peak + mean/64 + penalty for a wide top (p-norm-like term keeps a gradient)
static double cost(const int *pos, int *peak_out, double *mean_out)
{
    int diff[N + 2];
    int death[N];
    memset(diff, 0, sizeof diff);
    for (int p = 0; p < N; p++) death[p] = p;
    for (int r = 0; r < N; r++) {
        int ts[4], k = ntap[r];
        for (int i = 0; i < k; i++) ts[i] = pos[tap[r][i]];
        sort4(ts, k);
        int fold = k >= 3 ? ts[2] : ts[k - 1];
        for (int i = 0; i < k && i < 3; i++)
            if (death[ts[i]] < fold) death[ts[i]] = fold;
        int done = ts[k - 1];
        done_t[r] = done;
        if (k == 4 && ts[3] > ts[2]) {
            diff[ts[2]]++;
            diff[ts[3]]--;
        }
        int fr = pos[r] - lead_g;
        if (fr > done) {
            diff[done]++;
            diff[fr]--;
        }
    }
    for (int p = 0; p < N; p++)
        if (death[p] > p) {
            diff[p]++;
            diff[death[p]]--;
        }
    int lv = 0, peak = 0;
    long sum = 0;
    double p4 = 0;
    for (int t = 0; t < N; t++) {
        lv += diff[t];
        if (lv > peak) peak = lv;
        sum += lv;
        double x = lv / 64.0;
        p4 += x * x * x * x;
    }
    if (peak_out) *peak_out = peak;
    if (mean_out) *mean_out = sum / (double)N;
    double elig = 0;
    if (w_elig > 0)
        for (int r = 0; r < N; r++)
            if (done_t[r] > death[pos[r]]) elig += 1.0 - (double)(N - done_t[r] + pos[r]) / (2.0 * N);
    return peak + p4 / 64.0 - w_elig * elig;
}

int main(int argc, char **argv)
{
    long iters = argc > 1 ? atol(argv[1]) : 1000000;
    if (argc > 2) rng_s ^= (uint64_t)atol(argv[2]) * 0x9E3779B97F4A7C15ull;
    if (argc > 3) lead_g = atoi(argv[3]);
    if (argc > 4) w_elig = atof(argv[4]);
    for (int r = 0; r < N; r++) {
        if (scanf("%d", &ntap[r]) != 1) return 1;
        for (int i = 0; i < ntap[r]; i++)
            if (scanf("%d", &tap[r][i]) != 1) return 1;
    }
    int pos[N], best_pos[N];
    for (int i = 0; i < N; i++) pos[i] = i;
    // random start
    for (int i = N - 1; i > 0; i--) {
        int j = rnd64() % (i + 1), t = pos[i];
        pos[i] = pos[j];
        pos[j] = t;
    }
    double cur = cost(pos, 0, 0), best = cur;
    memcpy(best_pos, pos, sizeof pos);
    const double T0 = 3.0, T1 = 0.01;
    for (long it = 0; it < iters; it++) {
        double T = T0 * pow(T1 / T0, (double)it / iters);
        int a = rnd64() % N, b = rnd64() % N;
        if (a == b) continue;
        int t = pos[a];
        pos[a] = pos[b];
        pos[b] = t;
        double c = cost(pos, 0, 0), d = c - cur;
        if (d <= 0 || rndu() < exp(-d / T)) {
            cur = c;
            if (cur < best) {
                best = cur;
                memcpy(best_pos, pos, sizeof pos);
            }
        } else {
            t = pos[a];
            pos[a] = pos[b];
            pos[b] = t;
        }
        if ((it & 0xfffff) == 0) {
            int pk;
            double mn;
            cost(pos, &pk, &mn);
            fprintf(stderr, "it %ld cur %.3f best %.3f peak %d mean %.1f\n", it, cur, best, pk, mn);
        }
    }
    int pk;
    double mn;
    cost(best_pos, &pk, &mn);
    fprintf(stderr, "final peak %d mean %.2f\n", pk, mn);
    for (int i = 0; i < N; i++) printf("%d\n", best_pos[i]);
    return 0;
}
