// test_group.cpp -- drives a placement group of several GPUs through the C-ABI from C++ (no
// Python, no torch): k8sgpu_create with ndev devices, k8sgpu_place / k8sgpu_place_gang, and checks
// that every byte equals the one-GPU result.  This is what a cgo caller does.
//
//   test_group [ndev]      exit 0 = equal (or fewer than 2 GPUs: prints SKIP), 1 = mismatch / error
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "../../include/k8sgpu.h"

static uint64_t mix64(uint64_t z) {
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}
static uint32_t rnd(uint64_t stream, uint64_t i, uint32_t mod) {
    return (uint32_t)(mix64(stream * 0x9E3779B97F4A7C15ull + (i + 1) * 0xD1B54A32D192ED03ull) % mod);
}

struct Tables {
    uint32_t n, m;
    std::vector<uint32_t> nc[8], pc[8], gang, mins;
    k8sgpu_node_table nodes() const {
        return {n, nc[0].data(), nc[1].data(), nc[2].data(), nc[3].data(), nc[4].data(), nc[5].data(), nc[6].data(), nc[7].data()};
    }
    k8sgpu_pod_table pods() const {
        return {m, pc[0].data(), pc[1].data(), pc[2].data(), pc[3].data(), pc[4].data(), pc[5].data(), pc[6].data(), nullptr};
    }
};

static Tables make(uint32_t n, uint32_t m, uint64_t seed) {
    Tables t;
    t.n = n;
    t.m = m;
    for (auto &c : t.nc) c.resize(n);
    for (auto &c : t.pc) c.resize(m);
    for (uint32_t i = 0; i < n; ++i) {
        const bool big = rnd(seed + 1, i, 100) < 30;
        const uint32_t cap = big ? 8 : 1, fg = cap - rnd(seed + 2, i, cap + 1);
        t.nc[0][i] = fg;
        t.nc[1][i] = fg * (big ? 12000u : 4000u);
        t.nc[2][i] = fg * (big ? 243200u : 28672u);
        t.nc[3][i] = cap;
        t.nc[4][i] = big ? 1 : 0;
        t.nc[5][i] = rnd(seed + 3, i, 3);
        t.nc[6][i] = rnd(seed + 4, i, 100) < 5 ? 1 : 0;
        t.nc[7][i] = rnd(seed + 5, i, 100) < 99 ? 1 : 0;
    }
    for (uint32_t p = 0; p < m; ++p) {
        const uint32_t r = rnd(seed + 8, p, 100), g = r < 70 ? 1 : r < 85 ? 2 : r < 95 ? 4 : 8;
        t.pc[0][p] = g;
        t.pc[1][p] = g * 4000u;
        t.pc[2][p] = g * 28672u;
        t.pc[3][p] = rnd(seed + 9, p, 100) < 20 ? 2u : 0xFFFFFFFFu;
        t.pc[4][p] = rnd(seed + 10, p, 100) < 10 ? (1u << rnd(seed + 11, p, 3)) : 0xFFFFFFFFu;
        t.pc[5][p] = rnd(seed + 12, p, 100) < 50 ? 1 : 0;
        t.pc[6][p] = rnd(seed + 13, p, 100) < 50 ? rnd(seed + 14, p, 3) : 0xFFFFFFFFu;
    }
    const uint32_t ng = m / 6 + 1;
    t.gang.resize(m);
    t.mins.resize(ng);
    for (uint32_t p = 0; p < m; ++p) t.gang[p] = rnd(seed + 20, p, 100) < 30 ? 0xFFFFFFFFu : rnd(seed + 21, p, ng);
    for (uint32_t g = 0; g < ng; ++g) t.mins[g] = rnd(seed + 22, g, 7);
    return t;
}

struct Result {
    std::vector<int32_t> assign, gang_ok;
    std::vector<uint32_t> fg, fc, fm;
    k8sgpu_stats st;
    bool operator==(const Result &o) const {
        return assign == o.assign && gang_ok == o.gang_ok && fg == o.fg && fc == o.fc && fm == o.fm &&
               st.rounds == o.st.rounds && st.placed == o.st.placed && st.evals == o.st.evals;
    }
};

static int run(k8sgpu_ctx *ctx, const Tables &t, bool gangs, uint32_t flags, Result *r) {
    r->assign.assign(t.m, -99);
    r->fg.assign(t.n, 0);
    r->fc.assign(t.n, 0);
    r->fm.assign(t.n, 0);
    r->gang_ok.assign(gangs ? t.mins.size() : 0, -1);
    k8sgpu_node_table nt = t.nodes();
    k8sgpu_pod_table pt = t.pods();
    k8sgpu_policy pol = {flags, 0};
    k8sgpu_place_out out = {r->assign.data(), r->fg.data(), r->fc.data(), r->fm.data()};
    int rc;
    if (gangs) {
        k8sgpu_gang_table gt = {(uint32_t)t.mins.size(), t.gang.data(), t.mins.data()};
        rc = k8sgpu_place_gang(ctx, &nt, &pt, &gt, &pol, &out, r->gang_ok.data());
    } else {
        rc = k8sgpu_place(ctx, &nt, &pt, &pol, &out);
    }
    if (rc) fprintf(stderr, "k8sgpu error %d: %s\n", rc, k8sgpu_last_error(ctx));
    k8sgpu_get_stats(ctx, &r->st);
    return rc;
}

int main(int argc, char **argv) {
    int ndev = argc > 1 ? atoi(argv[1]) : 2;
    int devs[16];
    for (int i = 0; i < 16; ++i) devs[i] = i;
    k8sgpu_ctx *one = nullptr, *grp = nullptr;
    char err[512];
    if (k8sgpu_create_ex(devs, 1, &one, err, sizeof err)) {
        fprintf(stderr, "create(1): %s\n", err);
        return 1;
    }
    if (int rc = k8sgpu_create_ex(devs, ndev, &grp, err, sizeof err)) {
        if (rc == K8SGPU_E_ARG && strstr(err, "out of range")) {
            printf("SKIP: fewer than %d GPUs (%s)\n", ndev, err);
            k8sgpu_destroy(one);
            return 0;
        }
        fprintf(stderr, "create(%d): %s\n", ndev, err);
        return 1;
    }
    int world = 0, nlocal = 0, ver = 0;
    k8sgpu_group_info(grp, &world, nullptr, &nlocal, &ver);
    printf("group: world=%d nlocal=%d nccl=%d\n", world, nlocal, ver);
    int bad = 0, cases = 0;
    const struct {
        uint32_t n, m;
    } sizes[] = {{300, 2500}, {1, 7}, {1031, 20000}, {4000, 1}, {257, 0}, {2000, 60000}};
    for (auto sz : sizes) {
        const Tables t = make(sz.n, sz.m, 1000 + sz.n);
        for (int gangs = 0; gangs < 2; ++gangs)
            for (int variant = 0; variant < 5; ++variant) {
                // variant: 0 all-gather + one-launch admission, 1 peer stores, 2 all-gather + kernel chain,
                // 3 rot = 0 (all-gather, every GPU copies the whole pod table from the host), 4 defaults
                // (peer stores where peers map, pod-table upload shared out between the GPUs)
                k8sgpu_set_tuning(grp, "exchange", variant == 1 ? 1 : variant == 4 ? -1 : 0);
                k8sgpu_set_tuning(grp, "shard_upload", variant == 3 ? 0 : -1);
                k8sgpu_set_tuning(grp, "coop", variant == 2 ? 0 : -1);
                const uint32_t flags = variant == 3 ? K8SGPU_POLICY_ROT_ZERO : 0;
                Result a, b;
                if (run(one, t, gangs, flags, &a) || run(grp, t, gangs, flags, &b)) return 1;
                ++cases;
                if (!(a == b)) {
                    ++bad;
                    printf("MISMATCH n=%u m=%u gangs=%d variant=%d (rounds %u vs %u, placed %u vs %u)\n", sz.n, sz.m, gangs,
                           variant, a.st.rounds, b.st.rounds, a.st.placed, b.st.placed);
                }
            }
    }
    // same group context, tables of a different size, twice in a row (exchange buffers grow; reuse)
    printf("%s: %d cases on %d GPUs, %d mismatches\n", bad ? "FAIL" : "OK", cases, ndev, bad);
    k8sgpu_destroy(grp);
    k8sgpu_destroy(one);
    return bad ? 1 : 0;
}
