// TEST INFRASTRUCTURE ONLY — a concurrency stress model of the search loop of sph_b200/csrc/geodesic.cu (geo_search_kernel).
// tests/emu runs the unmodified kernel, but with the threads of a block serialised, so it cannot expose a lost flag or a
// premature end of the pass loop.  This file restates ONLY the synchronisation protocol of that loop — flag bitmap claimed by
// atomic exchange, distances relaxed by atomic min, flag set after the min, deferred / dropped / expanded vertices, the
// changed / min-pending / bound words, two barriers per pass — with 8 OS threads standing in for the 8 warps of a CTA, and
// checks every search bit for bit against a sequential Dijkstra on random graphs.  x86 is more strongly ordered than the GPU
// (the kernel adds __threadfence where this model relies on seq_cst atomics), so this checks the LOGIC of the protocol under
// real interleavings, not the fences.
//   g++ -O2 -std=c++17 -pthread tests/emu/geo_race_model.cpp -o /tmp/geo_race && /tmp/geo_race [rounds]
#include <algorithm>
#include <atomic>
#include <condition_variable>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <queue>
#include <random>
#include <thread>
#include <vector>

namespace {
constexpr int WARPS = 8;
constexpr uint32_t INF = 0x7f800000u;

uint32_t f2u(float f) { uint32_t u; std::memcpy(&u, &f, 4); return u; }
float u2f(uint32_t u) { float f; std::memcpy(&f, &u, 4); return f; }
void atomic_min(std::atomic<uint32_t>& a, uint32_t v, uint32_t* old_out = nullptr) {
    uint32_t old = a.load();
    while (v < old && !a.compare_exchange_weak(old, v)) {}
    if (old_out) *old_out = old;
}
void atomic_max(std::atomic<uint32_t>& a, uint32_t v) {
    uint32_t old = a.load();
    while (v > old && !a.compare_exchange_weak(old, v)) {}
}

struct Barrier {
    std::mutex m;
    std::condition_variable cv;
    int count = 0, gen = 0;
    void wait() {
        std::unique_lock<std::mutex> l(m);
        const int g = gen;
        if (++count == WARPS) {
            count = 0;
            ++gen;
            cv.notify_all();
        } else {
            cv.wait(l, [&] { return gen != g; });
        }
    }
};

struct Graph {
    int n, k;
    std::vector<int> ids;
    std::vector<float> w;
};

std::vector<float> dijkstra(const Graph& G, int s) {
    std::vector<float> d(G.n, u2f(INF));
    using QE = std::pair<float, int>;
    std::priority_queue<QE, std::vector<QE>, std::greater<QE>> q;
    d[s] = 0;
    q.emplace(0.f, s);
    std::vector<char> done(G.n, 0);
    while (!q.empty()) {
        auto [du, u] = q.top();
        q.pop();
        if (done[u]) continue;
        done[u] = 1;
        for (int j = 0; j < G.k; ++j) {
            const int v = G.ids[u * G.k + j];
            if (v == u) continue;
            const float nd = du + G.w[u * G.k + j];
            if (nd < d[v]) {
                d[v] = nd;
                q.emplace(nd, v);
            }
        }
    }
    return d;
}

struct Search {
    const Graph& G;
    int s;
    std::vector<int> targets;
    float delta;
    std::vector<std::atomic<uint32_t>> dist, bm;
    std::atomic<uint32_t> changed{0}, minpend{INF}, bound{0};
    Barrier bar;
    Search(const Graph& g, int src, std::vector<int> t, float dl) : G(g), s(src), targets(std::move(t)), delta(dl), dist(g.n), bm((g.n + 31) / 32) {
        for (auto& d : dist) d.store(INF);
        for (auto& b : bm) b.store(0);
        dist[s].store(0);
        bm[s >> 5].store(1u << (s & 31));
    }
    void warp(int wid) {  // the pass loop of geo_search_kernel, one "warp"
        const int nwords = static_cast<int>(bm.size());
        float thr = delta;
        for (;;) {
            bar.wait();
            if (wid == 0) {
                changed.store(0);
                minpend.store(INF);
                bound.store(0);
            }
            bar.wait();
            uint32_t bmax = 0;
            for (size_t i = wid; i < targets.size(); i += WARPS) bmax = std::max(bmax, dist[targets[i]].load());
            if (bmax) atomic_max(bound, bmax);
            bar.wait();
            const uint32_t bound_bits = bound.load();
            for (int base = wid * 32; base < nwords; base += WARPS * 32) {
                uint32_t snap[32];
                for (int l = 0; l < 32; ++l) snap[l] = base + l < nwords ? bm[base + l].load() : 0u;  // the ballot read
                for (int l = 0; l < 32; ++l) {
                    if (!snap[l]) continue;
                    const int w = base + l;
                    uint32_t bits = bm[w].exchange(0);
                    uint32_t keep = 0;
                    while (bits) {
                        const int bit = __builtin_ctz(bits);
                        bits &= bits - 1;
                        const int u = w * 32 + bit;
                        const uint32_t du_bits = dist[u].load();
                        const float du = u2f(du_bits);
                        if (du_bits >= bound_bits) continue;
                        if (!(du <= thr)) {
                            keep |= 1u << bit;
                            atomic_min(minpend, du_bits);
                            continue;
                        }
                        for (int j = 0; j < G.k; ++j) {
                            const int v = G.ids[u * G.k + j];
                            if (v == u) continue;
                            const float nd = du + G.w[u * G.k + j];
                            const uint32_t nd_bits = f2u(nd);
                            if (nd_bits >= bound_bits || nd_bits >= dist[v].load()) continue;  // the two filters in front of the atomic
                            uint32_t old;
                            atomic_min(dist[v], nd_bits, &old);
                            if (nd_bits < old) {
                                bm[v >> 5].fetch_or(1u << (v & 31));
                                if (nd <= thr) changed.store(1);
                                else atomic_min(minpend, nd_bits);
                            }
                        }
                    }
                    if (keep) bm[w].fetch_or(keep);
                }
            }
            bar.wait();
            if (changed.load() == 0) {
                if (minpend.load() == INF) break;
                thr = u2f(minpend.load()) + delta;
            }
        }
    }
};
}  // namespace

int main(int argc, char** argv) {
    const int rounds = argc > 1 ? std::atoi(argv[1]) : 40;
    std::mt19937 rng(7);
    long checked = 0;
    for (int r = 0; r < rounds; ++r) {
        Graph G;
        G.n = 300 + static_cast<int>(rng() % 3000);
        G.k = 3 + static_cast<int>(rng() % 10);
        G.ids.resize(static_cast<size_t>(G.n) * G.k);
        G.w.resize(G.ids.size());
        std::uniform_real_distribution<float> uw(0.f, 1.f);
        const bool local = r & 1;  // half of the graphs: neighbours near in index (long paths), half: random (short paths)
        for (int u = 0; u < G.n; ++u)
            for (int j = 0; j < G.k; ++j) {
                int v = local ? (u + 1 + static_cast<int>(rng() % 12)) % G.n : static_cast<int>(rng() % G.n);
                G.ids[u * G.k + j] = j == 0 ? u : v;
                G.w[u * G.k + j] = j == 0 ? 0.f : ((r % 5 == 0 && rng() % 4 == 0) ? 0.f : uw(rng));  // some exact zeros
            }
        double sum = 0;
        for (float x : G.w) sum += x;
        const float mean = static_cast<float>(sum / G.w.size());
        for (int q = 0; q < 12; ++q) {
            const int s = static_cast<int>(rng() % G.n);
            std::vector<int> targets;
            for (int t = 0; t < 1 + static_cast<int>(rng() % 4); ++t) targets.push_back(static_cast<int>(rng() % G.n));
            const float mul[4] = {0.1f, 1.f, 4.f, 1e-6f};
            Search S(G, s, targets, mean * mul[q & 3]);
            std::vector<std::thread> th;
            for (int w = 0; w < WARPS; ++w) th.emplace_back([&S, w] { S.warp(w); });
            for (auto& t : th) t.join();
            const auto ref = dijkstra(G, s);
            for (int t : targets) {
                const uint32_t got = S.dist[t].load();
                if (got != f2u(ref[t])) {
                    std::printf("MISMATCH round %d query %d source %d target %d: %g vs %g\n", r, q, s, t, u2f(got), ref[t]);
                    return 1;
                }
                ++checked;
            }
            for (auto& b : S.bm)
                if (b.load() != 0) {
                    std::printf("flag bitmap not empty at the end (round %d query %d)\n", r, q);
                    return 1;
                }
        }
    }
    std::printf("geo_race_model: %ld target distances bit-identical to Dijkstra, bitmaps empty at the end\n", checked);
    return 0;
}
