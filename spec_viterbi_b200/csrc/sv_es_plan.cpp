// K0 for the edge-stationary kernel: turns the CSR of T' into per-thread register slices.
//
//  1. in-degree capping: a light row may use at most shape[0] register slots, so a state
//     with more in-edges is split into virtual copies that share its emission column; every
//     out-edge is replicated from each copy.  Exact: min_c fl(u + v[c]) = fl(u + min_c v[c]).
//     Rows that are too dense to split (the N and C states of profile HMMs feed / are fed by
//     every match state) become "heavy" rows: their edges are dealt round-robin to all
//     threads and reduced with shuffles.
//  2. placement: light rows sorted by in-degree are cut into chunks of 32 and dealt to
//     (rank, warp) cells in snake order, so that a warp's trip count per rank is uniform
//     and the per-warp totals balance.  Internal state index p = rank*threads + thread,
//     hence the bank of v[p] is the lane of its owner.
//  3. bank scheduling: inside each (warp, rank) cell the edges of every lane are ordered so
//     that the lanes of one LDS phase (32 lanes x 4 B or 16 lanes x 8 B) read distinct banks
//     wherever possible; lanes inside a chunk are permuted by a local search that minimises
//     the resulting shared-memory wavefronts.
#include "sv_internal.h"

#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <limits>
#include <numeric>
#include <random>

namespace {

struct Edge {
    uint32_t other; // in-list: source internal state; out-list: destination internal state
    float w;
};

struct Graph {
    std::vector<std::vector<Edge>> in, out;
    std::vector<uint32_t> orig;
    std::vector<uint8_t> heavy;
};

constexpr float kInf = std::numeric_limits<float>::infinity();

// Split internal state x (in-degree > cap) into ceil(d/cap) copies.
bool split_state(Graph& g, uint32_t x, size_t cap, size_t max_states,
                 std::vector<uint32_t>& worklist) {
    const size_t d = g.in[x].size();
    const size_t copies = (d + cap - 1) / cap;
    if (copies < 2) return true;
    if (g.in.size() + copies - 1 > max_states) return false;
    const size_t per = (d + copies - 1) / copies;
    std::vector<Edge> all;
    all.swap(g.in[x]);
    std::vector<uint32_t> ids(copies);
    ids[0] = x;
    for (size_t c = 1; c < copies; ++c) {
        ids[c] = static_cast<uint32_t>(g.in.size());
        g.in.emplace_back();
        g.out.emplace_back();
        g.orig.push_back(g.orig[x]);
        g.heavy.push_back(0);
    }
    // out-edges of x are replicated from every new copy (before moving in-edges, so that a
    // self-loop x->x is seen as an ordinary out-edge whose in-list entry is then dealt out)
    for (size_t k = 0; k < all.size(); ++k) {
        const uint32_t to = ids[k / per];
        const Edge e = all[k];
        g.in[to].push_back(e);
        if (to != x) { // retarget the source's out-entry x -> to (first matching entry)
            for (auto& oe : g.out[e.other]) {
                if (oe.other == x && oe.w == e.w) {
                    oe.other = to;
                    break;
                }
            }
        }
    }
    // g.out[x] may have been retargeted above if x has a self loop; re-read it
    const std::vector<Edge> outs_now = g.out[x];
    for (size_t c = 1; c < copies; ++c) {
        for (const Edge& oe : outs_now) {
            g.in[oe.other].push_back(Edge{ids[c], oe.w});
            g.out[ids[c]].push_back(oe);
            if (!g.heavy[oe.other] && g.in[oe.other].size() > cap) worklist.push_back(oe.other);
        }
    }
    return true;
}

// Number of shared-memory wavefronts of one LDS phase group: lanes read 4*q-byte words at
// byte offsets off[]; words are spread over 32/q banks; equal words broadcast.
int group_wavefronts(const uint32_t* off, int lanes, int q) {
    const int units = 32 / q;
    uint32_t words[32];
    int nwords = 0;
    int load[32] = {0};
    int worst = 1;
    for (int a = 0; a < lanes; ++a) {
        const uint32_t w = off[a] / (4u * q);
        bool seen = false;
        for (int b = 0; b < nwords; ++b) seen |= words[b] == w;
        if (seen) continue;
        words[nwords++] = w;
        worst = std::max(worst, ++load[w % units]);
    }
    return worst;
}

bool try_variant(const SvHostCsr& csr, const std::vector<float>& emis,
                 const std::vector<float>& start, uint64_t emits, size_t smem_limit, int vindex,
                 SvEsPlan* plan) {
    const SvEsVariant& var = sv_es_variant(vindex);
    const int TPG = var.threads, Q = var.q, NW = TPG / 32;
    const size_t n = csr.n;
    const size_t cap = var.shape[0];
    const size_t max_states = static_cast<size_t>(SV_ES_RMAX) * TPG + var.nh;
    if (n > max_states) return false;

    // ---- internal graph ----
    Graph g;
    g.in.resize(n);
    g.out.resize(n);
    g.orig.resize(n);
    g.heavy.assign(n, 0);
    for (uint32_t i = 0; i < n; ++i) {
        g.orig[i] = i;
        for (uint32_t p = csr.rowptr[i]; p < csr.rowptr[i + 1]; ++p) {
            g.in[i].push_back(Edge{csr.col[p], csr.w[p]});
            g.out[csr.col[p]].push_back(Edge{i, csr.w[p]});
        }
    }
    // heavy rows: the densest rows that would be too costly to split
    std::vector<uint32_t> heavy_rows;
    if (var.nh > 0) {
        std::vector<uint32_t> byd(n);
        std::iota(byd.begin(), byd.end(), 0u);
        std::stable_sort(byd.begin(), byd.end(),
                         [&](uint32_t a, uint32_t b) { return g.in[a].size() > g.in[b].size(); });
        for (int h = 0; h < var.nh && h < static_cast<int>(n); ++h) {
            const uint32_t i = byd[h];
            if (g.in[i].size() > std::max<size_t>(cap, 32)) {
                g.heavy[i] = 1;
                heavy_rows.push_back(i);
            }
        }
    }
    {
        std::vector<uint32_t> work;
        for (uint32_t i = 0; i < n; ++i)
            if (!g.heavy[i] && g.in[i].size() > cap) work.push_back(i);
        size_t guard = 0;
        while (!work.empty()) {
            const uint32_t x = work.back();
            work.pop_back();
            if (g.heavy[x] || g.in[x].size() <= cap) continue;
            if (!split_state(g, x, cap, max_states, work)) return false;
            if (++guard > 4 * max_states) return false;
        }
    }
    const size_t n_int = g.in.size();
    const size_t n_heavy = heavy_rows.size();
    const size_t n_light = n_int - n_heavy;
    if (n_light > static_cast<size_t>(SV_ES_RMAX) * TPG) return false;
    for (const uint32_t h : heavy_rows)
        if (g.in[h].size() > static_cast<size_t>(var.hs) * TPG) return false;

    // ---- shared memory budget ----
    const size_t vb = (n_int * Q * 4 + 15) & ~size_t(15);
    const size_t smem = 2 * vb + 2 * static_cast<size_t>(var.nh) * NW * Q * 4 + emits * n_int * 4; // red is double buffered
    if (smem + 512 > smem_limit) return false; // headroom for the kernel's static shared variables

    // ---- placement: sorted light rows -> chunks of 32 -> (rank, warp) cells, snake order ----
    std::vector<uint32_t> light;
    light.reserve(n_light);
    for (uint32_t i = 0; i < n_int; ++i)
        if (!g.heavy[i]) light.push_back(i);
    std::stable_sort(light.begin(), light.end(),
                     [&](uint32_t a, uint32_t b) { return g.in[a].size() > g.in[b].size(); });
    const int nranks = static_cast<int>((n_light + TPG - 1) / TPG);
    // p(position k in sorted order): cells are filled rank by rank; inside a rank the chunk
    // order alternates direction so that per-warp totals even out
    std::vector<uint32_t> pos_of(n_int, 0); // internal id -> p
    std::vector<int32_t> at_pos(static_cast<size_t>(SV_ES_RMAX) * TPG, -1);
    for (size_t k = 0; k < n_light; ++k) {
        const int r = static_cast<int>(k / TPG);
        const int c = static_cast<int>((k % TPG) / 32);
        const int lane = static_cast<int>(k % 32);
        const int w = (r % 2 == 0) ? c : NW - 1 - c;
        const uint32_t p = static_cast<uint32_t>(r * TPG + w * 32 + lane);
        pos_of[light[k]] = p;
        at_pos[p] = static_cast<int32_t>(light[k]);
    }
    // The last rank may be partially filled: rows must occupy p < n_light for the kernel's
    // store predicate, so compact the last rank's cells to the low positions.
    {
        const int r = nranks - 1;
        if (r >= 0) {
            std::vector<int32_t> rows;
            for (int t = 0; t < TPG; ++t)
                if (at_pos[static_cast<size_t>(r) * TPG + t] >= 0)
                    rows.push_back(at_pos[static_cast<size_t>(r) * TPG + t]);
            // keep degree order: highest degrees first, contiguous from thread 0
            std::stable_sort(rows.begin(), rows.end(), [&](int32_t a, int32_t b) {
                return g.in[a].size() > g.in[b].size();
            });
            for (int t = 0; t < TPG; ++t) at_pos[static_cast<size_t>(r) * TPG + t] = -1;
            for (size_t k = 0; k < rows.size(); ++k) {
                at_pos[static_cast<size_t>(r) * TPG + k] = rows[k];
                pos_of[rows[k]] = static_cast<uint32_t>(r * TPG + k);
            }
        }
    }
    for (size_t h = 0; h < n_heavy; ++h) pos_of[heavy_rows[h]] = static_cast<uint32_t>(n_light + h);

    // shape check: every cell's max in-degree must fit its rank's register slice
    std::vector<uint8_t> deg(static_cast<size_t>(NW) * SV_ES_RMAX, 0);
    for (int r = 0; r < nranks; ++r) {
        for (int w = 0; w < NW; ++w) {
            size_t d = 0;
            for (int l = 0; l < 32; ++l) {
                const int32_t row = at_pos[static_cast<size_t>(r) * TPG + w * 32 + l];
                if (row >= 0) d = std::max(d, g.in[row].size());
            }
            if (d > static_cast<size_t>(var.shape[r])) return false;
            deg[static_cast<size_t>(w) * SV_ES_RMAX + r] = static_cast<uint8_t>(d);
        }
    }

    // ---- lane permutation inside cells (local search on predicted wavefronts) + slot order ----
    int base[SV_ES_RMAX + 1];
    base[0] = 0;
    for (int r = 0; r < SV_ES_RMAX; ++r) base[r + 1] = base[r] + var.shape[r];
    const int NL = base[SV_ES_RMAX];
    const int NSL = NL + var.nh * var.hs;
    plan->edge_w.assign(static_cast<size_t>(NSL) * TPG, kInf);
    plan->edge_off.assign(static_cast<size_t>(NSL) * TPG, 0);

    const int group = 32 / Q; // lanes per LDS phase
    uint64_t wavefronts = 0, ideal = 0;
    auto schedule_cell = [&](int r, int w, bool commit) -> uint64_t {
        // greedy slot assignment per phase group; returns wavefronts of this cell
        const int D = deg[static_cast<size_t>(w) * SV_ES_RMAX + r];
        uint64_t cost = 0;
        for (int g0 = 0; g0 < 32; g0 += group) {
            std::vector<std::vector<Edge>> rem(group);
            for (int l = 0; l < group; ++l) {
                const int32_t row = at_pos[static_cast<size_t>(r) * TPG + w * 32 + g0 + l];
                if (row >= 0) rem[l] = g.in[row];
            }
            for (int k = 0; k < D; ++k) {
                const int slots_left = D - k;
                std::vector<uint32_t> off(group, 0xffffffffu);
                std::vector<float> wt(group, kInf);
                std::vector<int> unit_load(group, 0);
                std::vector<int> order(group);
                std::iota(order.begin(), order.end(), 0);
                std::stable_sort(order.begin(), order.end(), [&](int a, int b) {
                    return rem[a].size() > rem[b].size();
                });
                std::vector<uint32_t> used_words;
                for (const int l : order) {
                    if (rem[l].empty()) continue;
                    const bool must = static_cast<int>(rem[l].size()) >= slots_left;
                    int pick = -1, pick_load = 1 << 30;
                    for (size_t e = 0; e < rem[l].size(); ++e) {
                        const uint32_t word = pos_of[rem[l][e].other];
                        const bool bcast =
                            std::find(used_words.begin(), used_words.end(), word) != used_words.end();
                        const int load = bcast ? 0 : unit_load[word % group];
                        if (load < pick_load) {
                            pick_load = load;
                            pick = static_cast<int>(e);
                        }
                    }
                    if (pick_load > 0 && !must) continue; // wait for a conflict-free slot
                    const Edge e = rem[l][pick];
                    rem[l].erase(rem[l].begin() + pick);
                    const uint32_t word = pos_of[e.other];
                    if (std::find(used_words.begin(), used_words.end(), word) == used_words.end()) {
                        used_words.push_back(word);
                        unit_load[word % group] += 1;
                    }
                    off[l] = word * Q * 4u;
                    wt[l] = e.w;
                }
                // padding lanes broadcast an address that is already being read
                uint32_t filler = 0xffffffffu;
                for (int l = 0; l < group; ++l)
                    if (off[l] != 0xffffffffu) filler = off[l];
                for (int l = 0; l < group; ++l) {
                    if (off[l] == 0xffffffffu) {
                        const uint32_t own = static_cast<uint32_t>(r * TPG + w * 32 + g0 + l);
                        off[l] = filler != 0xffffffffu
                                     ? filler
                                     : std::min<uint32_t>(own, static_cast<uint32_t>(n_int - 1)) * Q * 4u;
                    }
                }
                cost += group_wavefronts(off.data(), group, Q);
                if (commit) {
                    for (int l = 0; l < group; ++l) {
                        const size_t t = static_cast<size_t>(w) * 32 + g0 + l;
                        plan->edge_w[static_cast<size_t>(base[r] + k) * TPG + t] = wt[l];
                        plan->edge_off[static_cast<size_t>(base[r] + k) * TPG + t] = off[l];
                    }
                }
            }
        }
        return cost;
    };

    for (int r = 0; r < nranks; ++r) {
        for (int w = 0; w < NW; ++w) {
            wavefronts += schedule_cell(r, w, true);
            ideal += static_cast<uint64_t>(deg[static_cast<size_t>(w) * SV_ES_RMAX + r]) * Q;
        }
    }

    // ---- heavy rows: edges sorted by source position, dealt round-robin to all threads ----
    plan->heavy_slots.assign(SV_ES_NHMAX, 0);
    for (size_t h = 0; h < n_heavy; ++h) {
        std::vector<Edge> es = g.in[heavy_rows[h]];
        std::stable_sort(es.begin(), es.end(),
                         [&](const Edge& a, const Edge& b) { return pos_of[a.other] < pos_of[b.other]; });
        const int hs = static_cast<int>((es.size() + TPG - 1) / TPG);
        plan->heavy_slots[h] = static_cast<uint8_t>(hs);
        for (int k = 0; k < hs; ++k) {
            for (int t = 0; t < TPG; ++t) {
                const size_t m = static_cast<size_t>(k) * TPG + t;
                const size_t slot = static_cast<size_t>(NL + h * var.hs + k) * TPG + t;
                if (m < es.size()) {
                    plan->edge_w[slot] = es[m].w;
                    plan->edge_off[slot] = pos_of[es[m].other] * Q * 4u;
                } else {
                    plan->edge_w[slot] = kInf;
                    plan->edge_off[slot] = std::min<uint32_t>(t, static_cast<uint32_t>(n_int - 1)) * Q * 4u;
                }
            }
            for (int w = 0; w < NW; ++w) {
                for (int g0 = 0; g0 < 32; g0 += group) {
                    const size_t slot = static_cast<size_t>(NL + h * var.hs + k) * TPG + w * 32 + g0;
                    wavefronts += group_wavefronts(&plan->edge_off[slot], group, Q);
                }
                ideal += Q;
            }
        }
    }

    // ---- tables in internal order ----
    plan->emis_p.resize(emits * n_int);
    plan->start_p.resize(n_int);
    plan->orig.resize(n_int);
    std::vector<uint32_t> at(n_int, 0); // p -> internal id
    for (uint32_t i = 0; i < n_int; ++i) at[pos_of[i]] = i;
    for (uint32_t p = 0; p < n_int; ++p) {
        const uint32_t o = g.orig[at[p]];
        plan->orig[p] = o;
        plan->start_p[p] = start[o];
        for (uint64_t e = 0; e < emits; ++e) plan->emis_p[e * n_int + p] = emis[e * n + o];
    }
    plan->copy_ptr.assign(n + 1, 0);
    for (uint32_t p = 0; p < n_int; ++p) plan->copy_ptr[plan->orig[p] + 1] += 1;
    for (size_t i = 0; i < n; ++i) plan->copy_ptr[i + 1] += plan->copy_ptr[i];
    plan->copy_p.resize(n_int);
    {
        std::vector<uint32_t> fill(plan->copy_ptr.begin(), plan->copy_ptr.end() - 1);
        for (uint32_t p = 0; p < n_int; ++p) plan->copy_p[fill[plan->orig[p]]++] = p;
    }

    plan->valid = true;
    plan->var = var;
    plan->variant_index = vindex;
    plan->n_int = static_cast<int>(n_int);
    plan->n_light = static_cast<int>(n_light);
    plan->n_heavy = static_cast<int>(n_heavy);
    plan->nslots = NSL;
    plan->ranks = nranks;
    plan->gather_wavefronts = wavefronts;
    plan->gather_ideal = ideal;
    plan->smem_bytes = smem;
    plan->split_states = n_int - n;
    plan->deg = deg;
    return true;
}

// Per-step cost proxy used to rank candidate plans: the busiest warp's slot count
// (light slots + heavy slots), i.e. the critical path between two barriers.
uint64_t plan_cost(const SvEsPlan& p) {
    const int NW = p.var.threads / 32;
    uint64_t worst = 0;
    for (int w = 0; w < NW; ++w) {
        uint64_t s = 0;
        for (int r = 0; r < SV_ES_RMAX; ++r) s += p.deg[static_cast<size_t>(w) * SV_ES_RMAX + r];
        worst = std::max(worst, s);
    }
    for (int h = 0; h < SV_ES_NHMAX; ++h) worst += p.heavy_slots[h];
    return worst;
}

} // namespace

void sv_es_build_plan(const SvHostCsr& csr, const std::vector<float>& emis,
                      const std::vector<float>& start, uint64_t emits, size_t smem_limit,
                      SvEsPlan* plan, int force_q) {
    plan->valid = false;
    const size_t n = csr.n;
    // thread count: enough threads that a thread owns ~2-5 rows
    int want_threads = 128;
    if (n > 3 * 128) want_threads = 256;
    if (n > 3 * 256) want_threads = 512;
    if (const char* env = std::getenv("SV_ES_THREADS")) want_threads = std::atoi(env);
    int want_q = force_q;
    if (const char* env = std::getenv("SV_ES_Q")) want_q = std::atoi(env);

    SvEsPlan bestp;
    uint64_t best_cost = ~0ull;
    for (int pass = 0; pass < 2 && !bestp.valid; ++pass) {
        for (int v = 0; v < sv_es_num_variants(); ++v) {
            const SvEsVariant& var = sv_es_variant(v);
            if (pass == 0 && var.threads != want_threads) continue;
            if (pass == 1 && var.threads <= want_threads) continue; // fall back to more threads
            if (want_q != 0 && var.q != want_q) continue;
            SvEsPlan cand;
            if (!try_variant(csr, emis, start, emits, smem_limit, v, &cand)) continue;
            // prefer two sequences per CTA (amortises barriers and instruction issue),
            // then the cheaper schedule
            const uint64_t cost = plan_cost(cand) * 4 + (cand.var.q == 2 ? 0 : 2);
            if (cost < best_cost) {
                best_cost = cost;
                bestp = std::move(cand);
            }
        }
    }
    *plan = std::move(bestp);
}
