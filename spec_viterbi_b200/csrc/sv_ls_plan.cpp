// K0 for the lockstep (LS) kernel: CSR of T' -> per-thread register slices whose shared-memory
// gathers are bank-conflict free.
//
// The LS kernel keeps the score vectors of 4*nq4 sequences interleaved as float4 [state][4], so
// every relaxation slot is one LDS.128 per float4 group.  The hardware serves an LDS.128 in
// four phases of 8 consecutive lanes; a phase is a single wavefront iff its 8 addresses fall
// in 8 distinct 16-byte bank groups, i.e. iff the 8 source positions are distinct mod 8 (or
// equal: broadcast).  Position p = rank*threads + thread is both the owner of a row and the
// address of its score, so "colour" = p mod 8 = lane mod 8.  K0 therefore
//   1. drops +inf transitions (fl(fl(E + inf) + v) = +inf never wins the min), caps in-degrees
//      by splitting states into virtual copies (exact: min_c fl(u + v[c]) = fl(u + min_c v[c])),
//      and makes the densest rows of profile HMMs block-reduced "heavy" rows;
//   2. deals degree-sorted rows to (rank, warp) cells so that trip counts are warp-uniform;
//   3. searches (simulated annealing over row <-> position swaps) for a placement in which, for
//      every octet of lanes and every colour, at most D distinct sources of that colour feed
//      the octet's rows (D = the cell's trip count) -- by Koenig's edge-colouring theorem this
//      is exactly the condition for a conflict-free slot assignment to exist;
//   4. orders every row's edges over the D slots (small local search per octet) and emits the
//      register images.
// Everything here is host code run once per model; it only reorders work, so results stay
// bit-identical to the oracle for any outcome of the search.
#include "sv_internal.h"

#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <limits>
#include <numeric>
#include <random>
#include <thread>

namespace {

constexpr float kInf = std::numeric_limits<float>::infinity();

struct LEdge {
    uint32_t src, dst; // internal state ids
    float w;
};

struct LGraph {
    std::vector<LEdge> edges;                 // live edges (dead ones have dst == ~0u)
    std::vector<std::vector<uint32_t>> in, out; // edge ids per internal state
    std::vector<uint32_t> orig;
    std::vector<uint8_t> heavy;
    uint32_t add_state(uint32_t o) {
        in.emplace_back();
        out.emplace_back();
        orig.push_back(o);
        heavy.push_back(0);
        return static_cast<uint32_t>(orig.size() - 1);
    }
    uint32_t add_edge(uint32_t s, uint32_t d, float w) {
        edges.push_back(LEdge{s, d, w});
        const uint32_t id = static_cast<uint32_t>(edges.size() - 1);
        in[d].push_back(id);
        out[s].push_back(id);
        return id;
    }
};

// Splits state x (in-degree > cap) into copies with <= cap in-edges each; every out-edge of x
// is replicated from each copy.  Returns false if the state budget is exceeded.
bool split_state(LGraph& g, uint32_t x, size_t cap, size_t max_states, std::vector<uint32_t>& work) {
    const size_t d = g.in[x].size();
    const size_t copies = (d + cap - 1) / cap;
    if (copies < 2) return true;
    if (g.orig.size() + copies - 1 > max_states) return false;
    const size_t per = (d + copies - 1) / copies;
    std::vector<uint32_t> ids(copies);
    ids[0] = x;
    for (size_t c = 1; c < copies; ++c) ids[c] = g.add_state(g.orig[x]);
    // move in-edges k >= per to the copies
    std::vector<uint32_t> all;
    all.swap(g.in[x]);
    for (size_t k = 0; k < all.size(); ++k) {
        const uint32_t to = ids[k / per];
        g.edges[all[k]].dst = to;
        g.in[to].push_back(all[k]);
    }
    // replicate out-edges (self loops x->x were just retargeted like any other in-edge and
    // appear here as out-edges of x towards one of the copies)
    const std::vector<uint32_t> outs = g.out[x];
    for (size_t c = 1; c < copies; ++c) {
        for (const uint32_t e : outs) {
            const uint32_t d2 = g.edges[e].dst;
            g.add_edge(ids[c], d2, g.edges[e].w);
            if (!g.heavy[d2] && g.in[d2].size() > cap) work.push_back(d2);
        }
    }
    return true;
}

struct Rng {
    uint64_t s;
    explicit Rng(uint64_t seed) : s(seed * 0x9E3779B97F4A7C15ull + 0x1234567ull) {}
    uint32_t next() {
        s ^= s << 13;
        s ^= s >> 7;
        s ^= s << 17;
        return static_cast<uint32_t>(s >> 32);
    }
    uint32_t below(uint32_t n) { return static_cast<uint32_t>((static_cast<uint64_t>(next()) * n) >> 32); }
    double unit() { return (next() >> 8) * (1.0 / 16777216.0); }
};

} // namespace

namespace {

// ---------------------------------------------------------------- colouring state ----
// Positions p in [0, n_lpad): cell = p / 32 (a warp's rank), octet = p / 8, colour = p % 8.
// cnt[octet][colour] = number of in-edges of the octet's rows whose source has that colour
// (a heavy source counts once per octet: all its readers share one broadcast address).
struct Colouring {
    const LGraph* g = nullptr;
    int n_lpad = 0, n_heavy = 0;
    std::vector<int32_t> at_pos;    // position -> light state id, -1 = hole
    std::vector<uint32_t> pos_of;   // state id -> position (heavy: n_lpad + h)
    std::vector<uint8_t> cellD;     // [n_lpad / 32] trip count of the cell
    std::vector<uint16_t> cnt;      // [octets][8]
    std::vector<uint16_t> hrefs;    // [octets][n_heavy] readers of heavy source h in the octet
    int64_t cost = 0;               // sum over (octet, colour) of max(0, cnt - D)
    std::vector<uint32_t> hot;      // over-full (octet, colour) cells, as octet * 8 + colour
    std::vector<int32_t> hot_at;    // [octets * 8] index into hot, -1 = not over-full

    int octetD(int oct) const { return cellD[oct >> 2]; }
    void bump(int oct, int col, int delta) {
        uint16_t& c = cnt[static_cast<size_t>(oct) * 8 + col];
        const int D = octetD(oct);
        const uint32_t cell = static_cast<uint32_t>(oct) * 8u + static_cast<uint32_t>(col);
        if (delta > 0) {
            if (c >= D) {
                ++cost;
                if (c == D && !hot_at.empty()) { // becomes over-full
                    hot_at[cell] = static_cast<int32_t>(hot.size());
                    hot.push_back(cell);
                }
            }
            ++c;
        } else {
            if (c > D) {
                --cost;
                if (c == D + 1 && !hot_at.empty()) { // no longer over-full: swap-remove
                    const int32_t k = hot_at[cell];
                    hot_at[hot.back()] = k;
                    hot[static_cast<size_t>(k)] = hot.back();
                    hot.pop_back();
                    hot_at[cell] = -1;
                }
            }
            --c;
        }
    }
    void apply_edge(uint32_t e, int sign) {
        const LEdge& ed = g->edges[e];
        if (g->heavy[ed.dst]) return; // heavy rows gather with consecutive addresses
        const int oct = static_cast<int>(pos_of[ed.dst] >> 3);
        const uint32_t ps = pos_of[ed.src];
        const int col = static_cast<int>(ps & 7u);
        if (g->heavy[ed.src]) {
            uint16_t& r = hrefs[static_cast<size_t>(oct) * n_heavy + (ps - n_lpad)];
            if (sign > 0) {
                if (r++ == 0) bump(oct, col, +1);
            } else {
                if (--r == 0) bump(oct, col, -1);
            }
            return;
        }
        bump(oct, col, sign);
    }
};

// Collects the distinct edges touching states a and b (either may be -1 = hole).
void touching_edges(const LGraph& g, int32_t a, int32_t b, std::vector<uint32_t>& out) {
    out.clear();
    for (const int32_t x : {a, b}) {
        if (x < 0) continue;
        out.insert(out.end(), g.in[x].begin(), g.in[x].end());
        out.insert(out.end(), g.out[x].begin(), g.out[x].end());
    }
    std::sort(out.begin(), out.end());
    out.erase(std::unique(out.begin(), out.end()), out.end());
}

void swap_positions(Colouring& c, uint32_t pa, uint32_t pb, std::vector<uint32_t>& scratch) {
    const int32_t a = c.at_pos[pa], b = c.at_pos[pb];
    touching_edges(*c.g, a, b, scratch);
    for (const uint32_t e : scratch) c.apply_edge(e, -1);
    c.at_pos[pa] = b;
    c.at_pos[pb] = a;
    if (a >= 0) c.pos_of[a] = pb;
    if (b >= 0) c.pos_of[b] = pa;
    for (const uint32_t e : scratch) c.apply_edge(e, +1);
}

// Cost change of swapping the rows at positions pa and pb, evaluated WITHOUT touching the
// counters (most proposals are rejected).  Only for graphs without heavy sources.  Fills
// cells/deltas (merged per cell) so that an accepted move can be applied directly.
struct SwapDelta {
    static constexpr int kMax = 96;
    uint32_t cell[kMax];
    int16_t dv[kMax];
    int n = 0;
    bool overflow = false;
    void add(uint32_t c, int d) {
        for (int i = 0; i < n; ++i)
            if (cell[i] == c) {
                dv[i] = static_cast<int16_t>(dv[i] + d);
                return;
            }
        if (n == kMax) {
            overflow = true;
            return;
        }
        cell[n] = c;
        dv[n++] = static_cast<int16_t>(d);
    }
};

int64_t swap_delta(const Colouring& c, uint32_t pa, uint32_t pb, SwapDelta& sd) {
    const LGraph& g = *c.g;
    const int32_t a = c.at_pos[pa], b = c.at_pos[pb];
    sd.n = 0;
    sd.overflow = false;
    auto newpos = [&](uint32_t y) -> uint32_t {
        if (static_cast<int32_t>(y) == a) return pb;
        if (static_cast<int32_t>(y) == b) return pa;
        return c.pos_of[y];
    };
    for (int k = 0; k < 2; ++k) {
        const int32_t x = k == 0 ? a : b;
        if (x < 0) continue;
        const uint32_t px = k == 0 ? pa : pb, nx = k == 0 ? pb : pa;
        // x as destination: all its in-edges (both endpoints at their new positions)
        for (const uint32_t e : g.in[x]) {
            const uint32_t s = g.edges[e].src;
            sd.add((px >> 3) * 8u + (c.pos_of[s] & 7u), -1);
            sd.add((nx >> 3) * 8u + (newpos(s) & 7u), +1);
        }
        // x as source: out-edges to rows that do not move themselves
        for (const uint32_t e : g.out[x]) {
            const uint32_t t = g.edges[e].dst;
            if (static_cast<int32_t>(t) == a || static_cast<int32_t>(t) == b) continue;
            const uint32_t ot = c.pos_of[t] >> 3;
            sd.add(ot * 8u + (px & 7u), -1);
            sd.add(ot * 8u + (nx & 7u), +1);
        }
    }
    int64_t delta = 0;
    for (int i = 0; i < sd.n; ++i) {
        if (sd.dv[i] == 0) continue;
        const int D = c.octetD(static_cast<int>(sd.cell[i] >> 3));
        const int c0 = c.cnt[sd.cell[i]], c1 = c0 + sd.dv[i];
        delta += std::max(0, c1 - D) - std::max(0, c0 - D);
    }
    return delta;
}

void apply_swap_delta(Colouring& c, uint32_t pa, uint32_t pb, const SwapDelta& sd) {
    for (int i = 0; i < sd.n; ++i) {
        const int oct = static_cast<int>(sd.cell[i] >> 3), colr = static_cast<int>(sd.cell[i] & 7u);
        for (int d = sd.dv[i]; d > 0; --d) c.bump(oct, colr, +1);
        for (int d = sd.dv[i]; d < 0; ++d) c.bump(oct, colr, -1);
    }
    const int32_t a = c.at_pos[pa], b = c.at_pos[pb];
    c.at_pos[pa] = b;
    c.at_pos[pb] = a;
    if (a >= 0) c.pos_of[a] = pb;
    if (b >= 0) c.pos_of[b] = pa;
}

// Simulated annealing over position swaps; keeps every row inside a cell whose trip count
// covers its in-degree.  Deterministic (fixed seed).
void anneal(Colouring& c, uint64_t iters, uint64_t seed, const volatile int* cancel) {
    const LGraph& g = *c.g;
    if (c.cost == 0 || iters == 0) return;
    Rng rng(seed);
    std::vector<uint32_t> scratch;
    SwapDelta sd;
    const uint32_t npos = static_cast<uint32_t>(c.n_lpad);
    auto degree = [&](int32_t x) { return x < 0 ? size_t(0) : g.in[x].size(); };
    double t0 = 0.6, t1 = 0.03;
    if (const char* e0 = std::getenv("SV_LS_SA_T0")) t0 = std::atof(e0);
    if (const char* e1 = std::getenv("SV_LS_SA_T1")) t1 = std::atof(e1);
    double temp = t0;
    for (uint64_t it = 0; it < iters && c.cost > 0; ++it) {
        if ((it & 255u) == 0) {
            temp = t0 * std::pow(t1 / t0, static_cast<double>(it) / static_cast<double>(iters));
            if (cancel != nullptr && *cancel != 0) return;
        }
        // pick a row to move: an end of an edge sitting in an over-full (octet, colour), taken
        // from the explicit list of over-full cells (every iteration is a useful attempt)
        uint32_t pa;
        if ((rng.next() & 15u) != 0 && !c.hot.empty()) {
            const uint32_t cell = c.hot[rng.below(static_cast<uint32_t>(c.hot.size()))];
            const uint32_t oct = cell >> 3, colr = cell & 7u;
            uint32_t cand[64];
            uint32_t nc = 0;
            for (uint32_t l = 0; l < 8; ++l) {
                const int32_t row = c.at_pos[oct * 8 + l];
                if (row < 0) continue;
                for (const uint32_t e : g.in[row])
                    if ((c.pos_of[g.edges[e].src] & 7u) == colr && nc < 64) cand[nc++] = e;
            }
            if (nc == 0) continue;
            const LEdge& ed = g.edges[cand[rng.below(nc)]];
            pa = (rng.next() & 3u) ? c.pos_of[ed.src] : c.pos_of[ed.dst];
            if (pa >= npos) continue; // heavy source: not movable
        } else {
            pa = rng.below(npos);
        }
        // partner: same octet (pure recolouring) half of the time, anywhere otherwise
        uint32_t pb;
        if (rng.next() & 1u) {
            pb = (pa & ~7u) | (rng.next() & 7u);
        } else {
            pb = rng.below(npos);
        }
        if (pa == pb) continue;
        const int32_t a = c.at_pos[pa], b = c.at_pos[pb];
        if (a < 0 && b < 0) continue;
        if (degree(a) > c.cellD[pb >> 5] || degree(b) > c.cellD[pa >> 5]) continue;
        if (c.n_heavy == 0) {
            const int64_t delta = swap_delta(c, pa, pb, sd);
            if (!sd.overflow) {
                if (delta <= 0 || rng.unit() < std::exp(-static_cast<double>(delta) / temp))
                    apply_swap_delta(c, pa, pb, sd);
                continue;
            }
        }
        const int64_t before = c.cost;
        swap_positions(c, pa, pb, scratch);
        const int64_t delta = c.cost - before;
        if (delta > 0 && rng.unit() >= std::exp(-static_cast<double>(delta) / temp))
            swap_positions(c, pa, pb, scratch); // reject
    }
}

// End game of the colouring search ("min-conflicts"): the annealing above proposes random partners, and once few
// over-full (octet, colour) cells are left almost every proposal is rejected (8 M iterations leave ~40 of 1336,
// 40 M leave ~11).  Here every move is chosen: take a row at one end of an edge that sits in an over-full cell,
// evaluate the swap with EVERY degree-compatible position, apply the best one (ties broken at random; sideways
// moves allowed, uphill moves never).  Only for graphs without heavy sources (swap_delta's precondition).
void endgame(Colouring& c, uint64_t moves, uint64_t seed, const volatile int* cancel) {
    const LGraph& g = *c.g;
    if (c.cost == 0 || moves == 0 || c.n_heavy != 0) return;
    Rng rng(seed ^ 0xe4d9a3e5ull);
    SwapDelta sd, best_sd;
    const uint32_t npos = static_cast<uint32_t>(c.n_lpad);
    auto degree = [&](int32_t x) { return x < 0 ? size_t(0) : g.in[x].size(); };
    uint64_t stale = 0;
    for (uint64_t it = 0; it < moves && c.cost > 0 && !c.hot.empty(); ++it) {
        if (cancel != nullptr && *cancel != 0) return;
        const uint32_t cell = c.hot[rng.below(static_cast<uint32_t>(c.hot.size()))];
        const uint32_t oct = cell >> 3, colr = cell & 7u;
        uint32_t cand[64];
        uint32_t nc = 0;
        for (uint32_t l = 0; l < 8; ++l) {
            const int32_t row = c.at_pos[oct * 8 + l];
            if (row < 0) continue;
            for (const uint32_t e : g.in[row])
                if ((c.pos_of[g.edges[e].src] & 7u) == colr && nc < 64) cand[nc++] = e;
        }
        if (nc == 0) continue;
        const LEdge& ed = g.edges[cand[rng.below(nc)]];
        const uint32_t pa = (rng.next() & 1u) ? c.pos_of[ed.src] : c.pos_of[ed.dst];
        if (pa >= npos) continue;
        const int32_t a = c.at_pos[pa];
        int64_t best = 1; // accept only delta <= 0
        uint32_t best_pb = pa, nbest = 0;
        for (uint32_t pb = 0; pb < npos; ++pb) {
            if (pb == pa) continue;
            const int32_t b = c.at_pos[pb];
            if (degree(a) > c.cellD[pb >> 5] || degree(b) > c.cellD[pa >> 5]) continue;
            const int64_t delta = swap_delta(c, pa, pb, sd);
            if (sd.overflow || delta > 0 || delta > best) continue;
            if (delta < best) {
                best = delta;
                nbest = 0;
            }
            // reservoir sampling among the equally good partners
            if (rng.below(++nbest) == 0) {
                best_pb = pb;
                best_sd = sd;
            }
        }
        if (best_pb == pa) continue;
        apply_swap_delta(c, pa, best_pb, best_sd);
        stale = best < 0 ? 0 : stale + 1;
        if (stale > 4000) return; // plateau: sideways moves only for a long while
    }
}

// ------------------------------------------------------------- slot assignment ----
// One octet: rows[l] holds the (source position, weight) pairs of lane l.  Assigns every edge
// to one of D slots (at most one edge per lane and slot) minimising the number of wavefronts
// sum_slot max_bankgroup #distinct positions.  Greedy start + local search on swaps of two
// slots inside one lane.
struct OEdge {
    uint32_t pos;
    float w;
};

int slot_wavefronts(const int32_t (*tab)[8], int slot, const std::vector<std::vector<OEdge>>& rows) {
    uint32_t seen[8];
    int ns = 0, load[8] = {0, 0, 0, 0, 0, 0, 0, 0}, worst = 0;
    for (int l = 0; l < 8; ++l) {
        const int32_t k = tab[l][slot];
        if (k < 0) continue;
        const uint32_t p = rows[l][k].pos;
        bool dup = false;
        for (int a = 0; a < ns; ++a) dup |= seen[a] == p;
        if (dup) continue;
        seen[ns++] = p;
        worst = std::max(worst, ++load[p & 7u]);
    }
    return worst; // 0 for an empty slot
}

// Exact start for assign_octet: the edges of an octet form a bipartite multigraph lanes x colours (colour = source
// position mod 8 = bank group); a proper edge colouring with `ncol` = max degree colours exists and is found by
// alternating-path recolouring (Koenig).  A colour class is a slot in which every lane reads at most once and every
// bank group is hit at most once: conflict free.  When the colouring search left an over-full (octet, colour) cell,
// ncol > D and the classes beyond D are folded into the first D slots (least conflict first); the caller's local search
// then only has those few edges to improve.  Returns false (tab untouched) when two lanes read the same position --
// they could share a slot, which the plain edge colouring does not model; the greedy start handles that case.
bool koenig_start(const std::vector<std::vector<OEdge>>& rows, int D, int32_t (*tab)[8]) {
    struct E {
        int l, c, k, col;
    };
    std::vector<E> es;
    int dl[8] = {0}, dc[8] = {0};
    for (int l = 0; l < 8; ++l)
        for (size_t k = 0; k < rows[l].size(); ++k) {
            for (const E& o : es)
                if (rows[o.l][o.k].pos == rows[l][k].pos) return false;
            es.push_back(E{l, static_cast<int>(rows[l][k].pos & 7u), static_cast<int>(k), -1});
            ++dl[l];
            ++dc[rows[l][k].pos & 7u];
        }
    int ncol = D;
    for (int i = 0; i < 8; ++i) ncol = std::max(ncol, std::max(dl[i], dc[i]));
    if (ncol > 16) return false;
    int at_l[8][16], at_c[8][16]; // edge id holding colour x at lane l / colour node c, -1 = free
    for (int i = 0; i < 8; ++i)
        for (int x = 0; x < 16; ++x) at_l[i][x] = at_c[i][x] = -1;
    for (int e = 0; e < static_cast<int>(es.size()); ++e) {
        int a = 0, b = 0;
        while (at_l[es[e].l][a] >= 0) ++a; // free at the lane
        while (at_c[es[e].c][b] >= 0) ++b; // free at the colour node
        if (a != b) {
            // alternating a / b path from the colour node: it cannot reach the lane (which has no a-edge)
            std::vector<int> path;
            bool right = true;
            int node = es[e].c, want = a, other = b;
            for (;;) {
                const int f = right ? at_c[node][want] : at_l[node][want];
                if (f < 0) break;
                path.push_back(f);
                node = right ? es[f].l : es[f].c;
                right = !right;
                std::swap(want, other);
            }
            for (const int f : path) {
                at_l[es[f].l][es[f].col] = -1;
                at_c[es[f].c][es[f].col] = -1;
            }
            for (const int f : path) {
                es[f].col = es[f].col == a ? b : a;
                at_l[es[f].l][es[f].col] = f;
                at_c[es[f].c][es[f].col] = f;
            }
        }
        es[e].col = a;
        at_l[es[e].l][a] = e;
        at_c[es[e].c][a] = e;
    }
    for (int l = 0; l < 8; ++l)
        for (int x = 0; x < 8; ++x) tab[l][x] = -1;
    for (const E& e : es)
        if (e.col < D) tab[e.l][e.col] = e.k;
    for (const E& e : es) { // fold the classes beyond D
        if (e.col < D) continue;
        int best_s = -1, best_c = 1 << 30;
        for (int x = 0; x < D; ++x) {
            if (tab[e.l][x] >= 0) continue;
            const int before = slot_wavefronts(tab, x, rows);
            tab[e.l][x] = e.k;
            const int after = slot_wavefronts(tab, x, rows);
            tab[e.l][x] = -1;
            const int c = (after - std::max(before, 1)) * 16 + before;
            if (c < best_c) {
                best_c = c;
                best_s = x;
            }
        }
        if (best_s < 0) return false; // cannot happen: the lane has at most D edges
        tab[e.l][best_s] = e.k;
    }
    return true;
}

// tab[l][slot] = index into rows[l] or -1.  Returns total wavefronts (empty slots cost 1: the
// padded loads still issue).
int assign_octet(const std::vector<std::vector<OEdge>>& rows, int D, int32_t (*tab)[8], Rng& rng) {
    for (int l = 0; l < 8; ++l)
        for (int s = 0; s < 8; ++s) tab[l][s] = -1;
    static const bool kExact = [] { const char* e = std::getenv("SV_LS_SLOT_EXACT"); return !(e && e[0] == '0'); }();
    bool started = false;
    if (kExact && D >= 1) {
        int32_t t2[8][8];
        if (koenig_start(rows, D, t2)) {
            std::memcpy(tab, t2, sizeof(t2));
            started = true;
        }
    }
    // greedy: lanes with most edges first; each edge to the free slot with the least conflict
    int order[8] = {0, 1, 2, 3, 4, 5, 6, 7};
    std::stable_sort(order, order + 8, [&](int a, int b) { return rows[a].size() > rows[b].size(); });
    for (int oi = 0; oi < 8 && !started; ++oi) {
        const int l = order[oi];
        for (size_t k = 0; k < rows[l].size(); ++k) {
            int best_s = -1, best_c = 1 << 30;
            for (int s = 0; s < D; ++s) {
                if (tab[l][s] >= 0) continue;
                const int before = slot_wavefronts(tab, s, rows);
                tab[l][s] = static_cast<int32_t>(k);
                const int after = slot_wavefronts(tab, s, rows);
                tab[l][s] = -1;
                const int c = (after - std::max(before, 1)) * 16 + before; // prefer no growth, then emptier
                if (c < best_c) {
                    best_c = c;
                    best_s = s;
                }
            }
            tab[l][best_s] = static_cast<int32_t>(k);
        }
    }
    auto total = [&]() {
        int t = 0;
        for (int s = 0; s < D; ++s) t += std::max(1, slot_wavefronts(tab, s, rows));
        return t;
    };
    int cur = total();
    if (cur == D || D < 2) return cur;
    // lower bound: D slots, plus the overflow of the fullest colour (its distinct sources beyond one per slot; overflows of
    // DIFFERENT colours can share a slot's extra wavefronts, so they do not add up)
    int lb = D;
    {
        uint32_t seen[64];
        int ns = 0, cnt[8] = {0, 0, 0, 0, 0, 0, 0, 0};
        for (int l = 0; l < 8; ++l)
            for (const OEdge& e : rows[l]) {
                bool dup = false;
                for (int a = 0; a < ns; ++a) dup |= seen[a] == e.pos;
                if (dup || ns == 64) continue;
                seen[ns++] = e.pos;
                ++cnt[e.pos & 7u];
            }
        int over = 0;
        for (int c = 0; c < 8; ++c) over = std::max(over, cnt[c] - D);
        lb += over;
    }
    if (cur <= lb) return cur;
    static const int kSlotIters = [] { const char* e = std::getenv("SV_LS_SLOT_ITERS"); return e ? std::atoi(e) : 4000; }();
    for (int it = 0; it < kSlotIters && cur > lb; ++it) {
        const int l = static_cast<int>(rng.below(8));
        const int s1 = static_cast<int>(rng.below(D)), s2 = static_cast<int>(rng.below(D));
        if (s1 == s2 || (tab[l][s1] < 0 && tab[l][s2] < 0)) continue;
        const int b = std::max(1, slot_wavefronts(tab, s1, rows)) + std::max(1, slot_wavefronts(tab, s2, rows));
        std::swap(tab[l][s1], tab[l][s2]);
        const int a = std::max(1, slot_wavefronts(tab, s1, rows)) + std::max(1, slot_wavefronts(tab, s2, rows));
        if (a > b || (a == b && (rng.next() & 3u) == 0)) {
            std::swap(tab[l][s1], tab[l][s2]); // reject (sideways moves accepted 3 times in 4)
        } else {
            cur += a - b;
        }
    }
    return cur;
}

} // namespace

uint64_t ls_builtin_full_iters(uint64_t states);

namespace {

bool ls_try_variant(const SvHostCsr& csr, const std::vector<float>& emis,
                    const std::vector<float>& start, uint64_t emits, size_t smem_limit, int vindex,
                    uint64_t sa_iters, const SvLsPlanOpts& opts, SvLsPlan* plan) {
    const SvLsVariant& var = sv_ls_variant(vindex);
    const int TPG = var.threads, NW = TPG / 32;
    const size_t n = csr.n;
    const size_t cap = var.shape[0];
    const int R = var.ranks;
    const size_t max_light = static_cast<size_t>(R) * TPG;
    if (n > max_light + var.nh) return false;
    // tensor memory: one column per (warp group, symbol, rank)
    const size_t tcols_need = static_cast<size_t>(NW / 4) * emits * R;
    if (tcols_need > 512 || emits > 256) return false;
    int tcols = 32;
    while (static_cast<size_t>(tcols) < tcols_need) tcols *= 2;

    // ---- internal graph (without +inf transitions) ----
    LGraph g;
    for (uint32_t i = 0; i < n; ++i) g.add_state(i);
    for (uint32_t i = 0; i < n; ++i)
        for (uint32_t p = csr.rowptr[i]; p < csr.rowptr[i + 1]; ++p)
            if (csr.w[p] != kInf) g.add_edge(csr.col[p], i, csr.w[p]);
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
            if (!split_state(g, x, cap, max_light + heavy_rows.size(), work)) return false;
            if (++guard > 4 * max_light) return false;
        }
    }
    const size_t n_int = g.orig.size();
    const size_t n_heavy = heavy_rows.size();
    const size_t n_light = n_int - n_heavy;
    if (n_light > max_light) return false;
    for (const uint32_t h : heavy_rows)
        if (g.in[h].size() > static_cast<size_t>(var.hs) * TPG) return false;
    const int nranks = std::max<int>(1, static_cast<int>((n_light + TPG - 1) / TPG));
    const int n_lpad = nranks * TPG;
    const int n_pad = static_cast<int>((n_lpad + n_heavy + 7) / 8 * 8);

    // ---- shared memory: 2 buffers x nq4 groups x n_pad float4, heavy reduction scratch ----
    // (static layout: every float4 group is sized for the largest model the variant takes)
    const size_t npad_max = static_cast<size_t>(R) * TPG + 8;
    if (static_cast<size_t>(n_pad) > npad_max) return false;
    auto smem_for = [&](int nq4) {
        return static_cast<size_t>(2) * nq4 * npad_max * 16 + static_cast<size_t>(var.nh) * NW * (4 * nq4) * 4;
    };
    const size_t smem = smem_for(2);
    if (smem + 1024 > smem_limit) return false;

    // ---- placement: degree-sorted light rows -> cells of 32, snake order over warps ----
    Colouring col;
    col.g = &g;
    col.n_lpad = n_lpad;
    col.n_heavy = static_cast<int>(n_heavy);
    col.at_pos.assign(n_lpad, -1);
    col.pos_of.assign(n_int, 0);
    {
        std::vector<uint32_t> light;
        for (uint32_t i = 0; i < n_int; ++i)
            if (!g.heavy[i]) light.push_back(i);
        std::stable_sort(light.begin(), light.end(),
                         [&](uint32_t a, uint32_t b) { return g.in[a].size() > g.in[b].size(); });
        // cells of 32 consecutive rows of the degree-sorted list; the 16 cells of a rank are dealt
        // to the warps so that the per-warp slot totals stay level (heaviest cell -> lightest warp):
        // the step ends when the slowest warp reaches the barrier.  The manager warp (the last
        // one) starts with one slot of handicap for its bookkeeping.
        std::vector<int> load(NW, 0);
        load[NW - 1] = 1;
        if (const char* env = std::getenv("SV_LS_MGR_HANDICAP")) load[NW - 1] = std::atoi(env);
        for (int r = 0; r * static_cast<size_t>(TPG) < light.size(); ++r) {
            std::vector<int> cellD(NW, 0), cells(NW), warps(NW);
            for (int c = 0; c < NW; ++c) {
                const size_t k0 = static_cast<size_t>(r) * TPG + static_cast<size_t>(c) * 32;
                cellD[c] = k0 < light.size() ? static_cast<int>(g.in[light[k0]].size()) : 0;
                cells[c] = c;
                warps[c] = c;
            }
            std::stable_sort(cells.begin(), cells.end(), [&](int a, int b) { return cellD[a] > cellD[b]; });
            std::stable_sort(warps.begin(), warps.end(), [&](int a, int b) { return load[a] < load[b]; });
            for (int i = 0; i < NW; ++i) {
                const int c = cells[i], w = warps[i];
                load[w] += cellD[c];
                for (int l = 0; l < 32; ++l) {
                    const size_t k = static_cast<size_t>(r) * TPG + static_cast<size_t>(c) * 32 + l;
                    if (k >= light.size()) break;
                    const uint32_t p = static_cast<uint32_t>(r * TPG + w * 32 + l);
                    col.pos_of[light[k]] = p;
                    col.at_pos[p] = static_cast<int32_t>(light[k]);
                }
            }
        }
        for (size_t h = 0; h < n_heavy; ++h) col.pos_of[heavy_rows[h]] = static_cast<uint32_t>(n_lpad + h);
    }
    col.cellD.assign(n_lpad / 32, 0);
    std::vector<uint8_t> deg(static_cast<size_t>(NW) * R, 0);
    for (int r = 0; r < nranks; ++r) {
        for (int w = 0; w < NW; ++w) {
            size_t d = 0;
            for (int l = 0; l < 32; ++l) {
                const int32_t row = col.at_pos[static_cast<size_t>(r) * TPG + w * 32 + l];
                if (row >= 0) d = std::max(d, g.in[row].size());
            }
            if (d > static_cast<size_t>(var.shape[r])) return false;
            deg[static_cast<size_t>(w) * R + r] = static_cast<uint8_t>(d);
            col.cellD[(static_cast<size_t>(r) * TPG + w * 32) / 32] = static_cast<uint8_t>(d);
        }
    }

    // ---- colouring search ----
    col.cnt.assign(static_cast<size_t>(n_lpad / 8) * 8, 0);
    col.hrefs.assign(static_cast<size_t>(n_lpad / 8) * std::max<size_t>(n_heavy, 1), 0);
    col.cost = 0;
    col.hot.clear();
    col.hot_at.assign(col.cnt.size(), -1);
    for (uint32_t e = 0; e < g.edges.size(); ++e) col.apply_edge(e, +1);
    plan->colour_cost_before = static_cast<uint64_t>(col.cost);
    anneal(col, sa_iters, 0x5eed0000ull + n * 131u + vindex, opts.cancel);
    if (opts.cancel != nullptr && *opts.cancel != 0) return false;
    if (sa_iters >= ls_builtin_full_iters(n)) { // the full search only: the quick plan of a first batch skips it
        uint64_t moves = 30000;
        if (const char* env = std::getenv("SV_LS_ENDGAME_MOVES")) moves = std::strtoull(env, nullptr, 10);
        endgame(col, moves, 0x5eed0000ull + n * 131u + vindex, opts.cancel);
        if (opts.cancel != nullptr && *opts.cancel != 0) return false;
    }
    plan->colour_cost_after = static_cast<uint64_t>(col.cost);
    plan->sa_iters = sa_iters;

    // ---- slot assignment per octet, register images ----
    int base[SV_LS_RMAX + 1];
    base[0] = 0;
    for (int r = 0; r < R; ++r) base[r + 1] = base[r] + var.shape[r];
    const int NL = base[R];
    const int NSL = NL + var.nh * var.hs;
    plan->edge_w.assign(static_cast<size_t>(NSL) * TPG, kInf);
    plan->edge_off.assign(static_cast<size_t>(NSL) * TPG, 0);
    uint64_t wavefronts = 0, ideal = 0;
    {
        // one task per (rank, warp, octet); tasks are independent (each writes its own 8 threads' slots of
        // its own rank) and seeded by their index, so the result does not depend on the thread count
        const int ntask = nranks * NW * 4;
        std::vector<uint32_t> task_wf(ntask, 0), task_ideal(ntask, 0);
        auto run_tasks = [&](int t0, int t1) {
            for (int task = t0; task < t1; ++task) {
                const int r = task / (NW * 4), w = (task / 4) % NW, o = task % 4;
                const int D = deg[static_cast<size_t>(w) * R + r];
                Rng rng(977u + n * 4099u + static_cast<uint64_t>(task));
                const int p0 = r * TPG + w * 32 + o * 8;
                std::vector<std::vector<OEdge>> rows(8);
                for (int l = 0; l < 8; ++l) {
                    const int32_t row = col.at_pos[p0 + l];
                    if (row < 0) continue;
                    for (const uint32_t e : g.in[row])
                        rows[l].push_back(OEdge{col.pos_of[g.edges[e].src], g.edges[e].w});
                }
                int32_t tab[8][8];
                if (D > 0) task_wf[task] = static_cast<uint32_t>(assign_octet(rows, D, tab, rng));
                task_ideal[task] = static_cast<uint32_t>(D);
                for (int s = 0; s < D; ++s) {
                    // padded lanes re-read an address of this slot (broadcast, no extra wavefront)
                    uint32_t filler = static_cast<uint32_t>(p0);
                    for (int l = 0; l < 8; ++l)
                        if (tab[l][s] >= 0) filler = rows[l][tab[l][s]].pos;
                    for (int l = 0; l < 8; ++l) {
                        const size_t t = static_cast<size_t>(w) * 32 + o * 8 + l;
                        const size_t slot = static_cast<size_t>(base[r] + s) * TPG + t;
                        if (tab[l][s] >= 0) {
                            plan->edge_w[slot] = rows[l][tab[l][s]].w;
                            plan->edge_off[slot] = rows[l][tab[l][s]].pos * 16u;
                        } else {
                            plan->edge_off[slot] = filler * 16u;
                        }
                    }
                }
            }
        };
        const int nthr = std::max(1, std::min(opts.threads, ntask));
        if (nthr == 1) {
            run_tasks(0, ntask);
        } else {
            std::vector<std::thread> pool;
            for (int k = 0; k < nthr; ++k)
                pool.emplace_back(run_tasks, static_cast<int>(static_cast<int64_t>(ntask) * k / nthr),
                                  static_cast<int>(static_cast<int64_t>(ntask) * (k + 1) / nthr));
            for (auto& th : pool) th.join();
        }
        for (int task = 0; task < ntask; ++task) {
            wavefronts += task_wf[task];
            ideal += task_ideal[task];
        }
    }
    // heavy rows: sources sorted by position and dealt round-robin: consecutive lanes read
    // consecutive positions, i.e. distinct bank groups
    plan->heavy_slots.assign(SV_ES_NHMAX, 0);
    for (size_t h = 0; h < n_heavy; ++h) {
        std::vector<OEdge> es;
        for (const uint32_t e : g.in[heavy_rows[h]])
            es.push_back(OEdge{col.pos_of[g.edges[e].src], g.edges[e].w});
        std::stable_sort(es.begin(), es.end(), [](const OEdge& a, const OEdge& b) { return a.pos < b.pos; });
        const int hs = static_cast<int>((es.size() + TPG - 1) / TPG);
        plan->heavy_slots[h] = static_cast<uint8_t>(hs);
        for (int k = 0; k < hs; ++k) {
            for (int t = 0; t < TPG; ++t) {
                const size_t m = static_cast<size_t>(k) * TPG + t;
                const size_t slot = static_cast<size_t>(NL + h * var.hs + k) * TPG + t;
                if (m < es.size()) {
                    plan->edge_w[slot] = es[m].w;
                    plan->edge_off[slot] = es[m].pos * 16u;
                } else {
                    plan->edge_off[slot] = static_cast<uint32_t>(t) * 16u;
                }
            }
            for (int o = 0; o < TPG / 8; ++o) {
                uint32_t seen[8];
                int ns = 0, load[8] = {0}, worst = 1;
                for (int l = 0; l < 8; ++l) {
                    const uint32_t p = plan->edge_off[static_cast<size_t>(NL + h * var.hs + k) * TPG + o * 8 + l] / 16u;
                    bool dup = false;
                    for (int a = 0; a < ns; ++a) dup |= seen[a] == p;
                    if (dup) continue;
                    seen[ns++] = p;
                    worst = std::max(worst, ++load[p & 7u]);
                }
                wavefronts += worst;
                ideal += 1;
            }
        }
    }

    // ---- tables in position order ----
    plan->emis_p.assign(emits * static_cast<size_t>(n_pad), 0.f);
    plan->start_p.assign(n_pad, kInf);
    std::vector<int32_t> state_at(n_pad, -1);
    for (int p = 0; p < n_lpad; ++p) state_at[p] = col.at_pos[p];
    for (size_t h = 0; h < n_heavy; ++h) state_at[n_lpad + h] = static_cast<int32_t>(heavy_rows[h]);
    plan->copy_ptr.assign(n + 1, 0);
    for (int p = 0; p < n_pad; ++p) {
        if (state_at[p] < 0) continue;
        const uint32_t o = g.orig[state_at[p]];
        plan->start_p[p] = start[o];
        for (uint64_t e = 0; e < emits; ++e) plan->emis_p[e * n_pad + p] = emis[e * n + o];
        plan->copy_ptr[o + 1] += 1;
    }
    for (size_t i = 0; i < n; ++i) plan->copy_ptr[i + 1] += plan->copy_ptr[i];
    plan->copy_p.assign(n_int, 0);
    {
        std::vector<uint32_t> fill(plan->copy_ptr.begin(), plan->copy_ptr.end() - 1);
        for (int p = 0; p < n_pad; ++p)
            if (state_at[p] >= 0) plan->copy_p[fill[g.orig[state_at[p]]]++] = static_cast<uint32_t>(p);
    }

    plan->valid = true;
    plan->var = var;
    plan->variant_index = vindex;
    plan->n_int = static_cast<int>(n_int);
    plan->n_lpad = n_lpad;
    plan->n_pad = n_pad;
    plan->n_heavy = static_cast<int>(n_heavy);
    plan->nslots = NSL;
    plan->ranks = nranks;
    plan->tmem_cols = tcols;
    plan->gather_wavefronts = wavefronts;
    plan->gather_ideal = ideal;
    plan->smem_bytes = smem;
    plan->smem_bytes4 = smem_for(1);
    plan->split_states = n_int - n;
    plan->deg = deg;
    // Sums of scores that are all >= 0 (or -0) stay >= 0, so their float order is the signed-int
    // order of their bit patterns and warp-wide minima can run on the integer REDUX unit.
    {
        auto ok = [](float x) { return !(x < 0.0f) && x == x; };
        bool nn = true;
        for (const float x : csr.w) nn &= ok(x);
        for (const float x : emis) nn &= ok(x);
        for (const float x : start) nn &= ok(x);
        plan->nonneg = nn;
    }
    return true;
}

} // namespace

uint64_t ls_builtin_full_iters(uint64_t states) { return std::min<uint64_t>(2000000ull, 830ull * states); }

uint64_t sv_ls_default_sa_iters(uint64_t states) {
    // annealing length of the FULL search; the min-conflicts end game that follows it does the rest (2 M + end game:
    // 920 wavefronts per step and group on the synthetic 2405-state model, ideal 912; 8 M alone: 955, 40 M alone: 924)
    uint64_t sa_iters = ls_builtin_full_iters(states);
    if (const char* env = std::getenv("SV_LS_SA_ITERS")) sa_iters = std::strtoull(env, nullptr, 10);
    return sa_iters;
}

void sv_ls_build_plan(const SvHostCsr& csr, const std::vector<float>& emis,
                      const std::vector<float>& start, uint64_t emits, size_t smem_limit,
                      SvLsPlan* plan, const SvLsPlanOpts& opts) {
    plan->valid = false;
    const uint64_t sa_iters = opts.sa_iters ? opts.sa_iters : sv_ls_default_sa_iters(csr.n);
    // variants are listed in order of preference; the first that fits wins
    for (int v = 0; v < sv_ls_num_variants(); ++v) {
        SvLsPlan cand;
        if (ls_try_variant(csr, emis, start, emits, smem_limit, v, sa_iters, opts, &cand)) {
            *plan = std::move(cand);
            return;
        }
        if (opts.cancel != nullptr && *opts.cancel != 0) return;
    }
}

// ------------------------------------------------------------------ K0 self-check ----
// Host-only diagnostic (no GPU needed, not on the product path): builds the lockstep schedule
// for a model and replays its register images for ONE sequence exactly as a CTA would (same
// slots, same offsets, same virtual-state merge), so that tests can check the K0
// transformation -- splitting, placement, bank colouring, slot assignment -- against the
// oracle without a device.  stats = {valid, threads, Q, ranks, heavy rows, split states,
// wavefronts, ideal wavefronts, colour cost before, colour cost after, smem bytes, tmem cols}.
extern "C" int sv_k0_lockstep_check(uint64_t states, uint64_t emits, uint64_t trans,
                                    const uint64_t* trans_src, const uint64_t* trans_dst,
                                    const float* trans_w, const float* emissions, uint64_t nstart,
                                    const uint64_t* start_cols, const float* start_w,
                                    const uint64_t* seq, uint64_t len, float* out, uint64_t stats[12]) {
    SV_API_BEGIN
    if (states == 0 || emits == 0 || stats == nullptr) return sv_fail(SV_ERR_ARG, "sv_k0_lockstep_check: bad argument");
    // CSR of T' with FIRST-wins duplicates (same rule as sv_model_create)
    SvHostCsr csr;
    {
        std::vector<uint64_t> order(trans);
        std::iota(order.begin(), order.end(), 0ull);
        std::stable_sort(order.begin(), order.end(), [&](uint64_t a, uint64_t b) {
            return trans_dst[a] != trans_dst[b] ? trans_dst[a] < trans_dst[b] : trans_src[a] < trans_src[b];
        });
        csr.n = states;
        csr.rowptr.assign(states + 1, 0);
        for (size_t k = 0; k < order.size(); ++k) {
            const uint64_t e = order[k];
            if (trans_src[e] >= states || trans_dst[e] >= states) return sv_fail(SV_ERR_ARG, "index >= states");
            if (k > 0 && trans_dst[e] == trans_dst[order[k - 1]] && trans_src[e] == trans_src[order[k - 1]]) continue;
            csr.col.push_back(static_cast<uint32_t>(trans_src[e]));
            csr.w.push_back(trans_w[e]);
            csr.rowptr[trans_dst[e] + 1] += 1;
        }
        for (uint64_t i = 0; i < states; ++i) csr.rowptr[i + 1] += csr.rowptr[i];
    }
    std::vector<float> emis(emissions, emissions + emits * states);
    std::vector<float> start(states, kInf);
    {
        std::vector<unsigned char> seen(states, 0);
        for (uint64_t k = 0; k < nstart; ++k) {
            if (start_cols[k] >= states) return sv_fail(SV_ERR_ARG, "start index >= states");
            if (!seen[start_cols[k]]) {
                seen[start_cols[k]] = 1;
                start[start_cols[k]] = start_w[k];
            }
        }
    }
    SvLsPlan plan;
    SvLsPlanOpts opts;
    opts.threads = static_cast<int>(std::max(1u, std::min(8u, std::thread::hardware_concurrency())));
    // SV_K0_CHECK_USE_CACHE=1: go through the on-disk plan cache exactly as sv_model_ensure_ls does (load, else build
    // and store), so that the cache round trip can be tested without a device
    const char* use_cache = std::getenv("SV_K0_CHECK_USE_CACHE");
    const bool cached = use_cache != nullptr && use_cache[0] == '1';
    const uint64_t fp = cached ? sv_model_fingerprint(csr, emis, start, emits) : 0;
    if (!cached || !sv_plan_cache_load(fp, &plan)) {
        sv_ls_build_plan(csr, emis, start, emits, 227 * 1024, &plan, opts);
        if (cached) sv_plan_cache_store(fp, plan);
    }
    std::memset(stats, 0, 12 * sizeof(uint64_t));
    stats[0] = plan.valid ? 1 : 0;
    if (!plan.valid) return SV_OK;
    const int TPG = plan.var.threads, NW = TPG / 32;
    stats[1] = TPG;
    stats[2] = 8;
    stats[3] = plan.ranks;
    stats[4] = plan.n_heavy;
    stats[5] = plan.split_states;
    stats[6] = plan.gather_wavefronts;
    stats[7] = plan.gather_ideal;
    stats[8] = plan.colour_cost_before;
    stats[9] = plan.colour_cost_after;
    stats[10] = plan.smem_bytes;
    stats[11] = plan.tmem_cols;
    if (seq == nullptr || len == 0 || out == nullptr) return SV_OK;
    const int R = plan.var.ranks;
    int base[SV_LS_RMAX + 1];
    base[0] = 0;
    for (int r = 0; r < R; ++r) base[r + 1] = base[r] + plan.var.shape[r];
    const int NL = base[R];
    const int n_pad = plan.n_pad;
    std::vector<float> cur(n_pad, kInf), nxt(n_pad, kInf);
    for (uint64_t t = 0; t < len; ++t) {
        if (seq[t] >= emits) return sv_fail(SV_ERR_ARG, "symbol >= emit_num");
        const float* em = &plan.emis_p[seq[t] * n_pad];
        for (int p = 0; p < plan.n_lpad; ++p) {
            const int r = p / TPG, tid = p % TPG, w = tid / 32;
            const float e = em[p];
            float acc = kInf;
            const int D = plan.deg[static_cast<size_t>(w) * R + r];
            for (int k = 0; k < D; ++k) {
                const size_t slot = static_cast<size_t>(base[r] + k) * TPG + tid;
                acc = std::fmin(acc, (e + plan.edge_w[slot]) + cur[plan.edge_off[slot] / 16]);
            }
            nxt[p] = t == 0 ? e + plan.start_p[p] : acc;
        }
        for (int h = 0; h < plan.n_heavy; ++h) {
            const int ph = plan.n_lpad + h;
            const float e = em[ph];
            float acc = kInf;
            for (int k = 0; k < plan.heavy_slots[h]; ++k)
                for (int tid = 0; tid < TPG; ++tid) {
                    const size_t slot = static_cast<size_t>(NL + h * plan.var.hs + k) * TPG + tid;
                    acc = std::fmin(acc, (e + plan.edge_w[slot]) + cur[plan.edge_off[slot] / 16]);
                }
            nxt[ph] = t == 0 ? e + plan.start_p[ph] : acc;
        }
        cur.swap(nxt);
    }
    (void)NW;
    for (uint64_t i = 0; i < states; ++i) {
        float x = kInf;
        for (uint32_t c = plan.copy_ptr[i]; c < plan.copy_ptr[i + 1]; ++c) x = std::fmin(x, cur[plan.copy_p[c]]);
        out[i] = x;
    }
    return SV_OK;
    SV_API_END
}
