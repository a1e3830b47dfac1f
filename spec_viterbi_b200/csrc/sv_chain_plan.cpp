// K0 for the chain kernel (sv_plain_chain.cu): detects the "chain + specials" structure of profile
// HMMs and lays the model out for register-resident state vectors.
//
// Structure (never assumed, always detected -- SURVEY A.1): remove at most two SPECIAL states
// (states with more than kLightDegree in- or out-edges; N and C of a profile HMM) and drop +inf
// transitions (fl(fl(E + inf) + v) = +inf never wins the min).  What is left must be a union of
// simple paths: every remaining state has at most one predecessor and one successor among the
// remaining states, no cycles, no self loops.  Additionally
//   * only special 0 (the "hub") may have edges INTO chain states; any weights;
//   * a special row's in-edges from chain states are either absent or present from EVERY chain
//     state with one and the same weight u_s.  Then, because x -> fl(c + x) is monotone,
//         min_k fl(fl(E_o[s] + u_s) + v[k]) = fl(fl(E_o[s] + u_s) + min_k v[k]),
//     bit for bit: the dense row costs one min per chain state instead of two adds and a min;
//   * special -> special edges are arbitrary (a 2 x 2 block).
// Chain position p = thread * SV_CHAIN_RPT + rank follows the paths, so the band edge (p-1) -> p
// stays inside a thread except at rank 0, where it comes from the neighbouring lane.
//
// The reference's 24 shipped models all have this structure (states 0 = N, 1..L = M_k, L+1 = C;
// chmm_files/silent_hmm_to_chmm.py:93-144).  Anything else falls back to the other kernels.
#include "sv_internal.h"

#include <algorithm>
#include <cmath>
#include <cstring>
#include <limits>
#include <numeric>

namespace {
constexpr float kInf = std::numeric_limits<float>::infinity();
constexpr size_t kLightDegree = 3; // a chain state has <= 1 + NSP in-edges and <= 1 + NSP out-edges
} // namespace

void sv_chain_build_plan(const SvHostCsr& csr, const std::vector<float>& emis, const std::vector<float>& start,
                         uint64_t emits, SvChainPlan* plan) {
    *plan = SvChainPlan();
    const size_t n = csr.n;
    if (n < 3 || emits == 0 || emits > 256) return;
    // finite edges only
    std::vector<std::vector<std::pair<uint32_t, float>>> in(n), out(n); // (other end, weight)
    for (uint32_t i = 0; i < n; ++i)
        for (uint32_t p = csr.rowptr[i]; p < csr.rowptr[i + 1]; ++p) {
            if (csr.w[p] == kInf) continue;
            in[i].emplace_back(csr.col[p], csr.w[p]);
            out[csr.col[p]].emplace_back(i, csr.w[p]);
        }
    // specials: states that are not chain-like by degree
    std::vector<uint32_t> sp;
    for (uint32_t i = 0; i < n; ++i)
        if (in[i].size() > kLightDegree || out[i].size() > kLightDegree) {
            if (sp.size() == SV_CHAIN_NSP) return;
            sp.push_back(i);
        }
    if (sp.empty()) return; // a bare chain has no parallelism worth a kernel of its own
    std::vector<int> sidx(n, -1);
    // the hub (special with edges into non-special states) must be special 0
    auto feeds_chain = [&](uint32_t x) {
        for (const auto& e : out[x]) {
            bool to_special = false;
            for (const uint32_t y : sp) to_special |= y == e.first;
            if (!to_special) return true;
        }
        return false;
    };
    if (sp.size() == 2 && !feeds_chain(sp[0]) && feeds_chain(sp[1])) std::swap(sp[0], sp[1]);
    if (sp.size() == 2 && feeds_chain(sp[1])) return; // two hubs: not supported
    for (size_t k = 0; k < sp.size(); ++k) sidx[sp[k]] = static_cast<int>(k);

    // chain structure among the remaining states
    const size_t n_chain = n - sp.size();
    std::vector<int32_t> pred(n, -1), succ(n, -1);
    std::vector<float> wpred(n, kInf), whub(n, kInf);
    for (uint32_t i = 0; i < n; ++i) {
        if (sidx[i] >= 0) continue;
        for (const auto& e : in[i]) {
            const uint32_t j = e.first;
            if (sidx[j] == 0) {
                whub[i] = e.second;
            } else if (sidx[j] > 0) {
                return; // only the hub feeds the chain
            } else {
                if (j == i || pred[i] >= 0 || succ[j] >= 0) return; // self loop / second predecessor / fork
                pred[i] = static_cast<int32_t>(j);
                succ[j] = static_cast<int32_t>(i);
                wpred[i] = e.second;
            }
        }
    }
    // special rows: chain in-edges uniform over ALL chain states, or absent
    float u[SV_CHAIN_NSP], ss[SV_CHAIN_NSP][SV_CHAIN_NSP];
    for (int a = 0; a < SV_CHAIN_NSP; ++a) {
        u[a] = kInf;
        for (int b = 0; b < SV_CHAIN_NSP; ++b) ss[a][b] = kInf;
    }
    for (size_t k = 0; k < sp.size(); ++k) {
        size_t from_chain = 0;
        bool have_u = false;
        for (const auto& e : in[sp[k]]) {
            if (sidx[e.first] >= 0) {
                ss[k][sidx[e.first]] = e.second;
                continue;
            }
            if (!have_u) {
                u[k] = e.second;
                have_u = true;
            } else if (std::memcmp(&u[k], &e.second, sizeof(float)) != 0) {
                return; // two different weights: would need the general dense-row path
            }
            ++from_chain;
        }
        if (from_chain != 0 && from_chain != n_chain) return;
    }
    // order the chain states along their paths (path heads in ascending state order)
    std::vector<uint32_t> order;
    order.reserve(n_chain);
    for (uint32_t i = 0; i < n; ++i) {
        if (sidx[i] >= 0 || pred[i] >= 0) continue;
        for (int32_t x = static_cast<int32_t>(i); x >= 0; x = succ[x]) order.push_back(static_cast<uint32_t>(x));
    }
    if (order.size() != n_chain) return; // a cycle never gets a head

    int threads = 0;
    for (const int t : {128, 256, 512})
        if (static_cast<size_t>(t) * SV_CHAIN_RPT >= n_chain) {
            threads = t;
            break;
        }
    if (threads == 0) return;
    // + 2: the one- and two-sequence kernels fetch a symbol's six columns with one 8-column load
    const size_t tcols_need = static_cast<size_t>(threads / 128) * emits * SV_CHAIN_TMEM_CS + 2;
    if (tcols_need > 512) return;
    int tcols = 32;
    while (static_cast<size_t>(tcols) < tcols_need) tcols *= 2;

    SvChainPlan& p = *plan;
    p.threads = threads;
    p.n_chain = static_cast<int>(n_chain);
    p.n_pos = threads * SV_CHAIN_RPT;
    p.nsp = static_cast<int>(sp.size());
    p.tmem_cols = tcols;
    for (int a = 0; a < SV_CHAIN_NSP; ++a) {
        p.special[a] = a < p.nsp ? sp[a] : 0u;
        p.u[a] = u[a];
        p.start_s[a] = a < p.nsp ? start[sp[a]] : kInf;
        for (int b = 0; b < SV_CHAIN_NSP; ++b) p.ss[a][b] = ss[a][b];
    }
    p.w_pred.assign(p.n_pos, kInf);
    p.w_hub.assign(p.n_pos, kInf);
    p.start_p.assign(p.n_pos, kInf);
    p.emis_p.assign(emits * static_cast<size_t>(p.n_pos), 0.f);
    p.orig_of_pos.assign(p.n_pos, -1);
    for (size_t k = 0; k < order.size(); ++k) {
        const uint32_t x = order[k];
        p.orig_of_pos[k] = static_cast<int32_t>(x);
        // the band edge is "from position k-1": true exactly when the predecessor sits there
        p.w_pred[k] = (pred[x] >= 0 && k > 0 && order[k - 1] == static_cast<uint32_t>(pred[x])) ? wpred[x] : kInf;
        p.w_hub[k] = whub[x];
        p.start_p[k] = start[x];
        for (uint64_t e = 0; e < emits; ++e) p.emis_p[e * p.n_pos + k] = emis[e * n + x];
    }
    // rows 0 .. NSP-1: emissions of the special states; row NSP: min over the chain of fl(E[o][p] + w_hub[p]), the smallest
    // level-1 entry of the hub column per symbol -- K4c folds "special row <- chain state p <- hub" through it (monotonicity)
    p.emis_s.assign(static_cast<size_t>(SV_CHAIN_NSP + 1) * emits, 0.f);
    for (int a = 0; a < p.nsp; ++a)
        for (uint64_t e = 0; e < emits; ++e) p.emis_s[a * emits + e] = emis[e * n + sp[a]];
    for (uint64_t e = 0; e < emits; ++e) {
        float mh = kInf;
        for (int k = 0; k < p.n_pos; ++k) mh = std::fmin(mh, p.emis_p[e * p.n_pos + k] + p.w_hub[k]);
        p.emis_s[static_cast<size_t>(SV_CHAIN_NSP) * emits + e] = mh;
    }
    {
        auto ok = [](float x) { return !(x < 0.0f) && x == x; };
        bool nn = true;
        for (const float x : csr.w) nn &= ok(x);
        for (const float x : emis) nn &= ok(x);
        for (const float x : start) nn &= ok(x);
        p.nonneg = nn; // selects the kernel instantiation whose REDUX needs no order-preserving key
    }
    p.valid = true;
}

// ------------------------------------------------------------------ K0 self-check ----
// Host-only diagnostic (no GPU, not a compute path): builds the chain plan of a model and replays
// it for ONE sequence with exactly the algebra of the kernel -- band edge from position p-1, hub
// broadcast, special rows through the chain minimum -- so that tests can check the transformation
// against the oracle without a device.  stats = {valid, threads, chain states, specials, tmem cols,
// nonneg, special 0, special 1}.
namespace {
int chain_check_impl(int level, uint64_t states, uint64_t emits, uint64_t trans, const uint64_t* trans_src,
                     const uint64_t* trans_dst, const float* trans_w, const float* emissions,
                     uint64_t nstart, const uint64_t* start_cols, const float* start_w,
                     const uint64_t* seq, uint64_t len, float* out, uint64_t stats[8]) {
    if (states == 0 || emits == 0 || stats == nullptr) return sv_fail(SV_ERR_ARG, "sv_k0_chain_check: bad argument");
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
    SvChainPlan p;
    sv_chain_build_plan(csr, emis, start, emits, &p);
    std::memset(stats, 0, 8 * sizeof(uint64_t));
    stats[0] = p.valid ? 1 : 0;
    if (!p.valid) return SV_OK;
    stats[1] = p.threads;
    stats[2] = p.n_chain;
    stats[3] = p.nsp;
    stats[4] = p.tmem_cols;
    stats[5] = p.nonneg ? 1 : 0;
    stats[6] = p.special[0];
    stats[7] = p.special[1];
    if (seq == nullptr || len == 0 || out == nullptr) return SV_OK;
    const int np = p.n_pos;
    std::vector<float> v(np, kInf), nv(np, kInf);
    float vs[SV_CHAIN_NSP] = {kInf, kInf}, nvs[SV_CHAIN_NSP];
    for (uint64_t t = 0; t < len; ++t)
        if (seq[t] >= emits) return sv_fail(SV_ERR_ARG, "symbol >= emit_num");
    for (uint64_t t = 0; t < len; ++t) {
        if (level == 2 && t >= 1 && len - t >= 2) {
            // K4c (sv_spec_chain.cu): one level-2 chunk (a, b) = the min over the two-step paths i <- j <- k of
            // fl(fl(M_b[i][j] + M_a[j][k]) + v[k]), with the paths through the dense rows collapsed onto the chain
            // minimum (GraphBLAS_spec_impl.cpp:15-36,69-79 are what it must reproduce)
            const float* ea = &p.emis_p[seq[t] * np];
            const float* eb = &p.emis_p[seq[t + 1] * np];
            float mm = kInf;
            for (int k = 0; k < np; ++k) mm = std::fmin(mm, v[k]);
            auto at = [&](int k) { return k >= 0 ? v[k] : kInf; };
            float Ua[SV_CHAIN_NSP], Ub[SV_CHAIN_NSP], Sa[SV_CHAIN_NSP][SV_CHAIN_NSP], Sb[SV_CHAIN_NSP][SV_CHAIN_NSP];
            for (int x = 0; x < SV_CHAIN_NSP; ++x) {
                const float xa = p.emis_s[x * emits + seq[t]], xb = p.emis_s[x * emits + seq[t + 1]];
                Ua[x] = xa + p.u[x];
                Ub[x] = xb + p.u[x];
                for (int y = 0; y < SV_CHAIN_NSP; ++y) {
                    Sa[x][y] = xa + p.ss[x][y];
                    Sb[x][y] = xb + p.ss[x][y];
                }
            }
            for (int x = 0; x < SV_CHAIN_NSP; ++x) nvs[x] = kInf;
            for (int k = 0; k < np; ++k) {
                const float Pa = ea[k] + p.w_pred[k], Ha = ea[k] + p.w_hub[k];
                const float Pb = eb[k] + p.w_pred[k], Hb = eb[k] + p.w_hub[k];
                const float Pam1 = k > 0 ? ea[k - 1] + p.w_pred[k - 1] : kInf;
                const float Ham1 = k > 0 ? ea[k - 1] + p.w_hub[k - 1] : kInf;
                float x = (Pb + Pam1) + at(k - 2);
                x = std::fmin(x, (Pb + Ham1) + vs[0]);
                x = std::fmin(x, (Hb + Ua[0]) + mm);
                for (int y = 0; y < SV_CHAIN_NSP; ++y) x = std::fmin(x, (Hb + Sa[0][y]) + vs[y]);
                nv[k] = x;
                for (int y = 0; y < SV_CHAIN_NSP; ++y) {
                    nvs[y] = std::fmin(nvs[y], (Ub[y] + Pa) + at(k - 1));
                    nvs[y] = std::fmin(nvs[y], (Ub[y] + Ha) + vs[0]);
                }
            }
            for (int x = 0; x < SV_CHAIN_NSP; ++x)
                for (int y = 0; y < SV_CHAIN_NSP; ++y) {
                    nvs[x] = std::fmin(nvs[x], (Sb[x][y] + Ua[y]) + mm);
                    for (int z = 0; z < SV_CHAIN_NSP; ++z) nvs[x] = std::fmin(nvs[x], (Sb[x][y] + Sa[y][z]) + vs[z]);
                }
            v.swap(nv);
            for (int x = 0; x < SV_CHAIN_NSP; ++x) vs[x] = nvs[x];
            ++t; // two symbols consumed
            continue;
        }
        const float* em = &p.emis_p[seq[t] * np];
        float mm = kInf; // min over the chain of the PREVIOUS step's scores
        for (int k = 0; k < np; ++k) mm = std::fmin(mm, v[k]);
        for (int k = 0; k < np; ++k) {
            const float e = em[k];
            if (t == 0) {
                nv[k] = e + p.start_p[k];
                continue;
            }
            const float a = (e + p.w_pred[k]) + (k > 0 ? v[k - 1] : kInf);
            const float b = (e + p.w_hub[k]) + vs[0];
            nv[k] = std::fmin(a, b);
        }
        for (int s = 0; s < SV_CHAIN_NSP; ++s) {
            const float e = p.emis_s[s * emits + seq[t]];
            if (t == 0) {
                nvs[s] = e + p.start_s[s];
                continue;
            }
            float x = (e + p.u[s]) + mm;
            for (int s2 = 0; s2 < SV_CHAIN_NSP; ++s2) x = std::fmin(x, (e + p.ss[s][s2]) + vs[s2]);
            nvs[s] = x;
        }
        v.swap(nv);
        for (int s = 0; s < SV_CHAIN_NSP; ++s) vs[s] = nvs[s];
    }
    for (uint64_t i = 0; i < states; ++i) out[i] = kInf;
    for (int k = 0; k < np; ++k)
        if (p.orig_of_pos[k] >= 0) out[p.orig_of_pos[k]] = v[k];
    for (int s = 0; s < p.nsp; ++s) out[p.special[s]] = vs[s];
    return SV_OK;
}
} // namespace

extern "C" int sv_k0_chain_check(uint64_t states, uint64_t emits, uint64_t trans, const uint64_t* trans_src,
                                 const uint64_t* trans_dst, const float* trans_w, const float* emissions,
                                 uint64_t nstart, const uint64_t* start_cols, const float* start_w,
                                 const uint64_t* seq, uint64_t len, float* out, uint64_t stats[8]) {
    SV_API_BEGIN
    return chain_check_impl(1, states, emits, trans, trans_src, trans_dst, trans_w, emissions, nstart, start_cols, start_w, seq,
                            len, out, stats);
    SV_API_END
}

// The same for K4c (sv_spec_chain.cu): replays the level-2 specialised run -- first symbol, chunks of two symbols through
// the collapsed two-step paths, odd tail as one level-1 step -- so that the algebra can be checked against the level-2
// oracle without a device.
extern "C" int sv_k0_chain_pair_check(uint64_t states, uint64_t emits, uint64_t trans, const uint64_t* trans_src,
                                      const uint64_t* trans_dst, const float* trans_w, const float* emissions,
                                      uint64_t nstart, const uint64_t* start_cols, const float* start_w,
                                      const uint64_t* seq, uint64_t len, float* out, uint64_t stats[8]) {
    SV_API_BEGIN
    return chain_check_impl(2, states, emits, trans, trans_src, trans_dst, trans_w, emissions, nstart, start_cols, start_w, seq,
                            len, out, stats);
    SV_API_END
}
