// On-disk cache of finished K0 plans of the lockstep kernel.
//
// The bank-colouring search of sv_ls_plan.cpp is the only part of model creation that takes more
// than a few milliseconds (seconds for a random 2405-state model at full length).  Its result only
// reorders work -- any plan gives the oracle's bits -- so it is safe to reuse across processes:
// the plan is keyed by a fingerprint of everything it was derived from (T' in CSR form, the emission
// table, the start column) plus a format version that changes with the planner / kernel layout.
//
// Location: $SV_PLAN_CACHE_DIR, else $XDG_CACHE_HOME/spec_viterbi_b200, else
// $HOME/.cache/spec_viterbi_b200, else /tmp/spec_viterbi_b200.  SV_PLAN_CACHE=0 disables it.
// Files are written to a temporary name and renamed, so concurrent processes never see a torn file.
#include "sv_internal.h"

#include <cerrno>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <sys/stat.h>
#include <sys/types.h>
#include <unistd.h>

namespace {

constexpr char kMagic[8] = {'S', 'V', 'L', 'S', 'P', 'L', 'N', '3'};
// bump when SvLsPlan's meaning changes (planner, variant table, kernel register layout)
constexpr uint64_t kPlanVersion = 0x20261106ull; // end-game colouring search: plans cached by the older search are rebuilt

inline uint64_t mix(uint64_t h, uint64_t x) {
    h ^= x;
    h *= 0x9E3779B97F4A7C15ull;
    h ^= h >> 29;
    return h;
}

uint64_t hash_words(uint64_t h, const void* p, size_t bytes) {
    const unsigned char* b = static_cast<const unsigned char*>(p);
    size_t k = 0;
    for (; k + 8 <= bytes; k += 8) {
        uint64_t w;
        std::memcpy(&w, b + k, 8);
        h = mix(h, w);
    }
    uint64_t tail = 0;
    if (k < bytes) std::memcpy(&tail, b + k, bytes - k);
    return mix(mix(h, tail), bytes);
}

bool cache_enabled() {
    const char* e = std::getenv("SV_PLAN_CACHE");
    return !(e != nullptr && e[0] == '0');
}

std::string cache_dir() {
    if (const char* d = std::getenv("SV_PLAN_CACHE_DIR")) return d;
    if (const char* x = std::getenv("XDG_CACHE_HOME"))
        if (x[0]) return std::string(x) + "/spec_viterbi_b200";
    if (const char* h = std::getenv("HOME"))
        if (h[0]) return std::string(h) + "/.cache/spec_viterbi_b200";
    return "/tmp/spec_viterbi_b200";
}

bool make_dirs(const std::string& path) {
    std::string cur;
    for (size_t k = 0; k <= path.size(); ++k) {
        if (k == path.size() || path[k] == '/') {
            if (!cur.empty() && mkdir(cur.c_str(), 0755) != 0 && errno != EEXIST) return false;
        }
        if (k < path.size()) cur.push_back(path[k]);
    }
    return true;
}

std::string cache_path(uint64_t fp) {
    char name[64];
    std::snprintf(name, sizeof(name), "/ls-%016llx.svplan", static_cast<unsigned long long>(fp));
    return cache_dir() + name;
}

template <typename T> bool put_vec(std::FILE* f, const std::vector<T>& v) {
    const uint64_t n = v.size();
    return std::fwrite(&n, 8, 1, f) == 1 && (n == 0 || std::fwrite(v.data(), sizeof(T), n, f) == n);
}
template <typename T> bool get_vec(std::FILE* f, std::vector<T>* v, uint64_t max_elems) {
    uint64_t n = 0;
    if (std::fread(&n, 8, 1, f) != 1 || n > max_elems) return false;
    v->resize(n);
    return n == 0 || std::fread(v->data(), sizeof(T), n, f) == n;
}

struct PlanHeader {
    char magic[8];
    uint64_t version, fingerprint;
    int32_t variant_index, n_int, n_lpad, n_pad, n_heavy, nslots, ranks, tmem_cols, nonneg, pad;
    uint64_t gather_wavefronts, gather_ideal, split_states, smem_bytes, smem_bytes4, sa_iters, cost_before, cost_after;
    uint64_t payload_hash; // over every vector below: a damaged payload must never become a wrong schedule
};

uint64_t payload_hash(const SvLsPlan& p) {
    uint64_t h = 0x5061796c6f616421ull;
    h = hash_words(h, p.edge_w.data(), p.edge_w.size() * sizeof(float));
    h = hash_words(h, p.edge_off.data(), p.edge_off.size() * sizeof(uint32_t));
    h = hash_words(h, p.deg.data(), p.deg.size());
    h = hash_words(h, p.heavy_slots.data(), p.heavy_slots.size());
    h = hash_words(h, p.emis_p.data(), p.emis_p.size() * sizeof(float));
    h = hash_words(h, p.start_p.data(), p.start_p.size() * sizeof(float));
    h = hash_words(h, p.copy_ptr.data(), p.copy_ptr.size() * sizeof(uint32_t));
    h = hash_words(h, p.copy_p.data(), p.copy_p.size() * sizeof(uint32_t));
    return h;
}

} // namespace

uint64_t sv_model_fingerprint(const SvHostCsr& csr, const std::vector<float>& emis,
                              const std::vector<float>& start, uint64_t emits) {
    uint64_t h = 0x5356504C414E0001ull; // "SVPLAN"
    h = mix(h, csr.n);
    h = mix(h, emits);
    h = hash_words(h, csr.rowptr.data(), csr.rowptr.size() * sizeof(uint32_t));
    h = hash_words(h, csr.col.data(), csr.col.size() * sizeof(uint32_t));
    h = hash_words(h, csr.w.data(), csr.w.size() * sizeof(float));
    h = hash_words(h, emis.data(), emis.size() * sizeof(float));
    h = hash_words(h, start.data(), start.size() * sizeof(float));
    return h;
}

bool sv_plan_cache_load(uint64_t fingerprint, SvLsPlan* plan) {
    if (!cache_enabled()) return false;
    std::FILE* f = std::fopen(cache_path(fingerprint).c_str(), "rb");
    if (f == nullptr) return false;
    PlanHeader h;
    SvLsPlan p;
    bool ok = std::fread(&h, sizeof(h), 1, f) == 1 && std::memcmp(h.magic, kMagic, 8) == 0 &&
              h.version == kPlanVersion && h.fingerprint == fingerprint && h.variant_index >= 0 &&
              h.variant_index < sv_ls_num_variants();
    const uint64_t cap = 1ull << 26;
    ok = ok && get_vec(f, &p.edge_w, cap) && get_vec(f, &p.edge_off, cap) && get_vec(f, &p.deg, cap) &&
         get_vec(f, &p.heavy_slots, cap) && get_vec(f, &p.emis_p, cap) && get_vec(f, &p.start_p, cap) &&
         get_vec(f, &p.copy_ptr, cap) && get_vec(f, &p.copy_p, cap);
    std::fclose(f);
    if (!ok || payload_hash(p) != h.payload_hash) return false;
    p.var = sv_ls_variant(h.variant_index);
    p.variant_index = h.variant_index;
    p.n_int = h.n_int;
    p.n_lpad = h.n_lpad;
    p.n_pad = h.n_pad;
    p.n_heavy = h.n_heavy;
    p.nslots = h.nslots;
    p.ranks = h.ranks;
    p.tmem_cols = h.tmem_cols;
    p.nonneg = h.nonneg != 0;
    p.gather_wavefronts = h.gather_wavefronts;
    p.gather_ideal = h.gather_ideal;
    p.split_states = h.split_states;
    p.smem_bytes = h.smem_bytes;
    p.smem_bytes4 = h.smem_bytes4;
    p.sa_iters = h.sa_iters;
    p.colour_cost_before = h.cost_before;
    p.colour_cost_after = h.cost_after;
    // structural sanity: sizes must match what the kernel will index
    const size_t tpg = static_cast<size_t>(p.var.threads);
    if (p.n_lpad <= 0 || p.n_pad < p.n_lpad || p.ranks <= 0 || p.ranks > p.var.ranks ||
        p.edge_w.size() != static_cast<size_t>(p.nslots) * tpg || p.edge_off.size() != p.edge_w.size() ||
        p.deg.size() != tpg / 32 * static_cast<size_t>(p.var.ranks) || p.start_p.size() != static_cast<size_t>(p.n_pad) ||
        p.emis_p.size() % static_cast<size_t>(p.n_pad) != 0 || p.heavy_slots.size() != SV_ES_NHMAX ||
        p.copy_p.size() != static_cast<size_t>(p.n_int))
        return false;
    for (const uint32_t off : p.edge_off)
        if (off / 16u >= static_cast<uint32_t>(p.n_pad)) return false;
    for (const uint32_t cp : p.copy_p)
        if (cp >= static_cast<uint32_t>(p.n_pad)) return false;
    p.valid = true;
    *plan = std::move(p);
    return true;
}

void sv_plan_cache_store(uint64_t fingerprint, const SvLsPlan& p) {
    if (!cache_enabled() || !p.valid) return;
    const std::string dir = cache_dir();
    if (!make_dirs(dir)) return;
    const std::string path = cache_path(fingerprint);
    char suffix[48];
    std::snprintf(suffix, sizeof(suffix), ".tmp.%ld", static_cast<long>(getpid()));
    const std::string tmp = path + suffix;
    std::FILE* f = std::fopen(tmp.c_str(), "wb");
    if (f == nullptr) return;
    PlanHeader h;
    std::memset(&h, 0, sizeof(h));
    std::memcpy(h.magic, kMagic, 8);
    h.version = kPlanVersion;
    h.fingerprint = fingerprint;
    h.variant_index = p.variant_index;
    h.n_int = p.n_int;
    h.n_lpad = p.n_lpad;
    h.n_pad = p.n_pad;
    h.n_heavy = p.n_heavy;
    h.nslots = p.nslots;
    h.ranks = p.ranks;
    h.tmem_cols = p.tmem_cols;
    h.nonneg = p.nonneg ? 1 : 0;
    h.gather_wavefronts = p.gather_wavefronts;
    h.gather_ideal = p.gather_ideal;
    h.split_states = p.split_states;
    h.smem_bytes = p.smem_bytes;
    h.smem_bytes4 = p.smem_bytes4;
    h.sa_iters = p.sa_iters;
    h.cost_before = p.colour_cost_before;
    h.cost_after = p.colour_cost_after;
    h.payload_hash = payload_hash(p);
    bool ok = std::fwrite(&h, sizeof(h), 1, f) == 1 && put_vec(f, p.edge_w) && put_vec(f, p.edge_off) &&
              put_vec(f, p.deg) && put_vec(f, p.heavy_slots) && put_vec(f, p.emis_p) && put_vec(f, p.start_p) &&
              put_vec(f, p.copy_ptr) && put_vec(f, p.copy_p);
    ok = (std::fclose(f) == 0) && ok;
    if (!ok || std::rename(tmp.c_str(), path.c_str()) != 0) std::remove(tmp.c_str());
}
