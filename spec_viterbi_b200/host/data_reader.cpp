#include "data_reader.h"

#include <cerrno>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <iostream>

namespace {

// Whole-file tokenizer over whitespace separated numbers.  `operator>>` on an
// ifstream (what the reference uses) converts decimal text with the C-locale
// strtof/strtoull rules, so the parsed values are identical.
class Token_stream {
  public:
    explicit Token_stream(const std::string& name) {
        std::FILE* f = std::fopen(name.c_str(), "rb");
        if (f == nullptr) {
            return;
        }
        std::fseek(f, 0, SEEK_END);
        const long sz = std::ftell(f);
        std::fseek(f, 0, SEEK_SET);
        if (sz >= 0) {
            buf_.resize(static_cast<std::size_t>(sz) + 1, '\0');
            const std::size_t got = std::fread(buf_.data(), 1, static_cast<std::size_t>(sz), f);
            buf_[got] = '\0';
            cur_ = buf_.data();
            ok_ = true;
        }
        std::fclose(f);
    }

    bool opened() const { return ok_; }

    // Failed extraction leaves 0 in the target, like a failed stream extraction
    // value-initialises since C++11.
    std::size_t next_index() {
        char* end = nullptr;
        const unsigned long long v = std::strtoull(cur_, &end, 10);
        if (end == cur_) {
            return 0;
        }
        cur_ = end;
        return static_cast<std::size_t>(v);
    }

    float next_float() {
        char* end = nullptr;
        const float v = std::strtof(cur_, &end);
        if (end == cur_) {
            return 0.0F;
        }
        cur_ = end;
        return v;
    }

  private:
    std::string buf_;
    const char* cur_ = "";
    bool ok_ = false;
};

bool has_suffix(const std::string& s, const char* suffix) {
    const std::size_t n = std::strlen(suffix);
    return s.size() >= n && s.compare(s.size() - n, n, suffix) == 0;
}

constexpr char CHMMB_MAGIC[8] = {'S', 'V', 'C', 'H', 'M', 'M', 'B', '1'};
constexpr char ESSB_MAGIC[8] = {'S', 'V', 'E', 'S', 'S', 'B', '0', '1'};

template <typename T> bool put(std::FILE* f, const T* p, std::size_t n) {
    return n == 0 || std::fwrite(p, sizeof(T), n, f) == n;
}
template <typename T> bool get(std::FILE* f, T* p, std::size_t n) {
    return n == 0 || std::fread(p, sizeof(T), n, f) == n;
}

} // namespace

HMM read_HMM(const std::string& HMM_file_name) {
    Token_stream in(HMM_file_name);
    if (!in.opened()) {
        std::cerr << "Failed to open file with HMM: " << HMM_file_name << '\n';
        return HMM{};
    }

    HMM hmm;
    hmm.states_num = in.next_index();

    hmm.non_zero_start_probs = in.next_index();
    hmm.start_probabilities_cols.reserve(hmm.non_zero_start_probs);
    hmm.start_probabilities.reserve(hmm.non_zero_start_probs);
    for (std::size_t s = 0; s < hmm.non_zero_start_probs; ++s) {
        const std::size_t state = in.next_index();
        const float p = in.next_float();
        hmm.start_probabilities_cols.push_back(state);
        hmm.start_probabilities.push_back(HMM::to_modified_prob(p));
    }

    // File rows are states, columns are symbols; memory is [symbol][state].
    hmm.emit_num = in.next_index();
    hmm.emissions.assign(hmm.emit_num, HMM::Mod_prob_vec_t(hmm.states_num, HMM::zero_prob));
    for (std::size_t state = 0; state < hmm.states_num; ++state) {
        for (std::size_t sym = 0; sym < hmm.emit_num; ++sym) {
            hmm.emissions[sym][state] = HMM::to_modified_prob(in.next_float());
        }
    }

    hmm.trans_num = in.next_index();
    hmm.trans_rows.reserve(hmm.trans_num);
    hmm.trans_cols.reserve(hmm.trans_num);
    hmm.trans_probs.reserve(hmm.trans_num);
    for (std::size_t t = 0; t < hmm.trans_num; ++t) {
        const std::size_t src = in.next_index();
        const std::size_t dst = in.next_index();
        const float p = in.next_float();
        hmm.trans_rows.push_back(src);
        hmm.trans_cols.push_back(dst);
        hmm.trans_probs.push_back(HMM::to_modified_prob(p));
    }
    return hmm;
}

HMM::Emit_seq_vec_t read_emit_seq(const std::string& emit_seq_file_name) {
    Token_stream in(emit_seq_file_name);
    if (!in.opened()) {
        std::cerr << "Failed to open file with emitted sequences: " << emit_seq_file_name << '\n';
        return {};
    }

    const std::size_t count = in.next_index();
    HMM::Emit_seq_vec_t all;
    all.reserve(count);
    for (std::size_t k = 0; k < count; ++k) {
        const std::size_t idx = in.next_index();
        if (idx != k) {
            std::cerr << "Error in .ess file " << emit_seq_file_name
                      << ": expected sequence number is " << k << ", but read " << idx << '\n';
            return {};
        }
        const std::size_t len = in.next_index();
        HMM::Emit_seq_t seq;
        seq.reserve(len);
        for (std::size_t j = 0; j < len; ++j) {
            seq.push_back(in.next_index());
        }
        all.push_back(std::move(seq));
    }
    return all;
}

bool write_HMM_packed(const HMM& hmm, const std::string& file_name) {
    std::FILE* f = std::fopen(file_name.c_str(), "wb");
    if (f == nullptr) {
        return false;
    }
    const std::uint64_t hdr[4] = {hmm.states_num, hmm.emit_num, hmm.trans_num,
                                  hmm.non_zero_start_probs};
    bool ok = put(f, CHMMB_MAGIC, 8) && put(f, hdr, 4);
    std::vector<std::uint32_t> u32(hmm.non_zero_start_probs);
    for (std::size_t i = 0; i < u32.size(); ++i) {
        u32[i] = static_cast<std::uint32_t>(hmm.start_probabilities_cols[i]);
    }
    ok = ok && put(f, u32.data(), u32.size()) &&
         put(f, hmm.start_probabilities.data(), hmm.start_probabilities.size());
    for (std::size_t e = 0; ok && e < hmm.emit_num; ++e) {
        ok = put(f, hmm.emissions[e].data(), hmm.states_num);
    }
    u32.resize(hmm.trans_num);
    for (std::size_t i = 0; i < hmm.trans_num; ++i) {
        u32[i] = static_cast<std::uint32_t>(hmm.trans_rows[i]);
    }
    ok = ok && put(f, u32.data(), u32.size());
    for (std::size_t i = 0; i < hmm.trans_num; ++i) {
        u32[i] = static_cast<std::uint32_t>(hmm.trans_cols[i]);
    }
    ok = ok && put(f, u32.data(), u32.size()) && put(f, hmm.trans_probs.data(), hmm.trans_num);
    std::fclose(f);
    return ok;
}

namespace {
// bytes left in the file from the current position (packed readers check every size field against it
// before allocating: a corrupt header must not turn into a huge allocation or an out-of-bounds read)
std::uint64_t bytes_left(std::FILE* f) {
    const long at = std::ftell(f);
    if (at < 0 || std::fseek(f, 0, SEEK_END) != 0) return 0;
    const long end = std::ftell(f);
    std::fseek(f, at, SEEK_SET);
    return end > at ? static_cast<std::uint64_t>(end - at) : 0;
}
} // namespace

HMM read_HMM_packed(const std::string& file_name) {
    std::FILE* f = std::fopen(file_name.c_str(), "rb");
    if (f == nullptr) {
        std::cerr << "Failed to open file with HMM: " << file_name << '\n';
        return HMM{};
    }
    char magic[8];
    std::uint64_t hdr[4];
    HMM hmm;
    bool ok = get(f, magic, 8) && std::memcmp(magic, CHMMB_MAGIC, 8) == 0 && get(f, hdr, 4);
    if (ok) {
        // start: 8 B per entry; emissions: 4 B x emits x states; transitions: 12 B per entry
        const std::uint64_t left = bytes_left(f);
        const unsigned __int128 need = static_cast<unsigned __int128>(hdr[3]) * 8 +
                                       static_cast<unsigned __int128>(hdr[0]) * hdr[1] * 4 +
                                       static_cast<unsigned __int128>(hdr[2]) * 12;
        ok = hdr[0] > 0 && hdr[1] > 0 && need == left;
    }
    if (ok) {
        hmm.states_num = hdr[0];
        hmm.emit_num = hdr[1];
        hmm.trans_num = hdr[2];
        hmm.non_zero_start_probs = hdr[3];
        std::vector<std::uint32_t> u32(hdr[3]);
        hmm.start_probabilities.resize(hdr[3]);
        ok = get(f, u32.data(), u32.size()) && get(f, hmm.start_probabilities.data(), hdr[3]);
        hmm.start_probabilities_cols.assign(u32.begin(), u32.end());
        hmm.emissions.assign(hmm.emit_num, HMM::Mod_prob_vec_t(hmm.states_num));
        for (std::size_t e = 0; ok && e < hmm.emit_num; ++e) {
            ok = get(f, hmm.emissions[e].data(), hmm.states_num);
        }
        u32.resize(hmm.trans_num);
        ok = ok && get(f, u32.data(), u32.size());
        hmm.trans_rows.assign(u32.begin(), u32.end());
        ok = ok && get(f, u32.data(), u32.size());
        hmm.trans_cols.assign(u32.begin(), u32.end());
        hmm.trans_probs.resize(hmm.trans_num);
        ok = ok && get(f, hmm.trans_probs.data(), hmm.trans_num);
        for (std::size_t i = 0; ok && i < hmm.trans_num; ++i) {
            ok = hmm.trans_rows[i] < hmm.states_num && hmm.trans_cols[i] < hmm.states_num;
        }
        for (std::size_t i = 0; ok && i < hmm.non_zero_start_probs; ++i) {
            ok = hmm.start_probabilities_cols[i] < hmm.states_num;
        }
    }
    std::fclose(f);
    return ok ? hmm : HMM{};
}

bool write_emit_seq_packed(const HMM::Emit_seq_vec_t& seqs, const std::string& file_name) {
    std::FILE* f = std::fopen(file_name.c_str(), "wb");
    if (f == nullptr) {
        return false;
    }
    std::vector<std::uint64_t> offs(seqs.size() + 1, 0);
    std::size_t max_sym = 0;
    for (std::size_t i = 0; i < seqs.size(); ++i) {
        offs[i + 1] = offs[i] + seqs[i].size();
        for (const auto s : seqs[i]) {
            max_sym = s > max_sym ? s : max_sym;
        }
    }
    const std::uint64_t width = max_sym < 256 ? 1 : (max_sym < 65536 ? 2 : 4);
    const std::uint64_t hdr[3] = {seqs.size(), offs.back(), width};
    bool ok = put(f, ESSB_MAGIC, 8) && put(f, hdr, 3) && put(f, offs.data(), offs.size());
    for (std::size_t i = 0; ok && i < seqs.size(); ++i) {
        for (const auto s : seqs[i]) {
            const std::uint32_t v = static_cast<std::uint32_t>(s);
            ok = ok && std::fwrite(&v, width, 1, f) == 1; // little endian
        }
    }
    std::fclose(f);
    return ok;
}

HMM::Emit_seq_vec_t read_emit_seq_packed(const std::string& file_name) {
    std::FILE* f = std::fopen(file_name.c_str(), "rb");
    if (f == nullptr) {
        std::cerr << "Failed to open file with emitted sequences: " << file_name << '\n';
        return {};
    }
    char magic[8];
    std::uint64_t hdr[3];
    HMM::Emit_seq_vec_t seqs;
    bool ok = get(f, magic, 8) && std::memcmp(magic, ESSB_MAGIC, 8) == 0 && get(f, hdr, 3);
    if (ok) {
        // hdr = {sequences, total symbols, bytes per symbol}: everything is checked against the file size
        // before anything is allocated, the width before it is used as a memcpy length
        const std::uint64_t left = bytes_left(f);
        const unsigned __int128 need = (static_cast<unsigned __int128>(hdr[0]) + 1) * 8 +
                                       static_cast<unsigned __int128>(hdr[1]) * hdr[2];
        ok = (hdr[2] == 1 || hdr[2] == 2 || hdr[2] == 4) && need == left;
    }
    if (ok) {
        std::vector<std::uint64_t> offs(hdr[0] + 1);
        ok = get(f, offs.data(), offs.size());
        ok = ok && offs[0] == 0 && offs.back() == hdr[1];
        for (std::size_t i = 0; ok && i < hdr[0]; ++i) {
            ok = offs[i] <= offs[i + 1];
        }
        std::vector<unsigned char> raw;
        if (ok) {
            raw.resize(hdr[1] * hdr[2]);
            ok = get(f, raw.data(), raw.size());
        }
        if (ok) {
            seqs.resize(hdr[0]);
            for (std::size_t i = 0; i < hdr[0]; ++i) {
                seqs[i].reserve(offs[i + 1] - offs[i]);
                for (std::uint64_t k = offs[i]; k < offs[i + 1]; ++k) {
                    std::uint32_t v = 0;
                    std::memcpy(&v, raw.data() + k * hdr[2], hdr[2]);
                    seqs[i].push_back(v);
                }
            }
        }
    }
    std::fclose(f);
    return ok ? seqs : HMM::Emit_seq_vec_t{};
}

HMM read_HMM_any(const std::string& file_name) {
    return has_suffix(file_name, ".chmmb") ? read_HMM_packed(file_name) : read_HMM(file_name);
}

HMM::Emit_seq_vec_t read_emit_seq_any(const std::string& file_name) {
    return has_suffix(file_name, ".essb") ? read_emit_seq_packed(file_name)
                                          : read_emit_seq(file_name);
}
