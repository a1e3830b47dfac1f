// Clone of the reference's benchmark driver (main.cpp:5-21, benchmark/bench_Viterbi.h:17-66,
// bench_Viterbi_spec.h:16-90): for every model in a folder, the wall time of "all sequences of
// one .ess", median of 10 runs, written as a .dat table with one row per states_num.
//   Viterbi_bench_<ds>.dat       States  Oracle_CPU  B200  B200_batch  B200_batch_cold   (integer ms, as the reference)
//     B200 = one run_Viterbi call per sequence (the reference's loop), device model cached across calls;
//     B200_batch = all sequences in one call; B200_batch_cold = the same with a FRESH backend object in every
//     run, i.e. model upload + K0 schedule inside the timed region, as the reference's backends rebuild the
//     model in every call (GraphBLAS_impl.cpp:24-45)
//   Viterbi_spec_bench_<ds>.dat  States  B200_spec_1  B200_spec_1_prep  B200_spec_2  B200_spec_2_prep
// plus a _us.dat twin with microsecond resolution (the B200 columns round to 0 ms).
// usage: bench_viterbi <models dir> <ess file> <dataset name> [level-2 max states = 700] [--no-cpu]
#include "harness_common.h"

#include <dirent.h>

#include <fstream>
#include <iostream>
#include <map>

int main(int argc, char** argv) {
    if (argc < 4) {
        std::cerr << "usage: " << argv[0] << " <models dir> <ess file> <dataset> [lvl2 max states] [--no-cpu]\n";
        return 2;
    }
    const std::string dir = argv[1], ds = argv[3];
    size_t lvl2_max = 1000000; // level 2 runs implicitly (K4i); the table is still built by spec_with
    bool cpu = true;
    for (int i = 4; i < argc; ++i) {
        if (std::string(argv[i]) == "--no-cpu") cpu = false;
        else lvl2_max = std::stoul(argv[i]);
    }
    const auto seqs = read_emit_seq_any(argv[2]);
    if (seqs.empty()) return 2;
    std::vector<std::string> files;
    if (DIR* d = opendir(dir.c_str())) {
        while (dirent* e = readdir(d)) {
            const std::string n = e->d_name;
            if ((n.size() > 5 && n.substr(n.size() - 5) == ".chmm") || (n.size() > 6 && n.substr(n.size() - 6) == ".chmmb"))
                if (n.find("test") == std::string::npos) files.push_back(dir + "/" + n);
        }
        closedir(d);
    }
    std::map<size_t, std::vector<double>> plain_us, spec_us;
    for (const auto& f : files) {
        const HMM hmm = read_HMM_any(f);
        if (hmm.states_num == 0) continue;
        std::cout << "Benchmarking " << f << " (" << hmm.states_num << " states)" << std::endl;
        std::vector<double> row;
        row.push_back(cpu ? harness::median_us([&] { for (const auto& s : seqs) (void)harness::oracle_run(hmm, s); }) : 0.0);
        {
            const B200_impl impl;
            (void)impl.run_Viterbi(hmm, seqs[0]); // first call uploads the model (reported in DESIGN.md)
            row.push_back(harness::median_us([&] { for (const auto& s : seqs) (void)impl.run_Viterbi(hmm, s); }));
            row.push_back(harness::median_us([&] { (void)impl.run_Viterbi_batch(hmm, seqs); }));
        }
        row.push_back(harness::median_us([&] {
            const B200_impl cold;
            (void)cold.run_Viterbi_batch(hmm, seqs);
        }));
        plain_us[hmm.states_num] = row;
        std::vector<double> srow;
        for (size_t level = 1; level <= 2; ++level) {
            if (level == 2 && hmm.states_num > lvl2_max) {
                srow.push_back(0);
                srow.push_back(0);
                continue;
            }
            B200_spec_impl impl(level);
            const double prep = harness::median_us([&] { impl.spec_with(hmm); });
            srow.push_back(harness::median_us([&] { (void)impl.run_Viterbi_spec_batch(seqs); }));
            srow.push_back(prep);
        }
        spec_us[hmm.states_num] = srow;
    }
    auto write = [&](const std::string& name, const std::string& header, const std::map<size_t, std::vector<double>>& t,
                     double div) {
        std::ofstream o(name);
        o << header << "\n";
        for (const auto& kv : t) {
            o << kv.first;
            for (const double v : kv.second) o << "\t" << static_cast<long>(v / div + 0.5);
            o << "\n";
        }
    };
    write("Viterbi_bench_" + ds + ".dat", "States\tOracle_CPU\tB200\tB200_batch\tB200_batch_cold", plain_us, 1000.0);
    write("Viterbi_bench_" + ds + "_us.dat", "States\tOracle_CPU\tB200\tB200_batch\tB200_batch_cold", plain_us, 1.0);
    write("Viterbi_spec_bench_" + ds + ".dat", "States\tB200_spec_1\tB200_spec_1_prep\tB200_spec_2\tB200_spec_2_prep", spec_us, 1000.0);
    write("Viterbi_spec_bench_" + ds + "_us.dat", "States\tB200_spec_1\tB200_spec_1_prep\tB200_spec_2\tB200_spec_2_prep", spec_us, 1.0);
    return 0;
}
