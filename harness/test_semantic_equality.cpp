// Clone of the reference's cross-backend parity test (tests/test_semantic_equality.cpp:19-98)
// with the B200 backend pushed into both implementation vectors:
//   non-spec = [oracle (GraphBLAS_impl restatement), B200_impl]
//   spec     = [B200_spec_impl(1), B200_spec_impl(2) on the small models]
// For every model in the folder and every sequence: the reference's own criterion (abs <= 1.0
// or both inf) chained over neighbours, PLUS bit-equality B200 vs oracle on the same path.
// usage: test_semantic_equality <models dir> <sequence file> [max states for level 2 = 400]
#include "harness_common.h"

#include <cstdio>
#include <dirent.h>
#include <unistd.h>

#include <iostream>
#include <memory>

int main(int argc, char** argv) {
    if (argc < 3) {
        std::cerr << "usage: " << argv[0] << " <models dir> <ess file> [level-2 max states]\n";
        return 2;
    }
    const std::string dir = argv[1];
    const size_t lvl2_max = argc > 3 ? std::stoul(argv[3]) : 400;
    const auto sequences = read_emit_seq_any(argv[2]);
    if (sequences.empty()) {
        std::cerr << "cannot read sequences\n";
        return 2;
    }
    std::vector<std::string> files;
    if (DIR* d = opendir(dir.c_str())) {
        while (dirent* e = readdir(d)) {
            const std::string n = e->d_name;
            if ((n.size() > 5 && n.substr(n.size() - 5) == ".chmm") || (n.size() > 6 && n.substr(n.size() - 6) == ".chmmb"))
                if (n.find("test") == std::string::npos) files.push_back(dir + "/" + n);
        }
        closedir(d);
    }
    std::sort(files.begin(), files.end());
    int failures = 0;
    for (const auto& f : files) {
        const HMM hmm = read_HMM_any(f);
        if (hmm.states_num == 0) {
            std::cerr << "cannot read " << f << "\n";
            return 2;
        }
        const B200_impl b200;
        std::vector<std::unique_ptr<B200_spec_impl>> spec;
        spec.emplace_back(new B200_spec_impl(hmm, 1));
        if (hmm.states_num <= lvl2_max) spec.emplace_back(new B200_spec_impl(hmm, 2));
        const auto batch = b200.run_Viterbi_batch(hmm, sequences);
        for (size_t s = 0; s < sequences.size(); ++s) {
            const auto ref = harness::oracle_run(hmm, sequences[s]);
            const auto one = b200.run_Viterbi(hmm, sequences[s]);
            bool ok = harness::compare_two_answers(ref, one) && harness::bit_equal(ref, one) &&
                      harness::bit_equal(ref, batch[s]);
            for (const auto& sp : spec) {
                const auto r = sp->run_Viterbi_spec(sequences[s]);
                ok = ok && harness::compare_two_answers(one, r);           // the reference's criterion
                if (sp->get_level() == 1) ok = ok && harness::bit_equal(one, r); // level 1 == plain bitwise
            }
            if (!ok) {
                std::cout << "MISMATCH " << f << " sequence " << s << "\n";
                ++failures;
            }
        }
        // level 2 again, every way a chunk can be applied (spec[1] runs in AUTO mode): the precomputed handlers (K4, the
        // reference's GrB_mxm per chunk), the chunks recomputed from the sparse factors (K4i), and -- the shipped models
        // all have the chain + specials structure -- the one-launch chain run (K4c): same bits
        if (spec.size() > 1) {
            B200_spec_impl table(hmm, 2), implicit(hmm, 2), chain(hmm, 2);
            table.set_run_mode(SV_SPEC_RUN_TABLE);
            implicit.set_run_mode(SV_SPEC_RUN_IMPLICIT);
            bool has_chain = true;
            try {
                chain.set_run_mode(SV_SPEC_RUN_CHAIN);
            } catch (const std::exception&) {
                has_chain = false; // not a chain + specials model
            }
            for (const auto& sq : sequences) {
                const auto ref2 = table.run_Viterbi_spec(sq);
                if (!harness::bit_equal(ref2, spec[1]->run_Viterbi_spec(sq)) ||
                    !harness::bit_equal(ref2, implicit.run_Viterbi_spec(sq)) ||
                    (has_chain && !harness::bit_equal(ref2, chain.run_Viterbi_spec(sq)))) {
                    std::cout << "MISMATCH between level-2 run modes " << f << "\n";
                    ++failures;
                }
            }
        }
        // smallest model: the level-2 tables survive a save / attach round trip (not in the reference)
        if (&f == &files.front() && spec.size() > 1) {
            const std::string tmp = "/tmp/sv_tables_" + std::to_string(::getpid()) + ".svht";
            spec[1]->save_tables(tmp);
            B200_spec_impl again(2);
            again.spec_with_saved(hmm, tmp);
            std::remove(tmp.c_str());
            for (const auto& sq : sequences)
                if (!harness::bit_equal(spec[1]->run_Viterbi_spec(sq), again.run_Viterbi_spec(sq))) {
                    std::cout << "MISMATCH after save/load " << f << "\n";
                    ++failures;
                }
        }
        std::cout << hmm.states_num << " states: ok (" << spec.size() << " spec levels)" << std::endl;
    }
    std::cout << (failures ? "FAILED" : "PASSED") << ": " << files.size() << " models x " << sequences.size()
              << " sequences\n";
    return failures ? 1 : 0;
}
