// The reference's own known-answer tests (SA.rs:921-1221 / OCL.rs:1292-1448) through the C++ host mirror
// include/swb200.hpp, written the way the Rust tests are: build the matrix, `SequenceAlignment::new`, `align`, compare
// score and strings.  The cases come as tab-separated lines (made from tests/golden/rust_kats.json by the pytest that
// compiles and runs this):  test  mode  go  ge  retain_all  matrix  match  mismatch  scale  seq1  seq2  score  r1  r2
// ("-" in score / r1 marks "not asserted by the reference test"; partial = prefix lengths "p1,p2" or "-").
#include <cstdio>
#include <fstream>
#include <iostream>
#include <sstream>

#include "swb200.hpp"

using namespace swb200;

int main(int argc, char **argv)
{
    if (argc < 2) { fprintf(stderr, "usage: kats <cases.tsv>\n"); return 2; }
    std::ifstream f(argv[1]);
    std::string line;
    int done = 0, bad = 0;
    while (std::getline(f, line)) {
        std::vector<std::string> c; std::stringstream ss(line); std::string t;
        while (std::getline(ss, t, '\t')) c.push_back(t);
        if (c.size() < 14) continue;
        try {
            SubstitutionMatrix sm = c[5] == "blosum62" ? SubstitutionMatrix::get_blosum62_matrix()
                                                       : SubstitutionMatrix::get_mat_matrix(strtof(c[6].c_str(), nullptr), strtof(c[7].c_str(), nullptr));
            const float scale = strtof(c[8].c_str(), nullptr);
            for (float &v : sm.scores) v *= scale;
            SequenceAlignment sw(sm, strtof(c[2].c_str(), nullptr), strtof(c[3].c_str(), nullptr), atoi(c[1].c_str()));
            const SeqData s1 = SeqData::create("", "", c[9]), s2 = SeqData::create("", "", c[10]);
            auto [r1, r2, score] = sw.align(s1, s2, c[4] == "1");
            bool ok = true;
            if (c[11] != "-" && f32_display(score) != c[11]) ok = false;
            if (c[12] != "-" && (r1 != c[12] || r2 != c[13])) ok = false;
            // score_only agrees with the full alignment (align_partial, SA.rs:393-467)
            auto [e1, e2, sc2] = sw.align_partial(s1, s2, false, nullptr, true);
            if (sc2 != score || !e1.empty() || !e2.empty()) ok = false;
            if (!ok) { ++bad; std::cout << "FAIL " << c[0] << ": score " << f32_display(score) << " want " << c[11] << "\n" << r1 << "\n" << r2 << "\n"; }
            ++done;
        } catch (const Error &e) {
            ++bad; std::cout << "ERROR " << c[0] << ": " << e.what() << "\n";
        }
    }
    // error behaviour: an unknown letter without an X row is an error, not a crash (SA.rs:669-679)
    try {
        SubstitutionMatrix dna("ACGT", std::vector<float>{1, -1, -1, -1, -1, 1, -1, -1, -1, -1, 1, -1, -1, -1, -1, 1});
        SequenceAlignment sw(dna, 2.0f, 1.0f, ALIGN_LOCAL);
        sw.align(SeqData::create("", "", "ACGN"), SeqData::create("", "", "ACGT"), true);
        ++bad; std::cout << "FAIL unknown letter accepted\n";
    } catch (const Error &e) {
        if (e.code != SWB_ERR_LETTER) { ++bad; std::cout << "FAIL unknown letter: code " << e.code << "\n"; }
    }
    std::cout << "kats: " << done << " cases, " << bad << " failures\n";
    return bad ? 1 : 0;
}
