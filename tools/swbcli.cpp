// swbcli -- the reference's two command lines over the CUDA engine, in C++ (the reference's are Rust binaries):
//
//   swbcli [-global|-glocal|-local] <fasta1> <fasta2>                 seqalign   (SA-main.rs:10-46)
//   swbcli [-mode] -list <list file> [-out F]                          sa_opencl  (OCL-main.rs:267-316)
//   swbcli -cluster [-mode] [-identity x] [-coverage_short x] [-coverage_long x] -out F <fasta>
//                                                                      sa_opencl  (OCL-main.rs:149-265)
//
// Same flags and record format (`#score:` / `#type:` / `>name` / aligned string, SA-main.rs:40-43), BLOSUM62 with
// open 10 / extend 0.5 as both reference binaries hard-code (SA-main.rs:34, OCL-main.rs:147).  All pairs of a run go to
// the device in ONE batch; the clustering loop runs inside the library (swb_cluster).
//
// build:  g++ -O2 -std=c++17 -Iinclude tools/swbcli.cpp -Lsmithwaterman_b200 -lswb200 -Wl,-rpath,'$ORIGIN' -o smithwaterman_b200/swbcli
#include <cstdio>
#include <fstream>
#include <iostream>
#include <string>
#include <vector>

#include "swb200.hpp"

using namespace swb200;

static const char *USAGE =
    "usage: sa_opencl [(-global|-glocal|-local(default))] <infile1 (fasta file)>  <infile2 (fasta file)>\n"
    "usage: sa_opencl [(-global|-glocal|-local(default))] [-list] <list file>\n"
    "usage: sa_opencl -cluster[ing] [(-global|-glocal|-local(default))] [-identity 0.0-1.0] [-coverage_short 0.0-1.0] "
    "[-coverage_long 0.0-1.0] -out <output file> <fasta file>\n";

struct Options {                                               // AlignmentOptions::parse, OCL-main.rs:33-116
    int alignment_type = ALIGN_LOCAL;
    bool list = false, clustering = false;
    double identity = 0.8, coverage_short = 0.8, coverage_long = 0.8;
    std::string outfilename, file1, file2;

    explicit Options(const std::vector<std::string> &args)
    {
        std::vector<std::string> files;
        for (size_t k = 0; k < args.size(); ++k) {
            const std::string &a = args[k];
            if (a == "-glocal") alignment_type = ALIGN_GLOCAL;
            else if (a == "-global") alignment_type = ALIGN_GLOBAL;
            else if (a == "-local") alignment_type = ALIGN_LOCAL;
            else if (a == "-list") list = true;
            else if (a == "-cluster" || a == "-clustering") clustering = true;
            else if (a == "-coverage_short" || a == "-coverage_long" || a == "-identity" || a == "-out") {
                if (k + 1 >= args.size()) throw Error(SWB_ERR_ARG, "option " + a + " needs a value");
                const std::string &v = args[++k];
                if (a == "-out") outfilename = v;
                else {
                    char *end = nullptr;
                    const double x = strtod(v.c_str(), &end);
                    if (end == v.c_str() || *end) throw Error(SWB_ERR_ARG, "can not parse " + v);
                    (a == "-identity" ? identity : (a == "-coverage_short" ? coverage_short : coverage_long)) = x;
                }
            } else if (!a.empty() && a[0] == '-') throw Error(SWB_ERR_ARG, "Unknown option " + a);
            else files.push_back(a);
        }
        if (!clustering && !list) {
            if (files.size() != 2) throw Error(SWB_ERR_ARG, "2 files must be provided.");
            file1 = files[0]; file2 = files[1];
        } else {
            if (clustering && list) throw Error(SWB_ERR_ARG, "Incompatible option -list & -cluster(ing)");
            if (files.size() != 1) throw Error(SWB_ERR_ARG, "1 file must be provided.");
            file1 = files[0];
            if (clustering && outfilename.empty()) throw Error(SWB_ERR_ARG, "Clustering must have -out.");
        }
    }
};

// The list-file reader of OCL-main.rs:268-288: tab (else blank) separated pairs of FASTA paths.
static std::vector<std::pair<std::string, std::string>> read_list(const std::string &path)
{
    std::ifstream f(path);
    if (!f) throw Error(SWB_ERR_ARG, "can not open " + path);
    std::vector<std::pair<std::string, std::string>> pairs;
    std::string line;
    while (std::getline(f, line)) {
        if (!line.empty() && line.back() == '\r') line.pop_back();
        auto split = [&](char sep) {
            std::vector<std::string> out; size_t p = 0;
            for (;;) { size_t q = line.find(sep, p); out.push_back(line.substr(p, q - p)); if (q == std::string::npos) break; p = q + 1; }
            return out;
        };
        std::vector<std::string> spp = split('\t');
        if (spp.size() == 1) spp = split(' ');
        if (spp.size() > 2) std::cout << line << " \n^ Only " << spp[0] << " " << spp[1] << " are used.\n";   // (and, like the reference, not queued)
        else if (spp.size() < 2) std::cout << line << " \n is ignoed.\n";
        else pairs.push_back({spp[0], spp[1]});
    }
    return pairs;
}

int main(int argc, char **argv)
{
    std::vector<std::string> args(argv + 1, argv + argc);
    if (args.size() < 2) { fputs(USAGE, stderr); return 255; }
    try {
        // host-mirror checks that need no device (used by the CPU tests): the FASTA reader and the f32 formatter
        if (args[0] == "-dumpfasta") {
            for (const SeqData &s : SeqData::load_fasta(args[1], args.size() > 2 && args[2] == "retainws"))
                std::cout << "name=" << s.name << "|desc=" << s.desc << "|seq=" << s.seq << "\n";
            return 0;
        }
        if (args[0] == "-f32") {
            for (size_t k = 1; k < args.size(); ++k) std::cout << f32_display(strtof(args[k].c_str(), nullptr)) << "\n";
            return 0;
        }
        // EMBOSS srspair reports:  swbcli -emboss -local|-glocal|-global <file_list> <outdir>
        // writes res<k>.dat / needle_glocal_res<k>.dat / needle_res<k>.dat, diffable against test/emboss_results.
        if (args[0] == "-emboss") {
            if (args.size() != 4) throw Error(SWB_ERR_ARG, "usage: swbcli -emboss -local|-glocal|-global <file_list> <outdir>");
            const int type = args[1] == "-local" ? ALIGN_LOCAL : args[1] == "-glocal" ? ALIGN_GLOCAL : args[1] == "-global" ? ALIGN_GLOBAL : -1;
            if (type < 0) throw Error(SWB_ERR_ARG, "unknown alignment type " + args[1]);
            const char *pre = type == ALIGN_LOCAL ? "res" : type == ALIGN_GLOCAL ? "needle_glocal_res" : "needle_res";
            const SubstitutionMatrix bl = SubstitutionMatrix::get_blosum62_matrix();
            SequenceAlignment sw(bl, 10.0f, 0.5f, type);
            std::vector<std::vector<SeqData>> keep;
            for (const auto &fp : read_list(args[2])) { keep.push_back(SeqData::load_fasta(fp.first, false)); keep.push_back(SeqData::load_fasta(fp.second, false)); }
            std::vector<const SeqData *> sa, sb;
            for (size_t k = 0; k + 1 < keep.size(); k += 2) {
                if (keep[k].empty() || keep[k + 1].empty()) throw Error(SWB_ERR_ARG, "empty FASTA file in the list");
                sa.push_back(&keep[k][0]); sb.push_back(&keep[k + 1][0]);
            }
            std::vector<Alignment> res = sw.align_list(sa, sb, type != ALIGN_LOCAL);
            for (size_t k = 0; k < res.size(); ++k) {
                int64_t s1 = 1, s2 = 1;
                if (type == ALIGN_LOCAL) {
                    for (int64_t v : res[k].idx_a) if (v >= 0) { s1 = v + 1; break; }
                    for (int64_t v : res[k].idx_b) if (v >= 0) { s2 = v + 1; break; }
                }
                std::ofstream fo(args[3] + "/" + pre + std::to_string(k) + ".dat");
                if (!fo) throw Error(SWB_ERR_ARG, "can not write into " + args[3]);
                fo << srspair_body(res[k].aln_a, res[k].aln_b, res[k].score, bl, 10.0f, 0.5f, s1, s2);
            }
            std::cout << "wrote " << res.size() << " reports to " << args[3] << "\n";
            return 0;
        }
        Options opp(args);
        SequenceAlignment sw(SubstitutionMatrix::get_blosum62_matrix(), 10.0f, 0.5f, opp.alignment_type);
        if (opp.clustering) {
            std::vector<SeqData> seqs = SeqData::load_fasta(opp.file1, false);
            std::vector<int64_t> order, cluster_of, identical;
            sw.cluster(seqs, opp.identity, opp.coverage_short, opp.coverage_long, order, cluster_of, identical);
            const size_t n = seqs.size();
            for (size_t p = 0; p < n; ++p)                                    // progress lines of OCL-main.rs:240-242
                if (cluster_of[p] == (int64_t)p && identical[p] == (int64_t)p && (p + 1) % 10 == 0) std::cout << (p + 1) << " sequences were processed.\n";
            // member lists in the order the reference appends them (OCL-main.rs:187-233): a sequence is followed by its exact copies
            std::vector<std::vector<size_t>> copies(n), members(n);
            for (size_t p = 0; p < n; ++p) copies[(size_t)identical[p]].push_back(p);
            for (size_t p = 0; p < n; ++p)
                if (identical[p] == (int64_t)p) members[(size_t)cluster_of[p]].insert(members[(size_t)cluster_of[p]].end(), copies[p].begin(), copies[p].end());
            std::ofstream fa(opp.outfilename), fc(opp.outfilename + ".clstr");            // OCL-main.rs:244-265
            if (!fa || !fc) throw Error(SWB_ERR_ARG, "can not write " + opp.outfilename);
            for (size_t p = 0; p < n; ++p)
                if (cluster_of[p] == (int64_t)p) { const SeqData &s = seqs[(size_t)order[p]]; fa << ">" << s.name << " " << s.desc << "\n" << s.seq << "\n"; }
            for (size_t p = 0; p < n; ++p) {
                if (members[p].empty()) continue;
                for (size_t k = 0; k < members[p].size(); ++k) fc << (k ? " " : "") << seqs[(size_t)order[members[p][k]]].name;
                fc << "\n";
            }
            return 0;
        }
        std::vector<std::pair<std::string, std::string>> filelist;
        if (opp.list) filelist = read_list(opp.file1); else filelist.push_back({opp.file1, opp.file2});
        std::vector<std::vector<SeqData>> keep;                              // all-vs-all of every file pair, ONE device batch
        std::vector<const SeqData *> sa, sb;
        for (const auto &fp : filelist) {
            keep.push_back(SeqData::load_fasta(fp.first, false)); keep.push_back(SeqData::load_fasta(fp.second, false));
        }
        for (size_t k = 0; k + 1 < keep.size(); k += 2)
            for (const SeqData &x : keep[k]) for (const SeqData &y : keep[k + 1]) { sa.push_back(&x); sb.push_back(&y); }
        std::vector<Alignment> res = sw.align_list(sa, sb, true);
        std::ofstream fo;
        if (!opp.outfilename.empty()) { fo.open(opp.outfilename); if (!fo) throw Error(SWB_ERR_ARG, "can not write " + opp.outfilename); }
        for (size_t k = 0; k < res.size(); ++k) {
            if (fo.is_open())                                                // OCL-main.rs:303-307 (no newline after #score / #type there)
                fo << "#score:" << f32_display(res[k].score) << "#type:" << type_name(opp.alignment_type) << ">" << sa[k]->name << "\n"
                   << res[k].aln_a << "\n>" << sb[k]->name << "\n" << res[k].aln_b << "\n";
            else
                std::cout << "#score:" << f32_display(res[k].score) << "\n#type:" << type_name(opp.alignment_type) << "\n>" << sa[k]->name << "\n"
                          << res[k].aln_a << "\n\n>" << sb[k]->name << "\n" << res[k].aln_b << "\n\n";
        }
        return 0;
    } catch (const Error &e) {
        fprintf(stderr, "error %d: %s\n", e.code, e.what());
        return 1;
    }
}
