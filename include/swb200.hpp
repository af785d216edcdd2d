// swb200.hpp -- C++ host mirror of the reference's Rust API over the C ABI (swb200.h).
//
// The reference is a Rust library (rust/sequence_alignment/src/sequence_alignment.rs = "SA.rs") with two command line
// front ends (SA-main.rs, rust/sa_opencl/src/main.rs = "OCL-main.rs").  No Rust toolchain exists in this image, so the
// host side above the C ABI is written here in C++ with the reference's names, argument meaning and error behaviour:
//
//   swb200::SeqData               SA.rs:799-887   (create, line_to_seq, load_fasta)
//   swb200::SubstitutionMatrix    SA.rs:626-794   (get_blosum62_matrix, get_mat_matrix, load_from_lines, get_score_str)
//   swb200::SequenceAlignment     SA.rs:15-552    (new, align, align_partial(score_only), plus the batch calls)
//   swb200::f32_display           Rust `{}` of an f32 (SA-main.rs:40)
//
// Where the reference panics, these throw swb200::Error (code = SWB_ERR_*).  Header only; link with libswb200.so.
#pragma once
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <map>
#include <sstream>
#include <stdexcept>
#include <string>
#include <tuple>
#include <utility>
#include <vector>

#include "swb200.h"

namespace swb200 {

constexpr int ALIGN_GLOBAL = SWB_GLOBAL, ALIGN_GLOCAL = SWB_GLOCAL, ALIGN_LOCAL = SWB_LOCAL;   // SA.rs:11-13

struct Error : std::runtime_error {
    int code;
    Error(int c, const std::string &m) : std::runtime_error(m), code(c) {}
};

inline const char *type_name(int t)
{
    return t == ALIGN_GLOBAL ? "Global alignment" : (t == ALIGN_GLOCAL ? "Glocal alignment" : "Local alignment");
}

// Rust `{}` formatting of an f32: the shortest decimal that round-trips, positional (never an exponent), no trailing
// ".0" (485, 54.5, -1770.5, 0, 340282350000000000000000000000000000000).
inline std::string f32_display(float v)
{
    if (std::isnan(v)) return "NaN";
    if (std::isinf(v)) return v < 0 ? "-inf" : "inf";
    if (v == 0.0f) return std::signbit(v) ? "-0" : "0";
    char b[64];
    int p = 0;
    for (p = 0; p <= 8; ++p) {                                  // p+1 significant digits
        snprintf(b, sizeof b, "%.*e", p, (double)v);
        if (strtof(b, nullptr) == v) break;
    }
    std::string m(b);                                           // [-]d.ddddde[+-]xx
    const size_t epos = m.find('e');
    const int ex = atoi(m.c_str() + epos + 1);
    std::string digits;
    const bool neg = m[0] == '-';
    for (size_t k = neg ? 1 : 0; k < epos; ++k) if (m[k] != '.') digits.push_back(m[k]);
    while (digits.size() > 1 && digits.back() == '0') digits.pop_back();
    std::string r = neg ? "-" : "";
    const int nd = (int)digits.size();
    if (ex >= nd - 1) r += digits + std::string((size_t)(ex - (nd - 1)), '0');
    else if (ex >= 0) r += digits.substr(0, (size_t)ex + 1) + "." + digits.substr((size_t)ex + 1);
    else r += "0." + std::string((size_t)(-ex - 1), '0') + digits;
    return r;
}

struct SeqData {                                               // SA.rs:799-803 (seq: one byte per residue)
    std::string name, desc, seq;

    static std::string line_to_seq(const std::string &s, bool retainws)        // SA.rs:811-821
    {
        std::string r;
        for (char c : s) {
            if (retainws ? (c == '\r' || c == '\n') : (isspace((unsigned char)c) != 0)) continue;
            r.push_back(c);
        }
        return r;
    }
    static SeqData create(const std::string &n, const std::string &d, const std::string &s)   // SA.rs:808-810
    {
        return SeqData{n, d, line_to_seq(s, true)};
    }
    // SA.rs:830-887: any line holding '>' starts a record; name = first token, desc = what follows the first white
    // space after it; sequence lines lose their white space unless retainws.
    static std::vector<SeqData> load_fasta(const std::string &filename, bool retainws = false)
    {
        std::ifstream f(filename, std::ios::binary);
        if (!f) throw Error(SWB_ERR_ARG, "can not open " + filename);
        std::vector<SeqData> ret;
        std::string line, buf, name, desc;
        bool have = false;
        while (std::getline(f, line)) {
            if (!line.empty() && line.back() == '\r') line.pop_back();
            if (line.find('>') != std::string::npos) {
                if (!buf.empty() || !name.empty()) { ret.push_back(SeqData{name, desc, buf}); buf.clear(); }
                size_t b = 0, e = line.size();
                while (b < e && isspace((unsigned char)line[b])) ++b;
                while (e > b && isspace((unsigned char)line[e - 1])) --e;
                std::string head = line.substr(b, e - b);
                if (!head.empty() && head[0] == '>') head.erase(0, 1);
                size_t p = 0;
                while (p < head.size() && isspace((unsigned char)head[p])) ++p;
                size_t q = p;
                while (q < head.size() && !isspace((unsigned char)head[q])) ++q;
                name = head.substr(p, q - p);
                desc = (q < head.size()) ? head.substr(q + 1) : std::string();
                have = true;
            } else buf += line_to_seq(line, retainws);
        }
        if (!name.empty() || !buf.empty()) ret.push_back(SeqData{name, desc, buf});
        (void)have;
        return ret;
    }
};

class SubstitutionMatrix {                                     // SA.rs:626-631
public:
    std::string letters;                                       // index_to_string, one byte per letter
    std::vector<float> scores;                                 // [a][b], n*n

    SubstitutionMatrix() {}
    SubstitutionMatrix(const std::string &l, const std::vector<float> &s) : letters(l), scores(s)
    {
        if (scores.size() != letters.size() * letters.size()) throw Error(SWB_ERR_ARG, "matrix needs n*n scores");
    }
    static SubstitutionMatrix get_blosum62_matrix()            // SA.rs:697-733 (the table lives in the library)
    {
        char l[32]; std::vector<float> s(24 * 24);
        const int n = swb_matrix_blosum62(l, s.data());
        return SubstitutionMatrix(std::string(l, n), s);
    }
    static SubstitutionMatrix get_mat_matrix(float matchscore, float mismatchscore)   // SA.rs:681-695
    {
        char l[32]; std::vector<float> s(26 * 26);
        const int n = swb_matrix_match_mismatch(matchscore, mismatchscore, l, s.data());
        return SubstitutionMatrix(std::string(l, n), s);
    }
    // SA.rs:735-794: '#' lines are comments, the first other line names the columns, every following line is a row
    static SubstitutionMatrix load_from_lines(const std::vector<std::string> &lines)
    {
        std::vector<std::string> cols;
        std::map<std::pair<std::string, std::string>, float> table;
        int count = -1;
        for (const std::string &line : lines) {
            std::istringstream is(line);
            std::vector<std::string> tok; std::string t;
            while (is >> t) tok.push_back(t);
            if (tok.empty()) throw Error(SWB_ERR_ARG, "empty line in the matrix (the reference panics here)");
            if (tok[0][0] == '#') continue;
            if (++count == 0) {
                for (const std::string &c : tok) {
                    for (const std::string &d : cols) if (c == d) throw Error(SWB_ERR_ARG, c + " was already found.");
                    cols.push_back(c);
                }
                continue;
            }
            bool known = false;
            for (const std::string &c : cols) known = known || c == tok[0];
            if (!known) throw Error(SWB_ERR_ARG, tok[0] + " was not found in the row name.");
            for (size_t k = 1; k < tok.size() && k - 1 < cols.size(); ++k) {
                char *end = nullptr;
                float v = strtof(tok[k].c_str(), &end);
                if (end == tok[k].c_str() || *end) v = 0.0f;   // "can not be parsed! zero was assigned", SA.rs:769-772
                table[{tok[0].substr(0, 1), cols[k - 1]}] = v;
            }
        }
        SubstitutionMatrix m;
        for (const std::string &c : cols) {
            if (c.size() != 1) throw Error(SWB_ERR_ARG, "every matrix letter must be a single byte");
            m.letters += c;
        }
        const size_t n = cols.size();
        m.scores.assign(n * n, 0.0f);
        for (size_t i = 0; i < n; ++i)
            for (size_t j = 0; j < n; ++j) {
                auto it = table.find({cols[i], cols[j]});
                if (it == table.end()) throw Error(SWB_ERR_ARG, "score about " + cols[i] + " " + cols[j] + " is not defined.");
                m.scores[i * n + j] = it->second;
            }
        return m;
    }
    float get_score(size_t a, size_t b) const { return scores.at(a * letters.size() + b); }   // SA.rs:646-648
    float get_score_str(char a, char b) const                                                // SA.rs:633-645
    {
        const size_t ia = letters.find(a), ib = letters.find(b);
        if (ia == std::string::npos) throw Error(SWB_ERR_LETTER, std::string(1, a) + " was not found in scoring matrix!");
        if (ib == std::string::npos) throw Error(SWB_ERR_LETTER, std::string(1, b) + " was not found in scoring matrix!");
        return get_score(ia, ib);
    }
};

struct Alignment {                                             // align_partial's (Vec<String>, Vec<String>, f32) + the lists
    std::string aln_a, aln_b;
    float score = 0;
    int64_t end_a = -1, end_b = -1;
    std::vector<int64_t> idx_a, idx_b;
};

class SequenceAlignment {                                      // SA.rs:15-552 / OCL.rs:71-918
public:
    int alignment_type;

    SequenceAlignment(const SubstitutionMatrix &ss, float go, float ge, int alignment_type_, int device = 0)
        : alignment_type(alignment_type_), m_(ss)
    {
        swb_matrix m; m.n_letters = (int32_t)m_.letters.size(); m.letters = m_.letters.data(); m.scores = m_.scores.data();
        const int rc = swb_engine_create(&m, go, ge, alignment_type, device, &h_);
        if (rc != SWB_OK) { const char *msg = swb_last_error(nullptr); throw Error(rc, msg ? msg : "swb_engine_create failed"); }
    }
    ~SequenceAlignment() { if (h_) swb_engine_destroy(h_); }
    SequenceAlignment(const SequenceAlignment &) = delete;
    SequenceAlignment &operator=(const SequenceAlignment &) = delete;

    // SA.rs:389-392.  (The reference returns Vec<String> of one-letter strings; here one std::string per sequence.)
    std::tuple<std::string, std::string, float> align(const SeqData &s1, const SeqData &s2, bool retain_all)
    {
        Alignment r = align_full(s1.seq, s2.seq, retain_all);
        return {r.aln_a, r.aln_b, r.score};
    }
    // SA.rs:393-551: partial_region = the DP runs on the prefixes (SA.rs:396-405) while the retain_all padding spells the
    // whole sequences (SA.rs:512-549); score_only = no traceback, empty strings (SA.rs:438-467).
    std::tuple<std::string, std::string, float> align_partial(const SeqData &s1, const SeqData &s2, bool retain_all,
                                                              const std::pair<size_t, size_t> *partial_region, bool score_only)
    {
        const std::string &a = s1.seq, &b = s2.seq;
        if (partial_region && (partial_region->first > a.size() || partial_region->second > b.size()))
            throw Error(SWB_ERR_ARG, "partial_region exceeds the sequences");
        const int64_t pa = partial_region ? (int64_t)partial_region->first : (int64_t)a.size();
        const int64_t pb = partial_region ? (int64_t)partial_region->second : (int64_t)b.size();
        swb_alignment out;
        check(swb_align_partial(h_, a.data(), (int64_t)a.size(), b.data(), (int64_t)b.size(), pa, pb, retain_all ? 1 : 0, score_only ? 1 : 0, &out));
        Alignment r = take(out);
        swb_alignment_free(&out);
        return {r.aln_a, r.aln_b, r.score};
    }
    Alignment align_full(const std::string &a, const std::string &b, bool retain_all)
    {
        swb_alignment out;
        check(swb_align(h_, a.data(), (int64_t)a.size(), b.data(), (int64_t)b.size(), retain_all ? 1 : 0, &out));
        Alignment r = take(out);
        swb_alignment_free(&out);
        return r;
    }
    // The serial loops of the two CLIs as one device batch each (SA-main.rs:35-45, OCL-main.rs:295-316).
    std::vector<Alignment> align_list(const std::vector<const SeqData *> &sa, const std::vector<const SeqData *> &sb, bool retain_all)
    {
        if (sa.size() != sb.size()) throw Error(SWB_ERR_ARG, "align_list needs as many a as b");
        std::string a, b; std::vector<int64_t> oa, ob;
        pack(sa, a, oa); pack(sb, b, ob);
        std::vector<swb_alignment> outs(sa.size() ? sa.size() : 1);
        check(swb_align_batch(h_, a.data(), oa.data(), b.data(), ob.data(), (int64_t)sa.size(), retain_all ? SWB_RETAIN_ALL : 0u, outs.data()));
        std::vector<Alignment> r;
        for (size_t k = 0; k < sa.size(); ++k) { r.push_back(take(outs[k])); swb_alignment_free(&outs[k]); }
        return r;
    }
    // Scores + end cells of n pairs (align_partial(.., score_only) per pair, SA.rs:438-467).
    void score_batch(const std::vector<const SeqData *> &sa, const std::vector<const SeqData *> &sb, std::vector<float> &scores,
                     std::vector<int64_t> &end_a, std::vector<int64_t> &end_b)
    {
        if (sa.size() != sb.size()) throw Error(SWB_ERR_ARG, "score_batch needs as many a as b");
        std::string a, b; std::vector<int64_t> oa, ob;
        pack(sa, a, oa); pack(sb, b, ob);
        scores.assign(sa.size(), 0.0f); end_a.assign(sa.size(), -1); end_b.assign(sa.size(), -1);
        check(swb_score_batch(h_, a.data(), oa.data(), b.data(), ob.data(), (int64_t)sa.size(), scores.data(), end_a.data(), end_b.data()));
    }
    // The clustering loop of OCL-main.rs:149-243; arrays indexed by sorted position (see swb_cluster).
    void cluster(const std::vector<SeqData> &seqs, double identity, double coverage_short, double coverage_long,
                 std::vector<int64_t> &order, std::vector<int64_t> &cluster_of, std::vector<int64_t> &identical)
    {
        std::string s; std::vector<int64_t> off(1, 0);
        for (const SeqData &q : seqs) { s += q.seq; off.push_back((int64_t)s.size()); }
        const size_t n = seqs.size();
        order.assign(n, 0); cluster_of.assign(n, 0); identical.assign(n, 0);
        check(swb_cluster(h_, s.data(), off.data(), (int64_t)n, identity, coverage_short, coverage_long, order.data(), cluster_of.data(),
                          identical.data(), nullptr));
    }
    swb_engine *handle() { return h_; }

private:
    static void pack(const std::vector<const SeqData *> &v, std::string &s, std::vector<int64_t> &off)
    {
        off.assign(1, 0);
        for (const SeqData *q : v) { s += q->seq; off.push_back((int64_t)s.size()); }
    }
    static Alignment take(const swb_alignment &o)
    {
        Alignment r;
        r.score = o.score; r.end_a = o.end_a; r.end_b = o.end_b;
        if (o.aln_a) r.aln_a.assign(o.aln_a, (size_t)o.aln_len);
        if (o.aln_b) r.aln_b.assign(o.aln_b, (size_t)o.aln_len);
        if (o.n_cols > 0) { r.idx_a.assign(o.idx_a, o.idx_a + o.n_cols); r.idx_b.assign(o.idx_b, o.idx_b + o.n_cols); }
        return r;
    }
    void check(int rc)
    {
        if (rc != SWB_OK) { const char *msg = swb_last_error(h_); throw Error(rc, msg ? msg : "swb200 error"); }
    }
    SubstitutionMatrix m_;
    swb_engine *h_ = nullptr;
};

// EMBOSS `srspair` report body (format of test/emboss_results/*.dat, res0.dat:17-45): the text from the
// first `#====` rule to the closing `#----` rules.  Markup: `|` identical, `:` substitution score > 0, `.` other residue
// pairs, blank for gap columns; a block that holds residues shows the numbers of its first and last residue, a block of
// gaps only repeats the count so far.  start_a / start_b = 1-based number of the first residue of each row.
// (Same rules as smithwaterman_b200/report.py, which is pinned to all 300 golden reports.)
inline std::string srspair_body(const std::string &aln_a, const std::string &aln_b, float score, const SubstitutionMatrix &m,
                                float gap_open, float gap_extend, int64_t start_a = 1, int64_t start_b = 1,
                                const std::string &name_a = "s1", const std::string &name_b = "s2",
                                const std::string &matrix_name = "EBLOSUM62")
{
    if (aln_a.size() != aln_b.size()) throw Error(SWB_ERR_ARG, "aligned strings differ in length");
    const size_t n = aln_a.size(), nl = m.letters.size();
    auto index_of = [&](char c) { size_t k = m.letters.find(c); if (k == std::string::npos) k = m.letters.find('X'); return k; };
    std::string mk(n, ' ');
    int64_t ident = 0, simil = 0, gaps = 0;
    for (size_t k = 0; k < n; ++k) {
        const char x = aln_a[k], y = aln_b[k];
        if (x == '-' || y == '-') { ++gaps; continue; }
        if (x == y) { mk[k] = '|'; ++ident; ++simil; continue; }
        const size_t i = index_of(x), j = index_of(y);
        const float sc = (i != std::string::npos && j != std::string::npos) ? m.scores[i * nl + j] : 0.0f;
        if (sc > 0) { mk[k] = ':'; ++simil; } else mk[k] = '.';
    }
    char buf[256];
    std::string out = "#=======================================\n#\n# Aligned_sequences: 2\n";
    out += "# 1: " + name_a + "\n# 2: " + name_b + "\n# Matrix: " + matrix_name + "\n";
    snprintf(buf, sizeof buf, "# Gap_penalty: %.1f\n# Extend_penalty: %.1f\n#\n# Length: %zu\n", std::fabs(gap_open), std::fabs(gap_extend), n);
    out += buf;
    const double d = n ? 100.0 / (double)n : 0.0;
    snprintf(buf, sizeof buf, "# Identity:   %5lld/%zu (%4.1f%%)\n# Similarity: %5lld/%zu (%4.1f%%)\n# Gaps:       %5lld/%zu (%4.1f%%)\n",
             (long long)ident, n, ident * d, (long long)simil, n, simil * d, (long long)gaps, n, gaps * d);
    out += buf;
    snprintf(buf, sizeof buf, "# Score: %.1f\n# \n#\n#=======================================\n\n", score);
    out += buf;
    int64_t done_a = start_a - 1, done_b = start_b - 1;
    for (size_t lo = 0; lo < n; lo += 50) {
        const std::string ra = aln_a.substr(lo, 50), rb = aln_b.substr(lo, 50), rm = mk.substr(lo, 50);
        const int64_t na = (int64_t)ra.size() - std::count(ra.begin(), ra.end(), '-');
        const int64_t nb = (int64_t)rb.size() - std::count(rb.begin(), rb.end(), '-');
        snprintf(buf, sizeof buf, "%-13.13s %6lld %s %6lld\n", name_a.c_str(), (long long)(na ? done_a + 1 : done_a), ra.c_str(), (long long)(done_a + na));
        out += buf;
        snprintf(buf, sizeof buf, "%-13s %6s %s\n", "", "", rm.c_str());
        out += buf;
        snprintf(buf, sizeof buf, "%-13.13s %6lld %s %6lld\n\n", name_b.c_str(), (long long)(nb ? done_b + 1 : done_b), rb.c_str(), (long long)(done_b + nb));
        out += buf;
        done_a += na; done_b += nb;
    }
    out += "\n#---------------------------------------\n#---------------------------------------\n";
    return out;
}

}  // namespace swb200
