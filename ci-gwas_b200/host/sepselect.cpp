// sepselect.cpp -- host side of the separating-set selection / v-structure orientation step on a merged skeleton
// (SURVEY.md section 8f, row 2): `ci-gwas.py sepselect` and `orient-v-structs` (ci-gwas.py:467-476), i.e.
//
//   MergedCuskResults.__init__ / rm_collinear_markers      cusk_postprocessing/sepselect.py:429-483
//   get_unshielded_triples, get_rfci_relevant_...          sepselect.py:68-93, :139-160
//   find_maximal_and_min_pcorr_sepsets_incr                sepselect.py:262-329   -> the GPU (csrc/sepselect.cu)
//   mark_ambiguous_triples                                 sepselect.py:95-107
//   orient_v_structures                                    sepselect.py:485-511
//   to_file / max_sepset_to_file                           sepselect.py:545-573
//
// The reference turns the two MatrixMarket files into dense num_var^2 arrays (6.4 GB at 20 000 markers) and keeps
// Python sets of index tuples.  Here the adjacency stays a sorted neighbour list per row, the correlations a
// coordinate list, only rows 0..P of the correlations are densified (that is all the search reads), and unshielded
// triples are enumerated on the fly, restricted to those that can matter (middle variable <= P, at most one outer
// variable > P).  The greedy search itself, all the arithmetic of the step, runs on the device; there is no CPU path.
//
// Reproduced on purpose (they decide what is written): is_marker(ix) = ix > num_phen, so the FIRST marker counts as a
// trait in the triple filters while trait_neighbors uses ix < num_phen (sepselect.py:538-543); the "minimal partial
// correlation" set of a pair is the same list object as its maximal set (sepselect.py:289), so a triple is ambiguous
// iff its middle trait is in the maximal set of a pair that never saw its statistic rise after separation.
// Differences: triples in .ut / .atr are sorted (the reference writes them in Python set order); the .mtx files are
// written in "general" form with shortest round-trip values (scipy's writer picks symmetric/general itself); a step
// whose candidates are all singular ends that pair's search (the reference exits); no pair with a minimum gives
// level 0 instead of Python's ValueError.
#include <algorithm>
#include <array>
#include <charconv>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <iostream>
#include <map>

#include "host.hpp"

namespace cuskhost {
namespace {

struct Coo {
    long rows = 0, cols = 0;
    std::vector<int> r, c;  // 0-based
    std::vector<double> v;
};

// scipy.io.mmread(...).toarray() for the coordinate files of this pipeline: general or symmetric, integer / real /
// pattern; duplicates add up
Coo read_mtx(const std::string &path)
{
    check_path(path);
    const std::string txt = slurp(path);
    const char *p = txt.data(), *end = p + txt.size();
    auto next_line = [&](const char *&b, const char *&e) {
        if (p >= end) return false;
        b = p;
        e = (const char *)memchr(p, '\n', (size_t)(end - p));
        if (!e) e = end;
        p = e < end ? e + 1 : end;
        return true;
    };
    const char *b, *e;
    if (!next_line(b, e) || (size_t)(e - b) < 14 || strncmp(b, "%%MatrixMarket", 14) != 0) fail("not a MatrixMarket file: " + path);
    std::vector<std::string> h = split_ws(b, e);
    for (auto &s : h)
        for (auto &ch : s) ch = (char)tolower((unsigned char)ch);
    if (h.size() < 5 || h[1] != "matrix" || h[2] != "coordinate") fail("only coordinate MatrixMarket files are supported: " + path);
    const bool pattern = h[3] == "pattern", symmetric = h[4] == "symmetric";
    if (h[4] != "general" && !symmetric) fail("unsupported MatrixMarket symmetry in " + path);
    Coo m;
    long nnz = -1;
    while (next_line(b, e)) {
        if (b < e && *b == '%') continue;
        const std::vector<std::string> f = split_ws(b, e);
        if (f.empty()) continue;
        if (f.size() < 3) fail("malformed size line in " + path);
        m.rows = std::stol(f[0]);
        m.cols = std::stol(f[1]);
        nnz = std::stol(f[2]);
        break;
    }
    if (nnz < 0) fail("no size line in " + path);
    m.r.reserve((size_t)nnz);
    m.c.reserve((size_t)nnz);
    m.v.reserve((size_t)nnz);
    // entries: "i j [v]"; parsed in place
    const char *q = p;
    auto skip = [&]() {
        while (q < end && (*q == ' ' || *q == '\t' || *q == '\n' || *q == '\r')) ++q;
    };
    for (long k = 0; k < nnz; k++) {
        long i = 0, j = 0;
        double v = 1.0;
        skip();
        while (q < end && *q == '%') {  // comment line inside the body
            while (q < end && *q != '\n') ++q;
            skip();
        }
        auto r1 = std::from_chars(q, end, i);
        if (r1.ec != std::errc()) fail("malformed entry in " + path);
        q = r1.ptr;
        skip();
        auto r2 = std::from_chars(q, end, j);
        if (r2.ec != std::errc()) fail("malformed entry in " + path);
        q = r2.ptr;
        if (!pattern) {
            skip();
            const char *t = q;
            while (t < end && *t != ' ' && *t != '\t' && *t != '\n' && *t != '\r') ++t;
            const std::string tok(q, t);  // strtod: also takes nan / inf / 1E-3
            char *ep = nullptr;
            v = strtod(tok.c_str(), &ep);
            if (ep == tok.c_str()) fail("malformed value in " + path);
            q = t;
        }
        if (i < 1 || j < 1 || i > m.rows || j > m.cols) fail("index out of bounds in " + path);  // scipy raises too
        m.r.push_back((int)i - 1);
        m.c.push_back((int)j - 1);
        m.v.push_back(v);
        if (symmetric && i != j) {
            m.r.push_back((int)j - 1);
            m.c.push_back((int)i - 1);
            m.v.push_back(v);
        }
    }
    // row-major order, duplicates summed
    std::vector<size_t> ord(m.r.size());
    for (size_t k = 0; k < ord.size(); k++) ord[k] = k;
    std::sort(ord.begin(), ord.end(), [&](size_t a, size_t c) { return m.r[a] != m.r[c] ? m.r[a] < m.r[c] : (m.c[a] != m.c[c] ? m.c[a] < m.c[c] : a < c); });
    Coo s;
    s.rows = m.rows;
    s.cols = m.cols;
    for (size_t k : ord) {
        if (!s.r.empty() && s.r.back() == m.r[k] && s.c.back() == m.c[k]) {
            s.v.back() += m.v[k];
        } else {
            s.r.push_back(m.r[k]);
            s.c.push_back(m.c[k]);
            s.v.push_back(m.v[k]);
        }
    }
    return s;
}

void append_num(std::string &out, long long v)
{
    char buf[24];
    out.append(buf, std::to_chars(buf, buf + sizeof buf, v).ptr);
}
void append_num(std::string &out, int v) { append_num(out, (long long)v); }
void append_num(std::string &out, double v)
{
    if (std::isnan(v)) {
        out += "nan";
        return;
    }
    if (std::isinf(v)) {
        out += v < 0 ? "-inf" : "inf";
        return;
    }
    char buf[40];
    out.append(buf, std::to_chars(buf, buf + sizeof buf, v).ptr);
}

template <typename V>
void write_mtx(const std::string &path, const char *field, long n, const std::vector<int> &r, const std::vector<int> &c,
               const std::vector<V> &v)
{
    std::string out = std::string("%%MatrixMarket matrix coordinate ") + field + " general\n%\n";
    append_num(out, (long long)n);
    out += ' ';
    append_num(out, (long long)n);
    out += ' ';
    append_num(out, (long long)r.size());
    out += '\n';
    for (size_t k = 0; k < r.size(); k++) {
        append_num(out, (long long)r[k] + 1);
        out += ' ';
        append_num(out, (long long)c[k] + 1);
        out += ' ';
        append_num(out, v[k]);
        out += '\n';
    }
    FILE *f = fopen(path.c_str(), "wb");
    if (!f || fwrite(out.data(), 1, out.size(), f) != out.size() || fclose(f) != 0) fail("cannot write " + path);
}

struct Problem {
    long n = 0, P = 0, max_level_in = 0;
    std::vector<std::vector<int>> nb;  // neighbors(i): columns of row i of the adjacency, ascending (sepselect.py:133)
    std::vector<int> cr, cc;           // correlations: coordinate list, row-major, unit diagonal
    std::vector<double> cv;
    std::vector<double> Ct;  // rows 0..P dense, [(P+1) x n]
    size_t removed = 0;

    bool has(int a, int b) const { return std::binary_search(nb[(size_t)a].begin(), nb[(size_t)a].end(), b); }
    bool adjacent(int a, int b) const { return has(a, b) || has(b, a); }  // sepselect.py:136
    bool is_marker(int ix) const { return ix > P; }                       // sepselect.py:542-543

    // every unshielded triple (x, y, z) of get_unshielded_triples (sepselect.py:139-153) whose middle variable is
    // <= P and that has at most one outer variable > P; a triple may be reported more than once
    template <typename F>
    void for_each_relevant_triple(F &&f) const
    {
        const int lim = (int)std::min(n - 1, P);
        for (int a = 0; a <= lim; a++)  // (b, a, c): both neighbours of a
            for (int b : nb[(size_t)a])
                for (int c : nb[(size_t)a])
                    if (b != c && (b <= P || c <= P) && !adjacent(b, c)) f(b, a, c);
        for (int a = 0; a < (int)n; a++)  // (a, b, c): c a neighbour of b
            for (int b : nb[(size_t)a]) {
                if (b > P) break;  // ascending
                for (int c : nb[(size_t)b])
                    if (c != a && (a <= P || c <= P) && !adjacent(a, c)) f(a, b, c);
            }
    }
};

Problem load_problem(const std::string &stem)
{
    Problem pr;
    {  // sepselect.py:431-434
        check_path(stem + ".mdim");
        const std::string txt = slurp(stem + ".mdim");
        const std::vector<std::string> f = split_ws(txt.data(), txt.data() + txt.size());
        if (f.size() < 3) fail("malformed .mdim file: " + stem + ".mdim");
        pr.n = std::stol(f[0]);
        pr.P = std::stol(f[1]);
        pr.max_level_in = std::stol(f[2]);
    }
    if (pr.P < 1 || pr.n <= pr.P) fail("sepselect: needs at least one trait and one marker");
    if (pr.P >= CUSK_SS_MAXP) fail("sepselect: more than " + std::to_string(CUSK_SS_MAXP - 1) + " traits are not supported");
    Coo adj = read_mtx(stem + "_sam.mtx");
    Coo cor = read_mtx(stem + "_scm.mtx");
    if (adj.rows != adj.cols || cor.rows != cor.cols || adj.rows != cor.rows) fail("sepselect: matrix files disagree on the dimension");
    long n = adj.rows;  // the arrays are as large as the files say (mmread); .mdim's num_var must agree
    if (n != pr.n) fail("sepselect: .mdim and the matrix files disagree on the number of variables");
    const long P = pr.P;
    // np.fill_diagonal(corr, 1.0), then rm_collinear_markers (sepselect.py:469-483): going through the markers in
    // order, one whose row holds a second exact 1 in a column that is still there is deleted
    std::vector<std::vector<int>> ones((size_t)n);
    for (size_t k = 0; k < cor.r.size(); k++)
        if (cor.r[k] != cor.c[k] && cor.v[k] == 1.0) ones[(size_t)cor.r[k]].push_back(cor.c[k]);
    std::vector<uint8_t> gone((size_t)n, 0);
    for (long i = P; i < n; i++)
        for (int j : ones[(size_t)i])
            if (!gone[(size_t)j]) {
                gone[(size_t)i] = 1;
                ++pr.removed;
                break;
            }
    std::vector<int> renum((size_t)n, -1);
    long m = 0;
    for (long i = 0; i < n; i++)
        if (!gone[(size_t)i]) renum[(size_t)i] = (int)m++;
    pr.n = m;
    std::cout << "Removed " << pr.removed << " collinear markers" << std::endl;  // sepselect.py:483
    pr.nb.assign((size_t)m, {});
    for (size_t k = 0; k < adj.r.size(); k++) {
        const int a = renum[(size_t)adj.r[k]], b = renum[(size_t)adj.c[k]];
        if (a >= 0 && b >= 0 && adj.v[k] != 0) pr.nb[(size_t)a].push_back(b);  // astype(bool); input is row-major sorted
    }
    pr.Ct.assign((size_t)(P + 1) * (size_t)m, 0.0);
    std::vector<uint8_t> diag((size_t)m, 0);
    for (size_t k = 0; k < cor.r.size(); k++) {
        const int a = renum[(size_t)cor.r[k]], b = renum[(size_t)cor.c[k]];
        if (a < 0 || b < 0) continue;
        double v = cor.v[k];
        if (a == b) {
            v = 1.0;
            diag[(size_t)a] = 1;
        } else if (v == 0) {
            continue;
        }
        pr.cr.push_back(a);
        pr.cc.push_back(b);
        pr.cv.push_back(v);
        if (a <= P) pr.Ct[(size_t)a * (size_t)m + (size_t)b] = v;
    }
    for (long a = 0; a < m; a++)
        if (!diag[(size_t)a]) {  // absent diagonal entries become 1 as well; put them in place below
            pr.cr.push_back((int)a);
            pr.cc.push_back((int)a);
            pr.cv.push_back(1.0);
            if (a <= P) pr.Ct[(size_t)a * (size_t)m + (size_t)a] = 1.0;
        }
    {
        std::vector<size_t> ord(pr.cr.size());
        for (size_t k = 0; k < ord.size(); k++) ord[k] = k;
        std::sort(ord.begin(), ord.end(), [&](size_t x, size_t y) { return pr.cr[x] != pr.cr[y] ? pr.cr[x] < pr.cr[y] : pr.cc[x] < pr.cc[y]; });
        std::vector<int> r2(ord.size()), c2(ord.size());
        std::vector<double> v2(ord.size());
        for (size_t k = 0; k < ord.size(); k++) {
            r2[k] = pr.cr[ord[k]];
            c2[k] = pr.cc[ord[k]];
            v2[k] = pr.cv[ord[k]];
        }
        pr.cr.swap(r2);
        pr.cc.swap(c2);
        pr.cv.swap(v2);
    }
    // the search reads C[a, b] with a <= P only through row a: the matrix has to be symmetric there
    for (size_t k = 0; k < pr.cr.size(); k++) {
        const int a = pr.cr[k], b = pr.cc[k];
        if (b <= P && a > b && pr.Ct[(size_t)b * (size_t)m + (size_t)a] != pr.cv[k])
            fail("sepselect: the correlation matrix is not symmetric at (" + std::to_string(a + 1) + ", " + std::to_string(b + 1) + ")");
    }
    return pr;
}

}  // namespace

// orient_v_structures_merged(stem, alpha, num_samples, prior).to_file(out_stem) (sepselect.py:576-584, :545-573); with
// prior_path empty this is also everything sepselect_merged(...).to_file(...) writes before it fails on the missing PAG
int run_sepselect(cusk_ctx *ctx, const std::string &stem, double alpha, long num_samples, const std::string &prior_path,
                  const std::string &out_stem)
{
    if (!ctx) fail("sepselect needs a CUDA device: there is no CPU path");
    Problem pr = load_problem(stem);
    const long n = pr.n, P = pr.P;
    // orientation prior (sepselect.py:454-466): marker -> trait for every marker-trait link, then the optional P x P block
    std::vector<int> prior_pp;
    if (!prior_path.empty()) {
        check_path(prior_path);
        prior_pp = read_ints_from_binary(prior_path);
        if ((long)prior_pp.size() != P * P) fail("orientation prior has to have n_trait * n_trait entries");
    }
    auto prior = [&](int a, int b) -> int {  // orientation_prior[a, b]
        if (a < P && b < P) return prior_pp.empty() ? 0 : prior_pp[(size_t)a * (size_t)P + (size_t)b];
        if (a >= P && b < P) return pr.has(a, b) ? 1 : 0;
        return 0;
    };
    std::cout << "Orienting v-structures" << std::endl;

    // rfci-relevant unshielded triples (sepselect.py:68-86) and their outer pairs in both orders (:88-93)
    std::vector<std::array<int, 3>> ut;
    pr.for_each_relevant_triple([&](int x, int y, int z) {
        if (!pr.is_marker(y) && x < z && (int)pr.is_marker(x) + (int)pr.is_marker(y) + (int)pr.is_marker(z) < 2) ut.push_back({x, y, z});
    });
    std::sort(ut.begin(), ut.end());
    ut.erase(std::unique(ut.begin(), ut.end()), ut.end());
    std::vector<std::array<int, 2>> pairs;
    for (const auto &t : ut) {
        pairs.push_back({t[0], t[2]});
        pairs.push_back({t[2], t[0]});
    }
    std::sort(pairs.begin(), pairs.end());
    pairs.erase(std::unique(pairs.begin(), pairs.end()), pairs.end());
    if (pairs.empty()) fail("sepselect: the skeleton has no unshielded triple with a trait in the middle");  // max() of an empty sequence

    // the search, on the device
    std::vector<uint32_t> tmask((size_t)n, 0u);  // trait_neighbors(i): neighbours < num_phen (sepselect.py:538-540)
    for (long i = 0; i < n; i++)
        for (int b : pr.nb[(size_t)i]) {
            if (b >= P) break;
            tmask[(size_t)i] |= 1u << b;
        }
    std::vector<double> thr((size_t)P + 1);
    const double q = cusk_qnorm(1.0 - alpha / 2.0);  // sepselect.py:26-27
    for (long l = 0; l <= P; l++) thr[(size_t)l] = q / std::sqrt((double)(num_samples - l - 3));
    std::vector<cusk_sepselect_result> res(pairs.size());
    std::cout << "Processing " << pairs.size() << " outer pairs" << std::endl;
    const int rc = cusk_sepselect_pairs(ctx, pr.Ct.data(), n, (int)P, (int)n, &pairs[0][0], (int64_t)pairs.size(), tmask.data(),
                                        thr.data(), res.data());
    if (rc != CUSK_OK) return rc;
    auto find_pair = [&](int a, int c) -> const cusk_sepselect_result * {
        const std::array<int, 2> key{a, c};
        auto it = std::lower_bound(pairs.begin(), pairs.end(), key);
        return it != pairs.end() && *it == key ? &res[(size_t)(it - pairs.begin())] : nullptr;
    };
    auto in_set = [](const cusk_sepselect_result *r, int y) {
        for (int k = 0; r && k < r->len; k++)
            if (r->set[k] == y) return true;
        return false;
    };
    int max_level = 0;
    for (const auto &r : res) max_level = std::max(max_level, (int)r.len);

    // mark_ambiguous_triples (sepselect.py:95-107) over ALL unshielded triples; only those of a searched pair can qualify
    std::vector<std::array<int, 3>> atr;
    pr.for_each_relevant_triple([&](int a, int b, int c) {
        const cusk_sepselect_result *r = find_pair(a, c);
        if (r && !r->found_minimum && in_set(r, b)) atr.push_back({a, b, c});
    });
    std::sort(atr.begin(), atr.end());
    atr.erase(std::unique(atr.begin(), atr.end()), atr.end());

    // orient_v_structures (sepselect.py:485-511); the PAG as a map over the entries that are not 0
    std::map<std::pair<int, int>, int> pag;
    for (long a = 0; a < n; a++)
        for (int b : pr.nb[(size_t)a]) pag[{(int)a, b}] = 1;
    for (const auto &t : ut) {
        const int x = t[0], y = t[1], z = t[2];
        const bool orient = !in_set(find_pair(x, z), y) && !in_set(find_pair(z, x), y);
        for (int e : {x, z}) {
            if (prior(e, y) == 1) {
                pag[{e, y}] = 2;
                pag[{y, e}] = 3;
            } else if (prior(y, e) == 1) {
                pag[{y, e}] = 2;
                pag[{e, y}] = 3;
            } else if (orient) {
                pag[{e, y}] = 2;
            }
        }
    }

    // to_file (sepselect.py:545-573)
    {
        std::string s;
        for (long v : {n, P, (long)max_level, (long)atr.size(), (long)ut.size()}) {
            if (!s.empty()) s += '\t';
            s += std::to_string(v);
        }
        s += '\n';
        FILE *f = fopen((out_stem + ".mdim").c_str(), "wb");
        if (!f || fwrite(s.data(), 1, s.size(), f) != s.size() || fclose(f) != 0) fail("cannot write " + out_stem + ".mdim");
    }
    {
        std::vector<int> r, c, v;
        for (long a = 0; a < n; a++)
            for (int b : pr.nb[(size_t)a]) {
                r.push_back((int)a);
                c.push_back(b);
                v.push_back(1);
            }
        write_mtx(out_stem + "_sam.mtx", "integer", n, r, c, v);
        write_mtx(out_stem + "_scm.mtx", "real", n, pr.cr, pr.cc, pr.cv);
        r.clear();
        c.clear();
        v.clear();
        for (const auto &kv : pag) {
            r.push_back(kv.first.first);
            c.push_back(kv.first.second);
            v.push_back(kv.second);
        }
        write_mtx(out_stem + "_spm.mtx", "integer", n, r, c, v);
    }
    write_ints_to_binary(atr.empty() ? nullptr : &atr[0][0], atr.size() * 3, out_stem + ".atr");
    write_ints_to_binary(ut.empty() ? nullptr : &ut[0][0], ut.size() * 3, out_stem + ".ut");
    {  // max_sepset_to_file (sepselect.py:563-573): 1-based "i j s1 s2 ..." for every pair with a non-empty set
        std::string s;
        for (size_t k = 0; k < pairs.size(); k++) {
            if (res[k].len == 0) continue;
            s += std::to_string(pairs[k][0] + 1) + " " + std::to_string(pairs[k][1] + 1);
            for (int e = 0; e < res[k].len; e++) s += " " + std::to_string(res[k].set[e] + 1);
            s += '\n';
        }
        FILE *f = fopen((out_stem + ".ssm").c_str(), "wb");
        if (!f || fwrite(s.data(), 1, s.size(), f) != s.size() || fclose(f) != 0) fail("cannot write " + out_stem + ".ssm");
    }
    return CUSK_OK;
}

}  // namespace cuskhost
