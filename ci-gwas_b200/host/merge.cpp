// merge.cpp -- `merge-block-outputs` and the cuskss-merged reformatting step in native code
// (SURVEY.md section 8f, row 1).  Both are pure host work over the files the skeleton path writes:
//
//   merge_block_outputs        cusk_postprocessing/merge_blocks.py:361-403, called by ci-gwas.py:459-464
//   reformat_cuskss_merged     cusk_postprocessing/merge_blocks.py:406-425, called by ci-gwas.py:452-456
//   write_mm                   cusk_postprocessing/merge_blocks.py:300-325
//
// The reference keeps every non-zero of every block in Python dicts keyed by index tuples.  Here the
// blocks are streamed once into flat coordinate lists; only the trait x trait corner, the one place where
// the keys of different blocks collide, keeps an index so that the dict semantics of the reference
// (insertion order, in-place overwrite, deletion) are reproduced.  The four output files are byte-identical
// to what the reference writes, including the order of the coordinate entries and the text of the values
// (f"{numpy.float32}" formats the value as a Python float: shortest round-trip digits of the DOUBLE).
//
// Index convention of the merged files (make_dm_ix_to_sm_ix, merge_blocks.py:24-31): 1-based, traits first
// (trait t -> t + 1), then the retained markers of all blocks in block order
// (marker k of a block -> k + markers_before + num_phen + 1).
#include <algorithm>
#include <charconv>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <iostream>

#include <omp.h>

#include "host.hpp"

namespace cuskhost {
namespace {

struct Mdim {
    long num_var = 0, num_phen = 0, max_level = 0;
    long num_markers() const { return num_var - num_phen; }
};

// load_mdim (merge_blocks.py:17-20); false when the file is absent (the reference's FileNotFoundError)
bool read_mdim(const std::string &base, Mdim *m)
{
    const std::string p = base + ".mdim";
    if (!path_exists(p)) return false;
    const std::string txt = slurp(p);
    const char *e = (const char *)memchr(txt.data(), '\n', txt.size());
    const std::vector<std::string> f = split_ws(txt.data(), e ? e : txt.data() + txt.size());
    if (f.size() < 3) fail("malformed .mdim file: " + p);
    try {
        m->num_var = std::stol(f[0]);
        m->num_phen = std::stol(f[1]);
        m->max_level = std::stol(f[2]);
    } catch (const std::exception &) {
        fail("malformed .mdim file: " + p);
    }
    if (m->num_phen < 0 || m->num_var < m->num_phen) fail("inconsistent dimensions in " + p);
    return true;
}

// block_size (merge_blocks.py:356-358): from the last two '_' separated fields of the stem
long block_size_of(const std::string &base)
{
    const size_t b = base.rfind('_');
    const size_t a = b == std::string::npos || b == 0 ? std::string::npos : base.rfind('_', b - 1);
    if (a == std::string::npos) fail("block stem without <chr>_<first>_<last>: " + base);
    try {
        return std::stol(base.substr(b + 1)) - std::stol(base.substr(a + 1, b - a - 1)) + 1;
    } catch (const std::exception &) {
        fail("block stem without <chr>_<first>_<last>: " + base);
    }
}

// repr(float(x)) of CPython (float_repr_style 'short'): shortest digits that round-trip the double,
// exponent form when the decimal exponent is < -4 or >= 16, ".0" appended to integers
void append_py_float(std::string &out, float x)
{
    const double d = (double)x;
    if (std::isnan(d)) {
        out += "nan";
        return;
    }
    if (std::isinf(d)) {
        out += d < 0 ? "-inf" : "inf";
        return;
    }
    char buf[64];
    const auto r = std::to_chars(buf, buf + sizeof buf, d, std::chars_format::scientific);
    const char *p = buf, *end = r.ptr;
    if (*p == '-') {
        out += '-';
        ++p;
    }
    const char *epos = std::find(p, end, 'e');
    char digits[32];
    int nd = 0;
    for (const char *q = p; q < epos; ++q)
        if (*q != '.') digits[nd++] = *q;
    int e10 = 0;
    std::from_chars(epos + 1 + (epos[1] == '+'), end, e10);
    if (e10 < -4 || e10 >= 16) {
        out += digits[0];
        if (nd > 1) {
            out += '.';
            out.append(digits + 1, (size_t)nd - 1);
        }
        out += 'e';
        out += e10 < 0 ? '-' : '+';
        const int a = e10 < 0 ? -e10 : e10;
        if (a < 10) out += '0';
        out += std::to_string(a);
    } else if (e10 < 0) {
        out += "0.";
        out.append((size_t)(-e10 - 1), '0');
        out.append(digits, (size_t)nd);
    } else {
        const int ip = e10 + 1;  // digits before the point
        if (nd <= ip) {
            out.append(digits, (size_t)nd);
            out.append((size_t)(ip - nd), '0');
            out += ".0";
        } else {
            out.append(digits, (size_t)ip);
            out += '.';
            out.append(digits + ip, (size_t)(nd - ip));
        }
    }
}

void append_int(std::string &out, long long v)
{
    char buf[24];
    const auto r = std::to_chars(buf, buf + sizeof buf, v);
    out.append(buf, r.ptr);
}

// Ordered coordinate list with the dict semantics the merge needs.  Keys with both indices <= P (trait x
// trait) are indexed; all other keys are unique over the whole merge and only appended.
template <typename V>
struct OrderedCoo {
    std::vector<int64_t> row, col;
    std::vector<V> val;
    std::vector<uint8_t> dead;
    std::vector<int64_t> pos;  // [(P+1)*(P+1)], -1 = absent
    long P = -1;
    size_t live = 0;

    void set_num_phen(long p)
    {
        if (P == p) return;
        if (P >= 0) fail("merge-block-outputs: the blocks disagree on the number of traits");
        P = p;
        pos.assign((size_t)(P + 1) * (size_t)(P + 1), -1);
    }
    bool corner(int64_t i, int64_t j) const { return i <= P && j <= P; }
    int64_t &at(int64_t i, int64_t j) { return pos[(size_t)i * (size_t)(P + 1) + (size_t)j]; }
    void append(int64_t i, int64_t j, V v)
    {
        row.push_back(i);
        col.push_back(j);
        val.push_back(v);
        dead.push_back(0);
        ++live;
    }
    // a[(i, j)] = v
    void assign(int64_t i, int64_t j, V v)
    {
        if (!corner(i, j)) {
            append(i, j, v);
            return;
        }
        int64_t &p = at(i, j);
        if (p >= 0) {
            val[(size_t)p] = v;
        } else {
            p = (int64_t)row.size();
            append(i, j, v);
        }
    }
    // del a[(i, j)] when present
    void erase(int64_t i, int64_t j)
    {
        int64_t &p = at(i, j);
        if (p < 0) return;
        dead[(size_t)p] = 1;
        p = -1;
        --live;
    }
    bool contains(int64_t i, int64_t j) { return corner(i, j) && at(i, j) >= 0; }
};

// one block's .adj / .corr as the reference loads it (load_mat_sparse, merge_blocks.py:103-116):
// non-zeros in row-major order of the block's matrix, translated to merged indices
template <typename V, typename F>
void for_each_nonzero(const std::vector<V> &dm, const Mdim &md, long marker_offset, F &&f)
{
    const long n = md.num_var, m = md.num_markers(), p = md.num_phen;
    if ((long)dm.size() != n * n) fail("merge-block-outputs: matrix file does not hold num_var^2 entries");
    std::vector<int64_t> dm2sm((size_t)n);
    for (long k = 0; k < n; k++) dm2sm[(size_t)k] = k < m ? k + marker_offset + p + 1 : k - m + 1;
    for (long i = 0; i < n; i++) {
        const V *r = dm.data() + (size_t)i * (size_t)n;
        for (long j = 0; j < n; j++)
            if (r[j] != 0) f(dm2sm[(size_t)i], dm2sm[(size_t)j], r[j]);  // NaN != 0: kept, as numpy does
    }
}

struct Merged {  // GlobalBdpcResult (merge_blocks.py:289-297)
    OrderedCoo<int> sam;
    OrderedCoo<float> scm;
    std::vector<int> gmi;  // global marker indices (the dict's values; keys are unique)
    long num_var = 0, num_phen = 0, max_level = 0;
};

void write_text(const std::string &path, const std::string &s)
{
    FILE *f = fopen(path.c_str(), "wb");
    if (!f || fwrite(s.data(), 1, s.size(), f) != s.size()) fail("cannot write " + path);
    fclose(f);
}

// GlobalBdpcResult.write_mm (merge_blocks.py:300-325)
void write_mm(Merged &g, const std::string &base)
{
    if (g.sam.live == 0) fail("merge-block-outputs: the merged adjacency matrix is empty");  // max() of an empty sequence
    int64_t nmax = 0;
    for (size_t k = 0; k < g.sam.row.size(); k++)
        if (!g.sam.dead[k]) nmax = std::max(nmax, g.sam.row[k]);
    auto header = [&](std::string &out, const char *field, size_t L) {
        out += "%%MatrixMarket matrix coordinate ";
        out += field;
        out += " general\n";
        append_int(out, nmax);
        out += '\t';
        append_int(out, nmax);
        out += '\t';
        append_int(out, (long long)L);
        out += '\n';
    };
    // the entries are formatted in parallel, a contiguous slice per thread, and written in order
    auto write_coo = [&](const std::string &path, const char *field, auto &coo, auto &&append_value) {
        const size_t n = coo.row.size();
        const int T = std::max(1, omp_get_max_threads());
        std::vector<std::string> part((size_t)T);
#pragma omp parallel for schedule(static, 1) num_threads(T)
        for (int t = 0; t < T; t++) {
            const size_t b = n * (size_t)t / (size_t)T, e = n * (size_t)(t + 1) / (size_t)T;
            std::string &out = part[(size_t)t];
            out.reserve((e - b) * 32);
            for (size_t k = b; k < e; k++) {
                if (coo.dead[k]) continue;
                append_int(out, coo.row[k]);
                out += '\t';
                append_int(out, coo.col[k]);
                out += '\t';
                append_value(out, coo.val[k]);
                out += '\n';
            }
        }
        std::string head;
        header(head, field, coo.live);
        FILE *f = fopen(path.c_str(), "wb");
        bool ok = f && fwrite(head.data(), 1, head.size(), f) == head.size();
        for (int t = 0; ok && t < T; t++) ok = fwrite(part[(size_t)t].data(), 1, part[(size_t)t].size(), f) == part[(size_t)t].size();
        if (f) ok = fclose(f) == 0 && ok;
        if (!ok) fail("cannot write " + path);
    };
    write_coo(base + "_sam.mtx", "integer", g.sam, [](std::string &o, int v) { append_int(o, v); });
    write_coo(base + "_scm.mtx", "real", g.scm, [](std::string &o, float v) { append_py_float(o, v); });
    write_text(base + ".mdim",
               std::to_string(g.num_var) + "\t" + std::to_string(g.num_phen) + "\t" + std::to_string(g.max_level) + "\n");
    std::sort(g.gmi.begin(), g.gmi.end());
    write_ints_to_binary(g.gmi.data(), g.gmi.size(), base + ".ixs");
}

std::vector<int> read_block_ints(const std::string &path)
{
    if (!path_exists(path)) fail("No such file: " + path);
    return read_ints_from_binary(path);
}
std::vector<float> read_block_floats(const std::string &path)
{
    if (!path_exists(path)) fail("No such file: " + path);
    return read_floats_from_binary(path);
}

}  // namespace

// merge_block_outputs (merge_blocks.py:361-403) followed by write_mm(<out_base>) (ci-gwas.py:463-464)
int merge_block_outputs(const std::string &blockfile, const std::string &outdir_in, const std::string &out_base)
{
    check_path(blockfile);
    std::string outdir = outdir_in;
    if (outdir.empty() || outdir.back() != '/') outdir += '/';  // ci-gwas.py:460-462
    // get_block_out_stems (merge_blocks.py:8-15): tab-separated chr, first, last
    std::vector<std::string> bases;
    {
        const std::string txt = slurp(blockfile);
        size_t b = 0;
        while (b < txt.size()) {
            size_t e = txt.find('\n', b);
            if (e == std::string::npos) e = txt.size();
            const std::vector<std::string> f = split_ws(txt.data() + b, txt.data() + e);
            if (f.size() >= 3) bases.push_back(outdir + f[0] + "_" + f[1] + "_" + f[2]);
            b = e + 1;
        }
    }
    if (bases.empty()) fail("merge-block-outputs: no blocks in " + blockfile);

    Merged g;
    long marker_offset = 0, global_marker_offset = 0;
    bool any = false;
    Mdim last{};
    for (size_t bi = 0; bi < bases.size(); bi++) {
        const std::string &base = bases[bi];
        Mdim md;
        // first block: a missing .adj/.corr/.ixs counts as a missing block too (they are read inside the try,
        // merge_blocks.py:364-378); later blocks: only the .mdim decides (merge_blocks.py:381-386)
        bool present = read_mdim(base, &md);
        if (present && bi == 0)
            for (const char *sfx : {".adj", ".corr", ".ixs"}) present = present && path_exists(base + sfx);
        if (!present) {
            std::cout << "Missing: " << base << std::endl;
            global_marker_offset += block_size_of(base);
            continue;
        }
        const long P = md.num_phen, m = md.num_markers();
        g.sam.set_num_phen(P);
        g.scm.set_num_phen(P);
        const std::vector<int> adj = read_block_ints(base + ".adj");
        const std::vector<float> corr = read_block_floats(base + ".corr");
        const std::vector<int> rel = read_block_ints(base + ".ixs");
        if ((long)rel.size() < m) fail("merge-block-outputs: " + base + ".ixs is shorter than the marker count");
        if (bi == 0) {
            // sam = bo.sam(): the block's own dict
            for_each_nonzero(adj, md, marker_offset, [&](int64_t i, int64_t j, int v) { g.sam.assign(i, j, v); });
        } else {
            // add_sam (merge_blocks.py:337-347): trait x trait links must be present in every block; the loop runs
            // over range(num_p) on 1-based keys, so trait num_p is never intersected, only united
            std::vector<uint8_t> in_b((size_t)(P + 1) * (size_t)(P + 1), 0);
            for_each_nonzero(adj, md, marker_offset, [&](int64_t i, int64_t j, int) {
                if (i <= P && j <= P) in_b[(size_t)i * (size_t)(P + 1) + (size_t)j] = 1;
            });
            for (long i = 1; i < P; i++)
                for (long j = 1; j < P; j++)
                    if (!in_b[(size_t)i * (size_t)(P + 1) + (size_t)j]) g.sam.erase(i, j);
            for_each_nonzero(adj, md, marker_offset, [&](int64_t i, int64_t j, int v) {
                if (i >= P || j >= P) g.sam.assign(i, j, v);
            });
        }
        // add_scm = dict.update (merge_blocks.py:333-334)
        for_each_nonzero(corr, md, marker_offset, [&](int64_t i, int64_t j, float v) { g.scm.assign(i, j, v); });
        // load_global_marker_indices (merge_blocks.py:58-72)
        for (long k = 0; k < m; k++) g.gmi.push_back(rel[(size_t)k] + (int)global_marker_offset);
        marker_offset += m;
        global_marker_offset += block_size_of(base);
        last = md;
        any = true;
    }
    if (!any) fail("merge-block-outputs: none of the blocks of " + blockfile + " has outputs in " + outdir);
    g.num_var = marker_offset + last.num_phen;
    g.num_phen = last.num_phen;
    g.max_level = last.max_level;
    write_mm(g, out_base);
    return 0;
}

// reformat_cuskss_merged_output(cusk_dir).write_mm(cusk_dir/cuskss_merged) (merge_blocks.py:406-425, ci-gwas.py:452-456)
int reformat_cuskss_merged(const std::string &cusk_dir)
{
    const std::string base = cusk_dir + "/cuskss_merged";
    Mdim md;
    if (!read_mdim(base, &md)) fail("No such file: " + base + ".mdim");
    const std::vector<int> old_glob = read_block_ints(cusk_dir + "/merged_blocks.ixs");
    const std::vector<int> ixs = read_block_ints(base + ".ixs");
    if (md.num_phen < 1 || (long)ixs.size() < md.num_phen)
        fail("cuskss_merged.ixs does not end with the trait indices");  // ixs[:-num_trait]
    Merged g;
    g.num_var = md.num_var;
    g.num_phen = md.num_phen;
    g.max_level = md.max_level;
    for (size_t k = 0; k + (size_t)md.num_phen < ixs.size(); k++) {
        if (ixs[k] < 0 || (size_t)ixs[k] >= old_glob.size()) fail("cuskss_merged.ixs points outside merged_blocks.ixs");
        g.gmi.push_back(old_glob[(size_t)ixs[k]]);
    }
    g.sam.set_num_phen(md.num_phen);
    g.scm.set_num_phen(md.num_phen);
    const std::vector<int> adj = read_block_ints(base + ".adj");
    const std::vector<float> corr = read_block_floats(base + ".corr");
    for_each_nonzero(adj, md, 0, [&](int64_t i, int64_t j, int v) { g.sam.assign(i, j, v); });
    for_each_nonzero(corr, md, 0, [&](int64_t i, int64_t j, float v) { g.scm.assign(i, j, v); });
    write_mm(g, base);
    return 0;
}

}  // namespace cuskhost
