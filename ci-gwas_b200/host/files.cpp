// files.cpp -- on-disk formats at the boundary of the skeleton path: readers for the
// genotype file set, phenotypes, block descriptors and summary statistics; writers for
// .mdim/.ixs/.adj/.corr/.sep and the prep-bed column files.  Whole files are read in one call and
// tokenised in memory (the reference parses line by line through std::istringstream).
#include <fcntl.h>
#include <sys/stat.h>
#include <unistd.h>

#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <sstream>

#include "host.hpp"

namespace cuskhost {

void fail(const std::string &msg, int code) { throw HostError{msg, code}; }

bool path_exists(const std::string &p)
{
    struct stat sb;
    return stat(p.c_str(), &sb) == 0;
}

void check_path(const std::string &p)
{
    if (!path_exists(p)) fail("file or directory not found: " + p);
}

std::string make_path(const std::string &dir, const std::string &stem, const std::string &suffix)
{
    if (dir.empty() || dir.back() == '/') return dir + stem + suffix;
    return dir + "/" + stem + suffix;
}

std::string slurp(const std::string &path)
{
    FILE *f = fopen(path.c_str(), "rb");
    if (!f) fail("file or directory not found: " + path);
    std::string out;
    fseek(f, 0, SEEK_END);
    const long sz = ftell(f);
    fseek(f, 0, SEEK_SET);
    if (sz > 0) {
        out.resize((size_t)sz);
        const size_t got = fread(&out[0], 1, (size_t)sz, f);
        out.resize(got);
    }
    fclose(f);
    return out;
}

namespace {

// iterates the lines of a buffer the way std::getline does (a trailing '\n' ends the last line and
// does not start an empty one)
struct Lines {
    const char *p, *end;
    explicit Lines(const std::string &s) : p(s.data()), end(s.data() + s.size()) {}
    bool next(const char **b, const char **e)
    {
        if (p >= end) return false;
        const char *nl = (const char *)memchr(p, '\n', (size_t)(end - p));
        *b = p;
        *e = nl ? nl : end;
        p = nl ? nl + 1 : end;
        return true;
    }
};

inline bool is_ws(char c) { return c == ' ' || c == '\t' || c == '\r' || c == '\n' || c == '\v' || c == '\f'; }

// next whitespace-delimited token of [*p, e); false at the end of the line
inline bool next_tok(const char **p, const char *e, const char **tb, const char **te)
{
    const char *q = *p;
    while (q < e && is_ws(*q)) q++;
    if (q >= e) {
        *p = q;
        return false;
    }
    *tb = q;
    while (q < e && !is_ws(*q)) q++;
    *te = q;
    *p = q;
    return true;
}

inline bool tok_is(const char *b, const char *e, const char *lit)
{
    const size_t n = strlen(lit);
    return (size_t)(e - b) == n && memcmp(b, lit, n) == 0;
}

// std::stof on a token
inline float tok_stof(const char *b, const char *e)
{
    char buf[64];
    const size_t n = (size_t)(e - b) < sizeof(buf) - 1 ? (size_t)(e - b) : sizeof(buf) - 1;
    memcpy(buf, b, n);
    buf[n] = 0;
    char *endp = nullptr;
    const float v = strtof(buf, &endp);
    if (endp == buf) fail(std::string("cannot parse number: ") + buf);
    return v;
}

inline long tok_stol(const char *b, const char *e)
{
    char buf[64];
    const size_t n = (size_t)(e - b) < sizeof(buf) - 1 ? (size_t)(e - b) : sizeof(buf) - 1;
    memcpy(buf, b, n);
    buf[n] = 0;
    char *endp = nullptr;
    const long v = strtol(buf, &endp, 10);
    if (endp == buf) fail(std::string("cannot parse integer: ") + buf);
    return v;
}

inline float zero_if_nan(float v) { return std::isnan(v) ? 0.0f : v; }

}  // namespace

std::vector<std::string> split_ws(const char *b, const char *e)
{
    std::vector<std::string> out;
    const char *tb, *te;
    while (next_tok(&b, e, &tb, &te)) out.emplace_back(tb, te);
    return out;
}

size_t count_lines(const std::string &path)
{
    const std::string s = slurp(path);
    Lines L(s);
    const char *b, *e;
    size_t n = 0;
    while (L.next(&b, &e)) n++;
    return n;
}

std::vector<float> read_floats_from_line_range(const std::string &path, size_t first, size_t last)
{
    const std::string s = slurp(path);
    Lines L(s);
    const char *b, *e;
    std::vector<float> out;
    size_t ix = 0;
    while (L.next(&b, &e)) {
        if (ix > last) break;
        if (ix >= first) {
            const char *tb, *te, *p = b;
            if (!next_tok(&p, e, &tb, &te)) fail("empty line in " + path);
            out.push_back(tok_stof(tb, te));
        }
        ix++;
    }
    return out;
}

std::vector<int> read_ints_from_lines(const std::string &path)
{
    const std::string s = slurp(path);
    Lines L(s);
    const char *b, *e;
    std::vector<int> out;
    while (L.next(&b, &e)) {
        const char *tb, *te, *p = b;
        if (!next_tok(&p, e, &tb, &te)) fail("empty line in " + path);
        out.push_back((int)tok_stol(tb, te));
    }
    return out;
}

std::vector<int> read_ints_from_binary(const std::string &path)
{
    const std::string s = slurp(path);
    std::vector<int> out(s.size() / sizeof(int));
    memcpy(out.data(), s.data(), out.size() * sizeof(int));
    return out;
}

std::vector<float> read_floats_from_binary(const std::string &path)
{
    const std::string s = slurp(path);
    std::vector<float> out(s.size() / sizeof(float));
    memcpy(out.data(), s.data(), out.size() * sizeof(float));
    return out;
}

static void write_raw(const void *d, size_t bytes, const std::string &path)
{
    FILE *f = fopen(path.c_str(), "wb");
    if (!f) fail("cannot write " + path);
    if (bytes) fwrite(d, 1, bytes, f);
    fclose(f);
}
void write_ints_to_binary(const int *d, size_t n, const std::string &path) { write_raw(d, n * sizeof(int), path); }
void write_floats_to_binary(const float *d, size_t n, const std::string &path) { write_raw(d, n * sizeof(float), path); }

// operator<< of a float with the default stream precision (6 significant digits) == "%g"
void write_column(const std::vector<float> &v, const std::string &path)
{
    std::ofstream f(path, std::ios::out);
    if (!f) fail("cannot write " + path);
    for (float x : v) f << x << '\n';
}
void write_column(const std::vector<int> &v, const std::string &path)
{
    std::ofstream f(path, std::ios::out);
    if (!f) fail("cannot write " + path);
    for (int x : v) f << x << '\n';
}

// ---- genotype file set ---------------------------------------------------------------------------
std::vector<MarkerBlock> read_blocks(const std::string &path)
{
    const std::string s = slurp(path);
    Lines L(s);
    const char *b, *e;
    std::vector<MarkerBlock> out;
    size_t global_offset = 0, on_chr = 0;
    std::string cur;
    while (L.next(&b, &e)) {
        const std::vector<std::string> f = split_ws(b, e);
        if (f.size() < 3) fail("bad line in block file " + path);
        if (f[0] != cur) {
            cur = f[0];
            global_offset += on_chr;
            on_chr = 0;
        }
        MarkerBlock mb;
        mb.chr = f[0];
        mb.first = (size_t)std::stol(f[1]);
        mb.last = (size_t)std::stol(f[2]);
        mb.chr_global_offset = global_offset;
        out.push_back(mb);
        on_chr += mb.size();
    }
    return out;
}

BimInfo::BimInfo(const std::string &path)
{
    const std::string s = slurp(path);
    Lines L(s);
    const char *b, *e;
    while (L.next(&b, &e)) {
        const char *tb, *te, *p = b;
        if (!next_tok(&p, e, &tb, &te)) fail("empty line in " + path);
        const std::string chr(tb, te);
        if (num_lines == 0 || chr != chr_ids.back()) {
            chr_start.push_back(num_lines);
            ix[chr] = chr_ids.size();
            chr_ids.push_back(chr);
            num_on_chr.push_back(0);
        }
        num_on_chr.back()++;
        num_lines++;
    }
}

size_t BimInfo::chr_index(const std::string &chr) const
{
    auto it = ix.find(chr);
    if (it == ix.end()) fail("chr id does not match .bim content");
    return it->second;
}

BedDims read_dims(const std::string &path)
{
    const std::string s = slurp(path);
    const std::vector<std::string> f = split_ws(s.data(), s.data() + s.size());
    if (f.size() < 2) fail("bad .dim file " + path);
    BedDims d;
    d.num_samples = (size_t)std::stol(f[0]);
    d.num_markers = (size_t)std::stol(f[1]);
    return d;
}

bool has_valid_bed_prefix(const std::string &bed_path)
{
    FILE *f = fopen(bed_path.c_str(), "rb");
    if (!f) return false;
    unsigned char b[3] = {0, 0, 0};
    const size_t got = fread(b, 1, 3, f);
    fclose(f);
    return got == 3 && b[0] == 0x6c && b[1] == 0x1b && b[2] == 0x01;
}

Phen load_phen(const std::string &path)
{
    const std::string s = slurp(path);
    Lines L(s);
    const char *b, *e;
    Phen ph;
    if (!L.next(&b, &e)) return ph;  // header
    // pass 1: the lines; pass 2 (all host threads): their numbers.  A .phen of 458 000 samples x 17 traits is 80 MB of text.
    std::vector<std::pair<const char *, const char *>> lines;
    while (L.next(&b, &e)) lines.emplace_back(b, e);
    ph.num_samples = lines.size();
    if (lines.empty()) return ph;
    {
        const char *tb, *te, *p = lines[0].first;
        size_t col = 0;
        while (next_tok(&p, lines[0].second, &tb, &te)) col++;
        ph.num_phen = col >= 2 ? col - 2 : 0;
    }
    const size_t np = ph.num_phen, ns = ph.num_samples;
    ph.data.assign(np * ns, 0.0f);
    long long bad_line = -1;
#pragma omp parallel for schedule(static) if (ns > 4096)
    for (long long i = 0; i < (long long)ns; i++) {
        const char *tb, *te, *p = lines[(size_t)i].first, *le = lines[(size_t)i].second;
        size_t col = 0, nvals = 0;
        while (next_tok(&p, le, &tb, &te)) {
            if (col >= 2) {
                float v;
                if (tok_is(tb, te, "NA"))
                    v = std::nanf("");
                else {
                    char buf[64];
                    const size_t n = (size_t)(te - tb) < sizeof(buf) - 1 ? (size_t)(te - tb) : sizeof(buf) - 1;
                    memcpy(buf, tb, n);
                    buf[n] = 0;
                    v = (float)atof(buf);  // phen.cpp:22: (float)atof
                }
                if (nvals < np) ph.data[nvals * ns + (size_t)i] = v;  // column-major [num_phen][num_samples]
                nvals++;
            }
            col++;
        }
        if (nvals != np) bad_line = i;
    }
    if (bad_line >= 0)
        fail("Inconsistent row width in .phen file\nLine: " + std::string(lines[(size_t)bad_line].first, lines[(size_t)bad_line].second));
    return ph;
}

// reads [offset, offset + bytes) of a file into dst with several threads (pread on the page cache is memcpy-bound:
// one thread moves ~2 GB/s, a block of C3 is 1.1 GB)
void read_file_range(const std::string &path, size_t offset, size_t bytes, unsigned char *dst)
{
    const int fd = open(path.c_str(), O_RDONLY);
    if (fd < 0) fail("file or directory not found: " + path);
    const size_t chunk = (size_t)8 << 20;
    const long long nchunks = (long long)((bytes + chunk - 1) / chunk);
    bool short_read = false;
#pragma omp parallel for schedule(dynamic, 1) num_threads(8) if (nchunks > 2)
    for (long long c = 0; c < nchunks; c++) {
        size_t off = (size_t)c * chunk;
        const size_t end = std::min(bytes, off + chunk);
        while (off < end) {
            const ssize_t got = pread(fd, dst + off, end - off, (off_t)(offset + off));
            if (got <= 0) {
                short_read = true;
                break;
            }
            off += (size_t)got;
        }
    }
    close(fd);
    if (short_read) fail("short read of " + path);
}

// ---- summary statistics ------------------------------------------------------------------------
std::vector<float> load_mxm_square(const std::string &path, size_t *num_markers)
{
    const std::vector<float> tril = read_floats_from_binary(path);
    const size_t m = (size_t)((std::sqrt((double)(4 * 2 * tril.size() + 1)) - 1) / 2);
    std::vector<float> sq(m * m, 1.0f);
    size_t k = 0;
    for (size_t i = 0; i < m; i++)
        for (size_t j = 0; j <= i; j++) {
            const float v = zero_if_nan(tril[k++]);
            sq[i * m + j] = v;
            sq[j * m + i] = v;
        }
    *num_markers = m;
    return sq;
}

MxpRows load_mxp(const std::string &corr_path, const std::string &se_path, bool by_block, size_t first_global,
                 size_t last_global, const std::vector<int> &marker_ixs)
{
    const bool het = !se_path.empty();
    const std::string sc = slurp(corr_path);
    const std::string ss = het ? slurp(se_path) : std::string();
    Lines Lc(sc), Ls(ss);
    const char *b, *e, *sb = nullptr, *se = nullptr;
    if (!Lc.next(&b, &e)) fail(het ? "marker-trait corr file seems to be empty" : "marker-trait summary stat file seems to be empty");
    if (het && !Ls.next(&sb, &se)) fail("marker-trait se file seems to be empty");
    const std::vector<std::string> header = split_ws(b, e);
    if (header.size() < 3 || header[0] != "chr" || header[1] != "snp" || header[2] != "ref")
        fail("marker-trait summary stat file has bad header");
    MxpRows out;
    out.num_phen = header.size() - 3;
    // pass 1 (sequential, memchr speed): the lines to take.  pass 2 (all host threads): their numbers -- a
    // whole-genome table is 10^6 rows x 17 traits of text, and `cuskss-all` reads it once for all blocks.
    struct Span {
        const char *b, *e, *sb, *se;
    };
    std::vector<Span> take;
    size_t line = 0;
    while (Lc.next(&b, &e)) {
        if (het) Ls.next(&sb, &se);
        bool t;
        if (by_block) {
            if (line > last_global) break;
            t = line >= first_global;
        } else {
            if (take.size() >= marker_ixs.size()) break;
            t = line == (size_t)marker_ixs[take.size()];
        }
        if (t) take.push_back({b, e, sb, se});
        line++;
    }
    const size_t p = out.num_phen, nrows = take.size();
    out.num_markers = nrows;
    out.corr.assign(nrows * p, 0.0f);
    if (het) out.ess.assign(nrows * p, 0.0f);
    int bad = 0;  // 1: short corr line, 2: short se line, 3: unparsable number
#pragma omp parallel for schedule(static) if (nrows > 4096)
    for (long long r = 0; r < (long long)nrows; r++) {
        const Span &sp = take[(size_t)r];
        const char *q = sp.b, *tb, *te, *qs = sp.sb, *tsb = nullptr, *tse = nullptr;
        for (size_t j = 0; j < p + 3; j++) {
            if (!next_tok(&q, sp.e, &tb, &te)) {
                bad = 1;
                break;
            }
            if (het && !next_tok(&qs, sp.se, &tsb, &tse)) {
                bad = 2;
                break;
            }
            if (j < 3) continue;
            bool na = tok_is(tb, te, "NA") || tok_is(tb, te, "NaN") || tok_is(tb, te, "nan");
            if (het) na = na || tok_is(tb, te, "NAN");
            float rho = 0.0f, nn = std::nanf("");
            if (!na) {
                char buf[64], *endp = nullptr;
                size_t n = (size_t)(te - tb) < sizeof(buf) - 1 ? (size_t)(te - tb) : sizeof(buf) - 1;
                memcpy(buf, tb, n);
                buf[n] = 0;
                rho = strtof(buf, &endp);
                if (endp == buf) bad = 3;
                if (het) {
                    n = (size_t)(tse - tsb) < sizeof(buf) - 1 ? (size_t)(tse - tsb) : sizeof(buf) - 1;
                    memcpy(buf, tsb, n);
                    buf[n] = 0;
                    const float sev = strtof(buf, &endp);
                    if (endp == buf) bad = 3;
                    const float ss_sqrt = (float)((1.0 - (double)(rho * rho)) / (double)sev);
                    nn = ss_sqrt * ss_sqrt;
                }
            }
            out.corr[(size_t)r * p + (j - 3)] = rho;
            if (het) out.ess[(size_t)r * p + (j - 3)] = nn;
        }
    }
    if (bad == 1) fail("short line in " + corr_path);
    if (bad == 2) fail("short line in " + se_path);
    if (bad == 3) fail("cannot parse a number in " + corr_path + (het ? " / " + se_path : std::string()));
    return out;
}

Pxp load_pxp(const std::string &corr_path, const std::string &se_path, float sample_size)
{
    const bool het = !se_path.empty();
    const std::string sc = slurp(corr_path);
    const std::string ss = het ? slurp(se_path) : std::string();
    Lines Lc(sc), Ls(ss);
    const char *b, *e, *sb = nullptr, *se = nullptr;
    if (!Lc.next(&b, &e)) fail(het ? "trait summary corr file seems to be empty" : "trait summary stat file seems to be empty");
    if (het && !Ls.next(&sb, &se)) fail("trait summary se file seems to be empty");
    Pxp out;
    out.num_phen = split_ws(b, e).size();
    const size_t p = out.num_phen;
    out.corr.assign(p * p, 1.0f);
    out.ess.assign(p * p, het ? 0.0f : sample_size);
    size_t row = 0;
    while (Lc.next(&b, &e)) {
        if (het) Ls.next(&sb, &se);
        const char *q = b, *tb, *te, *qs = sb, *tsb = nullptr, *tse = nullptr;
        if (!next_tok(&q, e, &tb, &te)) break;  // empty line ends the table
        if (het) next_tok(&qs, se, &tsb, &tse);
        if (row >= p) fail("too many rows in " + corr_path);
        for (size_t j = 1; j <= p; j++) {
            if (!next_tok(&q, e, &tb, &te)) fail("short line in " + corr_path);
            if (het && !next_tok(&qs, se, &tsb, &tse)) fail("short line in " + se_path);
            float c = tok_stof(tb, te);
            if (het) {
                float nn;
                if (std::isnan(c)) {
                    c = 0.0f;
                    nn = std::nanf("");
                } else {
                    const float sev = tok_stof(tsb, tse);
                    const float ss_sqrt = (float)((1.0 - (double)(c * c)) / (double)sev);
                    nn = ss_sqrt * ss_sqrt;
                }
                out.ess[row * p + j - 1] = nn;
            } else {
                c = zero_if_nan(c);
            }
            out.corr[row * p + j - 1] = c;
        }
        row++;
    }
    for (size_t i = 0; i < p; i++)
        for (size_t j = i + 1; j < p; j++) {
            out.corr[j * p + i] = out.corr[i * p + j];
            if (het) out.ess[j * p + i] = out.ess[i * p + j];
        }
    return out;
}

SquareInputs make_square_cuskss_inputs(const std::vector<float> &mxm_sq, size_t m, const MxpRows &mxp, const Pxp &pxp,
                                       float sample_size, bool het)
{
    SquareInputs si;
    si.num_markers = m;
    si.num_phen = pxp.num_phen;
    const size_t p = pxp.num_phen, nv = m + p;
    si.C.assign(nv * nv, 1.0f);
    si.uniform = !het;
    si.n_uniform = sample_size;
    if (het) si.N.assign(nv * nv, sample_size);
    for (size_t i = 0; i < m; i++) memcpy(&si.C[i * nv], &mxm_sq[i * m], m * sizeof(float));
    for (size_t i = 0; i < m; i++)
        for (size_t t = 0; t < p; t++) {
            const float r = mxp.corr[i * p + t];
            si.C[i * nv + m + t] = r;
            si.C[(m + t) * nv + i] = r;
            if (het) {
                const float nn = mxp.ess[i * p + t];
                si.N[i * nv + m + t] = nn;
                si.N[(m + t) * nv + i] = nn;
            }
        }
    for (size_t a = 0; a < p; a++)
        for (size_t b = 0; b < p; b++) {
            si.C[(m + a) * nv + m + b] = pxp.corr[a * p + b];
            if (het) si.N[(m + a) * nv + m + b] = pxp.ess[a * p + b];
        }
    return si;
}

// ---- outputs -----------------------------------------------------------------------------------------
void Reduced::to_file(const std::string &base) const
{
    // .mdim is what `cusk-all` / `cuskss-all` take as "this block is done" when they resume (host/sched.cu), so it is
    // written LAST and moved into place in one step: a run killed mid-write leaves no .mdim and the block is redone.
    std::remove((base + ".mdim").c_str());
    write_ints_to_binary(new_to_old.data(), new_to_old.size(), base + ".ixs");
    write_ints_to_binary(G.data(), G.size(), base + ".adj");
    write_floats_to_binary(C.data(), C.size(), base + ".corr");
    if (has_sep) write_ints_to_binary(S.data(), S.size(), base + ".sep");
    const std::string tmp = base + ".mdim.tmp";
    {
        std::ofstream f(tmp, std::ios::out);
        if (!f) fail("cannot write " + tmp);
        f << num_var << "\t" << num_phen << "\t" << max_level << std::endl;
        if (!f) fail("cannot write " + tmp);
    }
    if (std::rename(tmp.c_str(), (base + ".mdim").c_str()) != 0) fail("cannot move " + tmp + " to " + base + ".mdim");
}

}  // namespace cuskhost
