// Host-side text writer for the GRM files.  agmat / dgmat_as end with
// np.savetxt(out_file + '0', mat)  (gmat/gmatrix/gmatrix.py:10-14: fmt '%.18e', ' ' delimiter, '\n' newline);
// at n = 10,000 that is 2.5 GB of text and minutes of single-threaded Python formatting -- far longer than
// the 1.5 s GRM kernel.  This writer produces the same bytes (C's "%.18e" and Python's '%.18e' are both
// correctly rounded, two-digit-minimum exponents) with all host threads formatting blocks of rows in parallel.
// No GPU work; plain pointers.
#include "common.cuh"
#include <algorithm>
#include <charconv>
#include <cmath>
#include <cstring>
#include <string>
#include <thread>
#include <vector>

namespace gb {

static void format_rows(const double *mat, long long cols, long long r0, long long r1, std::string &out) {
    out.clear();
    out.reserve((size_t)((r1 - r0) * cols * 26));
    char buf[64];
    for (long long r = r0; r < r1; ++r) {
        const double *row = mat + r * cols;
        for (long long c = 0; c < cols; ++c) {
            // "%.18e": std::to_chars with a precision is correctly rounded like printf and prints the same
            // two-digit-minimum exponent, several times faster; non-finite values keep printf's spelling
            int len;
            if (std::isfinite(row[c])) {
                len = (int)(std::to_chars(buf, buf + sizeof(buf), row[c], std::chars_format::scientific, 18).ptr - buf);
            } else {                                             // Python's '%.18e' % x: nan (never -nan), inf, -inf
                const char *t = std::isnan(row[c]) ? "nan" : (row[c] < 0 ? "-inf" : "inf");
                len = (int)strlen(t);
                memcpy(buf, t, (size_t)len);
            }
            out.append(buf, (size_t)len);
            out.push_back(c + 1 < cols ? ' ' : '\n');
        }
        if (cols == 0) out.push_back('\n');
    }
}

// repr(float) as pandas' to_csv writes a float64 column without float_format: the shortest digit string that
// round-trips (std::to_chars), laid out by CPython's rule (Python/pystrtod.c format_float_short, 'r'): exponent form
// when the decimal point falls outside (-4, 16], at least two exponent digits, ".0" after integers; NaN is pandas'
// na_rep (the empty string).  Returns the number of characters written to out (<= 32).
static int py_repr(double v, char *out) {
    if (std::isnan(v)) return 0;
    if (std::isinf(v)) { const char *t = v < 0 ? "-inf" : "inf"; const int l = (int)strlen(t); memcpy(out, t, (size_t)l); return l; }
    char sci[40];
    const auto res = std::to_chars(sci, sci + sizeof(sci), v, std::chars_format::scientific);   // [-]d[.ddd]e[+-]XX
    char *q = sci, *o = out;
    if (*q == '-') { *o++ = '-'; ++q; }
    char digits[24];
    int nd = 0;
    for (; q < res.ptr && *q != 'e'; ++q)
        if (*q != '.') digits[nd++] = *q;
    int e10 = 0;
    {
        ++q;                                                     // 'e'
        const bool neg = *q == '-';
        ++q;                                                     // sign
        for (; q < res.ptr; ++q) e10 = e10 * 10 + (*q - '0');
        if (neg) e10 = -e10;
    }
    const int decpt = e10 + 1;                                   // digits before the decimal point
    if (decpt > 16 || decpt < -3) {                              // exponent form
        *o++ = digits[0];
        if (nd > 1) { *o++ = '.'; memcpy(o, digits + 1, (size_t)(nd - 1)); o += nd - 1; }
        *o++ = 'e';
        *o++ = e10 < 0 ? '-' : '+';
        const int ae = e10 < 0 ? -e10 : e10;
        if (ae >= 100) *o++ = (char)('0' + ae / 100);
        *o++ = (char)('0' + ae / 10 % 10);
        *o++ = (char)('0' + ae % 10);
    } else if (decpt <= 0) {                                     // 0.000ddd
        *o++ = '0'; *o++ = '.';
        for (int z = 0; z < -decpt; ++z) *o++ = '0';
        memcpy(o, digits, (size_t)nd); o += nd;
    } else if (decpt >= nd) {                                    // ddd000.0
        memcpy(o, digits, (size_t)nd); o += nd;
        for (int z = nd; z < decpt; ++z) *o++ = '0';
        *o++ = '.'; *o++ = '0';
    } else {                                                     // dd.ddd
        memcpy(o, digits, (size_t)decpt); o += decpt;
        *o++ = '.';
        memcpy(o, digits + decpt, (size_t)(nd - decpt)); o += nd - decpt;
    }
    return (int)(o - out);
}

static int put_int(long long v, char *out) {
    char tmp[24];
    int l = 0;
    do { tmp[l++] = (char)('0' + v % 10); v /= 10; } while (v);
    for (int k = 0; k < l; ++k) out[k] = tmp[l - 1 - k];
    return l;
}

// lower-triangle rows [r0, r1) of the n x n matrix as "<row label> <col label> <value>\n"
static void format_tril(const double *mat, long long n, long long r0, long long r1, const char *ids,
                        const long long *id_off, std::string &out) {
    out.clear();
    char buf[96];
    for (long long r = r0; r < r1; ++r) {
        for (long long c = 0; c <= r; ++c) {
            if (ids) {
                out.append(ids + id_off[r], (size_t)(id_off[r + 1] - id_off[r]));
                out.push_back(' ');
                out.append(ids + id_off[c], (size_t)(id_off[c + 1] - id_off[c]));
                out.push_back(' ');
            } else {
                int l = put_int(r + 1, buf);
                buf[l++] = ' ';
                l += put_int(c + 1, buf + l);
                buf[l++] = ' ';
                out.append(buf, (size_t)l);
            }
            const int l = py_repr(mat[r * n + c], buf);
            out.append(buf, (size_t)l);
            out.push_back('\n');
        }
    }
}

// Row chunks [first[c], first[c+1]) are formatted by `nt` threads a round at a time and written in order; the
// threads of round k+1 format while the caller's thread writes round k (two sets of buffers).
template <class Fmt>
static bool write_chunks_pipelined(FILE *f, const std::vector<long long> &first, int nt, Fmt fmt) {
    const long long n_chunks = (long long)first.size() - 1;
    std::vector<std::string> bufs[2] = {std::vector<std::string>((size_t)nt), std::vector<std::string>((size_t)nt)};
    std::vector<std::thread> pool[2];
    auto launch = [&](long long round) {
        const int set = (int)(round & 1);
        for (int t = 0; t < nt; ++t) {
            const long long c = round * nt + t;
            if (c >= n_chunks) break;
            pool[set].emplace_back([&, c, t, set] { fmt(first[(size_t)c], first[(size_t)c + 1], bufs[set][(size_t)t]); });
        }
    };
    const long long rounds = (n_chunks + nt - 1) / nt;
    bool ok = true;
    if (rounds > 0) launch(0);
    for (long long r = 0; r < rounds; ++r) {
        const int set = (int)(r & 1);
        const size_t used = pool[set].size();
        for (auto &th : pool[set]) th.join();
        pool[set].clear();
        if (r + 1 < rounds) launch(r + 1);
        for (size_t t = 0; t < used && ok; ++t)
            ok = fwrite(bufs[set][t].data(), 1, bufs[set][t].size(), f) == bufs[set][t].size();
    }
    return ok;
}

}  // namespace gb

using namespace gb;

extern "C" int gmat_b200_repr_host(double v, char *out32) {
    if (!out32) { set_error("NULL pointer"); return GMAT_B200_ERR_ARG; }
    const int l = py_repr(v, out32);
    out32[l] = 0;
    return GMAT_B200_OK;
}

extern "C" int gmat_b200_write_tril_host(const char *path, const double *mat, long long n, const char *ids,
                                         const long long *id_off, int threads) {
    GB_CHECK_ARG(path != nullptr && (mat != nullptr || n == 0) && n >= 0, "NULL pointer or negative shape");
    GB_CHECK_ARG((ids == nullptr) == (id_off == nullptr), "ids and id_off go together");
    FILE *f = fopen(path, "wb");
    if (!f) {
        set_error("cannot open %s for writing", path);
        return GMAT_B200_ERR_IO;
    }
    int nt = threads > 0 ? threads : (int)std::thread::hardware_concurrency();
    nt = std::max(1, std::min(nt, 256));
    std::vector<long long> first{0};                            // chunks of ~100k lines each (row r holds r + 1 lines)
    for (long long r = 0; r < n;) {
        long long lines = 0;
        while (r < n && lines < 100000) { lines += r + 1; ++r; }
        first.push_back(r);
    }
    bool ok = write_chunks_pipelined(f, first, nt, [&](long long r0, long long r1, std::string &out) {
        format_tril(mat, n, r0, r1, ids, id_off, out);
    });
    if (fclose(f) != 0) ok = false;
    if (!ok) {
        set_error("write to %s failed", path);
        return GMAT_B200_ERR_IO;
    }
    return GMAT_B200_OK;
}

extern "C" int gmat_b200_savetxt_host(const char *path, const double *mat, long long rows, long long cols,
                                      int threads) {
    GB_CHECK_ARG(path != nullptr && (mat != nullptr || rows * cols == 0), "NULL pointer");
    GB_CHECK_ARG(rows >= 0 && cols >= 0, "shape");
    FILE *f = fopen(path, "wb");
    if (!f) {
        set_error("cannot open %s for writing", path);
        return GMAT_B200_ERR_IO;
    }
    int nt = threads > 0 ? threads : (int)std::thread::hardware_concurrency();
    nt = std::max(1, std::min(nt, 256));
    // rows per thread per round: ~4 MB of text each
    long long per = std::max(1LL, (4LL << 20) / std::max(1LL, cols * 26));
    std::vector<long long> first{0};
    for (long long r = 0; r < rows; r += per) first.push_back(std::min(rows, r + per));
    bool ok = write_chunks_pipelined(f, first, nt, [&](long long r0, long long r1, std::string &out) {
        format_rows(mat, cols, r0, r1, out);
    });
    if (fclose(f) != 0) ok = false;
    if (!ok) {
        set_error("write to %s failed", path);
        return GMAT_B200_ERR_IO;
    }
    return GMAT_B200_OK;
}
