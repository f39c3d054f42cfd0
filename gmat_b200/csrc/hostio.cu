// Host-side text writer for the GRM files.  agmat / dgmat_as end with
// np.savetxt(out_file + '0', mat)  (gmat/gmatrix/gmatrix.py:10-14: fmt '%.18e', ' ' delimiter, '\n' newline);
// at n = 10,000 that is 2.5 GB of text and minutes of single-threaded Python formatting -- far longer than
// the 1.5 s GRM kernel.  This writer produces the same bytes (C's "%.18e" and Python's '%.18e' are both
// correctly rounded, two-digit-minimum exponents) with all host threads formatting blocks of rows in parallel.
// No GPU work; plain pointers.
#include "common.cuh"
#include <algorithm>
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
            const int len = snprintf(buf, sizeof(buf), "%.18e", row[c]);
            out.append(buf, (size_t)len);
            out.push_back(c + 1 < cols ? ' ' : '\n');
        }
        if (cols == 0) out.push_back('\n');
    }
}

}  // namespace gb

using namespace gb;

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
    std::vector<std::string> bufs((size_t)nt);
    bool ok = true;
    for (long long base = 0; base < rows && ok; base += per * nt) {
        std::vector<std::thread> pool;
        int used = 0;
        for (int t = 0; t < nt; ++t) {
            const long long r0 = base + t * per, r1 = std::min(rows, r0 + per);
            if (r0 >= rows) break;
            ++used;
            pool.emplace_back(format_rows, mat, cols, r0, r1, std::ref(bufs[(size_t)t]));
        }
        for (auto &th : pool) th.join();
        for (int t = 0; t < used && ok; ++t)
            ok = fwrite(bufs[(size_t)t].data(), 1, bufs[(size_t)t].size(), f) == bufs[(size_t)t].size();
    }
    if (fclose(f) != 0) ok = false;
    if (!ok) {
        set_error("write to %s failed", path);
        return GMAT_B200_ERR_IO;
    }
    return GMAT_B200_OK;
}
