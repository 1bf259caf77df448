// pb200_io.cu -- output/<setup_name>.txt writer, host code only.
//
// Byte-identical replacement of the reference's result writer (prometheus.py:142-156):
//     orbphaseOutput = np.insert(orbphase / (2 pi), 0, nan)
//     output = np.vstack((orbphaseOutput, np.vstack((wavelength, R)).T))
//     np.savetxt(path, output, header=header)          # fmt '%.18e', delimiter ' ', newline '\n', comments '# '
// np.savetxt formats every cell with Python's '%.18e' % x, which like glibc's printf is correctly
// rounded, so snprintf("%.18e") yields the same characters (nan prints as "nan").  At C5 the table has
// 300 307 x 22 cells = 165 MB of text; np.savetxt needs tens of seconds for it, which would be the
// wall-clock floor of `python prometheus.py <setup>` once R itself takes 0.2 s.  Rows are formatted by
// a pool of host threads into per-thread buffers and written in order.
#include <cmath>
#include <cstdio>
#include <cstring>
#include <string>
#include <thread>
#include <vector>

#include "../../include/prometheus_b200.h"

namespace {

inline size_t put_cell(char* dst, double v) {
    if (std::isnan(v)) { memcpy(dst, "nan", 3); return 3; }                  // '%.18e' % nan == 'nan' (also for -nan)
    if (std::isinf(v)) { const char* s = v > 0 ? "inf" : "-inf"; size_t n = strlen(s); memcpy(dst, s, n); return n; }
    return (size_t)snprintf(dst, 32, "%.18e", v);
}

}  // namespace

extern "C" int pb200_write_output(const char* path, const char* header, const double* wavelength,
                                  const double* orbphase, const double* R, int n_orb, int64_t n_wav, int n_threads) {
    if (!path || !wavelength || !orbphase || !R || n_orb < 0 || n_wav < 0) return PB200_ERR_INVALID;
    FILE* f = fopen(path, "wb");
    if (!f) return PB200_ERR_INVALID;
    std::string head;
    if (header && *header) {                      // every header line gets the comment prefix "# "
        const char* p = header;
        while (true) {
            const char* nl = strchr(p, '\n');
            head += "# ";
            head.append(p, nl ? (size_t)(nl - p) : strlen(p));
            head += '\n';
            if (!nl) break;
            p = nl + 1;
        }
    }
    {   // first row: nan, orbphase / (2 pi)
        std::vector<char> row((size_t)(n_orb + 1) * 32 + 2);
        size_t n = put_cell(row.data(), NAN);
        const double two_pi = 2.0 * 3.141592653589793;
        for (int p = 0; p < n_orb; ++p) { row[n++] = ' '; n += put_cell(row.data() + n, orbphase[p] / two_pi); }
        row[n++] = '\n';
        head.append(row.data(), n);
    }
    bool ok = fwrite(head.data(), 1, head.size(), f) == head.size();

    if (n_threads <= 0) n_threads = (int)std::thread::hardware_concurrency();
    if (n_threads < 1) n_threads = 1;
    if (n_threads > 64) n_threads = 64;
    const size_t row_cap = (size_t)(n_orb + 1) * 26 + 1;          // a cell is at most 25 characters ("-1.234...e-308")
    const int64_t block = 4096;                                   // rows per work item
    std::vector<std::vector<char>> bufs(n_threads);
    for (auto& b : bufs) b.resize(row_cap * block);
    for (int64_t base = 0; base < n_wav && ok; base += block * n_threads) {
        std::vector<size_t> used(n_threads, 0);
        std::vector<std::thread> pool;
        for (int t = 0; t < n_threads; ++t) {
            const int64_t lo = base + (int64_t)t * block;
            if (lo >= n_wav) break;
            const int64_t hi = lo + block < n_wav ? lo + block : n_wav;
            pool.emplace_back([&, t, lo, hi] {
                char* dst = bufs[t].data();
                size_t n = 0;
                for (int64_t w = lo; w < hi; ++w) {
                    n += put_cell(dst + n, wavelength[w]);
                    for (int p = 0; p < n_orb; ++p) { dst[n++] = ' '; n += put_cell(dst + n, R[(int64_t)p * n_wav + w]); }
                    dst[n++] = '\n';
                }
                used[t] = n;
            });
        }
        for (auto& th : pool) th.join();
        for (size_t t = 0; t < pool.size() && ok; ++t) ok = fwrite(bufs[t].data(), 1, used[t], f) == used[t];
    }
    ok = (fclose(f) == 0) && ok;
    return ok ? PB200_OK : PB200_ERR_INVALID;
}
