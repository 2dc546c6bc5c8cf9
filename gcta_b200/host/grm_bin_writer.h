// grm_bin_writer.h - the .grm.bin of output_grm_vec / output_grm_MatrixXf (grm.cpp:307-318, 347-358): the full n x n
// matrix as float64, row by row, no header, filled symmetrically from the lower triangle.
//
// The file is n^2 * 8 bytes (20 GB at n = 50,000), so it is written by a pool of threads with pwrite, each owning
// bands of 16 rows.  The upper half of a row is a strided walk through the packed triangle (element (j, i) for j > i
// sits at j(j+1)/2 + i); a band therefore walks the columns once and scatters each column's 16 neighbouring elements -
// which are contiguous in the packed triangle - into its 16 row buffers.
#ifndef GCTA_B200_GRM_BIN_WRITER_H
#define GCTA_B200_GRM_BIN_WRITER_H

#include <fcntl.h>
#include <unistd.h>

#include <algorithm>
#include <atomic>
#include <cstdint>
#include <string>
#include <thread>
#include <vector>

// T = float (make_grm's result) or double (a GRM read back from disk).  Returns false if the file cannot be written.
template <typename T>
bool write_grm_bin_square(const std::string& path, const T* lower, int64_t n)
{
    const int fd = open(path.c_str(), O_WRONLY | O_CREAT | O_TRUNC, 0644);
    if (fd < 0) return false;
    constexpr int64_t BAND = 16;
    const int64_t nbands = (n + BAND - 1) / BAND;
    const int T_ = (int)std::max<int64_t>(1, std::min<int64_t>({(int64_t)32, (int64_t)std::thread::hardware_concurrency(), nbands}));
    std::atomic<int64_t> next(0);
    std::atomic<bool> failed(false);
    auto work = [&]() {
        std::vector<double> buf((size_t)(BAND * n));
        for (int64_t b = next++; b < nbands && !failed; b = next++) {
            const int64_t i0 = b * BAND, i1 = std::min(n, i0 + BAND), rows = i1 - i0;
            for (int64_t i = i0; i < i1; i++) {   // lower part and diagonal: contiguous in the packed triangle
                const T* src = lower + i * (i + 1) / 2;
                double* dst = &buf[(size_t)((i - i0) * n)];
                for (int64_t j = 0; j <= i; j++) dst[j] = (double)src[j];
            }
            for (int64_t j = i0 + 1; j < n; j++) {   // upper part: column j of the band = elements (j, i0 .. min(j, i1) - 1)
                const T* src = lower + j * (j + 1) / 2;
                const int64_t e = std::min(j, i1);
                for (int64_t i = i0; i < e; i++) buf[(size_t)((i - i0) * n + j)] = (double)src[i];
            }
            const char* p = (const char*)buf.data();
            size_t left = (size_t)(rows * n) * sizeof(double);
            off_t off = (off_t)(i0 * n) * (off_t)sizeof(double);
            while (left > 0) {
                const ssize_t w = pwrite(fd, p, left, off);
                if (w <= 0) { failed = true; break; }
                p += w; left -= (size_t)w; off += w;
            }
        }
    };
    std::vector<std::thread> th;
    for (int t = 1; t < T_; t++) th.emplace_back(work);
    work();
    for (auto& x : th) x.join();
    const bool ok = !failed && close(fd) == 0;
    if (failed) close(fd);
    return ok;
}

#endif
