// gz_member_writer.h - the ".grm.gz" file as a sequence of independent gzip members (grm_text_gz.h) produced by a pool
// of threads and written by the same threads.
//
// A member's place in the file is known once the sizes of all members before it are; placed members queue up and any
// worker pwrites them before it encodes another block, so formatting, compression and the copies into the page cache
// all run on all threads (one writing thread would cap the file at the ~1 GB/s a single core copies into the page
// cache).  At most `window` members are ahead of the first unplaced one.
#ifndef GCTA_B200_GZ_MEMBER_WRITER_H
#define GCTA_B200_GZ_MEMBER_WRITER_H

#include <fcntl.h>
#include <unistd.h>

#include <algorithm>
#include <condition_variable>
#include <cstdint>
#include <deque>
#include <exception>
#include <mutex>
#include <string>
#include <thread>
#include <utility>
#include <vector>

// produce(b, member) fills member b (it may throw std::string or std::exception).  Returns "" or the first error;
// `cannot_open` / `cannot_write` are the messages for the two I/O failures.
template <class Produce>
std::string write_gz_members(const std::string& path, size_t nb, Produce produce, const std::string& cannot_open, const std::string& cannot_write)
{
    const int fd = open(path.c_str(), O_WRONLY | O_CREAT | O_TRUNC, 0644);
    if (fd < 0) return cannot_open;
    const int T = (int)std::max<size_t>(1, std::min<size_t>({(size_t)32, (size_t)std::thread::hardware_concurrency(), nb}));
    const size_t window = (size_t)(4 * T);   // members of ~10 MB: a few hundred MB in flight
    std::vector<std::string> gz(nb);
    std::vector<char> done(nb, 0);
    std::deque<std::pair<size_t, uint64_t>> to_write;   // placed members: (index, file offset)
    std::mutex mu;
    std::condition_variable cv;
    size_t next_block = 0, next_place = 0, producing = 0;
    uint64_t offset = 0;
    std::string first_error;
    auto work = [&]() {
        std::unique_lock<std::mutex> lk(mu);
        for (;;) {
            cv.wait(lk, [&] {
                return !first_error.empty() || !to_write.empty() || (next_block < nb && next_block < next_place + window) ||
                       (next_block >= nb && producing == 0);
            });
            if (!first_error.empty()) return;
            if (!to_write.empty()) {   // writing first: it frees memory and never waits for anything
                const std::pair<size_t, uint64_t> job = to_write.front();
                to_write.pop_front();
                std::string data;
                data.swap(gz[job.first]);
                lk.unlock();
                const char* p = data.data();
                size_t left = data.size();
                off_t off = (off_t)job.second;
                while (left > 0) {
                    const ssize_t w = pwrite(fd, p, left, off);
                    if (w <= 0) break;
                    p += w; left -= (size_t)w; off += w;
                }
                lk.lock();
                if (left > 0) {
                    if (first_error.empty()) first_error = cannot_write;
                    cv.notify_all();
                    return;
                }
                continue;
            }
            if (next_block < nb && next_block < next_place + window) {
                const size_t b = next_block++;
                producing++;
                lk.unlock();
                std::string member, err;
                try { produce(b, member); }
                catch (const std::string& e) { err = e; }
                catch (const std::exception& e) { err = std::string("Error: ") + e.what(); }
                lk.lock();
                producing--;
                if (!err.empty()) {
                    if (first_error.empty()) first_error = err;
                    cv.notify_all();
                    return;
                }
                gz[b].swap(member);
                done[b] = 1;
                while (next_place < nb && done[next_place]) {   // every member whose offset is now settled
                    to_write.emplace_back(next_place, offset);
                    offset += gz[next_place].size();
                    next_place++;
                }
                cv.notify_all();
                continue;
            }
            return;   // nothing left to encode, nothing being encoded, nothing to write
        }
    };
    std::vector<std::thread> th;
    for (int t = 1; t < T; t++) th.emplace_back(work);
    work();
    for (auto& x : th) x.join();
    if (close(fd) != 0 && first_error.empty()) first_error = cannot_write;
    return first_error;
}

#endif
