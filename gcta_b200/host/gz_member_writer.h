// gz_member_writer.h - the ".grm.gz" file as a sequence of independent gzip members (grm_text_gz.h) produced by a pool
// of threads and written by the same threads.
//
// A member's place in the file is known once the sizes of all members before it are; placed members queue up, and
// whichever worker finds the queue non-empty and nobody writing becomes the writer until it has drained it, while the
// others go on encoding.  At most `window` members are ahead of the last one written; written-out buffers are reused.
#ifndef GCTA_B200_GZ_MEMBER_WRITER_H
#define GCTA_B200_GZ_MEMBER_WRITER_H

#include <fcntl.h>
#include <unistd.h>

#include <algorithm>
#include <condition_variable>
#include <cstdint>
#include <cstdlib>
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
    std::vector<std::string> pool;                      // written-out buffers, capacity kept: no fresh pages per member
    std::mutex mu;
    std::condition_variable cv;
    size_t next_block = 0, next_place = 0, producing = 0, written = 0;
    bool writing = false;
    const bool dry = getenv("GCTA_B200_GZ_DRY_RUN") != nullptr;   // measurements: everything but the writes
    uint64_t offset = 0;
    std::string first_error;
    auto work = [&]() {
        std::unique_lock<std::mutex> lk(mu);
        for (;;) {
            cv.wait(lk, [&] {
                return !first_error.empty() || (!writing && !to_write.empty()) || (next_block < nb && next_block < written + window) ||
                       (next_block >= nb && producing == 0 && !writing && to_write.empty());
            });
            if (!first_error.empty()) return;
            if (!writing && !to_write.empty()) {
                // One thread writes at a time and drains the queue; the others keep encoding.  (Buffered writes to one
                // file are serialised by the file system's inode lock: on a 16-core B200 host three or four threads
                // in pwrite at once left 12.1 GB of members at 7.4 s where the encoding alone takes 4.0 s.)
                writing = true;
                bool bad = false;
                while (!to_write.empty() && !bad) {
                    const std::pair<size_t, uint64_t> job = to_write.front();
                    to_write.pop_front();
                    std::string data;
                    data.swap(gz[job.first]);
                    lk.unlock();
                    const char* p = data.data();
                    size_t left = dry ? 0 : data.size();
                    off_t off = (off_t)job.second;
                    while (left > 0) {
                        const ssize_t w = pwrite(fd, p, left, off);
                        if (w <= 0) break;
                        p += w; left -= (size_t)w; off += w;
                    }
                    data.clear();
                    lk.lock();
                    bad = left > 0;
                    pool.emplace_back(std::move(data));
                    written++;
                    cv.notify_all();   // room for another block
                }
                writing = false;
                if (bad && first_error.empty()) first_error = cannot_write;
                cv.notify_all();
                if (bad) return;
                continue;
            }
            if (next_block < nb && next_block < written + window) {
                const size_t b = next_block++;
                producing++;
                std::string member, err;
                if (!pool.empty()) { member.swap(pool.back()); pool.pop_back(); }
                lk.unlock();
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
