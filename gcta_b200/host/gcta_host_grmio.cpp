// gcta_host_grmio.cpp - the consumer side of the .grm.gz: `--grm P` / `--mgrm list` with `--make-grm[-bin]`
// (save_grm -> manipulate_grm -> read_grm_gz / merge_grm -> output_grm_MatrixXf, grm.cpp:343-654), plus --keep / --remove /
// --grm-cutoff / --grm-adj / --dc on the way.  Host only, like the reference: O(n^2) text and arithmetic.
//
// Differences by design: the file is inflated in large blocks and parsed by a hand-written tokenizer instead of one
// stringstream per line; the matrices are kept as packed lower triangles (the reference allocates two full n x n); the
// output goes through the structure-aware gzip encoder.  Values, line format, log lines and errors follow the reference.
#include <zlib.h>

#include <algorithm>
#include <atomic>
#include <cmath>
#include <condition_variable>
#include <cstdio>
#include <cstring>
#include <deque>
#include <fstream>
#include <iostream>
#include <memory>
#include <mutex>
#include <thread>
#include <unordered_map>

#include "gcta_host.h"
#include "grm_bin_writer.h"
#include "gz_member_writer.h"
#include "grm_text_gz.h"

using std::cout;
using std::endl;
using std::string;
using std::vector;

double& gcta_host::grm_at(vector<double>& M, int64_t i, int64_t j)
{
    if (i < j) std::swap(i, j);
    if (j < 0 || i >= _grm_dim) throw string("Error: GRM index out of range (the IDs of the merged GRM files are not in the same order?).");
    return M[(size_t)(i * (i + 1) / 2 + j)];
}

namespace {

inline const char* skip_ws(const char* p, const char* e)
{
    while (p < e && (*p == ' ' || *p == '\t' || *p == '\r')) p++;
    return p;
}

// `ss >> int`: optional sign, digits; false if there is none
inline bool parse_int(const char*& p, const char* e, long& v)
{
    p = skip_ws(p, e);
    const char* q = p;
    bool neg = false;
    if (q < e && (*q == '-' || *q == '+')) { neg = *q == '-'; q++; }
    if (q >= e || *q < '0' || *q > '9') return false;
    long x = 0;
    while (q < e && *q >= '0' && *q <= '9') { x = x * 10 + (*q - '0'); q++; }
    v = neg ? -x : x;
    p = q;
    return true;
}

// `ss >> double`: decimal with optional fraction and exponent.  Up to 19 significant digits and a power of ten within
// 1e+-22 are converted exactly (one correctly rounded multiplication or division); anything else goes to strtod.
inline bool parse_double(const char*& p, const char* e, double& v)
{
    static const double P10[23] = {1e0, 1e1, 1e2, 1e3, 1e4, 1e5, 1e6, 1e7, 1e8, 1e9, 1e10, 1e11, 1e12, 1e13, 1e14, 1e15, 1e16, 1e17, 1e18, 1e19, 1e20, 1e21, 1e22};
    p = skip_ws(p, e);
    const char* q = p;
    bool neg = false;
    if (q < e && (*q == '-' || *q == '+')) { neg = *q == '-'; q++; }
    uint64_t m = 0;
    int nd = 0, dexp = 0;
    bool any = false, inexact = false;
    while (q < e && *q >= '0' && *q <= '9') { any = true; if (nd < 19) { m = m * 10 + (uint64_t)(*q - '0'); if (m) nd++; } else { dexp++; inexact = true; } q++; }
    if (q < e && *q == '.') {
        q++;
        while (q < e && *q >= '0' && *q <= '9') { any = true; if (nd < 19) { m = m * 10 + (uint64_t)(*q - '0'); if (m) nd++; dexp--; } else inexact = true; q++; }
    }
    if (!any) return false;
    if (q < e && (*q == 'e' || *q == 'E')) {
        const char* r = q + 1;
        bool eneg = false;
        if (r < e && (*r == '-' || *r == '+')) { eneg = *r == '-'; r++; }
        if (r < e && *r >= '0' && *r <= '9') {
            int x = 0;
            while (r < e && *r >= '0' && *r <= '9') { if (x < 100000) x = x * 10 + (*r - '0'); r++; }
            dexp += eneg ? -x : x;
            q = r;
        }
    }
    if (!inexact && m < (1ull << 53) && dexp >= -22 && dexp <= 22) {
        const double d = dexp >= 0 ? (double)m * P10[dexp] : (double)m / P10[-dexp];
        v = neg ? -d : d;
    } else {
        const string tok(p, q);
        v = strtod(tok.c_str(), nullptr);
        if (std::isinf(v)) return false;   // out of range: `ss >> double` sets failbit, the reference rejects the line
    }
    p = q;
    return true;
}

struct ParsedLine { long i1, i2; double N, A; };

// one line of the GRM text the way read_grm_gz takes it apart (grm.cpp:428-441): four tokens, nothing after them
inline bool parse_grm_line(const char* p, const char* e, ParsedLine& L)
{
    if (!parse_int(p, e, L.i1)) return false;
    if (!parse_int(p, e, L.i2)) return false;
    if (!parse_double(p, e, L.N)) return false;
    if (!parse_double(p, e, L.A)) return false;
    p = skip_ws(p, e);
    return p == e;
}

struct BlockResult {
    int64_t nline = 0;
    string err;                 // first bad line of the block (message as the reference throws it)
    vector<string> warnings;    // lines with N == 0 seen before it, in order
};

// Parses the complete lines of [p, end) into the packed triangles; `last` = nothing follows, so a final line without
// a newline counts too.  Returns where the unconsumed rest begins.  Stops at the first bad line (R.err).
const char* parse_block(const char* p, const char* end, bool last, int64_t n, double* grm, double* grm_N, const string& errmsg, BlockResult& R)
{
    for (;;) {
        const char* nl = (const char*)memchr(p, '\n', (size_t)(end - p));
        if (!nl) {
            if (!last || p == end) break;
            nl = end;
        }
        ParsedLine L;
        if (!parse_grm_line(p, nl, L) || L.i1 < L.i2 || L.i1 > n || L.i2 > n || L.i2 < 1) { R.err = errmsg + string(p, nl); return p; }
        if (L.N == 0) R.warnings.emplace_back(p, nl);
        const size_t o = (size_t)((L.i1 - 1) * L.i1 / 2 + (L.i2 - 1));
        grm_N[o] = L.N;
        grm[o] = L.A;
        R.nline++;
        p = nl < end ? nl + 1 : end;
    }
    return p;
}

// Member table of a file whose every gzip member announces its length (grm_text_gz.h); false for any other file.
bool scan_members(FILE* f, vector<std::pair<uint64_t, uint32_t>>& members)
{
    members.clear();
    if (fseeko(f, 0, SEEK_END) != 0) return false;
    const uint64_t total = (uint64_t)ftello(f);
    uint64_t pos = 0;
    unsigned char h[grm_text_gz::MEMBER_HEADER_BYTES];
    while (pos < total) {
        if (fseeko(f, (off_t)pos, SEEK_SET) != 0 || fread(h, 1, sizeof h, f) != sizeof h) { members.clear(); return false; }
        const uint32_t size = grm_text_gz::member_size_from_header(h, sizeof h);
        if (size < (uint32_t)sizeof h + 8 || pos + size > total) { members.clear(); return false; }
        members.emplace_back(pos, size);
        pos += size;
    }
    return !members.empty();
}

}  // namespace

// ---------------------------------------------------------------------------------------------
// read_grm_gz (grm.cpp:381-448)
// ---------------------------------------------------------------------------------------------
void gcta_host::read_grm_gz(const string& grm_file, vector<string>& grm_id, bool out_id_log, bool read_id_only)
{
    const string grm_id_file = grm_file + ".grm.id";
    if (out_id_log) cout << "Reading IDs of the genetic relationship matrix (GRM) from [" + grm_id_file + "]." << endl;
    std::ifstream i_grm_id(grm_id_file.c_str());
    if (!i_grm_id) throw("Error: can not open the file [" + grm_id_file + "] to read.");
    vector<string> fid, pid;
    grm_id.clear();
    // FID IID, rest of the line ignored.  One buffer for every extraction and the loop exit of the reference
    // (grm.cpp:392-402), so that a last line with a single token behaves the same: dropped when the file ends right
    // after it, otherwise completed with that same token
    string tok;
    while (i_grm_id) {
        i_grm_id >> tok;
        if (i_grm_id.eof()) break;
        fid.push_back(tok);
        string id = tok + ":";
        i_grm_id >> tok;
        pid.push_back(tok);
        id += tok;
        grm_id.push_back(id);
        std::getline(i_grm_id, tok);
    }
    i_grm_id.close();
    const int64_t n = (int64_t)grm_id.size();
    if (out_id_log) cout << n << " IDs read from [" + grm_id_file + "]." << endl;
    if (_id_map.empty()) {
        _fid = fid;
        _pid = pid;
        _indi_num = (int)_fid.size();
        _sex.resize(_fid.size());
        init_keep();
    }
    if (read_id_only) return;

    const string grm_gzfile = grm_file + ".grm.gz";
    FILE* raw = fopen(grm_gzfile.c_str(), "rb");
    if (!raw) throw("Error: can not open the file [" + grm_gzfile + "] to read.");
    const string errmsg = "Error: failed to read [" + grm_gzfile + "]. The format of the GRM file has been changed?\nError occurs in line:\n";
    cout << "Reading the GRM from [" + grm_gzfile + "]." << endl;
    _grm_dim = n;
    const size_t tri = (size_t)(n * (n + 1) / 2);
    _grm.assign(tri, 0.0);
    _grm_N.assign(tri, 0.0);
    int64_t nline = 0;
    // Files written by gcta64_b200 carry each gzip member's length in its header: then the members are inflated and
    // parsed by all cores.  Any other .grm.gz (the reference's single stream) is decoded sequentially.
    vector<std::pair<uint64_t, uint32_t>> members;
    const bool indexed = scan_members(raw, members);
    fclose(raw);
    if (indexed) {
        const int fd_T = (int)std::max(1u, std::min(32u, std::thread::hardware_concurrency()));
        vector<BlockResult> res(members.size());
        std::atomic<size_t> next(0);
        std::atomic<bool> not_line_aligned(false);
        vector<std::thread> th;
        for (int t = 0; t < fd_T && (size_t)t < members.size(); t++)
            th.emplace_back([&]() {
                FILE* f = fopen(grm_gzfile.c_str(), "rb");
                vector<unsigned char> comp;
                vector<char> text;
                for (size_t b = next++; f && b < members.size(); b = next++) {
                    BlockResult& R = res[b];
                    comp.resize(members[b].second);
                    if (fseeko(f, (off_t)members[b].first, SEEK_SET) != 0 || fread(comp.data(), 1, comp.size(), f) != comp.size()) { R.err = "Error: failed to read [" + grm_gzfile + "]."; continue; }
                    const unsigned char* tail = comp.data() + comp.size() - 4;
                    const size_t isize = (size_t)tail[0] | ((size_t)tail[1] << 8) | ((size_t)tail[2] << 16) | ((size_t)tail[3] << 24);
                    text.resize(isize + 1);
                    z_stream s;
                    memset(&s, 0, sizeof s);
                    if (inflateInit2(&s, 15 + 16) != Z_OK) { R.err = "Error: zlib inflateInit2 failed."; continue; }
                    s.next_in = comp.data(); s.avail_in = (uInt)comp.size();
                    s.next_out = (Bytef*)text.data(); s.avail_out = (uInt)text.size();
                    const int rc = inflate(&s, Z_FINISH);
                    const size_t got = (size_t)s.total_out;
                    inflateEnd(&s);
                    if (rc != Z_STREAM_END || got != isize) { R.err = "Error: failed to read [" + grm_gzfile + "] (corrupt gzip stream)."; continue; }
                    if (got > 0 && text[got - 1] != '\n') { not_line_aligned = true; continue; }
                    parse_block(text.data(), text.data() + got, true, n, _grm.data(), _grm_N.data(), errmsg, R);
                }
                if (f) fclose(f);
            });
        for (auto& x : th) x.join();
        if (!not_line_aligned) {
            for (const BlockResult& R : res) {
                for (const string& w : R.warnings) cout << "Warning: " << w << endl;
                if (!R.err.empty()) throw R.err;
                nline += R.nline;
            }
        } else {
            members.clear();   // members that split a line: start over, sequentially
            std::fill(_grm.begin(), _grm.end(), 0.0);
            std::fill(_grm_N.begin(), _grm_N.end(), 0.0);
        }
    }
    if (members.empty()) {
        gzFile zf = gzopen(grm_gzfile.c_str(), "rb");
        if (!zf) throw("Error: can not open the file [" + grm_gzfile + "] to read.");
        gzbuffer(zf, 1u << 20);
        // One stream can only be inflated by one thread; this thread does that and cuts the text into blocks at line
        // ends, a few workers parse the blocks.  Results are kept per block so that warnings and the first error come
        // out in file order.
        constexpr size_t BLOCK = 8u << 20;
        struct Blk { vector<char> data; size_t len = 0; int64_t idx = 0; bool last = false; BlockResult R; };
        std::deque<std::unique_ptr<Blk>> all;   // every block in file order (results live here)
        std::deque<Blk*> todo;
        std::mutex mu;
        std::condition_variable cv_todo, cv_room;
        bool done_reading = false;
        std::atomic<int64_t> first_bad(INT64_MAX);   // index of the earliest block with a bad line: later blocks are skipped
        size_t in_flight = 0;
        const int W = (int)std::max(1u, std::min(8u, std::thread::hardware_concurrency()));
        vector<std::thread> workers;
        for (int t = 0; t < W; t++)
            workers.emplace_back([&]() {
                for (;;) {
                    Blk* b;
                    {
                        std::unique_lock<std::mutex> lk(mu);
                        cv_todo.wait(lk, [&] { return !todo.empty() || done_reading; });
                        if (todo.empty()) return;
                        b = todo.front();
                        todo.pop_front();
                    }
                    if (b->idx < first_bad.load()) {
                        parse_block(b->data.data(), b->data.data() + b->len, b->last, n, _grm.data(), _grm_N.data(), errmsg, b->R);
                        if (!b->R.err.empty()) {
                            int64_t cur = first_bad.load();
                            while (b->idx < cur && !first_bad.compare_exchange_weak(cur, b->idx)) {}
                        }
                    }
                    vector<char>().swap(b->data);
                    { std::lock_guard<std::mutex> lk(mu); in_flight--; }
                    cv_room.notify_one();
                }
            });
        string read_error;
        vector<char> carry;
        bool eof = false;
        while (!eof && first_bad.load() == INT64_MAX) {
            std::unique_ptr<Blk> b(new Blk());
            b->data.resize(BLOCK + carry.size());
            if (!carry.empty()) memcpy(b->data.data(), carry.data(), carry.size());
            size_t have = carry.size();
            carry.clear();
            while (have < b->data.size()) {
                const int got = gzread(zf, b->data.data() + have, (unsigned)(b->data.size() - have));
                if (got < 0) { read_error = "Error: failed to read [" + grm_gzfile + "] (corrupt gzip stream)."; eof = true; break; }
                if (got == 0) { eof = true; break; }
                have += (size_t)got;
            }
            if (!read_error.empty()) break;
            size_t cutpos = have;
            if (!eof) {   // keep the unfinished last line for the next block
                while (cutpos > 0 && b->data[cutpos - 1] != '\n') cutpos--;
                if (cutpos == 0) { read_error = errmsg + string(b->data.data(), b->data.data() + 200); break; }   // a "line" longer than a block
                carry.assign(b->data.begin() + (std::ptrdiff_t)cutpos, b->data.begin() + (std::ptrdiff_t)have);
            }
            b->len = cutpos;
            b->last = eof;
            b->idx = (int64_t)all.size();
            {
                std::unique_lock<std::mutex> lk(mu);
                cv_room.wait(lk, [&] { return in_flight < (size_t)(2 * W); });
                in_flight++;
                todo.push_back(b.get());
                all.push_back(std::move(b));
            }
            cv_todo.notify_one();
        }
        {
            std::lock_guard<std::mutex> lk(mu);
            done_reading = true;
        }
        cv_todo.notify_all();
        for (auto& w : workers) w.join();
        gzclose(zf);
        for (const auto& b : all) {
            for (const string& w : b->R.warnings) cout << "Warning: " << w << endl;
            if (!b->R.err.empty()) throw b->R.err;
            nline += b->R.nline;
        }
        if (!read_error.empty()) throw read_error;
    }
    if (nline != (int64_t)tri) throw("Error: incorrect number of lines in the grm file. *.grm.gz file and *.grm.id file are mismatched?");
    cout << "Pairwise genetic relationships between " << n << " individuals are included from [" + grm_gzfile + "]." << endl;
}

// ---------------------------------------------------------------------------------------------
// merge_grm (grm.cpp:608-654): A = sum_f A_f N_f / sum_f N_f over the individuals common to all files
// ---------------------------------------------------------------------------------------------
static void read_grm_filenames(const string& merge_grm_file, vector<string>& grm_files)   // grm.cpp:656-672
{
    std::ifstream in(merge_grm_file.c_str());
    if (!in) throw("Error: can not open the file [" + merge_grm_file + "] to read.");
    grm_files.clear();
    string line;
    while (std::getline(in, line)) {
        // a line counts when it holds exactly one token (separators " ,\t;\n", StrFunc::split_string)
        vector<string> tok;
        string cur;
        for (char ch : line) {
            if (ch == ' ' || ch == ',' || ch == '\t' || ch == ';' || ch == '\n' || ch == '\r') { if (!cur.empty()) { tok.push_back(cur); cur.clear(); } }
            else cur += ch;
        }
        if (!cur.empty()) tok.push_back(cur);
        if (tok.size() == 1) grm_files.push_back(tok[0]);
    }
    cout << "There are " << grm_files.size() << " GRM file names specified in [" + merge_grm_file + "]." << endl;
    if (grm_files.size() > 1000) throw("Error: too many GRM file names specified in [" + merge_grm_file + "]. Maximum is 1000.");
    if (grm_files.size() < 1) throw("Error: no GRM file name is found in [" + merge_grm_file + "].");
}

void gcta_host::merge_grm(const string& merge_grm_file)
{
    vector<string> grm_files, grm_id;
    read_grm_filenames(merge_grm_file, grm_files);
    for (const string& f : grm_files) {
        read_grm_gz(f, grm_id, false, true);
        update_id_map_kp(grm_id, _id_map, _keep);
    }
    vector<string> uni_id;
    for (int i : _keep) uni_id.push_back(_fid[i] + ":" + _pid[i]);
    const int64_t n = (int64_t)uni_id.size();
    if (n == 0) throw("Error: no individual is in common in the GRM files.");
    cout << n << " individuals in common in the GRM files." << endl;

    const size_t tri = (size_t)(n * (n + 1) / 2);
    vector<double> grm(tri, 0.0), grm_N(tri, 0.0);
    for (size_t f = 0; f < grm_files.size(); f++) {
        cout << "Reading the GRM from the " << f + 1 << "th file ..." << endl;
        read_grm_gz(grm_files[f], grm_id);
        std::unordered_map<string, int> pos;
        for (size_t k = 0; k < grm_id.size(); k++) pos.emplace(grm_id[k], (int)k);   // first occurrence, like map::insert
        vector<int64_t> kp((size_t)n);
        for (int64_t i = 0; i < n; i++) kp[(size_t)i] = pos.at(uni_id[(size_t)i]);   // common IDs are in every file
        const int T = (int)std::max(1u, std::min(32u, std::thread::hardware_concurrency()));
        vector<std::thread> th;
        for (int t = 0; t < T; t++)
            th.emplace_back([&, t]() {
                for (int64_t i = t; i < n; i += T) {
                    const int64_t a = kp[(size_t)i];
                    size_t o = (size_t)(i * (i + 1) / 2);
                    for (int64_t j = 0; j <= i; j++, o++) {
                        const int64_t b = kp[(size_t)j];
                        const size_t s = a >= b ? (size_t)(a * (a + 1) / 2 + b) : (size_t)(b * (b + 1) / 2 + a);
                        grm[o] += _grm[s] * _grm_N[s];
                        grm_N[o] += _grm_N[s];
                    }
                }
            });
        for (auto& x : th) x.join();
    }
    for (size_t o = 0; o < tri; o++) {
        grm[o] = grm_N[o] == 0 ? 0.0 : grm[o] / grm_N[o];
    }
    _grm.swap(grm);
    _grm_N.swap(grm_N);
    _grm_dim = n;
    cout << "\n" << grm_files.size() << " GRMs have been merged together." << endl;
}

// ---------------------------------------------------------------------------------------------
// --grm-cutoff (grm.cpp:450-493): of every pair above the cutoff, drop the individual that is in more such pairs
// ---------------------------------------------------------------------------------------------
void gcta_host::rm_cor_indi(double grm_cutoff)
{
    cout << "Pruning the GRM with a cutoff of " << grm_cutoff << " ..." << endl;
    vector<int> id1, id2;
    for (size_t i = 0; i < _keep.size(); i++)
        for (size_t j = 0; j < i; j++)
            if (grm_at(_grm, _keep[i], _keep[j]) > grm_cutoff) { id1.push_back(_keep[i]); id2.push_back(_keep[j]); }
    std::unordered_map<int, int> count;   // pairs an individual takes part in (count() over both lists in the reference)
    for (int v : id1) count[v]++;
    for (int v : id2) count[v]++;
    for (size_t k = 0; k < id1.size(); k++)
        if (count[id1[k]] < count[id2[k]]) std::swap(id1[k], id2[k]);
    std::stable_sort(id1.begin(), id1.end());
    id1.erase(std::unique(id1.begin(), id1.end()), id1.end());
    vector<string> removed;
    for (int v : id1) removed.push_back(_fid[v] + ":" + _pid[v]);
    update_id_map_rm(removed, _id_map, _keep);
    cout << "After pruning the GRM, there are " << _keep.size() << " individuals (" << removed.size() << " individuals removed)." << endl;
}

// ---------------------------------------------------------------------------------------------
// --grm-adj (grm.cpp:495-522): shrink towards the expectation under sampling error
// ---------------------------------------------------------------------------------------------
void gcta_host::adj_grm(double adj_grm_fac)
{
    cout << "Adjusting the GRM for sampling errors ..." << endl;
    const int64_t n = (int64_t)_keep.size();
    double off_mean = 0.0, diag_mean = 0.0, off_var = 0.0, diag_var = 0.0;
    for (int64_t i = 0; i < n; i++) {   // same summation order as the reference: the sums are rounded in this order
        diag_mean += grm_at(_grm, _keep[i], _keep[i]);
        for (int64_t j = 0; j < i; j++) off_mean += grm_at(_grm, _keep[i], _keep[j]);
    }
    diag_mean /= n;
    off_mean /= 0.5 * n * (n - 1.0);
    for (int64_t i = 0; i < n; i++) {
        double d = grm_at(_grm, _keep[i], _keep[i]) - diag_mean;
        diag_var += d * d;
        for (int64_t j = 0; j < i; j++) {
            d = grm_at(_grm, _keep[i], _keep[j]) - off_mean;
            off_var += d * d;
        }
    }
    diag_var /= n - 1.0;
    off_var /= 0.5 * n * (n - 1.0) - 1.0;
    for (int64_t i = 0; i < n; i++) {
        double& gii = grm_at(_grm, _keep[i], _keep[i]);
        const double d = 1.0 - (adj_grm_fac + 1.0 / grm_at(_grm_N, _keep[i], _keep[i])) / diag_var;
        if (gii > 0) gii = 1.0 + d * (gii - 1.0);
        for (int64_t j = 0; j < i; j++) {
            const double nij = grm_at(_grm_N, _keep[i], _keep[j]);
            if (nij > 0) grm_at(_grm, _keep[i], _keep[j]) *= 1.0 - (adj_grm_fac + 1.0 / nij) / off_var;
        }
    }
}

// ---------------------------------------------------------------------------------------------
// --dc (grm.cpp:524-541).  Like the reference, elements are addressed by position (i, j) while the sexes are taken
// from _keep[i], _keep[j]; the two agree unless individuals were dropped before this step.
// ---------------------------------------------------------------------------------------------
void gcta_host::dc(int dosage_compen)
{
    cout << "Parameterizing the GRM under the assumption of ";
    if (dosage_compen == 1) cout << "full dosage compensation ..." << endl;
    else if (dosage_compen == 0) cout << "no dosage compensation ..." << endl;
    double c1 = 1.0, c2 = 1.0;
    if (dosage_compen == 1) { c1 = 2.0; c2 = sqrt(2.0); }
    else if (dosage_compen == 0) { c1 = 0.5; c2 = sqrt(0.5); }
    for (size_t i = 0; i < _keep.size(); i++)
        for (size_t j = 0; j <= i; j++) {
            const int s = _sex[_keep[i]] * _sex[_keep[j]];
            if (s == 1) grm_at(_grm, (int64_t)i, (int64_t)j) *= c1;
            else if (s == 2) grm_at(_grm, (int64_t)i, (int64_t)j) *= c2;
        }
}

// ---------------------------------------------------------------------------------------------
// manipulate_grm / save_grm (grm.cpp:544-577)
// ---------------------------------------------------------------------------------------------
void gcta_host::manipulate_grm(const string& grm_file, const string& keep_indi_file, const string& remove_indi_file, const string& sex_file,
                               double grm_cutoff, double adj_grm_fac, int dosage_compen, bool merge_grm_flag)
{
    vector<string> grm_id;
    if (merge_grm_flag) merge_grm(grm_file);
    else read_grm_gz(grm_file, grm_id);
    if (!keep_indi_file.empty()) keep_indi(keep_indi_file);
    if (!remove_indi_file.empty()) remove_indi(remove_indi_file);
    if (grm_cutoff > -1.0) rm_cor_indi(grm_cutoff);
    if (!sex_file.empty()) update_sex(sex_file);
    if (adj_grm_fac > -1.0) adj_grm(adj_grm_fac);
    if (dosage_compen > -1) dc(dosage_compen);
    if (grm_cutoff > -1.0 || !keep_indi_file.empty() || !remove_indi_file.empty()) {
        // keep the rows / columns of the individuals that are left (ascending, so the result is again a lower triangle)
        const int64_t nk = (int64_t)_keep.size();
        vector<double> g((size_t)(nk * (nk + 1) / 2)), gn(g.size());
        size_t o = 0;
        for (int64_t i = 0; i < nk; i++)
            for (int64_t j = 0; j <= i; j++, o++) {
                g[o] = grm_at(_grm, _keep[i], _keep[j]);
                gn[o] = grm_at(_grm_N, _keep[i], _keep[j]);
            }
        _grm.swap(g);
        _grm_N.swap(gn);
        _grm_dim = nk;
    }
}

void gcta_host::save_grm(const string& grm_file, const string& keep_indi_file, const string& remove_indi_file, const string& sex_file,
                         double grm_cutoff, double adj_grm_fac, int dosage_compen, bool merge_grm_flag, bool output_grm_bin)
{
    if (dosage_compen > -1) check_sex();   // nobody is loaded yet at this point in the reference either (grm.cpp:574)
    manipulate_grm(grm_file, keep_indi_file, remove_indi_file, sex_file, grm_cutoff, adj_grm_fac, dosage_compen, merge_grm_flag);
    output_grm_g15(output_grm_bin);
}

// ---------------------------------------------------------------------------------------------
// output_grm_MatrixXf (grm.cpp:343-380): elements by position, "i \t j \t (int)N \t A" with A at precision 15
// ---------------------------------------------------------------------------------------------
void gcta_host::output_grm_g15(bool output_grm_bin)
{
    const int64_t n = (int64_t)_keep.size();
    if (n > _grm_dim) throw string("Error: GRM index out of range (the IDs of the merged GRM files are not in the same order?).");
    string grm_file;
    if (output_grm_bin) {
        grm_file = _out + ".grm.bin";
        if (!write_grm_bin_square(grm_file, _grm.data(), n)) throw("Error: can not open the file [" + grm_file + "] to write.");
        cout << "GRM of " << n << " individuals has been saved in the file [" + grm_file + "] (in binary format)." << endl;
    } else {
        grm_file = _out + ".grm.gz";
        {
            FILE* f = fopen(grm_file.c_str(), "wb");
            if (!f) throw("Error: can not open the file [" + grm_file + "] to write.");
            fclose(f);
        }
        cout << "Saving the genetic relationship matrix to the file [" + grm_file + "] (in compressed text format)." << endl;
        // same scheme as gcta_host::output_grm: blocks of ~1M lines encoded and written by a pool of threads
        vector<std::pair<int64_t, int64_t>> blocks;
        for (int64_t i0 = 0; i0 < n;) {
            int64_t i1 = i0, lines = 0;
            while (i1 < n && (lines == 0 || lines + i1 + 1 <= (1 << 20))) { lines += i1 + 1; i1++; }
            blocks.emplace_back(i0, i1);
            i0 = i1;
        }
        const string err = write_gz_members(
            grm_file, blocks.size(),
            [&](size_t b, string& member) { grm_text_gz::encode_rows_g15(_grm.data(), _grm_N.data(), blocks[b].first, blocks[b].second, member); },
            "Error: can not open the file [" + grm_file + "] to write.", "Error: can not write to the file [" + grm_file + "].");
        if (!err.empty()) throw err;
        cout << "The genetic relationship matrix has been saved in the file [" + grm_file + "] (in compressed text format)." << endl;
    }
    const string famfile = _out + ".grm.id";
    std::ofstream fam(famfile.c_str());
    if (!fam) throw("Error: can not open the file [" + famfile + "] to write.");
    for (int i : _keep) fam << _fid[i] + "\t" + _pid[i] << endl;
    fam.close();
    cout << "IDs for the GRM file [" + grm_file + "] have been saved in the file [" + famfile + "]." << endl;
}
