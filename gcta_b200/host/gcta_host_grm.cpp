// gcta_host_grm.cpp - allele frequencies, MAF filters, make_grm dispatch to the device, writers.
#include <zlib.h>

#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <fstream>
#include <iomanip>
#include <iostream>
#include <chrono>
#include <thread>

#include "../../include/gcta_b200.h"
#include "gcta_host.h"

using std::cout;
using std::endl;
using std::string;
using std::vector;

// ---------------------------------------------------------------------------------------------
// allele frequencies (data.cpp:759-803): counting on the device, the division here
// ---------------------------------------------------------------------------------------------
void gcta_host::calcu_mu()
{
    cout << "Calculating allele frequencies ..." << endl;
    const int64_t m = (int64_t)_include.size();
    vector<int32_t> keep(_keep.begin(), _keep.end()), inc(_include.begin(), _include.end()), sex(_sex.begin(), _sex.end());
    vector<int32_t> cnt(m), sum(m), cM(m), sM(m), cF(m), sF(m);
    gcta_b200_args a;
    memset(&a, 0, sizeof a);
    a.bed = _bed.data(); a.n_file = _indi_num; a.m_file = _snp_num;
    a.keep = keep.data(); a.n = (int64_t)keep.size();
    a.include = inc.data(); a.m = m;
    a.sex = sex.data();
    char err[512];
    if (gcta_b200_allele_counts(&a, cnt.data(), sum.data(), cM.data(), sM.data(), cF.data(), sF.data(), err, sizeof err) != 0)
        throw string(err);
    _mu.assign(_snp_num, 0.0);
    for (int64_t j = 0; j < m; j++) {
        const int id = _include[j];
        const bool flip = _allele2[id] == _ref_A[id];   // mu_func counts 2 - x then (data.cpp:794)
        double num, den;
        if (_chr[id] == _autosome_num + 1) {            // X: males weigh 0.5, females 1.0, others 0 (data.cpp:765-767, 776)
            const double xm = flip ? 2.0 * cM[j] - sM[j] : sM[j], xf = flip ? 2.0 * cF[j] - sF[j] : sF[j];
            num = 0.5 * xm + xf;
            den = 0.5 * cM[j] + cF[j];
        } else {                                        // autosomes weigh 1.0; chr > autosome_num+1 weighs 0.5 each: same quotient
            num = flip ? 2.0 * cnt[j] - sum[j] : sum[j];
            den = cnt[j];
        }
        if (den > 0.0) _mu[id] = num / den;
    }
}

void gcta_host::update_freq(const string& freq)
{
    std::ifstream in(freq.c_str());
    if (!in) throw("Error: can not open the file [" + freq + "] to read.");
    cout << "Reading allele frequencies of the SNPs from [" + freq + "]." << endl;
    _mu.assign(_snp_num, 0.0);
    vector<char> confirm(_snp_num, 0);
    int icount = 0;
    string snp, rest;
    char a = 0;
    double f = 0.0;
    while (in >> snp) {
        in >> a >> f;
        auto it = _snp_name_map.find(snp);
        if (it != _snp_name_map.end()) {
            if (f > 1.0 || f < 0.0) throw("Error: invalid value of allele frequency for the SNP " + snp + ".");
            if (a != _allele1[it->second] && a != _allele2[it->second])
                throw("Invalid allele type \"" + string(1, a) + "\" for the SNP " + _snp_name[it->second] + ".");
            _mu[it->second] = (a == _ref_A[it->second]) ? f * 2.0 : (1.0 - f) * 2.0;   // data.cpp:915-916
            confirm[it->second] = 1;
            icount++;
        }
        std::getline(in, rest);
    }
    for (int j : _include)
        if (!confirm[j]) throw("Error: frequencies of all of the included SNPs should be updated.");
    _mu_from_file = true;
    cout << "Allele frequency for " << icount << " SNPs are update from [" + freq + "]." << endl;
}

void gcta_host::filter_snp_maf(double maf)
{
    if (_mu.empty()) calcu_mu();
    cout << "Pruning SNPs with MAF > " << maf << " ..." << endl;
    const size_t prev = _include.size();
    std::map<string, int> kept;
    _include.clear();
    for (const auto& kv : _snp_name_map) {
        const double p = _mu[kv.second] * 0.5;
        if (p < maf || 1.0 - p < maf) continue;   // data.cpp:600-601
        kept.insert(kv);
        _include.push_back(kv.second);
    }
    _snp_name_map.swap(kept);
    if (_include.empty()) throw("Error: No SNP is retained for analysis.");
    std::stable_sort(_include.begin(), _include.end());
    cout << "After pruning SNPs with MAF > " << maf << ", there are " << _include.size() << " SNPs (" << prev - _include.size() << " SNPs with MAF < " << maf << ")." << endl;
}

void gcta_host::filter_snp_max_maf(double max_maf)
{
    if (_mu.empty()) calcu_mu();
    cout << "Pruning SNPs with MAF < " << max_maf << " ..." << endl;
    const size_t prev = _include.size();
    std::map<string, int> kept;
    _include.clear();
    for (const auto& kv : _snp_name_map) {
        const double p = _mu[kv.second] * 0.5;
        if (p > max_maf && 1.0 - p > max_maf) continue;   // data.cpp:624-625
        kept.insert(kv);
        _include.push_back(kv.second);
    }
    _snp_name_map.swap(kept);
    if (_include.empty()) throw("Error: No SNP is retained for analysis.");
    std::stable_sort(_include.begin(), _include.end());
    cout << "After pruning SNPs with MAF < " << max_maf << ", there are " << _include.size() << " SNPs (" << prev - _include.size() << " SNPs with MAF > " << max_maf << ")." << endl;
}

void gcta_host::save_freq()
{
    if (_mu.empty()) calcu_mu();
    const string file = _out + ".freq";
    std::ofstream o(file.c_str());
    if (!o) throw("Error: can not open the file [" + file + "] to write.");
    cout << "Writing allele frequencies of " << _include.size() << " SNPs to [" + file + "]." << endl;
    for (int j : _include) o << _snp_name[j] << "\t" << _ref_A[j] << "\t" << std::setprecision(15) << _mu[j] * 0.5 << endl;   // data.cpp:938-941
    o.close();
    cout << "Allele frequencies of " << _include.size() << " SNPs have been saved in the file [" + file + "]." << endl;
}

// ---------------------------------------------------------------------------------------------
// guards (grm.cpp:15-34, data.cpp:1194-1196)
// ---------------------------------------------------------------------------------------------
void gcta_host::check_autosome() const
{
    for (int j : _include)
        if (_chr[j] > _autosome_num) throw("Error: this option is for the autosomal SNPs only.");
}

void gcta_host::check_chrX() const
{
    for (int j : _include)
        if (_chr[j] != _autosome_num + 1) throw("Error: this option is for SNPs on the X chromosome only.");
}

void gcta_host::check_sex() const
{
    for (int i : _keep)
        if (_sex[i] != 1 && _sex[i] != 2)
            throw("Error: Sex information of the individual \"" + _fid[i] + " " + _pid[i] + "\" is missing.\nUse --update-sex option to update the sex information of the individuals.");
}

void gcta_host::rm_high_ld()
{
    if (_rm_ld_cutoff < 0.0 || _include.size() == 1) return;
    throw("Error: --rm-high-ld is not available in gcta64_b200 (outside the accelerated --make-grm path).");
}

// ---------------------------------------------------------------------------------------------
// make_grm (grm.cpp:219-301): everything between the guards and the writer runs on the GPU(s)
// ---------------------------------------------------------------------------------------------
void gcta_host::make_grm(bool grm_xchr_flag, bool output_bin, int grm_mtd)
{
    if (grm_xchr_flag) check_chrX();
    else check_autosome();
    rm_high_ld();
    cout << "Recoding genotypes (individual major mode) ..." << endl;   // make_XMat's line, data.cpp:986
    const bool have_mu = !_mu.empty();
    if (!have_mu) cout << "Calculating allele frequencies ..." << endl;  // std_XMat -> calcu_mu, data.cpp:1018, 771
    if (grm_xchr_flag) check_sex();                                      // data.cpp:1047

    const int64_t n = (int64_t)_keep.size(), m = (int64_t)_include.size();
    vector<int32_t> keep(_keep.begin(), _keep.end()), inc(_include.begin(), _include.end()), sex(_sex.begin(), _sex.end());
    // 2p handed down only when it came from --update-freq; frequencies computed from the data are
    // recomputed on the device from the same individuals (identical integers -> identical doubles)
    vector<double> mu;
    if (_mu_from_file) {
        mu.resize(m);
        for (int64_t j = 0; j < m; j++) {
            const int id = _include[j];
            mu[j] = (_allele1[id] == _ref_A[id]) ? _mu[id] : 2.0 - _mu[id];   // the ABI counts allele 1
        }
    }
    if (grm_xchr_flag) cout << "\nCalculating the genetic relationship matrix for the X chromosome ... ((NOTE: default speed-optimized mode, may use huge RAM))" << endl;
    else cout << "\nCalculating the genetic relationship matrix ... (NOTE: default speed-optimized mode, may use huge RAM)" << endl;

    const size_t tri = (size_t)n * (size_t)(n + 1) / 2;
    vector<float> A(tri);
    vector<int32_t> N(output_bin ? 0 : tri);   // --make-grm-bin drops N (grm.cpp:307-318)
    const int G = std::max(1, std::min(n_gpus, gcta_b200_device_count() > 0 ? gcta_b200_device_count() : 1));
    vector<string> errs(G);
    vector<int> rcs(G, 0);
    vector<gcta_b200_out> outs(G);
    const auto t_c0 = std::chrono::steady_clock::now();
    auto work = [&](int g) {
        gcta_b200_args a;
        memset(&a, 0, sizeof a);
        a.bed = _bed.data(); a.n_file = _indi_num; a.m_file = _snp_num;
        a.keep = keep.data(); a.n = n; a.include = inc.data(); a.m = m;
        a.mu = mu.empty() ? nullptr : mu.data();
        a.sex = sex.data(); a.xchr = grm_xchr_flag; a.alg = grm_mtd; a.operand = operand;
        a.device = g; a.part_rank = g; a.part_world = G; a.chunk_snps = chunk_snps;
        gcta_b200_out& o = outs[g];
        memset(&o, 0, sizeof o);
        o.A_lower = A.data();
        o.N_lower = N.empty() ? nullptr : N.data();
        char err[512] = {0};
        rcs[g] = gcta_b200_make_grm(&a, &o, err, sizeof err);
        errs[g] = err;
    };
    if (G == 1) work(0);
    else {
        vector<std::thread> th;
        for (int g = 0; g < G; g++) th.emplace_back(work, g);
        for (auto& t : th) t.join();
    }
    for (int g = 0; g < G; g++)
        if (rcs[g]) throw errs[g];
    cout << n << " of " << n << " individuals.\r";   // the reference prints one such line per row (grm.cpp:297)
    cout << endl << endl;
    if (getenv("GCTA_B200_TIMING")) {
        std::cerr << "{\"compute_wall_s\": " << std::chrono::duration<double>(std::chrono::steady_clock::now() - t_c0).count() << "}" << endl;
        for (int g = 0; g < G; g++)
            std::cerr << "{\"gpu\": " << g << ", \"rows\": [" << outs[g].row_begin << ", " << outs[g].row_end << "], \"ms_h2d\": " << outs[g].ms_h2d
                      << ", \"ms_decode\": " << outs[g].ms_decode << ", \"ms_syrk\": " << outs[g].ms_syrk << ", \"ms_paircount\": " << outs[g].ms_paircount
                      << ", \"ms_finalize\": " << outs[g].ms_finalize << ", \"ms_d2h\": " << outs[g].ms_d2h << ", \"launches\": " << outs[g].launches << "}" << endl;
    }
    const auto t_w0 = std::chrono::steady_clock::now();
    output_grm(A, N, output_bin);
    if (getenv("GCTA_B200_TIMING"))
        std::cerr << "{\"write_s\": " << std::chrono::duration<double>(std::chrono::steady_clock::now() - t_w0).count() << "}" << endl;
}

// ---------------------------------------------------------------------------------------------
// writers (grm.cpp:303-341)
// ---------------------------------------------------------------------------------------------
// Fast formatters for the n(n+1)/2 text lines.  put_e6 writes exactly what printf("%.6e") writes (the reference's
// `setf(scientific); precision(6)`, grm.cpp:326-327): the value is scaled by an exact power of ten to 7 significant
// digits; whenever the scaled value is close enough to a rounding boundary for the one floating-point rounding of the
// scaling to matter (or the exponent is outside the exact-power range) it falls back to snprintf itself.
static inline char* put_uint(char* q, uint64_t v)
{
    char tmp[24];
    int k = 0;
    do { tmp[k++] = (char)('0' + v % 10); v /= 10; } while (v);
    while (k) *q++ = tmp[--k];
    return q;
}

static int put_e6(char* q, double v)
{
    static const double P10[23] = {1e0, 1e1, 1e2, 1e3, 1e4, 1e5, 1e6, 1e7, 1e8, 1e9, 1e10, 1e11, 1e12, 1e13, 1e14, 1e15, 1e16, 1e17, 1e18, 1e19, 1e20, 1e21, 1e22};
    char* const q0 = q;
    double a = std::fabs(v);
    if (!(a >= 1e-15 && a < 1e15)) {
        if (v == 0.0 && !std::signbit(v)) { memcpy(q, "0.000000e+00", 12); return 12; }
        return snprintf(q, 32, "%.6e", v);
    }
    int e = (int)std::floor(std::log10(a));
    for (int attempt = 0; attempt < 3; attempt++) {
        const int k = 6 - e;   // a * 10^k should lie in [1e6, 1e7)
        const double scaled = k >= 0 ? a * P10[k] : a / P10[-k];
        if (scaled < 1e6) { e--; continue; }
        if (scaled >= 1e7) { e++; continue; }
        const double fl = std::floor(scaled), frac = scaled - fl;
        if (std::fabs(frac - 0.5) < 1e-5) return snprintf(q0, 32, "%.6e", v);   // too close to a tie to decide here
        uint64_t d = (uint64_t)fl + (frac > 0.5 ? 1 : 0);
        int ee = e;
        if (d >= 10000000ull) { d = 1000000ull; ee++; }
        if (std::signbit(v)) *q++ = '-';
        char dig[7];
        for (int t = 6; t >= 0; t--) { dig[t] = (char)('0' + d % 10); d /= 10; }
        *q++ = dig[0]; *q++ = '.';
        for (int t = 1; t < 7; t++) *q++ = dig[t];
        *q++ = 'e';
        if (ee < 0) { *q++ = '-'; ee = -ee; } else *q++ = '+';
        *q++ = (char)('0' + ee / 10); *q++ = (char)('0' + ee % 10);
        return (int)(q - q0);
    }
    return snprintf(q0, 32, "%.6e", v);
}

// `gcta64_b200 --selftest-format`: put_e6 against snprintf on many values (tests/test_host_cli.py runs it)
int gcta_host_selftest_format()
{
    uint64_t s = 0x9E3779B97F4A7C15ull;
    auto next = [&]() { s ^= s << 13; s ^= s >> 7; s ^= s << 17; return s; };
    long bad = 0, total = 0;
    char a[40], b[40];
    auto check = [&](double v) {
        const int la = put_e6(a, v);
        a[la] = 0;
        snprintf(b, sizeof b, "%.6e", v);
        total++;
        if (strcmp(a, b) != 0 && bad++ < 5) fprintf(stderr, "mismatch: %s vs %s\n", a, b);
    };
    const double specials[] = {0.0, -0.0, 1.0, -1.0, 0.5, 1.5, 9.9999995, 9.99999949, 0.99999995, 1e-3, 9.999999e-4, 1e-15, 9.9e14, 1e15, 1e-16,
                               123456.75, 1234567.5, 12345675.0, 2.5e-7, 1.0000005, 1.00000049999, 3.0000015e-3};
    for (double v : specials) check(v);
    for (int i = 0; i < 4000000; i++) {
        const uint64_t r = next();
        const float f = (float)((double)(r >> 11) * (1.0 / 9007199254740992.0));          // U(0,1) as float
        const int ex = (int)(next() % 24) - 16;
        double v = (double)(float)std::ldexp((double)f, ex) * ((r & 1) ? -1.0 : 1.0);      // float-valued, like A
        check(v);
        if ((i & 7) == 0) check((double)(int64_t)(next() % 20000000) / 1e7 * ((r & 2) ? 1e-3 : 1.0));   // near 7-digit decimals
    }
    printf("format selftest: %ld values, %ld mismatches\n", total, bad);
    return bad ? 1 : 0;
}

static void gz_member(const string& text, string& out)
{
    z_stream s;
    memset(&s, 0, sizeof s);
    // Level 1 by default: deflate is ~75 % of the writer's time at zlib's default level 6 (what the reference's
    // gzopen(name, "w") uses, zfstream.cpp:51-76) and 3.4x faster at level 1 for ~23 % larger files; the decompressed text is the
    // same.  GCTA_B200_GZ_LEVEL=6 restores the reference's level.
    static const int level = getenv("GCTA_B200_GZ_LEVEL") ? std::max(1, std::min(9, atoi(getenv("GCTA_B200_GZ_LEVEL")))) : 1;
    if (deflateInit2(&s, level, Z_DEFLATED, 15 + 16 /* gzip wrapper */, 8, Z_DEFAULT_STRATEGY) != Z_OK)
        throw string("Error: zlib deflateInit2 failed.");
    out.resize(deflateBound(&s, (uLong)text.size()) + 64);
    s.next_in = (Bytef*)text.data();
    s.avail_in = (uInt)text.size();
    s.next_out = (Bytef*)&out[0];
    s.avail_out = (uInt)out.size();
    const int rc = deflate(&s, Z_FINISH);
    out.resize(s.total_out);
    deflateEnd(&s);
    if (rc != Z_STREAM_END) throw string("Error: zlib deflate failed.");
}

void gcta_host::output_grm(const vector<float>& A, const vector<int32_t>& N, bool output_bin)
{
    const int64_t n = (int64_t)_keep.size();
    string grm_file;
    if (output_bin) {
        // full n x n float64, row-major, symmetric fill, no header, no N (grm.cpp:309-317)
        grm_file = _out + ".grm.bin";
        FILE* f = fopen(grm_file.c_str(), "wb");
        if (!f) throw("Error: can not open the file [" + grm_file + "] to write.");
        vector<double> row(n);
        for (int64_t i = 0; i < n; i++) {
            for (int64_t j = 0; j <= i; j++) row[j] = A[(size_t)(i * (i + 1) / 2 + j)];
            for (int64_t j = i + 1; j < n; j++) row[j] = A[(size_t)(j * (j + 1) / 2 + i)];
            fwrite(row.data(), sizeof(double), (size_t)n, f);
        }
        fclose(f);
        cout << "GRM of " << n << " individuals has been saved in the file [" + grm_file + "] (in binary format)." << endl;
    } else {
        // gzip text "i \t j \t N \t A" (scientific, 6 digits), 1-based, j <= i (grm.cpp:321-331).  Written as
        // concatenated gzip members compressed in parallel; gzread - hence the reference's own reader
        // (read_grm_gz, grm.cpp:381) - decodes a multi-member file transparently.
        grm_file = _out + ".grm.gz";
        FILE* f = fopen(grm_file.c_str(), "wb");
        if (!f) throw("Error: can not open the file [" + grm_file + "] to write.");
        cout << "Saving the genetic relationship matrix to the file [" + grm_file + "] (in compressed text format)." << endl;
        const int T = (int)std::max(1u, std::min(32u, std::thread::hardware_concurrency()));
        const int64_t lines_per_block = 1 << 20;
        vector<std::pair<int64_t, int64_t>> blocks;   // row ranges of ~1M lines
        for (int64_t i0 = 0; i0 < n;) {
            int64_t i1 = i0, lines = 0;
            while (i1 < n && (lines == 0 || lines + i1 + 1 <= lines_per_block)) { lines += i1 + 1; i1++; }
            blocks.emplace_back(i0, i1);
            i0 = i1;
        }
        for (size_t b0 = 0; b0 < blocks.size(); b0 += T) {
            const size_t b1 = std::min(blocks.size(), b0 + T);
            vector<string> gz(b1 - b0);
            vector<string> errs(b1 - b0);
            vector<std::thread> th;
            for (size_t b = b0; b < b1; b++)
                th.emplace_back([&, b]() {
                    try {
                        string text;
                        char buf[96];
                        for (int64_t i = blocks[b].first; i < blocks[b].second; i++)
                            for (int64_t j = 0; j <= i; j++) {
                                const size_t o = (size_t)(i * (i + 1) / 2 + j);
                                char* q = buf;
                                q = put_uint(q, (uint64_t)(i + 1)); *q++ = '\t';
                                q = put_uint(q, (uint64_t)(j + 1)); *q++ = '\t';
                                if (N[o] < 0) { *q++ = '-'; q = put_uint(q, (uint64_t)(-(int64_t)N[o])); } else q = put_uint(q, (uint64_t)N[o]);
                                *q++ = '\t';
                                q += put_e6(q, (double)A[o]);
                                *q++ = '\n';
                                text.append(buf, (size_t)(q - buf));
                            }
                        gz_member(text, gz[b - b0]);
                    } catch (const string& e) { errs[b - b0] = e; }
                });
            for (auto& t : th) t.join();
            for (size_t k = 0; k < gz.size(); k++) {
                if (!errs[k].empty()) { fclose(f); throw errs[k]; }
                fwrite(gz[k].data(), 1, gz[k].size(), f);
            }
        }
        fclose(f);
        cout << "The genetic relationship matrix has been saved in the file [" + grm_file + "] (in compressed text format)." << endl;
    }
    const string famfile = _out + ".grm.id";
    std::ofstream fam(famfile.c_str());
    if (!fam) throw("Error: can not open the file [" + famfile + "] to write.");
    for (int i : _keep) fam << _fid[i] + "\t" + _pid[i] << endl;
    fam.close();
    cout << "IDs for the GRM file [" + grm_file + "] have been saved in the file [" + famfile + "]." << endl;
}
