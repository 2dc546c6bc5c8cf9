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
#include <condition_variable>
#include <memory>
#include <mutex>
#include <thread>

#include "../../include/gcta_b200.h"
#include "gcta_host.h"
#include "grm_bin_writer.h"
#include "gz_member_writer.h"
#include "grm_text_gz.h"

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
                throw("Invalid allele type \"\" for the SNP " + _snp_name[it->second] + ".");   // the reference's text: it means to quote the allele but quotes an empty string (data.cpp:912-913)
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
    // not value-initialised: at n = 50,000 zero-filling these 10 GB would cost seconds before the first kernel starts,
    // and every element of the band(s) is written by the library
    big_alloc::Array<float> A(tri);
    big_alloc::Array<int32_t> N(output_bin ? 0 : tri);   // --make-grm-bin drops N (grm.cpp:307-318)
    const int G = std::max(1, std::min(n_gpus, gcta_b200_device_count() > 0 ? gcta_b200_device_count() : 1));
    vector<gcta_b200_out> outs(G);
    const auto t_c0 = std::chrono::steady_clock::now();
    {
        // one call from the main thread, like the reference's (option.cpp:620-621).  With several GPUs the library cuts the
        // triangle into row bands, uploads 1/G of the packed rows per device and spreads them with peer copies.
        gcta_b200_args a;
        memset(&a, 0, sizeof a);
        a.bed = _bed.data(); a.n_file = _indi_num; a.m_file = _snp_num;
        a.keep = keep.data(); a.n = n; a.include = inc.data(); a.m = m;
        a.mu = mu.empty() ? nullptr : mu.data();
        a.sex = sex.data(); a.xchr = grm_xchr_flag; a.alg = grm_mtd; a.operand = operand;
        a.device = 0; a.part_rank = 0; a.part_world = 1; a.chunk_snps = chunk_snps;
        gcta_b200_out o;
        memset(&o, 0, sizeof o);
        o.A_lower = A.get();
        o.N_lower = N.get();
        char err[512] = {0};
        if (gcta_b200_make_grm_multi(&a, G, &o, outs.data(), err, sizeof err) != 0) throw string(err);
    }
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
    output_grm(A.get(), N.get(), output_bin);
    if (getenv("GCTA_B200_TIMING"))
        std::cerr << "{\"write_s\": " << std::chrono::duration<double>(std::chrono::steady_clock::now() - t_w0).count() << "}" << endl;
}

// ---------------------------------------------------------------------------------------------
// writers (grm.cpp:303-341)
// ---------------------------------------------------------------------------------------------
using grm_text_gz::put_e6;   // exact "%.6e" formatter (grm_text_gz.cpp)

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
    // put_g15 (the re-save path's "%.15g"): decimals of up to 7 digits parsed by strtod (what a .grm.gz holds), full
    // doubles (merged GRMs), and the notation switches at 1e-5 and 1e15
    auto check15 = [&](double v) {
        const int la = grm_text_gz::put_g15(a, v);
        a[la] = 0;
        snprintf(b, sizeof b, "%.15g", v);
        total++;
        if (strcmp(a, b) != 0 && bad++ < 5) fprintf(stderr, "mismatch (%%.15g): %s vs %s\n", a, b);
    };
    const double specials15[] = {0.0, -0.0, 1.0, -1.0, 10.0, 100.0, 1e14, 9.999999e14, 1e15, 1e-4, 1e-5, 9.999999e-5, 1.5, 0.1, 0.3, 1234567.0, 12345670.0,
                                 1.234567e-3, 123456.7, 1e-15, 1e-16, 1.0 / 3.0, 2.0 / 3.0, 1e22, 123456789012345.0, 1234567890123456.0};
    for (double v : specials15) check15(v);
    for (int i = 0; i < 2000000; i++) {
        char txt[40];
        const uint64_t r = next();
        snprintf(txt, sizeof txt, "%s%d.%06de%+03d", (r & 1) ? "-" : "", (int)(1 + next() % 9), (int)(next() % 1000000), (int)(next() % 30) - 15);
        check15(strtod(txt, nullptr));                                             // a value as read from a .grm.gz
        if ((i & 3) == 0) check15(std::ldexp((double)(next() >> 11), (int)(next() % 100) - 103));   // arbitrary doubles
        if ((i & 15) == 0) check15((double)(int64_t)(next() % 20000000) * 1e-3);   // short decimals in fixed notation
    }
    printf("format selftest: %ld values, %ld mismatches\n", total, bad);
    return bad ? 1 : 0;
}

// `gcta64_b200 --selftest-gz`: the structured encoder against zlib's inflate on awkward inputs (one line, constant
// fields that merge into long runs, NaN / inf / zero / huge values, negative and extreme N, row ranges that start
// mid-triangle); every member must inflate - CRC and length checked by zlib - to the formatter's text.
int gcta_host_selftest_gz()
{
    auto inflate_all = [](const string& gz, string& text) -> bool {
        z_stream s;
        memset(&s, 0, sizeof s);
        if (inflateInit2(&s, 15 + 16) != Z_OK) return false;
        s.next_in = (Bytef*)gz.data();
        s.avail_in = (uInt)gz.size();
        vector<char> buf(1 << 16);
        int rc = Z_OK;
        while (rc == Z_OK) {
            s.next_out = (Bytef*)buf.data();
            s.avail_out = (uInt)buf.size();
            rc = inflate(&s, Z_NO_FLUSH);
            text.append(buf.data(), buf.size() - s.avail_out);
        }
        const bool ok = rc == Z_STREAM_END && s.avail_in == 0;
        inflateEnd(&s);
        return ok;
    };
    uint64_t st = 0x2545F4914F6CDD1Dull;
    auto next = [&]() { st ^= st << 13; st ^= st >> 7; st ^= st << 17; return st; };
    int bad = 0, cases = 0;
    auto run = [&](const char* name, const vector<float>& A, const vector<int32_t>& N, int64_t r0, int64_t r1) {
        string text, gz, back;
        grm_text_gz::format_rows(A.data(), N.data(), r0, r1, text);
        grm_text_gz::encode_rows(A.data(), N.data(), r0, r1, gz);
        cases++;
        if (!inflate_all(gz, back) || back != text) { bad++; fprintf(stderr, "gz selftest: case %s rows [%lld,%lld) failed\n", name, (long long)r0, (long long)r1); }
    };
    const float specials[] = {0.0f, -0.0f, 1.0f, -1.0f, NAN, INFINITY, -INFINITY, 1e30f, -1e-30f, 1.17549435e-38f, 1e-45f, 3.4028235e38f, 9.9999995e-1f, 0.5f};
    const int32_t nspecial[] = {0, 1, -1, 9, 10, 99, 100, 2147483647, -2147483647 - 1, 1000000, 999999};
    for (int64_t n : {1, 2, 3, 17, 60, 300, 1500}) {
        const size_t tri = (size_t)(n * (n + 1) / 2);
        vector<float> A(tri);
        vector<int32_t> N(tri);
        std::fill(A.begin(), A.end(), 0.0f); std::fill(N.begin(), N.end(), 1000000);
        run("constant", A, N, 0, n);
        for (size_t o = 0; o < tri; o++) { A[o] = specials[next() % (sizeof specials / sizeof specials[0])]; N[o] = nspecial[next() % (sizeof nspecial / sizeof nspecial[0])]; }
        run("specials", A, N, 0, n);
        for (size_t o = 0; o < tri; o++) { A[o] = (float)((double)(int64_t)(next() % 2000001 - 1000000) * 1e-9); N[o] = 500000 - (int32_t)(next() % 3); }
        run("grm-like", A, N, 0, n);
        if (n >= 17) { run("mid-triangle", A, N, n / 2, n / 2 + 1); run("tail", A, N, n - 3, n); run("empty", A, N, 5, 5); }
    }
    {   // i and j passing 9,999 -> 10,000; values with and without sign, N constant, then changing in width
        const int64_t n = 10050, r0 = 9995, r1 = 10005;
        const size_t tri = (size_t)(n * (n + 1) / 2);
        vector<float> A(tri, 0.0f);
        vector<int32_t> N(tri, 1000000);
        for (size_t o = (size_t)(r0 * (r0 + 1) / 2); o < (size_t)(r1 * (r1 + 1) / 2); o++) A[o] = (float)((double)(int64_t)(next() % 2000001 - 1000000) * 1.7e-9);
        run("five-digit indices", A, N, r0, r1);
        for (size_t o = (size_t)(r0 * (r0 + 1) / 2); o < (size_t)(r1 * (r1 + 1) / 2); o++) N[o] = 100005 - (int32_t)(next() % 10);
        run("five-digit indices, N 99,996 .. 100,005", A, N, r0, r1);
    }
    printf("gz selftest: %d cases, %d failures\n", cases, bad);
    return bad ? 1 : 0;
}

static void gz_member(const string& text, string& out)
{
    z_stream s;
    memset(&s, 0, sizeof s);
    // Level 1 by default: deflate is ~75 % of the writer's time at zlib's default level 6 (what the reference's
    // gzopen(name, "w") uses, zfstream.cpp:51-76) and 3.4x faster at level 1 for ~23 % larger files; the decompressed text is the
    // same.  GCTA_B200_GZ_LEVEL=6 restores the reference's level.
    static const int level = getenv("GCTA_B200_GZ_LEVEL") ? std::max(0, std::min(9, atoi(getenv("GCTA_B200_GZ_LEVEL")))) : 1;
    if (deflateInit2(&s, level, Z_DEFLATED, 15 + 16 /* gzip wrapper */, 8, Z_DEFAULT_STRATEGY) != Z_OK)
        throw string("Error: zlib deflateInit2 failed.");
    // the same member-length subfield the built-in encoder writes (grm_text_gz.h), so these files are read in parallel too
    gz_header head;
    memset(&head, 0, sizeof head);
    Bytef extra[8] = {'G', 'M', 4, 0, 0, 0, 0, 0};
    head.extra = extra;
    head.extra_len = 8;
    head.os = 3;
    deflateSetHeader(&s, &head);
    out.resize(deflateBound(&s, (uLong)text.size()) + 64);
    s.next_in = (Bytef*)text.data();
    s.avail_in = (uInt)text.size();
    s.next_out = (Bytef*)&out[0];
    s.avail_out = (uInt)out.size();
    const int rc = deflate(&s, Z_FINISH);
    out.resize(s.total_out);
    deflateEnd(&s);
    if (rc != Z_STREAM_END) throw string("Error: zlib deflate failed.");
    if (grm_text_gz::member_size_from_header((const unsigned char*)out.data(), out.size()) == 0 && out.size() >= 20 && (unsigned char)out[3] == 4)
        grm_text_gz::patch_member_size(out, 0);
}

void gcta_host::output_grm(const float* A, const int32_t* N, bool output_bin)
{
    const int64_t n = (int64_t)_keep.size();
    string grm_file;
    if (output_bin) {
        // full n x n float64, row-major, symmetric fill, no header, no N (grm.cpp:309-317)
        grm_file = _out + ".grm.bin";
        if (!write_grm_bin_square(grm_file, A, n)) throw("Error: can not open the file [" + grm_file + "] to write.");
        cout << "GRM of " << n << " individuals has been saved in the file [" + grm_file + "] (in binary format)." << endl;
    } else {
        // gzip text "i \t j \t N \t A" (scientific, 6 digits), 1-based, j <= i (grm.cpp:321-331).  Written as
        // concatenated gzip members compressed in parallel; gzread - hence the reference's own reader
        // (read_grm_gz, grm.cpp:381) - decodes a multi-member file transparently.
        grm_file = _out + ".grm.gz";
        {   // the reference opens the file before it prints the line below (grm.cpp:322-325)
            FILE* f = fopen(grm_file.c_str(), "wb");
            if (!f) throw("Error: can not open the file [" + grm_file + "] to write.");
            fclose(f);
        }
        cout << "Saving the genetic relationship matrix to the file [" + grm_file + "] (in compressed text format)." << endl;
        const int64_t lines_per_block = 1 << 20;
        vector<std::pair<int64_t, int64_t>> blocks;   // row ranges of ~1M lines
        for (int64_t i0 = 0; i0 < n;) {
            int64_t i1 = i0, lines = 0;
            while (i1 < n && (lines == 0 || lines + i1 + 1 <= lines_per_block)) { lines += i1 + 1; i1++; }
            blocks.emplace_back(i0, i1);
            i0 = i1;
        }
        // Blocks are formatted, compressed and written by a pool of threads (gz_member_writer.h).
        // Default codec: the structure-aware encoder of grm_text_gz.cpp; GCTA_B200_GZ_LEVEL=<0..9> formats the text and
        // runs zlib's deflate at that level instead (6 = what the reference's gzopen(name, "w") uses).
        const bool use_zlib = getenv("GCTA_B200_GZ_LEVEL") != nullptr;
        const string err = write_gz_members(
            grm_file, blocks.size(),
            [&](size_t b, string& member) {
                if (use_zlib) {
                    string text;
                    grm_text_gz::format_rows(A, N, blocks[b].first, blocks[b].second, text);
                    gz_member(text, member);
                } else
                    grm_text_gz::encode_rows(A, N, blocks[b].first, blocks[b].second, member);
            },
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

// `gcta64_b200 --selftest-writer n --out prefix [--make-grm-bin]`: time output_grm on a fabricated GRM of n individuals
// (values shaped like a real one: diagonal near 1, off-diagonal ~ N(0, 1e-3), N a little below a constant).  Needs no GPU.
int gcta_host::selftest_writer(int64_t n, bool output_bin)
{
    _keep.resize((size_t)n);
    _fid.resize((size_t)n);
    _pid.resize((size_t)n);
    for (int64_t i = 0; i < n; i++) { _keep[(size_t)i] = (int)i; _fid[(size_t)i] = "F" + std::to_string(i); _pid[(size_t)i] = "I" + std::to_string(i); }
    const size_t tri = (size_t)(n * (n + 1) / 2);
    vector<float> A(tri);
    vector<int32_t> N(tri);
    uint64_t s = 0x9E3779B97F4A7C15ull;
    auto next = [&]() { s ^= s << 13; s ^= s >> 7; s ^= s << 17; return s; };
    const bool const_n = getenv("GCTA_B200_SELFTEST_CONST_N") != nullptr;   // no missing genotypes: N = M everywhere
    size_t o = 0;
    for (int64_t i = 0; i < n; i++)
        for (int64_t j = 0; j <= i; j++, o++) {
            const uint64_t r = next();
            // sum of four uniforms, centred: close enough to a normal for the digit statistics
            const double u = ((double)(r & 0xFFFF) + (double)((r >> 16) & 0xFFFF) + (double)((r >> 32) & 0xFFFF) + (double)(r >> 48)) / 65536.0 - 2.0;
            A[o] = (float)(i == j ? 1.0 + 0.02 * u : 1.7e-3 * u);
            N[o] = const_n ? 1000000 : 500000 - (int32_t)(next() % 20000);
        }
    const auto t0 = std::chrono::steady_clock::now();
    output_grm(A.data(), N.data(), output_bin);
    const double sec = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    printf("writer selftest: n=%lld, %zu pairs, %.3f s, %.1f M pairs/s\n", (long long)n, tri, sec, tri / sec / 1e6);
    return 0;
}
