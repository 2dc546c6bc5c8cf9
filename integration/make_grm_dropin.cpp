// make_grm_dropin.cpp - INTEGRATION.md section 1 made real: the body of gcta::make_grm (grm.cpp:219-301) replaced by one
// call of gcta_b200_make_grm, compiled INTO the reference.  integration/Makefile builds the reference's own translation
// units where they lie in /root/reference (grm.cpp with -Dmake_grm=make_grm_cpu, so its CPU loop keeps existing under
// another name), this file, and links libgcta_b200.so: the result is the reference program - its option parser, readers,
// filters, writers, log lines - with the hot path on the GPU.  tests/test_gpu_dropin.py runs it against the goldens.
//
// This file is written against the reference's class (gcta.h) and uses its private state the way make_XMat does
// (data.cpp:982-1014); it contains no reference code.
#include "gcta.h"

#include "gcta_b200.h"

void gcta::make_grm(bool grm_xchr_flag, bool output_bin, int grm_mtd)
{
    if (grm_xchr_flag) check_chrX();
    else check_autosome();
    rm_high_ld();
    if (_dosage_flag) throw("Error: the GPU build of --make-grm reads PLINK binary genotypes only (no dosage data).");

    // After read_bedfile the two bit planes hold one row per SNP that was read and one column per kept individual
    // (data.cpp:171-185): column i belongs to individual _keep[i].  Re-pack them into .bed payload rows.
    const int64_t n = (int64_t)_keep.size(), m_rows = (int64_t)_snp_num, stride = (n + 3) / 4;
    std::vector<uint8_t> bed((size_t)(m_rows * stride), 0);
    for (int64_t s = 0; s < m_rows; s++) {
        uint8_t* row = &bed[(size_t)(s * stride)];
        for (int64_t i = 0; i < n; i++) {
            const unsigned lo = !_snp_2[s][i], hi = !_snp_1[s][i];   // read_bedfile stored the negated bits
            row[i >> 2] |= (uint8_t)((lo | (hi << 1)) << ((i & 3) * 2));
        }
    }
    std::vector<int32_t> include(_include.begin(), _include.end()), sex((size_t)n, 0);
    for (int64_t i = 0; i < n; i++) sex[(size_t)i] = _sex[_keep[i]];
    // --update-freq / an earlier --maf left 2p of the REFERENCE allele in _mu, indexed by SNP (data.cpp:915-916, 779, 794);
    // the library centres the dosage of allele 1 (.bim column 5), so a SNP whose reference allele was flipped by
    // --update-ref-allele hands over 2 - 2p
    std::vector<double> mu;
    if (!_mu.empty())
        for (size_t j = 0; j < _include.size(); j++) {
            const int id = _include[j];
            mu.push_back(_allele1[id] == _ref_A[id] ? _mu[id] : 2.0 - _mu[id]);
        }

    cout << "Recoding genotypes (individual major mode) ..." << endl;               // make_XMat's and std_XMat's lines,
    if (_mu.empty()) cout << "Calculating allele frequencies ..." << endl;         // so that logs stay comparable
    if (grm_xchr_flag) cout << "\nCalculating the genetic relationship matrix for the X chromosome ... ((NOTE: default speed-optimized mode, may use huge RAM))" << endl;
    else cout << "\nCalculating the genetic relationship matrix ... (NOTE: default speed-optimized mode, may use huge RAM)" << endl;

    const size_t tri = (size_t)(n * (n + 1) / 2);
    std::vector<float> A(tri);
    std::vector<int32_t> N(tri);
    gcta_b200_args a;
    memset(&a, 0, sizeof a);
    a.bed = bed.data();
    a.n_file = n;
    a.m_file = m_rows;
    a.keep = NULL;               // the planes are already restricted to the kept individuals
    a.n = n;
    a.include = include.data();
    a.m = (int64_t)include.size();
    a.mu = mu.empty() ? NULL : mu.data();
    a.sex = sex.data();
    a.xchr = grm_xchr_flag ? 1 : 0;
    a.alg = grm_mtd;
    a.operand = GCTA_B200_FP16;
    a.part_world = 1;
    gcta_b200_out o;
    memset(&o, 0, sizeof o);
    o.A_lower = A.data();
    o.N_lower = N.data();
    char err[512];
    if (gcta_b200_make_grm(&a, &o, err, sizeof err) != 0) throw string(err);   // caught in main like any GCTA error

    // hand the result to the reference's own writer (grm.cpp:303-341): it reads A[i][j], A_N[i][j] for j <= i
    vector< vector<double> > Am((size_t)n);
    vector< vector<int> > Nm((size_t)n);
    for (int64_t i = 0; i < n; i++) {
        Am[(size_t)i].resize((size_t)i + 1);
        Nm[(size_t)i].resize((size_t)i + 1);
        for (int64_t j = 0; j <= i; j++) {
            Am[(size_t)i][(size_t)j] = A[(size_t)(i * (i + 1) / 2 + j)];
            Nm[(size_t)i][(size_t)j] = N[(size_t)(i * (i + 1) / 2 + j)];
        }
    }
    output_grm_vec(Am, Nm, output_bin);
}
