// gcta_host.h - C++ host side of the --bfile ... --make-grm path, above the C ABI.
//
// Mirrors the part of the reference's `class gcta` (gcta.h:49-279) that the path touches: same
// method names, argument meaning, log lines and error strings, so that option handling reads
// like option.cpp:575-628.  State differs on purpose: genotypes stay as the packed .bed payload
// (the reference expands them into two vector<vector<bool>> planes, gcta.h:242-243) because the
// device decodes the 2-bit codes itself.  All O(nM) and O(n^2 M) work happens in
// libgcta_b200.so; this file only parses, filters index lists and writes results.
#pragma once
#include "big_alloc.h"
#include <cstdint>
#include <map>
#include <memory>
#include <string>
#include <utility>
#include <vector>

class gcta_host {
public:
    gcta_host(int autosome_num, double rm_ld_cutoff, const std::string& out);   // data.cpp:15-24

    // readers (data.cpp:39-194)
    void read_famfile(const std::string& famfile);
    void read_bimfile(const std::string& bimfile);
    void read_bedfile(const std::string& bedfile);

    // selection (data.cpp:535-586, 675-690), run BEFORE read_bedfile like option.cpp:588-599
    void keep_indi(const std::string& indi_list_file);
    void remove_indi(const std::string& indi_list_file);
    void extract_snp(const std::string& snplistfile);
    void exclude_snp(const std::string& snplistfile);
    void extract_single_snp(const std::string& snpname);
    void exclude_single_snp(const std::string& snpname);
    void extract_chr(int chr_start, int chr_end);
    void update_sex(const std::string& sex_file);         // data.cpp:692-724
    void update_ref_A(const std::string& ref_A_file);     // data.cpp:726-757

    // after read_bedfile (option.cpp:601-605)
    void update_freq(const std::string& freq);            // data.cpp:889-928
    void filter_snp_maf(double maf);                      // data.cpp:588-610
    void filter_snp_max_maf(double max_maf);              // data.cpp:612-634

    // analyses
    void save_freq();                                     // data.cpp:930-945
    void make_grm(bool grm_xchr_flag, bool output_bin, int grm_mtd);   // grm.cpp:219-301 -> gcta_b200_make_grm

    int selftest_writer(int64_t n, bool output_bin);      // times output_grm on a fabricated GRM (no GPU)

    // consumer side of the .grm.gz: --grm / --mgrm with --make-grm[-bin], grm.cpp:343-654.  Host only,
    // like the reference (O(n^2) text and arithmetic); gcta_host_grmio.cpp
    void save_grm(const std::string& grm_file, const std::string& keep_indi_file, const std::string& remove_indi_file,
                  const std::string& sex_file, double grm_cutoff, double adj_grm_fac, int dosage_compen, bool merge_grm_flag,
                  bool output_grm_bin);                    // grm.cpp:572-577

    // device options (not in the reference)
    int operand = 0;      // GCTA_B200_FP16
    int n_gpus = 1;       // row bands computed one after another on devices 0..n_gpus-1 of this process
    int64_t chunk_snps = 0;

private:
    void init_keep();
    void init_include();
    void calcu_mu();                                      // data.cpp:759-779 via gcta_b200_allele_counts
    void check_autosome() const;                          // grm.cpp:15-20
    void check_chrX() const;                              // grm.cpp:22-27
    void check_sex() const;                               // grm.cpp:29-34
    void rm_high_ld();                                    // data.cpp:1194-1196 (default: no-op)
    void output_grm(const float* A, const int32_t* N, bool output_bin);   // N == NULL with output_bin   // grm.cpp:303-341
    void read_grm_gz(const std::string& grm_file, std::vector<std::string>& grm_id, bool out_id_log = true, bool read_id_only = false);   // grm.cpp:381-448
    void merge_grm(const std::string& merge_grm_file);    // grm.cpp:608-654
    void rm_cor_indi(double grm_cutoff);                  // grm.cpp:450-493
    void adj_grm(double adj_grm_fac);                     // grm.cpp:495-522
    void dc(int dosage_compen);                           // grm.cpp:524-541
    void manipulate_grm(const std::string& grm_file, const std::string& keep_indi_file, const std::string& remove_indi_file,
                        const std::string& sex_file, double grm_cutoff, double adj_grm_fac, int dosage_compen, bool merge_grm_flag);   // grm.cpp:544-570
    void output_grm_g15(bool output_grm_bin);             // output_grm_MatrixXf, grm.cpp:343-380
    // the GRM as read back: packed lower triangle of dimension _grm_dim, doubles like the reference's default build
    // (gcta.h:43 eigenMatrix = MatrixXd); element (i, j), i >= j, at i(i+1)/2 + j
    std::vector<double> _grm, _grm_N;
    int64_t _grm_dim = 0;
    double& grm_at(std::vector<double>& M, int64_t i, int64_t j);
    static void read_snplist(const std::string& file, std::vector<std::string>& list, const std::string& msg = "SNPs");
    static void read_indi_list(const std::string& file, std::vector<std::string>& list);
    static void update_id_map_kp(const std::vector<std::string>& ids, std::map<std::string, int>& id_map, std::vector<int>& keep);
    static void update_id_map_rm(const std::vector<std::string>& ids, std::map<std::string, int>& id_map, std::vector<int>& keep);

    // .fam (gcta.h:230-238)
    int _indi_num = 0;
    std::vector<std::string> _fid, _pid, _fa_id, _mo_id;
    std::vector<int> _sex, _pheno, _keep;
    std::map<std::string, int> _id_map;
    // .bim (gcta.h:216-227); after read_bedfile these are compacted to the SNPs that were read
    int _snp_num = 0;
    std::vector<int> _chr, _bp, _include;
    std::vector<std::string> _snp_name;
    std::vector<double> _genet_dst;
    std::vector<char> _allele1, _allele2, _ref_A;
    std::map<std::string, int> _snp_name_map;
    // genotypes: payload rows of the SNPs read (compact SNP index), all file individuals
    // (12.5 GB at the headline size: huge-page backed and never zero-filled - every byte comes from the file)
    std::vector<uint8_t, big_alloc::allocator<uint8_t>> _bed;
    int64_t _bed_stride = 0;
    std::vector<double> _mu;   // 2p in the _ref_A orientation, indexed by compact SNP id (gcta.h:273)
    bool _mu_from_file = false; // set by update_freq: only then is 2p passed down instead of recomputed
    int _autosome_num;
    double _rm_ld_cutoff;
    std::string _out;
};
