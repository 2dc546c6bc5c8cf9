// main.cpp - gcta64_b200: command-line front end for the accelerated `--bfile ... --make-grm` path.
//
// Drop-in for the flag subset of SURVEY.md 8a-1 (option.cpp:114-543): same flag names, value
// parsing (atoi / atof), range errors, option echo, banner, ordering of the calls
// (option.cpp:580-621) and error convention (message on stderr, "Analysis finished", exit code 0;
// option.cpp:42-50).  Flags of GCTA analyses outside the path are rejected like unknown flags.
// Extra flags (not in the reference): --gpus N, --operand fp16|bf16, --chunk-snps K.
#include <cstdlib>
#include <chrono>
#include <cstring>
#include <ctime>
#include <iostream>
#include <sstream>
#include <string>

#include "../../include/gcta_b200.h"
#include "gcta_host.h"

using std::cerr;
using std::cout;
using std::endl;
using std::string;

static void option(int argc, char* argv_in[])
{
    string bfile, update_sex_file, update_freq_file, update_refA_file, kp_indi_file, rm_indi_file, extract_snp_file,
        exclude_snp_file, extract_snp_name, exclude_snp_name, out = "gcta";
    bool bfile_flag = false, out_freq_flag = false, autosome_flag = false, make_grm_flag = false, make_grm_xchr_flag = false,
         grm_out_bin_flag = false, grm_flag = false, m_grm_flag = false, pca_flag = false;
    string grm_file;
    double grm_adj_fac = -2.0, grm_cutoff = -2.0;
    int dosage_compen = -2;
    int autosome_num = 22, extract_chr_start = 0, extract_chr_end = 0, grm_mtd = 0, n_gpus = 1, operand = GCTA_B200_FP16;
    long long chunk_snps = 0;
    double maf = 0.0, max_maf = 0.0, rm_high_ld_cutoff = -1.0;

    // like the reference, a sentinel after the last argument keeps `argv[++i]` in range (option.cpp:111-113)
    std::vector<const char*> argv(argv_in, argv_in + argc);
    argv.push_back("gcta");
    argv.push_back("gcta");
    for (int i = 1; i < argc; i++) {
        const char* o = argv[i];
        auto is = [&](const char* name) { return strcmp(o, name) == 0; };
        if (is("--bfile")) { bfile_flag = true; bfile = argv[++i]; cout << "--bfile " << argv[i] << endl; }
        else if (is("--update-freq")) { update_freq_file = argv[++i]; cout << "--update-freq " << update_freq_file << endl; }
        else if (is("--update-ref-allele")) { update_refA_file = argv[++i]; cout << "--update-ref-allele " << update_refA_file << endl; }
        else if (is("--keep")) { kp_indi_file = argv[++i]; cout << "--keep " << kp_indi_file << endl; }
        else if (is("--remove")) { rm_indi_file = argv[++i]; cout << "--remove " << rm_indi_file << endl; }
        else if (is("--update-sex")) { update_sex_file = argv[++i]; cout << "--update-sex " << update_sex_file << endl; }
        else if (is("--chr")) {
            extract_chr_start = extract_chr_end = atoi(argv[++i]);
            cout << "--chr " << extract_chr_start << endl;
            if (extract_chr_start < 1 || extract_chr_start > 100) throw("\nError: --chr should be within the range from 1 to 100.\n");
        }
        else if (is("--autosome-num")) {
            autosome_num = atoi(argv[++i]);
            cout << "--autosome-num " << autosome_num << endl;
            if (autosome_num < 1 || autosome_num > 100) throw("\nError: invalid number specified after the option --autosome-num.\n");
        }
        else if (is("--autosome")) { autosome_flag = true; cout << "--autosome" << endl; }
        else if (is("--extract")) { extract_snp_file = argv[++i]; cout << "--extract " << extract_snp_file << endl; }
        else if (is("--exclude")) { exclude_snp_file = argv[++i]; cout << "--exclude " << exclude_snp_file << endl; }
        else if (is("--extract-snp")) { extract_snp_name = argv[++i]; cout << "--extract-snp " << extract_snp_name << endl; }
        else if (is("--exclude-snp")) { exclude_snp_name = argv[++i]; cout << "--exclude-snp " << exclude_snp_name << endl; }
        else if (is("--maf")) {
            maf = atof(argv[++i]);
            cout << "--maf " << maf << endl;
            if (maf < 0 || maf > 0.5) throw("\nError: --maf should be within the range from 0 to 0.5.\n");
        }
        else if (is("--max-maf")) {
            max_maf = atof(argv[++i]);
            cout << "--max-maf " << max_maf << endl;
            if (max_maf < 0 || max_maf > 0.5) throw("\nError: --max-maf should be within the range from 0 to 0.5.\n");
        }
        else if (is("--out")) { out = argv[++i]; cout << "--out " << out << endl; }
        else if (is("--freq")) { out_freq_flag = true; cout << "--freq" << endl; }
        else if (is("--save-ram")) { cout << "--save-ram" << endl; }   // only read by the LD analysis in the reference
        else if (is("--grm-alg")) {
            grm_mtd = atoi(argv[++i]);
            cout << "--grm-alg " << grm_mtd << endl;
            if (grm_mtd < 0 || grm_mtd > 1) throw("\nError: --grm-alg should be 0 or 1.\n");
        }
        else if (is("--rm-high-ld")) {
            rm_high_ld_cutoff = atof(argv[++i]);
            cout << "--rm-high-ld " << rm_high_ld_cutoff << endl;
            if (rm_high_ld_cutoff <= 0 || rm_high_ld_cutoff >= 1) throw("\nError: the value to be specified after --rm-high-ld should be within the range from 0 to 1.\n");
        }
        // ---- consumer side of the GRM files (option.cpp:267-276, 305-327) ----
        else if (is("--mgrm")) { m_grm_flag = true; grm_file = argv[++i]; cout << "--mgrm " << grm_file << endl; }
        else if (is("--grm")) { grm_flag = true; grm_file = argv[++i]; cout << "--grm " << grm_file << endl; }
        else if (is("--grm-adj")) {
            grm_adj_fac = atof(argv[++i]);
            cout << "--grm-adj " << grm_adj_fac << endl;
            if (grm_adj_fac < 0 || grm_adj_fac > 1) throw("\nError: the value to be specified after --grm-adj should be within the range from 0 to 1.\n");
        }
        else if (is("--dc")) {
            dosage_compen = atoi(argv[++i]);
            cout << "--dc " << dosage_compen << endl;
            if (dosage_compen != 0 && dosage_compen != 1) throw("\nError: the value to be specified after --dc should be 0 or 1.\n");
        }
        else if (is("--grm-cutoff")) {
            grm_cutoff = atof(argv[++i]);
            cout << "--grm-cutoff " << grm_cutoff << endl;
            if (grm_cutoff < 0 || grm_cutoff > 1) throw("\nError: the value to be specified after --grm-cutoff should be within the range from 0 to 1.\n");
        }
        else if (is("--pca")) {   // optional count (option.cpp:320-327); the analysis itself is not part of this build
            pca_flag = true;
            int out_pc_num = 20;
            i++;
            if (strcmp(argv[i], "gcta") == 0 || strncmp(argv[i], "--", 2) == 0) i--;
            else out_pc_num = atoi(argv[i]);
            cout << "--pca " << out_pc_num << endl;
            if (out_pc_num < 1) throw("\nError: the value to be specified after --pca should be positive.\n");
        }
        else if (is("--make-grm")) { make_grm_flag = true; cout << "--make-grm" << endl; }
        else if (is("--make-grm-xchr")) { make_grm_xchr_flag = true; cout << "--make-grm-xchr" << endl; }
        else if (is("--make-grm-bin")) { make_grm_flag = true; grm_out_bin_flag = true; cout << "--make-grm-bin" << endl; }
        else if (is("--make-grm-xchr-bin")) { make_grm_xchr_flag = true; grm_out_bin_flag = true; cout << "--make-grm-xchr-bin" << endl; }
        // ---- device options, not in the reference ----
        else if (is("--gpus")) { n_gpus = atoi(argv[++i]); cout << "--gpus " << n_gpus << endl; if (n_gpus < 1 || n_gpus > 64) throw("\nError: --gpus should be within the range from 1 to 64.\n"); }
        else if (is("--operand")) {
            const string v = argv[++i];
            cout << "--operand " << v << endl;
            if (v == "fp16") operand = GCTA_B200_FP16;
            else if (v == "bf16") operand = GCTA_B200_BF16;
            else if (v == "fp16x2") operand = GCTA_B200_FP16X2;
            else throw("\nError: --operand should be fp16, bf16 or fp16x2.\n");
        }
        else if (is("--chunk-snps")) { chunk_snps = atoll(argv[++i]); cout << "--chunk-snps " << chunk_snps << endl; }
        else if (is("gcta")) break;
        else { std::stringstream e; e << "\nError: invalid option \"" << o << "\".\n"; throw(e.str()); }
    }
    cout << endl;
    // conflicted options (option.cpp:547-563)
    if (m_grm_flag) {
        if (grm_flag) { grm_flag = false; cout << "Warning: --grm option suppressed by the --mgrm option." << endl; }
        if (grm_cutoff > -1.0) { grm_cutoff = -2.0; cout << "Warning: --grm-cutoff option suppressed by the --mgrm option." << endl; }
    }
    if (pca_flag) throw("Error: --pca is not available in gcta64_b200 (outside the accelerated --make-grm path).");
    if (dosage_compen > -1 && update_sex_file.empty())
        throw("Error: you need to specify the sex information for the individuals by the option --update-sex because of the option --dc.");
    // set autosome (option.cpp:566-571)
    if (autosome_flag) { extract_chr_start = 1; extract_chr_end = autosome_num; }
    if (make_grm_xchr_flag) extract_chr_start = extract_chr_end = autosome_num + 1;

    cout << endl;
    gcta_host g(autosome_num, rm_high_ld_cutoff, out);
    g.n_gpus = n_gpus; g.operand = operand; g.chunk_snps = chunk_snps;
    if (bfile_flag) {   // option.cpp:580-621
        g.read_famfile(bfile + ".fam");
        if (!kp_indi_file.empty()) g.keep_indi(kp_indi_file);
        if (!rm_indi_file.empty()) g.remove_indi(rm_indi_file);
        if (!update_sex_file.empty()) g.update_sex(update_sex_file);
        g.read_bimfile(bfile + ".bim");
        if (!extract_snp_file.empty()) g.extract_snp(extract_snp_file);
        if (!exclude_snp_file.empty()) g.exclude_snp(exclude_snp_file);
        if (extract_chr_start > 0) g.extract_chr(extract_chr_start, extract_chr_end);
        if (!extract_snp_name.empty()) g.extract_single_snp(extract_snp_name);
        if (!exclude_snp_name.empty()) g.exclude_single_snp(exclude_snp_name);
        if (!update_refA_file.empty()) g.update_ref_A(update_refA_file);
        {
            const auto t_r0 = std::chrono::steady_clock::now();
            g.read_bedfile(bfile + ".bed");
            if (getenv("GCTA_B200_TIMING"))   // phase split for tools/cli_cfg3.sh; the .fam / .bim parsing came before
                std::cerr << "{\"read_bed_s\": " << std::chrono::duration<double>(std::chrono::steady_clock::now() - t_r0).count() << "}" << std::endl;
        }
        if (!update_freq_file.empty()) g.update_freq(update_freq_file);
        if (maf > 0) g.filter_snp_maf(maf);
        if (max_maf > 0.0) g.filter_snp_max_maf(max_maf);
        if (out_freq_flag) g.save_freq();
        else if (make_grm_flag) g.make_grm(false, grm_out_bin_flag, grm_mtd);
        else if (make_grm_xchr_flag) g.make_grm(true, grm_out_bin_flag, grm_mtd);
    }
    else if (grm_flag || m_grm_flag) {   // option.cpp:657-660
        if (make_grm_flag)
            g.save_grm(grm_file, kp_indi_file, rm_indi_file, update_sex_file, grm_cutoff, grm_adj_fac, dosage_compen, m_grm_flag, grm_out_bin_flag);
    }
    else throw("Error: no analysis has been launched by the option(s).\n");
}

int gcta_host_selftest_format();   // gcta_host_grm.cpp
int gcta_host_selftest_gz();

int main(int argc, char* argv[])
{
    if (argc == 2 && strcmp(argv[1], "--selftest-format") == 0) return gcta_host_selftest_format();
    if (argc == 2 && strcmp(argv[1], "--selftest-gz") == 0) return gcta_host_selftest_gz();
    if (argc >= 5 && strcmp(argv[1], "--selftest-writer") == 0 && strcmp(argv[3], "--out") == 0) {
        // gcta64_b200 --selftest-writer n --out prefix [--make-grm-bin]
        try {
            gcta_host g(22, -1.0, argv[4]);
            return g.selftest_writer(atoll(argv[2]), argc >= 6 && strcmp(argv[5], "--make-grm-bin") == 0);
        } catch (const string& err_msg) { cerr << err_msg << endl; return 1; }
    }
    cout << "*******************************************************************" << endl;
    cout << "* Genome-wide Complex Trait Analysis (GCTA)" << endl;
    cout << "* version 0.93.9 - B200 build of the --make-grm path (" << gcta_b200_version() << ")" << endl;
    cout << "* (C) 2010 Jian Yang, Hong Lee, Michael Goddard and Peter Visscher" << endl;
    cout << "* GNU General Public License, v2" << endl;
    cout << "* Queensland Institute of Medical Research" << endl;
    cout << "*******************************************************************" << endl;
    const time_t start = time(NULL);
    time_t curr = time(0);
    cout << "Analysis started: " << ctime(&curr) << endl;
    cout << "Options:" << endl;
    try { option(argc, argv); }
    catch (const string& err_msg) { cerr << err_msg << endl; }
    catch (const char* err_msg) { cerr << err_msg << endl; }
    curr = time(0);
    cout << "\nAnalysis finished: " << ctime(&curr);
    const long t = (long)(time(NULL) - start);
    cout << "Computational time: " << t / 3600 << ":" << (t % 3600) / 60 << ":" << t % 60 << endl;
    return 0;   // the reference exits 0 even after an error (option.cpp:50)
}
