// gcta_host.cpp - readers, filters and allele-frequency bookkeeping of the --make-grm path.
// See gcta_host.h.  Log lines and error strings follow the reference (data.cpp) because
// pipelines grep them; the code itself is new.
#include "gcta_host.h"

#include <fcntl.h>
#include <unistd.h>

#include <algorithm>
#include <atomic>
#include <cmath>
#include <thread>
#include <fstream>
#include <iomanip>
#include <iostream>
#include <sstream>

#include "../../include/gcta_b200.h"

using std::cout;
using std::endl;
using std::string;
using std::vector;

gcta_host::gcta_host(int autosome_num, double rm_ld_cutoff, const string& out)
    : _autosome_num(autosome_num), _rm_ld_cutoff(rm_ld_cutoff), _out(out) {}

// ---------------------------------------------------------------------------------------------
// readers
// ---------------------------------------------------------------------------------------------
void gcta_host::read_famfile(const string& famfile)
{
    std::ifstream fam(famfile.c_str());
    if (!fam) throw("Error: can not open the file [" + famfile + "] to read.");
    cout << "Reading PLINK FAM file from [" + famfile + "]." << endl;
    _fid.clear(); _pid.clear(); _fa_id.clear(); _mo_id.clear(); _sex.clear(); _pheno.clear();
    // six whitespace-separated tokens per individual; SEX and PHENO go through atoi (data.cpp:53-67).  Same loop
    // exit as the reference: end-of-file seen while reading a first field ends the loop, and a record cut short by the
    // end of the file is completed with the last token read
    string tok;
    while (fam) {
        fam >> tok;
        if (fam.eof()) break;
        _fid.push_back(tok);
        fam >> tok;
        _pid.push_back(tok);
        fam >> tok;
        _fa_id.push_back(tok);
        fam >> tok;
        _mo_id.push_back(tok);
        fam >> tok;
        _sex.push_back(atoi(tok.c_str()));
        fam >> tok;
        _pheno.push_back(atoi(tok.c_str()));
    }
    _indi_num = (int)_fid.size();
    cout << _indi_num << " individuals to be included from [" + famfile + "]." << endl;
    init_keep();
}

void gcta_host::init_keep()
{
    _keep.resize(_indi_num);
    _id_map.clear();
    for (int i = 0; i < _indi_num; i++) {
        _keep[i] = i;
        _id_map.insert(std::make_pair(_fid[i] + ":" + _pid[i], i));
    }
}

void gcta_host::read_bimfile(const string& bimfile)
{
    std::ifstream bim(bimfile.c_str());
    if (!bim) throw("Error: can not open the file [" + bimfile + "] to read.");
    cout << "Reading PLINK BIM file from [" + bimfile + "]." << endl;
    _chr.clear(); _snp_name.clear(); _genet_dst.clear(); _bp.clear(); _allele1.clear(); _allele2.clear();
    // int chr, name, double cM, int bp, then ONE character per allele (data.cpp:104-118).  The loop ends like the
    // reference's: only end-of-file on the first field stops it cleanly; a field that does not parse (a chromosome "X",
    // a two-letter allele followed by text, a fractional position) still appends one last record - built from whatever
    // the failed extractions leave in the variables - before the failed stream ends the loop.  Pipelines that feed
    // such files get the same SNP count and the same downstream messages.
    int ival = 0;
    double dval = 0.0;
    char cval = '0';
    string sval;
    while (bim) {
        bim >> ival;
        if (bim.eof()) break;
        _chr.push_back(ival);
        bim >> sval;
        _snp_name.push_back(sval);
        bim >> dval;
        _genet_dst.push_back(dval);
        bim >> ival;
        _bp.push_back(ival);
        bim >> cval;
        _allele1.push_back(cval);
        bim >> cval;
        _allele2.push_back(cval);
    }
    _snp_num = (int)_chr.size();
    _ref_A = _allele1;   // data.cpp:121
    cout << _snp_num << " SNPs to be included from [" + bimfile + "]." << endl;
    init_include();
}

void gcta_host::init_include()
{
    _include.resize(_snp_num);
    _snp_name_map.clear();
    for (int i = 0; i < _snp_num; i++) {
        _include[i] = i;
        _snp_name_map.insert(std::make_pair(_snp_name[i], i));
    }
}

void gcta_host::read_bedfile(const string& bedfile)
{
    if (_include.empty()) throw("Error: No SNP is retained for analysis.");          // data.cpp:149
    if (_keep.empty()) throw("Error: No individual is retained for analysis.");      // data.cpp:150
    // a SNP / individual is read iff it is still in the name maps (get_rsnp / get_rindi, data.cpp:196-214)
    // (a map as large as the file means no filter has removed anything and no name repeats: every look-up would succeed -
    // a million string searches saved on the headline .bim)
    vector<char> rsnp(_snp_num, 1);
    if ((int64_t)_snp_name_map.size() != (int64_t)_snp_num)
        for (int j = 0; j < _snp_num; j++) rsnp[j] = _snp_name_map.find(_snp_name[j]) != _snp_name_map.end();
    vector<int> keep;
    const bool all_indi = (int64_t)_id_map.size() == (int64_t)_indi_num;
    for (int i = 0; i < _indi_num; i++)
        if (all_indi || _id_map.find(_fid[i] + ":" + _pid[i]) != _id_map.end()) keep.push_back(i);
    const int fd = open(bedfile.c_str(), O_RDONLY);
    if (fd < 0) throw("Error: can not open the file [" + bedfile + "] to read.");
    cout << "Reading PLINK BED file from [" + bedfile + "] in SNP-major format ..." << endl;
    // the 3 magic bytes are skipped unchecked (data.cpp:164); the rows wanted are read straight into the packed buffer,
    // runs of consecutive rows with one pread each, by a few threads (the file is 12.5 GB at the headline size)
    _bed_stride = ((int64_t)_indi_num + 3) / 4;
    vector<std::pair<int64_t, int64_t>> runs;   // (first file row, rows) of consecutive wanted rows
    int64_t m_read = 0;
    for (int j = 0; j < _snp_num; j++) {
        if (!rsnp[j]) continue;
        if (!runs.empty() && runs.back().first + runs.back().second == j) runs.back().second++;
        else runs.emplace_back(j, 1);
        m_read++;
    }
    _bed.resize((size_t)(m_read * _bed_stride));   // not zero-filled: every byte comes from the file
    // split long runs so that every thread gets work
    const int64_t piece = std::max<int64_t>(1, (64ll << 20) / std::max<int64_t>(1, _bed_stride));
    struct Piece { int64_t file_row, dst_row, rows; };
    vector<Piece> pieces;
    int64_t dst = 0;
    for (const auto& r : runs)
        for (int64_t o = 0; o < r.second; o += piece) {
            const int64_t k = std::min(piece, r.second - o);
            pieces.push_back({r.first + o, dst, k});
            dst += k;
        }
    std::atomic<size_t> next(0);
    std::atomic<bool> bad(false);
    auto work = [&]() {
        for (size_t p = next++; p < pieces.size() && !bad; p = next++) {
            char* out = reinterpret_cast<char*>(&_bed[(size_t)(pieces[p].dst_row * _bed_stride)]);
            size_t left = (size_t)(pieces[p].rows * _bed_stride);
            off_t off = (off_t)3 + (off_t)pieces[p].file_row * (off_t)_bed_stride;
            while (left > 0) {
                const ssize_t g = pread(fd, out, left, off);
                if (g <= 0) { bad = true; break; }   // short file
                out += g; left -= (size_t)g; off += g;
            }
        }
    };
    const int T = (int)std::max<size_t>(1, std::min<size_t>({(size_t)8, (size_t)std::thread::hardware_concurrency(), pieces.size()}));
    vector<std::thread> th;
    for (int t = 1; t < T; t++) th.emplace_back(work);
    work();
    for (auto& x : th) x.join();
    close(fd);
    if (bad) throw("Error: problem with the BED file...has the FAM/BIM file been changed?");   // data.cpp:173
    _keep = keep;
    cout << "Genotype data for " << _keep.size() << " individuals and " << m_read << " SNPs to be included from [" + bedfile + "]." << endl;

    // update_bim (data.cpp:216-245): SNP tables shrink to the rows that were read and are renumbered.  Every row read
    // and the name map complete (see above): the tables, _include and the map already are what update_bim + init_include
    // would rebuild - a million string copies and map insertions saved.
    if (m_read == _snp_num && (int64_t)_snp_name_map.size() == (int64_t)_snp_num && (int64_t)_include.size() == (int64_t)_snp_num) return;
    vector<int> chr, bp;
    vector<char> a1, a2, ra;
    vector<string> nm;
    vector<double> gd;
    for (int j = 0; j < _snp_num; j++) {
        if (!rsnp[j]) continue;
        chr.push_back(_chr[j]); nm.push_back(_snp_name[j]); gd.push_back(_genet_dst[j]); bp.push_back(_bp[j]);
        a1.push_back(_allele1[j]); a2.push_back(_allele2[j]); ra.push_back(_ref_A[j]);
    }
    _chr.swap(chr); _snp_name.swap(nm); _genet_dst.swap(gd); _bp.swap(bp); _allele1.swap(a1); _allele2.swap(a2); _ref_A.swap(ra);
    _snp_num = (int)_chr.size();
    init_include();
}

// ---------------------------------------------------------------------------------------------
// selection lists
// ---------------------------------------------------------------------------------------------
void gcta_host::update_id_map_kp(const vector<string>& ids, std::map<string, int>& id_map, vector<int>& keep)
{
    std::map<string, int> kept;
    for (const string& id : ids) {
        auto it = id_map.find(id);
        if (it != id_map.end()) kept.insert(*it);
    }
    id_map.swap(kept);
    keep.clear();
    for (const auto& kv : id_map) keep.push_back(kv.second);
    std::stable_sort(keep.begin(), keep.end());
}

void gcta_host::update_id_map_rm(const vector<string>& ids, std::map<string, int>& id_map, vector<int>& keep)
{
    for (const string& id : ids) id_map.erase(id);
    keep.clear();
    for (const auto& kv : id_map) keep.push_back(kv.second);
    std::stable_sort(keep.begin(), keep.end());
}

void gcta_host::read_snplist(const string& file, vector<string>& list, const string& msg)
{
    list.clear();
    std::ifstream in(file.c_str());
    if (!in) throw("Error: can not open the file [" + file + "] to read.");
    cout << "Reading a list of " << msg << " from [" + file + "]." << endl;
    string tok, rest;
    while (in >> tok) {   // first token of each line (data.cpp:520-533)
        list.push_back(tok);
        std::getline(in, rest);
    }
}

void gcta_host::read_indi_list(const string& file, vector<string>& list)
{
    std::ifstream in(file.c_str());
    if (!in) throw("Error: can not open the file [" + file + "] to read.");
    list.clear();
    string fid, pid, rest;
    while (in >> fid) {   // FID IID, rest of the line ignored (data.cpp:657-673)
        in >> pid;
        list.push_back(fid + ":" + pid);
        std::getline(in, rest);
    }
}

void gcta_host::keep_indi(const string& f)
{
    vector<string> list;
    read_indi_list(f, list);
    update_id_map_kp(list, _id_map, _keep);
    cout << _keep.size() << " individuals are kept from [" + f + "]." << endl;
}

void gcta_host::remove_indi(const string& f)
{
    vector<string> list;
    read_indi_list(f, list);
    const size_t prev = _keep.size();
    update_id_map_rm(list, _id_map, _keep);
    cout << prev - _keep.size() << " individuals are removed from [" + f + "] and there are " << _keep.size() << " individuals remaining." << endl;
}

void gcta_host::extract_snp(const string& f)
{
    vector<string> list;
    read_snplist(f, list);
    update_id_map_kp(list, _snp_name_map, _include);
    cout << _include.size() << " SNPs are extracted from [" + f + "]." << endl;
}

void gcta_host::exclude_snp(const string& f)
{
    vector<string> list;
    read_snplist(f, list);
    const size_t prev = _include.size();
    update_id_map_rm(list, _snp_name_map, _include);
    cout << prev - _include.size() << " SNPs are excluded from [" + f + "] and there are " << _include.size() << " SNPs remaining." << endl;
}

void gcta_host::extract_single_snp(const string& snp)
{
    update_id_map_kp(vector<string>(1, snp), _snp_name_map, _include);
    if (_include.empty()) throw("Error: can not find the SNP [" + snp + "] in the data.");
    cout << "Only the SNP[" + snp + "] is included in the analysis." << endl;
}

void gcta_host::exclude_single_snp(const string& snp)
{
    const size_t prev = _include.size();
    update_id_map_rm(vector<string>(1, snp), _snp_name_map, _include);
    if (_include.size() == prev) throw("Error: can not find the SNP [" + snp + "] in the data.");
    cout << "The SNP[" + snp + "] has been excluded from the analysis." << endl;
}

void gcta_host::extract_chr(int chr_start, int chr_end)
{
    std::map<string, int> kept;
    _include.clear();
    for (const auto& kv : _snp_name_map)
        if (_chr[kv.second] >= chr_start && _chr[kv.second] <= chr_end) { kept.insert(kv); _include.push_back(kv.second); }
    _snp_name_map.swap(kept);
    std::stable_sort(_include.begin(), _include.end());
    if (chr_start != chr_end) cout << _include.size() << " SNPs from chromosome " << chr_start << " to chromosome " << chr_end << " are included in the analysis." << endl;
    else cout << _include.size() << " SNPs on chromosome " << chr_start << " are included in the analysis." << endl;
}

void gcta_host::update_sex(const string& sex_file)
{
    std::ifstream in(sex_file.c_str());
    if (!in) throw("Error: can not open the file [" + sex_file + "] to read.");
    cout << "Reading sex information from [" + sex_file + "]." << endl;
    _sex.assign(_indi_num, 0);
    vector<char> confirm(_indi_num, 0);
    int icount = 0;
    // `code` also receives the rest of each line (as the reference's one buffer does, data.cpp:704-716): a last record
    // without a sex code is then reported with an empty code instead of silently taking the previous one
    string fid, pid, code;
    while (in) {
        in >> fid;
        if (in.eof()) break;
        in >> pid;
        in >> code;
        if (code != "1" && code != "2" && code != "M" && code != "F")
            throw("Error: unrecognized sex code: \"" + fid + " " + pid + " " + code + "\" in [" + sex_file + "].");
        auto it = _id_map.find(fid + ":" + pid);
        if (it != _id_map.end()) {
            _sex[it->second] = (code == "M" || code == "1") ? 1 : 2;
            confirm[it->second] = 1;
            icount++;
        }
        std::getline(in, code);
    }
    for (int i : _keep)
        if (!confirm[i]) throw("Error: sex information for all of the included individuals should be updated.");
    cout << "Sex information for " << icount << " individuals are update from [" + sex_file + "]." << endl;
}

void gcta_host::update_ref_A(const string& file)
{
    std::ifstream in(file.c_str());
    if (!in) throw("Error: can not open the file [" + file + "] to read.");
    cout << "Reading reference alleles of SNPs from [" + file + "]." << endl;
    _ref_A.assign(_snp_num, 0);
    vector<char> confirm(_snp_num, 0);
    int icount = 0;
    string snp, rest;
    char a = 0;
    while (in >> snp) {
        in >> a;
        auto it = _snp_name_map.find(snp);
        if (it != _snp_name_map.end()) {
            if (a != _allele1[it->second] && a != _allele2[it->second])
                throw("Error: invalid reference allele for SNP \"" + _snp_name[it->second] + "\".");
            _ref_A[it->second] = a;
            confirm[it->second] = 1;
            icount++;
        }
        std::getline(in, rest);
    }
    for (int j : _include)
        if (!confirm[j]) throw("Error: The reference alleles of all of the included SNPs should be updated.");
    cout << "Reference alleles for " << icount << " SNPs are update from [" + file + "]." << endl;
}
