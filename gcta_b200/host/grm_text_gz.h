// grm_text_gz.h - the ".grm.gz" text of output_grm_vec (grm.cpp:321-331) written as gzip members by an encoder that knows
// the line structure, instead of formatting into a buffer and handing that to zlib's deflate.
//
// A line is "i \t j \t N \t A \n".  Against the line before it, every field either repeats, shares a prefix (i, the
// leading digits of j and N) or a suffix (the exponent of A), and the rest is close to random digits.  A general LZ77
// match search spends nearly all of its time discovering exactly that, line after line.  The encoder below knows it:
// the formatter reports, per line, how many trailing digits of j changed and whether N did, so in the usual case
// (same row, j as wide as before) the line becomes, without a single byte compare,
//     <run: newline [+ exponent] of the line before, "i \t", unchanged digits of j>  <literals: changed digits of j>
//     <run: "\t N \t">  <literals: sign, seven digits [, exponent] of A>
// with the distances given by the two lines' lengths and the literal codes of the value taken from tables indexed by
// its digit pairs (no data-dependent branch: the digits are random, a branch on them mispredicts).  Other lines (first
// of a row, j gaining a digit, nan/inf, the %.15g values of the re-save path) are compared field by field with the
// line before.  One dynamic Huffman code per member, built from a sample of the member's own lines; no hash chains.
// Measured (one core of the build container, lines of config 3's shape): 138 -> 68 ns per line against the round-1
// encoder, 7.9 -> 8.2 bytes per line (zlib level 6: 7.6 bytes, 1,700 ns; level 1: 9.3 bytes, 520 ns).
// The output is plain RFC 1951/1952: gzread, zcat and the reference's reader (read_grm_gz, grm.cpp:381-448, through
// zfstream.cpp's gzifstream) decode it like any other gzip file.
#ifndef GCTA_B200_GRM_TEXT_GZ_H
#define GCTA_B200_GRM_TEXT_GZ_H

#include <cstdint>
#include <string>
#include <vector>

namespace grm_text_gz {

// Formats "%.6e" exactly like printf (falls back to snprintf near rounding ties); returns the length.
int put_e6(char* q, double v);

// One gzip member holding the lines of rows [row_begin, row_end) of the packed lower triangle (1-based indices in the
// text, j <= i; A and N indexed by i(i+1)/2 + j).  Appends to `out`.
void encode_rows(const float* A, const int32_t* N, int64_t row_begin, int64_t row_end, std::string& out);

// The same text, uncompressed (tests; the zlib route of the writer).
void format_rows(const float* A, const int32_t* N, int64_t row_begin, int64_t row_end, std::string& text);

// Every member starts with this 20-byte gzip header: FEXTRA with one subfield 'G','M' holding the member's byte length
// (BGZF's idea with a 32-bit size).  patch_member_size fills it in once the member is complete;
// member_size_from_header returns it, or 0 when the bytes are not such a header (then: decode sequentially).
constexpr int MEMBER_HEADER_BYTES = 20;
extern const char MEMBER_HEADER[MEMBER_HEADER_BYTES];
void patch_member_size(std::string& out, size_t member_start);
uint32_t member_size_from_header(const unsigned char* h, size_t avail);

// The GRM re-save path (output_grm_MatrixXf, grm.cpp:343-380): matrices held in double, line "i \t j \t (int)N \t A"
// with A printed like `setprecision(15) << double`, i.e. printf("%.15g").
int put_g15(char* q, double v);
void encode_rows_g15(const double* A, const double* N, int64_t row_begin, int64_t row_end, std::string& out);
void format_rows_g15(const double* A, const double* N, int64_t row_begin, int64_t row_end, std::string& text);

}  // namespace grm_text_gz

#endif
