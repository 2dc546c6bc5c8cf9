// grm_text_gz.cpp - structure-aware gzip encoder for the .grm.gz text (see grm_text_gz.h).
#include "grm_text_gz.h"

#include <zlib.h>   // crc32 only

#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <functional>
#include <queue>
#if defined(__SSE2__)
#include <emmintrin.h>
#endif
#if defined(__x86_64__) && defined(__GNUC__)
#include <immintrin.h>
#endif

namespace grm_text_gz {

// ---------------------------------------------------------------------------------------------
// number formatting
// ---------------------------------------------------------------------------------------------
static const char DIGIT_PAIRS[201] =
    "00010203040506070809101112131415161718192021222324252627282930313233343536373839404142434445464748495051525354555657585960616263646566676869"
    "707172737475767778798081828384858687888990919293949596979899";

static inline char* put_uint(char* q, uint64_t v)
{
    char tmp[24];
    int k = 24;
    while (v >= 100) {   // two digits at a time, from the right
        const uint64_t r = v % 100;
        v /= 100;
        k -= 2;
        memcpy(tmp + k, DIGIT_PAIRS + 2 * r, 2);
    }
    if (v >= 10) { k -= 2; memcpy(tmp + k, DIGIT_PAIRS + 2 * v, 2); }
    else tmp[--k] = (char)('0' + v);
    memcpy(q, tmp + k, (size_t)(24 - k));
    return q + (24 - k);
}

// printf("%.6e") (the reference's `setf(scientific); precision(6)`, grm.cpp:326-327): the value is scaled by an exact
// power of ten to 7 significant digits; whenever the scaled value is close enough to a rounding boundary for the one
// floating-point rounding of the scaling to matter (or the exponent is outside the exact-power range) snprintf decides.
// put_e6_std does the first part: "[-]d.dddddde[+-]XX", 12 characters after the sign, or 0 if it cannot decide; it
// also leaves the digits as numbers (the leading one, three pairs, the exponent) for the encoder's tables.
struct E6Digits { uint8_t h0 = 0, p[3] = {0, 0, 0}, ee = 0, eneg = 0; };
static inline int put_e6_std(char* q, double v, E6Digits& dg)
{
    static const double P10[23] = {1e0, 1e1, 1e2, 1e3, 1e4, 1e5, 1e6, 1e7, 1e8, 1e9, 1e10, 1e11, 1e12, 1e13, 1e14, 1e15, 1e16, 1e17, 1e18, 1e19, 1e20, 1e21, 1e22};
    const double a = std::fabs(v);
    if (!(a >= 1e-15 && a < 1e15)) {
        if (v == 0.0 && !std::signbit(v)) { memcpy(q, "0.000000e+00", 12); dg = E6Digits(); return 12; }
        return 0;
    }
    uint64_t bits;
    memcpy(&bits, &a, 8);
    const int e2 = (int)((bits >> 52) & 0x7ff) - 1022;     // a = f * 2^e2, f in [0.5, 1) (a is normal in this range)
    int e = ((e2 - 1) * 1233) >> 12;                       // floor((e2 - 1) * log10(2)), possibly one off: fixed below
    for (int attempt = 0; attempt < 3; attempt++) {
        const int k = 6 - e;   // a * 10^k should lie in [1e6, 1e7)
        const double scaled = k >= 0 ? a * P10[k] : a / P10[-k];
        if (scaled < 1e6) { e--; continue; }
        if (scaled >= 1e7) { e++; continue; }
        const uint64_t fl = (uint64_t)scaled;
        const double frac = scaled - (double)fl;
        if (std::fabs(frac - 0.5) < 1e-5) return 0;   // too close to a tie to decide here
        uint32_t d = (uint32_t)fl + (frac > 0.5 ? 1u : 0u);
        int ee = e;
        if (d >= 10000000u) { d = 1000000u; ee++; }
        const int neg = std::signbit(v) ? 1 : 0;
        *q = '-';
        q += neg;
        // 7 digits: the leading one, then three pairs out of a 2-digit table
        const uint32_t hi = d / 10000u, lo = d - hi * 10000u;      // hi: 3 digits, lo: 4 digits
        const uint32_t h0 = hi / 100u, h12 = hi - h0 * 100u, l01 = lo / 100u, l23 = lo - l01 * 100u;
        q[0] = (char)('0' + h0);
        q[1] = '.';
        memcpy(q + 2, DIGIT_PAIRS + 2 * h12, 2);
        memcpy(q + 4, DIGIT_PAIRS + 2 * l01, 2);
        memcpy(q + 6, DIGIT_PAIRS + 2 * l23, 2);
        q[8] = 'e';
        q[9] = ee < 0 ? '-' : '+';
        const int ea = ee < 0 ? -ee : ee;   // <= 15
        memcpy(q + 10, DIGIT_PAIRS + 2 * ea, 2);
        dg.h0 = (uint8_t)h0; dg.p[0] = (uint8_t)h12; dg.p[1] = (uint8_t)l01; dg.p[2] = (uint8_t)l23;
        dg.ee = (uint8_t)ea; dg.eneg = (uint8_t)(ee < 0);
        return 12 + neg;
    }
    return 0;
}

int put_e6(char* q, double v)
{
    E6Digits dg;
    const int len = put_e6_std(q, v, dg);
    return len ? len : snprintf(q, 32, "%.6e", v);
}

// printf("%.15g") (iostream's default float format at precision 15, which is what the reference's GRM re-save prints,
// grm.cpp:367).  Fast path for the common case there - a double that was itself read from a decimal of at most seven
// significant digits: such a value prints as that decimal (the double is within 1.2e-16 relative of it, far less than
// half a unit of the 15th digit), so the digits are recovered by one exact scaling and checked by scaling back.
// Everything else goes to snprintf.
int put_g15(char* q, double v)
{
    static const double P10[23] = {1e0, 1e1, 1e2, 1e3, 1e4, 1e5, 1e6, 1e7, 1e8, 1e9, 1e10, 1e11, 1e12, 1e13, 1e14, 1e15, 1e16, 1e17, 1e18, 1e19, 1e20, 1e21, 1e22};
    char* const q0 = q;
    const double a = std::fabs(v);
    if (v == 0.0 && !std::signbit(v)) { *q = '0'; return 1; }
    if (!(a >= 1e-15 && a < 1e15)) return snprintf(q, 40, "%.15g", v);
    uint64_t bits;
    memcpy(&bits, &a, 8);
    const int e2 = (int)((bits >> 52) & 0x7ff) - 1022;
    int e = ((e2 - 1) * 1233) >> 12;
    for (int attempt = 0; attempt < 3; attempt++) {
        const int k = 6 - e;
        const double scaled = k >= 0 ? a * P10[k] : a / P10[-k];
        if (scaled < 1e6) { e--; continue; }
        if (scaled >= 1e7) { e++; continue; }
        uint64_t m = (uint64_t)(scaled + 0.5);
        int X = e;
        if (m >= 10000000ull) { m = 1000000ull; X++; }
        const int kk = 6 - X;
        const double back = kk >= 0 ? (double)m / P10[kk] : (double)m * P10[-kk];
        if (back != a) break;   // not a short decimal: let snprintf do it
        char dig[7];
        for (int t = 6; t >= 0; t--) { dig[t] = (char)('0' + m % 10); m /= 10; }
        int nd = 7;
        while (nd > 1 && dig[nd - 1] == '0') nd--;
        if (std::signbit(v)) *q++ = '-';
        if (X < -4 || X >= 15) {
            *q++ = dig[0];
            if (nd > 1) { *q++ = '.'; for (int t = 1; t < nd; t++) *q++ = dig[t]; }
            *q++ = 'e';
            int ee = X;
            if (ee < 0) { *q++ = '-'; ee = -ee; } else *q++ = '+';
            *q++ = (char)('0' + ee / 10); *q++ = (char)('0' + ee % 10);
        } else if (X >= 0) {
            for (int t = 0; t <= X; t++) *q++ = t < nd ? dig[t] : '0';
            if (nd > X + 1) { *q++ = '.'; for (int t = X + 1; t < nd; t++) *q++ = dig[t]; }
        } else {
            *q++ = '0'; *q++ = '.';
            for (int t = 0; t < -X - 1; t++) *q++ = '0';
            for (int t = 0; t < nd; t++) *q++ = dig[t];
        }
        return (int)(q - q0);
    }
    return snprintf(q0, 40, "%.15g", v);
}

// ---------------------------------------------------------------------------------------------
// member header with the member's own length (RFC 1952 FEXTRA subfield 'G','M', 4 bytes little endian)
// ---------------------------------------------------------------------------------------------
const char MEMBER_HEADER[MEMBER_HEADER_BYTES] = {(char)0x1f, (char)0x8b, 8, 4 /* FEXTRA */, 0, 0, 0, 0, 0, 3,
                                                 8, 0 /* XLEN */, 'G', 'M', 4, 0 /* LEN */, 0, 0, 0, 0};

void patch_member_size(std::string& out, size_t member_start)
{
    const uint64_t size = out.size() - member_start;
    // a member that does not fit 32 bits keeps 0 = "unknown": readers then fall back to sequential decoding
    const uint32_t v = size <= 0xFFFFFFFFull ? (uint32_t)size : 0;
    for (int k = 0; k < 4; k++) out[member_start + 16 + k] = (char)((v >> (8 * k)) & 255);
}

uint32_t member_size_from_header(const unsigned char* h, size_t avail)
{
    if (avail < (size_t)MEMBER_HEADER_BYTES) return 0;
    if (h[0] != 0x1f || h[1] != 0x8b || h[2] != 8 || h[3] != 4) return 0;
    if (h[10] != 8 || h[11] != 0 || h[12] != 'G' || h[13] != 'M' || h[14] != 4 || h[15] != 0) return 0;
    return (uint32_t)h[16] | ((uint32_t)h[17] << 8) | ((uint32_t)h[18] << 16) | ((uint32_t)h[19] << 24);
}

// ---------------------------------------------------------------------------------------------
// CRC-32 of the text (RFC 1952)
// ---------------------------------------------------------------------------------------------
#if defined(__x86_64__) && defined(__GNUC__)
// CRC-32 (gzip polynomial, reflected) of a multiple of 16 bytes, >= 64, by carry-less multiplication: four 128-bit lanes
// folded 64 bytes at a time, then reduced (Gopal et al., "Fast CRC computation for generic polynomials using PCLMULQDQ").
__attribute__((target("pclmul,sse4.1")))
static uint32_t crc32_clmul(uint32_t crc, const unsigned char* buf, size_t len)   // crc: the register itself (zlib's value inverted)
{
    const __m128i k1k2 = _mm_set_epi64x(0x01c6e41596, 0x0154442bd4);
    const __m128i k3k4 = _mm_set_epi64x(0x00ccaa009e, 0x01751997d0);
    const __m128i k5k0 = _mm_set_epi64x(0, 0x0163cd6124);
    const __m128i poly = _mm_set_epi64x(0x01f7011641, 0x01db710641);
    __m128i x1 = _mm_loadu_si128((const __m128i*)(buf + 0)), x2 = _mm_loadu_si128((const __m128i*)(buf + 16));
    __m128i x3 = _mm_loadu_si128((const __m128i*)(buf + 32)), x4 = _mm_loadu_si128((const __m128i*)(buf + 48));
    x1 = _mm_xor_si128(x1, _mm_cvtsi32_si128((int)crc));
    buf += 64; len -= 64;
    while (len >= 64) {
        __m128i x5 = _mm_clmulepi64_si128(x1, k1k2, 0x00), x6 = _mm_clmulepi64_si128(x2, k1k2, 0x00);
        __m128i x7 = _mm_clmulepi64_si128(x3, k1k2, 0x00), x8 = _mm_clmulepi64_si128(x4, k1k2, 0x00);
        x1 = _mm_clmulepi64_si128(x1, k1k2, 0x11); x2 = _mm_clmulepi64_si128(x2, k1k2, 0x11);
        x3 = _mm_clmulepi64_si128(x3, k1k2, 0x11); x4 = _mm_clmulepi64_si128(x4, k1k2, 0x11);
        x1 = _mm_xor_si128(_mm_xor_si128(x1, x5), _mm_loadu_si128((const __m128i*)(buf + 0)));
        x2 = _mm_xor_si128(_mm_xor_si128(x2, x6), _mm_loadu_si128((const __m128i*)(buf + 16)));
        x3 = _mm_xor_si128(_mm_xor_si128(x3, x7), _mm_loadu_si128((const __m128i*)(buf + 32)));
        x4 = _mm_xor_si128(_mm_xor_si128(x4, x8), _mm_loadu_si128((const __m128i*)(buf + 48)));
        buf += 64; len -= 64;
    }
#define GRM_CRC_FOLD(a, b) _mm_xor_si128(_mm_xor_si128(_mm_clmulepi64_si128(a, k3k4, 0x11), _mm_clmulepi64_si128(a, k3k4, 0x00)), b)
    x1 = GRM_CRC_FOLD(x1, x2); x1 = GRM_CRC_FOLD(x1, x3); x1 = GRM_CRC_FOLD(x1, x4);
    while (len >= 16) { x1 = GRM_CRC_FOLD(x1, _mm_loadu_si128((const __m128i*)buf)); buf += 16; len -= 16; }
#undef GRM_CRC_FOLD
    // 128 -> 64 bits
    const __m128i mask32 = _mm_setr_epi32(~0, 0, ~0, 0);
    __m128i t = _mm_clmulepi64_si128(x1, k3k4, 0x10);
    x1 = _mm_xor_si128(_mm_srli_si128(x1, 8), t);
    t = _mm_srli_si128(x1, 4);
    x1 = _mm_and_si128(x1, mask32);
    x1 = _mm_xor_si128(_mm_clmulepi64_si128(x1, k5k0, 0x00), t);
    // Barrett reduction to 32 bits
    t = _mm_and_si128(x1, mask32);
    t = _mm_clmulepi64_si128(t, poly, 0x10);
    t = _mm_and_si128(t, mask32);
    t = _mm_clmulepi64_si128(t, poly, 0x00);
    x1 = _mm_xor_si128(x1, t);
    return (uint32_t)_mm_extract_epi32(x1, 1);
}
#endif

// zlib's crc32() semantics; the bulk by carry-less multiplication where the CPU has it (about 8x zlib's table walk)
static uint32_t text_crc32(uint32_t crc, const unsigned char* buf, size_t len)
{
#if defined(__x86_64__) && defined(__GNUC__)
    static const bool have = __builtin_cpu_supports("pclmul") && __builtin_cpu_supports("sse4.1");
    if (have && len >= 64) {
        const size_t bulk = len & ~(size_t)15;
        crc = ~crc32_clmul(~crc, buf, bulk);
        buf += bulk;
        len -= bulk;
    }
#endif
    return len ? (uint32_t)crc32(crc, (const Bytef*)buf, (uInt)len) : crc;
}

namespace {

// One formatted line and where its four fields start; a field includes the separator that ends it.  The longest
// line is 49 bytes ("-2147483648" for N, "%.6e" of a double), the longest field 16; the text sits between two pads so
// that 16-byte loads and stores around any field stay inside the object.
struct Line {
    char front[16];
    char s[96];
    int fs[5] = {0, 0, 0, 0, 0};   // fs[4] = length of the line
    int std_value = 0;             // the value is "[-]d.dddddde[+-]XX": 12 + neg characters (put_e6_std)
    int neg = 0;
    E6Digits dg;                   // its digits, then
    // against the line the same formatter produced just before: same row and j as wide (then only the last `jchanged`
    // digits of j differ up to the N field), and N the same
    int head_known = 0, jchanged = 0, same_n = 0;
};

static inline void copy16(char* dst, const char* src) { memcpy(dst, src, 16); }   // one unaligned 16-byte move

// Formats the lines of one row after another: the row index is formatted once, the column index counts up in place,
// and the text of N is kept while N does not change (without missing genotypes it never does).
struct RowFormatter {
    char ipre[32] = {}; int ilen = 0;   // "i \t"
    char jdec[32] = {}; int jlen = 0;   // decimal digits of j + 1, most significant first, then '\t'; jlen counts the digits
    char nbuf[32] = {}; int nlen = 0;   // "N \t"
    int32_t ncur = 0;
    bool nvalid = false, first = true;
    void start_row(int64_t i)
    {
        first = true;
        char* q = put_uint(ipre, (uint64_t)(i + 1));
        *q++ = '\t';
        ilen = (int)(q - ipre);
        jdec[0] = '0'; jdec[1] = '\t'; jlen = 1;   // incremented to "1" by the first next()
    }
    inline int bump_j()   // number of digits that changed, 0 if the number grew by a digit
    {
        int k = jlen - 1;
        while (k >= 0 && jdec[k] == '9') jdec[k--] = '0';
        if (k >= 0) { jdec[k]++; return jlen - k; }
        memmove(jdec + 1, jdec, (size_t)jlen + 1); jdec[0] = '1'; jlen++;
        return 0;
    }
    inline void set_n(int32_t Nv)
    {
        char* q = nbuf;
        if (Nv < 0) { *q++ = '-'; q = put_uint(q, (uint64_t)(-(int64_t)Nv)); } else q = put_uint(q, (uint64_t)Nv);
        *q++ = '\t';
        nlen = (int)(q - nbuf);
        ncur = Nv;
        nvalid = true;
    }
    // fields 0..2 ("i \t", "j \t", "N \t"; each at most 12 bytes); returns where the value field starts
    inline char* head(Line& L, int32_t Nv)
    {
        L.jchanged = bump_j();
        L.head_known = !first && L.jchanged > 0;
        first = false;
        L.same_n = nvalid && Nv == ncur;
        if (!L.same_n) set_n(Nv);
        char* q = L.s;
        copy16(q, ipre); q += ilen;
        L.fs[0] = 0; L.fs[1] = ilen;
        copy16(q, jdec); q += jlen + 1;
        L.fs[2] = (int)(q - L.s);
        copy16(q, nbuf); q += nlen;
        L.fs[3] = (int)(q - L.s);
        return q;
    }
    // output_grm_vec's line (grm.cpp:328-330): A in scientific notation, 6 decimals
    inline void next(Line& L, int32_t Nv, float Av)
    {
        char* q = head(L, Nv);
        const int len = put_e6_std(q, (double)Av, L.dg);
        L.std_value = len;
        L.neg = len - 12;
        q += len ? len : snprintf(q, 32, "%.6e", (double)Av);
        *q++ = '\n';
        L.fs[4] = (int)(q - L.s);
    }
    // output_grm_MatrixXf's line (grm.cpp:367): `setprecision(15) << double`, i.e. printf("%.15g")
    inline void next15(Line& L, int32_t Nv, double Av)
    {
        char* q = head(L, Nv);
        L.std_value = 0;
        q += put_g15(q, Av);
        *q++ = '\n';
        L.fs[4] = (int)(q - L.s);
    }
};

// the two sources of lines: the packed float/int32 triangle of make_grm, and the packed double/double one of the
// GRM re-save path (N is printed as `(int)N`, grm.cpp:367)
struct SrcE6 {
    const float* A; const int32_t* N;
    inline void fill(RowFormatter& rf, Line& L, size_t o) const { rf.next(L, N[o], A[o]); }
};
struct SrcG15 {
    const double* A; const double* N;
    inline void fill(RowFormatter& rf, Line& L, size_t o) const { rf.next15(L, (int32_t)N[o], A[o]); }
};

// ---------------------------------------------------------------------------------------------
// deflate pieces (RFC 1951)
// ---------------------------------------------------------------------------------------------
constexpr int NLIT = 286, NDIST = 30, MIN_MATCH = 4, MAX_MATCH = 258;
const uint16_t LEN_BASE[29] = {3, 4, 5, 6, 7, 8, 9, 10, 11, 13, 15, 17, 19, 23, 27, 31, 35, 43, 51, 59, 67, 83, 99, 115, 131, 163, 195, 227, 258};
const uint8_t LEN_EXTRA[29] = {0, 0, 0, 0, 0, 0, 0, 0, 1, 1, 1, 1, 2, 2, 2, 2, 3, 3, 3, 3, 4, 4, 4, 4, 5, 5, 5, 5, 0};
const uint16_t DIST_BASE[30] = {1, 2, 3, 4, 5, 7, 9, 13, 17, 25, 33, 49, 65, 97, 129, 193, 257, 385, 513, 769, 1025, 1537, 2049, 3073, 4097, 6145, 8193, 12289, 16385, 24577};
const uint8_t DIST_EXTRA[30] = {0, 0, 0, 0, 1, 1, 2, 2, 3, 3, 4, 4, 5, 5, 6, 6, 7, 7, 8, 8, 9, 9, 10, 10, 11, 11, 12, 12, 13, 13};

inline int len_symbol(int len)   // index into LEN_BASE
{
    int k = 28;
    while (LEN_BASE[k] > len) k--;
    return k;
}
inline int dist_symbol(int dist)
{
    int k = 29;
    while (DIST_BASE[k] > dist) k--;
    return k;
}

// Huffman code lengths (<= limit) for freq[0..n); symbols with freq 0 get length 0.
void huffman_lengths(const uint32_t* freq, int n, int limit, uint8_t* len)
{
    std::vector<uint64_t> f(freq, freq + n);
    for (;;) {
        struct Node { uint64_t w; int a, b; };
        std::vector<Node> nodes;
        typedef std::pair<uint64_t, int> QE;
        std::priority_queue<QE, std::vector<QE>, std::greater<QE>> pq;
        for (int i = 0; i < n; i++)
            if (f[i]) { nodes.push_back({f[i], -1 - i, 0}); pq.push({f[i], (int)nodes.size() - 1}); }
        std::fill(len, len + n, 0);
        if (nodes.empty()) return;
        if (nodes.size() == 1) { len[-1 - nodes[0].a] = 1; return; }
        while (pq.size() > 1) {
            QE x = pq.top(); pq.pop();
            QE y = pq.top(); pq.pop();
            nodes.push_back({x.first + y.first, x.second, y.second});
            pq.push({x.first + y.first, (int)nodes.size() - 1});
        }
        // depths, root = last node
        std::vector<int> depth(nodes.size(), 0);
        int maxd = 0;
        for (int k = (int)nodes.size() - 1; k >= 0; k--) {
            if (nodes[k].a < 0) { len[-1 - nodes[k].a] = (uint8_t)std::min(depth[k], 255); maxd = std::max(maxd, depth[k]); }
            else { depth[nodes[k].a] = depth[k] + 1; depth[nodes[k].b] = depth[k] + 1; }
        }
        if (maxd <= limit) return;
        for (int i = 0; i < n; i++)
            if (f[i]) f[i] = (f[i] + 1) / 2;   // flatten and retry
    }
}

// canonical codes, bit-reversed for the LSB-first bit stream
void canonical_codes(const uint8_t* len, int n, uint16_t* code)
{
    int bl_count[16] = {0};
    for (int i = 0; i < n; i++) bl_count[len[i]]++;
    bl_count[0] = 0;
    int next_code[16] = {0}, c = 0;
    for (int b = 1; b <= 15; b++) { c = (c + bl_count[b - 1]) << 1; next_code[b] = c; }
    for (int i = 0; i < n; i++) {
        if (!len[i]) { code[i] = 0; continue; }
        int v = next_code[len[i]]++, r = 0;
        for (int b = 0; b < len[i]; b++) { r = (r << 1) | (v & 1); v >>= 1; }
        code[i] = (uint16_t)r;
    }
}

// LSB-first bit stream into a 1 MiB staging buffer that is appended to `out` whenever it runs low (the caller calls
// room() once per line; a line never needs more than a few hundred bytes).  put() takes up to 56 bits: the accumulator
// never holds more than 7 bits between calls, its whole bytes are stored with one 8-byte write and p moves past them.
#if defined(__BYTE_ORDER__) && __BYTE_ORDER__ != __ORDER_LITTLE_ENDIAN__
#error "BitWriter stores the accumulator as a little-endian word"
#endif
struct BitWriter {
    std::string& out;
    std::vector<unsigned char> buf;
    unsigned char* p;
    uint64_t acc = 0;
    int nbits = 0;
    explicit BitWriter(std::string& o) : out(o), buf(1 << 20), p(buf.data()) {}
    inline void put(uint64_t v, int n)
    {
        acc |= v << nbits;
        nbits += n;
        memcpy(p, &acc, 8);
        const int whole = nbits >> 3;
        p += whole;
        acc >>= 8 * whole;
        nbits &= 7;
    }
    inline void room()
    {
        if ((size_t)(p - buf.data()) + 4096 > buf.size()) { out.append((const char*)buf.data(), (size_t)(p - buf.data())); p = buf.data(); }
    }
    void finish()
    {
        if (nbits > 0) { *p++ = (unsigned char)(acc & 255); acc = 0; nbits = 0; }
        out.append((const char*)buf.data(), (size_t)(p - buf.data()));
        p = buf.data();
    }
};

// ---------------------------------------------------------------------------------------------
// the tokeniser: line -> shared runs with the previous line + literals
// ---------------------------------------------------------------------------------------------
struct CountSink {
    uint32_t lit[NLIT] = {0}, dist[NDIST] = {0};
    inline void literals(const uint8_t* b, int n) { for (int k = 0; k < n; k++) lit[b[k]]++; }
    inline void match(int len, int d) { lit[257 + len_symbol(len)]++; dist[dist_symbol(d)]++; }
    inline void value_std(const uint8_t* v, int neg, const E6Digits&, bool with_exp) { lit['-'] += (uint32_t)neg; literals(v, with_exp ? 12 : 8); }
};

struct EmitSink {
    BitWriter& bw;
    const uint16_t* lcode; const uint8_t* llen; const uint16_t* dcode; const uint8_t* dlen;
    // per match length / small distance: symbol + extra bits folded into one (value, nbits) pair
    uint32_t mlen_bits[MAX_MATCH + 1]; uint8_t mlen_n[MAX_MATCH + 1];
    uint32_t mdist_bits[513]; uint8_t mdist_n[513];
    EmitSink(BitWriter& b, const uint16_t* lc, const uint8_t* ll, const uint16_t* dc, const uint8_t* dl) : bw(b), lcode(lc), llen(ll), dcode(dc), dlen(dl)
    {
        for (int len = 3; len <= MAX_MATCH; len++) {
            const int k = len_symbol(len);
            mlen_bits[len] = lcode[257 + k] | ((uint32_t)(len - LEN_BASE[k]) << llen[257 + k]);
            mlen_n[len] = (uint8_t)(llen[257 + k] + LEN_EXTRA[k]);
        }
        for (int d = 1; d <= 512; d++) {
            const int k = dist_symbol(d);
            mdist_bits[d] = dcode[k] | ((uint32_t)(d - DIST_BASE[k]) << dlen[k]);
            mdist_n[d] = (uint8_t)(dlen[k] + DIST_EXTRA[k]);
        }
        build_value_tables();
    }
    inline void literals(const uint8_t* b, int n)   // three codes (<= 15 bits each) per put
    {
        int k = 0;
        for (; k + 2 < n; k += 3) {
            const int l0 = llen[b[k]], l1 = llen[b[k + 1]];
            bw.put((uint64_t)lcode[b[k]] | ((uint64_t)lcode[b[k + 1]] << l0) | ((uint64_t)lcode[b[k + 2]] << (l0 + l1)), l0 + l1 + llen[b[k + 2]]);
        }
        for (; k < n; k++) bw.put(lcode[b[k]], llen[b[k]]);
    }
    // "[-]d.dddddd" and, if with_exp, "e[+-]XX" as literals, from the digits as numbers: [sign, leading digit, '.'],
    // three digit pairs and ['e', sign, exponent pair] each have their codes ready in a table, and what is absent goes
    // out as zero bits - five look-ups, three puts, no data-dependent branch.  (Code lengths are whatever the Huffman
    // construction gave; should three of these groups ever exceed a put's 56 bits, the plain route below takes over.)
    inline void value_std(const uint8_t* v, int neg, const E6Digits& dg, bool with_exp)
    {
        if (!tables_fit) { value_chars(v, neg, with_exp); return; }
        const uint64_t em = with_exp ? ~0ull : 0ull;
        const int ln = lead_n[neg][dg.h0], p0 = pair_n[dg.p[0]], p1 = pair_n[dg.p[1]];
        bw.put(lead_bits[neg][dg.h0] | ((uint64_t)pair_bits[dg.p[0]] << ln), ln + p0);
        bw.put((uint64_t)pair_bits[dg.p[1]] | ((uint64_t)pair_bits[dg.p[2]] << p1), p1 + pair_n[dg.p[2]]);
        bw.put(exp_bits[dg.eneg][dg.ee] & em, exp_n[dg.eneg][dg.ee] & (int)em);
    }
    inline void value_chars(const uint8_t* v, int neg, bool with_exp)
    {
        if (neg) bw.put(lcode['-'], llen['-']);
        literals(v, with_exp ? 12 : 8);
    }
    uint64_t lead_bits[2][10], exp_bits[2][16];
    uint32_t pair_bits[100];
    uint8_t lead_n[2][10], exp_n[2][16], pair_n[100];
    bool tables_fit = true;
    void build_value_tables()
    {
        int max_lead = 0, max_pair = 0, max_exp = 0;
        auto cat = [&](uint64_t& bits, int& n, uint8_t ch) { bits |= (uint64_t)lcode[ch] << n; n += llen[ch]; };
        for (int a = 0; a < 10; a++)
            for (int b = 0; b < 10; b++) {
                uint64_t bits = 0; int n = 0;
                cat(bits, n, (uint8_t)('0' + a)); cat(bits, n, (uint8_t)('0' + b));
                pair_bits[10 * a + b] = (uint32_t)bits; pair_n[10 * a + b] = (uint8_t)n;
                max_pair = std::max(max_pair, n);
            }
        for (int neg = 0; neg < 2; neg++) {
            for (int d = 0; d < 10; d++) {
                uint64_t bits = 0; int n = 0;
                if (neg) cat(bits, n, '-');
                cat(bits, n, (uint8_t)('0' + d)); cat(bits, n, '.');
                lead_bits[neg][d] = bits; lead_n[neg][d] = (uint8_t)n;
                max_lead = std::max(max_lead, n);
            }
            for (int e = 0; e < 16; e++) {
                uint64_t bits = 0; int n = 0;
                cat(bits, n, 'e'); cat(bits, n, neg ? '-' : '+');
                cat(bits, n, (uint8_t)('0' + e / 10)); cat(bits, n, (uint8_t)('0' + e % 10));
                exp_bits[neg][e] = bits; exp_n[neg][e] = (uint8_t)n;
                max_exp = std::max(max_exp, n);
            }
        }
        tables_fit = max_lead + max_pair <= 56 && 2 * max_pair <= 56 && max_exp <= 56;
    }
    inline void match(int len, int d)
    {
        if (d <= 512)   // <= 20 bits of length, <= 23 of distance
            bw.put((uint64_t)mlen_bits[len] | ((uint64_t)mdist_bits[d] << mlen_n[len]), mlen_n[len] + mdist_n[d]);
        else {
            const int k = dist_symbol(d);
            bw.put(mlen_bits[len], mlen_n[len]);
            bw.put((uint64_t)dcode[k] | ((uint64_t)(d - DIST_BASE[k]) << dlen[k]), dlen[k] + DIST_EXTRA[k]);
        }
    }
};

// Length of the common prefix of a[0..n) and b[0..n), and of the common suffix of a_end[-n..0) and b_end[-n..0), looking
// at 16 bytes at most (a longer field just gets a shorter run); both touch 16 bytes from the pointer (Line is padded for it).
#if defined(__SSE2__)
static inline int common_prefix16(const char* a, const char* b, int n)
{
    const unsigned ne = ~(unsigned)_mm_movemask_epi8(_mm_cmpeq_epi8(_mm_loadu_si128((const __m128i*)a), _mm_loadu_si128((const __m128i*)b))) | (1u << n);
    return __builtin_ctz(ne);
}
static inline int common_suffix16(const char* a_end, const char* b_end, int n)
{
    if (n <= 0) return 0;
    const unsigned eq = (unsigned)_mm_movemask_epi8(_mm_cmpeq_epi8(_mm_loadu_si128((const __m128i*)(a_end - 16)), _mm_loadu_si128((const __m128i*)(b_end - 16))));
    // bit 15 of the mask = the last byte: count equal bytes from the top; bit 31 - n stops the count at n, bit 15 at 16
    const unsigned ne = ((~eq) << 16) | (0x80000000u >> n) | 0x8000u;
    return __builtin_clz(ne);
}
#else
static inline int common_prefix16(const char* a, const char* b, int n)
{
    int c = 0;
    while (c < n && a[c] == b[c]) c++;
    return c;
}
static inline int common_suffix16(const char* a_end, const char* b_end, int n)
{
    int t = 0;
    while (t < n && a_end[-1 - t] == b_end[-1 - t]) t++;
    return t;
}
#endif

// bit x = (a[x] == b[x]) for x < 64
static inline uint64_t equal_mask64(const char* a, const char* b)
{
#if defined(__SSE2__)
    uint64_t m = 0;
    for (int k = 0; k < 4; k++)
        m |= (uint64_t)(unsigned)_mm_movemask_epi8(_mm_cmpeq_epi8(_mm_loadu_si128((const __m128i*)(a + 16 * k)), _mm_loadu_si128((const __m128i*)(b + 16 * k)))) << (16 * k);
    return m;
#else
    uint64_t m = 0;
    for (int x = 0; x < 64; x++) m |= (uint64_t)(a[x] == b[x]) << x;
    return m;
#endif
}

template <class Sink>
struct Tokeniser {
    Sink& sink;
    int run_len = 0, run_dist = 0;
    uint8_t run_head[MIN_MATCH];   // first bytes of the pending run, in case it ends up too short for a match
    explicit Tokeniser(Sink& s) : sink(s) {}

    inline void flush_run()
    {
        if (run_len >= MIN_MATCH) sink.match(run_len, run_dist);
        else sink.literals(run_head, run_len);
        run_len = 0;
    }
    inline void add_run(const char* bytes, int len, int dist)
    {
        if (len <= 0) return;
        if (run_len > 0 && (run_dist != dist || run_len + len > MAX_MATCH)) flush_run();
        for (int k = 0; k < len && run_len + k < MIN_MATCH; k++) run_head[run_len + k] = (uint8_t)bytes[k];
        run_len += len;
        run_dist = dist;
    }
    inline void add_literals(const char* bytes, int len)
    {
        if (len <= 0) return;
        flush_run();
        sink.literals((const uint8_t*)bytes, len);
    }
    // cur[0, lim) against prev[0, lim), byte for byte (distance = the previous line's length): equal stretches become
    // runs when they are MIN_MATCH bytes or more, or touch either end (there they may join the run of the neighbouring
    // line or zone); everything else is literals.  eq: bit x = (cur[x] == prev[x]).
    inline void aligned(const char* cur, uint64_t eq, int lim, int dist)
    {
        const uint64_t all = lim >= 64 ? ~0ull : (1ull << lim) - 1;
        eq &= all;
        // keep the stretches that become runs, drop the others from the mask up front (the digits of the value agree
        // with the line before in one place out of ten, at random: a loop over those would mispredict its branches)
        const uint64_t s4 = eq & (eq >> 1) & (eq >> 2) & (eq >> 3);          // a stretch of >= MIN_MATCH (= 4) starts or continues here
        uint64_t keep = s4 | (s4 << 1) | (s4 << 2) | (s4 << 3);
        keep |= eq & ~(eq + 1);                                               // the stretch that begins at byte 0
        const uint64_t ne = ~eq & all;
        keep |= ne ? all & ~((2ull << (63 - __builtin_clzll(ne))) - 1) : all;   // the stretch that ends at byte lim - 1
        int pos = 0;
        while (keep) {
            const int r0 = __builtin_ctzll(keep);
            const uint64_t m = ~(keep >> r0);
            const int run = m ? __builtin_ctzll(m) : 64 - r0;
            add_literals(cur + pos, r0 - pos);
            add_run(cur + r0, run, dist);
            pos = r0 + run;
            keep = pos >= 64 ? 0 : keep & ~((1ull << pos) - 1);
        }
        add_literals(cur + pos, lim - pos);
    }
    // `prev` == nullptr: first line of a member (nothing to refer to)
    inline void line(const Line& cur, const Line* prev)
    {
        if (!prev) { add_literals(cur.s, cur.fs[4]); return; }
        const int pl = prev->fs[4], cl = cur.fs[4];
        // The usual case, told by the formatter: the same row as the line before and j as wide, so up to the N field
        // only the last digits of j differ; and a value in the standard form.  The digits of the value are close to
        // random and go out as literals; its exponent, when it is the one of the line before, and the newline start
        // the run that the head of the next line continues (distance = this line's length).
        if (cur.head_known && cur.std_value) {
            const int n0 = cur.fs[2] - 1, a = n0 - cur.jchanged;   // n0: the tab that ends j
            add_run(cur.s, a, pl);
            flush_run();
            sink.literals((const uint8_t*)cur.s + a, cur.jchanged);
            if (cur.same_n) {
                add_run(cur.s + n0, cur.fs[3] - n0, pl);
                flush_run();
            } else {   // the tab and the leading digits N shares with the line before, then literals up to and including its tab
                const int nl = cur.fs[3] - cur.fs[2], c = common_prefix16(cur.s + cur.fs[2], prev->s + cur.fs[2], std::min(nl, prev->fs[3] - prev->fs[2]) - 1);
                add_run(cur.s + n0, 1 + c, pl);
                flush_run();
                sink.literals((const uint8_t*)cur.s + cur.fs[2] + c, nl - c);
            }
            uint32_t ce4, pe4;
            memcpy(&ce4, cur.s + cl - 5, 4);
            memcpy(&pe4, prev->s + pl - 5, 4);
            const bool same_exp = ce4 == pe4;   // prev need not be in the standard form: its last five bytes are what counts
            sink.value_std((const uint8_t*)cur.s + cur.fs[3] + cur.neg, cur.neg, cur.dg, !same_exp);
            run_len = same_exp ? 5 : 1;
            run_dist = cl;
            memcpy(run_head, cur.s + cl - run_len, 4);   // only looked at while the run is shorter than MIN_MATCH
            return;
        }
        // i, j and N as wide as on the line before and the lines equally long: they line up byte for byte
        if (cur.fs[1] == prev->fs[1] && cur.fs[2] == prev->fs[2] && cur.fs[3] == prev->fs[3] && cl == pl && cl <= 64) {
            aligned(cur.s, equal_mask64(cur.s, prev->s), cl, pl);
            return;
        }
        for (int f = 0; f < 4; f++) field(cur, *prev, f);
    }
    // one field against the same field of the line before: shared prefix, literals, shared suffix
    inline void field(const Line& cur, const Line& prev, int f)
    {
        const int pl = prev.fs[4];
        const int cs = cur.fs[f], ce = cur.fs[f + 1], ps = prev.fs[f], pe = prev.fs[f + 1];
        const int clen = ce - cs, plen = pe - ps, mn = std::min(clen, plen);
        const int c = common_prefix16(cur.s + cs, prev.s + ps, mn);   // looks at 16 bytes at most
        add_run(cur.s + cs, c, pl + cs - ps);
        if (c == clen && c == plen) return;   // the field repeats as a whole
        const int t = common_suffix16(cur.s + ce, prev.s + pe, mn - c);
        add_literals(cur.s + cs + c, clen - c - t);
        add_run(cur.s + ce - t, t, pl + ce - pe);
    }
};

inline void put_le32(std::string& out, uint32_t v)
{
    char b[4] = {(char)(v & 255), (char)((v >> 8) & 255), (char)((v >> 16) & 255), (char)(v >> 24)};
    out.append(b, 4);
}

}  // namespace

template <class Src>
static void format_impl(const Src& src, int64_t row_begin, int64_t row_end, std::string& text)
{
    RowFormatter rf;
    Line L;
    for (int64_t i = row_begin; i < row_end; i++) {
        rf.start_row(i);
        const size_t base = (size_t)(i * (i + 1) / 2);
        for (int64_t j = 0; j <= i; j++) {
            src.fill(rf, L, base + j);
            text.append(L.s, (size_t)L.fs[4]);
        }
    }
}

template <class Src>
static void encode_impl(const Src& src, int64_t row_begin, int64_t row_end, std::string& out)
{
    // ---- pass 1: symbol statistics from a sample of the member's lines (runs of consecutive lines from 16 rows) ----
    CountSink cs;
    {
        Tokeniser<CountSink> tk(cs);
        RowFormatter rf;
        Line buf[2];
        const int64_t rows = row_end - row_begin;
        const int64_t nsample = std::min<int64_t>(16, rows);
        for (int64_t s = 0; s < nsample; s++) {
            const int64_t i = row_begin + (nsample > 1 ? s * (rows - 1) / (nsample - 1) : 0);
            rf.start_row(i);
            const size_t base = (size_t)(i * (i + 1) / 2);
            const int64_t cnt = std::min<int64_t>(i + 1, 512);
            const Line* prev = nullptr;
            // the last `cnt` lines of the row (the widest column indices) would be as good; the first are simplest
            for (int64_t j = 0; j < cnt; j++) {
                Line& cur = buf[j & 1];
                src.fill(rf, cur, base + j);
                tk.line(cur, prev);
                prev = &cur;
            }
            tk.flush_run();
        }
    }
    // every symbol keeps a code: the sample need not have seen everything the member holds
    for (int k = 0; k < NLIT; k++) cs.lit[k] = cs.lit[k] * 16 + 1;
    for (int k = 0; k < NDIST; k++) cs.dist[k] = cs.dist[k] * 16 + 1;
    uint8_t llen[NLIT], dlen[NDIST];
    uint16_t lcode[NLIT], dcode[NDIST];
    huffman_lengths(cs.lit, NLIT, 15, llen);
    huffman_lengths(cs.dist, NDIST, 15, dlen);
    canonical_codes(llen, NLIT, lcode);
    canonical_codes(dlen, NDIST, dcode);

    // ---- gzip member header (RFC 1952) and the dynamic-Huffman block header (RFC 1951, 3.2.7) ----
    // FEXTRA carries one subfield 'G','M' = the byte length of this member (patched in below), the way BGZF does it:
    // a reader that knows it can hop from member to member and inflate them in parallel; every other gzip reader
    // skips the field.
    const size_t member_start = out.size();
    out.append(MEMBER_HEADER, MEMBER_HEADER_BYTES);
    BitWriter bw(out);
    bw.put(1, 1);   // BFINAL
    bw.put(2, 2);   // BTYPE = dynamic Huffman
    bw.put(NLIT - 257, 5);
    bw.put(NDIST - 1, 5);
    bw.put(19 - 4, 4);
    uint32_t clfreq[19] = {0};
    for (int k = 0; k < NLIT; k++) clfreq[llen[k]]++;
    for (int k = 0; k < NDIST; k++) clfreq[dlen[k]]++;
    uint8_t cllen[19];
    uint16_t clcode[19];
    huffman_lengths(clfreq, 19, 7, cllen);
    canonical_codes(cllen, 19, clcode);
    static const int CL_ORDER[19] = {16, 17, 18, 0, 8, 7, 9, 6, 10, 5, 11, 4, 12, 3, 13, 2, 14, 1, 15};
    for (int k = 0; k < 19; k++) bw.put(cllen[CL_ORDER[k]], 3);
    for (int k = 0; k < NLIT; k++) bw.put(clcode[llen[k]], cllen[llen[k]]);     // plain lengths, no repeat codes:
    for (int k = 0; k < NDIST; k++) bw.put(clcode[dlen[k]], cllen[dlen[k]]);    // ~250 bytes per member

    // ---- pass 2: format, tokenise and emit every line; CRC of the text in 32 KiB pieces ----
    EmitSink es(bw, lcode, llen, dcode, dlen);
    Tokeniser<EmitSink> tk(es);
    RowFormatter rf;
    Line buf[2];
    int which = 0;
    const Line* prev = nullptr;
    constexpr size_t STAGE = 1 << 15;
    std::vector<char> stage(STAGE + sizeof(Line::s));
    size_t staged = 0;
    uint32_t crc = 0;
    uint64_t isize = 0;
    for (int64_t i = row_begin; i < row_end; i++) {
        rf.start_row(i);
        const size_t base = (size_t)(i * (i + 1) / 2);
        for (int64_t j = 0; j <= i; j++) {
            Line& cur = buf[which];
            which ^= 1;
            src.fill(rf, cur, base + j);
            bw.room();
            tk.line(cur, prev);
            prev = &cur;
            memcpy(stage.data() + staged, cur.s, 64);   // a line is at most 49 bytes; fixed size = four 16-byte moves
            staged += (size_t)cur.fs[4];
            if (staged >= STAGE) {
                crc = text_crc32(crc, (const unsigned char*)stage.data(), staged);
                isize += staged;
                staged = 0;
            }
        }
    }
    tk.flush_run();
    crc = text_crc32(crc, (const unsigned char*)stage.data(), staged);
    isize += staged;
    bw.put(lcode[256], llen[256]);   // end of block
    bw.finish();
    put_le32(out, crc);
    put_le32(out, (uint32_t)(isize & 0xFFFFFFFFull));
    patch_member_size(out, member_start);
}

void format_rows(const float* A, const int32_t* N, int64_t row_begin, int64_t row_end, std::string& text)
{
    format_impl(SrcE6{A, N}, row_begin, row_end, text);
}
void encode_rows(const float* A, const int32_t* N, int64_t row_begin, int64_t row_end, std::string& out)
{
    encode_impl(SrcE6{A, N}, row_begin, row_end, out);
}
void format_rows_g15(const double* A, const double* N, int64_t row_begin, int64_t row_end, std::string& text)
{
    format_impl(SrcG15{A, N}, row_begin, row_end, text);
}
void encode_rows_g15(const double* A, const double* N, int64_t row_begin, int64_t row_end, std::string& out)
{
    encode_impl(SrcG15{A, N}, row_begin, row_end, out);
}

}  // namespace grm_text_gz
