// strtod.cuh - correctly rounded decimal text -> binary64, the `float(val[1])` of the reference
// (gpu_util_stats.py:48, :74) as device code.
//
// CPython's float() is David Gay's correctly rounded strtod; to be bit-identical this has to be
// correctly rounded too, for every input.  Three steps:
//   1. scan the token with the float() grammar of Python 2, the reference's interpreter (sign, digits
//      with '.', exponent, inf / infinity / nan in any case, blanks around; the underscores Python 3.6
//      added are refused, as Python 2 refuses them) into at most 19
//      significant digits w, a decimal exponent q and a "digits were dropped" flag;
//   2. Eisel-Lemire: w * 5^q from a 128-bit table of powers of five, one or two 64x64->128
//      multiplications; for w of at most 19 digits this is always decisive (Mushtak & Lemire,
//      "Fast number parsing without fallback", 2023);
//   3. only if digits were dropped AND w, w+1 round to different doubles: exact comparison of the
//      full digit string (768 digits + sticky) with the halfway point in multi-word integers.
//      Prometheus prints shortest round-trip values (<= 17 digits), so live answers never get here.
// Compiles for the device (nvcc) and, for the CPU unit test of the arithmetic only
// (tests/test_strtod_host.py), as plain C++ - libkgs itself has no host decode path.
#pragma once
#include <cstdint>

#ifdef __CUDACC__
#define KGS_HD __host__ __device__
#else
#define KGS_HD
#endif

namespace kgs {

#if defined(__CUDACC__)
__device__ const unsigned long long kPow5Dev[] = {
#include "pow5_table.inc"
};
#endif
#if !defined(__CUDACC__) || defined(KGS_STRTOD_HOST_TABLE)
static const unsigned long long kPow5Host[] = {
#include "pow5_table.inc"
};
#endif

constexpr int kPow5MinQ = -342, kPow5MaxQ = 308;

KGS_HD inline unsigned long long pow5_word(int idx) {
#ifdef __CUDA_ARCH__
    return kPow5Dev[idx];
#else
#if !defined(__CUDACC__) || defined(KGS_STRTOD_HOST_TABLE)
    return kPow5Host[idx];
#else
    return 0;
#endif
#endif
}

KGS_HD inline void mul64(unsigned long long a, unsigned long long b, unsigned long long& hi, unsigned long long& lo) {
#ifdef __CUDA_ARCH__
    lo = a * b;
    hi = __umul64hi(a, b);
#else
    const unsigned __int128 p = (unsigned __int128)a * b;
    lo = (unsigned long long)p;
    hi = (unsigned long long)(p >> 64);
#endif
}

KGS_HD inline int clz64(unsigned long long x) {
#ifdef __CUDA_ARCH__
    return __clzll((long long)x);
#else
    return __builtin_clzll(x);
#endif
}

// w * 10^q -> nearest binary64 (bits, sign not included); w != 0, w < 2^64.
KGS_HD inline unsigned long long eisel_lemire(unsigned long long w, long long q) {
    constexpr unsigned long long kInfBits = 0x7FF0000000000000ull;
    if (q < kPow5MinQ) return 0;
    if (q > kPow5MaxQ) return kInfBits;
    const int lz = clz64(w);
    w <<= lz;
    const int idx = 2 * (int)(q - kPow5MinQ);
    unsigned long long hi, lo;
    mul64(w, pow5_word(idx), hi, lo);
    if ((hi & 0x1FFull) == 0x1FFull) {   // the low 9 bits are all ones: the second table word can carry in
        unsigned long long hi2, lo2;
        mul64(w, pow5_word(idx + 1), hi2, lo2);
        lo += hi2;
        if (hi2 > lo) ++hi;
    }
    const int upperbit = (int)(hi >> 63);
    const int shift = upperbit + 64 - 52 - 3;
    unsigned long long mant = hi >> shift;
    long long power2 = ((217706 * q) >> 16) + 63 + upperbit - lz + 1023;
    if (power2 <= 0) {   // subnormal
        if (-power2 + 1 >= 64) return 0;
        mant >>= (-power2 + 1);
        mant += mant & 1;
        mant >>= 1;
        return mant;     // a carry into bit 52 is exactly the smallest normal number
    }
    if (lo <= 1 && q >= -4 && q <= 23 && (mant & 3) == 1 && (mant << shift) == hi) mant &= ~1ull;   // exact tie: to even
    mant += mant & 1;
    mant >>= 1;
    if (mant >= (2ull << 52)) {
        mant = 1ull << 52;
        ++power2;
    }
    mant &= ~(1ull << 52);
    if (power2 >= 0x7FF) return kInfBits;
    return ((unsigned long long)power2 << 52) | mant;
}

// ---- the common case first -------------------------------------------------------------------------
// "-?digits[.digits]" with at most 15 digits in all (what a utilisation or an epoch timestamp looks
// like): the digits are an integer below 2^53 and 10^k is exact, so ONE correctly rounded division
// gives the correctly rounded value (Clinger's fast path) - a few dozen instructions instead of the
// full grammar + Eisel-Lemire.  Returns the number of characters consumed (the caller checks what
// follows), 0 if the text at s does not have this form; s must be readable up to a non-digit byte.
KGS_HD inline uint32_t parse_short_decimal(const char* s, double* out) {
    uint32_t i = 0, nd = 0, nf = 0;
    const bool neg = s[0] == '-';
    if (neg) i = 1;
    unsigned long long w = 0;
    uint32_t c;
    while ((c = (uint32_t)(unsigned char)s[i] - (uint32_t)'0') <= 9u && nd <= 15u) {
        w = w * 10 + c;
        ++i;
        ++nd;
    }
    if (s[i] == '.') {
        ++i;
        while ((c = (uint32_t)(unsigned char)s[i] - (uint32_t)'0') <= 9u && nd + nf <= 15u) {
            w = w * 10 + c;
            ++i;
            ++nf;
        }
    }
    if (nd + nf == 0 || nd + nf > 15u) return 0;
    double v = (double)(long long)w;
    if (nf) {
        double p = 1.0;
        switch (nf) {   // exact powers of ten
            case 1: p = 1e1; break;   case 2: p = 1e2; break;   case 3: p = 1e3; break;   case 4: p = 1e4; break;
            case 5: p = 1e5; break;   case 6: p = 1e6; break;   case 7: p = 1e7; break;   case 8: p = 1e8; break;
            case 9: p = 1e9; break;   case 10: p = 1e10; break; case 11: p = 1e11; break; case 12: p = 1e12; break;
            case 13: p = 1e13; break; case 14: p = 1e14; break; default: p = 1e15; break;
        }
#ifdef __CUDA_ARCH__
        v = __ddiv_rn(v, p);
#else
        v = v / p;
#endif
    }
    *out = neg ? -v : v;
    return i;
}

// ---- exact tie-break for tokens with more than 19 significant digits --------------------------------
struct BigNat {
    static constexpr int kLimbs = 136;   // 4352 bits: 769 digits, or 2^54 * 5^1100, with room to shift
    uint32_t n = 0;
    bool overflow = false;
    uint32_t d[kLimbs];
    KGS_HD void mul_add(uint32_t m, uint32_t a) {   // *this = *this * m + a
        unsigned long long carry = a;
        for (uint32_t i = 0; i < n; ++i) {
            const unsigned long long t = (unsigned long long)d[i] * m + carry;
            d[i] = (uint32_t)t;
            carry = t >> 32;
        }
        if (carry) {
            if (n == (uint32_t)kLimbs) { overflow = true; return; }
            d[n++] = (uint32_t)carry;
        }
    }
    KGS_HD void mul_pow5(uint32_t e) {
        while (e >= 13) { mul_add(1220703125u, 0); e -= 13; }
        uint32_t m = 1;
        while (e--) m *= 5;
        if (m > 1) mul_add(m, 0);
    }
    KGS_HD void shl(uint32_t bits) {
        if (n == 0 || bits == 0) return;
        const uint32_t words = bits >> 5, b = bits & 31;
        if (n + words + 1 > (uint32_t)kLimbs) { overflow = true; return; }
        uint32_t top = 0;
        if (b) {
            for (uint32_t i = 0; i < n; ++i) {
                const uint32_t v = d[i];
                d[i] = (v << b) | top;
                top = v >> (32 - b);
            }
            if (top) d[n++] = top;
        }
        if (words) {
            for (uint32_t i = n; i-- > 0;) d[i + words] = d[i];
            for (uint32_t i = 0; i < words; ++i) d[i] = 0;
            n += words;
        }
    }
};

KGS_HD inline int big_cmp(const BigNat& a, const BigNat& b) {
    if (a.n != b.n) return a.n < b.n ? -1 : 1;
    for (uint32_t i = a.n; i-- > 0;)
        if (a.d[i] != b.d[i]) return a.d[i] < b.d[i] ? -1 : 1;
    return 0;
}

// The token's mantissa characters are s[m0, m1) (digits, one '.', underscores), its explicit exponent e.
// lower = the double (bits) that the truncated 19-digit prefix rounds to; the answer is lower or
// lower + 1 ulp, decided against the halfway point between them.  Returns false if a number did not fit.
#ifdef __CUDACC__
__noinline__
#endif
KGS_HD inline bool exact_round(const char* s, uint32_t m0, uint32_t m1, long long e, unsigned long long lower,
                               unsigned long long* out) {
    constexpr uint32_t kKeep = 768;
    BigNat A, B;
    uint32_t kept = 0;
    bool point = false, sticky = false;
    long long e10 = e;
    for (uint32_t i = m0; i < m1; ++i) {
        const char c = s[i];
        if (c == '.') { point = true; continue; }
        if (c < '0' || c > '9') continue;
        const uint32_t dg = (uint32_t)(c - '0');
        if (kept == 0 && dg == 0) {
            if (point) --e10;
            continue;
        }
        if (kept < kKeep) {
            A.mul_add(10, dg);
            ++kept;
            if (point) --e10;
        } else {
            sticky |= dg != 0;
            if (!point) ++e10;
        }
    }
    if (sticky) {
        A.mul_add(10, 1);
        --e10;
    }
    const unsigned long long expf = lower >> 52, frac = lower & ((1ull << 52) - 1);
    const unsigned long long mant = expf ? (frac | (1ull << 52)) : frac;
    const long long e2 = (expf ? (long long)expf - 1075 : -1074) - 1;   // exponent of half an ulp
    const unsigned long long H = 2 * mant + 1;
    B.n = 2;
    B.d[0] = (uint32_t)H;
    B.d[1] = (uint32_t)(H >> 32);
    if (B.d[1] == 0) B.n = 1;
    if (e10 > 4000 || e10 < -4000) return false;
    if (e10 >= 0) A.mul_pow5((uint32_t)e10);
    else B.mul_pow5((uint32_t)(-e10));
    const long long sh = e10 - e2;
    if (sh > 8000 || sh < -8000) return false;
    if (sh > 0) A.shl((uint32_t)sh);
    else if (sh < 0) B.shl((uint32_t)(-sh));
    if (A.overflow || B.overflow) return false;
    const int c = big_cmp(A, B);
    const bool up = c > 0 || (c == 0 && (mant & 1));
    *out = lower + (up ? 1 : 0);
    return true;
}

KGS_HD inline bool is_blank(char c) { return c == ' ' || (c >= '\t' && c <= '\r'); }
KGS_HD inline bool is_digit(char c) { return c >= '0' && c <= '9'; }
KGS_HD inline char lower_ascii(char c) { return (c >= 'A' && c <= 'Z') ? (char)(c + 32) : c; }

KGS_HD inline bool word_is(const char* s, uint32_t a, uint32_t b, const char* w, uint32_t wl) {
    if (b - a != wl) return false;
    for (uint32_t i = 0; i < wl; ++i)
        if (lower_ascii(s[a + i]) != w[i]) return false;
    return true;
}

// A BARE number of a JSON text as json.loads reads it (the timestamps of an answer, and a value that is
// not a string): -?(0|[1-9][0-9]*)(.[0-9]+)?([eE][+-]?[0-9]+)? or one of NaN, Infinity, -Infinity,
// which Python's json accepts beyond the standard.  What float() would also take - '+1', '.5', '5.',
// '007', 'nan' - makes the reference's .json() raise, so it is refused here too.
KGS_HD inline bool json_number_ok(const char* s, uint32_t len) {
    uint32_t i = 0;
    if (len == 3 && s[0] == 'N' && s[1] == 'a' && s[2] == 'N') return true;
    if (i < len && s[i] == '-') ++i;
    if (len - i == 8) {
        const char* w = "Infinity";
        bool same = true;
        for (uint32_t k = 0; k < 8; ++k) same = same && s[i + k] == w[k];
        if (same) return true;
    }
    if (i >= len || s[i] < '0' || s[i] > '9') return false;
    if (s[i] == '0') {
        ++i;
    } else {
        while (i < len && s[i] >= '0' && s[i] <= '9') ++i;
    }
    if (i < len && s[i] == '.') {
        ++i;
        if (i >= len || s[i] < '0' || s[i] > '9') return false;
        while (i < len && s[i] >= '0' && s[i] <= '9') ++i;
    }
    if (i < len && (s[i] == 'e' || s[i] == 'E')) {
        ++i;
        if (i < len && (s[i] == '+' || s[i] == '-')) ++i;
        if (i >= len || s[i] < '0' || s[i] > '9') return false;
        while (i < len && s[i] >= '0' && s[i] <= '9') ++i;
    }
    return i == len;
}

// float(str) of Python 2 for the ASCII token s[0, len).  false = float() would raise ValueError (or a
// multi-word number did not fit, which no finite input reaches).
KGS_HD inline bool parse_float(const char* s, uint32_t len, double* out) {
    uint32_t i = 0, end = len;
    while (i < end && is_blank(s[i])) ++i;
    while (end > i && is_blank(s[end - 1])) --end;
    bool neg = false;
    if (i < end && (s[i] == '+' || s[i] == '-')) {
        neg = s[i] == '-';
        ++i;
    }
    unsigned long long bits;
    if (word_is(s, i, end, "inf", 3) || word_is(s, i, end, "infinity", 8)) {
        bits = 0x7FF0000000000000ull;
    } else if (word_is(s, i, end, "nan", 3)) {
        bits = 0x7FF8000000000000ull;
    } else {
        const uint32_t m0 = i;
        unsigned long long w = 0;
        uint32_t nd = 0;
        long long e10 = 0;
        bool trunc = false, seen = false, point = false;
        for (; i < end; ++i) {
            const char c = s[i];
            if (is_digit(c)) {
                seen = true;
                const uint32_t dg = (uint32_t)(c - '0');
                if (nd == 0 && dg == 0) {
                    if (point) --e10;
                } else if (nd < 19) {
                    w = w * 10 + dg;
                    ++nd;
                    if (point) --e10;
                } else {
                    trunc |= dg != 0;
                    if (!point) ++e10;
                }
            } else if (c == '.' && !point) {
                point = true;
            } else {
                break;
            }
        }
        if (!seen) return false;
        const uint32_t m1 = i;
        long long e = 0;
        if (i < end && (s[i] == 'e' || s[i] == 'E')) {
            ++i;
            bool eneg = false;
            if (i < end && (s[i] == '+' || s[i] == '-')) {
                eneg = s[i] == '-';
                ++i;
            }
            bool eseen = false;
            for (; i < end; ++i) {
                const char c = s[i];
                if (is_digit(c)) {
                    eseen = true;
                    if (e < 100000000) e = e * 10 + (c - '0');
                } else {
                    break;
                }
            }
            if (!eseen) return false;
            if (eneg) e = -e;
        }
        if (i != end) return false;
        if (w == 0) {
            bits = 0;
        } else {
            const long long q = e10 + e;
            bits = eisel_lemire(w, q);
            if (trunc) {
                const unsigned long long up = eisel_lemire(w + 1, q);
                if (up != bits && !exact_round(s, m0, m1, e, bits, &bits)) return false;
            }
        }
    }
    if (neg) bits |= 0x8000000000000000ull;
#ifdef __CUDA_ARCH__
    *out = __longlong_as_double((long long)bits);
#else
    union { unsigned long long u; double d; } cv;
    cv.u = bits;
    *out = cv.d;
#endif
    return true;
}

}  // namespace kgs
