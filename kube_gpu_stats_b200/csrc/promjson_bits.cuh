// promjson_bits.cuh - the lexical work of promjson.cuh on 64-byte chunks as bit arithmetic.
//
// A byte loop costs ~26 instructions per byte of answer text (ncu: k_mx_classify was bound by the
// integer pipe).  Here a chunk's 16 words are turned into five 64-bit masks with four-bytes-at-a-time
// arithmetic (exact zero-byte test per class, then a multiply gathers the four flag bits), and
// everything after that is a few dozen 64-bit operations:
//   bs  backslashes            qt  '"'            op  '[' or '{'          cl  ']' or '}'
//   b5  bytes with bit 0x20 set (tells '{' from '[' and '}' from ']')
// Escapes: a character behind an odd run of backslashes (the carry-less adder trick of simdjson's
// stage 1, restated); strings: prefix XOR of the unescaped quotes.
// Plain C++ as well as device code: tests/test_promjson_bits_host.py checks every function against a
// byte-by-byte loop on the CPU.  libkgs only ever runs it on the GPU.
#pragma once
#include <cstdint>

#ifdef __CUDACC__
#define KGS_HD __host__ __device__
#else
#ifndef KGS_HD
#define KGS_HD
#endif
#endif

namespace kgs {

struct MxMasks {
    unsigned long long bs, qt, op, cl, b5;
};

KGS_HD inline int mx_popc(unsigned long long x) {
#ifdef __CUDA_ARCH__
    return __popcll(x);
#else
    return __builtin_popcountll(x);
#endif
}

KGS_HD inline int mx_ctz(unsigned long long x) {   // x != 0
#ifdef __CUDA_ARCH__
    return __ffsll((long long)x) - 1;
#else
    return __builtin_ctzll(x);
#endif
}

// 0x80 in every byte of y that is zero (exact: no borrow crosses a byte)
KGS_HD inline uint32_t mx_zero_bytes(uint32_t y) {
    const uint32_t t = ((y & 0x7F7F7F7Fu) + 0x7F7F7F7Fu) | y;
    return ~t & 0x80808080u;
}

// bit k of the result = bit 8k of f (f has at most bits 0, 8, 16, 24 set)
KGS_HD inline uint32_t mx_gather4(uint32_t f) { return (f * 0x10204080u) >> 28; }

KGS_HD inline MxMasks mx_masks(const uint32_t (&w)[16]) {
    uint32_t m[5][2] = {{0, 0}, {0, 0}, {0, 0}, {0, 0}, {0, 0}};
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
    for (int i = 0; i < 16; ++i) {
        const uint32_t x = w[i], z = x | 0x20202020u;
        const uint32_t f0 = mx_zero_bytes(x ^ 0x5C5C5C5Cu) >> 7;
        const uint32_t f1 = mx_zero_bytes(x ^ 0x22222222u) >> 7;
        const uint32_t f2 = mx_zero_bytes(z ^ 0x7B7B7B7Bu) >> 7;
        const uint32_t f3 = mx_zero_bytes(z ^ 0x7D7D7D7Du) >> 7;
        const uint32_t f4 = (x >> 5) & 0x01010101u;
        const int half = i >> 3, sh = 4 * (i & 7);
        m[0][half] |= mx_gather4(f0) << sh;
        m[1][half] |= mx_gather4(f1) << sh;
        m[2][half] |= mx_gather4(f2) << sh;
        m[3][half] |= mx_gather4(f3) << sh;
        m[4][half] |= mx_gather4(f4) << sh;
    }
    MxMasks r;
    r.bs = ((unsigned long long)m[0][1] << 32) | m[0][0];
    r.qt = ((unsigned long long)m[1][1] << 32) | m[1][0];
    r.op = ((unsigned long long)m[2][1] << 32) | m[2][0];
    r.cl = ((unsigned long long)m[3][1] << 32) | m[3][0];
    r.b5 = ((unsigned long long)m[4][1] << 32) | m[4][0];
    return r;
}

// Mask of the characters that are escaped (preceded by an odd run of backslashes).  `first_escaped`:
// the chunk's first byte is escaped by the text before it.  Backslashes that are themselves escaped
// do not start an escape.
KGS_HD inline unsigned long long mx_escaped(unsigned long long bs, bool first_escaped) {
    const unsigned long long even = 0x5555555555555555ull, carry = first_escaped ? 1ull : 0ull;
    bs &= ~carry;
    const unsigned long long follows = (bs << 1) | carry;
    const unsigned long long odd_starts = bs & ~even & ~follows;
    const unsigned long long seq_even = odd_starts + bs;          // runs that start on an odd bit flip their phase
    const unsigned long long invert = seq_even << 1;
    return (even ^ invert) & follows;
}

KGS_HD inline unsigned long long mx_prefix_xor(unsigned long long x) {
    x ^= x << 1;
    x ^= x << 2;
    x ^= x << 4;
    x ^= x << 8;
    x ^= x << 16;
    x ^= x << 32;
    return x;
}

struct MxLex {                 // a chunk after lexing
    unsigned long long inside; // bit i: byte i lies inside a string IF the chunk starts outside one
                               // (an opening quote counts as inside, its closing quote as outside)
    unsigned long long open, close, b5;   // unescaped brackets, string or not
    uint32_t flip;             // quote parity of the chunk
};

KGS_HD inline MxLex mx_lex(const MxMasks& m, bool first_escaped) {
    const unsigned long long esc = mx_escaped(m.bs, first_escaped);
    const unsigned long long q = m.qt & ~esc;
    MxLex x;
    x.inside = mx_prefix_xor(q);
    x.open = m.op & ~esc;
    x.close = m.cl & ~esc;
    x.b5 = m.b5;
    x.flip = (uint32_t)(mx_popc(q) & 1);
    return x;
}

// bracket-depth change of the chunk if it starts outside (d0) / inside (d1) a string
KGS_HD inline void mx_depth_change(const MxLex& x, int& d0, int& d1) {
    d0 = mx_popc(x.open & ~x.inside) - mx_popc(x.close & ~x.inside);
    d1 = mx_popc(x.open & x.inside) - mx_popc(x.close & x.inside);
}

// Walk the structural brackets of a chunk whose first byte is at (in_string, depth); f(kind, index):
// kind 0 = '[' that opens at pair_depth (a sample), 1 = '{' opening at depth 3 (a series),
// 2 = '[' opening at depth 4 (values array, matrix only), 3 = ']' closing to depth 4 (matrix only),
// 4 = '}' closing to depth 3.
template <class F>
KGS_HD inline void mx_brackets(const MxLex& x, uint32_t in_string, int depth, int pair_depth, F&& f) {
    const unsigned long long outside = in_string ? x.inside : ~x.inside;
    unsigned long long s = (x.open | x.close) & outside;
    while (s) {
        const int i = mx_ctz(s);
        s &= s - 1;
        const bool curly = (x.b5 >> i) & 1ull;
        if ((x.open >> i) & 1ull) {
            if (!curly) {
                if (depth == pair_depth) f(0, (uint32_t)i);
                else if (depth == 4) f(2, (uint32_t)i);
            } else if (depth == 3) {
                f(1, (uint32_t)i);
            }
            ++depth;
        } else {
            --depth;
            if (!curly) {
                if (depth == 4 && pair_depth == 5) f(3, (uint32_t)i);
            } else if (depth == 3) {
                f(4, (uint32_t)i);
            }
        }
    }
}

}  // namespace kgs
