// Test harness (CPU, tests only): the chunk lexer of promjson_bits.cuh as plain C++.
// stdin: one case per line "<128 hex digits = 64 bytes> <first_escaped> <in_string> <depth> <pair_depth>".
// stdout per case: "esc inside open close flip d0 d1 | kind:index ..." (masks in hex).
#include <cstdio>
#include <cstring>
#include <iostream>
#include <string>
#include "../../kube_gpu_stats_b200/csrc/promjson_bits.cuh"

int main() {
    std::string hex;
    int fe, ins, depth, pd;
    while (std::cin >> hex >> fe >> ins >> depth >> pd) {
        unsigned char b[64];
        for (int i = 0; i < 64; ++i) b[i] = (unsigned char)std::stoi(hex.substr(2 * i, 2), nullptr, 16);
        uint32_t w[16];
        std::memcpy(w, b, 64);
        const kgs::MxMasks m = kgs::mx_masks(w);
        const unsigned long long esc = kgs::mx_escaped(m.bs, fe != 0);
        const kgs::MxLex x = kgs::mx_lex(m, fe != 0);
        int d0, d1;
        kgs::mx_depth_change(x, d0, d1);
        std::printf("%016llx %016llx %016llx %016llx %u %d %d |", esc, x.inside, x.open, x.close, x.flip, d0, d1);
        kgs::mx_brackets(x, (uint32_t)ins, depth, pd, [&](int kind, uint32_t i) { std::printf(" %d:%u", kind, i); });
        std::printf("\n");
    }
    return 0;
}
