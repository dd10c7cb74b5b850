// Test harness (CPU, tests only): runs the decimal->binary64 routine of strtod.cuh as plain C++ on
// every line of stdin and prints "<ok> <bits as hex>" per line.  tests/test_strtod_host.py compares
// the output with Python's float().  libkgs never runs this code on the host.
#include <cstdio>
#include <cstring>
#include <string>
#include <iostream>
#include "../../kube_gpu_stats_b200/csrc/strtod.cuh"

int main() {
    std::string line;
    while (std::getline(std::cin, line)) {
        double v = 0;
        const bool ok = kgs::parse_float(line.data(), (uint32_t)line.size(), &v);
        unsigned long long bits;
        std::memcpy(&bits, &v, 8);
        // the short-decimal fast path on the same token (zero-terminated by std::string)
        double f = 0;
        const uint32_t used = kgs::parse_short_decimal(line.c_str(), &f);
        unsigned long long fbits;
        std::memcpy(&fbits, &f, 8);
        std::printf("%d %016llx %u %016llx\n", ok ? 1 : 0, ok ? bits : 0ull, used, used ? fbits : 0ull);
    }
    return 0;
}
