// hf_divmagic.h -- exact division of a 31-bit numerator by an invariant divisor without a divide instruction
// (prim_index -> (cx, cy) in every kernel that starts from a stored hit: one IMAD.HI + shift instead of the ~20-instruction
// 32-bit division sequence). Granlund & Montgomery, "Division by invariant integers using multiplication", theorem 4.2 with N = 31:
// for l = ceil(log2 d) and m = ceil(2^(31+l) / d) one has 2^(31+l) <= m d <= 2^(31+l) + 2^l, hence floor(n / d) = (n m) >> (31 + l)
// for every 0 <= n < 2^31, and m < 2^32. Compiles as CUDA and as plain C++ (tests/trav_host.cpp checks it against `/`).
#pragma once
#include <stdint.h>

struct HfDivMagic { uint32_t m; int32_t s; };      // s < 0: divisor 1

static inline HfDivMagic hf_divmagic(uint32_t d) {
    HfDivMagic r;
    if (d <= 1u) { r.m = 0u; r.s = -1; return r; }
    int l = 0;
    while ((1ull << l) < (unsigned long long) d) ++l;
    const unsigned long long num = 1ull << (31 + l);
    r.m = (uint32_t) ((num + d - 1ull) / d);
    r.s = l - 1;                                    // (n m) >> (31 + l) == umulhi(n, m) >> (l - 1)
    return r;
}

#ifdef __CUDACC__
__device__ __forceinline__ uint32_t hf_divide(uint32_t n, uint32_t m, int32_t s) { return s < 0 ? n : (__umulhi(n, m) >> s); }
#else
static inline uint32_t hf_divide(uint32_t n, uint32_t m, int32_t s) {
    return s < 0 ? n : (uint32_t) ((((unsigned long long) n * m) >> 32) >> s);
}
#endif
