// pack.cuh -- K1: ASCII DNA -> two 1-bit planes.
//
// Replaces sse3_convert2bit1 (GASMA/bit_convert.cpp:248-369): plane0 bit i = (c_i in {C,T}),
// plane1 bit i = (c_i in {G,T}), every other byte (N, lower case, NUL ...) -> 00, LSB-first.
//
// B200 shape of the work: a CTA owns a TILE of consecutive strings, i.e. one contiguous byte range of
// the CSR buffer.  Each thread takes 16-byte chunks of that range (one coalesced uint4 load, no matter
// where string boundaries fall), turns 16 characters into 16+16 plane bits with SWAR byte compares and
// a multiply-gather, and drops them into a shared-memory bit stream (2 bits per base: 4x smaller than
// staging ASCII).  A string's planes are then a funnel-shifted window of that stream.
#pragma once
#include "common.cuh"

namespace gasma {

// bit 7 of every byte of the result is set iff that byte of x equals the byte replicated in c4 (exact,
// no inter-byte borrow: every byte of (y|0x80) is >= 0x80 before 1 is subtracted).
GASMA_HD uint32_t eq_bytes(uint32_t x, uint32_t c4) {
    uint32_t y = x ^ c4;
    uint32_t t = (y | 0x80808080u) - 0x01010101u;
    return ~(t | y) & 0x80808080u;
}

// flags at bit 7 of each byte -> 4-bit value (byte p -> bit p).  0x00204081 = 2^0+2^7+2^14+2^21 moves
// bit 8p (after >>7) to bit 21+p; all partial products land on distinct bits, so no carries.
GASMA_HD uint32_t gather4(uint32_t flags7) { return (((flags7 >> 7) * 0x00204081u) >> 21) & 0xFu; }

// 4 characters -> plane nibbles
GASMA_HD void pack_word(uint32_t x, uint32_t& n0, uint32_t& n1) {
    uint32_t c = eq_bytes(x, 0x43434343u);  // 'C'
    uint32_t g = eq_bytes(x, 0x47474747u);  // 'G'
    uint32_t t = eq_bytes(x, 0x54545454u);  // 'T'
    n0 = gather4(c | t);
    n1 = gather4(g | t);
}

// 4 characters -> the top nibbles of two products, assuming they are all in {A,C,G,T}; `bad` collects the proof that
// they are.  ASCII A=0x41 C=0x43 G=0x47 T=0x54: bit 2 is the G/T plane, bit 1 (C or G) XOR bit 2 the C/T plane.
// The masked bits stay where they are and ONE multiply per plane moves bit 8j+2 (8j+1) of byte j to bit 28+j:
// 0x04081020 = 2^26+2^19+2^12+2^5 (0x08102040 = twice that); the 16 partial products land on 16 distinct bits, so no
// carry reaches bits 28..31.  The caller appends the nibbles with funnel shifts and forms C/T = bit1 ^ bit2 once per
// chunk.  Validation: the character is rebuilt from its two bits -- 0x41 + (bits 1,2 in place) + 15 where the pattern
// is "bit 2 without bit 1" (T = 0x41 + 4 + 15) -- and compared with the input: any other byte leaves a mark in `bad`.
GASMA_HD void pack_word_acgt(uint32_t x, uint32_t& prod_b1, uint32_t& prod_b2, uint32_t& bad) {
    const uint32_t h = x & 0x02020202u, g = x & 0x04040404u;
    const uint32_t t = (g >> 2) & ~(h >> 1);                       // 1 per byte that looks like 'T'
    const uint32_t rebuilt = 0x41414141u + (h | g) + 15u * t;      // bytes never carry: at most 0x41 + 4 + 15
    bad |= rebuilt ^ x;
    prod_b1 = h * 0x08102040u;
    prod_b2 = g * 0x04081020u;
}

// (acc << 4) | top nibble of prod
GASMA_HD uint32_t append_nibble(uint32_t acc, uint32_t prod) {
#if defined(__CUDA_ARCH__)
    return __funnelshift_l(prod, acc, 4);
#else
    return (acc << 4) | (prod >> 28);
#endif
}

// 16 characters -> 16 + 16 plane bits
GASMA_HD void pack_chunk(uint32_t w0, uint32_t w1, uint32_t w2, uint32_t w3, uint32_t& p0, uint32_t& p1) {
    uint32_t a1, a2, b1, b2, c1, c2, d1, d2, bad = 0u;
    pack_word_acgt(w0, a1, a2, bad);
    pack_word_acgt(w1, b1, b2, bad);
    pack_word_acgt(w2, c1, c2, bad);
    pack_word_acgt(w3, d1, d2, bad);
    if (bad) {   // some byte is not A/C/G/T (N, lower case, NUL ...): exact compares, such bytes pack as 00
        uint32_t a0, b0, c0, d0;
        pack_word(w0, a0, a1);
        pack_word(w1, b0, b1);
        pack_word(w2, c0, c1);
        pack_word(w3, d0, d1);
        p0 = a0 | (b0 << 4) | (c0 << 8) | (d0 << 12);
        p1 = a1 | (b1 << 4) | (c1 << 8) | (d1 << 12);
        return;
    }
    const uint32_t bit1 = append_nibble(append_nibble(append_nibble(d1 >> 28, c1), b1), a1);   // C or G
    p1 = append_nibble(append_nibble(append_nibble(d2 >> 28, c2), b2), a2);                     // G or T
    p0 = bit1 ^ p1;                                                                           // C or T
}

// Window of W 64-bit words starting at bit `bit` of a 32-bit-word stream, bits >= len cleared
// (canonical 'A' tail: SURVEY.md section 8c).  The stream must be readable for 2W+1 words from bit>>5.
template <int W>
GASMA_HD void extract_plane(const uint32_t* stream32, int bit, int len, uint64_t out[W]) {
    const int idx = bit >> 5, sh = bit & 31;
    uint32_t w[2 * W + 1];
#pragma unroll
    for (int j = 0; j <= 2 * W; ++j) w[j] = stream32[idx + j];
#pragma unroll
    for (int j = 0; j < W; ++j) {
        uint64_t lo = ((uint64_t)w[2 * j + 1] << 32) | w[2 * j];
        uint64_t v = sh ? ((lo >> sh) | ((uint64_t)w[2 * j + 2] << (64 - sh))) : lo;
        uint64_t keep = ~ones_from(len, j);  // bits [64j, len)
        out[j] = v & keep;
    }
}

}  // namespace gasma
