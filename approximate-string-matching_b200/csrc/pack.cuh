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

// 4 characters -> plane nibbles, assuming they are all in {A,C,G,T}; `bad` collects the proof that they are.
// ASCII A=0x41 C=0x43 G=0x47 T=0x54: bit 2 is the G/T plane and bit1^bit2 the C/T plane.  The character is then
// rebuilt from its two bits (0x41 + 2*ct + 6*gt + 11*(ct&gt) gives 0x41/0x43/0x47/0x54; bytes never carry) and
// compared with the input, so any other byte leaves a non-zero mark in `bad`.
GASMA_HD void pack_word_acgt(uint32_t x, uint32_t& n0, uint32_t& n1, uint32_t& bad) {
    const uint32_t gt = (x >> 2) & 0x01010101u;
    const uint32_t ct = ((x >> 1) ^ (x >> 2)) & 0x01010101u;
    const uint32_t rebuilt = 0x41414141u + 2u * ct + 6u * gt + 11u * (ct & gt);
    bad |= rebuilt ^ x;
    // bit 8j -> bit 28+j: 0x10204080 = 2^28+2^21+2^14+2^7; the cross terms fall below bit 28 on distinct bits or overflow
    n0 = (ct * 0x10204080u) >> 28;
    n1 = (gt * 0x10204080u) >> 28;
}

// 16 characters -> 16 + 16 plane bits
GASMA_HD void pack_chunk(uint32_t w0, uint32_t w1, uint32_t w2, uint32_t w3, uint32_t& p0, uint32_t& p1) {
    uint32_t a0, a1, b0, b1, c0, c1, d0, d1, bad = 0u;
    pack_word_acgt(w0, a0, a1, bad);
    pack_word_acgt(w1, b0, b1, bad);
    pack_word_acgt(w2, c0, c1, bad);
    pack_word_acgt(w3, d0, d1, bad);
    if (bad) {   // some byte is not A/C/G/T (N, lower case, NUL ...): exact compares, such bytes pack as 00
        pack_word(w0, a0, a1);
        pack_word(w1, b0, b1);
        pack_word(w2, c0, c1);
        pack_word(w3, d0, d1);
    }
    p0 = a0 | (b0 << 4) | (c0 << 8) | (d0 << 12);
    p1 = a1 | (b1 << 4) | (c1 << 8) | (d1 << 12);
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
