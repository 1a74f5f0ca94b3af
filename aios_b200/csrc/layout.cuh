// layout.cuh — how quantised weight matrices live in HBM on the B200.
//
// GGUF stores a matrix W[N rows][K cols] as N rows of K/256 (K/32 for Q8_0) self-contained
// blocks (gguf-py quants.py:396-572; sizes constants.py:4231-4246).  144-, 210- and 34-byte
// blocks cannot be read with aligned 128-bit loads, so at load time each matrix is re-tiled
// ON THE GPU into separate, 16-byte-aligned "planes".  The bytes are only permuted — no value
// is changed, so dequantised weights stay bit-exact, and the byte count (the algorithmic
// bytes of the roofline) is identical to the GGUF tensor's.
//
// The unit of work of one lane is a "unit" of 64 weights (32 for Q8_0):
//
//  Q4_K  unit (b,p), p=0..3   = sub-blocks 2p (low nibbles) and 2p+1 (high nibbles) of block b
//        plane QSa[N][U][16] = qs[32p    .. 32p+16)      U = K/64
//        plane QSb[N][U][16] = qs[32p+16 .. 32p+32)
//        plane META[N][nb][16] = 4 words, one per unit j of the block (nb = K/256):
//                                sc(2j) | sc(2j+1)<<6 | m(2j)<<12 | m(2j+1)<<18 | byte j of (d | dmin<<16) << 24
//                                (the twelve packed 6-bit scale bytes, re-ordered per unit; still 16 bytes)
//        activations used: a_k = q8[64p+16k .. +16), k=0..3  (lo: a0,a1  hi: a2,a3)
//
//  Q6_K  unit (b,h,t), h,t in {0,1}; unit index = 4b + 2h + t
//        plane QLa[N][U][16] = ql[64h+16t    .. +16)
//        plane QLb[N][U][16] = ql[64h+32+16t .. +16)
//        plane QH [N][U][16] = qh[32h+16t    .. +16)
//        plane SC [N][U][4]  = scales[8h+t+{0,2,4,6}]
//        plane D  [N][nb][2] = d
//        activations used: a_k = q8[128h+16t+32k .. +16), k=0..3
//
//  Q8_0  unit = one block of 32;  U8 = K/32
//        plane QSa[N][U8][16] = qs[0..16), plane QSb[N][U8][16] = qs[16..32), plane D[N][U8][2]
//
//  Q5_K  as Q4_K plus plane QH5[N][U][8]: bit l of the first u32 = 5th bit of element l of
//        sub-block 2p, second u32 likewise for sub-block 2p+1.
#pragma once
#include <stdint.h>

namespace aios {

enum : int { T_F32 = 0, T_F16 = 1, T_Q8_0 = 8, T_Q4_K = 12, T_Q5_K = 13, T_Q6_K = 14 };

struct WPlanes {
    const uint8_t *p[5];
    int32_t type;
    int32_t n_rows;  // N
    int32_t k;       // K
    int32_t pad;
};

inline int type_block_elems(int t) { return (t == T_F32 || t == T_F16) ? 1 : (t == T_Q8_0 ? 32 : 256); }
inline int type_block_bytes(int t) {
    switch (t) {
        case T_F32: return 4;
        case T_F16: return 2;
        case T_Q8_0: return 34;
        case T_Q4_K: return 144;
        case T_Q5_K: return 176;
        case T_Q6_K: return 210;
    }
    return 0;
}
inline size_t tensor_bytes(int t, int64_t n_rows, int64_t k) {
    return (size_t)n_rows * (size_t)(k / type_block_elems(t)) * (size_t)type_block_bytes(t);
}
// bytes of plane i for a [n_rows][k] matrix (0 if the plane does not exist)
inline size_t plane_bytes(int t, int i, int64_t n_rows, int64_t k) {
    const size_t U = (size_t)k / 64, nb = (size_t)k / 256, U8 = (size_t)k / 32, N = (size_t)n_rows;
    switch (t) {
        case T_Q4_K: { const size_t s[5] = {N * U * 16, N * U * 16, N * nb * 16, 0, 0}; return s[i]; }
        case T_Q5_K: { const size_t s[5] = {N * U * 16, N * U * 16, N * nb * 16, N * U * 8, 0}; return s[i]; }
        case T_Q6_K: { const size_t s[5] = {N * U * 16, N * U * 16, N * U * 16, N * U * 4, N * nb * 2}; return s[i]; }
        case T_Q8_0: { const size_t s[5] = {N * U8 * 16, N * U8 * 16, N * U8 * 2, 0, 0}; return s[i]; }
    }
    return 0;
}

}  // namespace aios
