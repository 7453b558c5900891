// Warp-cooperative writers: per-game bit strings staged in shared memory ->
// fp32 tensors in HBM with coalesced 128-bit streaming stores.
//
// One lane owns one game for the integer rules; when the 32 games of a warp have
// produced their observation / legal bits, the whole warp streams the tile out:
// 32 games x NB floats are contiguous in the (N, ...) output, so lane q writes
// float4 number q of the tile and the warp covers 512 contiguous bytes per store.
#pragma once
#include "common.cuh"

// s_bits: [32][RW] words, game-major, bit j of game g = s_bits[g*RW + j/32] >> (j%32).
// out_tile: pointer to the first float of the tile (16B aligned), ng = valid games.
// bit set -> v1, clear -> v0.
template <int NB, int RW>
__device__ __forceinline__ void emit_tile_f32(const uint32_t* __restrict__ s_bits, float* __restrict__ out_tile,
                                              int ng, int lane, float v0, float v1) {
    const int total = ng * NB;
    const int n4 = total >> 2;
    float4* o4 = reinterpret_cast<float4*>(out_tile);
    for (int q = lane; q < n4; q += 32) {
        const int e0 = q << 2;
        int g = e0 / NB;
        int a = e0 - g * NB;
        float f[4];
        if ((NB & 3) == 0) {
            // the four bits never straddle two games; they may straddle two words
            const uint32_t* w = s_bits + g * RW + (a >> 5);
            uint32_t lo = w[0];
            uint32_t nib = (lo >> (a & 31)) & 0xF;  // NB%4==0 and a%4==0 -> nibble aligned
#pragma unroll
            for (int j = 0; j < 4; ++j) f[j] = ((nib >> j) & 1) ? v1 : v0;
        } else {
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                int aj = a + j, gj = g;
                if (aj >= NB) { aj -= NB; gj += 1; }
                uint32_t b = (s_bits[gj * RW + (aj >> 5)] >> (aj & 31)) & 1;
                f[j] = b ? v1 : v0;
            }
        }
        st_stream_f4(o4 + q, make_float4(f[0], f[1], f[2], f[3]));
    }
    // tail (< 4 floats) of a partial tile
    for (int e = (n4 << 2) + lane; e < total; e += 32) {
        int g = e / NB, a = e - g * NB;
        out_tile[e] = ((s_bits[g * RW + (a >> 5)] >> (a & 31)) & 1) ? v1 : v0;
    }
}
