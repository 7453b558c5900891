// Warp-cooperative writers: per-row bit strings -> fp32 tensors in HBM with
// coalesced 128-bit streaming stores.
//
// One lane owns one game (or one batch cell) for the integer rules.  Every output of
// the hot path (observation planes, scalar one-hots, action_mask, the RandomModel
// policy) is a two-valued fp32 image of a bit string, and the 32 rows of a warp are
// contiguous in the (N, ...) output.  So the warp first packs its rows into ONE flat
// bit string in shared memory (row r at bit offset r*NB), and then streams the tile
// out: float4 number q of the tile is the image of the nibble at flat bit 4q, looked
// up in a 16-entry float4 table.  That is ~6 instructions per 16 bytes stored — no
// per-element division, no straddling cases — which keeps the kernels HBM-bound
// instead of issue-bound (profiles/r1_rollout_*.md).
#pragma once
#include "common.cuh"

// 16 expansions of a nibble: bit j set -> v1, clear -> v0.
__device__ __forceinline__ void nibble_table_init(float4* tbl, float v0, float v1, int tid) {
    if (tid < 16) tbl[tid] = make_float4((tid & 1) ? v1 : v0, (tid & 2) ? v1 : v0, (tid & 4) ? v1 : v0, (tid & 8) ? v1 : v0);
}

constexpr int flat_words(int rows, int nb) { return (rows * nb + 31) / 32 + 2; }

// Zero the flat words that are shared between neighbouring rows (the first word of
// every row and the word after the last row); the words in between are owned by one
// row and are written with plain stores by flat_put.  Call for every row, then
// __syncwarp(), then flat_put.
template <int NB>
__device__ __forceinline__ void flat_clear_row(uint32_t* flat, int row, bool is_last_row) {
    flat[(row * NB) >> 5] = 0;
    if (is_last_row) {
        flat[((row + 1) * NB) >> 5] = 0;
        flat[(((row + 1) * NB) >> 5) + 1] = 0;
    }
}

// OR row `row` (NB bits in RW words, bits >= NB must be clear) into the flat string.
template <int NB, int RW>
__device__ __forceinline__ void flat_put(uint32_t* flat, int row, const uint32_t (&bits)[RW]) {
    const int pos = row * NB;
    const int w0 = pos >> 5, sh = pos & 31;
    const int lo_excl = (pos + 31) >> 5;            // first word fully inside the row
    const int hi_excl = ((pos + NB) >> 5);          // one past the last word fully inside the row
#pragma unroll
    for (int k = 0; k <= RW; ++k) {
        const uint32_t cur = k < RW ? bits[k] : 0u;
        const uint32_t prev = k > 0 ? bits[k - 1] : 0u;
        const uint32_t v = __funnelshift_l(prev, cur, sh);  // (cur << sh) | (prev >> (32 - sh)), sh == 0 -> cur
        const int w = w0 + k;
        if (32 * w >= pos + NB) break;              // past the row
        if (w >= lo_excl && w < hi_excl) flat[w] = v;
        else if (v) atomicOr(&flat[w], v);
    }
}

// Stream `rows` x NB floats (tile base 16-byte aligned) from the flat bit string.
template <int NB>
__device__ __forceinline__ void emit_flat_f32(const uint32_t* __restrict__ flat, const float4* __restrict__ tbl,
                                              float* __restrict__ out_tile, int rows, int lane, float v0, float v1) {
    const int total = rows * NB;
    const int n4 = total >> 2;
    float4* o4 = reinterpret_cast<float4*>(out_tile);
    int q = lane;
#pragma unroll 1
    for (; q + 96 < n4; q += 128) {
        uint32_t w[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) w[u] = flat[(q + 32 * u) >> 3];
        float4 f[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) f[u] = tbl[(w[u] >> (((q + 32 * u) & 7) << 2)) & 15u];
#pragma unroll
        for (int u = 0; u < 4; ++u) st_stream_f4(o4 + q + 32 * u, f[u]);
    }
    for (; q < n4; q += 32) st_stream_f4(o4 + q, tbl[(flat[q >> 3] >> ((q & 7) << 2)) & 15u]);
    // tail (< 4 floats) of a partial tile
    for (int e = (n4 << 2) + lane; e < total; e += 32) out_tile[e] = ((flat[e >> 5] >> (e & 31)) & 1u) ? v1 : v0;
}
