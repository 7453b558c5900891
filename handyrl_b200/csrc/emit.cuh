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

// Atomic-free packing for NB >= 32: a flat word then holds bits of at most two neighbouring rows, so the
// partial last word ("tail") of row r is handed to row r+1 (by shuffle when the rows sit in different lanes)
// and OR-ed into that row's first word.  No clearing pass; one __syncwarp() after the stores.
template <int NB, int RW>
__device__ __forceinline__ uint32_t flat_row_tail(int row, const uint32_t (&bits)[RW]) {
    static_assert(NB >= 32, "shuffle packing needs rows of at least one word");
    const int pos = row * NB, end = pos + NB;
    if ((end & 31) == 0) return 0u;
    // bits [32*wt - pos, NB) of the row, right-aligned in word wt = end >> 5
    const int off = ((end >> 5) << 5) - pos;  // first row bit that lands in the tail word (multiple of ... any)
    uint32_t t = 0;
#pragma unroll
    for (int k = 0; k < RW; ++k) {
        const int lo = 32 * k;  // row bits [lo, lo+32) live in bits[k]
        if (off >= lo && off < lo + 32) {
            t = bits[k] >> (off - lo);
            if (k + 1 < RW && off - lo > 0) t |= bits[k + 1] << (32 - (off - lo));
        }
    }
    return t;
}

// Store row `row`; `carry` = tail of row-1 (0 for the first row of the tile).  The last row also stores its tail.
template <int NB, int RW>
__device__ __forceinline__ void flat_row_store(uint32_t* flat, int row, const uint32_t (&bits)[RW], uint32_t carry,
                                               bool is_last_row, uint32_t tail) {
    const int pos = row * NB;
    const int w0 = pos >> 5, sh = pos & 31;
    const int wt = (pos + NB) >> 5;
    flat[w0] = (bits[0] << sh) | carry;
#pragma unroll
    for (int k = 1; k <= RW; ++k) {
        const uint32_t cur = k < RW ? bits[k] : 0u;
        if (w0 + k < wt) flat[w0 + k] = __funnelshift_l(bits[k - 1], cur, sh);
    }
    if (is_last_row && ((pos + NB) & 31) != 0) flat[wt] = tail;
}

// One row per lane (the env kernels).  All 32 lanes must call; lanes >= n_rows pass zero bits.
template <int NB, int RW>
__device__ __forceinline__ void flat_pack_warp(uint32_t* flat, int lane, int n_rows, const uint32_t (&bits)[RW]) {
    const uint32_t tail = flat_row_tail<NB, RW>(lane, bits);
    const uint32_t carry = __shfl_up_sync(0xffffffffu, tail, 1);
    if (lane < n_rows) flat_row_store<NB, RW>(flat, lane, bits, lane > 0 ? carry : 0u, lane == n_rows - 1, tail);
}

// nibble -> float4 without touching shared memory (B200RL_EMIT_ALU, default): 4 predicated selects on the
// ALU pipes instead of one LDS.128 per 16 bytes — the L1/LSU pipe, which also carries the global stores, was
// the busiest unit with the shared-memory table (profiles/r1_step_full.md: L1/TEX 66 %, issue 29 %).
#ifndef B200RL_EMIT_TABLE
#define B200RL_EMIT_ALU 1
#endif
__device__ __forceinline__ float4 nibble_expand(uint32_t nib, const float4* __restrict__ tbl, float v0, float v1) {
#ifdef B200RL_EMIT_ALU
    return make_float4((nib & 1u) ? v1 : v0, (nib & 2u) ? v1 : v0, (nib & 4u) ? v1 : v0, (nib & 8u) ? v1 : v0);
#else
    return tbl[nib & 15u];
#endif
}

// Stream `rows` x NB floats (tile base 16-byte aligned) from the flat bit string.
template <int NB>
__device__ __forceinline__ void emit_flat_f32(const uint32_t* __restrict__ flat, const float4* __restrict__ tbl,
                                              float* __restrict__ out_tile, int rows, int lane, float v0, float v1) {
    const int total = rows * NB;
    const int n4 = total >> 2;
    float4* o4 = reinterpret_cast<float4*>(out_tile);
    // lane q always reads nibble (q & 7) of word (q >> 3) + 4*j: constant shift, constant address stride
    const int sh = (lane & 7) << 2;
    const uint32_t* fw = flat + (lane >> 3);
    if (NB * 8 <= 256 && rows <= 32) {
        // small tiles (TicTacToe, one row per lane: 216 and 72 float4): fully unrolled, constant word offsets,
        // predicated stores; make_batch's two-player cells (up to 64 rows) take the general loops below
        constexpr int IT = (NB * 8 + 31) / 32;
#pragma unroll
        for (int u = 0; u < IT; ++u) {
            const int qq = lane + 32 * u;
            if (qq < n4) st_stream_f4(o4 + qq, nibble_expand(fw[4 * u] >> sh, tbl, v0, v1));
        }
        for (int e = (n4 << 2) + lane; e < total; e += 32) out_tile[e] = ((flat[e >> 5] >> (e & 31)) & 1u) ? v1 : v0;
        return;
    }
    int q = lane;
#pragma unroll 1
    for (; q + 96 < n4; q += 128) {
        uint32_t w[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) w[u] = fw[(q >> 3 << 0) - (lane >> 3) + 4 * u];
        float4 f[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) f[u] = nibble_expand(w[u] >> sh, tbl, v0, v1);
#pragma unroll
        for (int u = 0; u < 4; ++u) st_stream_f4(o4 + q + 32 * u, f[u]);
    }
    for (; q < n4; q += 32) st_stream_f4(o4 + q, nibble_expand(flat[q >> 3] >> sh, tbl, v0, v1));
    // tail (< 4 floats) of a partial tile
    for (int e = (n4 << 2) + lane; e < total; e += 32) out_tile[e] = ((flat[e >> 5] >> (e & 31)) & 1u) ? v1 : v0;
}

// ---------------------------------------------------------------------------------------------------------------
// Rows shorter than one word, one row per lane (TicTacToe: 27 observation bits and 9 mask bits per game).  The 32
// rows of a warp are only NB words of flat string, so the whole tile fits in ONE register per lane: lane w builds
// flat word w from the <= SPAN rows that overlap bits [32w, 32w + 32) with shuffles, and float4 q = lane + 32u of the
// tile is nibble (lane & 7) of word (lane >> 3) + 4u, fetched by one more shuffle.  No shared memory, no atomicOr, no
// __syncwarp, and the store predicate of a full tile is a compile-time constant.  The TicTacToe rollout is bound by
// instruction latency at low occupancy (2,048 warps for 65,536 games), not by HBM: the flat-string-in-shared path above
// cost it ~300 of its 656 warp instructions per ply (profiles/r2_ttt_rollout.md).  Same bytes as emit_flat_f32.
// All 32 lanes must call; lanes >= rows pass zero bits.
// ---------------------------------------------------------------------------------------------------------------
template <int NB>
__device__ __forceinline__ uint32_t short_rows_word(uint32_t bits, int lane) {
    static_assert(NB >= 1 && NB < 32, "rows shorter than one word");
    constexpr int SPAN = (NB + 30) / NB + 1;   // rows a 32-bit window starting inside a row can touch
    const int p = 32 * lane, r0 = p / NB, o = p - r0 * NB;
    uint32_t w = 0;
#pragma unroll
    for (int j = 0; j < SPAN; ++j) {
        const int r = r0 + j;
        uint32_t b = __shfl_sync(0xffffffffu, bits, r & 31);
        const int sh = j * NB - o;             // where bit 0 of row r lands in this word (negative only for j = 0)
        // sh <= j * NB, so only the last rows of the span can start past the word; rows >= 32 (which wrap around in
        // the shuffle) only ever reach the words of lanes >= NB, which nobody reads
        if (j * NB >= 32 && sh >= 32) b = 0u;
        w |= j == 0 ? b >> o : b << (sh & 31);
    }
    return w;
}

// `tbl` = the 16 nibble images of (v0, v1) in shared memory (nibble_table_init): one LDS.128 per float4 instead of
// the ~8 ALU instructions of nibble_expand - this path is bound by instruction issue, its LSU pipe is at 13 %.
template <int NB>
__device__ __forceinline__ void emit_short_rows_f32(uint32_t bits, const float4* __restrict__ tbl,
                                                    float* __restrict__ out_tile, int rows, int lane, float v0,
                                                    float v1) {
    const uint32_t word = short_rows_word<NB>(bits, lane);
    constexpr int N4 = 8 * NB;                 // float4 of a full tile of 32 rows
    constexpr int IT = (N4 + 31) / 32;
    static_assert(3 + 4 * (IT - 1) < 32, "source lanes of the word shuffles");
    const int sh = (lane & 7) << 2, src = lane >> 3;
    float4* o4 = reinterpret_cast<float4*>(out_tile);
    if (rows == 32) {
#pragma unroll
        for (int u = 0; u < IT; ++u) {
            const float4 f = tbl[(__shfl_sync(0xffffffffu, word, src + 4 * u) >> sh) & 15u];
            if (32 * u + 32 <= N4) st_stream_f4(o4 + lane + 32 * u, f);
            else if (lane < N4 - 32 * u) st_stream_f4(o4 + lane + 32 * u, f);
        }
        return;
    }
    // the ragged last group of a batch
    const int total = rows * NB, n4 = total >> 2;
#pragma unroll
    for (int u = 0; u < IT; ++u) {
        const float4 f = tbl[(__shfl_sync(0xffffffffu, word, src + 4 * u) >> sh) & 15u];
        if (lane + 32 * u < n4) st_stream_f4(o4 + lane + 32 * u, f);
    }
    const int e = (n4 << 2) + lane;            // < 4 floats left
    const uint32_t wv = __shfl_sync(0xffffffffu, word, (e >> 5) & 31);
    if (e < total) out_tile[e] = ((wv >> (e & 31)) & 1u) ? v1 : v0;
}

// ---------------------------------------------------------------------------------------------------------------
// TMA variant of the streaming loop (profiles/r2_probe_write_tma.txt: 7.04 -> 7.3-7.45 TB/s on the same tiles).
// The warp expands the tile chunk by chunk into a private ring of NBUF staging buffers in shared memory
// (st.shared.v4, lane q -> float4 q: conflict-free) and ONE lane hands each finished chunk to the copy engine:
//     cp.async.bulk.global.shared::cta.bulk_group   (SASS UBLKCP.G.S)
// so the global stores leave the LSU pipe and reach L2 as CH-byte bursts.  A buffer is reused once the bulk group
// that read it has completed (cp.async.bulk.wait_group.read).  Bulk groups are tracked per thread: lane 0 issues and
// waits for all of them, the other lanes follow through __syncwarp.
// ---------------------------------------------------------------------------------------------------------------
template <int CH, int NBUF>
struct TmaStage {
    static_assert(CH % 16 == 0 && NBUF >= 1, "chunks are multiples of 16 bytes");
    float4* buf;   // NBUF x CH bytes, 128-byte aligned, private to the warp
    int c;         // chunks committed so far (all lanes keep the same count)

    __device__ __forceinline__ float4* acquire(int lane) {
        if (c >= NBUF) {   // the group that last read this buffer is NBUF commits old
            if (lane == 0) asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(NBUF - 1) : "memory");
            __syncwarp();
        }
        return buf + (c % NBUF) * (CH / 16);
    }
    // all lanes have written their part of `b` (generic proxy) -> visible to the async proxy -> bulk copy
    __device__ __forceinline__ void commit(int lane, void* gdst, const float4* b, int bytes) {
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        __syncwarp();
        if (lane == 0) {
            asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(gdst),
                         "r"((uint32_t)__cvta_generic_to_shared(b)), "r"(bytes)
                         : "memory");
            asm volatile("cp.async.bulk.commit_group;" ::: "memory");
        }
        ++c;
    }
    // before the CTA exits (or the buffers are used for anything else): every bulk read of shared memory is done
    __device__ __forceinline__ void drain(int lane) {
        if (lane == 0) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
        __syncwarp();
    }
};

// Same bytes as emit_flat_f32 (tile base 16-byte aligned).
template <int NB, int CH, int NBUF>
__device__ __forceinline__ void emit_flat_f32_tma(const uint32_t* __restrict__ flat, float* __restrict__ out_tile,
                                                  int rows, int lane, float v0, float v1, TmaStage<CH, NBUF>& st) {
    constexpr int CH4 = CH / 16;
    static_assert(CH4 % 32 == 0, "a chunk is a whole number of warp-wide float4 rows");
    const int total = rows * NB;
    const int n4 = total >> 2;
    const int sh = (lane & 7) << 2;                      // q0 and the lane stride are multiples of 8 float4
#pragma unroll 1
    for (int q0 = 0; q0 < n4; q0 += CH4) {
        const int nq = min(CH4, n4 - q0);
        float4* b = st.acquire(lane);
        const uint32_t* fw = flat + ((q0 + lane) >> 3);
#pragma unroll 4
        for (int q = lane; q < nq; q += 32) b[q] = nibble_expand(fw[(q - lane) >> 3] >> sh, nullptr, v0, v1);
        st.commit(lane, out_tile + (size_t)q0 * 4, b, nq * 16);
    }
    for (int e = (n4 << 2) + lane; e < total; e += 32) out_tile[e] = ((flat[e >> 5] >> (e & 31)) & 1u) ? v1 : v0;
}
