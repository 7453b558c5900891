// Shared helpers for the b200rl CUDA library (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdarg.h>

#define B200RL_OK 0
#define B200RL_ERR_ARG 1
#define B200RL_ERR_CUDA 2
#define B200RL_ERR_UNSUPPORTED 3

// Thread-local last-error message, read through b200rl_last_error().
void b200rl_set_error(const char* fmt, ...);

// tuning knobs (abi.cu): environment default read once, overridable through b200rl_set_tuning
enum {
    B200RL_TUNE_ROLLOUT_CHUNK = 0, B200RL_TUNE_TARGET_WS_STAGES, B200RL_TUNE_TARGET_WS_ROW,
    B200RL_TUNE_TARGET_WS_WIDE_ROWS, B200RL_TUNE_TARGET_GENERIC, B200RL_TUNE_TARGET_NO_DUAL,
    B200RL_TUNE_TARGET_PERSISTENT, B200RL_TUNE_EMIT_TMA, B200RL_TUNE_TARGET_WS_ROWLANE, B200RL_TUNE_COUNT
};
bool b200rl_tuning(int knob, long long* out);   // true when the knob is set (env or set_tuning)

static inline int b200rl_check_launch(const char* what) {
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) {
        b200rl_set_error("%s: %s", what, cudaGetErrorString(e));
        return B200RL_ERR_CUDA;
    }
    return B200RL_OK;
}

#define B200RL_REQUIRE(cond, ...)            \
    do {                                     \
        if (!(cond)) {                       \
            b200rl_set_error(__VA_ARGS__);   \
            return B200RL_ERR_ARG;           \
        }                                    \
    } while (0)

// 128-bit streaming store: output tensors are written once and consumed by a
// later kernel (the net / the loss), so do not allocate them in L1.
__device__ __forceinline__ void st_stream_f4(float4* p, float4 v) {
    asm volatile("st.global.L1::no_allocate.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p), "f"(v.x), "f"(v.y),
                 "f"(v.z), "f"(v.w)
                 : "memory");
}

__device__ __forceinline__ float4 ld_stream_f4(const float4* p) {
    float4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0, %1, %2, %3}, [%4];"
                 : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w)
                 : "l"(p));
    return r;
}

__device__ __forceinline__ uint4 ld_stream_u4(const uint4* p) {
    uint4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0, %1, %2, %3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
                 : "l"(p));
    return r;
}

// Trajectory ring: [N][D] records of 32 bytes (8 words), slot-major: the record of (iteration it, slot i) is at
// (i * D + it % D) * 32 bytes, so the T plies of a training window are T x 32 B CONTIGUOUS and a warp of
// make_batch reads them as one coalesced 1 KB request (the ply-major [D][N] layout cost make_batch 14 %: 2 M
// isolated 32-byte sector reads, 4x read amplification).  The rollout pays for it with strided writes, which it
// combines into full 128-byte lines (4 plies) through shared memory.  One 256-bit access per record.
// x mod m for ring indices: rings are normally sized to powers of two, where the 64-bit remainder (~70
// instructions, 13 % of the TicTacToe rollout's instruction stream before this) is one AND.
__device__ __forceinline__ uint64_t ring_mod(uint64_t x, uint64_t m) {
    return (m & (m - 1)) == 0 ? (x & (m - 1)) : (x % m);
}

// Ring address of the record of (iteration it, game slot i), in 32-byte records.
//   B200RL_RING_TILE == 0: slot-major [N][D] - the D plies of a slot are contiguous (a training window = T x 32 B).
//   B200RL_RING_TILE == TP: tiled [N/32][D/TP][32 games][TP plies] - a warp's 32 games x TP plies are one
//       contiguous 32 x TP x 32 B block, a window is T/TP pieces of TP x 32 B (N padded to 32, D % TP == 0).
#ifndef B200RL_RING_TILE
#define B200RL_RING_TILE 0
#endif
__device__ __forceinline__ int64_t ring_index(int64_t D, uint64_t it, int64_t i) {
    const int64_t r = (int64_t)ring_mod(it, (uint64_t)D);
#if B200RL_RING_TILE == 0
    return i * D + r;
#else
    constexpr int64_t TP = B200RL_RING_TILE;
    return (((i >> 5) * (D / TP) + r / TP) * 32 + (i & 31)) * TP + (r % TP);
#endif
}

__device__ __forceinline__ void st_record(void* traj, int64_t D, int64_t N, uint64_t it, int64_t i,
                                          const uint32_t (&r)[8]) {
    uint32_t* p = reinterpret_cast<uint32_t*>(traj) + ring_index(D, it, i) * 8;
    asm volatile("st.global.v8.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"l"(p), "r"(r[0]), "r"(r[1]), "r"(r[2]),
                 "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7])
                 : "memory");
}

__device__ __forceinline__ void ld_record(const void* traj, int64_t D, int64_t N, uint64_t it, int64_t i,
                                          uint32_t (&r)[8]) {
    const uint32_t* p = reinterpret_cast<const uint32_t*>(traj) + ring_index(D, it, i) * 8;
    asm volatile("ld.global.nc.v8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7])
                 : "l"(p));
}

// Counter-based RNG: splitmix64 finaliser over (seed, stream id, counter).
// Used for the device-side uniform policy and ParallelTicTacToe's selected player;
// parity with the reference is defined on action tapes, never on RNG streams.
__host__ __device__ __forceinline__ uint64_t b200rl_mix64(uint64_t z) {
    z += 0x9E3779B97F4A7C15ull;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}

// 64 random bits per (seed, stream, counter): window selection takes two 32-bit draws from each
__host__ __device__ __forceinline__ uint64_t b200rl_rand64(uint64_t seed, uint64_t stream, uint64_t counter) {
    return b200rl_mix64(seed ^ b200rl_mix64(stream * 0xD6E8FEB86659FD93ull + counter));
}

__host__ __device__ __forceinline__ uint32_t b200rl_rand32(uint64_t seed, uint64_t stream, uint64_t counter) {
    uint64_t h = b200rl_mix64(seed ^ b200rl_mix64(stream * 0xD6E8FEB86659FD93ull + counter));
    return (uint32_t)(h >> 32);
}
