// Shared device helpers for the rl4co_b200 kernels (sm_100a).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <math.h>

#include "../../include/rl4co_b200.h"

namespace r4 {

constexpr int kE = 128;        // embed_dim   (am/policy.py:54)
constexpr int kH = 8;          // num_heads   (am/policy.py:56)
constexpr int kD = kE / kH;    // head dim 16
constexpr int kMaxNodes = 128; // v1 limit on n_nodes (4 x 32-lane blocks)
constexpr int kNB = kMaxNodes / 32;
constexpr unsigned kFull = 0xffffffffu;

// host-side error text
void set_error(const char* fmt, ...);

#define R4_CHECK_ARG(cond, ...)              \
    do {                                     \
        if (!(cond)) {                       \
            r4::set_error(__VA_ARGS__);      \
            return RL4CO_E_ARG;              \
        }                                    \
    } while (0)

#define R4_CHECK_LAUNCH(what)                                                   \
    do {                                                                        \
        cudaError_t e__ = cudaGetLastError();                                   \
        if (e__ != cudaSuccess) {                                               \
            r4::set_error("%s: %s", what, cudaGetErrorString(e__));             \
            return RL4CO_E_LAUNCH;                                              \
        }                                                                       \
    } while (0)

// ---- device-side sticky error word -------------------------------------------------
__device__ __forceinline__ void raise_err(int32_t* err, int bits, int64_t row) {
    if (err == nullptr) return;
    atomicOr(err, bits);
    atomicCAS(err + 1, 0, (int)(row + 1));
}

// ---- counter-based noise (must match oracle/rollout_oracle.py:_mix32/noise_uniform) ---
__device__ __host__ __forceinline__ uint32_t mix32(uint32_t x) {
    x ^= x >> 16;
    x *= 0x7FEB352Du;
    x ^= x >> 15;
    x *= 0x846CA68Bu;
    x ^= x >> 16;
    return x;
}
// key for (seed, b, s, t); node mixed in by noise_u
__device__ __host__ __forceinline__ uint32_t noise_key(uint32_t seed, int64_t b, int64_t s, int32_t t) {
    uint32_t h = mix32(seed ^ ((uint32_t)b * 0x9E3779B1u));
    h = mix32(h ^ ((uint32_t)s * 0x85EBCA77u));
    h = mix32(h ^ ((uint32_t)t * 0xC2B2AE3Du));
    return h;
}
__device__ __forceinline__ float noise_u(uint32_t key, int32_t node) {
    uint32_t h = mix32(key ^ ((uint32_t)node * 0x27D4EB2Fu));
    return ((float)(h >> 9) + 0.5f) * 1.1920928955078125e-07f;  // 2^-23, exact in fp32
}
__device__ __forceinline__ float noise_gumbel(uint32_t key, int32_t node) {
    return -logf(-logf(noise_u(key, node)));
}

// ---- warp reductions ---------------------------------------------------------------
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(kFull, v, o));
    return v;
}
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(kFull, v, o);
    return v;
}
// argmax with first-index tie rule (torch.argmax, decoding.py:395). NaN never wins
// unless all are NaN; callers flag NaN separately.
__device__ __forceinline__ void warp_argmax(float& v, int& idx) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        float ov = __shfl_xor_sync(kFull, v, o);
        int oi = __shfl_xor_sync(kFull, idx, o);
        if (ov > v || (ov == v && oi < idx)) {
            v = ov;
            idx = oi;
        }
    }
}

__device__ __forceinline__ float4 ld4(const float* p) { return *reinterpret_cast<const float4*>(p); }

}  // namespace r4
