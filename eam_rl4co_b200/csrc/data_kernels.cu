// On-device instance generation: the input spec of tsp/generator.py:53-62 and
// cvrp/generator.py:120-147 (uniform locations in [0,1)^2, depot = one more location, integer demand 1..9 divided by
// the capacity) drawn from the counter hash of common.cuh, keyed by (seed, GLOBAL instance index, element): a shard
// generates exactly the instances it would have received from a host-side dataset, whatever the world size.
#include "common.cuh"

namespace r4 {

__device__ __host__ __forceinline__ float unit_uniform(uint32_t key, uint32_t elem) {
    const uint32_t h = mix32(key ^ (elem * 0x85EBCA77u));
    return (float)(h >> 8) * 5.9604644775390625e-08f;   // 24 random bits * 2^-24: [0, 1), exact in fp32
}

// one thread per (instance, element): elements [0, 2*(n_loc + has_depot)) are coordinates, then n_loc demands
__global__ void generate_instances_kernel(uint32_t seed, int64_t inst_offset, int64_t B, int n_loc, int has_depot,
                                          float capacity, float* __restrict__ locs, float* __restrict__ depot,
                                          float* __restrict__ demand) {
    const int n_xy = 2 * (n_loc + has_depot);
    const int per = n_xy + (demand != nullptr ? n_loc : 0);
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= B * per) return;
    const int64_t b = i / per;
    const int e = (int)(i - b * per);
    const uint32_t key = mix32(seed ^ ((uint32_t)(b + inst_offset) * 0x9E3779B1u) ^ 0x5bd1e995u);
    const float u = unit_uniform(key, (uint32_t)e);
    if (e < n_xy) {
        if (has_depot) {   // cvrp/generator.py:126-131: the depot is the first sampled location
            if (e < 2) depot[b * 2 + e] = u;
            else locs[b * (int64_t)(2 * n_loc) + (e - 2)] = u;
        } else {
            locs[b * (int64_t)(2 * n_loc) + e] = u;
        }
    } else {               // :134-140: (Uniform(0, 9).int() + 1) / capacity
        demand[b * (int64_t)n_loc + (e - n_xy)] = (float)((int)(u * 9.0f) + 1) / capacity;
    }
}

}  // namespace r4

extern "C" int rl4co_generate_instances(int32_t env, uint32_t seed, int64_t inst_offset, int64_t B, int32_t n_loc,
                                        float capacity, float* locs, float* depot, float* demand, void* stream) {
    R4_CHECK_ARG(locs && B >= 0 && n_loc >= 1, "generate_instances: bad args");
    const bool cvrp = env == RL4CO_ENV_CVRP;
    R4_CHECK_ARG(env == RL4CO_ENV_TSP || cvrp, "generate_instances: env must be tsp or cvrp");
    R4_CHECK_ARG(!cvrp || (depot && demand && capacity > 0.f), "generate_instances: cvrp needs depot, demand, capacity");
    if (B == 0) return RL4CO_OK;
    const int per = 2 * (n_loc + (cvrp ? 1 : 0)) + (cvrp ? n_loc : 0);
    const int64_t total = B * per;
    r4::generate_instances_kernel<<<(unsigned)((total + 255) / 256), 256, 0, (cudaStream_t)stream>>>(
        seed, inst_offset, B, n_loc, cvrp ? 1 : 0, capacity, locs, cvrp ? depot : nullptr, cvrp ? demand : nullptr);
    R4_CHECK_LAUNCH("generate_instances");
    return RL4CO_OK;
}
