// Issue rate of the instruction kinds the softmax / selection phases are made of, one SM sub-partition view:
// 16 warps per SM (4 per scheduler), 16 independent chains per thread.  Result = warp-instructions per clock
// per scheduler (1.0 = full rate, 0.5 = a 16-lane pipe, 0.125 = MUFU).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -o profiles/ubench_pipes.bin profiles/ubench_pipes.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

constexpr int kIters = 512;

#define CHAIN16(BODY)                               \
    for (int it = 0; it < kIters; ++it) {           \
        _Pragma("unroll") for (int i = 0; i < 16; ++i) { BODY; } \
    }

template <int MODE>
__global__ void __launch_bounds__(512, 1) k(long long* out, float* sink, float seed, uint32_t useed) {
    float x[16], y[16];
    uint32_t u[16];
#pragma unroll
    for (int i = 0; i < 16; ++i) {
        x[i] = seed * (threadIdx.x + i) + 1.0f;
        y[i] = seed * i + 0.5f;
        u[i] = useed + threadIdx.x * 16 + i;
    }
    __syncthreads();
    const long long t0 = clock64();
    if (MODE == 0) CHAIN16(asm volatile("fma.rn.f32 %0, %0, %1, %1;" : "+f"(x[i]) : "f"(y[i])))
    if (MODE == 1) CHAIN16(asm volatile("add.f32 %0, %0, %1;" : "+f"(x[i]) : "f"(y[i])))
    if (MODE == 2) CHAIN16(asm volatile("max.f32 %0, %0, %1;" : "+f"(x[i]) : "f"(y[i])))
    if (MODE == 3) CHAIN16(asm volatile("max.f32 %0, %0, %1, %2;" : "+f"(x[i]) : "f"(y[i]), "f"(y[(i + 1) & 15])))
    if (MODE == 4) CHAIN16(asm volatile("{.reg .pred p; setp.ne.b32 p, %2, 0; selp.f32 %0, %0, %1, p;}" : "+f"(x[i]) : "f"(y[i]), "r"(u[i] & 1u)))
    if (MODE == 5) CHAIN16(asm volatile("cvt.rn.f16x2.f32 %0, %1, %2;" : "=r"(u[i]) : "f"(x[i]), "f"(__uint_as_float(u[i]))))
    if (MODE == 6) CHAIN16(asm volatile("{.reg .b16 a, b; mov.b32 {a, b}, %1; add.rn.f32.f16 %0, a, %0;}" : "+f"(x[i]) : "r"(u[i])))
    if (MODE == 7) CHAIN16(asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(u[i]) : "r"(u[(i + 1) & 15]), "r"(useed)))
    if (MODE == 8) CHAIN16(asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(x[i])))
    if (MODE == 9) CHAIN16(asm volatile("{.reg .b16 a, b; mov.b32 {a, b}, %1; cvt.f32.f16 %0, a;}" : "=f"(x[i]) : "r"(u[i])); u[i] += __float_as_uint(x[i]))
    if (MODE == 10) CHAIN16(asm volatile("ex2.approx.ftz.f32 %0, %0;\n\tfma.rn.f32 %1, %1, %2, %2;\n\tfma.rn.f32 %1, %1, %2, %2;\n\tfma.rn.f32 %1, %1, %2, %2;\n\tfma.rn.f32 %1, %1, %2, %2;" : "+f"(x[i]), "+f"(y[i]) : "f"(seed)))
    if (MODE == 11) CHAIN16(asm volatile("ex2.approx.ftz.f32 %0, %0;\n\tmax.f32 %1, %1, %2;\n\tmax.f32 %1, %1, %3;" : "+f"(x[i]), "+f"(y[i]) : "f"(seed), "f"(y[(i + 1) & 15])))
    if (MODE == 12) CHAIN16(asm volatile("fma.rn.f32 %0, %0, %2, %2;\n\tmax.f32 %1, %1, %2;" : "+f"(x[i]), "+f"(y[i]) : "f"(seed)))
    if (MODE == 13) CHAIN16(asm volatile("cvt.rn.f16x2.f32 %0, %1, %2;\n\tmax.f32 %1, %1, %2;" : "=r"(u[i]), "+f"(x[i]) : "f"(y[i])))
    if (MODE == 14) CHAIN16(asm volatile("{.reg .b16 a, b; mov.b32 {a, b}, %2; add.rn.f32.f16 %0, a, %0;\n\tmax.f32 %1, %1, %3;}" : "+f"(x[i]), "+f"(y[i]) : "r"(u[i]), "f"(seed)))
    if (MODE == 15) CHAIN16(asm volatile("{.reg .b16 a, b; mov.b32 {a, b}, %2; add.rn.f32.f16 %0, a, %0;\n\tfma.rn.f32 %1, %1, %3, %3;}" : "+f"(x[i]), "+f"(y[i]) : "r"(u[i]), "f"(seed)))
    const long long t1 = clock64();
    float acc = 0.f;
#pragma unroll
    for (int i = 0; i < 16; ++i) acc += x[i] + y[i] + __uint_as_float(u[i]);
    sink[blockIdx.x * 512 + threadIdx.x] = acc;
    if (threadIdx.x == 0) out[blockIdx.x] = t1 - t0;
}

template <int MODE>
static void run(const char* name, int instr_per_body) {
    long long* out;
    float* sink;
    cudaMalloc(&out, 148 * sizeof(long long));
    cudaMalloc(&sink, 148 * 512 * sizeof(float));
    k<MODE><<<148, 512>>>(out, sink, 1e-9f, 12345u);
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { printf("%s: %s\n", name, cudaGetErrorString(e)); return; }
    long long h;
    cudaMemcpy(&h, out, sizeof(h), cudaMemcpyDeviceToHost);
    // 4 warps per scheduler, 16 bodies per iteration
    printf("%-36s %9lld cycles  %.3f warp-instr/clk/scheduler\n", name, h, 4.0 * 16 * kIters * instr_per_body / (double)h);
    cudaFree(out);
    cudaFree(sink);
}

int main() {
    run<0>("FFMA", 1);
    run<1>("FADD", 1);
    run<2>("FMNMX", 1);
    run<3>("FMNMX3", 1);
    run<4>("ISETP + FSEL", 2);
    run<5>("F2FP.F16.F32.PACK_AB", 1);
    run<6>("FHADD (f32 + f16)", 1);
    run<7>("LOP3", 1);
    run<8>("MUFU.EX2", 1);
    run<9>("HADD2.F32 (f16 -> f32) + IADD", 2);
    run<10>("MUFU.EX2 + 4 FFMA", 5);
    run<11>("MUFU.EX2 + 2 FMNMX", 3);
    run<12>("FFMA + FMNMX", 2);
    run<13>("F2FP + FMNMX", 2);
    run<14>("FHADD + FMNMX", 2);
    run<15>("FHADD + FFMA", 2);
    return 0;
}
