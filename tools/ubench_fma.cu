// tools/ubench_fma.cu -- issue-rate micro-benchmarks of the sm_100a FP32 pipes (scalar vs packed f32x2, operand reuse):
// which instruction forms the pair sweep should be written in.  nvcc -arch=sm_100a -O3 -o build/ubench_fma tools/ubench_fma.cu
#include <cstdio>
#include <cuda_runtime.h>
typedef unsigned long long u64;
#define NCH 12
#define ITER 2048
__device__ __forceinline__ u64 pk(float a, float b) { u64 r; asm("mov.b64 %0, {%1,%2};" : "=l"(r) : "f"(a), "f"(b)); return r; }
__device__ __forceinline__ float lo(u64 v) { float a, b; asm("mov.b64 {%0,%1}, %2;" : "=f"(a), "=f"(b) : "l"(v)); return a + b; }

template <int MODE> __global__ void __launch_bounds__(512) k(float *out, long long *cyc, float seed) {
    float x[NCH], y[NCH], z[NCH];
    u64 X[NCH], Y[NCH], Z[NCH];
    const float a = seed * 0.999f + threadIdx.x * 1e-7f, b = seed * 0.001f + threadIdx.x * 1e-8f;
    const u64 A = pk(a, a * 1.00001f), B = pk(b, b * 1.1f);
#pragma unroll
    for (int i = 0; i < NCH; i++) {
        x[i] = threadIdx.x * 1e-3f + i; y[i] = 0.999f + (i + threadIdx.x * 0.01f) * 1e-4f * seed; z[i] = 1e-3f * (i + 1 + threadIdx.x) * seed;
        X[i] = pk(x[i], x[i] + 1); Y[i] = pk(y[i], y[i] * 1.0001f); Z[i] = pk(z[i], z[i] * 0.9f);
    }
    __syncthreads();
    long long t0 = clock64();
    for (int it = 0; it < ITER; it++) {
#pragma unroll
        for (int i = 0; i < NCH; i++) {
            if (MODE == 0) asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(x[i]) : "f"(a), "f"(b));                 // 2 shared operands
            if (MODE == 1) asm volatile("fma.rn.f32 %0, %1, %2, %0;" : "+f"(x[i]) : "f"(y[i]), "f"(z[i]));            // 3 distinct registers
            if (MODE == 2) asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(X[i]) : "l"(A), "l"(B));
            if (MODE == 3) asm volatile("fma.rn.f32x2 %0, %1, %2, %0;" : "+l"(X[i]) : "l"(Y[i]), "l"(Z[i]));
            if (MODE == 4) asm volatile("add.rn.f32 %0, %0, %1;" : "+f"(x[i]) : "f"(y[i]));
            if (MODE == 5) asm volatile("add.rn.f32x2 %0, %0, %1;" : "+l"(X[i]) : "l"(Y[i]));
            if (MODE == 6) asm volatile("mul.rn.f32 %0, %0, %1;" : "+f"(x[i]) : "f"(y[i]));
            if (MODE == 7) asm volatile("mul.rn.f32x2 %0, %0, %1;" : "+l"(X[i]) : "l"(Y[i]));
            if (MODE == 8) asm volatile("max.f32 %0, %0, %1, %2;" : "+f"(x[i]) : "f"(y[i]), "f"(z[i]));               // FMNMX3 (ALU pipe)
            if (MODE == 9) asm volatile("rcp.approx.ftz.f32 %0, %0;" : "+f"(x[i]));                                  // MUFU
            if (MODE == 10) asm volatile("fma.rn.f32 %0, %1, %2, %0;" : "+f"(x[i]) : "f"(y[i]), "f"(a));              // 2 distinct + 1 shared
            if (MODE == 11) asm volatile("fma.rn.f32x2 %0, %1, %2, %0;" : "+l"(X[i]) : "l"(Y[i]), "l"(A));
            if (MODE == 12) { unsigned g = __float_as_uint(x[i]); asm volatile("shf.l.wrap.b32 %0, %1, %0, 1;" : "+r"(g) : "r"(__float_as_uint(y[i]))); x[i] = __uint_as_float(g); }
        }
    }
    long long t1 = clock64();
    float s = 0;
#pragma unroll
    for (int i = 0; i < NCH; i++) s += x[i] + lo(X[i]);
    if (s == 12345.678f) out[0] = s;
    if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}

template <int MODE> void run(const char *name, int flop_per_inst, int threads) {
    float *out; long long *cyc, h[148];
    cudaMalloc(&out, 4); cudaMalloc(&cyc, 148 * 8);
    k<MODE><<<148, threads>>>(out, cyc, 1.0f);
    k<MODE><<<148, threads>>>(out, cyc, 1.0f);
    cudaDeviceSynchronize();
    cudaMemcpy(h, cyc, 148 * 8, cudaMemcpyDeviceToHost);
    double c = 0; for (int i = 0; i < 148; i++) c += h[i]; c /= 148;
    const double warps_per_smsp = threads / 32.0 / 4.0;
    const double inst_per_smsp = (double)ITER * NCH * warps_per_smsp;
    printf("%-44s threads %4d  cycles/warp-inst/SMSP %.3f   (%.1f flop/clk/SM)\n", name, threads, c / inst_per_smsp,
           flop_per_inst * 32.0 * 4.0 * inst_per_smsp / c);
    cudaFree(out); cudaFree(cyc);
}

int main() {
    for (int threads : {128, 512}) {
        run<0>("FFMA  x = x*a + b (2 shared operands)", 2, threads);
        run<10>("FFMA  x = y*a + x (1 shared operand)", 2, threads);
        run<1>("FFMA  x = y*z + x (3 distinct registers)", 2, threads);
        run<2>("FFMA2 X = X*A + B (2 shared operands)", 4, threads);
        run<11>("FFMA2 X = Y*A + X (1 shared operand)", 4, threads);
        run<3>("FFMA2 X = Y*Z + X (3 distinct pairs)", 4, threads);
        run<4>("FADD  x = x + y", 1, threads);
        run<5>("FADD2 X = X + Y", 2, threads);
        run<6>("FMUL  x = x * y", 1, threads);
        run<7>("FMUL2 X = X * Y", 2, threads);
        run<8>("FMNMX3 x = max(x, y, z)", 2, threads);
        run<12>("SHF.L.W", 1, threads);
        run<9>("MUFU.RCP", 1, threads);
    }
    return 0;
}
