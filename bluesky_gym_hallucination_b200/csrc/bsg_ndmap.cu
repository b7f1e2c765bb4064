// bsg_ndmap.cu -- ND-map: 4D (t, z, y, x) spatial-temporal occupancy / conflict grid over straight-line
// look-ahead trajectories.  No reference implementation exists (SURVEY.md 0.3 tier C): the spec is this
// repo's own (DESIGN.md "ND-map"); the CPU restatement of the same spec is oracle/c/bsgo_ndmap.c.
// One CTA rasterises one (env, time-slab): aircraft samples are counted with shared-memory atomics into
// a u16 grid (two counters per 32-bit word), a separable 3x3x3 box sum marks cells whose neighbourhood
// holds >= 2 aircraft, and the slab is streamed to HBM with 16-byte stores (the kernel is HBM-write bound).
// Also holds the FP32 / FP64 FMA-chain micro-benchmarks (the measured non-tensor roofline denominators).
#include <cuda_runtime.h>
#include <cstdint>
#include "../../include/bsgcore.h"

namespace {

constexpr int kNdThreads = 512;

// POW2: nx / 8, ny and nx * ny are powers of two (the C4 grid 64 x 64 x 8): the index arithmetic of the two box-sum passes
// is shifts and masks instead of three runtime integer divisions per item (they were ~40 % of the instructions).
template <bool POW2>
__global__ void __launch_bounds__(kNdThreads, 2) ndmap_kernel(const bsg_ndmap_params p, float icx, float icy, float icz,
                                                           int lgx, int lplane,
                                                           const float *__restrict__ rec, uint8_t *__restrict__ occ,
                                                           uint8_t *__restrict__ conf, int32_t *__restrict__ ncell) {
    extern __shared__ __align__(16) unsigned char smem[];
    const int nx = p.nx, ny = p.ny, nz = p.nz, n = p.num_ac;
    const int cells = nx * ny * nz;
    uint16_t *cnt = reinterpret_cast<uint16_t *>(smem);          // [cells]
    uint8_t *sxy8 = smem + (size_t)cells * 2;                    // [cells] u8: (y, x) box sums of counts clipped to 2
    __shared__ int s_ncf;
    const int k = blockIdx.x, e = blockIdx.y;
    uint32_t *cw = reinterpret_cast<uint32_t *>(cnt);
    for (int c = threadIdx.x; c < cells / 8; c += kNdThreads) reinterpret_cast<uint4 *>(cw)[c] = make_uint4(0u, 0u, 0u, 0u);
    if (threadIdx.x == 0) s_ncf = 0;
    __syncthreads();
    const float *r = rec + (size_t)e * 6 * n;
    const float t = (float)(k + 1) * p.dt;
    for (int a = threadIdx.x; a < n; a += kNdThreads) {
        const float px = fmaf(r[3 * n + a], t, r[a]);
        const float py = fmaf(r[4 * n + a], t, r[n + a]);
        const float pz = fmaf(r[5 * n + a], t, r[2 * n + a]);
        const int ix = (int)floorf((px - p.x0) * icx);
        const int iy = (int)floorf((py - p.y0) * icy);
        const int iz = (int)floorf((pz - p.z0) * icz);
        if (ix < 0 || ix >= nx || iy < 0 || iy >= ny || iz < 0 || iz >= nz) continue;
        const int c = (iz * ny + iy) * nx + ix;
        atomicAdd(&cw[c >> 1], (c & 1) ? 0x10000u : 1u);         // u16 lanes: no carry for n <= 65535
    }
    __syncthreads();
    // pass 1: 3x3 box sum in the (y, x) plane, 8 cells (one 16-byte vector of packed u16 pairs) per item.  The counts are
    // clipped to 2 first (sum >= 2 <=> clipped sum >= 2), so a 27-cell sum is at most 54 and the plane sums are kept as
    // BYTES: the grid costs 3 B per cell of shared memory instead of 4 and two CTAs share an SM -- the HBM stores of one
    // slab overlap the counting / summing of another.  The +-1 neighbours in x come from funnel shifts.
    const int gx = nx >> 3;                                   // 16-byte groups per row
    for (int g = threadIdx.x; g < cells / 8; g += kNdThreads) {
        const int x8 = POW2 ? (g & (gx - 1)) : g % gx, yz = POW2 ? (g >> lgx) : g / gx, y = POW2 ? (yz & (ny - 1)) : yz % ny;
        uint32_t rs[6] = { 0u, 0u, 0u, 0u, 0u, 0u };
#pragma unroll
        for (int dy = -1; dy <= 1; dy++) {
            const int yy = y + dy;
            if (yy < 0 || yy >= ny) continue;
            const int wb = ((yz + dy) * nx + x8 * 8) >> 1;    // first word of the group in row yy
            const uint4 m = *reinterpret_cast<const uint4 *>(cw + wb);
            if constexpr (!POW2) rs[0] += x8 > 0 ? __vminu2(cw[wb - 1], 0x00020002u) : 0u;
            rs[1] += __vminu2(m.x, 0x00020002u); rs[2] += __vminu2(m.y, 0x00020002u);
            rs[3] += __vminu2(m.z, 0x00020002u); rs[4] += __vminu2(m.w, 0x00020002u);
            if constexpr (!POW2) rs[5] += x8 < gx - 1 ? __vminu2(cw[wb + 4], 0x00020002u) : 0u;
        }
        if constexpr (POW2) {
            // the groups of a row sit in neighbouring lanes (gx divides 32, whole warps in the loop): the column sums of the
            // words next to this group come from the lanes beside it instead of 6 more (bank-conflicting) loads
            const uint32_t lft = __shfl_up_sync(0xffffffffu, rs[4], 1), rgt = __shfl_down_sync(0xffffffffu, rs[1], 1);
            rs[0] = x8 > 0 ? lft : 0u;
            rs[5] = x8 < gx - 1 ? rgt : 0u;
        }
        uint4 o;
        o.x = rs[1] + __funnelshift_l(rs[0], rs[1], 16) + __funnelshift_r(rs[1], rs[2], 16);
        o.y = rs[2] + __funnelshift_l(rs[1], rs[2], 16) + __funnelshift_r(rs[2], rs[3], 16);
        o.z = rs[3] + __funnelshift_l(rs[2], rs[3], 16) + __funnelshift_r(rs[3], rs[4], 16);
        o.w = rs[4] + __funnelshift_l(rs[3], rs[4], 16) + __funnelshift_r(rs[4], rs[5], 16);
        *reinterpret_cast<uint2 *>(sxy8 + ((size_t)g << 3)) = make_uint2(__byte_perm(o.x, o.y, 0x6420), __byte_perm(o.z, o.w, 0x6420));
    }
    __syncthreads();
    // pass 2: sum over z (byte lanes, <= 54), emit 16 cells per item as one 16-byte store each for occ and conf
    const size_t slab = ((size_t)e * p.nt + k) * cells;
    const int plane = nx * ny;                                // bytes per z plane of sxy8
    int ncf = 0;
    for (int g = threadIdx.x; g < cells / 16; g += kNdThreads) {
        const int w0 = g << 3, b0 = g << 4, z = POW2 ? (b0 >> lplane) : b0 / plane;
        uint4 a = *reinterpret_cast<const uint4 *>(sxy8 + b0);
        if (z > 0) { const uint4 q = *reinterpret_cast<const uint4 *>(sxy8 + b0 - plane); a.x += q.x; a.y += q.y; a.z += q.z; a.w += q.w; }
        if (z < nz - 1) { const uint4 q = *reinterpret_cast<const uint4 *>(sxy8 + b0 + plane); a.x += q.x; a.y += q.y; a.z += q.z; a.w += q.w; }
        // byte >= 2 <=> bit 7 of byte + 0x7E (no carry: byte <= 54); the flags come out in cell order
        uint32_t fw[4] = { ((a.x + 0x7E7E7E7Eu) >> 7) & 0x01010101u, ((a.y + 0x7E7E7E7Eu) >> 7) & 0x01010101u,
                           ((a.z + 0x7E7E7E7Eu) >> 7) & 0x01010101u, ((a.w + 0x7E7E7E7Eu) >> 7) & 0x01010101u };
        const uint4 c0 = *reinterpret_cast<const uint4 *>(cw + w0), c1 = *reinterpret_cast<const uint4 *>(cw + w0 + 4);
        const uint32_t cv[8] = { c0.x, c0.y, c0.z, c0.w, c1.x, c1.y, c1.z, c1.w };
        uint32_t ow[4];
#pragma unroll
        for (int w = 0; w < 4; w++) {
            ow[w] = __byte_perm(__vminu2(cv[2 * w], 0x00ff00ffu), __vminu2(cv[2 * w + 1], 0x00ff00ffu), 0x6420);
            ncf += __popc(fw[w]);
        }
        *reinterpret_cast<uint4 *>(occ + slab + ((size_t)g << 4)) = make_uint4(ow[0], ow[1], ow[2], ow[3]);
        *reinterpret_cast<uint4 *>(conf + slab + ((size_t)g << 4)) = make_uint4(fw[0], fw[1], fw[2], fw[3]);
    }
#pragma unroll
    for (int m = 16; m > 0; m >>= 1) ncf += __shfl_xor_sync(0xffffffffu, ncf, m);
    if ((threadIdx.x & 31) == 0 && ncf) atomicAdd(&s_ncf, ncf);
    __syncthreads();
    if (threadIdx.x == 0) ncell[(size_t)e * p.nt + k] = s_ncf;
}

template <typename T> __global__ void __launch_bounds__(256) fma_peak_kernel(T *out, int iters, T a, T b) {
    T x0 = T(threadIdx.x) * T(1e-3), x1 = x0 + T(1), x2 = x0 + T(2), x3 = x0 + T(3);
    T x4 = x0 + T(4), x5 = x0 + T(5), x6 = x0 + T(6), x7 = x0 + T(7);
    for (int i = 0; i < iters; i++) {
#pragma unroll
        for (int u = 0; u < 8; u++) {
            x0 = x0 * a + b; x1 = x1 * a + b; x2 = x2 * a + b; x3 = x3 * a + b;
            x4 = x4 * a + b; x5 = x5 * a + b; x6 = x6 * a + b; x7 = x7 * a + b;
        }
    }
    T s = x0 + x1 + x2 + x3 + x4 + x5 + x6 + x7;
    if (s == T(12345.678)) out[blockIdx.x * blockDim.x + threadIdx.x] = s;   // never true: keeps the chains live
}

template <typename T> int run_peak(int device, int iters, double *tflops, double *ms) {
    if (!tflops || iters < 1) return BSG_ERR_ARG;
    if (cudaSetDevice(device) != cudaSuccess) return BSG_ERR_CUDA;
    T *d = nullptr;
    const int blocks = 148 * 8, threads = 256;
    if (cudaMalloc(&d, sizeof(T) * blocks * threads) != cudaSuccess) return BSG_ERR_NOMEM;
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    fma_peak_kernel<T><<<blocks, threads>>>(d, iters / 8 + 1, T(0.999), T(0.001));   // warm-up
    float best = 1e30f;
    for (int rep = 0; rep < 5; rep++) {
        cudaEventRecord(e0);
        fma_peak_kernel<T><<<blocks, threads>>>(d, iters, T(0.999), T(0.001));
        cudaEventRecord(e1);
        cudaEventSynchronize(e1);
        float t = 0;
        cudaEventElapsedTime(&t, e0, e1);
        best = t < best ? t : best;
    }
    cudaError_t err = cudaGetLastError();
    cudaEventDestroy(e0); cudaEventDestroy(e1); cudaFree(d);
    if (err != cudaSuccess) return BSG_ERR_CUDA;
    const double flop = 2.0 * 64.0 * (double)iters * blocks * threads;
    *tflops = flop / (best * 1e-3) / 1e12;
    if (ms) *ms = best;
    return BSG_OK;
}

}  // namespace

extern "C" {

int bsg_ndmap(const bsg_ndmap_params *p, const float *rec, uint8_t *occ, uint8_t *conf, int32_t *ncell, void *stream) {
    if (!p || !rec || !occ || !conf || !ncell) return BSG_ERR_ARG;
    const long cells = (long)p->nx * p->ny * p->nz;
    if (p->num_envs < 1 || p->num_ac < 1 || p->num_ac > 65535 || p->nt < 1 || cells < 16 || (p->nx % 16) != 0) return BSG_ERR_ARG;
    const size_t smem = (size_t)cells * 3;                    // u16 counts + u8 plane sums
    if (smem > 227 * 1024) return BSG_ERR_UNSUPPORTED;
    const float icx = 1.0f / p->cell_x, icy = 1.0f / p->cell_y, icz = 1.0f / p->cell_z;
    dim3 grid(p->nt, p->num_envs);
    auto ispow2 = [](int v) { return v > 0 && (v & (v - 1)) == 0; };
    auto lg2 = [](int v) { int l = 0; while ((1 << l) < v) l++; return l; };
    const int gx = p->nx / 8, plane = p->nx * p->ny;
    if (ispow2(gx) && ispow2(p->ny) && gx <= 32 && (cells / 8) % 32 == 0) {
        if (cudaFuncSetAttribute(ndmap_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess) return BSG_ERR_CUDA;
        ndmap_kernel<true><<<grid, kNdThreads, smem, (cudaStream_t)stream>>>(*p, icx, icy, icz, lg2(gx), lg2(plane), rec, occ, conf, ncell);
    } else {
        if (cudaFuncSetAttribute(ndmap_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess) return BSG_ERR_CUDA;
        ndmap_kernel<false><<<grid, kNdThreads, smem, (cudaStream_t)stream>>>(*p, icx, icy, icz, 0, 0, rec, occ, conf, ncell);
    }
    return cudaGetLastError() == cudaSuccess ? BSG_OK : BSG_ERR_CUDA;
}

int bsg_peak_fp32(int device, int iters, double *tflops, double *ms) { return run_peak<float>(device, iters, tflops, ms); }
int bsg_peak_fp64(int device, int iters, double *tflops, double *ms) { return run_peak<double>(device, iters, tflops, ms); }

}  // extern "C"
