// bsg_ndmap.cu -- ND-map: 4D (t, z, y, x) spatial-temporal occupancy / conflict grid over straight-line
// look-ahead trajectories.  No reference implementation exists (SURVEY.md 0.3 tier C): the spec is this
// repo's own (DESIGN.md "ND-map"); the CPU restatement of the same spec is oracle/c/bsgo_ndmap.c.
//
// Outputs per (env, time slab): occ u8 [cells] = min(count, 255); conf = BITMASK (1 bit per cell, cell c <-> bit c & 31
// of word c >> 5) of the cells whose 3x3x3 neighbourhood (clipped at the faces) holds >= 2 aircraft; ncell = popcount.
// The kernel is HBM-write bound by design (1.125 B per cell), so everything else is kept off the per-cell path:
//  * persistent CTAs (two per SM) walk (env, chunk of time slabs) items; the env's aircraft stay in registers;
//  * counting: one shared-memory atomic per sample straight into the u8 slab image that is stored (byte lanes of
//    32-bit words; saturating CAS form when more than 255 aircraft could meet in a cell), plus one atomicOr into the
//    bit-planes A = [count >= 1] / B2 = [count >= 2] -- so no dense pass ever reads the counts;
//  * the 3x3x3 threshold is BIT-PARALLEL on those planes, 32 cells per operation: for a set S of cells keep
//    o(S) = [sum >= 1], t(S) = [sum >= 2]; disjoint sets combine as o = o1 | o2, t = t1 | t2 | (o1 & o2).  Three
//    separable stages (x: shifts inside a row of nx / 32 words, y: neighbouring rows, z: neighbouring planes), one
//    thread per grid row, ~1 instruction per cell in total (the byte-sum form it replaces spent ~8);
//  * the finished slab (u8 image + conflict bits) leaves as two TMA bulk stores (cp.async.bulk.global.shared::cta)
//    issued by one thread; images and bit stages are double-buffered so the stores of slab k overlap the counting of
//    slab k+1, and an image is cleaned SPARSELY (the cells its samples touched) once its store has read it.
// Also holds the FP32 / FP64 FMA-chain micro-benchmarks (the measured non-tensor roofline denominators).
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdlib>
#include "bsg_handle.h"

namespace {

constexpr int kNdThreads = 512;

__device__ __forceinline__ uint32_t nd_smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void bulk_store(void *gdst, const void *ssrc, uint32_t bytes) {
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(gdst), "r"(nd_smem_u32(ssrc)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait_read0() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait0() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }
__device__ __forceinline__ void fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

// a grid row of nx = 32 RW cells as RW words; cell x <-> bit x & 31 of word x >> 5
template <int RW> struct Row {
    uint32_t w[RW];
    __device__ __forceinline__ void load(const uint32_t *p) {
#pragma unroll
        for (int i = 0; i < RW; i++) w[i] = p[i];
    }
    __device__ __forceinline__ void store(uint32_t *p) const {
#pragma unroll
        for (int i = 0; i < RW; i++) p[i] = w[i];
    }
    __device__ __forceinline__ Row up() const {        // cell x takes the value of x - 1 (0 enters at the face)
        Row r;
#pragma unroll
        for (int i = 0; i < RW; i++) r.w[i] = (w[i] << 1) | (i > 0 ? w[i - 1] >> 31 : 0u);
        return r;
    }
    __device__ __forceinline__ Row down() const {      // cell x takes the value of x + 1
        Row r;
#pragma unroll
        for (int i = 0; i < RW; i++) r.w[i] = (w[i] >> 1) | (i < RW - 1 ? w[i + 1] << 31 : 0u);
        return r;
    }
    __device__ __forceinline__ Row operator|(const Row &o) const { Row r;
#pragma unroll
        for (int i = 0; i < RW; i++) r.w[i] = w[i] | o.w[i];
        return r; }
    __device__ __forceinline__ Row operator&(const Row &o) const { Row r;
#pragma unroll
        for (int i = 0; i < RW; i++) r.w[i] = w[i] & o.w[i];
        return r; }
};
// (o, t) of the union of three disjoint cell sets
template <int RW> __device__ __forceinline__ void combine3(const Row<RW> &oa, const Row<RW> &ta, const Row<RW> &ob, const Row<RW> &tb,
                                                           const Row<RW> &oc, const Row<RW> &tc, Row<RW> &o, Row<RW> &t) {
    o = oa | ob | oc;
    t = ta | tb | tc | (oa & ob) | ((oa | ob) & oc);
}

// RW = nx / 32.  REGS: num_ac <= kNdThreads, every thread keeps its aircraft in registers for the whole item.
// SAT: more than 255 aircraft exist, the byte counters saturate through a CAS loop.
template <int RW, bool REGS, bool SAT>
__global__ void __launch_bounds__(kNdThreads, 2) ndmap_kernel(const bsg_ndmap_params p, float icx, float icy, float icz, int ks,
                                                              const float *__restrict__ rec, uint8_t *__restrict__ occ,
                                                              uint32_t *__restrict__ conf, int32_t *__restrict__ ncell) {
    extern __shared__ __align__(128) unsigned char smem[];
    __shared__ int s_ncf;
    const int nx = p.nx, ny = p.ny, nz = p.nz, n = p.num_ac, nt = p.nt;
    const int cells = nx * ny * nz, nw = cells >> 5, rows = ny * nz;
    uint8_t *img[2] = { smem, smem + cells };
    uint32_t *A = reinterpret_cast<uint32_t *>(smem + 2 * (size_t)cells), *B2 = A + nw, *O1 = B2 + nw, *T1 = O1 + nw, *O2 = T1 + nw,
             *T2 = O2 + nw;
    uint32_t *CF[2] = { T2 + nw, T2 + 2 * nw };
    const int tid = threadIdx.x;
    for (int c = tid; c < (3 * cells) >> 4; c += kNdThreads) reinterpret_cast<uint4 *>(smem)[c] = make_uint4(0u, 0u, 0u, 0u);
    if (tid == 0) s_ncf = 0;
    const int chunks = (nt + ks - 1) / ks, nitems = p.num_envs * chunks;
    const int y_tid = tid % ny, z_tid = tid / ny;         // grid row of this thread's first pass (usually its only one)
    int par = 0;
    for (int item = blockIdx.x; item < nitems; item += gridDim.x) {
        const int e = item / chunks, k0 = (item - e * chunks) * ks, k1 = min(k0 + ks, nt);
        const float *r = rec + (size_t)e * 6 * n;
        float ax = 0, ay = 0, az = 0, au = 0, av = 0, aw = 0;
        if (REGS && tid < n) { ax = r[tid]; ay = r[n + tid]; az = r[2 * n + tid]; au = r[3 * n + tid]; av = r[4 * n + tid]; aw = r[5 * n + tid]; }
        // cell of aircraft a in slab k, or -1 (outside the grid): the spec's fp32 sequence (oracle/c/bsgo_ndmap.c)
        auto cell_of = [&](int a, int k) -> int {
            const float t = (float)(k + 1) * p.dt;
            float x, y, z;
            if constexpr (REGS) { x = fmaf(au, t, ax); y = fmaf(av, t, ay); z = fmaf(aw, t, az); }
            else { x = fmaf(r[3 * n + a], t, r[a]); y = fmaf(r[4 * n + a], t, r[n + a]); z = fmaf(r[5 * n + a], t, r[2 * n + a]); }
            const int ix = (int)floorf((x - p.x0) * icx), iy = (int)floorf((y - p.y0) * icy), iz = (int)floorf((z - p.z0) * icz);
            if (ix < 0 || ix >= nx || iy < 0 || iy >= ny || iz < 0 || iz >= nz) return -1;
            return (iz * ny + iy) * nx + ix;
        };
        // sparse cleaning: the cells slab k's samples touched.  REGS: one sample per thread, its cell of the last two
        // slabs is remembered (cprev[parity]) instead of being recomputed
        int cprev[2] = { -1, -1 };
        auto clean = [&](int k, uint8_t *im, int pq) {
            if (REGS) { if (cprev[pq] >= 0) im[cprev[pq]] = 0; }                  // (REGS is a template constant)
            else { for (int a = tid; a < n; a += kNdThreads) { const int c = cell_of(a, k); if (c >= 0) im[c] = 0; } }
        };
        __syncthreads();                                 // (first item: the zero fill; later: the previous item's cleaning)
        for (int k = k0; k < k1; k++) {
            uint8_t *im = img[par];
            // ---- count: one atomic per sample into the byte lanes of the slab image + the [>= 1] / [>= 2] bit-planes
            if constexpr (REGS) cprev[par] = -1;
            for (int a = tid; a < n; a += kNdThreads) {
                const int c = cell_of(a, k);
                if (c < 0) continue;
                if constexpr (REGS) cprev[par] = c;
                uint32_t *wp = reinterpret_cast<uint32_t *>(im) + (c >> 2);
                const int sh = (c & 3) << 3;
                unsigned lane;
                if constexpr (!SAT) lane = (atomicAdd(wp, 1u << sh) >> sh) & 255u;
                else {
                    unsigned old = *wp;
                    for (;;) {
                        lane = (old >> sh) & 255u;
                        if (lane == 255u) break;
                        const unsigned prev = atomicCAS(wp, old, old + (1u << sh));
                        if (prev == old) break;
                        old = prev;
                    }
                }
                if (lane == 0u) atomicOr(&A[c >> 5], 1u << (c & 31));
                else if (lane == 1u) atomicOr(&B2[c >> 5], 1u << (c & 31));
            }
            __syncthreads();
            // ---- x stage: three cells of a row -> (o, t); the planes are cleared for the next slab
            for (int rr = tid; rr < rows; rr += kNdThreads) {
                Row<RW> a1, b2;
                a1.load(A + rr * RW); b2.load(B2 + rr * RW);
                const Row<RW> au1 = a1.up(), ad1 = a1.down();
                const Row<RW> o = a1 | au1 | ad1;
                const Row<RW> t = b2 | b2.up() | b2.down() | (a1 & au1) | ((a1 | au1) & ad1);
                o.store(O1 + rr * RW); t.store(T1 + rr * RW);
                Row<RW> zero = {};
                zero.store(A + rr * RW); zero.store(B2 + rr * RW);
            }
            __syncthreads();
            // ---- y stage (rows y - 1, y, y + 1 of one plane); the image of the previous slab is cleaned meanwhile
            for (int rr = tid; rr < rows; rr += kNdThreads) {
                const int y = rr == tid ? y_tid : rr % ny;
                Row<RW> oa = {}, ta = {}, ob, tb, oc = {}, tc = {}, o, t;
                ob.load(O1 + rr * RW); tb.load(T1 + rr * RW);
                if (y > 0) { oa.load(O1 + (rr - 1) * RW); ta.load(T1 + (rr - 1) * RW); }
                if (y < ny - 1) { oc.load(O1 + (rr + 1) * RW); tc.load(T1 + (rr + 1) * RW); }
                combine3<RW>(oa, ta, ob, tb, oc, tc, o, t);
                o.store(O2 + rr * RW); t.store(T2 + rr * RW);
            }
            // the previous slab's stores have had the count, x and y stages to read their image / bit stage out of shared
            // memory; from here on the image may be cleaned (z stage below) and the bit stage rewritten (next slab)
            if (tid == 0) bulk_wait_read0();
            __syncthreads();
            // ---- z stage (planes z - 1, z, z + 1) -> conflict bits of the slab
            int ncf = 0;
            uint32_t *cf = CF[par];
            if (k > k0) clean(k - 1, img[par ^ 1], par ^ 1);
            for (int rr = tid; rr < rows; rr += kNdThreads) {
                const int z = rr == tid ? z_tid : rr / ny;
                Row<RW> oa = {}, ta = {}, ob, tb, oc = {}, tc = {}, o, t;
                ob.load(O2 + rr * RW); tb.load(T2 + rr * RW);
                if (z > 0) { oa.load(O2 + (rr - ny) * RW); ta.load(T2 + (rr - ny) * RW); }
                if (z < nz - 1) { oc.load(O2 + (rr + ny) * RW); tc.load(T2 + (rr + ny) * RW); }
                combine3<RW>(oa, ta, ob, tb, oc, tc, o, t);
                t.store(cf + rr * RW);
#pragma unroll
                for (int i = 0; i < RW; i++) ncf += __popc(t.w[i]);
            }
            ncf = __reduce_add_sync(0xffffffffu, ncf);       // redux.sync
            if ((tid & 31) == 0 && ncf) atomicAdd(&s_ncf, ncf);
            fence_async_smem();                          // generic-proxy writes (atomics, stores) before the TMA reads them
            __syncthreads();
            if (tid == 0) {
                const size_t slab = (size_t)e * nt + k;
                bulk_store(occ + slab * cells, im, (uint32_t)cells);
                bulk_store(conf + slab * nw, cf, (uint32_t)nw * 4u);
                bulk_commit();
                ncell[slab] = s_ncf;
                s_ncf = 0;
            }
            par ^= 1;
        }
        // the item's last slab: wait for its store, clean its image (the records change with the next item)
        if (tid == 0) bulk_wait_read0();
        __syncthreads();
        clean(k1 - 1, img[par ^ 1], par ^ 1);
    }
    if (tid == 0) bulk_wait0();
}

// conflict bits -> one byte per cell (for callers that want the dense 0/1 form)
__global__ void ndmap_unpack_kernel(size_t nwords, const uint32_t *__restrict__ bits, uint8_t *__restrict__ bytes) {
    const size_t g = (size_t)blockIdx.x * blockDim.x + threadIdx.x;          // 8 cells per thread
    if (g >= nwords * 4) return;
    const uint32_t b = (bits[g >> 2] >> ((g & 3) << 3)) & 0xffu;
    uint2 o;
    o.x = (b & 1u) | ((b & 2u) << 7) | ((b & 4u) << 14) | ((b & 8u) << 21);
    o.y = ((b >> 4) & 1u) | ((b & 32u) << 3) | ((b & 64u) << 10) | ((b & 128u) << 17);
    reinterpret_cast<uint2 *>(bytes)[g] = o;
}

// straight-line records (x, y, z, u, v, w) of the LIVE env state, flat-earth offsets from (lat0, lon0); dead slots are
// parked far outside any grid
template <typename R>
__global__ void ndmap_rec_kernel(int E, int N, const double *__restrict__ pos, const R *__restrict__ st, const int *__restrict__ nac,
                                 int ext, double lat0, double lon0, double m_lat, double m_lon, float *__restrict__ rec) {
    const size_t EN = (size_t)E * N, k = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= EN) return;
    const int e = (int)(k / N), i = (int)(k - (size_t)e * N);
    float *o = rec + (size_t)e * 6 * N + i;
    if (i >= nac[e]) { o[0] = -3.0e38f; o[N] = -3.0e38f; o[2 * N] = -3.0e38f; o[3 * N] = 0.0f; o[4 * N] = 0.0f; o[5 * N] = 0.0f; return; }
    const double lat = pos[k], lon = pos[EN + k];
    o[0] = (float)((lon - lon0) * m_lon);
    o[N] = (float)((lat - lat0) * m_lat);
    o[2 * N] = (float)st[(size_t)bsg::S_ALT * EN + k];
    double gsn, gse;
    if (ext) { gsn = (double)st[(size_t)bsg::S_GSN * EN + k]; gse = (double)st[(size_t)bsg::S_GSE * EN + k]; }
    else {
        const double tas = (double)st[(size_t)bsg::S_TAS * EN + k], hdg = (double)st[(size_t)bsg::S_HDG * EN + k] * bsg::kD2R;
        gsn = tas * cos(hdg); gse = tas * sin(hdg);
    }
    o[3 * N] = (float)gse; o[4 * N] = (float)gsn; o[5 * N] = (float)st[(size_t)bsg::S_VS * EN + k];
}

template <int RW>
cudaError_t launch_ndmap(const bsg_ndmap_params &p, const float *rec, uint8_t *occ, uint32_t *conf, int32_t *ncell, cudaStream_t s) {
    const size_t cells = (size_t)p.nx * p.ny * p.nz, smem = 3 * cells;
    const float icx = 1.0f / p.cell_x, icy = 1.0f / p.cell_y, icz = 1.0f / p.cell_z;
    int dev = 0, sms = 148;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    const int slots = 2 * sms;
    int ks = p.nt;                                       // slabs per work item: enough items for an even deal
    while (ks > 1 && (long)p.num_envs * ((p.nt + ks - 1) / ks) < 16L * slots) ks = (ks + 1) / 2;
    const long nitems = (long)p.num_envs * ((p.nt + ks - 1) / ks);
    const int grid = (int)(nitems < slots ? nitems : slots);
    const bool regs = p.num_ac <= kNdThreads, sat = p.num_ac > 255;
#define ND_LAUNCH(REGS, SAT)                                                                                              \
    do {                                                                                                                  \
        cudaError_t e_ = cudaFuncSetAttribute(ndmap_kernel<RW, REGS, SAT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem); \
        if (e_ != cudaSuccess) return e_;                                                                                 \
        ndmap_kernel<RW, REGS, SAT><<<grid, kNdThreads, smem, s>>>(p, icx, icy, icz, ks, rec, occ, conf, ncell);           \
    } while (0)
    if (regs) { if (sat) ND_LAUNCH(true, true); else ND_LAUNCH(true, false); }
    else { if (sat) ND_LAUNCH(false, true); else ND_LAUNCH(false, false); }
#undef ND_LAUNCH
    return cudaGetLastError();
}

static int ndmap_check(const bsg_ndmap_params *p) {
    if (!p) return BSG_ERR_ARG;
    const long cells = (long)p->nx * p->ny * p->nz;
    if (p->num_envs < 1 || p->num_ac < 1 || p->nt < 1 || p->nx < 32 || p->ny < 1 || p->nz < 1 || !(p->cell_x > 0) || !(p->cell_y > 0) ||
        !(p->cell_z > 0))
        return BSG_ERR_ARG;
    // rows of whole 32-bit words (nx in {32, 64, 128}), stages 16-byte aligned for the bulk stores, grid in shared memory
    if ((p->nx != 32 && p->nx != 64 && p->nx != 128) || (cells % 128) != 0 || 3 * cells > 227 * 1024) return BSG_ERR_UNSUPPORTED;
    return BSG_OK;
}

static int ndmap_run(const bsg_ndmap_params *p, const float *rec, uint8_t *occ, uint8_t *conf, int32_t *ncell, cudaStream_t s) {
    cudaError_t e;
    uint32_t *cb = reinterpret_cast<uint32_t *>(conf);
    if (p->nx == 32) e = launch_ndmap<1>(*p, rec, occ, cb, ncell, s);
    else if (p->nx == 64) e = launch_ndmap<2>(*p, rec, occ, cb, ncell, s);
    else e = launch_ndmap<4>(*p, rec, occ, cb, ncell, s);
    return e == cudaSuccess ? BSG_OK : BSG_ERR_CUDA;
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

int bsg_ndmap(const bsg_ndmap_params *p, const float *rec, uint8_t *occ, uint8_t *conf_bits, int32_t *ncell, void *stream) {
    if (!p || !rec || !occ || !conf_bits || !ncell) return BSG_ERR_ARG;
    const int rc = ndmap_check(p);
    if (rc != BSG_OK) return rc;
    return ndmap_run(p, rec, occ, conf_bits, ncell, (cudaStream_t)stream);
}

int bsg_ndmap_unpack(const bsg_ndmap_params *p, const uint8_t *conf_bits, uint8_t *conf_bytes, void *stream) {
    if (!p || !conf_bits || !conf_bytes || ndmap_check(p) != BSG_OK) return BSG_ERR_ARG;
    const size_t nwords = (size_t)p->num_envs * p->nt * ((size_t)p->nx * p->ny * p->nz / 32);
    ndmap_unpack_kernel<<<(unsigned)((nwords * 4 + 255) / 256), 256, 0, (cudaStream_t)stream>>>(nwords, reinterpret_cast<const uint32_t *>(conf_bits),
                                                                                                conf_bytes);
    return cudaGetLastError() == cudaSuccess ? BSG_OK : BSG_ERR_CUDA;
}

int bsg_ndmap_state(bsg_handle *h, const bsg_ndmap_params *p, const bsg_ndmap_geo *geo, uint8_t *occ, uint8_t *conf_bits,
                    int32_t *ncell, float *rec_out, void *stream) {
    if (!h || !p || !geo || !occ || !conf_bits || !ncell) return BSG_ERR_ARG;
    const bsg_config &g = h->cfg;
    if (p->num_envs != g.num_envs || p->num_ac != g.max_ac) { h->err = "bsg_ndmap_state: num_envs / num_ac must be the handle's"; return BSG_ERR_ARG; }
    const int rc = ndmap_check(p);
    if (rc != BSG_OK) return rc;
    CK(h, cudaSetDevice(g.device));
    const size_t EN = (size_t)g.num_envs * g.max_ac;
    float *rec = rec_out;
    if (!rec) {
        if (!h->nd_rec) CK(h, cudaMalloc(&h->nd_rec, 6 * EN * sizeof(float)));
        rec = h->nd_rec;
    }
    cudaStream_t s = (cudaStream_t)stream;
    const int ext = (g.env_kind == BSG_KIND_MERGE || h->wind.n > 0 || h->noise_sigma > 0.0) ? 1 : 0;   // the EXT kernels keep gsn / gse
    const unsigned gb = (unsigned)((EN + 255) / 256);
    if (g.precision == BSG_FP64)
        ndmap_rec_kernel<double><<<gb, 256, 0, s>>>(g.num_envs, g.max_ac, h->pos, (const double *)h->st, h->nac, ext, geo->lat0, geo->lon0,
                                                    geo->m_per_deg_lat, geo->m_per_deg_lon, rec);
    else
        ndmap_rec_kernel<float><<<gb, 256, 0, s>>>(g.num_envs, g.max_ac, h->pos, (const float *)h->st, h->nac, ext, geo->lat0, geo->lon0,
                                                   geo->m_per_deg_lat, geo->m_per_deg_lon, rec);
    h->launches += 2;
    CK(h, cudaGetLastError());
    // the record kernel is the only reader of the handle's state: a host-buffer step (the handle's own stream) that is
    // begun after this call must not overwrite the state before it has been read
    if (h->hs && s != h->hs && h->nd_ev) {
        CK(h, cudaEventRecord(h->nd_ev, s));
        CK(h, cudaStreamWaitEvent(h->hs, h->nd_ev, 0));
    }
    return ndmap_run(p, rec, occ, conf_bits, ncell, s);
}

int bsg_peak_fp32(int device, int iters, double *tflops, double *ms) { return run_peak<float>(device, iters, tflops, ms); }
int bsg_peak_fp64(int device, int iters, double *tflops, double *ms) { return run_peak<double>(device, iters, tflops, ms); }

}  // extern "C"
