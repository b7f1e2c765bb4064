// bsg_sweep.cu -- stand-alone all-pairs state-based CD sweep for large airspaces (giant config:
// 16 envs x 10 000 aircraft; pair matrix row-block [row0,row1) per GPU) and the record packer.
// What: StateBased.detect row outputs.  How: every CTA owns RPB ownship rows, each row
// is split over JS adjacent lanes (j = q, q+JS, ...), the intruder columns stream through shared
// memory in 32 KB tiles moved by TMA bulk copies (cp.async.bulk + mbarrier, double-buffered), the
// row outputs are combined with warp-shuffle sum / max / lexicographic-min reductions.
#include <cuda_runtime.h>
#include <cstdlib>
#include <cstdint>
#include "bsg_pair.cuh"

using namespace bsg;

namespace {

constexpr int kTile = 1024;            // records per j-tile: 32 KB
constexpr int kThreads = 256;

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t *bar, int count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t *bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity) {
    asm volatile(
        "{\n .reg .pred p;\n WAIT_%=:\n mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n @p bra DONE_%=;\n"
        " bra WAIT_%=;\n DONE_%=:\n}" ::"r"(smem_u32(bar)), "r"(parity) : "memory");
}
__device__ __forceinline__ void tma_bulk_g2s(void *dst, const void *src, uint32_t bytes, uint64_t *bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                     smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}

template <int JS>
__global__ void __launch_bounds__(kThreads) cd_sweep_kernel(const bsg_cd_params p, const RecFast *__restrict__ rec,
                                                            uint8_t *inconf, int32_t *nconf, int32_t *nlos,
                                                            int32_t *near_idx, float *near_tcpa, float *near_dcpa,
                                                            float *near_tin, float *tcpamax) {
    extern __shared__ __align__(128) unsigned char smem[];
    RecFast *tiles = reinterpret_cast<RecFast *>(smem);                     // [2][kTile]
    uint64_t *bar = reinterpret_cast<uint64_t *>(smem + 2 * kTile * sizeof(RecFast));
    constexpr int RPB = kThreads / JS;
    const int N = p.num_ac, e = blockIdx.y;
    const int r = threadIdx.x / JS, q = threadIdx.x % JS;
    const int i = p.row0 + blockIdx.x * RPB + r;
    const bool rowok = i < p.row1;
    const RecFast *base = rec + (size_t)e * N;
    const int ntiles = (N + kTile - 1) / kTile;
    if (threadIdx.x == 0) {
        mbar_init(&bar[0], 1); mbar_init(&bar[1], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        uint32_t bytes = (uint32_t)(min(kTile, N) * sizeof(RecFast));
        mbar_expect_tx(&bar[0], bytes);
        tma_bulk_g2s(tiles, base, bytes, &bar[0]);
    }
    RecFast me = {};
    if (rowok) {
        const float4 *m = reinterpret_cast<const float4 *>(base + i);
        float4 a = __ldg(m), b = __ldg(m + 1);
        me = { a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w };
    }
    CdParamsF P;
    P.R2 = p.rpz_m * p.rpz_m; P.hpz = p.hpz_m; P.dtlook = p.dtlookahead; P.rpz = p.rpz_m; P.rpz_m = 0; P.hpz_m = 0;
    RowAcc<float> acc;
    for (int t = 0; t < ntiles; t++) {
        const int buf = t & 1;
        if (threadIdx.x == 0 && t + 1 < ntiles) {
            const int n1 = min(kTile, N - (t + 1) * kTile);
            uint32_t bytes = (uint32_t)(n1 * sizeof(RecFast));
            mbar_expect_tx(&bar[buf ^ 1], bytes);
            tma_bulk_g2s(tiles + (size_t)(buf ^ 1) * kTile, base + (size_t)(t + 1) * kTile, bytes, &bar[buf ^ 1]);
        }
        mbar_wait(&bar[buf], (t >> 1) & 1);
        const int nt = min(kTile, N - t * kTile), j0 = t * kTile;
        const RecFast *tb = tiles + (size_t)buf * kTile;
        // 4 pairs per trip, flags first and one rare branch behind them (conflicts are rare: the row reductions
        // stay off the hot path and the four dependency chains interleave)
        int jj = q;
        for (; jj + 3 * JS < nt; jj += 4 * JS) {
            PairCore pc[4];
            bool sw[4], any = false;
#pragma unroll
            for (int u = 0; u < 4; u++) {
                const float4 *s = reinterpret_cast<const float4 *>(tb + jj + u * JS);
                const float4 a = s[0], b = s[1];
                const RecFast o = { a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w };
                sw[u] = pair_fast_core(me, o, P, (j0 + jj + u * JS) != i, pc[u]);
                any |= sw[u];
            }
            if (any) {
#pragma unroll
                for (int u = 0; u < 4; u++)
                    if (sw[u]) acc.add_conflict(j0 + jj + u * JS, pc[u].tcpa, pc[u].dcpa2, pc[u].tin,
                                                (pc[u].dist2 < P.R2) && (fabsf(pc[u].dalt) < P.hpz));
            }
        }
        for (; jj < nt; jj += JS) {
            const float4 *s = reinterpret_cast<const float4 *>(tb + jj);
            const float4 a = s[0], b = s[1];
            const RecFast o = { a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w };
            PairCore pc;
            if (pair_fast_core(me, o, P, (j0 + jj) != i, pc))
                acc.add_conflict(j0 + jj, pc.tcpa, pc.dcpa2, pc.tin, (pc.dist2 < P.R2) && (fabsf(pc.dalt) < P.hpz));
        }
        __syncthreads();      // everyone is done with tiles[buf] before it is refilled two iterations later
    }
    // combine the JS partial rows (adjacent lanes) -- warp-shuffle sum / max / lexicographic (tcpa, j) min
    float ntc = acc.near_tcpa;          // +inf while the partial row has no conflict
#pragma unroll
    for (int m = 1; m < JS; m <<= 1) {
        acc.nconf += __shfl_xor_sync(0xffffffffu, acc.nconf, m);
        acc.nlos += __shfl_xor_sync(0xffffffffu, acc.nlos, m);
        acc.tcpamax = fmaxf(acc.tcpamax, __shfl_xor_sync(0xffffffffu, acc.tcpamax, m));
        const float otc = __shfl_xor_sync(0xffffffffu, ntc, m);
        const int oid = __shfl_xor_sync(0xffffffffu, acc.near_idx, m);
        const float od2 = __shfl_xor_sync(0xffffffffu, acc.near_dcpa2, m);
        const float oti = __shfl_xor_sync(0xffffffffu, acc.near_tin, m);
        const bool take = oid >= 0 && (acc.near_idx < 0 || otc < ntc || (otc == ntc && oid < acc.near_idx));
        if (take) { ntc = otc; acc.near_idx = oid; acc.near_dcpa2 = od2; acc.near_tin = oti; }
    }
    if (rowok && q == 0) {
        const size_t k = (size_t)e * N + i;
        if (inconf) inconf[k] = acc.nconf > 0;
        if (nconf) nconf[k] = acc.nconf;
        if (nlos) nlos[k] = acc.nlos;
        if (near_idx) near_idx[k] = acc.near_idx;
        if (near_tcpa) near_tcpa[k] = acc.near_idx < 0 ? 0.0f : ntc;
        if (near_dcpa) near_dcpa[k] = sqrtf(acc.near_dcpa2);
        if (near_tin) near_tin[k] = acc.near_tin;
        if (tcpamax) tcpamax[k] = acc.tcpamax;
    }
}

// ---------------------------------------------------------------------------------------------------------------
// Symmetric sweep: every UNORDERED pair is evaluated once.  The fp32 pair form is exactly symmetric under i <-> j
// (bsg_pair.cuh), so the ordered N x N matrix it stands for is bit-identical to the full sweep's.  The pair matrix
// of an env is cut into kSymT x kSymT tiles; a CTA owns one tile (bi <= bj) of the upper triangle, thread r holds
// ownship row bi*T + r in registers and the tile's columns stream from shared memory as broadcast reads.  Conflicts
// are rare: a conflict updates the row accumulators (registers) and the column accumulators (shared-memory
// atomics); both are flushed with global atomics that are order-independent -- integer add / max / 64-bit min on
// (tCPA as an order-preserving int32, intruder index) -- so the result is deterministic.  dCPA / tin of the winner
// are recomputed by the finalize kernel (pure functions of the pair).  tile_rank/tile_stride deal the tiles to
// GPUs round-robin; the partial outputs are then combined with one all-reduce (sum, max, min) before finalize.
constexpr int kSymT = 256;

__device__ __forceinline__ long long sym_key(float tcpa, int j) {
    int b = __float_as_int(tcpa);
    b ^= (b >> 31) & 0x7fffffff;                         // signed-int order == float order
    return ((long long)b << 32) | (unsigned)j;
}
__device__ __forceinline__ float sym_key_tcpa(long long key) {
    int b = (int)(key >> 32);
    b ^= (b >> 31) & 0x7fffffff;
    return __int_as_float(b);
}

__global__ void cd_sym_init_kernel(size_t n, int32_t *nconf, int32_t *nlos, int32_t *tcpamax_bits, long long *key) {
    const size_t k = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= n) return;
    nconf[k] = 0; nlos[k] = 0; tcpamax_bits[k] = 0; key[k] = 0x7fffffffffffffffLL;
}

// Peer tables: aircraft j of env e lives on rank g = j / rows at local slot j - g*rows; records are read from, and
// row outputs flushed to, the owner's buffers (plain device pointers at G = 1, NVLink peer mappings of symmetric
// memory at G > 1: the all-gather of the state and the reduce of the outputs become loads / atomics over NVLink
// inside this one kernel).  System-scope atomics: the targets may live on another GPU.
struct SymPeers {
    int G, rows;
    const RecFast *rec[BSG_MAX_PEERS];
    int32_t *nconf[BSG_MAX_PEERS], *nlos[BSG_MAX_PEERS], *tmax[BSG_MAX_PEERS];
    long long *key[BSG_MAX_PEERS];
};
__device__ __forceinline__ void sym_locate(const SymPeers &Q, int e, int j, int &g, size_t &k) {
    g = Q.G == 1 ? 0 : j / Q.rows;
    k = (size_t)e * Q.rows + (j - g * Q.rows);
}

// `cs` column slices per tile (1, 2 or 4): a CTA sweeps its 256 ownship rows against kSymT / cs columns of the tile, so
// that the launch has enough CTAs for an even number of waves on every rank (the row / column accumulators are
// combined with order-independent atomics: the result does not depend on the slicing).
__global__ void __launch_bounds__(kSymT) cd_sym_sweep_kernel(const bsg_cd_params p, const SymPeers Q, int tile_rank,
                                                            int tile_stride, int cs) {
    __shared__ float4 tA[kSymT], tB[kSymT];
    __shared__ int c_nconf[kSymT], c_nlos[kSymT], c_tmax[kSymT];
    __shared__ long long c_key[kSymT];
    const int N = p.num_ac, e = blockIdx.y, nb = (N + kSymT - 1) / kSymT;
    const int T = blockIdx.x * tile_stride + tile_rank;    // (tile of the upper triangle, row-major) * cs + column slice
    int t = T / cs;
    const int sl = T - t * cs, cw = kSymT / cs;
    if (t >= nb * (nb + 1) / 2) return;
    int bi = 0;
    while (t >= nb - bi) { t -= nb - bi; bi++; }
    const int bj = bi + t, r = threadIdx.x;
    const int i = bi * kSymT + r, j0 = bj * kSymT + sl * cw, nt = min(cw, N - j0);
    if (nt <= 0) return;
    const bool rowok = i < N;
    int g; size_t kk;
    if (r < nt) {
        sym_locate(Q, e, j0 + r, g, kk);
        const float4 *m = reinterpret_cast<const float4 *>(Q.rec[g] + kk);
        tA[r] = m[0]; tB[r] = m[1];
    }
    c_nconf[r] = 0; c_nlos[r] = 0; c_tmax[r] = 0; c_key[r] = 0x7fffffffffffffffLL;
    RecFast me = {};
    if (rowok) {
        sym_locate(Q, e, i, g, kk);
        const float4 *m = reinterpret_cast<const float4 *>(Q.rec[g] + kk);
        const float4 a = m[0], b = m[1];
        me = { a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w };
    }
    CdParamsF P;
    P.R2 = p.rpz_m * p.rpz_m; P.hpz = p.hpz_m; P.dtlook = p.dtlookahead; P.rpz = p.rpz_m; P.rpz_m = 0; P.hpz_m = 0;
    __syncthreads();
    RowAcc<float> acc;
    const bool diag = bi == bj;
    auto hit = [&](int jj, const PairCore &pc) {             // rare
        const bool los = (pc.dist2 < P.R2) && (fabsf(pc.dalt) < P.hpz);
        acc.add_conflict(j0 + jj, pc.tcpa, pc.dcpa2, pc.tin, los);
        atomicAdd(&c_nconf[jj], 1);
        if (los) atomicAdd(&c_nlos[jj], 1);
        atomicMax(&c_tmax[jj], __float_as_int(fmaxf(pc.tcpa, 0.0f)));
        atomicMin(&c_key[jj], sym_key(pc.tcpa, i));
    };
    if (rowok) {
        int jj = 0;
        for (; jj + 3 < nt; jj += 4) {
            PairCore pc[4];
            bool sw[4], any = false;
#pragma unroll
            for (int u = 0; u < 4; u++) {
                const float4 a = tA[jj + u], b = tB[jj + u];
                const RecFast o = { a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w };
                sw[u] = pair_fast_core(me, o, P, !diag || (j0 + jj + u > i), pc[u]);
                any |= sw[u];
            }
            if (any) {
#pragma unroll
                for (int u = 0; u < 4; u++) if (sw[u]) hit(jj + u, pc[u]);
            }
        }
        for (; jj < nt; jj++) {
            const float4 a = tA[jj], b = tB[jj];
            const RecFast o = { a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w };
            PairCore pc;
            if (pair_fast_core(me, o, P, !diag || (j0 + jj > i), pc)) hit(jj, pc);
        }
    }
    __syncthreads();
    if (rowok && acc.nconf > 0) {                            // rows of this tile
        sym_locate(Q, e, i, g, kk);
        atomicAdd_system(&Q.nconf[g][kk], acc.nconf);
        if (acc.nlos) atomicAdd_system(&Q.nlos[g][kk], acc.nlos);
        atomicMax_system(&Q.tmax[g][kk], __float_as_int(acc.tcpamax));
        atomicMin_system(&Q.key[g][kk], sym_key(acc.near_tcpa, acc.near_idx));
    }
    if (r < nt && c_nconf[r] > 0) {                          // columns of this tile
        sym_locate(Q, e, j0 + r, g, kk);
        atomicAdd_system(&Q.nconf[g][kk], c_nconf[r]);
        if (c_nlos[r]) atomicAdd_system(&Q.nlos[g][kk], c_nlos[r]);
        atomicMax_system(&Q.tmax[g][kk], c_tmax[r]);
        atomicMin_system(&Q.key[g][kk], c_key[r]);
    }
}

__global__ void cd_sym_finalize_kernel(const bsg_cd_params p, const SymPeers Q, int rank, uint8_t *inconf,
                                       int32_t *near_idx, float *near_tcpa, float *near_dcpa, float *near_tin) {
    // one thread per OWN row (e, il): outputs are [E][rows] of this rank
    const size_t k = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= (size_t)p.num_envs * Q.rows) return;
    const long long ky = Q.key[rank][k];
    const int nc = Q.nconf[rank][k];
    int j = -1;
    float tc = 0.0f, dc = 0.0f, ti = 0.0f;
    if (nc > 0) {
        j = (int)(unsigned)(ky & 0xffffffffLL);
        tc = sym_key_tcpa(ky);
        const int e = (int)(k / Q.rows);
        int g; size_t kj;
        sym_locate(Q, e, j, g, kj);
        const float4 *m = reinterpret_cast<const float4 *>(Q.rec[rank] + k), *n = reinterpret_cast<const float4 *>(Q.rec[g] + kj);
        const float4 a = m[0], b = m[1], c = n[0], d = n[1];
        const RecFast me = { a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w }, o = { c.x, c.y, c.z, c.w, d.x, d.y, d.z, d.w };
        CdParamsF P;
        P.R2 = p.rpz_m * p.rpz_m; P.hpz = p.hpz_m; P.dtlook = p.dtlookahead; P.rpz = p.rpz_m; P.rpz_m = 0; P.hpz_m = 0;
        PairCore pc;
        pair_fast_core(me, o, P, true, pc);
        dc = sqrtf(pc.dcpa2); ti = pc.tin;
    }
    if (inconf) inconf[k] = nc > 0;
    if (near_idx) near_idx[k] = j;
    if (near_tcpa) near_tcpa[k] = tc;
    if (near_dcpa) near_dcpa[k] = dc;
    if (near_tin) near_tin[k] = ti;
}

// INIT: the row accumulators of the symmetric sweep are reset by the same launch (one kernel fewer in a peer-sweep step)
template <bool INIT>
__global__ void cd_pack_kernel(int E, int N, const double *lat, const double *lon, const float *alt, const float *trk,
                               const float *gs, const float *vs, double lat0, double lon0, RecFast *rec,
                               int32_t *nconf = nullptr, int32_t *nlos = nullptr, int32_t *tcpamax_bits = nullptr, long long *key = nullptr) {
    size_t k = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= (size_t)E * N) return;
    if (INIT) { nconf[k] = 0; nlos[k] = 0; tcpamax_bits[k] = 0; key[k] = 0x7fffffffffffffffLL; }
    RecFast r;
    r.x = (float)((lon[k] - lon0) * (kD2R * kRe));
    r.y = (float)((lat[k] - lat0) * (kD2R * kRe));
    r.alt = alt[k];
    float s, c;
    sincosf(trk[k] * float(kD2R), &s, &c);
    r.u = gs[k] * s; r.v = gs[k] * c; r.vs = vs[k];
    sincosf((float)(lat[k] * (kD2R * 0.5)), &r.sa, &r.ca);
    float4 *o = reinterpret_cast<float4 *>(rec + k);
    o[0] = make_float4(r.x, r.y, r.alt, r.u);
    o[1] = make_float4(r.v, r.vs, r.ca, r.sa);
}

template <int JS>
cudaError_t launch_sweep(const bsg_cd_params *p, const float *rec, uint8_t *inconf, int32_t *nconf, int32_t *nlos,
                         int32_t *near_idx, float *near_tcpa, float *near_dcpa, float *near_tin, float *tcpamax,
                         cudaStream_t s) {
    const size_t smem = 2 * kTile * sizeof(RecFast) + 64;
    cudaError_t e = cudaFuncSetAttribute(cd_sweep_kernel<JS>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    const int rows = p->row1 - p->row0, rpb = kThreads / JS;
    dim3 grid((rows + rpb - 1) / rpb, p->num_envs);
    cd_sweep_kernel<JS><<<grid, kThreads, smem, s>>>(*p, reinterpret_cast<const RecFast *>(rec), inconf, nconf, nlos,
                                                     near_idx, near_tcpa, near_dcpa, near_tin, tcpamax);
    return cudaGetLastError();
}

}  // namespace

extern "C" {

int bsg_cd_pack(int32_t E, int32_t N, const double *lat, const double *lon, const float *alt, const float *trk,
                const float *gs, const float *vs, double lat0, double lon0, float *rec, void *stream) {
    if (E < 1 || N < 1 || !lat || !lon || !alt || !trk || !gs || !vs || !rec) return BSG_ERR_ARG;
    const size_t EN = (size_t)E * N;
    cd_pack_kernel<false><<<(unsigned)((EN + 255) / 256), 256, 0, (cudaStream_t)stream>>>(E, N, lat, lon, alt, trk, gs, vs, lat0,
                                                                                         lon0, reinterpret_cast<RecFast *>(rec));
    return cudaGetLastError() == cudaSuccess ? BSG_OK : BSG_ERR_CUDA;
}

int bsg_cd_pack_init(int32_t E, int32_t N, const double *lat, const double *lon, const float *alt, const float *trk,
                     const float *gs, const float *vs, double lat0, double lon0, float *rec, int32_t *nconf, int32_t *nlos,
                     float *tcpamax, int64_t *key, void *stream) {
    if (E < 1 || N < 1 || !lat || !lon || !alt || !trk || !gs || !vs || !rec || !nconf || !nlos || !tcpamax || !key) return BSG_ERR_ARG;
    const size_t EN = (size_t)E * N;
    cd_pack_kernel<true><<<(unsigned)((EN + 255) / 256), 256, 0, (cudaStream_t)stream>>>(
        E, N, lat, lon, alt, trk, gs, vs, lat0, lon0, reinterpret_cast<RecFast *>(rec), nconf, nlos, (int32_t *)tcpamax, (long long *)key);
    return cudaGetLastError() == cudaSuccess ? BSG_OK : BSG_ERR_CUDA;
}

int bsg_cd_sweep(const bsg_cd_params *p, const float *rec, uint8_t *inconf, int32_t *nconf, int32_t *nlos,
                 int32_t *near_idx, float *near_tcpa, float *near_dcpa, float *near_tin, float *tcpamax, void *stream) {
    if (!p || !rec || p->num_envs < 1 || p->num_ac < 1 || p->row0 < 0 || p->row1 > p->num_ac || p->row1 <= p->row0)
        return BSG_ERR_ARG;
    // pick the row split so the grid covers the 148 SMs several times over
    const long rows = (long)(p->row1 - p->row0) * p->num_envs;
    int js = 1;
    while (js < 32 && rows * js < 148L * 8 * kThreads) js <<= 1;
    cudaStream_t s = (cudaStream_t)stream;
    cudaError_t e;
#define SW(J) e = launch_sweep<J>(p, rec, inconf, nconf, nlos, near_idx, near_tcpa, near_dcpa, near_tin, tcpamax, s)
    switch (js) {
        case 1: SW(1); break;
        case 2: SW(2); break;
        case 4: SW(4); break;
        case 8: SW(8); break;
        case 16: SW(16); break;
        default: SW(32); break;
    }
#undef SW
    return e == cudaSuccess ? BSG_OK : BSG_ERR_CUDA;
}

int bsg_cd_sym_init(const bsg_cd_params *p, int32_t *nconf, int32_t *nlos, float *tcpamax, int64_t *key, void *stream) {
    if (!p || !nconf || !nlos || !tcpamax || !key || p->num_envs < 1 || p->num_ac < 1) return BSG_ERR_ARG;
    const size_t n = (size_t)p->num_envs * p->num_ac;
    cd_sym_init_kernel<<<(unsigned)((n + 255) / 256), 256, 0, (cudaStream_t)stream>>>(n, nconf, nlos, (int32_t *)tcpamax,
                                                                                       (long long *)key);
    return cudaGetLastError() == cudaSuccess ? BSG_OK : BSG_ERR_CUDA;
}

static bool sym_peers_from(const bsg_cd_params *p, const bsg_cd_peers *q, SymPeers &Q) {
    if (!q || q->num_ranks < 1 || q->num_ranks > BSG_MAX_PEERS || q->rows_per_rank < 1 ||
        (long)q->rows_per_rank * q->num_ranks != p->num_ac)
        return false;
    Q.G = q->num_ranks; Q.rows = q->rows_per_rank;
    for (int g = 0; g < Q.G; g++) {
        if (!q->rec[g] || !q->nconf[g] || !q->nlos[g] || !q->tcpamax[g] || !q->key[g]) return false;
        Q.rec[g] = reinterpret_cast<const RecFast *>(q->rec[g]);
        Q.nconf[g] = q->nconf[g]; Q.nlos[g] = q->nlos[g]; Q.tmax[g] = (int32_t *)q->tcpamax[g]; Q.key[g] = (long long *)q->key[g];
    }
    return true;
}

int bsg_cd_sym_sweep_peers(const bsg_cd_params *p, const bsg_cd_peers *q, int32_t tile_rank, int32_t tile_stride, void *stream) {
    SymPeers Q;
    if (!p || p->num_envs < 1 || p->num_ac < 1 || tile_stride < 1 || tile_rank < 0 || tile_rank >= tile_stride ||
        !sym_peers_from(p, q, Q))
        return BSG_ERR_ARG;
    const int nb = (p->num_ac + kSymT - 1) / kSymT, ntile = nb * (nb + 1) / 2;
    // column slices per tile: the smallest of 1 / 2 / 4 whose CTA count fills >= 90 % of its last wave on this GPU
    // (8 ranks x 16 envs x 10 000 aircraft: 1640 tile CTAs over 1184 resident slots = 1.39 waves -> 2 slices, 2.77)
    static int slots = 0;
    if (!slots) {
        int dev = 0, sms = 148, per = 8;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per, cd_sym_sweep_kernel, kSymT, 0);
        slots = sms * (per > 0 ? per : 1);
    }
    int cs = 1;
    double best = 0.0;
    if (const char *f = getenv("BSG_SYM_SLICES")) { int v = atoi(f); if (v == 1 || v == 2 || v == 4) { cs = v; best = 2.0; } }
    for (int c = 1; c <= 4 && best < 0.9; c *= 2) {
        const double w = (double)(((long)ntile * c + tile_stride - 1) / tile_stride) * p->num_envs / slots;
        const double eff = w / (double)(long)(w + 0.999999);
        if (eff > best + 0.02) { best = eff; cs = c; }
    }
    dim3 grid((unsigned)(((long)ntile * cs + tile_stride - 1) / tile_stride), p->num_envs);
    cd_sym_sweep_kernel<<<grid, kSymT, 0, (cudaStream_t)stream>>>(*p, Q, tile_rank, tile_stride, cs);
    return cudaGetLastError() == cudaSuccess ? BSG_OK : BSG_ERR_CUDA;
}

int bsg_cd_sym_finalize_peers(const bsg_cd_params *p, const bsg_cd_peers *q, int32_t rank, uint8_t *inconf,
                              int32_t *near_idx, float *near_tcpa, float *near_dcpa, float *near_tin, void *stream) {
    SymPeers Q;
    if (!p || p->num_envs < 1 || p->num_ac < 1 || !sym_peers_from(p, q, Q) || rank < 0 || rank >= Q.G) return BSG_ERR_ARG;
    const size_t n = (size_t)p->num_envs * Q.rows;
    cd_sym_finalize_kernel<<<(unsigned)((n + 255) / 256), 256, 0, (cudaStream_t)stream>>>(*p, Q, rank, inconf, near_idx,
                                                                                           near_tcpa, near_dcpa, near_tin);
    return cudaGetLastError() == cudaSuccess ? BSG_OK : BSG_ERR_CUDA;
}

int bsg_cd_sym_sweep(const bsg_cd_params *p, const float *rec, int32_t tile_rank, int32_t tile_stride, int32_t *nconf,
                     int32_t *nlos, float *tcpamax, int64_t *key, void *stream) {
    if (!p) return BSG_ERR_ARG;
    bsg_cd_peers q = {};
    q.num_ranks = 1; q.rows_per_rank = p->num_ac;
    q.rec[0] = rec; q.nconf[0] = nconf; q.nlos[0] = nlos; q.tcpamax[0] = tcpamax; q.key[0] = key;
    return bsg_cd_sym_sweep_peers(p, &q, tile_rank, tile_stride, stream);
}

int bsg_cd_sym_finalize(const bsg_cd_params *p, const float *rec, const int32_t *nconf, const int64_t *key,
                        uint8_t *inconf, int32_t *near_idx, float *near_tcpa, float *near_dcpa, float *near_tin,
                        void *stream) {
    if (!p || !rec || !nconf || !key) return BSG_ERR_ARG;
    bsg_cd_peers q = {};
    q.num_ranks = 1; q.rows_per_rank = p->num_ac;
    q.rec[0] = rec; q.nconf[0] = const_cast<int32_t *>(nconf); q.nlos[0] = const_cast<int32_t *>(nconf);
    q.tcpamax[0] = reinterpret_cast<float *>(const_cast<int32_t *>(nconf)); q.key[0] = const_cast<int64_t *>(key);
    return bsg_cd_sym_finalize_peers(p, &q, 0, inconf, near_idx, near_tcpa, near_dcpa, near_tin, stream);
}

}  // extern "C"

// ---------------------------------------------------------------------------------------------------------------
// bsg_cd_sym_static: the symmetric sweep for callers that must not allocate (declaration + rationale: bsgcore.h).
// The sweep's row accumulators -- the packed (tCPA, intruder) minimum key, the conflict and loss-of-separation counts
// and the tCPA maximum, 20 bytes per aircraft row -- are global-memory scratch that bsg_cd_sym_init / _sweep / _finalize
// take from the caller.  Here they live in ONE file-scope device array instead: 5.25 MB that every GPU this library is
// loaded on holds until the process exits, shared by every call (calls that overlap on the GPU race on it).
// Host code may not hand `bsg_static_scratch` to a kernel: in host code that name is only the host-side handle of the
// symbol.  The device address is looked up once per device with cudaGetSymbolAddress, with that device current.
static __device__ __align__(256) unsigned char bsg_static_scratch[(size_t)BSG_CD_STATIC_MAX_ROWS * 20 + 4 * 256];

namespace {

constexpr int kStaticMaxDev = 64;
unsigned char *g_static_addr[kStaticMaxDev] = {};      // device address of bsg_static_scratch, per device ordinal
int g_static_lock = 0;                                  // spin lock around the table (first call per device only)

// Device address of the scratch array on the CURRENT device.  The first call on a device resolves the symbol (which
// also makes the CUDA runtime load this module onto that device); later calls read the table.
int static_scratch_addr(int dev, unsigned char **out) {
    if (dev < 0 || dev >= kStaticMaxDev) return BSG_ERR_UNSUPPORTED;
    unsigned char *a = __atomic_load_n(&g_static_addr[dev], __ATOMIC_ACQUIRE);
    if (!a) {
        while (__atomic_exchange_n(&g_static_lock, 1, __ATOMIC_ACQUIRE)) {}
        a = g_static_addr[dev];
        cudaError_t e = cudaSuccess;
        if (!a) {
            void *sym = nullptr;
            e = cudaGetSymbolAddress(&sym, bsg_static_scratch);
            if (e == cudaSuccess && sym) {
                a = static_cast<unsigned char *>(sym);
                __atomic_store_n(&g_static_addr[dev], a, __ATOMIC_RELEASE);
            }
        }
        __atomic_store_n(&g_static_lock, 0, __ATOMIC_RELEASE);
        if (e != cudaSuccess || !a) return BSG_ERR_CUDA;
    }
    *out = a;
    return BSG_OK;
}

// counts and tCPA maximum of the rows: scratch -> the caller's outputs (the ordered sweep writes them itself)
__global__ void cd_static_export_kernel(size_t n, const int32_t *__restrict__ s_nconf, const int32_t *__restrict__ s_nlos,
                                        const int32_t *__restrict__ s_tmax, int32_t *nconf, int32_t *nlos, float *tcpamax) {
    const size_t k = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= n) return;
    if (nconf) nconf[k] = s_nconf[k];
    if (nlos) nlos[k] = s_nlos[k];
    if (tcpamax) tcpamax[k] = __int_as_float(s_tmax[k]);
}

}  // namespace

extern "C" int bsg_cd_sym_static(const bsg_cd_params *p, const float *rec, uint8_t *inconf, int32_t *nconf, int32_t *nlos,
                                 int32_t *near_idx, float *near_tcpa, float *near_dcpa, float *near_tin, float *tcpamax,
                                 void *stream) {
    if (!p || !rec || p->num_envs < 1 || p->num_ac < 1) return BSG_ERR_ARG;
    const size_t n = (size_t)p->num_envs * (size_t)p->num_ac;
    if (n > (size_t)BSG_CD_STATIC_MAX_ROWS) return BSG_ERR_TOOLARGE;      // nothing was launched, nothing was touched
    int dev = -1;
    if (cudaGetDevice(&dev) != cudaSuccess) return BSG_ERR_CUDA;
    // the records must live on the device whose copy of the scratch array the kernels will use (the current one)
    cudaPointerAttributes at;
    if (cudaPointerGetAttributes(&at, rec) != cudaSuccess) { cudaGetLastError(); return BSG_ERR_ARG; }
    if (at.type == cudaMemoryTypeUnregistered || (at.type == cudaMemoryTypeDevice && at.device != dev)) return BSG_ERR_ARG;
    unsigned char *base = nullptr;
    if (int rc = static_scratch_addr(dev, &base)) return rc;
    // carve the array: key (8 B / row), then nconf, nlos, tcpamax bits (4 B / row each), every part 256-byte aligned
    const size_t r8 = (n * 8 + 255) & ~(size_t)255, r4 = (n * 4 + 255) & ~(size_t)255;
    int64_t *s_key = reinterpret_cast<int64_t *>(base);
    int32_t *s_nconf = reinterpret_cast<int32_t *>(base + r8);
    int32_t *s_nlos = reinterpret_cast<int32_t *>(base + r8 + r4);
    int32_t *s_tmax = reinterpret_cast<int32_t *>(base + r8 + 2 * r4);
    // the array keeps whatever the last call left in it: the init stage rewrites every row this call reads
    int rc = bsg_cd_sym_init(p, s_nconf, s_nlos, reinterpret_cast<float *>(s_tmax), s_key, stream);
    if (rc) return rc;
    rc = bsg_cd_sym_sweep(p, rec, 0, 1, s_nconf, s_nlos, reinterpret_cast<float *>(s_tmax), s_key, stream);
    if (rc) return rc;
    rc = bsg_cd_sym_finalize(p, rec, s_nconf, s_key, inconf, near_idx, near_tcpa, near_dcpa, near_tin, stream);
    if (rc) return rc;
    if (nconf || nlos || tcpamax) {
        cd_static_export_kernel<<<(unsigned)((n + 255) / 256), 256, 0, (cudaStream_t)stream>>>(n, s_nconf, s_nlos, s_tmax,
                                                                                               nconf, nlos, tcpamax);
        if (cudaGetLastError() != cudaSuccess) return BSG_ERR_CUDA;
    }
    return BSG_OK;
}
