// bsg_api.cu -- C-ABI of libbsgcore.so (include/bsgcore.h): handle, state interchange, fused step launch.
// No torch types cross this boundary; device pointers are plain void*/T*.
#include <cuda_runtime.h>
#include <cstdio>
#include <cstring>
#include <cstdlib>
#include <string>
#include <new>
#include "bsg_step_kernel.cuh"
#include "bsg_handle.h"

using namespace bsg;

static size_t rsize(const bsg_handle *h) { return h->cfg.precision == BSG_FP64 ? 8 : 4; }

// ------------------------------------------------------------------ state interchange kernels
template <typename R>
__global__ void import_state_kernel(const double *ac, int E, int N, double *pos, R *st) {
    size_t EN = (size_t)E * N, k = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= EN) return;
    size_t e = k / N, i = k - e * N;
    const double *a = ac + e * BSG_NF * N + i;
    pos[k] = a[BSG_F_LAT * N]; pos[EN + k] = a[BSG_F_LON * N];
    const int map[16] = { BSG_F_ALT, BSG_F_HDG, BSG_F_TAS, BSG_F_VS, BSG_F_CAS, BSG_F_SELSPD, BSG_F_SELALT, BSG_F_SELVS,
                          BSG_F_APTRK, BSG_F_AX, BSG_F_ASAS_TRK, BSG_F_ASAS_TAS, BSG_F_ASAS_VS, BSG_F_ASAS_ALT,
                          BSG_F_ASAS_ACTIVE, BSG_F_PHASE };
#pragma unroll
    for (int f = 0; f < 16; f++) st[(size_t)f * EN + k] = (R)a[(size_t)map[f] * N];
    for (int f = S_C_ALT; f < S_NS; f++) st[(size_t)f * EN + k] = (R)NAN;     // invalidate the memo keys
}
template <typename R>
__global__ void export_state_kernel(double *ac, int E, int N, const double *pos, const R *st) {
    size_t EN = (size_t)E * N, k = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= EN) return;
    size_t e = k / N, i = k - e * N;
    double *a = ac + e * BSG_NF * N + i;
    a[BSG_F_LAT * N] = pos[k]; a[BSG_F_LON * N] = pos[EN + k];
    const int map[16] = { BSG_F_ALT, BSG_F_HDG, BSG_F_TAS, BSG_F_VS, BSG_F_CAS, BSG_F_SELSPD, BSG_F_SELALT, BSG_F_SELVS,
                          BSG_F_APTRK, BSG_F_AX, BSG_F_ASAS_TRK, BSG_F_ASAS_TAS, BSG_F_ASAS_VS, BSG_F_ASAS_ALT,
                          BSG_F_ASAS_ACTIVE, BSG_F_PHASE };
#pragma unroll
    for (int f = 0; f < 16; f++) a[(size_t)map[f] * N] = (double)st[(size_t)f * EN + k];
}

extern "C" {

int bsg_abi_version(void) { return BSG_ABI_VERSION; }

const char *bsg_status_string(int s) {
    switch (s) {
        case BSG_OK: return "ok";
        case BSG_ERR_ARG: return "invalid argument";
        case BSG_ERR_CUDA: return "CUDA error (see bsg_last_error)";
        case BSG_ERR_NOMEM: return "out of memory";
        case BSG_ERR_UNSUPPORTED: return "unsupported configuration";
        case BSG_ERR_NODEVICE: return "no CUDA device";
        default: return "unknown status";
    }
}

int bsg_device_count(int *count) {
    if (!count) return BSG_ERR_ARG;
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess) { *count = 0; cudaGetLastError(); return BSG_ERR_NODEVICE; }
    *count = n;
    return BSG_OK;
}

const char *bsg_last_error(const bsg_handle *h) { return h ? h->err.c_str() : "null handle"; }
int bsg_obs_dim(const bsg_handle *h) { return h ? h->obs_dim : BSG_ERR_ARG; }
int bsg_act_dim(const bsg_handle *h) { return h ? h->act_dim : BSG_ERR_ARG; }
int bsg_launch_count(const bsg_handle *h, int64_t *n) {
    if (!h || !n) return BSG_ERR_ARG;
    *n = h->launches;
    return BSG_OK;
}

int bsg_create(const bsg_config *c, bsg_handle **out) {
    if (!c || !out || c->struct_size != (int32_t)sizeof(bsg_config)) return BSG_ERR_ARG;
    if (c->env_kind < 0 || c->env_kind > BSG_KIND_RAW || c->num_envs < 1 || c->max_ac < 1) return BSG_ERR_ARG;
    if (c->max_ac > 256) return BSG_ERR_UNSUPPORTED;          // fused step: one env's rows live in one CTA
    if (c->reso_on && (c->max_ac > 64 || !c->asas_on)) return BSG_ERR_UNSUPPORTED;
    int ndev = 0;
    if (bsg_device_count(&ndev) != BSG_OK || ndev < 1) return BSG_ERR_NODEVICE;
    if (c->device < 0 || c->device >= ndev) return BSG_ERR_ARG;
    bsg_handle *h = new (std::nothrow) bsg_handle();
    if (!h) return BSG_ERR_NOMEM;
    *out = h;
    h->cfg = *c;
    bsg_config &g = h->cfg;
    static const int dsub[4] = { 10, 30, 5, 1 };
    static const double ddt[4] = { 5.0, 1.0, 1.0, 1.0 };
    if (g.substeps <= 0) g.substeps = dsub[g.env_kind];
    if (g.simdt <= 0) g.simdt = ddt[g.env_kind];
    if (g.rpz_m <= 0) g.rpz_m = 5 * kNm;
    if (g.hpz_m <= 0) g.hpz_m = 1000 * kFt;
    if (g.dtlookahead <= 0) g.dtlookahead = 300.0;
    if (g.resofach <= 0) g.resofach = 1.05;
    if (g.resofacv <= 0) g.resofacv = 1.05;
    const int N = g.max_ac, K = N - 1, E = g.num_envs;
    h->obs_dim = g.env_kind == BSG_KIND_HORIZONTAL ? 5 * K + 3 : g.env_kind == BSG_KIND_VERTICAL ? 4 + 7 * K
                 : g.env_kind == BSG_KIND_SECTOR ? 31 : 0;
    h->act_dim = g.env_kind == BSG_KIND_SECTOR ? 2 : (g.env_kind == BSG_KIND_RAW ? 0 : 1);
    // CTA size: the one of {128, 256} that wastes fewer lanes on whole envs (ties -> 128: better balance)
    {   // ISA conversions of the fixed creation speeds (host fp64 replica of vcas2tas / vtas2cas)
        auto atmos = [](double hh, double &p, double &rho) {
            const double T = std::fmax(288.15 - 0.0065 * hh, 216.65);
            rho = 1.225 * std::pow(T / 288.15, 4.256848030018761) * std::exp(-std::fmax(0.0, hh - 11000.0) / 6341.552161);
            p = rho * kRgas * T;
        };
        auto cas2tas = [&](double cas, double hh) {
            double p, rho; atmos(hh, p, rho);
            const double q = kP0 * (std::pow(1.0 + kRho0 * cas * cas / (7.0 * kP0), 3.5) - 1.0);
            return std::sqrt(7.0 * p / rho * (std::pow(1.0 + q / p, 2.0 / 7.0) - 1.0));
        };
        auto tas2cas = [&](double tas, double hh) {
            double p, rho; atmos(hh, p, rho);
            const double q = p * (std::pow(1.0 + rho * tas * tas / (7.0 * p), 3.5) - 1.0);
            return std::sqrt(7.0 * kP0 / kRho0 * (std::pow(q / kP0 + 1.0, 2.0 / 7.0) - 1.0));
        };
        h->rc.tas_h = cas2tas(150.0, 0.0);
        h->rc.cas_h = tas2cas(h->rc.tas_h, 0.0);
        h->rc.tas_s = cas2tas(150.0, 350.0);
        h->rc.coslat_hv = std::cos(52.0 * kD2R);
        h->rc.ccl = std::cos(kCenterLat * kD2R);
    }
    h->tpb = 128;
    if (N > 128 || (256 / N) * N * 128 > (128 / N) * N * 256) h->tpb = 256;
    if (const char *f = getenv("BSG_FORCE_TPB")) { int t = atoi(f); if ((t == 128 || t == 256 || (t == 320 && g.precision == BSG_FP32)) && t >= N) h->tpb = t; }
    h->epb = h->tpb / N;
    if (const char *f = getenv("BSG_FORCE_EPB")) { int t = atoi(f); if (t >= 1 && t * N <= h->tpb) h->epb = t; }
    const size_t rs = rsize(h);
    h->smem = g.precision == BSG_FP64 ? step_smem<double>(N, h->epb).total : step_smem<float>(N, h->epb).total;
    CK(h, cudaSetDevice(g.device));
    const size_t EN = (size_t)E * N;
    CK(h, cudaMalloc(&h->pos, 2 * EN * sizeof(double)));
    CK(h, cudaMalloc(&h->st, (size_t)S_NS * EN * rs));
    CK(h, cudaMalloc(&h->reso, EN * sizeof(unsigned long long)));
    CK(h, cudaMalloc(&h->nac, E * sizeof(int)));
    CK(h, cudaMalloc(&h->ev, (size_t)E * BSG_NE * sizeof(double)));
    CK(h, cudaMalloc(&h->stage64, (size_t)E * BSG_NF * N * sizeof(double)));
    CK(h, cudaMemset(h->pos, 0, 2 * EN * sizeof(double)));
    CK(h, cudaMemset(h->st, 0, (size_t)S_NS * EN * rs));
    CK(h, cudaMemset(h->reso, 0, EN * sizeof(unsigned long long)));
    CK(h, cudaMemset(h->nac, 0, E * sizeof(int)));
    CK(h, cudaMemset(h->ev, 0, (size_t)E * BSG_NE * sizeof(double)));
    // host-buffer step: [obs | reward | info | terminated | truncated] in one pinned block
    h->act_bytes = (size_t)E * (h->act_dim > 0 ? h->act_dim : 1) * rs;
    h->out_bytes = (size_t)E * (h->obs_dim + 1 + 4) * rs + (size_t)E * 2 * sizeof(int32_t);
    CK(h, cudaMallocHost(&h->h_act, h->act_bytes));
    CK(h, cudaMallocHost(&h->h_out, h->out_bytes));
    CK(h, cudaMalloc(&h->d_act, h->act_bytes));
    CK(h, cudaMalloc(&h->d_out, h->out_bytes));
    CK(h, cudaStreamCreateWithFlags(&h->hs, cudaStreamNonBlocking));
    if (h->smem > 48 * 1024) { h->err = "shared memory per CTA above 48 KB"; return BSG_ERR_UNSUPPORTED; }
    return BSG_OK;
}

int bsg_destroy(bsg_handle *h) {
    if (!h) return BSG_ERR_ARG;
    cudaSetDevice(h->cfg.device);
    cudaFree(h->pos); cudaFree(h->st); cudaFree(h->reso); cudaFree(h->nac); cudaFree(h->ev); cudaFree(h->stage64);
    cudaFree(h->d_act); cudaFree(h->d_out); cudaFree(h->episode); cudaFree(h->done_buf[0]); cudaFree(h->done_buf[1]);
    if (h->h_act) cudaFreeHost(h->h_act);
    if (h->h_out) cudaFreeHost(h->h_out);
    if (h->hs) cudaStreamDestroy(h->hs);
    delete h;
    return BSG_OK;
}

}  // extern "C"

#include "bsg_api_step.inl"

#ifdef BSG_TRACE
extern "C" int bsg_debug_trace(unsigned long long *out_host, int n_words) {
    return cudaMemcpyFromSymbol(out_host, bsg::g_trace, (size_t)n_words * sizeof(unsigned long long)) == cudaSuccess ? 0 : -1;
}
#endif
