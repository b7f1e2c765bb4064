// bsg_api.cu -- C-ABI of libbsgcore.so (include/bsgcore.h): handle, state interchange, fused step launch.
// No torch types cross this boundary; device pointers are plain void*/T*.
#include <vector>
#include <cuda_runtime.h>
#include <cstdio>
#include <cstring>
#include <cstdlib>
#include <string>
#include <new>
#include "bsg_step_kernel.cuh"
#include "bsg_step_launch.h"
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
    const int map[19] = { BSG_F_ALT, BSG_F_HDG, BSG_F_TAS, BSG_F_VS, BSG_F_CAS, BSG_F_SELSPD, BSG_F_SELALT, BSG_F_SELVS,
                          BSG_F_APTRK, BSG_F_AX, BSG_F_ASAS_TRK, BSG_F_ASAS_TAS, BSG_F_ASAS_VS, BSG_F_ASAS_ALT,
                          BSG_F_ASAS_ACTIVE, BSG_F_PHASE, BSG_F_GSN, BSG_F_GSE, BSG_F_LNAV };
#pragma unroll
    for (int f = 0; f < 19; f++) st[(size_t)f * EN + k] = (R)a[(size_t)map[f] * N];
    for (int f = S_C_ALT; f < S_NS; f++) st[(size_t)f * EN + k] = (R)NAN;     // invalidate the memo keys
}
template <typename R>
__global__ void export_state_kernel(double *ac, int E, int N, const double *pos, const R *st, int derive_gs) {
    size_t EN = (size_t)E * N, k = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= EN) return;
    size_t e = k / N, i = k - e * N;
    double *a = ac + e * BSG_NF * N + i;
    a[BSG_F_LAT * N] = pos[k]; a[BSG_F_LON * N] = pos[EN + k];
    const int map[19] = { BSG_F_ALT, BSG_F_HDG, BSG_F_TAS, BSG_F_VS, BSG_F_CAS, BSG_F_SELSPD, BSG_F_SELALT, BSG_F_SELVS,
                          BSG_F_APTRK, BSG_F_AX, BSG_F_ASAS_TRK, BSG_F_ASAS_TAS, BSG_F_ASAS_VS, BSG_F_ASAS_ALT,
                          BSG_F_ASAS_ACTIVE, BSG_F_PHASE, BSG_F_GSN, BSG_F_GSE, BSG_F_LNAV };
#pragma unroll
    for (int f = 0; f < 19; f++) a[(size_t)map[f] * N] = (double)st[(size_t)f * EN + k];
    if (derive_gs) {     // only the EXT kernels keep the ground-speed vector in the state: without wind it is tas * (cos, sin)(hdg)
        const double tas = (double)st[(size_t)S_TAS * EN + k], hdg = (double)st[(size_t)S_HDG * EN + k] * kD2R;
        a[(size_t)BSG_F_GSN * N] = tas * cos(hdg); a[(size_t)BSG_F_GSE * N] = tas * sin(hdg);
    }
}

// the plain kernels do not keep the ground-speed vector in the state: refresh it when the EXT features switch on
template <typename R>
__global__ void derive_gs_kernel(int E, int N, R *st) {
    size_t EN = (size_t)E * N, k = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= EN) return;
    const double tas = (double)st[(size_t)S_TAS * EN + k], hdg = (double)st[(size_t)S_HDG * EN + k] * kD2R;
    st[(size_t)S_GSN * EN + k] = (R)(tas * cos(hdg)); st[(size_t)S_GSE * EN + k] = (R)(tas * sin(hdg));
}
static bool is_ext(const bsg_handle *h) { return h->cfg.env_kind == BSG_KIND_MERGE || h->wind.n > 0 || h->noise_sigma > 0.0; }
static int refresh_gs(bsg_handle *h) {
    const size_t EN = (size_t)h->cfg.num_envs * h->cfg.max_ac;
    const int tb = 256, gb = (int)((EN + tb - 1) / tb);
    if (h->cfg.precision == BSG_FP64) derive_gs_kernel<double><<<gb, tb>>>(h->cfg.num_envs, h->cfg.max_ac, (double *)h->st);
    else derive_gs_kernel<float><<<gb, tb>>>(h->cfg.num_envs, h->cfg.max_ac, (float *)h->st);
    h->launches++;
    CK(h, cudaGetLastError());
    CK(h, cudaDeviceSynchronize());
    return BSG_OK;
}

// the pinned / device output block of bsg_step_host; sized by obs_dim (which bsg_set_wind can grow by two slots)
static int alloc_out(bsg_handle *h) {
    h->gen++;                                      // captured host-step graphs hold the old output pointers
    if (h->h_out) { cudaFreeHost(h->h_out); h->h_out = nullptr; }
    if (h->h_outs[1]) { cudaFreeHost(h->h_outs[1]); h->h_outs[1] = nullptr; }
    if (h->d_out) { cudaFree(h->d_out); h->d_out = nullptr; }
    h->out_bytes = (size_t)h->cfg.num_envs * (h->obs_dim + 1 + 4) * rsize(h) + (size_t)h->cfg.num_envs * 2 * sizeof(int32_t);
    CK(h, cudaMallocHost(&h->h_out, h->out_bytes));
    CK(h, cudaMallocHost(&h->h_outs[1], h->out_bytes));
    h->h_outs[0] = h->h_out;
    CK(h, cudaMalloc(&h->d_out, h->out_bytes));
    return BSG_OK;
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
        case BSG_ERR_TOOLARGE: return "more rows than the built-in scratch array holds (BSG_CD_STATIC_MAX_ROWS); use bsg_cd_sym_init/_sweep/_finalize with your own buffers";
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
    if (c->env_kind < 0 || c->env_kind > BSG_KIND_MERGE || c->num_envs < 1 || c->max_ac < 1) return BSG_ERR_ARG;
    if (c->env_kind == BSG_KIND_MERGE && (c->max_ac < 2 || c->max_ac > 64)) return BSG_ERR_UNSUPPORTED;   // EXT kernels: N <= 64
    if (c->max_ac > 256) return BSG_ERR_UNSUPPORTED;          // fused step: one env's rows live in one CTA
    if (c->reso_on && !c->asas_on) return BSG_ERR_UNSUPPORTED;
    if (c->reso_mode < 0 || c->reso_mode > 2 || (c->reso_mode != 0 && !c->reso_on)) return BSG_ERR_ARG;
    if (c->reso_mode != 0 && c->env_kind == BSG_KIND_MERGE && c->precision != BSG_FP64) return BSG_ERR_UNSUPPORTED;   // fp32 EXT kernels: symmetric sweep only
    int ndev = 0;
    if (bsg_device_count(&ndev) != BSG_OK || ndev < 1) return BSG_ERR_NODEVICE;
    if (c->device < 0 || c->device >= ndev) return BSG_ERR_ARG;
    bsg_handle *h = new (std::nothrow) bsg_handle();
    if (!h) return BSG_ERR_NOMEM;
    *out = h;
    h->cfg = *c;
    bsg_config &g = h->cfg;
    static const int dsub[5] = { 10, 30, 5, 1, 10 };
    static const double ddt[5] = { 5.0, 1.0, 1.0, 1.0, 5.0 };
    if (g.substeps <= 0) g.substeps = dsub[g.env_kind];
    if (g.simdt <= 0) g.simdt = ddt[g.env_kind];
    if (g.rpz_m <= 0) g.rpz_m = 5 * kNm;
    if (g.hpz_m <= 0) g.hpz_m = 1000 * kFt;
    if (g.dtlookahead <= 0) g.dtlookahead = 300.0;
    if (g.resofach <= 0) g.resofach = 1.05;
    if (g.resofacv <= 0) g.resofacv = 1.05;
    const int N = g.max_ac, K = N - 1, E = g.num_envs;
    (void)K;
    h->obs_dim = env_obs_dim(g.env_kind, N);
    h->act_dim = (g.env_kind == BSG_KIND_SECTOR || g.env_kind == BSG_KIND_MERGE) ? 2 : (g.env_kind == BSG_KIND_RAW ? 0 : 1);
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
        // Merge (merge_env.py:39-45): runway, FAF = get_point_at_distance(runway, 200 km, bearing 0) (functions.py:24-42)
        h->rc.tas_m = cas2tas(100.0, 10000.0);
        h->rc.rwy_lat = 52.36239301495972; h->rc.rwy_lon = 4.713195734579777;
        const double lat1 = h->rc.rwy_lat * kD2R, lon1 = h->rc.rwy_lon * kD2R, dr = 200.0 / 6371.0, al = 0.0;
        const double lat2 = std::asin(std::sin(lat1) * std::cos(dr) + std::cos(lat1) * std::sin(dr) * std::cos(al));
        const double lon2 = lon1 + std::atan2(std::sin(al) * std::sin(dr) * std::cos(lat1), std::cos(dr) - std::sin(lat1) * std::sin(lat2));
        h->rc.faf_lat = lat2 * kR2D; h->rc.faf_lon = lon2 * kR2D;
    }
    h->tpb = 128;
    if (N > 128 || (256 / N) * N * 128 > (128 / N) * N * 256) h->tpb = 256;
    if (const char *f = getenv("BSG_FORCE_TPB")) { int t = atoi(f); if ((t == 128 || t == 256) && t >= N) h->tpb = t; }
    h->epb = h->tpb / N;
    if (const char *f = getenv("BSG_FORCE_EPB")) { int t = atoi(f); if (t >= 1 && t * N <= h->tpb) h->epb = t; }
    const size_t rs = rsize(h);
    {
        const bool reso = g.reso_on && g.asas_on;
        h->nbuf = g.precision == BSG_FP64 ? step_nbuf<double>(N, h->epb, reso) : step_nbuf<float>(N, h->epb, reso);
        if (const char *f = getenv("BSG_FORCE_NBUF")) { int t = atoi(f); if (t >= 2 && t <= 16) h->nbuf = t; }
        const bool pairs = reso && g.precision != BSG_FP64 && N <= 64 && !getenv("BSG_NO_SYM") && g.reso_mode == 0;   // the symmetric sweep's pair scratch
        h->smem = g.precision == BSG_FP64 ? step_smem<double>(N, h->epb, h->nbuf, false).total : step_smem<float>(N, h->epb, h->nbuf, pairs).total;
    }
    CK(h, cudaSetDevice(g.device));
    // detection-only symmetric fp32 kernel: a one-wave launch shape that balances the warps over the SMs, when there is one
    if (g.precision != BSG_FP64 && N <= 64 && !(g.reso_on && g.asas_on) && !getenv("BSG_NO_SYM") && !getenv("BSG_NO_WIDE")) {
        int we = 0, wn = 0; size_t ws = 0;
        if (const char *f = getenv("BSG_WIDE_EPB")) {          // A/B hook: force the envs per CTA of the wide shape
            we = atoi(f);
            if (we >= 1 && we * N <= 640) {
                wn = BSG_CH < g.substeps ? BSG_CH : (g.substeps > 2 ? g.substeps : 2);
                while (wn > 2 && step_smem<float>(N, we, wn, false).total > 100 * 1024) wn--;
                h->wepb = we; h->wnbuf = wn; h->wsmem = step_smem<float>(N, we, wn, false).total;
            }
        } else if (pick_wide_shape(E, N, g.substeps, h->epb, ((h->epb * N + 31) / 32) * 32, &we, &wn, &ws)) {
            h->wepb = we; h->wnbuf = wn; h->wsmem = ws;
        }
    }
    const size_t EN = (size_t)E * N;
    CK(h, cudaMalloc(&h->pos, 2 * EN * sizeof(double)));
    CK(h, cudaMalloc(&h->st, (size_t)S_NS * EN * rs));
    const size_t RW = (size_t)(N + 63) / 64;          // resopairs words per aircraft, word-major: reso[RW][E*N]
    CK(h, cudaMalloc(&h->reso, RW * EN * sizeof(unsigned long long)));
    CK(h, cudaMalloc(&h->nac, E * sizeof(int)));
    CK(h, cudaMalloc(&h->ev, (size_t)E * BSG_NE * sizeof(double)));
    CK(h, cudaMalloc(&h->stage64, (size_t)E * BSG_NF * N * sizeof(double)));
    CK(h, cudaMemset(h->pos, 0, 2 * EN * sizeof(double)));
    CK(h, cudaMemset(h->st, 0, (size_t)S_NS * EN * rs));
    CK(h, cudaMemset(h->reso, 0, RW * EN * sizeof(unsigned long long)));
    CK(h, cudaMemset(h->nac, 0, E * sizeof(int)));
    CK(h, cudaMemset(h->ev, 0, (size_t)E * BSG_NE * sizeof(double)));
    // host-buffer step: [obs | reward | info | terminated | truncated] in one pinned block
    h->act_bytes = (size_t)E * (h->act_dim > 0 ? h->act_dim : 1) * rs;
    CK(h, cudaMallocHost(&h->h_act, h->act_bytes));
    CK(h, cudaMallocHost(&h->h_acts[1], h->act_bytes));
    h->h_acts[0] = h->h_act;
    for (int q = 0; q < 2; q++) CK(h, cudaEventCreateWithFlags(&h->hev[q], cudaEventDisableTiming));
    CK(h, cudaEventCreateWithFlags(&h->nd_ev, cudaEventDisableTiming));
    CK(h, cudaMalloc(&h->d_act, h->act_bytes));
    { int rc = alloc_out(h); if (rc != BSG_OK) return rc; }
    CK(h, cudaStreamCreateWithFlags(&h->hs, cudaStreamNonBlocking));
    if (h->smem > 227 * 1024) { h->err = "shared memory per CTA above 227 KB"; return BSG_ERR_UNSUPPORTED; }
    return BSG_OK;
}

int bsg_set_wind(bsg_handle *h, int32_t npoints, const double *lat, const double *lon, const double *vnorth,
                 const double *veast, int32_t augment_obs) {
    if (!h || npoints < 0 || npoints > BSG_MAX_WIND) return BSG_ERR_ARG;
    if (npoints > 0 && (!lat || !lon || !vnorth || !veast)) return BSG_ERR_ARG;
    if (h->cfg.env_kind == BSG_KIND_RAW) return BSG_ERR_UNSUPPORTED;
    if (npoints > 0 && h->cfg.precision == BSG_FP32 && h->cfg.max_ac > 64) {
        h->err = "bsg_set_wind: the fp32 EXT kernels hold N <= 64 aircraft per env"; return BSG_ERR_UNSUPPORTED;
    }
    CK(h, cudaSetDevice(h->cfg.device));
    CK(h, cudaDeviceSynchronize());
    const bool was_ext = is_ext(h);
    h->gen++;
    memset(&h->wind, 0, sizeof(h->wind));
    h->wind.n = npoints;
    h->wind.obs = (npoints > 0 && augment_obs) ? 1 : 0;
    for (int k = 0; k < npoints; k++) { h->wind.lat[k] = lat[k]; h->wind.lon[k] = lon[k]; h->wind.vn[k] = vnorth[k]; h->wind.ve[k] = veast[k]; }
    if (!was_ext && is_ext(h)) { int rc = refresh_gs(h); if (rc != BSG_OK) return rc; }
    const int od = env_obs_dim(h->cfg.env_kind, h->cfg.max_ac) + (h->wind.obs ? 2 : 0);
    if (od != h->obs_dim) { h->obs_dim = od; return alloc_out(h); }
    return BSG_OK;
}

int bsg_set_wind_profile(bsg_handle *h, double lat, double lon, int32_t nknots, const double *alt_m, const double *vnorth,
                         const double *veast, int32_t augment_obs) {
    if (!h || nknots < 1 || nknots > BSG_MAX_WIND || !alt_m || !vnorth || !veast) return BSG_ERR_ARG;
    for (int k = 1; k < nknots; k++) if (!(alt_m[k] > alt_m[k - 1])) return BSG_ERR_ARG;     // np.interp wants increasing knots
    int rc = bsg_set_wind(h, 1, &lat, &lon, vnorth, veast, augment_obs);                    // one point: ext switch, obs slots
    if (rc != BSG_OK) return rc;
    h->gen++;
    h->wind.nk = nknots;
    for (int k = 0; k < nknots; k++) { h->wind.kalt[k] = alt_m[k]; h->wind.vn[k] = vnorth[k]; h->wind.ve[k] = veast[k]; }
    return BSG_OK;
}

int bsg_set_obs_noise(bsg_handle *h, double sigma, uint64_t seed) {
    if (!h || !(sigma >= 0.0)) return BSG_ERR_ARG;
    if (h->cfg.env_kind == BSG_KIND_RAW) return BSG_ERR_UNSUPPORTED;
    if (sigma > 0.0 && h->cfg.precision == BSG_FP32 && h->cfg.max_ac > 64) {
        h->err = "bsg_set_obs_noise: the fp32 EXT kernels hold N <= 64 aircraft per env"; return BSG_ERR_UNSUPPORTED;
    }
    const bool was_ext = is_ext(h);
    h->noise_sigma = sigma; h->noise_seed = seed; h->gen++;
    if (!was_ext && is_ext(h)) { CK(h, cudaSetDevice(h->cfg.device)); return refresh_gs(h); }
    return BSG_OK;
}

int bsg_destroy(bsg_handle *h) {
    if (!h) return BSG_ERR_ARG;
    cudaSetDevice(h->cfg.device);
    cudaFree(h->pos); cudaFree(h->st); cudaFree(h->reso); cudaFree(h->nac); cudaFree(h->ev); cudaFree(h->stage64);
    for (int a = 0; a < 2; a++) for (int b = 0; b < 2; b++) if (h->hgraph[a][b]) cudaGraphExecDestroy(h->hgraph[a][b]);
    cudaFree(h->d_act); cudaFree(h->d_out); cudaFree(h->episode); cudaFree(h->done_buf[0]); cudaFree(h->done_buf[1]); cudaFree(h->nd_rec);
    if (h->h_act) cudaFreeHost(h->h_act);
    if (h->h_acts[1]) cudaFreeHost(h->h_acts[1]);
    if (h->h_outs[1]) cudaFreeHost(h->h_outs[1]);
    for (int q = 0; q < 2; q++) if (h->hev[q]) cudaEventDestroy(h->hev[q]);
    if (h->nd_ev) cudaEventDestroy(h->nd_ev);
    if (h->h_out) cudaFreeHost(h->h_out);
    if (h->hs) cudaStreamDestroy(h->hs);
    delete h;
    return BSG_OK;
}

}  // extern "C"

#include "bsg_api_step.inl"

#ifdef BSG_TRACE
// (the stamps live in the translation unit that holds the fp32 instantiations: bsg_step_f32.cu)
namespace bsg { int trace_read_f32(unsigned long long *out_host, int n_words); }
extern "C" int bsg_debug_trace(unsigned long long *out_host, int n_words) { return bsg::trace_read_f32(out_host, n_words); }
#endif
