// bsg_api_step.inl -- state interchange and the step entry points (included by bsg_api.cu).

template <typename R>
static int launch_step(bsg_handle *h, const void *actions, const bsg_step_out *out, int substeps, int observe_only,
                       cudaStream_t s) {
    const bsg_config &g = h->cfg;
    StepArgs<R> A;
    A.kind = g.env_kind; A.E = g.num_envs; A.N = g.max_ac; A.substeps = substeps; A.asas_on = g.asas_on;
    A.reso_on = g.reso_on; A.reso_off_own = g.reso_off_own; A.max_steps = g.max_episode_steps; A.epb = h->epb;
    A.observe_only = observe_only; A.nbuf = h->nbuf; A.force_walk = getenv("BSG_FORCE_WALK") ? 1 : 0; A.reso_mode = g.reso_mode;
    A.dt = (R)g.simdt;
    A.P.rpz = g.rpz_m; A.P.hpz = g.hpz_m; A.P.dtlook = g.dtlookahead; A.P.fach = g.resofach; A.P.facv = g.resofacv;
    cd_params_fast(A.P, A.PF);
    A.pos = h->pos; A.st = (R *)h->st; A.reso = h->reso; A.nac = h->nac; A.ev = h->ev;
    A.actions = (const R *)actions;
    if (out) A.out = *out; else memset(&A.out, 0, sizeof(A.out));
    const bool ar = h->autoreset && !observe_only && g.env_kind != BSG_KIND_RAW;
    A.skip = (ar && h->autoreset == 1) ? h->done_buf[h->flip] : nullptr;
    // next-step mode: the reset of the envs that ended in the previous step is fused into this launch
    A.fuse_reset = (A.skip != nullptr) ? 1 : 0; A.fixed_nac = g.fixed_nac; A.density_mode = g.density_mode; A.RC = h->rc; A.ar_seed = h->ar_seed; A.episode = h->episode;
    A.W = h->wind; A.noise_sigma = (R)h->noise_sigma; A.noise_seed = h->noise_seed; A.noise_ctr = (unsigned int)h->noise_steps;
    // EXT instantiations: Merge env, wind field, observation noise (fp32: the symmetric sweep only, N <= 64)
    const bool ext = is_ext(h);
    int32_t *done_now = nullptr;
    // auto-reset: the kernel stores `done` twice -- to the caller's array and to the library's flag buffer of the next
    // launch -- instead of a device-to-device copy node behind every step
    A.done_next = nullptr;
    if (ar) { done_now = h->done_buf[h->flip ^ 1]; if (!A.out.done) A.out.done = done_now; else if (A.out.done != done_now) A.done_next = done_now; }
    const int grid = (g.num_envs + h->epb - 1) / h->epb;
    const bool reso = g.reso_on && g.asas_on;
    const dim3 gd(grid), bd(((h->epb * g.max_ac + 31) / 32) * 32);      // whole warps covering the CTA's envs
    cudaError_t le;
    if constexpr (sizeof(R) == 4) {
        if (ext) {
            if (g.max_ac > 64) { h->err = "EXT kernels (fp32): N <= 64"; return BSG_ERR_UNSUPPORTED; }
            if (reso && g.reso_mode != 0) { h->err = "EXT kernels (fp32): horizontal resolution method only (reso_mode 0)"; return BSG_ERR_UNSUPPORTED; }
            le = launch_step_f32_ext(h->tpb, reso, gd, bd, h->smem, s, g.device, A);
        } else {
            // symmetric half-ring sweep (fp32 fast path); the resolution methods with a vertical term run the ordered sweep
            const bool sym = g.max_ac <= 64 && !getenv("BSG_NO_SYM") && !(reso && g.reso_mode != 0);
            if (sym && !reso && h->wepb > 0) {
                // detection only: the one-wave shape picked at creation (whole warps balanced over the SMs)
                A.epb = h->wepb; A.nbuf = h->wnbuf;
                const dim3 wg((g.num_envs + h->wepb - 1) / h->wepb), wb(((h->wepb * g.max_ac + 31) / 32) * 32);
                le = launch_step_f32(640, false, true, wg, wb, h->wsmem, s, g.device, A);
            } else le = launch_step_f32(h->tpb, reso, sym, gd, bd, h->smem, s, g.device, A);
        }
    } else le = launch_step_f64(h->tpb, reso, ext, gd, bd, h->smem, s, g.device, A);
    if (le == cudaErrorInvalidValue && h->tpb != 128 && h->tpb != 256) { h->err = "step kernel: CTA of 128 or 256 threads"; return BSG_ERR_UNSUPPORTED; }
    h->launches++; h->noise_steps++;
    CK(h, le);
    if (ar) {
        h->flip ^= 1;
        if (h->autoreset == 2) {                       // same-step mode: reset now (separate masked launch)
            int rc = bsg_reset_impl(h, done_now, h->ar_seed, (void *)s, true);
            if (rc != BSG_OK) return rc;
        }
    }
    return BSG_OK;
}

extern "C" {

int bsg_set_autoreset(bsg_handle *h, int32_t enable, uint64_t seed) {
    if (!h) return BSG_ERR_ARG;
    CK(h, cudaSetDevice(h->cfg.device));
    for (int b = 0; b < 2; b++) {
        if (!h->done_buf[b]) CK(h, cudaMalloc(&h->done_buf[b], h->cfg.num_envs * sizeof(int32_t)));
        CK(h, cudaMemset(h->done_buf[b], 0, h->cfg.num_envs * sizeof(int32_t)));
    }
    if (!h->episode) {
        CK(h, cudaMalloc(&h->episode, h->cfg.num_envs * sizeof(uint32_t)));
        CK(h, cudaMemset(h->episode, 0, h->cfg.num_envs * sizeof(uint32_t)));
    }
    h->autoreset = enable == 2 ? 2 : (enable ? 1 : 0);
    h->ar_seed = seed;
    h->flip = 0;
    h->gen++;
    return BSG_OK;
}

int bsg_set_state(bsg_handle *h, const double *ac, const uint64_t *reso, const int32_t *nac, const double *ev) {
    if (!h) return BSG_ERR_ARG;
    const bsg_config &g = h->cfg;
    const size_t EN = (size_t)g.num_envs * g.max_ac;
    CK(h, cudaSetDevice(g.device));
    if (ac) {
        CK(h, cudaMemcpy(h->stage64, ac, (size_t)g.num_envs * BSG_NF * g.max_ac * sizeof(double), cudaMemcpyHostToDevice));
        const int tb = 256, gb = (int)((EN + tb - 1) / tb);
        if (g.precision == BSG_FP64) import_state_kernel<double><<<gb, tb>>>(h->stage64, g.num_envs, g.max_ac, h->pos, (double *)h->st);
        else import_state_kernel<float><<<gb, tb>>>(h->stage64, g.num_envs, g.max_ac, h->pos, (float *)h->st);
        h->launches++;
        CK(h, cudaGetLastError());
    }
    if (reso) {
        if (g.precision != BSG_FP64 && g.max_ac <= 64) {
            // the fp32 symmetric sweep evaluates every unordered pair once for both aircraft: it keeps the resopairs
            // matrix symmetric (conflicts are, and so is resumenav's decision) and expects it symmetric on intake.
            // BlueSky's own resopairs are symmetric up to rounding at a threshold; an imported row bit (i, j) without
            // its mirror is mirrored here.
            std::vector<uint64_t> sym(reso, reso + EN);
            const int N = g.max_ac;
            for (size_t e = 0; e < (size_t)g.num_envs; e++)
                for (int i = 0; i < N; i++)
                    for (uint64_t r = reso[e * N + i] & (N < 64 ? (1ull << N) - 1ull : ~0ull); r; r &= r - 1) {
                        const int j = __builtin_ctzll(r);
                        if (j != i) sym[e * N + j] |= 1ull << i;
                    }
            CK(h, cudaMemcpy(h->reso, sym.data(), EN * sizeof(uint64_t), cudaMemcpyHostToDevice));
        } else CK(h, cudaMemcpy(h->reso, reso, (size_t)((g.max_ac + 63) / 64) * EN * sizeof(uint64_t), cudaMemcpyHostToDevice));
    }
    if (nac) CK(h, cudaMemcpy(h->nac, nac, g.num_envs * sizeof(int32_t), cudaMemcpyHostToDevice));
    if (ev) CK(h, cudaMemcpy(h->ev, ev, (size_t)g.num_envs * BSG_NE * sizeof(double), cudaMemcpyHostToDevice));
    CK(h, cudaDeviceSynchronize());
    return BSG_OK;
}

int bsg_get_state(bsg_handle *h, double *ac, uint64_t *reso, int32_t *nac, double *ev) {
    if (!h) return BSG_ERR_ARG;
    const bsg_config &g = h->cfg;
    const size_t EN = (size_t)g.num_envs * g.max_ac;
    CK(h, cudaSetDevice(g.device));
    CK(h, cudaDeviceSynchronize());
    if (ac) {
        CK(h, cudaMemset(h->stage64, 0, (size_t)g.num_envs * BSG_NF * g.max_ac * sizeof(double)));
        const int tb = 256, gb = (int)((EN + tb - 1) / tb);
        const int derive_gs = !is_ext(h);
        if (g.precision == BSG_FP64) export_state_kernel<double><<<gb, tb>>>(h->stage64, g.num_envs, g.max_ac, h->pos, (const double *)h->st, derive_gs);
        else export_state_kernel<float><<<gb, tb>>>(h->stage64, g.num_envs, g.max_ac, h->pos, (const float *)h->st, derive_gs);
        h->launches++;
        CK(h, cudaGetLastError());
        CK(h, cudaMemcpy(ac, h->stage64, (size_t)g.num_envs * BSG_NF * g.max_ac * sizeof(double), cudaMemcpyDeviceToHost));
    }
    if (reso) CK(h, cudaMemcpy(reso, h->reso, (size_t)((g.max_ac + 63) / 64) * EN * sizeof(uint64_t), cudaMemcpyDeviceToHost));
    if (nac) CK(h, cudaMemcpy(nac, h->nac, g.num_envs * sizeof(int32_t), cudaMemcpyDeviceToHost));
    if (ev) CK(h, cudaMemcpy(ev, h->ev, (size_t)g.num_envs * BSG_NE * sizeof(double), cudaMemcpyDeviceToHost));
    return BSG_OK;
}

int bsg_step(bsg_handle *h, const void *actions_dev, const bsg_step_out *out, void *stream) {
    if (!h) return BSG_ERR_ARG;
    if (h->cfg.env_kind != BSG_KIND_RAW && !actions_dev) { h->err = "bsg_step: actions_dev is NULL"; return BSG_ERR_ARG; }
    CK(h, cudaSetDevice(h->cfg.device));
    if (h->cfg.precision == BSG_FP64) return launch_step<double>(h, actions_dev, out, h->cfg.substeps, 0, (cudaStream_t)stream);
    return launch_step<float>(h, actions_dev, out, h->cfg.substeps, 0, (cudaStream_t)stream);
}

int bsg_observe(bsg_handle *h, void *obs_dev, void *stream) {
    if (!h || !obs_dev) return BSG_ERR_ARG;
    if (h->cfg.env_kind == BSG_KIND_RAW) return BSG_ERR_UNSUPPORTED;
    CK(h, cudaSetDevice(h->cfg.device));
    bsg_step_out o;
    memset(&o, 0, sizeof(o));
    o.obs = obs_dev;
    if (h->cfg.precision == BSG_FP64) return launch_step<double>(h, nullptr, &o, 0, 1, (cudaStream_t)stream);
    return launch_step<float>(h, nullptr, &o, 0, 1, (cudaStream_t)stream);
}

static void host_block_views(bsg_handle *h, void *blk, void **obs, void **reward, void **info, int32_t **terminated, int32_t **truncated) {
    const size_t rs = rsize(h), E = h->cfg.num_envs;
    char *s = (char *)blk;
    if (obs) *obs = s;
    s += E * h->obs_dim * rs;
    if (reward) *reward = s;
    s += E * rs;
    if (info) *info = s;
    s += E * 4 * rs;
    if (terminated) *terminated = (int32_t *)s;
    s += E * sizeof(int32_t);
    if (truncated) *truncated = (int32_t *)s;
}

int bsg_host_views(bsg_handle *h, void **obs, void **reward, void **info, int32_t **terminated, int32_t **truncated) {
    if (!h) return BSG_ERR_ARG;
    host_block_views(h, h->h_outs[h->h_last], obs, reward, info, terminated, truncated);
    return BSG_OK;
}

int bsg_step_host_begin(bsg_handle *h, const void *actions) {
    if (!h || !actions) return BSG_ERR_ARG;
    if (h->h_pending >= 2) { h->err = "bsg_step_host_begin: two steps already in flight (call bsg_step_host_wait)"; return BSG_ERR_ARG; }
    const bsg_config &g = h->cfg;
    const size_t rs = rsize(h), E = g.num_envs;
    CK(h, cudaSetDevice(g.device));
    const int q = h->h_head;
    memcpy(h->h_acts[q], actions, h->act_bytes);
    // direct mode: the kernel reads the actions from / stores the outputs to the pinned host block itself (unified
    // addressing); the stores are posted over PCIe while later waves of CTAs still compute.  It wins when the launch
    // has many waves (C3: +5 %) and loses on a one-wave launch (C2: -10 %, the stores all land at the end), so the
    // default follows the grid size; BSG_HOST_DIRECT=0/1 overrides.
    const char *dir_env = getenv("BSG_HOST_DIRECT");          // read per call (tests flip it)
    const int n_cta = (int)((E + h->epb - 1) / h->epb);
    const bool direct = dir_env ? atoi(dir_env) != 0 : n_cta >= 4096;
    const bool dir = direct && !(h->noise_sigma > 0);          // the noise epilogue reads the observation back
    // copy mode, steady configuration: the three stream operations of a step (H2D, kernel, D2H) are captured once per
    // (pinned block, auto-reset flag buffer) and replayed as ONE graph launch -- less host work per step, no gaps between
    // the three on the device.  Anything that changes what a captured launch holds bumps h->gen and forces a re-capture.
    // (Not with observation noise: its Philox counter is a kernel argument that changes every launch.)
    // the actions (4-8 bytes per env) are read by the kernel straight from the pinned block (unified addressing) instead of
    // a host-to-device copy node in front of it: e2e +3.2 % on C2 (profiles/ab_r2_host_graph.log); BSG_HOST_ACT_DIRECT=0/1
    const char *ad_env = getenv("BSG_HOST_ACT_DIRECT");
    const bool act_dir = dir || (ad_env ? atoi(ad_env) != 0 : true);
    if ((int)act_dir != h->act_dir_last) { h->act_dir_last = (int)act_dir; h->gen++; }       // (captured graphs hold the choice)
    const bool ar_next = h->autoreset == 1 && g.env_kind != BSG_KIND_RAW;
    // Measured (profiles/ab_r2_host_graph.log): e2e +3.2 % on C2, +2.4 % on the Vertical batch, -5 % on the Merge (EXT)
    // kernels -- so the plain instantiations replay a graph and the EXT ones keep the three calls (BSG_HOST_GRAPH=0/1 overrides)
    const char *gr_env = getenv("BSG_HOST_GRAPH");
    const bool use_graph = !dir && !(h->noise_sigma > 0) && h->autoreset != 2 && (gr_env ? atoi(gr_env) != 0 : !is_ext(h));
    const int fk = ar_next ? h->flip : 0;
    if (use_graph && h->hgraph[q][fk] && h->hgraph_gen[q][fk] == h->gen) {
        CK(h, cudaGraphLaunch(h->hgraph[q][fk], h->hs));
        h->launches++; h->noise_steps++;
        if (ar_next) h->flip ^= 1;                             // what launch_step does on the host side of a step
        CK(h, cudaEventRecord(h->hev[q], h->hs));
        h->h_head = q ^ 1; h->h_pending++;
        return BSG_OK;
    }
    bool capturing = false;
    if (use_graph) {
        if (h->hgraph[q][fk]) { cudaGraphExecDestroy(h->hgraph[q][fk]); h->hgraph[q][fk] = nullptr; }
        // (the kernel's shared-memory attribute is set by its first plain launch: capture only once the handle has stepped)
        capturing = h->launches > 0 && cudaStreamBeginCapture(h->hs, cudaStreamCaptureModeThreadLocal) == cudaSuccess;
    }
    // the stream operations of one step; an error inside returns from the lambda only (a capture is always ended below)
    auto enqueue = [&]() -> int {
        if (!act_dir) CK(h, cudaMemcpyAsync(h->d_act, h->h_acts[q], h->act_bytes, cudaMemcpyHostToDevice, h->hs));
        char *d = (char *)(dir ? h->h_outs[q] : h->d_out);
        bsg_step_out o;
        memset(&o, 0, sizeof(o));
        o.obs = d; d += E * h->obs_dim * rs;
        o.reward = d; d += E * rs;
        o.info = d; d += E * 4 * rs;
        o.terminated = (int32_t *)d; d += E * sizeof(int32_t);
        o.truncated = (int32_t *)d;
        const void *act = act_dir ? h->h_acts[q] : h->d_act;
        const int rc = g.precision == BSG_FP64 ? launch_step<double>(h, act, &o, g.substeps, 0, h->hs)
                                               : launch_step<float>(h, act, &o, g.substeps, 0, h->hs);
        if (rc != BSG_OK) return rc;
        if (!dir) CK(h, cudaMemcpyAsync(h->h_outs[q], h->d_out, h->out_bytes, cudaMemcpyDeviceToHost, h->hs));
        return BSG_OK;
    };
    const int flip0 = h->flip;
    const int64_t launches0 = h->launches, noise0 = h->noise_steps;
    int rc = enqueue();
    if (capturing) {                                           // nothing has run yet: instantiate and launch the capture
        cudaGraph_t gr = nullptr;
        cudaError_t ce = cudaStreamEndCapture(h->hs, &gr);     // whatever happened above, the stream leaves capture mode
        cudaGraphExec_t ex = nullptr;
        if (rc == BSG_OK && ce == cudaSuccess && gr) ce = cudaGraphInstantiate(&ex, gr, 0);
        if (gr) cudaGraphDestroy(gr);
        if (rc == BSG_OK && ce == cudaSuccess && ex) {
            h->hgraph[q][fk] = ex; h->hgraph_gen[q][fk] = h->gen;
            CK(h, cudaGraphLaunch(ex, h->hs));
        } else {
            // the capture failed and nothing was enqueued: forget it and issue the step as the plain calls
            if (ex) cudaGraphExecDestroy(ex);
            cudaGetLastError();
            h->flip = flip0; h->launches = launches0; h->noise_steps = noise0;
            rc = enqueue();
            if (rc != BSG_OK) return rc;
        }
    } else if (rc != BSG_OK) return rc;
    CK(h, cudaEventRecord(h->hev[q], h->hs));
    h->h_head = q ^ 1; h->h_pending++;
    return BSG_OK;
}

int bsg_step_host_wait(bsg_handle *h, void **obs, void **reward, void **info, int32_t **terminated, int32_t **truncated) {
    if (!h) return BSG_ERR_ARG;
    if (h->h_pending <= 0) { h->err = "bsg_step_host_wait: no step in flight"; return BSG_ERR_ARG; }
    const int q = h->h_pending == 2 ? h->h_head : h->h_head ^ 1;      // the oldest step in flight
    CK(h, cudaEventSynchronize(h->hev[q]));
    h->h_pending--; h->h_last = q;
    host_block_views(h, h->h_outs[q], obs, reward, info, terminated, truncated);
    return BSG_OK;
}

// begin + wait in one call; returns the index (0 / 1) of the pinned block that holds the results, or a negative status.
// For bindings whose per-call overhead matters (ctypes): no pointer arguments to marshal, the views of the two blocks are
// built once from bsg_host_views.
int bsg_step_host_block(bsg_handle *h, const void *actions) {
    if (!h || !actions) return BSG_ERR_ARG;
    if (h->h_pending) { h->err = "bsg_step_host_block: steps begun with bsg_step_host_begin are still in flight"; return BSG_ERR_ARG; }
    int rc = bsg_step_host_begin(h, actions);
    if (rc != BSG_OK) return rc;
    rc = bsg_step_host_wait(h, nullptr, nullptr, nullptr, nullptr, nullptr);
    return rc != BSG_OK ? rc : h->h_last;
}
// bsg_step_host_wait that returns the block index instead of five pointers
int bsg_step_host_wait_block(bsg_handle *h) {
    const int rc = bsg_step_host_wait(h, nullptr, nullptr, nullptr, nullptr, nullptr);
    return rc != BSG_OK ? rc : h->h_last;
}

int bsg_step_host(bsg_handle *h, const void *actions, void *obs, void *reward, int32_t *term, int32_t *trunc, void *info) {
    if (!h || !actions) return BSG_ERR_ARG;
    if (h->h_pending) { h->err = "bsg_step_host: steps begun with bsg_step_host_begin are still in flight"; return BSG_ERR_ARG; }
    int rc = bsg_step_host_begin(h, actions);
    if (rc != BSG_OK) return rc;
    void *po, *pr, *pi; int32_t *pt, *pu;
    rc = bsg_step_host_wait(h, &po, &pr, &pi, &pt, &pu);
    if (rc != BSG_OK) return rc;
    const size_t rs = rsize(h), E = h->cfg.num_envs;
    if (obs) memcpy(obs, po, E * h->obs_dim * rs);
    if (reward) memcpy(reward, pr, E * rs);
    if (info) memcpy(info, pi, E * 4 * rs);
    if (term) memcpy(term, pt, E * sizeof(int32_t));
    if (trunc) memcpy(trunc, pu, E * sizeof(int32_t));
    return BSG_OK;
}

}  // extern "C"
