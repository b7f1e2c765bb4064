/* ORACLE (test infrastructure, not product code).
 * ND-map: 4D (t, z, y, x) spatial-temporal occupancy / conflict grid.  No reference
 * implementation exists (SURVEY.md 0.3 tier C; the package that would hold it is absent
 * from /root/reference): this restates the builder-owned spec of DESIGN.md "ND-map".
 * parity unpinned.  Sample arithmetic is fp32 with explicit fmaf so host and device
 * produce identical cell indices. */
#include <math.h>
#include <stdlib.h>
#include <string.h>
#include "bsgo.h"

void bsgo_ndmap(const bsgo_ndcfg *g, int E, int n, const float *rec, uint8_t *occ, uint8_t *conf,
                int32_t *ncell, int nthreads) {
    const int nx = g->nx, ny = g->ny, nz = g->nz, nt = g->nt;
    const size_t slab = (size_t)nx * ny * nz;
    if (nthreads < 1) nthreads = 1;
#pragma omp parallel for num_threads(nthreads) schedule(dynamic, 1) collapse(2)
    for (int e = 0; e < E; e++)
        for (int k = 0; k < nt; k++) {
            uint16_t *cnt = calloc(slab, sizeof(uint16_t));
            uint16_t *sx = malloc(slab * sizeof(uint16_t)), *sy = malloc(slab * sizeof(uint16_t));
            const float *r = rec + (size_t)e * 6 * n;
            float t = (float)(k + 1) * g->dt;
            for (int a = 0; a < n; a++) {
                float px = fmaf(r[3 * n + a], t, r[0 * n + a]);
                float py = fmaf(r[4 * n + a], t, r[1 * n + a]);
                float pz = fmaf(r[5 * n + a], t, r[2 * n + a]);
                int ix = (int)floorf((px - g->x0) * g->inv_cx);
                int iy = (int)floorf((py - g->y0) * g->inv_cy);
                int iz = (int)floorf((pz - g->z0) * g->inv_cz);
                if (ix < 0 || ix >= nx || iy < 0 || iy >= ny || iz < 0 || iz >= nz) continue;
                cnt[((size_t)iz * ny + iy) * nx + ix]++;
            }
            /* separable 3x3x3 box sum, clipped at the grid faces */
            for (int z = 0; z < nz; z++) for (int y = 0; y < ny; y++) for (int x = 0; x < nx; x++) {
                size_t c = ((size_t)z * ny + y) * nx + x;
                sx[c] = cnt[c] + (x > 0 ? cnt[c - 1] : 0) + (x < nx - 1 ? cnt[c + 1] : 0);
            }
            for (int z = 0; z < nz; z++) for (int y = 0; y < ny; y++) for (int x = 0; x < nx; x++) {
                size_t c = ((size_t)z * ny + y) * nx + x;
                sy[c] = sx[c] + (y > 0 ? sx[c - nx] : 0) + (y < ny - 1 ? sx[c + nx] : 0);
            }
            int32_t ncf = 0;
            uint8_t *o = occ + ((size_t)e * nt + k) * slab, *cf = conf + ((size_t)e * nt + k) * slab;
            for (int z = 0; z < nz; z++) for (int y = 0; y < ny; y++) for (int x = 0; x < nx; x++) {
                size_t c = ((size_t)z * ny + y) * nx + x;
                unsigned s = sy[c] + (z > 0 ? sy[c - (size_t)nx * ny] : 0) + (z < nz - 1 ? sy[c + (size_t)nx * ny] : 0);
                o[c] = cnt[c] > 255 ? 255 : (uint8_t)cnt[c];
                cf[c] = s >= 2;
                ncf += s >= 2;
            }
            ncell[(size_t)e * nt + k] = ncf;
            free(cnt); free(sx); free(sy);
        }
}
