/* CPU restatement of the world-coordinate voxel neighbour query of Point-NeRF.
 *
 * TEST INFRASTRUCTURE (see oracle/__init__.py): only tests/, __graft_entry__.smoke() and the
 * cpu_baseline / --impl reference legs of bench.py may load this.  It restates, for B = 1 (the only
 * batch size any caller uses, neural_points.py:648), the six pycuda kernels of
 *   models/neural_points/query_point_indices_worldcoords.py
 * executed under the CANONICAL SCHEDULE "threads run in ascending global thread index":
 *   pn_port_claim_occ      <- claim_occ                      :243-303
 *   pn_port_map_coor2occ   <- map_coor2occ                   :305-340
 *   pn_port_fill_occ2pnts  <- fill_occ2pnts                  :342-387
 *   pn_port_mask_raypos    <- mask_raypos                    :390-414
 *   pn_port_get_shadingloc <- get_shadingloc                 :417-439
 *   pn_port_query_along_ray<- query_neigh_along_ray_layered  :442-528
 * Arithmetic model (what nvcc emits for those lines, checked in the PTX for sm_100a):
 *   voxel coordinate = (int) floorf( (p - shift) / vsize )   sub.f32, div.rn.f32, cvt.rmi, cvt.rzi
 *   d^2              = fmaf(z, z, fmaf(x, x, y*y))            default -fmad=true contraction
 * The time-seeded reservoir branches (:290-298, :378-383) are NOT restated: a slot beyond max_o
 * or a bucket entry beyond P is dropped and counted in *overflow (documented divergence).
 * Checked against oracle/_ref (the reference's own kernel string compiled for the host) by
 * tests/test_oracle_querier.py.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>

static inline int vox_of(float p, float shift, float vsize) { return (int)floorf((p - shift) / vsize); }

/* returns the linear cell index of point/sample p, or -1 if outside the grid */
static inline int64_t cell_of(const float* p, const float* shift, const float* vsz, const int* gsz, int c[3]) {
    for (int a = 0; a < 3; ++a) {
        c[a] = vox_of(p[a], shift[a], vsz[a]);
        if (c[a] < 0 || c[a] >= gsz[a]) return -1;
    }
    return ((int64_t)c[0] * gsz[1] + c[1]) * gsz[2] + c[2];
}

/* Occupied voxels get slots in order of their first point (ascending point index). */
void pn_port_claim_occ(const float* xyz, int n_actual, const float* shift, const float* vsz, const int* gsz,
                       int max_o, int* occ_idx, int* coor_2_occ, int* occ_2_coor, int* overflow) {
    for (int i = 0; i < n_actual; ++i) {
        int c[3];
        int64_t cell = cell_of(xyz + 3 * (int64_t)i, shift, vsz, gsz, c);
        if (cell < 0 || coor_2_occ[cell] != -1) continue;
        coor_2_occ[cell] = 0; /* claimed marker, as the CAS(-1 -> 0) leaves it */
        int slot = (*occ_idx)++;
        if (slot < max_o) {
            occ_2_coor[3 * slot + 0] = c[0];
            occ_2_coor[3 * slot + 1] = c[1];
            occ_2_coor[3 * slot + 2] = c[2];
        } else if (overflow) {
            ++*overflow;
        }
    }
}

/* coor_2_occ was refilled with -1 by the host (worldcoords.py:575); write slot ids and dilate. */
void pn_port_map_coor2occ(const int* gsz, const int* dil, int max_o, const int* occ_idx, int* coor_occ,
                          int* coor_2_occ, const int* occ_2_coor) {
    int n = *occ_idx < max_o ? *occ_idx : max_o;
    for (int s = 0; s < n; ++s) {
        const int* c = occ_2_coor + 3 * s;
        if (c[0] < 0) continue;
        coor_2_occ[((int64_t)c[0] * gsz[1] + c[1]) * gsz[2] + c[2]] = s;
        int lo[3], hi[3];
        for (int a = 0; a < 3; ++a) {
            lo[a] = c[a] - dil[a] / 2;
            if (lo[a] < 0) lo[a] = 0;
            hi[a] = c[a] + (dil[a] + 1) / 2;
            if (hi[a] > gsz[a]) hi[a] = gsz[a];
        }
        for (int x = lo[0]; x < hi[0]; ++x)
            for (int y = lo[1]; y < hi[1]; ++y)
                for (int z = lo[2]; z < hi[2]; ++z) coor_occ[((int64_t)x * gsz[1] + y) * gsz[2] + z] = 1;
    }
}

/* Buckets hold point indices in ascending order, first P only.  NOTE `slot > 0` (:372): the voxel
 * that owns slot 0 never receives points - reference behaviour, kept. */
void pn_port_fill_occ2pnts(const float* xyz, int n_actual, int P, const float* shift, const float* vsz,
                           const int* gsz, const int* coor_2_occ, int* occ_2_pnts, int* occ_numpnts,
                           int* overflow) {
    for (int i = 0; i < n_actual; ++i) {
        int c[3];
        int64_t cell = cell_of(xyz + 3 * (int64_t)i, shift, vsz, gsz, c);
        if (cell < 0) continue;
        int slot = coor_2_occ[cell];
        if (slot <= 0) continue;
        int pos = occ_numpnts[slot]++;
        if (pos < P)
            occ_2_pnts[(int64_t)slot * P + pos] = i;
        else if (overflow)
            ++*overflow;
    }
}

/* raypos [R*D,3]; raypos_mask pre-zeroed; samples outside the grid keep 0. */
void pn_port_mask_raypos(const float* raypos, const int* coor_occ, int64_t n_samples, const float* shift,
                         const int* gsz, const float* vsz, int* raypos_mask) {
    for (int64_t i = 0; i < n_samples; ++i) {
        int c[3];
        int64_t cell = cell_of(raypos + 3 * i, shift, vsz, gsz, c);
        if (cell >= 0) raypos_mask[i] = coor_occ[cell];
    }
}

/* slot_of_sample [R,D] = rank among occupied samples of the ray (0-based) or -1 (host cumsum, :664-665) */
void pn_port_get_shadingloc(const float* raypos, const int* slot_of_sample, int R, int D, int SR,
                            float* sample_loc, int* sample_loc_mask) {
    for (int r = 0; r < R; ++r)
        for (int d = 0; d < D; ++d) {
            int s = slot_of_sample[(int64_t)r * D + d];
            if (s < 0) continue;
            const float* p = raypos + 3 * ((int64_t)r * D + d);
            float* q = sample_loc + 3 * ((int64_t)r * SR + s);
            q[0] = p[0]; q[1] = p[1]; q[2] = p[2];
            sample_loc_mask[(int64_t)r * SR + s] = 1;
        }
}

/* Layered, radius-limited K "nearest" of every kept sample.  Visit order: layer (Chebyshev ring),
 * then x, y, z ascending, then bucket position.  Slot arrangement and the early exit
 * (`kid >= K` tested only after a whole layer, kid counting every in-radius candidate) follow
 * :482-527 exactly, so the result is NOT the exact K-NN - it is what the reference returns.
 * n_candidates (optional) accumulates the bucket entries visited (for the roofline byte count). */
void pn_port_query_along_ray(const float* xyz, int64_t n_slots, int P, int K, float radius_limit2,
                             const float* shift, const int* gsz, const float* vsz, const int* ksz,
                             const int* occ_numpnts, const int* occ_2_pnts, const int* coor_2_occ,
                             const float* sample_loc, const int* sample_loc_mask, int* sample_pidx,
                             int64_t* n_candidates) {
    float* best = (float*)malloc(sizeof(float) * (size_t)(K > 0 ? K : 1));
    int layers = (ksz[0] + 1) / 2; /* only kernel_size[0] drives all three axes (:482) */
    int64_t visited = 0;
    for (int64_t s = 0; s < n_slots; ++s) {
        if (sample_loc_mask[s] <= 0) continue;
        const float* q = sample_loc + 3 * s;
        int f[3];
        for (int a = 0; a < 3; ++a) f[a] = vox_of(q[a], shift[a], vsz[a]);
        int* out = sample_pidx + s * K;
        int seen = 0, far_slot = 0;
        float far2 = 0.0f;
        for (int layer = 0; layer < layers; ++layer) {
            int lo[3], hi[3];
            for (int a = 0; a < 3; ++a) {
                lo[a] = -f[a] > -layer ? -f[a] : -layer;
                hi[a] = gsz[a] - f[a] < layer + 1 ? gsz[a] - f[a] : layer + 1;
            }
            for (int dx = lo[0]; dx < hi[0]; ++dx)
                for (int dy = lo[1]; dy < hi[1]; ++dy)
                    for (int dz = lo[2]; dz < hi[2]; ++dz) {
                        int ring = abs(dx) > abs(dy) ? abs(dx) : abs(dy);
                        if (abs(dz) > ring) ring = abs(dz);
                        if (ring != layer) continue;
                        int slot = coor_2_occ[((int64_t)(f[0] + dx) * gsz[1] + (f[1] + dy)) * gsz[2] + (f[2] + dz)];
                        if (slot < 0) continue;
                        int cnt = occ_numpnts[slot] < P ? occ_numpnts[slot] : P;
                        for (int g = 0; g < cnt; ++g) {
                            int pid = occ_2_pnts[(int64_t)slot * P + g];
                            const float* p = xyz + 3 * (int64_t)pid;
                            float x = p[0] - q[0], y = p[1] - q[1], z = p[2] - q[2];
                            float d2 = fmaf(z, z, fmaf(x, x, y * y));
                            ++visited;
                            if (!(radius_limit2 == 0.0f || d2 <= radius_limit2)) continue;
                            if (seen < K) {
                                out[seen] = pid;
                                best[seen] = d2;
                                if (d2 > far2) { far2 = d2; far_slot = seen; }
                                ++seen;
                            } else {
                                ++seen;
                                if (d2 < far2) {
                                    out[far_slot] = pid;
                                    best[far_slot] = d2;
                                    far2 = d2;
                                    for (int i = 0; i < K; ++i)
                                        if (best[i] > far2) { far2 = best[i]; far_slot = i; }
                                }
                            }
                        }
                    }
            if (seen >= K) break;
        }
    }
    if (n_candidates) *n_candidates += visited;
    free(best);
}
