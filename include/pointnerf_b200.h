/* pointnerf_b200 - C ABI of the B200-native Point-NeRF per-ray hot path.
 *
 * One shared library (pointnerf_b200/lib/libpointnerf_b200.so, nvcc -arch=sm_100a) replaces the
 * pycuda-JIT kernels and the torch op chains of the reference (yjcaimeow/pointnerf) on the path
 *   lighting_fast_querier.query_points -> NeuralPoints.forward -> PointAggregator.forward -> ray_march.
 * The reference has no FFI: pycuda launches are inline Python
 * (models/neural_points/query_point_indices_worldcoords.py:558-708).  Each entry point below names the
 * reference function(s) whose work it does; INTEGRATION.md shows the ctypes binding a maintainer adds.
 *
 * Conventions
 *   - plain C: raw DEVICE pointers, extents, POD structs; no torch / C++ types cross the boundary;
 *   - the CALLER owns every buffer, including workspaces (sized by the *_workspace_bytes queries);
 *     the library keeps no state besides a thread-local error string;
 *   - all work is enqueued asynchronously on `stream` (a cudaStream_t passed as void*; torch's
 *     current stream in the Python host); no call synchronises or allocates;
 *   - return value: 0 = ok, negative = error (pn_last_error() has the message);
 *   - B = 1 everywhere (the reference's only batch size, neural_points.py:648);
 *   - float = IEEE binary32, int = int32; layouts are C-contiguous, innermost dimension last.
 */
#ifndef POINTNERF_B200_H
#define POINTNERF_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define PN_OK 0
#define PN_ERR_ARG (-1)
#define PN_ERR_CUDA (-2)
#define PN_ERR_UNSUPPORTED (-3)

const char* pn_last_error(void);
/* ABI version; bumped on any signature change. */
int pn_abi_version(void);
/* Number of kernel launches this library has enqueued since load / since the last reset
 * (process-wide counter; bench.py reports it as `gpu_launches`). */
int64_t pn_launch_count(void);
void pn_launch_count_reset(void);

/* The two GEMM forms of the aggregator MLP (they replace the cuBLAS SGEMMs behind nn.Linear in point_aggregators.py:547,571,
 * 605,635 and their autograd), exposed for kernel-level tests and the roofline leg of bench.py.  tcgen05 tensor cores,
 * 3xTF32 split products (hi*lo + lo*hi + hi*hi) with fp32 accumulation in TMEM: fp32-class accuracy (~2e-6 of max).
 *   nt: C[M,N] = act(A[M,Kd] * B[N,Kd]^T + bias)       (forward layers, data gradients; persistent CTA pairs)
 *   tn: C[N,Kb] = A[M,N]^T * B[M,Kb], colsum[N] = sum_m A[m,:]   (weight / bias gradients; split over M, fixed-order sum)
 * M_dev (device int32, may be NULL): the row count is min(*M_dev, M) and never reaches the host; M is then the capacity.
 * Operands 16-byte aligned, leading dimensions multiples of 4.  image_scratch: pn_gemm_nt_image_bytes bytes (B is packed into
 * the pre-split, pre-swizzled operand image that TMA bulk copies stream); scratch: pn_gemm_tn_scratch_floats floats. */
int64_t pn_gemm_nt_image_bytes(int32_t N, int32_t Kd);
int pn_gemm_nt(const float* A, int64_t lda, const float* B, int64_t ldb, float* C, int64_t ldc, int64_t M, const int32_t* M_dev,
               int32_t N, int32_t Kd, const float* bias, int32_t leaky, float slope, void* image_scratch, void* stream);
int64_t pn_gemm_tn_scratch_floats(int64_t M, int32_t N, int32_t Kb);
int pn_gemm_tn(const float* A, int64_t lda, const float* B, int64_t ldb, float* C, int64_t ldc, int32_t Kout,
               float* colsum, int64_t M, const int32_t* M_dev, int32_t N, int32_t Kb, float* scratch, void* stream);

/* ------------------------------------------------------------------------------------------------
 * 1. Voxel grid  (replaces claim_occ :243-303, map_coor2occ :305-340, fill_occ2pnts :342-387 and the
 *    fills/launches of build_occ_vox :546-611 of query_point_indices_worldcoords.py)
 *
 * Representation in HBM (V = dims[0]*dims[1]*dims[2] cells, linear index (x*dims[1]+y)*dims[2]+z):
 *   cell_start int32[V+1]  exclusive prefix of points per cell: cell c owns
 *                          bucket_pidx[cell_start[c] .. cell_start[c+1])  (three z-neighbour cells = 4
 *                          consecutive ints);
 *   bucket_pidx int32[N]   point indices grouped by cell, ASCENDING point index inside a cell
 *                          (= bucket order of the reference under the canonical schedule);
 *   occ_bits uint32[(V+31)/32]  occupancy dilated by `dilate` (the reference's coor_occ);
 *   info int32[PN_GRID_INFO_LEN]  see enum below.
 * The reference's quirk `if (voxel_idx > 0)` (:372) - the voxel that owns slot 0 (canonically the voxel
 * of the lowest-index in-grid point) never receives points - is kept through info[PN_GI_SLOT0_CELL].
 * Buckets are exposed to queries truncated to the first P points by index; occupied voxels beyond
 * max_o are kept (the reference drops/replaces them at random; documented divergence) and flagged.
 * ---------------------------------------------------------------------------------------------- */
typedef struct pn_grid_params {
    float shift[3];  /* ranges[:3] of get_hyperparameters (:71)                     */
    float vsize[3];  /* scaled voxel size  vsize*vscale (:60)                       */
    int32_t dims[3]; /* scaled_vdim (:75)                                           */
    int32_t dilate[3]; /* query_size: dilation box of map_coor2occ (:330-332, :626) */
    int32_t P;       /* bucket capacity                                             */
    int32_t max_o;   /* occupied-voxel budget of the reference (only reported)      */
} pn_grid_params;

enum {
    PN_GI_N_OCC = 0,      /* occupied cells                                              */
    PN_GI_FIRST_POINT,    /* lowest point index that falls inside the grid (INT_MAX: none) */
    PN_GI_SLOT0_CELL,     /* linear cell index that owns the reference's slot 0 (-1: none) */
    PN_GI_N_INSIDE,       /* points inside the grid                                      */
    PN_GI_MAX_COUNT,      /* largest per-cell point count                                */
    PN_GI_OVER_P,         /* cells holding more than P points                            */
    PN_GI_OVER_MAX_O,     /* 1 if n_occ > max_o                                          */
    PN_GI_RESERVED,
    PN_GRID_INFO_LEN
};

/* bytes of scratch pn_grid_build needs for N points in a grid of V cells */
int64_t pn_grid_workspace_bytes(int64_t n_points, int64_t n_cells);

int pn_grid_build(const float* xyz, int64_t n_points, const pn_grid_params* params,
                  int32_t* cell_start, int32_t* bucket_pidx, uint32_t* occ_bits, int32_t* info,
                  void* workspace, int64_t workspace_bytes, void* stream);

/* ------------------------------------------------------------------------------------------------
 * 2. Ray sampling  (replaces the raypos materialisation of near_far_linear_ray_generation
 *    diff_ray_marching.py:387, mask_raypos :390-414, the host max/cumsum/select :655-666 and
 *    get_shadingloc :417-439)
 *
 * One warp per ray walks the D candidate depths `ts` (middle_point_ts, computed by the caller with
 * the reference's own linspace/cumsum chain so that sample positions are bit-identical), forms
 * pos = campos + raydir * t with separately rounded mul and add, tests the dilated occupancy bit and
 * keeps the first SR hits.
 *   ts            float[ts_rows, D], ts_rows = 1 (no jitter: same depths for every ray) or R
 *   sample_loc_w  float[R, SR, 3]  world positions of kept samples, 0 for unused slots (as :657)
 *   sample_t      float[R, SR] or NULL  depth t of kept samples (0 unused)
 *   hit_count     int32[R]  number of kept samples (<= SR); ray_mask of :655 is hit_count > 0
 * ---------------------------------------------------------------------------------------------- */
/* Jittered candidate depths (the per-ray part of near_far_linear_ray_generation, diff_ray_marching.py:369-386):
 *   seg = seg_base[d] * (1 + jitter * (uniform[r,d] - 0.5)); ends = near + cumsum(seg), the running sum accumulated in
 *   DOUBLE left to right and rounded to float per element - what torch.cumsum computes on the host, the canonical chain;
 *   ts[r,d] = (ends[d-1] + ends[d]) / 2 with ends[-1] = near.  seg_base float[D] = t[1:] - t[:-1] of the linspace chain
 *   (caller, D+1 elements); uniform float[R,D] stands for the torch.rand draw of :373 (NULL: no jitter); ts float[R,D]. */
int pn_sample_depths(const float* seg_base, const float* uniform, int64_t n_rays, int32_t D, float near_depth,
                     float jitter, float* ts, void* stream);
int pn_ray_sample(const float* campos, const float* raydir, const float* ts, int64_t ts_rows,
                  int64_t n_rays, int32_t D, int32_t SR, const pn_grid_params* params,
                  const uint32_t* occ_bits, float* sample_loc_w, float* sample_t, int32_t* hit_count,
                  void* stream);

/* ------------------------------------------------------------------------------------------------
 * 3. Neighbour query  (replaces query_neigh_along_ray_layered :442-528)
 *
 * One warp per kept sample.  Candidates are visited in the reference's order - Chebyshev layer, then
 * x, y, z ascending, then bucket position - with d^2 = fma(z,z,fma(x,x,y*y)) (what nvcc emits for :498);
 * the K slots are filled and replaced exactly as :499-519 (so slot ARRANGEMENT matches, not only
 * the set), and the search stops after the first layer that has seen >= K in-radius candidates (:526).
 *   n_layers     (kernel_size[0]+1)/2   (:482 - only the x kernel size is used by the reference)
 *   sample_pidx  int32[R, SR, K], -1 = empty (also written for unused slots)
 *   ray_any      int32[R]  1 if the ray has at least one neighbour (masked_valid_ray of :715)
 *   n_candidates uint64[1] or NULL: += bucket entries visited (roofline byte count)
 * ---------------------------------------------------------------------------------------------- */
int pn_knn_query(const float* xyz, const pn_grid_params* params, const int32_t* cell_start,
                 const int32_t* bucket_pidx, const int32_t* grid_info, const float* sample_loc_w,
                 const int32_t* hit_count, int64_t n_rays, int32_t SR, int32_t K, int32_t n_layers,
                 float radius_limit2, int32_t* sample_pidx, int32_t* ray_any,
                 unsigned long long* n_candidates, void* stream);

/* ------------------------------------------------------------------------------------------------
 * 3b. Perspective-frustum querier  (replaces get_occ_vox :265-314, near_vox_full :316-365,
 *     insert_vox_points :368-413, query_neigh_along_ray_layered :495-600 and the fills / cumsum / launches of
 *     query_grid_point_index :617-786 of models/neural_points/query_point_indices.py - `wcoord_query = 0`)
 *
 * The grid lives in the camera's perspective space (x/z, y/z, z): dims = ceil([w, h, z_depth_dim] / vscale), rebuilt for
 * every view.  Same CSR representation as section 1 (pn_grid_params: shift = ranges[:3], vsize = scaled voxel size,
 * dilate = query_size), with the two rules of this querier: occupancy comes from floor() voxel coordinates (:290-297),
 * bucket membership from (int) truncation (:394-399) restricted to voxels that hold a floor()-valid point.  The
 * reference's per-column slot table ([w, h, max_o] buckets of the voxels "needed" by this batch's rays, :697-727) is
 * replaced by the CSR over all claimed voxels; that is equivalent whenever query_size covers the cube the query visits
 * (query_size[0], [1] >= kernel_size[0] and query_size[2] >= kernel_size[2]; the host class refuses anything else,
 * because the reference's result then depends on which other rays are in the batch).  Buckets are the first P points
 * by index (the reference: arrival order, then a time-seeded reservoir); `inverse` (disparity-linear z) is not offered.
 *   pixel_idx  int32[R,2]  (x, y) of every ray
 *   sample_z   int32[R,SR] depth-voxel index of the kept samples (first SR occupied depths of the ray's column), -1 unused;
 *                          a ray whose mask is 0 keeps none (the reference drops such rays before the query, :691-695)
 *   hit_count  int32[R]    kept samples
 *   ray_mask   int32[R]    the reference's `far_id > 0` (:339)
 *   sample_loc float[R,SR,3] perspective coordinates of the sample centres, written for every slot as :532-540
 *   sample_pidx int32[R,SR,K], -1 = empty
 * ---------------------------------------------------------------------------------------------- */
typedef struct pn_frustum_query {
    float ray_vsize[2];   /* scaled voxel size / vscale, x and y (:707)                          */
    int32_t vscale[2];    /* opt.vscale x and y                                                  */
    float radius_limit2;  /* (radius_limit_scale * max(vsize_x, vsize_y))^2, 0 = off (:571)      */
    float depth_limit2;   /* (depth_limit_scale * vsize_z)^2, 0 = off                            */
    int32_t n_layers;     /* (kernel_size[0]+1)/2 (:548)                                         */
    int32_t z_layers;     /* (kernel_size[2]+1)/2 (:549)                                         */
    int32_t nn;           /* opt.NN: 1 = distances in perspective space, >= 2 = x, y scaled by z */
} pn_frustum_query;

int64_t pn_frustum_grid_workspace_bytes(int64_t n_points, int64_t n_cells);
int pn_frustum_grid_build(const float* xyz_pers, int64_t n_points, const pn_grid_params* params, int32_t* cell_start,
                          int32_t* bucket_pidx, uint32_t* occ_bits, int32_t* info, void* workspace,
                          int64_t workspace_bytes, void* stream);
/* vscale_xy: HOST int32[2] */
int pn_frustum_sample(const int32_t* pixel_idx, int64_t n_rays, int32_t SR, const pn_grid_params* params,
                      const int32_t* vscale_xy, const uint32_t* occ_bits, int32_t* sample_z, int32_t* hit_count,
                      int32_t* ray_mask, void* stream);
int pn_frustum_knn_query(const float* xyz_pers, const pn_grid_params* params, const pn_frustum_query* q,
                         const int32_t* cell_start, const int32_t* bucket_pidx, const int32_t* pixel_idx,
                         const int32_t* sample_z, const int32_t* hit_count, int64_t n_rays, int32_t SR, int32_t K,
                         int32_t* sample_pidx, float* sample_loc, unsigned long long* n_candidates, void* stream);

/* Camera-frame "perspective" coordinates (x/z, y/z, z) of world positions
 * (w2pers: worldcoords.py:102-109 for samples, neural_points.py:687-693 for points).
 * camrot = camrotc2w row-major 3x3, campos 3 floats, both on the device. */
int pn_w2pers(const float* xyz_w, int64_t n, const float* camrot, const float* campos, float* xyz_pers,
              void* stream);

/* Work lists for the aggregator: valid samples (>= 1 neighbour) and valid (sample, neighbour) pairs,
 * both in ascending (ray, slot, k) order - replaces the boolean-mask compactions (nonzero syncs) of
 * point_aggregators.py:521-538.
 *   sample_list int32[n_slots]    indices r*SR+s of valid samples (first S entries used)
 *   pair_start  int32[n_slots+1]  first compact pair of each valid sample; pair_start[S] = M
 *   pair_src    int32[n_slots*K]  slot*K+k of each compact pair (first M entries used)
 *   counts      int32[2]          [0] = S, [1] = M   (the caller reads these back: one host sync)
 *   slot_row    int32[n_slots] or NULL: inverse of sample_list (-1 for samples without neighbours)
 * workspace: pn_compact_workspace_bytes(n_slots). */
int64_t pn_compact_workspace_bytes(int64_t n_slots);
int pn_compact_samples(const int32_t* sample_pidx, int64_t n_slots, int32_t K, int32_t* sample_list,
                       int32_t* pair_start, int32_t* pair_src, int32_t* counts, int32_t* slot_row, void* workspace,
                       int64_t workspace_bytes, void* stream);

/* pn_compact_samples under CAPACITIES, for callers that keep the counts on the device (pn_agg_inputs.counts): buffers and
 * launches downstream are sized by S_cap samples / M_cap pairs instead of the read-back counts.  The lists keep the longest
 * prefix of valid samples (slot order) that fits both capacities; later samples are dropped whole (slot_row = -1: they
 * composite as empty space).  sample_list int32[min(n_slots, S_cap)], pair_start int32[min(n_slots, S_cap) + 1], pair_src
 * int32[min(n_slots * K, M_cap)]; counts4 int32[4] = {S kept, M kept, S total, M total}: kept < total means the batch
 * overflowed the capacities (the host checks it after the fact, e.g. from an asynchronous copy, and grows them). */
int pn_compact_samples_capped(const int32_t* sample_pidx, int64_t n_slots, int32_t K, int64_t S_cap, int64_t M_cap,
                              int32_t* sample_list, int32_t* pair_start, int32_t* pair_src, int32_t* counts4,
                              int32_t* slot_row, void* workspace, int64_t workspace_bytes, void* stream);

/* pn_compact_samples for callers that cannot allocate (code under CUDA-graph capture, tight serving loops, harnesses that
 * forbid cudaMalloc).  No counterpart in the reference.  Same inputs, outputs and
 * results as pn_compact_samples, but the workspace is a file-scope `static __device__` array inside the library (8.4 MB,
 * held by every GPU that loads the library until the process exits) instead of a caller-owned buffer; the call allocates
 * nothing and calls no allocating routine.  n_slots <= pn_compact_samples_static_max_slots() = 524 288 (4096 rays x 80
 * samples = 327 680), larger batches are rejected with PN_ERR_ARG.  All calls on a device share the one array: calls that
 * overlap on the GPU (different streams or host threads) race; serialise them or use pn_compact_samples. */
int64_t pn_compact_samples_static_max_slots(void);
int pn_compact_samples_static(const int32_t* sample_pidx, int64_t n_slots, int32_t K, int32_t* sample_list,
                              int32_t* pair_start, int32_t* pair_src, int32_t* counts, int32_t* slot_row, void* stream);

/* ------------------------------------------------------------------------------------------------
 * 4. Point aggregator  (replaces the index_select gather of NeuralPoints.forward neural_points.py:788-799
 *    and PointAggregator.forward / linear / viewmlp, point_aggregators.py:729-816, 421-429, 488-646,
 *    for the shipped configuration: viewmlp, agg_intrp_order=2, linear kernel, agg_dist_pers=20,
 *    apply_pnt_mask=1, LeakyReLU(0.01), act_super=1)
 *
 * The kernels read per-point rows THROUGH an index (`sample_pidx`): with the neural point tables and
 * point indices this is the fused gather path; with the already gathered [R,SR,K,*] tensors of the
 * reference API and identity indices it is the drop-in PointAggregator.forward.
 * Work lists come from pn_compact_samples; S, M are read back by the caller (one host sync per call,
 * where the reference has eight).
 * ---------------------------------------------------------------------------------------------- */
typedef struct pn_agg_params {
    int32_t F;          /* point_features_dim (32)                    */
    int32_t feat_freqs; /* num_feat_freqs (3)                         */
    int32_t dist_freqs; /* dist_xyz_freq (5); distance vector is 6-D  */
    int32_t view_freqs; /* num_viewdir_freqs (4)                      */
    int32_t W;          /* shading_feature_num (256)                  */
    int32_t C;          /* shading_color_channel_num (128 in this fork) */
    int32_t K;          /* neighbours per sample                      */
    int32_t dirs_per_ray; /* 1: ray_dirs is [R,3] (row = slot / SR); 0: [n_slots,3] */
    int32_t SR;
    float leaky_slope;  /* nn.LeakyReLU default 0.01                  */
    float dist_scale;   /* 1 if dist_xyz_deno == 0 else 1/(deno*|vsize|) (:524-525) */
    int32_t compact_out; /* 0: decoded_features / grad_decoded are [n_slots, C+1] (reference layout, zero rows for
                            invalid samples); 1: [S, C+1], row i = sample sample_list[i] (fused path, see
                            pn_composite_forward slot_row) */
    int32_t intrp_order; /* agg_intrp_order: 2 (or 0) = one density per neighbour, interpolated over K (:603-646, every shipped
                            per-scene script); 1 = features interpolated over K, then one density per sample and
                            raw2out_color (sigmoid, act_super scaling) on the colour (:573-599, the DTU feed-forward configs) */
    int32_t forward_only; /* 1: no pn_agg_backward follows on this `saved` buffer (inference): the forward skips what only the
                             backward reads (the LeakyReLU sign bits its GEMM epilogues pack, ~6 % of the forward) */
} pn_agg_params;

/* nn.Linear parameters, weight [out,in] row-major fp32 (state_dict names of the reference) */
typedef struct pn_agg_weights {
    float *block1_0_w, *block1_0_b; /* [W, F+2*ff*F+12*df]  */
    float *block1_2_w, *block1_2_b; /* [W, W]               */
    float *block3_0_w, *block3_0_b; /* [W, W+7]             */
    float *block3_2_w, *block3_2_b; /* [W, W]               */
    float *alpha_0_w, *alpha_0_b;   /* [1, W]               */
    float *color_0_w, *color_0_b;   /* [C, W+6*vf]          */
    float *color_2_w, *color_2_b;   /* [C, C]               */
    float *color_4_w, *color_4_b;   /* [C, C]               */
} pn_agg_weights;

typedef struct pn_agg_inputs {
    const float* xyz;       /* [T,3] world position of row g                          */
    const float* xyz_pers;  /* [T,3] (x/z, y/z, z) in the camera frame                */
    const float* embedding; /* [T,F]                                                  */
    const float* color;     /* [T,3]                                                  */
    const float* dir;       /* [T,3]                                                  */
    const float* conf;      /* [T]                                                    */
    const float* Rw2c;      /* [9] device, row-major                                  */
    const int32_t* sample_pidx; /* [n_slots,K] row index g or -1                      */
    const float* sample_loc_w;  /* [n_slots,3]                                        */
    const float* sample_loc;    /* [n_slots,3] perspective coords of the samples      */
    const float* ray_dirs;      /* see dirs_per_ray                                   */
    const int32_t* sample_list; /* [S] valid slots                                    */
    const int32_t* pair_start;  /* [S+1] first compact pair of each valid sample      */
    const int32_t* pair_src;    /* [M] slot*K+k of each compact pair                  */
    int64_t n_slots, S, M, T;
    /* counts: NULL - S and M above are the exact numbers of valid samples / pairs (the caller read them back);
     *         device int32[2] = {S, M} as written by pn_compact_samples - the counts stay on the device, S and M above are
     *         CAPACITIES (>= the true counts; every buffer sized by S / M is sized by them) and the whole pass is a fixed
     *         launch sequence with no host read-back. */
    const int32_t* counts;
} pn_agg_inputs;

int64_t pn_agg_saved_bytes(const pn_agg_params* p, int64_t S, int64_t M);
int64_t pn_agg_workspace_bytes(const pn_agg_params* p, int64_t S, int64_t M, int64_t T, int backward);

/* decoded_features [n_slots, C+1]: column 0 density, 1..C colour feature; rows of invalid samples are
 * zero-filled.  weight_out [n_slots,K] (normalised inverse-distance weights) may be NULL.
 * saved: activations kept for pn_agg_backward (NULL = inference). */
int pn_agg_forward(const pn_agg_params* p, const pn_agg_inputs* in, const pn_agg_weights* w,
                   float* decoded_features, float* weight_out, void* saved, int64_t saved_bytes,
                   void* workspace, int64_t workspace_bytes, void* stream);

/* grad_decoded [n_slots, C+1].  grad_w receives d/d(MLP parameters) (overwritten, deterministic
 * fixed-order reductions).  Per-row gradients of embedding/color/dir/conf:
 *   scatter_rows == 0: written per PAIR into [n_slots*K, *] arrays (zeros for empty slots) - the
 *                      gradients of the reference's sampled_* tensors;
 *   scatter_rows == 1: summed per table row g into [T, *] arrays by a sorted segmented reduction (no
 *                      atomics; replaces the index_add_ backward of index_select), rows never referenced
 *                      are zero-filled;
 *   scatter_rows == 2: as 1, but the [T, *] arrays are NOT zero-filled: the caller guarantees they are zero outside the
 *                      rows this call writes (pn_rows_zero). */
/* Streams: with the default schedule (pn_set_backward_schedule) part of the work runs on a library-owned stream that is
 * forked from and joined back into `stream` by events, so for the caller everything is still ordered on `stream`: work
 * queued on it afterwards sees every output, and the call never blocks the host. */
int pn_agg_backward(const pn_agg_params* p, const pn_agg_inputs* in, const pn_agg_weights* w,
                    const float* grad_decoded, const void* saved, int64_t saved_bytes,
                    const pn_agg_weights* grad_w, int scatter_rows, float* grad_embedding,
                    float* grad_color, float* grad_dir, float* grad_conf, int32_t* touched_ids,
                    int32_t* n_touched, void* workspace, int64_t workspace_bytes, void* stream);

/* LeakyReLU activation pattern of a saved forward (what pn_agg_backward will differentiate through): 1 where the unit's
 * output is > 0 (derivative 1), 0 where the derivative is leaky_slope.  pair_signs uint8[4, M, W]: block1.0, block1.2,
 * block3.0, block3.2 (compact pair rows, ascending (slot, k)); sample_signs uint8[3, S, C]: color_branch.0 / .2 / .4 (compact
 * sample rows).  S_cap / M_cap: the S / M the saved buffer was sized with.  The MLP is piecewise linear, so a reference backward evaluated under the SAME pattern is comparable at the
 * full tolerance on every ray, units at a kink included (point_aggregators.py:547-571, 635). */
int pn_agg_activation_signs(const pn_agg_params* p, const void* saved, int64_t saved_bytes, int64_t S_cap, int64_t M_cap,
                            int64_t S, int64_t M, uint8_t* pair_signs, uint8_t* sample_signs, void* stream);

/* Makes `stream` wait for the point where the most recent pn_agg_backward on this device had produced its per-row
 * gradients (grad_embedding / grad_color / grad_dir / grad_conf, touched_ids, n_touched) - BEFORE the weight-gradient GEMMs
 * that trail on the caller's stream.  A multi-GPU host starts the exchange of the touched rows there (ray_shard.py), so
 * that it runs under the remaining GEMMs.  Returns 1 if the wait was queued, 0 if no such point exists (single-stream
 * backward: order `stream` after the caller's stream instead), < 0 on error. */
int pn_agg_backward_wait_rows(void* stream);
/* Kernel schedule of pn_agg_backward (process-wide; results are bit-identical in all three):
 *   0  everything on the caller's stream, each layer's weight gradient next to its data gradient;
 *   1  (default) the data-gradient chain on a library-owned high-priority stream, the weight-gradient GEMMs trailing it on
 *      the caller's stream: tails and prologues of the persistent GEMMs overlap (fastest on one GPU);
 *   2  the whole chain first, then the seven weight-gradient GEMMs: pn_agg_backward_wait_rows marks the end of the chain,
 *      and a multi-GPU host runs the gradient-row exchange under those GEMMs. */
int pn_set_backward_schedule(int schedule);
int pn_get_backward_schedule(void);

/* conf_coefficient of the aggregator output (point_aggregators.py:809-811 applied to neural_points.py's index_select of
 * points_conf with sample_pidx.clamp(min=0)): out[i] = clamp(table[max(idx[i],0)], lo, hi) for the whole query result in one
 * pass (evaluated as x - (x - clamp(x)) like the reference, bit-identical).  table float[n_points]; idx int32[n] (-1 = empty slot -> row 0, as in the reference); idx/out 16-byte aligned.
 * The clamp is straight-through for gradients (gradiant_clamp): the host scatters d(out) back to the table rows. */
int pn_gather_clamp(const float* table, int64_t n_points, const int32_t* idx, int64_t n, float lo, float hi, float* out,
                    void* stream);

/* ------------------------------------------------------------------------------------------------
 * 4b. Sparse exchange of per-point gradient rows between ranks (new: the reference has no multi-GPU path).
 * With scatter_rows == 1, pn_agg_backward can also return touched_ids int32[M] - the table rows that received
 * gradient, ascending and unique, padded with -1 - and n_touched int32[1].
 *   pn_rows_pack : out[cap, F+8] = [id bits | embedding F | conf | dir 3 | color 3] of rows ids[0..n_ids);
 *                  rows n_ids..cap are padding (id -1, zeros)
 *   pn_rows_merge: gathered [n_ranks, cap, F+8] (all-gather of the packed buffers): every table row named there
 *                  becomes the sum over ranks IN RANK ORDER of the rows carrying its id (bit-identical on every
 *                  replica, no atomics; at two ranks the v0 + v1 of a dense all-reduce).  One pass, rows are written,
 *                  not accumulated.  slots: int32[n_points * n_ranks] owned by the caller, ALL -1 on entry; it is
 *                  all -1 again on return (allocate and fill once, reuse every step).  n_ranks <= 32.
 * ---------------------------------------------------------------------------------------------- */
/* count[0] = number of distinct table rows referenced by sample_pidx int32[n] (-1 = empty): the touched-row count of the
 * coming backward, available as soon as the neighbour query is - ranks agree on the exchange capacity without a host
 * sync after the backward.  mark: int32[n_points] scratch. */
int pn_count_touched_rows(const int32_t* sample_pidx, int64_t n, int64_t n_points, int32_t* mark, int32_t* count,
                          void* stream);
/* Zeroes the table rows named by ids[i * stride], i < n (negative: skipped), of the four [n_points, *] gradient tables: with
 * pn_agg_backward(scatter_rows = 2) the caller keeps persistent gradient tables that are zero outside the rows a step
 * touched, and clears those rows before the next step instead of four whole-table memsets (3.1 GB per step at 20 M points).
 * stride: distance in int32 between ids (1 for touched_ids; F + 8 for the id column of a gathered row buffer). */
int pn_rows_zero(const int32_t* ids, int64_t n, int64_t stride, int32_t F, float* g_emb, float* g_conf, float* g_dir,
                 float* g_color, void* stream);
int pn_rows_pack(const int32_t* ids, int64_t n_ids, int64_t cap, int32_t F, const float* g_emb,
                 const float* g_conf, const float* g_dir, const float* g_color, float* out, void* stream);
int pn_rows_merge(const float* gathered, int32_t n_ranks, int64_t cap, int32_t F, int64_t n_points, int32_t* slots,
                  float* g_emb, float* g_conf, float* g_dir, float* g_color, void* stream);

/* ------------------------------------------------------------------------------------------------
 * 5. Compositing  (replaces ray_march / alpha_ray_march diff_ray_marching.py:509-573 with
 *    radiance_render and alpha_blend, diff_render_func.py:36-37,48-50, and the ray_dist block of
 *    neural_points_volumetric_model.py:293-301)
 *
 * One warp per ray: opacity = 1-exp(-sigma*dist), transmittance = exclusive prefix product of
 * (1-opacity+1e-10) by warp shuffles, blend weight = opacity*T, colour = sum_s w_s*feat_s[1:] (+bg*T_bg).
 * Feature rows whose weight cannot matter (invalid sample, or T below 1e-8: early ray termination) are
 * not read.  All per-sample outputs are still written in full.
 *   ray_features [R,SR,C+1]; ray_dist [R,SR]; ray_valid uint8[R,SR]; bg_color [C] or NULL
 *   slot_row int32[R,SR] or NULL: COMPACT feature layout of the fused path - ray_features (and g_features) are then
 *   [S, C+1] with the row of sample (r,s) at slot_row[r,s] (-1: the sample has no row and no density), so that the
 *   dense, mostly empty [R,SR,C+1] tensor of the reference API is never materialised
 *   ray_color [R,C] (NULL for alpha_ray_march); opacity, acc_T, blend_w [R,SR]; bg_T [R]
 * ---------------------------------------------------------------------------------------------- */
int pn_ray_dist(const float* sample_loc, const uint8_t* sample_valid, int64_t n_rays, int32_t SR,
                float vsize_z, int32_t raydist_mode_unit, float* ray_dist, void* stream);

int pn_composite_forward(const float* ray_features, const float* ray_dist, const uint8_t* ray_valid,
                         const int32_t* slot_row, const float* bg_color, int64_t n_rays, int32_t SR, int32_t C, float* ray_color,
                         float* opacity, float* acc_T, float* blend_w, float* bg_T, void* stream);

/* Gradients w.r.t. ray_features [R,SR,C+1] and bg_color [C] (deterministic two-stage reduction over
 * rays).  Upstream gradients g_opacity/g_acc_T/g_blend_w [R,SR], g_bg_T [R], g_ray_color [R,C] may each be
 * NULL (= zero).  n_rows: rows of the compact ray_features (with slot_row; 0 / ignored for the dense layout) - the
 * per-row part of the backward then runs one warp per feature row instead of walking each ray.  n_rows_dev (device int32,
 * may be NULL): the row count is min(*n_rows_dev, n_rows) and stays on the device; n_rows is then the capacity.
 * workspace: pn_composite_workspace_bytes(n_rays, C, n_rows). */
int64_t pn_composite_workspace_bytes(int64_t n_rays, int32_t C, int64_t n_rows);
int pn_composite_backward(const float* ray_features, const float* ray_dist, const uint8_t* ray_valid,
                          const int32_t* slot_row, const float* bg_color, const float* opacity, const float* acc_T,
                          const float* bg_T, const float* g_ray_color, const float* g_opacity,
                          const float* g_acc_T, const float* g_blend_w, const float* g_bg_T,
                          int64_t n_rays, int32_t SR, int32_t C, int64_t n_rows, const int32_t* n_rows_dev,
                          float* g_features, float* g_bg_color, void* workspace, int64_t workspace_bytes, void* stream);

/* Colour MSE of a batch of rays and its gradient in one pass (replaces torch.nn.MSELoss on coarse_raycolor,
 * base_rendering_model.py:543-550 `compute_losses`, and its autograd chain).
 *   loss[0] = inv_norm * sum_r w_r * sum_c (pred[r,c]-target[r,c])^2      grad[r,c] = 2*inv_norm*w_r*(pred-target)
 *   ray_weight float[n_rays] or NULL (= 1); ray_mask int32[n_rays] or NULL: rays with ray_mask <= 0 get weight 0 - the
 *   reference's `ray_masked_coarse_raycolor` item (:543-550) is ray_mask = the query's per-ray neighbour flag with
 *   inv_norm * *inv_norm_dev = 1 / (n_masked * C); inv_norm_dev (device float, may be NULL) lets that count - and, with ray
 *   sharding, its sum over ranks - stay on the device.
 *   grad float[n_rays,C] or NULL.  Fixed-tree reduction: deterministic.  workspace >= pn_color_mse_workspace_bytes(). */
int64_t pn_color_mse_workspace_bytes(void);
int pn_color_mse(const float* pred, const float* target, const float* ray_weight, const int32_t* ray_mask, int64_t n_rays,
                 int32_t C, float inv_norm, const float* inv_norm_dev, float* loss, float* grad, void* workspace,
                 int64_t workspace_bytes, void* stream);

/* ------------------------------------------------------------------------------------------------
 * 5. Point-cloud maintenance between passes
 *
 * pn_probe_rays - the probe outputs of one batch of rays (models/neural_points_volumetric_model_ori.py:351-372,
 *   consumed by probe_hole, run/train_ft.py:420-533): per ray the sample of largest opacity (first among ties, like
 *   torch.max), its world location, the distance to the nearest of its K neighbour rows (empty neighbour slots read
 *   table row 0, as the reference's clamp(min=0) gather does) and the (weight * clamp(conf, lo, hi))-weighted sums of
 *   the neighbours' colour / dir / conf / embedding.  Uncompacted ray layout: rays with ray_any <= 0 get zeros and
 *   max_slot = -1, which is what `unmask` (:131-137) leaves for them.
 *     opacity [R,SR]; ray_any int32[R]; sample_loc_w [R,SR,3]; sample_pidx int32[R,SR,K]; weight [R,SR,K] (weight_out
 *     of pn_agg_forward); tables xyz [N,3], embedding [N,F], color/dir [N,3], conf [N]
 *     -> max_opacity [R], max_slot int32[R], max_loc_w [R,3], far_dist [R], avg_color/avg_dir [R,3], avg_conf [R],
 *        avg_embedding [R,F]
 *
 * pn_prune_mark + pn_points_compact - NeuralPoints.prune (models/neural_points/neural_points.py:350-373):
 *   dst_index int32[N+1] = exclusive scan of (conf[i] >= thresh), dst_index[N] = rows kept (the caller reads it to size
 *   the new tables); pn_points_compact then moves every kept row of all five tables in ONE pass (a warp per row) where
 *   the reference runs a boolean-mask gather per table.  conf/dir/color (and their outputs) may be NULL.
 *   The output tables may be LARGER than the kept count: NeuralPoints.grow_points (:376-402) appends its rows behind
 *   them in the same allocation instead of a second torch.cat pass over every table.
 * ---------------------------------------------------------------------------------------------- */
int pn_probe_rays(const float* opacity, const int32_t* ray_any, const float* sample_loc_w, const int32_t* sample_pidx,
                  const float* weight, const float* xyz, const float* embedding, const float* color, const float* dir,
                  const float* conf, int64_t n_rays, int32_t SR, int32_t K, int32_t F, float conf_lo, float conf_hi,
                  float* max_opacity, int32_t* max_slot, float* max_loc_w, float* far_dist, float* avg_color,
                  float* avg_dir, float* avg_conf, float* avg_embedding, void* stream);

int64_t pn_prune_workspace_bytes(int64_t n_points);
int pn_prune_mark(const float* conf, int64_t n_points, float thresh, int32_t* dst_index, void* workspace,
                  int64_t workspace_bytes, void* stream);
int pn_points_compact(const int32_t* dst_index, int64_t n_points, int32_t F, const float* xyz, const float* embedding,
                      const float* conf, const float* dir, const float* color, float* out_xyz, float* out_embedding,
                      float* out_conf, float* out_dir, float* out_color, void* stream);

/* ------------------------------------------------------------------------------------------------
 * 6. Voxel down-sampling of a point cloud
 *
 * pn_voxel_downsample - models/mvs/mvs_utils.py:537-561 `construct_vox_points_closest` (called when points are
 *   (re)voxelised: run/train_ft.py:141-146,693, train_waymo.py:531): voxel coordinate floor((p - space_min) / vox_size)
 *   per point, the distinct voxels in lexicographic (x, y, z) order (= torch.unique(dim=0)), each voxel's centroid
 *   (= torch_scatter.scatter_mean, summed in ascending point index) and the index of the point nearest to that centroid
 *   (= scatter_min's argmin; the smallest index among ties).  One stable LSD radix sort of (voxel key, point index)
 *   replaces unique + the two atomic scatters.
 *     xyz [N,3] device; space_min, vox_size: HOST float[3]; coord_base: HOST int32[3], the smallest voxel coordinate per
 *     axis (coordinates may be negative: the reference's box filter `prod(...) > 0` lets points through that are outside
 *     on two axes); bits: key bits per axis (coordinate - coord_base must lie in [0, 2^bits); 3*bits <= 63) -
 *     coordinates outside are clamped and counted in n_outside (device int32[1], may be NULL)
 *     -> centroid [N,3], grid_idx int32[N,3], min_idx int64[N], inv_idx int32[N] (voxel of every point) - only the
 *        first n_voxels[0] (device int32[1]) rows of the first three are written
 * ---------------------------------------------------------------------------------------------- */
int64_t pn_voxel_downsample_workspace_bytes(int64_t n_points);
int pn_voxel_downsample(const float* xyz, int64_t n_points, const float* space_min, const float* vox_size,
                        const int32_t* coord_base, int32_t bits, float* centroid, int32_t* grid_idx, int64_t* min_idx, int32_t* inv_idx, int32_t* n_voxels,
                        int32_t* n_outside, void* workspace, int64_t workspace_bytes, void* stream);

/* ------------------------------------------------------------------------------------------------
 * 7. Optimiser step
 *
 * pn_adam_step - torch.optim.Adam(lr, betas=(0.9, 0.999), eps=1e-8; no weight decay, no amsgrad) as the reference sets
 *   it up for the MLP (`optimizer`) and for the point tables + bg_color (`neural_point_optimizer`),
 *   models/mvs_points_volumetric_model.py:47-100, stepped at :103-125.  ONE launch per 32 tensors; the arithmetic
 *   follows torch's single-tensor Adam op by op in fp32 (lerp, mul+addcmul, sqrt/div/add, addcdiv).
 *   segs: HOST array; every tensor is a contiguous fp32 device buffer of n elements; step = 1 for the first update.
 * ---------------------------------------------------------------------------------------------- */
typedef struct pn_adam_segment {
    float* param;
    const float* grad;
    float* exp_avg;
    float* exp_avg_sq;
    int64_t n;
} pn_adam_segment;

int pn_adam_step(const pn_adam_segment* segs, int32_t n_segs, double lr, double beta1, double beta2, double eps, int64_t step,
                 void* stream);

/* Lazy (touched-row) Adam for the per-point tables: only rows ids[i * stride] >= 0, i < n, are updated, with the arithmetic of
 * torch.optim.SparseAdam (moments of untouched rows stay as they are; bias corrections use the global `step`).  An explicit
 * option with DIFFERENT semantics from the reference's dense torch.optim.Adam (mvs_points_volumetric_model.py:47-100), for
 * clouds where the dense update (28 B x 39 floats x N per step: 21.8 GB at 20 M points) costs more than the step itself.
 * Up to four tables [n_points, width] share one id list. */
typedef struct pn_adam_rows_table {
    float* param;
    const float* grad;
    float* exp_avg;
    float* exp_avg_sq;
    int32_t width;
} pn_adam_rows_table;
int pn_adam_rows(const int32_t* ids, int64_t n, int64_t stride, const pn_adam_rows_table* tabs, int32_t n_tabs, double lr,
                 double beta1, double beta2, double eps, int64_t step, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* POINTNERF_B200_H */
