"""Recipe: compile the reference's OWN querier kernels as serial host C++ -> oracle/_ref/.

TEST INFRASTRUCTURE (see oracle/__init__.py).  Runs only where ``/root/reference`` exists (the
build container); the resulting ``oracle/_ref/libpn_ref_querier.so`` is git-ignored but travels to
the GPU box with the snapshot.  No reference source is copied into the repository: the CUDA C
string is read from where it lies
(models/neural_points/query_point_indices_worldcoords.py:114-530, the first argument of
``SourceModule``), written to a temporary directory, compiled and discarded.

How the string becomes host code:
  * ``oracle/ref_shim/{cuda.h,cuda_runtime.h,curand_kernel.h}`` shadow the CUDA headers it includes
    (``__global__``, ``blockIdx``…, serial atomics, nvcc's fma contraction of d^2);
  * ``KN`` (string-spliced from ``opt.K`` at :116) only sizes a local buffer; the kernels use the
    runtime ``K`` argument, so one build with KN=128 serves every K <= 128;
  * a launcher per kernel (appended below, ours) loops blocks and threads in ascending order.
"""
import ast
import os
import subprocess
import sys
import tempfile

HERE = os.path.dirname(os.path.abspath(__file__))
REF_ROOT = os.environ.get("POINTNERF_REFERENCE", "/root/reference")
REF_PY = os.path.join(REF_ROOT, "models", "neural_points", "query_point_indices_worldcoords.py")
OUT_DIR = os.path.join(HERE, "_ref")
OUT_SO = os.path.join(OUT_DIR, "libpn_ref_querier.so")
# the perspective-frustum querier: same recipe, its own library because both kernel strings define
# ``query_neigh_along_ray_layered`` and ``namespace cuda``
REF_PY_FRUSTUM = os.path.join(REF_ROOT, "models", "neural_points", "query_point_indices.py")
OUT_SO_FRUSTUM = os.path.join(OUT_DIR, "libpn_ref_querier_frustum.so")
KN = 128

LAUNCHERS = r'''
#undef float
#define PN_LOOP(grid, block) \
    blockDim.x = (unsigned)(block); gridDim.x = (unsigned)(grid); \
    for (unsigned _b = 0; _b < (unsigned)(grid); ++_b) for (unsigned _t = 0; _t < (unsigned)(block); ++_t) \
        if ((blockIdx.x = _b, threadIdx.x = _t, true))
typedef pn_gfloat gf;
extern "C" {
int pn_ref_kn() { return KN; }
void pn_ref_claim_occ(int grid, int block, const float* in_data, const int* in_actual_numpoints, int B, int N,
                      const float* shift, const float* vsz, const int* gsz, int vol, int max_o,
                      int* occ_idx, int* coor_2_occ, int* occ_2_coor) {
    PN_LOOP(grid, block) claim_occ((const gf*)in_data, in_actual_numpoints, B, N, (const gf*)shift, (const gf*)vsz,
                                   gsz, vol, max_o, occ_idx, coor_2_occ, occ_2_coor, 0ul);
}
void pn_ref_map_coor2occ(int grid, int block, int B, const int* gsz, const int* ksz, int vol, int max_o,
                         int* occ_idx, int* coor_occ, int* coor_2_occ, int* occ_2_coor) {
    PN_LOOP(grid, block) map_coor2occ(B, gsz, ksz, vol, max_o, occ_idx, coor_occ, coor_2_occ, occ_2_coor);
}
void pn_ref_fill_occ2pnts(int grid, int block, const float* in_data, const int* in_actual_numpoints, int B, int N,
                          int P, const float* shift, const float* vsz, const int* gsz, int vol, int max_o,
                          int* coor_2_occ, int* occ_2_pnts, int* occ_numpnts) {
    PN_LOOP(grid, block) fill_occ2pnts((const gf*)in_data, in_actual_numpoints, B, N, P, (const gf*)shift,
                                       (const gf*)vsz, gsz, vol, max_o, coor_2_occ, occ_2_pnts, occ_numpnts, 0ul);
}
void pn_ref_mask_raypos(int grid, int block, float* raypos, int* coor_occ, int B, int R, int D, int vol,
                        const float* shift, const int* gsz, const float* vsz, int* raypos_mask) {
    PN_LOOP(grid, block) mask_raypos((gf*)raypos, coor_occ, B, R, D, vol, (const gf*)shift, gsz, (const gf*)vsz,
                                     raypos_mask);
}
void pn_ref_get_shadingloc(int grid, int block, const float* raypos, const int* raypos_mask, int B, int R, int D,
                           int SR, float* sample_loc, int* sample_loc_mask) {
    PN_LOOP(grid, block) get_shadingloc((const gf*)raypos, raypos_mask, B, R, D, SR, (gf*)sample_loc, sample_loc_mask);
}
void pn_ref_query_along_ray(int grid, int block, const float* in_data, int B, int SR, int R, int max_o, int P, int K,
                            int vol, float radius_limit2, const float* shift, const int* gsz, const float* vsz,
                            const int* ksz, const int* occ_numpnts, const int* occ_2_pnts, const int* coor_2_occ,
                            const float* sample_loc, const int* sample_loc_mask, int* sample_pidx, int NN) {
    PN_LOOP(grid, block) query_neigh_along_ray_layered((const gf*)in_data, B, SR, R, max_o, P, K, vol, gf(radius_limit2),
                                                       (const gf*)shift, gsz, (const gf*)vsz, ksz, occ_numpnts, occ_2_pnts,
                                                       coor_2_occ, (const gf*)sample_loc, sample_loc_mask, sample_pidx, 0ul, NN);
}
}
'''


LAUNCHERS_FRUSTUM = r'''
#undef float
#define PN_LOOP(grid, block) \
    blockDim.x = (unsigned)(block); gridDim.x = (unsigned)(grid); \
    for (unsigned _b = 0; _b < (unsigned)(grid); ++_b) for (unsigned _t = 0; _t < (unsigned)(block); ++_t) \
        if ((blockIdx.x = _b, threadIdx.x = _t, true))
typedef pn_gfloat gf;
extern "C" {
int pn_ref_kn() { return KN; }
// query_point_indices.py:265-314
void pn_ref_get_occ_vox(int grid, int block, const float* in_data, const int* n_act, int B, int N, const float* shift,
                        const float* vsz, const int* gsz, const int* ksz, int pixel_size, int vol, uint8_t* coor_occ,
                        int8_t* loc_coor_counter, int* near_id, int* far_id, int inverse) {
    PN_LOOP(grid, block) get_occ_vox((const gf*)in_data, n_act, B, N, (const gf*)shift, (const gf*)vsz, gsz, ksz, pixel_size,
                                     vol, coor_occ, loc_coor_counter, near_id, far_id, inverse);
}
// :316-365
void pn_ref_near_vox_full(int grid, int block, int B, int SR, const int* pixel_idx, int R, const int* vscale, const int* gsz,
                          int pixel_size, int vol, const int* ksz, uint8_t* pixel_map, int8_t* ray_mask,
                          const uint8_t* coor_occ, int8_t* loc_coor_counter, const int* near_id, const int* far_id,
                          short* voxel_to_coorz_idx) {
    PN_LOOP(grid, block) near_vox_full(B, SR, pixel_idx, R, vscale, gsz, pixel_size, vol, ksz, pixel_map, ray_mask, coor_occ,
                                       loc_coor_counter, near_id, far_id, voxel_to_coorz_idx);
}
// :368-413
void pn_ref_insert_vox_points(int grid, int block, float* in_data, int* n_act, int B, int N, int P, int max_o,
                              int pixel_size, int vol, const float* shift, const int* gsz, const float* vsz,
                              const int8_t* loc_coor_counter, short* voxel_pnt_counter, int* voxel_to_pntidx, int inverse) {
    PN_LOOP(grid, block) insert_vox_points((gf*)in_data, n_act, B, N, P, max_o, pixel_size, vol, (const gf*)shift, gsz,
                                           (const gf*)vsz, loc_coor_counter, voxel_pnt_counter, voxel_to_pntidx, 0ul, inverse);
}
// :495-600
void pn_ref_query_along_ray(int grid, int block, const float* in_data, int B, int SR, int R, int max_o, int P, int K,
                            int pixel_size, int vol, float radius_limit2, float depth_limit2, const float* shift,
                            const int* gsz, const float* vsz, const float* ray_vsz, const int* vscale, const int* ksz,
                            const int* pixel_idx, const int8_t* loc_coor_counter, const short* voxel_to_coorz_idx,
                            const short* voxel_pnt_counter, const int* voxel_to_pntidx, int* sample_pidx, float* sample_loc,
                            int NN, int inverse) {
    PN_LOOP(grid, block) query_neigh_along_ray_layered((const gf*)in_data, B, SR, R, max_o, P, K, pixel_size, vol,
                                                       gf(radius_limit2), gf(depth_limit2), (const gf*)shift, gsz,
                                                       (const gf*)vsz, (const gf*)ray_vsz, vscale, ksz, pixel_idx,
                                                       loc_coor_counter, voxel_to_coorz_idx, voxel_pnt_counter,
                                                       voxel_to_pntidx, sample_pidx, (gf*)sample_loc, 0ul, NN, inverse);
}
}
'''


def extract_kernel_source(k_define: int, ref_py: str = None) -> str:
    """Evaluates the string expression passed to ``SourceModule(...)`` with ``str(self.opt.K)`` -> k_define."""
    ref_py = ref_py or REF_PY
    tree = ast.parse(open(ref_py).read())
    for node in ast.walk(tree):
        if isinstance(node, ast.Call) and getattr(node.func, "id", None) == "SourceModule":
            expr = node.args[0]

            def ev(e):
                if isinstance(e, ast.Constant):
                    return e.value
                if isinstance(e, ast.BinOp) and isinstance(e.op, ast.Add):
                    return ev(e.left) + ev(e.right)
                if isinstance(e, ast.Call) and getattr(e.func, "id", None) == "str":
                    return str(k_define)
                raise ValueError("unexpected node in SourceModule argument: %s" % ast.dump(e)[:80])
            return ev(expr)
    raise RuntimeError("SourceModule(...) call not found in %s" % ref_py)


def available() -> bool:
    return os.path.isfile(REF_PY)


def _build_one(ref_py, out_so, launchers, force):
    if os.path.isfile(out_so) and not force:
        if not os.path.isfile(ref_py) or os.path.getmtime(out_so) >= max(
                os.path.getmtime(ref_py), os.path.getmtime(__file__),
                *(os.path.getmtime(os.path.join(HERE, "ref_shim", f)) for f in os.listdir(os.path.join(HERE, "ref_shim")))):
            return out_so
    if not os.path.isfile(ref_py):
        raise RuntimeError("reference tree not present; oracle/_ref cannot be (re)built here")
    os.makedirs(OUT_DIR, exist_ok=True)
    src = extract_kernel_source(KN, ref_py) + launchers
    with tempfile.TemporaryDirectory() as tmp:
        cpp = os.path.join(tmp, "ref_querier_host.cpp")
        with open(cpp, "w") as f:
            f.write(src)
        cmd = ["g++", "-O2", "-std=c++17", "-shared", "-fPIC", "-mfma", "-ffp-contract=off", "-w",
               "-I", os.path.join(HERE, "ref_shim"), cpp, "-o", out_so]
        subprocess.check_call(cmd)
    return out_so


def build_gpu(K: int, force: bool = False) -> str:
    """The SAME kernel string, unmodified (``KN`` = K as the reference splices it, :116), compiled for the GPU:
    ``nvcc -arch=sm_100a -cubin`` -> oracle/_ref/ref_querier_sm100a_K<K>.cubin, loaded by oracle/ref_gpu.py through the
    driver API (what pycuda's SourceModule does).  The reference's GPU-side baseline on the B200."""
    out = os.path.join(OUT_DIR, "ref_querier_sm100a_K%d.cubin" % K)
    if os.path.isfile(out) and not force:
        return out
    os.makedirs(OUT_DIR, exist_ok=True)
    with tempfile.TemporaryDirectory() as tmp:
        cu = os.path.join(tmp, "ref_querier.cu")
        with open(cu, "w") as f:
            f.write(extract_kernel_source(K))
        r = subprocess.run(["nvcc", "-arch=sm_100a", "-O3", "-cubin", "-w", "-o", out, cu], capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError("nvcc failed on the reference kernel string:\n" + r.stderr[-2000:])
    return out


def build(force: bool = False) -> str:
    """Builds both host libraries and the GPU cubins; returns the world-coordinate host library (the hot path's oracle)."""
    _build_one(REF_PY_FRUSTUM, OUT_SO_FRUSTUM, LAUNCHERS_FRUSTUM, force)
    for K in (8, 16):
        build_gpu(K, force)
    return _build_one(REF_PY, OUT_SO, LAUNCHERS, force)


if __name__ == "__main__":
    print(build(force="--force" in sys.argv))
