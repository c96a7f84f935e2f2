"""numpy model of the data structure the CUDA frustum querier uses (pointnerf_b200.h section 3b).

TEST INFRASTRUCTURE (see oracle/__init__.py).  The reference keeps, per view, dense uint8 / int8 tables, a per-column
list of the voxels "needed" by this batch's rays and [w, h, max_o, P] buckets filled in arrival order
(query_point_indices.py:617-786); the product keeps one CSR over all claimed frustum cells.  This file restates the
product's formulation in plain numpy / Python so that the CPU suite can show the two formulations give the same
neighbours (tests/test_oracle.py, against oracle/frustum_ref.py = the reference's own kernels), independent of the
CUDA transcription that the GPU suite checks.  Small cases only (Python loops).
"""
import numpy as np

f32 = np.float32


def build(xyz_pers, hp):
    """-> cells dict {linear cell: ascending point indices}, occ bool [dx,dy,dz] (dilated), dims."""
    xyz = np.asarray(xyz_pers, f32)
    shift, vs = hp["ranges"][:3].astype(f32), hp["scaled_vsize"].astype(f32)
    dims = [int(v) for v in hp["scaled_vdim"]]
    t = (xyz - shift[None, :]) / vs[None, :]                       # float32 sub, div
    fl = np.floor(t).astype(np.int64)
    tr = np.trunc(t).astype(np.int64)
    d = np.asarray(dims)[None, :]
    fl_ok = np.all((fl >= 0) & (fl < d), axis=1)
    tr_ok = np.all((tr >= 0) & (tr < d), axis=1)
    lin = lambda c: (c[:, 0] * dims[1] + c[:, 1]) * dims[2] + c[:, 2]
    src = set(lin(fl[fl_ok]).tolist())                             # cells claimed by a floor()-valid point
    cells = {}
    for i in np.nonzero(tr_ok)[0]:
        c = int(lin(tr[i:i + 1])[0])
        if c in src:
            cells.setdefault(c, []).append(int(i))
    occ = np.zeros(dims, bool)
    q = [int(v) for v in hp["query_size"]]
    for c in src:
        z, y, x = c % dims[2], (c // dims[2]) % dims[1], c // (dims[1] * dims[2])
        occ[max(0, x - q[0] // 2):min(dims[0], x + (q[0] + 1) // 2), max(0, y - q[1] // 2):min(dims[1], y + (q[1] + 1) // 2),
            max(0, z - q[2] // 2):min(dims[2], z + (q[2] + 1) // 2)] = True
    return cells, occ, dims


def sample(pix, occ, dims, vscale, SR):
    """-> sample_z int32 [R,SR] (-1 unused), ray_mask [R]."""
    R = len(pix)
    sz = np.full((R, SR), -1, np.int32)
    mask = np.zeros(R, np.int8)
    for r in range(R):
        col = occ[pix[r, 0] // vscale[0], pix[r, 1] // vscale[1]]
        zs = np.nonzero(col)[0]
        mask[r] = 1 if (len(zs) and zs[-1] > 0) else 0
        if mask[r]:                                                 # rays the reference drops keep no samples
            sz[r, :min(SR, len(zs))] = zs[:SR]
    return sz, mask


def knn(xyz_pers, cells, dims, hp, pix, sample_z, opt):
    """-> sample_pidx int32 [R,SR,K], sample_loc float32 [R,SR,3]; float arithmetic as the kernel's PTX."""
    xyz = np.asarray(xyz_pers, f32)
    R, SR = sample_z.shape
    K, P, NN = int(opt.K), int(opt.P), int(opt.NN)
    shift, vs = hp["ranges"][:3].astype(f32), hp["scaled_vsize"].astype(f32)
    vscale = hp["vscale"]
    rv = (hp["scaled_vsize"] / hp["vscale"]).astype(f32)
    r2, dl2 = f32(hp["radius_limit"] ** 2), f32(hp["depth_limit"] ** 2)
    ks = hp["kernel_size"]
    n_layers, z_layers = (int(ks[0]) + 1) // 2, (int(ks[2]) + 1) // 2
    pidx = np.full((R, SR, K), -1, np.int32)
    loc = np.zeros((R, SR, 3), f32)

    def fma32(a, b, c):
        return f32(np.float64(a) * np.float64(b) + np.float64(c))   # exact product fits in float64; one rounding to float32

    for r in range(R):
        px, py = int(pix[r, 0]), int(pix[r, 1])
        fx, fy = px // vscale[0], py // vscale[1]
        cx = f32((px % vscale[0] + 0.5) * np.float64(rv[0]) + np.float64(fma32(vs[0], f32(fx), shift[0])))
        cy = f32((py % vscale[1] + 0.5) * np.float64(rv[1]) + np.float64(fma32(vs[1], f32(fy), shift[1])))
        for s in range(SR):
            fz = int(sample_z[r, s])
            cz = f32((fz + 0.5) * np.float64(vs[2]) + np.float64(shift[2]))
            loc[r, s] = (cx, cy, cz)
            if fz < 0:
                continue
            cxz, cyz = f32(cz * cx), f32(cz * cy)
            ids, d2s = [], []
            far2, far_ind, kid = f32(0), 0, 0
            for layer in range(n_layers):
                zl = min(z_layers - 1, layer)
                for x in range(max(-fx, -layer), min(dims[0] - fx, layer + 1)):
                    for y in range(max(-fy, -layer), min(dims[1] - fy, layer + 1)):
                        for z in range(max(-fz, -zl), min(dims[2] - fz, zl + 1)):
                            if max(abs(x), abs(y)) != layer and ((abs(z) != zl) if zl == layer else True):
                                continue
                            c = ((fx + x) * dims[1] + fy + y) * dims[2] + fz + z
                            for p in cells.get(c, ())[:P]:
                                pp = xyz[p]
                                if NN < 2:
                                    xv, yv = f32(pp[0] - cx), f32(pp[1] - cy)
                                else:
                                    xv, yv = f32(f32(pp[0] * pp[2]) - cxz), f32(f32(pp[1] * pp[2]) - cyz)
                                xy2 = fma32(xv, xv, f32(yv * yv))
                                dz = f32(pp[2] - cz)
                                z2 = f32(dz * dz)
                                d2 = f32(xy2 + z2)
                                if (r2 == 0 or xy2 <= r2) and (dl2 == 0 or z2 <= dl2):
                                    if kid < K:
                                        ids.append(p)
                                        d2s.append(d2)
                                        if d2 > far2:
                                            far2, far_ind = d2, kid
                                    elif d2 < far2:
                                        ids[far_ind], d2s[far_ind] = p, d2
                                        far2 = d2
                                        for i in range(K):
                                            if d2s[i] > far2:
                                                far2, far_ind = d2s[i], i
                                    kid += 1
            pidx[r, s, :len(ids)] = ids
    return pidx, loc
