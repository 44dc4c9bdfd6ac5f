// voxel.cuh — voxel_down_sample_and_trace on the device.
// Replaces pcd.voxel_down_sample_and_trace(voxel_size, min_bound, max_bound, False)
// (reference PointCloud.py:338-341; Open3D PointCloud::VoxelDownSampleAndTrace):
//   key    = (ix,iy,iz) = floor((p - min_bound) / voxel_size)      FP64 subtract + divide
//   sort   = stable LSD radix sort of the packed key, payload = original index
//   rows   = segment heads of the sorted keys (cell-range table), ascending (ix,iy,iz)
//   reduce = one thread per row walks its points in ascending original index, exactly the
//            order in which Open3D's serial loop calls AddPoint, so sums round identically
#pragma once
#include "cloud.cuh"

struct VoxPack { int bx, by, bz; };  // bits per axis in the packed key
struct VoxOrigin {                   // grid origin (min_bound) and the smallest voxel index per axis: keys are index - lo >= 0
    double o[3];
    int64_t lo[3];
};

// one point: its 32-byte record (position + octant) and its key
__device__ __forceinline__ void voxel_key_one(double px, double py, double pz, int64_t i, const VoxOrigin &og, double voxel_size,
                                              const VoxPack &pk, int wide, int pack, uint64_t *__restrict__ key_lo,
                                              uint32_t *__restrict__ key_ix, double4 *__restrict__ p4)
{
    double r0 = __ddiv_rn(__dsub_rn(px, og.o[0]), voxel_size);
    double r1 = __ddiv_rn(__dsub_rn(py, og.o[1]), voxel_size);
    double r2 = __ddiv_rn(__dsub_rn(pz, og.o[2]), voxel_size);
    double f0 = floor(r0), f1 = floor(r1), f2 = floor(r2);
    int o = 0;
    if (__dsub_rn(r0, f0) >= 0.5) o |= 1;
    if (__dsub_rn(r1, f1) >= 0.5) o |= 2;
    if (__dsub_rn(r2, f2) >= 0.5) o |= 4;
    p4[i] = make_double4(px, py, pz, (double)o);      // position + octant in one 32-byte record: the reduction gathers one sector per point
    // voxel index relative to the smallest index of the cloud: non-negative, row order unchanged
    uint64_t ix = (uint64_t)((int64_t)f0 - og.lo[0]), iy = (uint64_t)((int64_t)f1 - og.lo[1]), iz = (uint64_t)((int64_t)f2 - og.lo[2]);
    if (wide) {
        key_lo[i] = (iy << pk.bz) | iz;
        key_ix[i] = (uint32_t)ix;
    } else {
        const uint64_t k = (ix << (pk.by + pk.bz)) | (iy << pk.bz) | iz;
        key_lo[i] = pack ? (k << pack) | (uint64_t)i : k;          // packed: the original index rides below the key (radix_sort_packed)
    }
}

// HBM bound (24 B read, 40 B written per point): two points per thread and step, so that six loads are in flight per thread
// (one point per step left the kernel latency bound at 62 % of the copy rate, ncu r02y)
__global__ void __launch_bounds__(256)
voxel_key_kernel(const double *__restrict__ x, const double *__restrict__ y, const double *__restrict__ z,
                 int64_t n, const VoxOrigin og, double voxel_size, VoxPack pk, int wide, int pack,
                 uint64_t *__restrict__ key_lo, uint32_t *__restrict__ key_ix, double4 *__restrict__ p4)
{
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += 2 * stride) {
        const int64_t i2 = i + stride;
        const bool two = i2 < n;
        const double ax = x[i], ay = y[i], az = z[i];
        const double bx = two ? x[i2] : 0.0, by = two ? y[i2] : 0.0, bz = two ? z[i2] : 0.0;
        voxel_key_one(ax, ay, az, i, og, voxel_size, pk, wide, pack, key_lo, key_ix, p4);
        if (two) voxel_key_one(bx, by, bz, i2, og, voxel_size, pk, wide, pack, key_lo, key_ix, p4);
    }
}

__global__ void gather_u32_as_key_kernel(const uint32_t *__restrict__ src, const uint32_t *__restrict__ perm,
                                         int64_t n, uint64_t *__restrict__ dst)
{
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
        dst[i] = src[perm[i]];
}

struct HeadPred {  // sorted position j starts a row
    const uint64_t *keys;
    int shift;          // packed keys: the index bits below the key are not part of it
    __device__ bool operator()(int64_t j) const { return j == 0 || (keys[j] >> shift) != (keys[j - 1] >> shift); }
};
struct HeadPredWide {
    const uint64_t *key_lo;   // per ORIGINAL point
    const uint32_t *key_ix;   // per ORIGINAL point
    const uint32_t *order;
    __device__ bool operator()(int64_t j) const
    {
        if (j == 0) return true;
        uint32_t a = order[j], b = order[j - 1];
        return key_lo[a] != key_lo[b] || key_ix[a] != key_ix[b];
    }
};

__global__ void __launch_bounds__(128)
voxel_reduce_kernel(int64_t V, int64_t n, const int32_t *__restrict__ heads, const uint32_t *__restrict__ order,
                    const double4 *__restrict__ p4,
                    const double *__restrict__ nx, const double *__restrict__ ny, const double *__restrict__ nz,
                    const double *__restrict__ cr, const double *__restrict__ cg, const double *__restrict__ cb,
                    int32_t *__restrict__ rep, int32_t *__restrict__ count, double *__restrict__ centroid,
                    double *__restrict__ onormal, double *__restrict__ ocolor, int32_t *__restrict__ cubic,
                    uint8_t *__restrict__ rep_flags)
{
    for (int64_t v = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; v < V; v += (int64_t)gridDim.x * blockDim.x) {
        int64_t j0 = heads[v], j1 = (v + 1 < V) ? (int64_t)heads[v + 1] : n;
        double sp0 = 0, sp1 = 0, sp2 = 0, sn0 = 0, sn1 = 0, sn2 = 0, sc0 = 0, sc1 = 0, sc2 = 0;
        int32_t cub[8];
#pragma unroll
        for (int k = 0; k < 8; ++k) cub[k] = -1;
        int32_t last = -1;
        for (int64_t j = j0; j < j1; ++j) {
            uint32_t i = order[j];
            const double4 q = p4[i];                 // two 16-byte loads: the single 256-bit load (ld_rec) costs this divergent loop 0.2 ms at 10 M
            sp0 = __dadd_rn(sp0, q.x); sp1 = __dadd_rn(sp1, q.y); sp2 = __dadd_rn(sp2, q.z);
            if (nx) { sn0 = __dadd_rn(sn0, nx[i]); sn1 = __dadd_rn(sn1, ny[i]); sn2 = __dadd_rn(sn2, nz[i]); }
            if (cr) { sc0 = __dadd_rn(sc0, cr[i]); sc1 = __dadd_rn(sc1, cg[i]); sc2 = __dadd_rn(sc2, cb[i]); }
            int o = (int)q.w;
#pragma unroll
            for (int k = 0; k < 8; ++k) if (k == o) cub[k] = (int32_t)i;
            last = (int32_t)i;
        }
        double cnt = (double)(j1 - j0);
        rep[v] = last;
        rep_flags[last] = 1;
        count[v] = (int32_t)(j1 - j0);
        centroid[3 * v] = __ddiv_rn(sp0, cnt);
        centroid[3 * v + 1] = __ddiv_rn(sp1, cnt);
        centroid[3 * v + 2] = __ddiv_rn(sp2, cnt);
        if (nx) {
            double nn = sqrt(__dadd_rn(__dadd_rn(__dmul_rn(sn0, sn0), __dmul_rn(sn1, sn1)), __dmul_rn(sn2, sn2)));
            bool ok = nn > 0.0;
            onormal[3 * v] = ok ? __ddiv_rn(sn0, nn) : sn0;
            onormal[3 * v + 1] = ok ? __ddiv_rn(sn1, nn) : sn1;
            onormal[3 * v + 2] = ok ? __ddiv_rn(sn2, nn) : sn2;
        }
        if (cr) {
            ocolor[3 * v] = __ddiv_rn(sc0, cnt);
            ocolor[3 * v + 1] = __ddiv_rn(sc1, cnt);
            ocolor[3 * v + 2] = __ddiv_rn(sc2, cnt);
        }
#pragma unroll
        for (int k = 0; k < 8; ++k) cubic[8 * v + k] = cub[k];
    }
}

static int bit_length_u32(uint32_t v)
{
    int b = 0;
    while (v) { ++b; v >>= 1; }
    return b < 1 ? 1 : b;
}

// min3 / max3 (nullable): the explicit min_bound / max_bound of the call being replaced (PointCloud.py:339-340); Open3D's
// trace variant takes min_bound itself as the grid origin and uses max_bound only in the size check.  nullptr = the
// cloud's own bounds (what the reference passes).  A slab-sharded scan passes the GLOBAL bounds on every rank.
static int voxel_impl(o3dpp_ctx *c, double voxel_size, const double *min3, const double *max3, int64_t *Vout)
{
    if (Vout) *Vout = 0;
    if (!(voxel_size > 0.0)) return set_err(O3DPP_E_VOXEL_SIZE, "voxel_size <= 0.");
    int64_t n = c->n;
    c->vox_valid = false;
    double mn[3], mx[3];
    cudaEventRecord(c->ev[0], c->stream);
    O3_TRY(bounds_impl(c, mn, mx));
    VoxOrigin og;
    double far3[3];
    for (int a = 0; a < 3; ++a) {
        og.o[a] = min3 ? min3[a] : mn[a];
        far3[a] = max3 ? max3[a] : mx[a];
        if (!(og.o[a] == og.o[a]) || !(far3[a] == far3[a])) return set_err(O3DPP_E_ARG, "min_bound / max_bound must not be NaN");
    }
    double ext = far3[0] - og.o[0];
    if (far3[1] - og.o[1] > ext) ext = far3[1] - og.o[1];
    if (far3[2] - og.o[2] > ext) ext = far3[2] - og.o[2];
    if (voxel_size * 2147483647.0 < ext) return set_err(O3DPP_E_VOXEL_SMALL, "voxel_size is too small.");
    if (n == 0) { c->vox_valid = true; c->vox_V = 0; c->vox_n = 0; return 0; }
    // the smallest / largest index per axis come from the cloud's own bounds through the very same expression
    // (subtraction and division by a positive number are monotone under rounding)
    VoxPack pk;
    uint32_t mi[3];
    for (int a = 0; a < 3; ++a) {
        const double flo = floor((mn[a] - og.o[a]) / voxel_size), fhi = floor((mx[a] - og.o[a]) / voxel_size);
        if (!(flo >= -2147483648.0) || !(fhi <= 2147483647.0))      // points outside the given bounds: Open3D's int cast would overflow
            return set_err(O3DPP_E_VOXEL_SMALL, "voxel_size is too small.");
        og.lo[a] = (int64_t)flo;
        mi[a] = (uint32_t)((int64_t)fhi - og.lo[a]);
    }
    pk.bx = bit_length_u32(mi[0]); pk.by = bit_length_u32(mi[1]); pk.bz = bit_length_u32(mi[2]);
    int wide = (pk.bx + pk.by + pk.bz) > 64;

    size_t nn = (size_t)n + 64;
    O3_TRY(c->s_keys0.ensure(nn * 8)); O3_TRY(c->s_keys1.ensure(nn * 8));
    O3_TRY(c->s_vals0.ensure(nn * 4)); O3_TRY(c->s_vals1.ensure(nn * 4));
    O3_TRY(c->g_p4.ensure(nn * 32)); O3_TRY(c->v_order.ensure(nn * 4)); O3_TRY(c->v_seg.ensure(nn * 4));
    DevBuf wide_lo, wide_ix;
    uint64_t *key_lo = c->s_keys0.as<uint64_t>();
    uint32_t *key_ix = nullptr;
    if (wide) {
        O3_TRY(wide_lo.ensure(nn * 8)); O3_TRY(wide_ix.ensure(nn * 4));
        key_lo = wide_lo.as<uint64_t>();
        key_ix = wide_ix.as<uint32_t>();
    }
    int grid = grid_for(c, n, 256, 16);
    const int pack = (!wide && n <= ((int64_t)1 << RS_PACK_BITS) && pk.bx + pk.by + pk.bz + RS_PACK_BITS <= 64) ? RS_PACK_BITS : 0;
    LAUNCH(c, voxel_key_kernel, grid, 256, 0, c->pts.x(), c->pts.y(), c->pts.z(), n, og,
           voxel_size, pk, wide, pack, key_lo, key_ix, c->g_p4.as<double4>());
    O3_TRY(check_launch("voxel_key"));
    uint64_t *ks = nullptr;
    uint32_t *vs = nullptr;
    int64_t V = 0;
    int32_t *heads = c->v_seg.as<int32_t>();
    if (pack) {
        O3_TRY(radix_sort_packed(c, c->s_keys0.as<uint64_t>(), c->s_keys1.as<uint64_t>(), n, pack, pk.bx + pk.by + pk.bz, &ks));
        HeadPred hp{ks, pack};
        O3_TRY(compact_indices(c, hp, n, heads, &V));
        LAUNCH(c, unpack_index_kernel, grid, 256, 0, ks, n, c->v_order.as<uint32_t>());
    } else if (!wide) {
        O3_TRY(radix_sort_pairs(c, c->s_keys0.as<uint64_t>(), c->s_keys1.as<uint64_t>(), c->s_vals0.as<uint32_t>(),
                                c->s_vals1.as<uint32_t>(), n, pk.bx + pk.by + pk.bz, &ks, &vs));
        HeadPred hp{ks, 0};
        O3_TRY(compact_indices(c, hp, n, heads, &V));
    } else {
        // stage 1: (iy,iz) ; stage 2: ix, stable, payload carried over
        CU_TRY(cudaMemcpyAsync(c->s_keys0.p, key_lo, (size_t)n * 8, cudaMemcpyDeviceToDevice, c->stream));
        O3_TRY(radix_sort_pairs(c, c->s_keys0.as<uint64_t>(), c->s_keys1.as<uint64_t>(), c->s_vals0.as<uint32_t>(),
                                c->s_vals1.as<uint32_t>(), n, pk.by + pk.bz, &ks, &vs));
        if (vs != c->s_vals0.as<uint32_t>())
            CU_TRY(cudaMemcpyAsync(c->s_vals0.p, vs, (size_t)n * 4, cudaMemcpyDeviceToDevice, c->stream));
        LAUNCH(c, gather_u32_as_key_kernel, grid, 256, 0, key_ix, c->s_vals0.as<uint32_t>(), n, c->s_keys0.as<uint64_t>());
        O3_TRY(radix_sort_pairs(c, c->s_keys0.as<uint64_t>(), c->s_keys1.as<uint64_t>(), c->s_vals0.as<uint32_t>(),
                                c->s_vals1.as<uint32_t>(), n, pk.bx, &ks, &vs, true));
        HeadPredWide hp{key_lo, key_ix, vs};
        O3_TRY(compact_indices(c, hp, n, heads, &V));
    }
    if (!pack) CU_TRY(cudaMemcpyAsync(c->v_order.p, vs, (size_t)n * 4, cudaMemcpyDeviceToDevice, c->stream));
    size_t vv = (size_t)V + 64;
    O3_TRY(c->v_rep.ensure(vv * 4)); O3_TRY(c->v_rep_asc.ensure(vv * 4)); O3_TRY(c->v_count.ensure(vv * 4));
    O3_TRY(c->v_centroid.ensure(vv * 24)); O3_TRY(c->v_normal.ensure(vv * 24)); O3_TRY(c->v_color.ensure(vv * 24));
    O3_TRY(c->v_cubic.ensure(vv * 32));
    DevBuf &flags = c->v_flags;
    O3_TRY(flags.ensure(nn));
    CU_TRY(cudaMemsetAsync(flags.p, 0, nn, c->stream));
    bool hn = c->nrm.present, hc = c->col.present;
    LAUNCH(c, voxel_reduce_kernel, grid_for(c, V, 128, 32), 128, 0, V, n, heads, c->v_order.as<uint32_t>(),
           c->g_p4.as<double4>(),
           hn ? c->nrm.x() : nullptr, hn ? c->nrm.y() : nullptr, hn ? c->nrm.z() : nullptr,
           hc ? c->col.x() : nullptr, hc ? c->col.y() : nullptr, hc ? c->col.z() : nullptr,
           c->v_rep.as<int32_t>(), c->v_count.as<int32_t>(), c->v_centroid.as<double>(),
           c->v_normal.as<double>(), c->v_color.as<double>(), c->v_cubic.as<int32_t>(), flags.as<uint8_t>());
    int rc = check_launch("voxel_reduce");
    int64_t V2 = 0;
    if (!rc) {
        FlagPred fp{flags.as<uint8_t>(), 1};
        rc = compact_indices(c, fp, n, c->v_rep_asc.as<int32_t>(), &V2);
    }
    cudaEventRecord(c->ev[1], c->stream);
    cudaStreamSynchronize(c->stream);
    wide_lo.release(); wide_ix.release();
    if (rc) return rc;
    if (V2 != V) return set_err(O3DPP_E_CUDA, "internal: representative count differs from row count");
    c->vox_valid = true;
    c->vox_V = V;
    c->vox_n = n;
    if (Vout) *Vout = V;
    float ms = 0;
    cudaEventElapsedTime(&ms, c->ev[0], c->ev[1]);
    c->timings[0] = ms;
    return 0;
}

static int d2h(o3dpp_ctx *c, void *dst, const void *src, size_t bytes)
{
    if (!dst || bytes == 0) return 0;
    CU_TRY(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToHost, c->stream));
    return 0;
}

static int voxel_fetch_impl(o3dpp_ctx *c, int32_t *rep_asc, int32_t *rep, int32_t *count, double *centroid,
                            double *normal, double *color, int32_t *cubic, int32_t *order, int64_t *seg)
{
    if (!c->vox_valid) return set_err(O3DPP_E_STATE, "o3dpp_voxel has not been run on the current cloud");
    size_t V = (size_t)c->vox_V, n = (size_t)c->vox_n;
    if (normal && !c->nrm.present) return set_err(O3DPP_E_STATE, "cloud has no normals");
    if (color && !c->col.present) return set_err(O3DPP_E_STATE, "cloud has no colours");
    O3_TRY(d2h(c, rep_asc, c->v_rep_asc.p, V * 4));
    O3_TRY(d2h(c, rep, c->v_rep.p, V * 4));
    O3_TRY(d2h(c, count, c->v_count.p, V * 4));
    O3_TRY(d2h(c, centroid, c->v_centroid.p, V * 24));
    O3_TRY(d2h(c, normal, c->v_normal.p, V * 24));
    O3_TRY(d2h(c, color, c->v_color.p, V * 24));
    O3_TRY(d2h(c, cubic, c->v_cubic.p, V * 32));
    O3_TRY(d2h(c, order, c->v_order.p, n * 4));
    std::vector<int32_t> h;
    if (seg) {
        h.resize(V + 1);
        O3_TRY(d2h(c, h.data(), c->v_seg.p, V * 4));
    }
    CU_TRY(cudaStreamSynchronize(c->stream));
    if (seg) {
        for (size_t v = 0; v < V; ++v) seg[v] = h[v];
        seg[V] = (int64_t)n;
    }
    return 0;
}

static int voxel_select_impl(o3dpp_ctx *c)
{
    if (!c->vox_valid) return set_err(O3DPP_E_STATE, "o3dpp_voxel has not been run on the current cloud");
    // apply_selection invalidates the voxel state; keep the list alive in a private buffer
    DevBuf sel;
    int64_t V = c->vox_V;
    O3_TRY(sel.ensure((size_t)(V + 64) * 4));
    cudaError_t e = cudaMemcpyAsync(sel.p, c->v_rep_asc.p, (size_t)V * 4, cudaMemcpyDeviceToDevice, c->stream);
    int rc = e == cudaSuccess ? apply_selection(c, sel.as<int32_t>(), V) : set_err(O3DPP_E_CUDA, "copy failed");
    cudaStreamSynchronize(c->stream);
    sel.release();
    return rc;
}
