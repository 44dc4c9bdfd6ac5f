/*
 * o3d_oracle.c — CPU ORACLE for the open3dpypro point-cloud hot path.
 *
 * THIS FILE IS TEST INFRASTRUCTURE, NOT PRODUCT CODE.  Only tests/, the
 * smoke() check in __graft_entry__.py and bench.py's cpu_baseline /
 * --impl reference legs may load it.  The product (libo3dpp_b200.so) never
 * links, loads or falls back to anything in oracle/.
 *
 * PARITY UNPINNED: the arithmetic of this path is not in /root/reference.  The
 * reference delegates it to the un-vendored binary wheel open3d==0.19.0
 * (reference uv.lock:3108-3109; call sites open3dpypro/PointCloud.py:61, :72,
 * :76, :339-340), which can be neither imported nor installed here, and the
 * reference ships no golden vectors or assertions for the path
 * (SURVEY.md §4).  What follows restates Open3D 0.19.0's published algorithms
 * (cpp/open3d/geometry/{PointCloud,EstimateNormals,PointCloudSegmentation,
 * KDTreeFlann,Geometry3D}.cpp, utility/{Eigen.h,Random.cpp}, nanoflann's
 * L2 metric and KNN result set) from memory, in plain C, FP64, built with
 * -ffp-contract=off so that a*b+c is never fused.  It is cross-checked in
 * tests/ against independent implementations (scipy cKDTree, numpy eigh,
 * numpy plane scoring), not against Open3D itself.
 *
 * Deliberate, documented deviations from upstream (both deterministic
 * definitions of behaviour that upstream leaves implementation-defined):
 *   - KNN ties are ordered by (d^2, index)           (north_star; upstream:
 *     kd-tree traversal order, NANOFLANN_FIRST_MATCH undefined)
 *   - voxel rows are emitted in ascending (ix,iy,iz)  (upstream: hash order)
 *   - RANSAC selection is the single-thread replay    (upstream: OpenMP order)
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#define ORC_API __attribute__((visibility("default")))

/* ------------------------------------------------------------------------- */
/* transform — PointCloud.py:60-62 -> Geometry3D::TransformPoints/Normals    */
/* p' = (T [p,1]).xyz / w ;  n' = (T [n,0]).xyz  (no renormalisation)        */
/* ------------------------------------------------------------------------- */
ORC_API int orc_transform(double *pts, double *normals, int64_t n, const double *T)
{
    for (int64_t i = 0; i < n; ++i) {
        double x = pts[3 * i], y = pts[3 * i + 1], z = pts[3 * i + 2];
        double r[4];
        for (int a = 0; a < 4; ++a)
            r[a] = ((T[4 * a] * x + T[4 * a + 1] * y) + T[4 * a + 2] * z) + T[4 * a + 3] * 1.0;
        pts[3 * i] = r[0] / r[3];
        pts[3 * i + 1] = r[1] / r[3];
        pts[3 * i + 2] = r[2] / r[3];
    }
    if (normals) {
        for (int64_t i = 0; i < n; ++i) {
            double x = normals[3 * i], y = normals[3 * i + 1], z = normals[3 * i + 2];
            double r[3];
            for (int a = 0; a < 3; ++a)
                r[a] = ((T[4 * a] * x + T[4 * a + 1] * y) + T[4 * a + 2] * z) + T[4 * a + 3] * 0.0;
            normals[3 * i] = r[0];
            normals[3 * i + 1] = r[1];
            normals[3 * i + 2] = r[2];
        }
    }
    return 0;
}

/* ------------------------------------------------------------------------- */
/* voxel_down_sample_and_trace — PointCloud.py:338-341 ->                    */
/* PointCloud::VoxelDownSampleAndTrace(voxel, min_bound, max_bound, false)   */
/* ------------------------------------------------------------------------- */
ORC_API int orc_bounds(const double *pts, int64_t n, double *mn, double *mx)
{
    if (n <= 0) { for (int a = 0; a < 3; ++a) mn[a] = mx[a] = 0.0; return 0; }
    for (int a = 0; a < 3; ++a) mn[a] = mx[a] = pts[a];
    for (int64_t i = 1; i < n; ++i)
        for (int a = 0; a < 3; ++a) {
            double v = pts[3 * i + a];
            if (v < mn[a]) mn[a] = v;
            if (v > mx[a]) mx[a] = v;
        }
    return 0;
}

typedef struct { int32_t ix, iy, iz; int32_t i; } vox_ent;

static int vox_cmp(const void *pa, const void *pb)
{
    const vox_ent *a = (const vox_ent *)pa, *b = (const vox_ent *)pb;
    if (a->ix != b->ix) return a->ix < b->ix ? -1 : 1;
    if (a->iy != b->iy) return a->iy < b->iy ? -1 : 1;
    if (a->iz != b->iz) return a->iz < b->iz ? -1 : 1;
    return a->i < b->i ? -1 : (a->i > b->i);
}

/* Phase 1: voxel index + octant per point, sorted by (ix,iy,iz,i).
 * Returns V (>=0) or a negative error:
 *   -1 voxel_size <= 0            (upstream: "voxel_size <= 0.")
 *   -2 voxel_size*INT_MAX < extent (upstream: "voxel_size is too small.")
 * out_vidx : n*3 int32 voxel index per ORIGINAL point
 * out_oct  : n   int32 octant (cubic id) per ORIGINAL point
 * out_order: n   int32 original indices sorted by (voxel, index)
 * out_seg  : n+1 int64 segment offsets into out_order (first V+1 valid)      */
/* min3 / max3 (nullable): the explicit grid origin and far corner of PointCloud::VoxelDownSampleAndTrace(voxel_size,
 * min_bound, max_bound, ...) as the reference passes them (PointCloud.py:339-340); NULL = the cloud's own bounds, which
 * is what the reference always passes.  The "trace" variant uses min_bound itself as the origin (no half-voxel shift)
 * and max_bound only in the size check. */
ORC_API int64_t orc_voxel_index_bounded(const double *pts, int64_t n, double voxel_size, const double *min3,
                                        const double *max3, int32_t *out_vidx, int32_t *out_oct,
                                        int32_t *out_order, int64_t *out_seg);

ORC_API int64_t orc_voxel_index(const double *pts, int64_t n, double voxel_size,
                                int32_t *out_vidx, int32_t *out_oct,
                                int32_t *out_order, int64_t *out_seg)
{
    return orc_voxel_index_bounded(pts, n, voxel_size, NULL, NULL, out_vidx, out_oct, out_order, out_seg);
}

ORC_API int64_t orc_voxel_index_bounded(const double *pts, int64_t n, double voxel_size, const double *min3,
                                        const double *max3, int32_t *out_vidx, int32_t *out_oct,
                                        int32_t *out_order, int64_t *out_seg)
{
    if (!(voxel_size > 0.0)) return -1;
    double mn[3], mx[3];
    orc_bounds(pts, n, mn, mx);
    if (min3) { mn[0] = min3[0]; mn[1] = min3[1]; mn[2] = min3[2]; }
    if (max3) { mx[0] = max3[0]; mx[1] = max3[1]; mx[2] = max3[2]; }
    double ext = mx[0] - mn[0];
    if (mx[1] - mn[1] > ext) ext = mx[1] - mn[1];
    if (mx[2] - mn[2] > ext) ext = mx[2] - mn[2];
    if (voxel_size * 2147483647.0 < ext) return -2;
    vox_ent *e = (vox_ent *)malloc(sizeof(vox_ent) * (size_t)(n > 0 ? n : 1));
    for (int64_t i = 0; i < n; ++i) {
        int32_t v[3];
        int oct = 0;
        for (int a = 0; a < 3; ++a) {
            double ref = (pts[3 * i + a] - mn[a]) / voxel_size;
            v[a] = (int32_t)floor(ref);
            if ((ref - (double)v[a]) >= 0.5) oct += (1 << a);
            out_vidx[3 * i + a] = v[a];
        }
        out_oct[i] = oct;
        e[i].ix = v[0]; e[i].iy = v[1]; e[i].iz = v[2]; e[i].i = (int32_t)i;
    }
    qsort(e, (size_t)n, sizeof(vox_ent), vox_cmp);
    int64_t V = 0;
    for (int64_t j = 0; j < n; ++j) {
        if (j == 0 || e[j].ix != e[j - 1].ix || e[j].iy != e[j - 1].iy || e[j].iz != e[j - 1].iz)
            out_seg[V++] = j;
        out_order[j] = e[j].i;
    }
    out_seg[V] = n;
    free(e);
    return V;
}

/* Phase 2: per-voxel accumulation in ascending original index (the order in
 * which upstream's serial loop calls AddPoint).  Rows in ascending voxel key.
 * cubic_id[row][oct] = last (= largest) original index seen in that octant.  */
ORC_API int orc_voxel_reduce(const double *pts, const double *normals, const double *colors,
                             int64_t V, const int32_t *order, const int64_t *seg,
                             const int32_t *oct,
                             int32_t *cubic_id, int32_t *rep, int32_t *count,
                             double *centroid, double *out_normal, double *out_color)
{
    for (int64_t v = 0; v < V; ++v) {
        double sp[3] = {0, 0, 0}, sn[3] = {0, 0, 0}, sc[3] = {0, 0, 0};
        for (int c = 0; c < 8; ++c) cubic_id[8 * v + c] = -1;
        int64_t cnt = seg[v + 1] - seg[v];
        for (int64_t j = seg[v]; j < seg[v + 1]; ++j) {
            int32_t i = order[j];
            for (int a = 0; a < 3; ++a) {
                sp[a] += pts[3 * (int64_t)i + a];
                if (normals) sn[a] += normals[3 * (int64_t)i + a];
                if (colors) sc[a] += colors[3 * (int64_t)i + a];
            }
            cubic_id[8 * v + oct[i]] = i;
        }
        int32_t m = -1;
        for (int c = 0; c < 8; ++c) if (cubic_id[8 * v + c] > m) m = cubic_id[8 * v + c];
        rep[v] = m;
        count[v] = (int32_t)cnt;
        for (int a = 0; a < 3; ++a) centroid[3 * v + a] = sp[a] / (double)cnt;
        if (normals && out_normal) {
            double nn = sqrt((sn[0] * sn[0] + sn[1] * sn[1]) + sn[2] * sn[2]);
            /* Eigen normalized(): divides by the norm when it is > 0 */
            for (int a = 0; a < 3; ++a) out_normal[3 * v + a] = nn > 0.0 ? sn[a] / nn : sn[a];
        }
        if (colors && out_color)
            for (int a = 0; a < 3; ++a) out_color[3 * v + a] = sc[a] / (double)cnt;
    }
    return 0;
}

/* ------------------------------------------------------------------------- */
/* k-d tree with (d^2, index) ordered results — stands in for KDTreeFlann    */
/* d^2 = (dx*dx + dy*dy) + dz*dz in FP64, unfused  (nanoflann L2 adaptor)    */
/* ------------------------------------------------------------------------- */
#define KD_LEAF 16

typedef struct {
    const double *pts;
    int64_t n;
    int32_t *perm;     /* n */
    int32_t *node_lo;  /* per node: first perm slot */
    int32_t *node_hi;  /* per node: one past last   */
    int8_t *node_dim;  /* -1 = leaf */
    double *node_split;
    int32_t *node_left, *node_right;
    int32_t n_nodes, cap_nodes;
} kdtree;

static inline double dist2(const double *a, const double *b)
{
    double dx = a[0] - b[0], dy = a[1] - b[1], dz = a[2] - b[2];
    return (dx * dx + dy * dy) + dz * dz;
}

static void kd_select(const double *pts, int32_t *perm, int32_t lo, int32_t hi, int32_t mid, int dim)
{
    /* quickselect on coordinate `dim`, ties by index for determinism */
    while (hi - lo > 1) {
        int32_t p = perm[lo + (hi - lo) / 2];
        double pv = pts[3 * (int64_t)p + dim];
        int32_t i = lo, j = hi - 1;
        while (i <= j) {
            while (1) { int32_t q = perm[i]; double v = pts[3 * (int64_t)q + dim];
                        if (v < pv || (v == pv && q < p)) ++i; else break; }
            while (1) { int32_t q = perm[j]; double v = pts[3 * (int64_t)q + dim];
                        if (v > pv || (v == pv && q > p)) --j; else break; }
            if (i <= j) { int32_t t = perm[i]; perm[i] = perm[j]; perm[j] = t; ++i; --j; }
        }
        if (mid <= j) hi = j + 1;
        else if (mid >= i) lo = i;
        else return;
    }
}

static int32_t kd_build_rec(kdtree *t, int32_t lo, int32_t hi)
{
    int32_t id = t->n_nodes++;
    t->node_lo[id] = lo; t->node_hi[id] = hi;
    if (hi - lo <= KD_LEAF) { t->node_dim[id] = -1; t->node_left[id] = t->node_right[id] = -1; return id; }
    double mn[3], mx[3];
    for (int a = 0; a < 3; ++a) mn[a] = mx[a] = t->pts[3 * (int64_t)t->perm[lo] + a];
    for (int32_t j = lo + 1; j < hi; ++j)
        for (int a = 0; a < 3; ++a) {
            double v = t->pts[3 * (int64_t)t->perm[j] + a];
            if (v < mn[a]) mn[a] = v;
            if (v > mx[a]) mx[a] = v;
        }
    int dim = 0;
    if (mx[1] - mn[1] > mx[dim] - mn[dim]) dim = 1;
    if (mx[2] - mn[2] > mx[dim] - mn[dim]) dim = 2;
    int32_t mid = lo + (hi - lo) / 2;
    kd_select(t->pts, t->perm, lo, hi, mid, dim);
    t->node_dim[id] = (int8_t)dim;
    t->node_split[id] = t->pts[3 * (int64_t)t->perm[mid] + dim];
    int32_t l = kd_build_rec(t, lo, mid);
    int32_t r = kd_build_rec(t, mid, hi);
    t->node_left[id] = l; t->node_right[id] = r;
    return id;
}

static kdtree *kd_build(const double *pts, int64_t n)
{
    kdtree *t = (kdtree *)calloc(1, sizeof(kdtree));
    t->pts = pts; t->n = n;
    t->perm = (int32_t *)malloc(sizeof(int32_t) * (size_t)(n > 0 ? n : 1));
    for (int64_t i = 0; i < n; ++i) t->perm[i] = (int32_t)i;
    int64_t cap = 2 * (n / (KD_LEAF / 2) + 2) + 8;
    t->cap_nodes = (int32_t)cap;
    t->node_lo = (int32_t *)malloc(sizeof(int32_t) * cap);
    t->node_hi = (int32_t *)malloc(sizeof(int32_t) * cap);
    t->node_dim = (int8_t *)malloc(cap);
    t->node_split = (double *)malloc(sizeof(double) * cap);
    t->node_left = (int32_t *)malloc(sizeof(int32_t) * cap);
    t->node_right = (int32_t *)malloc(sizeof(int32_t) * cap);
    t->n_nodes = 0;
    if (n > 0) kd_build_rec(t, 0, (int32_t)n);
    return t;
}

static void kd_free(kdtree *t)
{
    if (!t) return;
    free(t->perm); free(t->node_lo); free(t->node_hi); free(t->node_dim);
    free(t->node_split); free(t->node_left); free(t->node_right); free(t);
}

/* bounded result list sorted ascending by (d2, idx) */
typedef struct { double *d2; int32_t *idx; int cap; int cnt; double limit2; /* strict d2 < limit2, <0: none */ } knn_set;

static inline int key_less(double d2a, int32_t ia, double d2b, int32_t ib)
{
    return d2a < d2b || (d2a == d2b && ia < ib);
}

static inline void knn_offer(knn_set *s, double d2, int32_t i)
{
    if (s->limit2 >= 0.0 && !(d2 < s->limit2)) return;
    if (s->cnt == s->cap && !key_less(d2, i, s->d2[s->cap - 1], s->idx[s->cap - 1])) return;
    int j = s->cnt < s->cap ? s->cnt : s->cap - 1;
    while (j > 0 && key_less(d2, i, s->d2[j - 1], s->idx[j - 1])) {
        s->d2[j] = s->d2[j - 1]; s->idx[j] = s->idx[j - 1]; --j;
    }
    s->d2[j] = d2; s->idx[j] = i;
    if (s->cnt < s->cap) s->cnt++;
}

static void kd_knn_rec(const kdtree *t, int32_t node, const double *q, knn_set *s)
{
    if (t->node_dim[node] < 0) {
        for (int32_t j = t->node_lo[node]; j < t->node_hi[node]; ++j) {
            int32_t i = t->perm[j];
            knn_offer(s, dist2(q, t->pts + 3 * (int64_t)i), i);
        }
        return;
    }
    int dim = t->node_dim[node];
    double diff = q[dim] - t->node_split[node];
    int32_t nearc = diff < 0 ? t->node_left[node] : t->node_right[node];
    int32_t farc = diff < 0 ? t->node_right[node] : t->node_left[node];
    kd_knn_rec(t, nearc, q, s);
    double b = diff * diff; /* lower bound on any d2 behind the split plane */
    int visit = 1;
    if (s->limit2 >= 0.0 && !(b < s->limit2)) visit = 0;
    if (visit && s->cnt == s->cap && b > s->d2[s->cap - 1]) visit = 0; /* equal: index may still win */
    if (visit) kd_knn_rec(t, farc, q, s);
}

/* radius search: visit every point with d2 < r2, in tree order */
typedef void (*radius_cb)(void *ctx, int32_t i, double d2);
static void kd_radius_rec(const kdtree *t, int32_t node, const double *q, double r2, radius_cb cb, void *ctx)
{
    if (t->node_dim[node] < 0) {
        for (int32_t j = t->node_lo[node]; j < t->node_hi[node]; ++j) {
            int32_t i = t->perm[j];
            double d2 = dist2(q, t->pts + 3 * (int64_t)i);
            if (d2 < r2) cb(ctx, i, d2);
        }
        return;
    }
    int dim = t->node_dim[node];
    double diff = q[dim] - t->node_split[node];
    int32_t nearc = diff < 0 ? t->node_left[node] : t->node_right[node];
    int32_t farc = diff < 0 ? t->node_right[node] : t->node_left[node];
    kd_radius_rec(t, nearc, q, r2, cb, ctx);
    if (diff * diff < r2) kd_radius_rec(t, farc, q, r2, cb, ctx);
}

/* Public KNN. mode 0 = KNN(k); mode 2 = Hybrid(radius,k): KNN(k) then keep the
 * prefix with d2 < radius^2 (KDTreeFlann::SearchHybrid).  Rows padded with -1.
 * brute != 0 forces the O(N*Q) scan (used to validate the tree).             */
ORC_API int orc_knn(const double *pts, int64_t n, const double *queries, int64_t nq,
                    int mode, int k, double radius, int brute,
                    int32_t *out_idx, double *out_d2, int32_t *out_cnt, int nthreads)
{
    if (k <= 0) return -1;
    kdtree *t = brute ? NULL : kd_build(pts, n);
#ifdef _OPENMP
    if (nthreads > 0) omp_set_num_threads(nthreads);
#endif
#pragma omp parallel
    {
        double *d2 = (double *)malloc(sizeof(double) * (size_t)k);
        int32_t *ix = (int32_t *)malloc(sizeof(int32_t) * (size_t)k);
#pragma omp for schedule(dynamic, 256)
        for (int64_t qi = 0; qi < nq; ++qi) {
            knn_set s = { d2, ix, k, 0, -1.0 };
            const double *q = queries + 3 * qi;
            if (brute || n == 0) {
                for (int64_t i = 0; i < n; ++i) knn_offer(&s, dist2(q, pts + 3 * i), (int32_t)i);
            } else {
                kd_knn_rec(t, 0, q, &s);
            }
            int cnt = s.cnt;
            if (mode == 2) {
                double r2 = radius * radius;
                int c = 0;
                while (c < cnt && d2[c] < r2) ++c;
                cnt = c;
            }
            for (int j = 0; j < k; ++j) {
                out_idx[qi * k + j] = j < cnt ? ix[j] : -1;
                if (out_d2) out_d2[qi * k + j] = j < cnt ? d2[j] : -1.0;
            }
            if (out_cnt) out_cnt[qi] = cnt;
        }
        free(d2); free(ix);
    }
    kd_free(t);
    return 0;
}

/* ------------------------------------------------------------------------- */
/* covariance + closed-form smallest eigenvector                             */
/* utility::ComputeCovariance, FastEigen3x3 (EstimateNormals.cpp)            */
/* ------------------------------------------------------------------------- */
typedef struct { double c[9]; int64_t n; } cumul;

static inline void cumul_add(cumul *u, const double *p)
{
    u->c[0] += p[0]; u->c[1] += p[1]; u->c[2] += p[2];
    u->c[3] += p[0] * p[0]; u->c[4] += p[0] * p[1]; u->c[5] += p[0] * p[2];
    u->c[6] += p[1] * p[1]; u->c[7] += p[1] * p[2]; u->c[8] += p[2] * p[2];
    u->n++;
}

/* C stored as [c00 c01 c02 c11 c12 c22] */
static inline void cumul_cov(const cumul *u, double *C)
{
    double m[9];
    for (int a = 0; a < 9; ++a) m[a] = u->c[a] / (double)u->n;
    C[0] = m[3] - m[0] * m[0];
    C[1] = m[4] - m[0] * m[1];
    C[2] = m[5] - m[0] * m[2];
    C[3] = m[6] - m[1] * m[1];
    C[4] = m[7] - m[1] * m[2];
    C[5] = m[8] - m[2] * m[2];
}

static inline void cross3(const double *a, const double *b, double *o)
{
    o[0] = a[1] * b[2] - a[2] * b[1];
    o[1] = a[2] * b[0] - a[0] * b[2];
    o[2] = a[0] * b[1] - a[1] * b[0];
}
static inline double dot3(const double *a, const double *b) { return (a[0] * b[0] + a[1] * b[1]) + a[2] * b[2]; }

/* eigenvector for the best-separated eigenvalue: largest row cross product */
static void eigvec0(const double *A, double ev, double *out)
{
    double r0[3] = { A[0] - ev, A[1], A[2] };
    double r1[3] = { A[1], A[3] - ev, A[4] };
    double r2[3] = { A[2], A[4], A[5] - ev };
    double x01[3], x02[3], x12[3];
    cross3(r0, r1, x01); cross3(r0, r2, x02); cross3(r1, r2, x12);
    double d0 = dot3(x01, x01), d1 = dot3(x02, x02), d2 = dot3(x12, x12);
    double dmax = d0; int imax = 0;
    if (d1 > dmax) { dmax = d1; imax = 1; }
    if (d2 > dmax) { imax = 2; }
    const double *src = imax == 0 ? x01 : (imax == 1 ? x02 : x12);
    double dd = imax == 0 ? d0 : (imax == 1 ? d1 : d2);
    double s = sqrt(dd);
    out[0] = src[0] / s; out[1] = src[1] / s; out[2] = src[2] / s;
}

/* second eigenvector in the plane orthogonal to e0 */
static void eigvec1(const double *A, const double *e0, double ev1, double *out)
{
    double U[3], V[3];
    if (fabs(e0[0]) > fabs(e0[1])) {
        double il = 1.0 / sqrt(e0[0] * e0[0] + e0[2] * e0[2]);
        U[0] = -e0[2] * il; U[1] = 0.0; U[2] = e0[0] * il;
    } else {
        double il = 1.0 / sqrt(e0[1] * e0[1] + e0[2] * e0[2]);
        U[0] = 0.0; U[1] = e0[2] * il; U[2] = -e0[1] * il;
    }
    cross3(e0, U, V);
    double AU[3] = { (A[0] * U[0] + A[1] * U[1]) + A[2] * U[2],
                     (A[1] * U[0] + A[3] * U[1]) + A[4] * U[2],
                     (A[2] * U[0] + A[4] * U[1]) + A[5] * U[2] };
    double AV[3] = { (A[0] * V[0] + A[1] * V[1]) + A[2] * V[2],
                     (A[1] * V[0] + A[3] * V[1]) + A[4] * V[2],
                     (A[2] * V[0] + A[4] * V[1]) + A[5] * V[2] };
    double m00 = dot3(U, AU) - ev1;
    double m01 = dot3(U, AV);
    double m11 = dot3(V, AV) - ev1;
    double a00 = fabs(m00), a01 = fabs(m01), a11 = fabs(m11);
    if (a00 >= a11) {
        double mx = a00 > a01 ? a00 : a01;
        if (mx > 0) {
            if (a00 >= a01) { m01 /= m00; m00 = 1.0 / sqrt(1.0 + m01 * m01); m01 *= m00; }
            else            { m00 /= m01; m01 = 1.0 / sqrt(1.0 + m00 * m00); m00 *= m01; }
            for (int a = 0; a < 3; ++a) out[a] = m01 * U[a] - m00 * V[a];
        } else { for (int a = 0; a < 3; ++a) out[a] = U[a]; }
    } else {
        double mx = a11 > a01 ? a11 : a01;
        if (mx > 0) {
            if (a11 >= a01) { m01 /= m11; m11 = 1.0 / sqrt(1.0 + m01 * m01); m01 *= m11; }
            else            { m11 /= m01; m01 = 1.0 / sqrt(1.0 + m11 * m11); m11 *= m01; }
            for (int a = 0; a < 3; ++a) out[a] = m11 * U[a] - m01 * V[a];
        } else { for (int a = 0; a < 3; ++a) out[a] = U[a]; }
    }
}

/* C = [c00 c01 c02 c11 c12 c22] -> eigenvector of the smallest eigenvalue */
ORC_API void orc_fast_eigen3x3(const double *C, double *out)
{
    double mc = C[0];
    for (int a = 1; a < 6; ++a) if (C[a] > mc) mc = C[a];   /* maxCoeff, not abs-max */
    if (mc == 0.0) { out[0] = out[1] = out[2] = 0.0; return; }
    double A[6];
    for (int a = 0; a < 6; ++a) A[a] = C[a] / mc;
    double norm = (A[1] * A[1] + A[2] * A[2]) + A[4] * A[4];
    if (norm > 0.0) {
        double q = ((A[0] + A[3]) + A[5]) / 3.0;
        double b00 = A[0] - q, b11 = A[3] - q, b22 = A[5] - q;
        double p = sqrt((((b00 * b00 + b11 * b11) + b22 * b22) + norm * 2.0) / 6.0);
        double c00 = b11 * b22 - A[4] * A[4];
        double c01 = A[1] * b22 - A[4] * A[2];
        double c02 = A[1] * A[4] - b11 * A[2];
        double det = ((b00 * c00 - A[1] * c01) + A[2] * c02) / (p * p * p);
        double half_det = det * 0.5;
        if (half_det < -1.0) half_det = -1.0;
        if (half_det > 1.0) half_det = 1.0;
        double angle = acos(half_det) / 3.0;
        const double two_thirds_pi = 2.09439510239319549;
        double beta2 = cos(angle) * 2.0;
        double beta0 = cos(angle + two_thirds_pi) * 2.0;
        double beta1 = -(beta0 + beta2);
        double e0 = q + p * beta0, e1 = q + p * beta1, e2 = q + p * beta2;
        double v0[3], v1[3], v2[3];
        if (half_det >= 0.0) {
            eigvec0(A, e2, v2);
            if (e2 < e0 && e2 < e1) { memcpy(out, v2, sizeof v2); return; }
            eigvec1(A, v2, e1, v1);
            if (e1 < e0 && e1 < e2) { memcpy(out, v1, sizeof v1); return; }
            cross3(v1, v2, v0);
            memcpy(out, v0, sizeof v0);
        } else {
            eigvec0(A, e0, v0);
            if (e0 < e1 && e0 < e2) { memcpy(out, v0, sizeof v0); return; }
            eigvec1(A, v0, e1, v1);
            if (e1 < e0 && e1 < e2) { memcpy(out, v1, sizeof v1); return; }
            cross3(v0, v1, v2);
            memcpy(out, v2, sizeof v2);
        }
    } else {
        /* diagonal matrix: axis of the smallest diagonal entry, default +z */
        if (C[0] < C[3] && C[0] < C[5]) { out[0] = 1; out[1] = 0; out[2] = 0; }
        else if (C[3] < C[0] && C[3] < C[5]) { out[0] = 0; out[1] = 1; out[2] = 0; }
        else { out[0] = 0; out[1] = 0; out[2] = 1; }
    }
}

typedef struct { cumul u; const double *pts; } rad_ctx;
static void rad_acc(void *ctx, int32_t i, double d2) { (void)d2; rad_ctx *r = (rad_ctx *)ctx; cumul_add(&r->u, r->pts + 3 * (int64_t)i); }

/* estimate_normals — PointCloud.py:68-73 -> PointCloud::EstimateNormals with
 * fast_normal_computation = true.  mode 0 KNN(k), 1 Radius(r), 2 Hybrid(r,k).
 * old_normals (nullable): pre-existing normals used for sign alignment.
 * out_cov (nullable): 6 doubles per point [c00 c01 c02 c11 c12 c22].
 * Radius mode accumulates in tree order (upstream: sorted by d2); the set is
 * the same, only the FP64 summation order differs.                          */
ORC_API int orc_estimate_normals(const double *pts, int64_t n, int mode, int k, double radius,
                                 const double *old_normals, double *out_normals, double *out_cov,
                                 int nthreads)
{
    if (n <= 0) return 0;
    if ((mode == 0 || mode == 2) && k <= 0) return -1;
    kdtree *t = kd_build(pts, n);
#ifdef _OPENMP
    if (nthreads > 0) omp_set_num_threads(nthreads);
#endif
#pragma omp parallel
    {
        int kk = k > 0 ? k : 1;
        double *d2 = (double *)malloc(sizeof(double) * (size_t)kk);
        int32_t *ix = (int32_t *)malloc(sizeof(int32_t) * (size_t)kk);
#pragma omp for schedule(dynamic, 256)
        for (int64_t i = 0; i < n; ++i) {
            const double *q = pts + 3 * i;
            double C[6];
            cumul u; memset(&u, 0, sizeof u);
            if (mode == 1) {
                rad_ctx rc; memset(&rc, 0, sizeof rc); rc.pts = pts;
                kd_radius_rec(t, 0, q, radius * radius, rad_acc, &rc);
                u = rc.u;
            } else {
                knn_set s = { d2, ix, k, 0, -1.0 };
                kd_knn_rec(t, 0, q, &s);
                int cnt = s.cnt;
                if (mode == 2) { double r2 = radius * radius; int c = 0; while (c < cnt && d2[c] < r2) ++c; cnt = c; }
                for (int j = 0; j < cnt; ++j) cumul_add(&u, pts + 3 * (int64_t)ix[j]);
            }
            if (u.n >= 3) cumul_cov(&u, C);
            else { C[0] = 1; C[1] = 0; C[2] = 0; C[3] = 1; C[4] = 0; C[5] = 1; }
            if (out_cov) memcpy(out_cov + 6 * i, C, sizeof C);
            double nv[3];
            orc_fast_eigen3x3(C, nv);
            double nn = sqrt((nv[0] * nv[0] + nv[1] * nv[1]) + nv[2] * nv[2]);
            if (nn == 0.0) {
                if (old_normals) { nv[0] = old_normals[3 * i]; nv[1] = old_normals[3 * i + 1]; nv[2] = old_normals[3 * i + 2]; }
                else { nv[0] = 0; nv[1] = 0; nv[2] = 1; }
            }
            if (old_normals && dot3(nv, old_normals + 3 * i) < 0.0) { nv[0] = -nv[0]; nv[1] = -nv[1]; nv[2] = -nv[2]; }
            out_normals[3 * i] = nv[0]; out_normals[3 * i + 1] = nv[1]; out_normals[3 * i + 2] = nv[2];
        }
        free(d2); free(ix);
    }
    kd_free(t);
    return 0;
}

/* ------------------------------------------------------------------------- */
/* RANSAC samplers                                                           */
/* ------------------------------------------------------------------------- */
/* (1) std::mt19937 replay of utility::random::RandUint32() % N with
 *     rejection of duplicates inside a sample (RandomSampler<size_t>).      */
typedef struct { uint32_t mt[624]; int idx; } mt19937_t;
static void mt_seed(mt19937_t *m, uint32_t s)
{
    m->mt[0] = s;
    for (int i = 1; i < 624; ++i) m->mt[i] = 1812433253u * (m->mt[i - 1] ^ (m->mt[i - 1] >> 30)) + (uint32_t)i;
    m->idx = 624;
}
static uint32_t mt_next(mt19937_t *m)
{
    if (m->idx >= 624) {
        for (int i = 0; i < 624; ++i) {
            uint32_t y = (m->mt[i] & 0x80000000u) | (m->mt[(i + 1) % 624] & 0x7fffffffu);
            m->mt[i] = m->mt[(i + 397) % 624] ^ (y >> 1) ^ ((y & 1u) ? 0x9908b0dfu : 0u);
        }
        m->idx = 0;
    }
    uint32_t y = m->mt[m->idx++];
    y ^= y >> 11; y ^= (y << 7) & 0x9d2c5680u; y ^= (y << 15) & 0xefc60000u; y ^= y >> 18;
    return y;
}
ORC_API uint32_t orc_mt19937_nth(uint32_t seed, int nth)
{
    mt19937_t m; mt_seed(&m, seed);
    uint32_t v = 0;
    for (int i = 0; i <= nth; ++i) v = mt_next(&m);
    return v;
}
ORC_API int orc_sample_mt19937(uint32_t seed, int64_t n, int ransac_n, int H, int64_t *out)
{
    if (n < ransac_n || ransac_n <= 0) return -1;
    mt19937_t m; mt_seed(&m, seed);
    for (int h = 0; h < H; ++h) {
        int got = 0;
        while (got < ransac_n) {
            int64_t c = (int64_t)(mt_next(&m) % (uint64_t)n);
            int dup = 0;
            for (int j = 0; j < got; ++j) if (out[(int64_t)h * ransac_n + j] == c) dup = 1;
            if (!dup) out[(int64_t)h * ransac_n + got++] = c;
        }
    }
    return 0;
}

/* (2) counter-based Philox4x32-10: key = seed, counter = (draw, 0, iteration,
 *     tag).  idx = (x1:x0 as u64) % N, duplicates rejected by bumping `draw`.
 *     The CUDA library generates the same stream on the device.              */
static inline void philox4x32_10(uint32_t c[4], uint32_t k0, uint32_t k1)
{
    for (int r = 0; r < 10; ++r) {
        uint64_t p0 = (uint64_t)0xD2511F53u * c[0];
        uint64_t p1 = (uint64_t)0xCD9E8D57u * c[2];
        uint32_t n0 = (uint32_t)(p1 >> 32) ^ c[1] ^ k0;
        uint32_t n1 = (uint32_t)p1;
        uint32_t n2 = (uint32_t)(p0 >> 32) ^ c[3] ^ k1;
        uint32_t n3 = (uint32_t)p0;
        c[0] = n0; c[1] = n1; c[2] = n2; c[3] = n3;
        k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
    }
}
ORC_API void orc_philox(uint64_t seed, uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t *out4)
{
    uint32_t c[4] = { c0, c1, c2, c3 };
    philox4x32_10(c, (uint32_t)seed, (uint32_t)(seed >> 32));
    memcpy(out4, c, sizeof c);
}
#define ORC_PHILOX_TAG 0x52414E53u /* "RANS" */
ORC_API int orc_sample_philox(uint64_t seed, int64_t n, int ransac_n, int H, int64_t *out)
{
    if (n < ransac_n || ransac_n <= 0) return -1;
    for (int h = 0; h < H; ++h) {
        int got = 0;
        uint32_t draw = 0;
        while (got < ransac_n) {
            uint32_t c[4] = { draw++, 0u, (uint32_t)h, ORC_PHILOX_TAG };
            philox4x32_10(c, (uint32_t)seed, (uint32_t)(seed >> 32));
            uint64_t r = ((uint64_t)c[1] << 32) | c[0];
            int64_t cand = (int64_t)(r % (uint64_t)n);
            int dup = 0;
            for (int j = 0; j < got; ++j) if (out[(int64_t)h * ransac_n + j] == cand) dup = 1;
            if (!dup) out[(int64_t)h * ransac_n + got++] = cand;
        }
    }
    return 0;
}

/* ------------------------------------------------------------------------- */
/* segment_plane — PointCloud.py:75-77 -> PointCloud::SegmentPlane           */
/* ------------------------------------------------------------------------- */
static void triangle_plane(const double *p0, const double *p1, const double *p2, double *pl)
{
    double e0[3] = { p1[0] - p0[0], p1[1] - p0[1], p1[2] - p0[2] };
    double e1[3] = { p2[0] - p0[0], p2[1] - p0[1], p2[2] - p0[2] };
    double abc[3];
    cross3(e0, e1, abc);
    double nn = sqrt(dot3(abc, abc));
    if (nn == 0.0) { pl[0] = pl[1] = pl[2] = pl[3] = 0.0; return; }
    abc[0] /= nn; abc[1] /= nn; abc[2] /= nn;
    pl[0] = abc[0]; pl[1] = abc[1]; pl[2] = abc[2];
    pl[3] = -dot3(abc, p0);
}

/* GetPlaneFromPoints: centroid, centred second moments, largest 2x2 minor */
static void plane_from_points(const double *pts, const int64_t *idx, const int32_t *idx32, int64_t m, double *pl)
{
    double c[3] = { 0, 0, 0 };
    for (int64_t j = 0; j < m; ++j) {
        const double *p = pts + 3 * (idx ? idx[j] : (int64_t)idx32[j]);
        c[0] += p[0]; c[1] += p[1]; c[2] += p[2];
    }
    c[0] /= (double)m; c[1] /= (double)m; c[2] /= (double)m;
    double xx = 0, xy = 0, xz = 0, yy = 0, yz = 0, zz = 0;
    for (int64_t j = 0; j < m; ++j) {
        const double *p = pts + 3 * (idx ? idx[j] : (int64_t)idx32[j]);
        double rx = p[0] - c[0], ry = p[1] - c[1], rz = p[2] - c[2];
        xx += rx * rx; xy += rx * ry; xz += rx * rz;
        yy += ry * ry; yz += ry * rz; zz += rz * rz;
    }
    double det_x = yy * zz - yz * yz;
    double det_y = xx * zz - xz * xz;
    double det_z = xx * yy - xy * xy;
    double abc[3];
    if (det_x > det_y && det_x > det_z) { abc[0] = det_x; abc[1] = xz * yz - xy * zz; abc[2] = xy * yz - xz * yy; }
    else if (det_y > det_z)             { abc[0] = xz * yz - xy * zz; abc[1] = det_y; abc[2] = xy * xz - yz * xx; }
    else                                { abc[0] = xy * yz - xz * yy; abc[1] = xy * xz - yz * xx; abc[2] = det_z; }
    double nn = sqrt(dot3(abc, abc));
    if (nn == 0.0 || m == 0) { pl[0] = pl[1] = pl[2] = pl[3] = 0.0; return; }
    abc[0] /= nn; abc[1] /= nn; abc[2] /= nn;
    pl[0] = abc[0]; pl[1] = abc[1]; pl[2] = abc[2];
    pl[3] = -dot3(abc, c);
}

static inline double plane_dist(const double *pl, const double *p)
{
    return fabs(((pl[0] * p[0] + pl[1] * p[1]) + pl[2] * p[2]) + pl[3]);
}

/* plane model of hypothesis h (exposed so tests can compare the GPU's models) */
ORC_API void orc_plane_model(const double *pts, const int64_t *sample, int ransac_n, double *pl)
{
    if (ransac_n == 3) triangle_plane(pts + 3 * sample[0], pts + 3 * sample[1], pts + 3 * sample[2], pl);
    else plane_from_points(pts, sample, NULL, ransac_n, pl);
}

/* score one plane over all points: inlier count and sum of distances */
ORC_API void orc_score_plane(const double *pts, int64_t n, const double *pl, double thr,
                             int64_t *count, double *err)
{
    int64_t c = 0; double e = 0.0;
    for (int64_t i = 0; i < n; ++i) {
        double d = plane_dist(pl, pts + 3 * i);
        if (d < thr) { e += d; ++c; }
    }
    *count = c; *err = e;
}

/* Error codes: -3 ransac_n < 3 ; -4 n < ransac_n  (upstream RuntimeError).
 * samples: H*ransac_n int64 (the hypothesis sequence, an explicit input).
 * Selection = single-thread replay of upstream's loop incl. early exit.
 * counts_out/err_out (nullable, H each): per hypothesis inlier count (-1 if
 * skipped by early exit, -2 if degenerate) and distance sum.
 * nthreads > 1 scores `nthreads` hypotheses at a time (like the OpenMP
 * upstream does) but applies them in order, so the result is identical.     */
ORC_API int orc_segment_plane(const double *pts, int64_t n, double thr, int ransac_n, int H,
                              double prob, const int64_t *samples,
                              double *plane_out, int32_t *inliers_out, int64_t *n_inliers,
                              double *best_plane_out, int32_t *best_iter_out,
                              int64_t *counts_out, double *err_out, int nthreads)
{
    if (ransac_n < 3) return -3;
    if (n < ransac_n) return -4;
    if (prob <= 0.0 || prob > 1.0) return -5;
    double best_fit = 0.0, best_rmse = 0.0, best_pl[4] = { 0, 0, 0, 0 };
    int best_iter = -1;
    double break_iteration = 1.7976931348623157e308; /* numeric_limits<size_t>::max() upstream */
    int64_t iteration_count = 0;
    if (counts_out) for (int h = 0; h < H; ++h) { counts_out[h] = -1; if (err_out) err_out[h] = 0.0; }
    if (nthreads < 1) nthreads = 1;
#ifdef _OPENMP
    omp_set_num_threads(nthreads);
#endif
    int64_t *bc = (int64_t *)malloc(sizeof(int64_t) * (size_t)nthreads);
    double *be = (double *)malloc(sizeof(double) * (size_t)nthreads);
    double *bp = (double *)malloc(sizeof(double) * 4 * (size_t)nthreads);
    for (int h0 = 0; h0 < H; h0 += nthreads) {
        int hb = H - h0 < nthreads ? H - h0 : nthreads;
        if ((double)iteration_count > break_iteration) break;
#pragma omp parallel for schedule(static, 1)
        for (int b = 0; b < hb; ++b) {
            orc_plane_model(pts, samples + (int64_t)(h0 + b) * ransac_n, ransac_n, bp + 4 * b);
            const double *pl = bp + 4 * b;
            if (pl[0] == 0.0 && pl[1] == 0.0 && pl[2] == 0.0 && pl[3] == 0.0) { bc[b] = -2; be[b] = 0.0; }
            else orc_score_plane(pts, n, pl, thr, &bc[b], &be[b]);
        }
        for (int b = 0; b < hb; ++b) {
            int h = h0 + b;
            if ((double)iteration_count > break_iteration) continue;   /* skipped */
            if (bc[b] == -2) { if (counts_out) counts_out[h] = -2; continue; } /* degenerate sample */
            if (counts_out) { counts_out[h] = bc[b]; if (err_out) err_out[h] = be[b]; }
            double fit, rmse;
            if (bc[b] == 0) { fit = 0.0; rmse = 0.0; }
            else { fit = (double)bc[b] / (double)n; rmse = be[b] / sqrt((double)bc[b]); }
            if (fit > best_fit || (fit == best_fit && rmse < best_rmse)) {
                best_fit = fit; best_rmse = rmse; best_iter = h;
                memcpy(best_pl, bp + 4 * b, sizeof best_pl);
                if (fit < 1.0) {
                    double bi = log(1.0 - prob) / log(1.0 - pow(fit, (double)ransac_n));
                    /* upstream stores min(bi, H) in a size_t.  When fit^n underflows against 1.0
                     * the quotient is -inf, whose x86-64 conversion is 2^63: "never break". */
                    if (!(bi >= 0.0)) break_iteration = 1.7976931348623157e308;
                    else break_iteration = floor(bi < (double)H ? bi : (double)H);
                } else break_iteration = 0.0;
            }
            iteration_count++;
        }
    }
    free(bc); free(be); free(bp);
    /* final inliers of the (pre-refit) best model, ascending index */
    int64_t m = 0;
    int zero = best_pl[0] == 0.0 && best_pl[1] == 0.0 && best_pl[2] == 0.0 && best_pl[3] == 0.0;
    if (!zero)
        for (int64_t i = 0; i < n; ++i)
            if (plane_dist(best_pl, pts + 3 * i) < thr) inliers_out[m++] = (int32_t)i;
    *n_inliers = m;
    plane_from_points(pts, NULL, inliers_out, m, plane_out);
    if (best_plane_out) memcpy(best_plane_out, best_pl, sizeof best_pl);
    if (best_iter_out) *best_iter_out = best_iter;
    return 0;
}

ORC_API int orc_num_threads(void)
{
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}
