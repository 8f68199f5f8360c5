/*
 * oracle/geom_oracle.c -- TEST INFRASTRUCTURE ONLY (never linked into the product).
 *
 * Plain-C (double, no FMA: build with -ffp-contract=off) CPU restatement of the reference's own
 * fp64 geometry on the stereo hot path.  Every function cites the reference lines it follows
 * (paths relative to /root/reference/multiview/multiview_cpp/src/perceive/):
 *   geometry/projective/binocular-camera.cpp:682-785  (BinocularCamera::init -> KRR_inv_, C_)
 *   geometry/projective/binocular-camera.cpp:908-927  (solve3d)
 *   geometry/projective/triangulation.cpp:55-95       (intersect_rays, intersect_rays_2)
 *   geometry/vector-4.hpp:374-390,437-450             (Plane::side/image, plane_from_3_points)
 *   geometry/vector-3.hpp:64-75,127-137               (normalise, dot, cross)
 *   geometry/point-cloud.cpp:10-72                    (PointCloud::init)
 *   geometry/euclidean-transform.hpp:121-124, geometry/quaternion.hpp:126-156,205-213,274-277
 *   cost-functions/floor-histogram.hpp:90-96, cost-functions/floor-histogram.cpp:19-35,197-300
 *
 * PARITY UNPINNED: the reference ships no golden vector / known-answer test for these functions
 *  and its C++ cannot be compiled here (needs Eigen + OpenCV headers, both absent),
 * so this restatement is checked only against independent numpy formulas (tests/test_oracle_geom.py).
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

typedef struct { double krr_inv[2][9]; double c[2][3]; } orc_bcam;

typedef struct { double x, y, z; } v3;
static inline v3 V(double x, double y, double z) { v3 r = {x, y, z}; return r; }
static inline v3 add(v3 a, v3 b) { return V(a.x + b.x, a.y + b.y, a.z + b.z); }
static inline v3 sub(v3 a, v3 b) { return V(a.x - b.x, a.y - b.y, a.z - b.z); }
static inline v3 mul(v3 a, double s) { return V(a.x * s, a.y * s, a.z * s); }
static inline double dot(v3 a, v3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
/* vector-3.hpp:131-137 */
static inline v3 cross(v3 a, v3 v) { return V(a.y * v.z - a.z * v.y, a.z * v.x - a.x * v.z, a.x * v.y - a.y * v.x); }
/* vector-3.hpp:64-75 */
static inline v3 normalise(v3 a)
{
    double mag2 = a.x * a.x + a.y * a.y + a.z * a.z;
    if (fabs(mag2 - 1.0) > 1e-9) {
        double inv = 1.0 / sqrt(mag2);
        a.x *= inv; a.y *= inv; a.z *= inv;
    }
    return a;
}
static inline v3 matvec(const double* m, v3 p)
{
    return V(m[0] * p.x + m[1] * p.y + m[2] * p.z, m[3] * p.x + m[4] * p.y + m[5] * p.z,
             m[6] * p.x + m[7] * p.y + m[8] * p.z);
}
static void matmul3(const double* a, const double* b, double* c)
{
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j) c[i * 3 + j] = a[i * 3] * b[j] + a[i * 3 + 1] * b[3 + j] + a[i * 3 + 2] * b[6 + j];
}
static void inv3(const double* m, double* o)
{
    double c00 = m[4] * m[8] - m[5] * m[7], c01 = m[5] * m[6] - m[3] * m[8], c02 = m[3] * m[7] - m[4] * m[6];
    double det = m[0] * c00 + m[1] * c01 + m[2] * c02, id = 1.0 / det;
    o[0] = c00 * id; o[1] = (m[2] * m[7] - m[1] * m[8]) * id; o[2] = (m[1] * m[5] - m[2] * m[4]) * id;
    o[3] = c01 * id; o[4] = (m[0] * m[8] - m[2] * m[6]) * id; o[5] = (m[2] * m[3] - m[0] * m[5]) * id;
    o[6] = c02 * id; o[7] = (m[1] * m[6] - m[0] * m[7]) * id; o[8] = (m[0] * m[4] - m[1] * m[3]) * id;
}

/* binocular-camera.cpp:682-785 (use_calib_roi=false path). q = (x,y,z,w), K row-major 3x3. */
void orc_bcam_init(const double q[4], const double t[3], double baseline, const double K[9], orc_bcam* out)
{
    double x = q[0], y = q[1], z = q[2], w = q[3];
    double R[9] = {1 - 2 * (y * y + z * z), 2 * (x * y - z * w), 2 * (x * z + y * w),
                   2 * (x * y + z * w), 1 - 2 * (x * x + z * z), 2 * (y * z - x * w),
                   2 * (x * z - y * w), 2 * (y * z + x * w), 1 - 2 * (x * x + y * y)};
    double Rinv[9];
    inv3(R, Rinv);
    v3 tn = normalise(V(t[0], t[1], t[2]));
    v3 C0 = V(0, 0, 0), C1 = mul(matvec(Rinv, mul(tn, baseline)), -1.0);
    v3 z0 = V(R[6], R[7], R[8]);
    v3 X = normalise(sub(C1, C0)), Y = normalise(cross(z0, X)), Z = normalise(cross(X, Y));
    double Rn[9] = {X.x, X.y, X.z, Y.x, Y.y, Y.z, Z.x, Z.y, Z.z};
    double HR1[9], KR0[9], KR1[9], KRR1[9];
    matmul3(Rn, Rinv, HR1);
    matmul3(K, Rn, KR0);
    matmul3(K, HR1, KR1);
    matmul3(KR1, R, KRR1);
    inv3(KR0, out->krr_inv[0]);
    inv3(KRR1, out->krr_inv[1]);
    out->c[0][0] = C0.x; out->c[0][1] = C0.y; out->c[0][2] = C0.z;
    out->c[1][0] = C1.x; out->c[1][1] = C1.y; out->c[1][2] = C1.z;
}

/* triangulation.cpp:55-82 */
static v3 intersect_rays(v3 u0, v3 u1, v3 v0, v3 v1)
{
    v3 u = sub(u1, u0), v = sub(v1, v0), w = sub(u0, v0);
    double a = dot(u, u), b = dot(u, v), c = dot(v, v), d = dot(u, w), e = dot(v, w);
    double D = a * c - b * b, sc, tc;
    if (D < 1e-9) {
        sc = 0.0;
        tc = (b > c ? d / b : e / c);
    } else {
        sc = (b * e - c * d) / D;
        tc = (a * e - b * d) / D;
    }
    /* 0.5 * (u0 + (sc * u) + v0 + (tc * v)), evaluated left to right */
    return mul(add(add(add(u0, mul(u, sc)), v0), mul(v, tc)), 0.5);
}

/* binocular-camera.cpp:917-927 + triangulation.cpp:84-95 + vector-4.hpp:374-390,437-450 */
static v3 solve3d(const orc_bcam* b, double x0, double x1, double y)
{
    v3 o0 = matvec(b->krr_inv[0], V(x0, y, 1.0)), o1 = matvec(b->krr_inv[1], V(x1, y, 1.0));
    v3 C0 = V(b->c[0][0], b->c[0][1], b->c[0][2]), C1 = V(b->c[1][0], b->c[1][1], b->c[1][2]);
    v3 u1 = add(C0, o0), v1 = add(C1, o1);
    v3 c = mul(add(u1, v1), 0.5);
    v3 n = normalise(cross(sub(C0, C1), sub(C0, c)));
    double d = -1.0 * dot(C0, n);
    /* image(p) = p - n * side(p), side(p) = p.x*x + p.y*y + p.z*z + d */
    double su = u1.x * n.x + u1.y * n.y + u1.z * n.z + d;
    double sv = v1.x * n.x + v1.y * n.y + v1.z * n.z + d;
    return intersect_rays(C0, sub(u1, mul(n, su)), C1, sub(v1, mul(n, sv)));
}

void orc_solve3d(const orc_bcam* b, double x0, double x1, double y, double out[3])
{
    v3 r = solve3d(b, x0, x1, y);
    out[0] = r.x; out[1] = r.y; out[2] = r.z;
}

/* point-cloud.cpp:10-72.  Xs cap w*h*3, xy cap w*h*2, lookup w*h. Returns N. */
int orc_point_cloud(const orc_bcam* b, const float* disp, size_t step, int w, int h, int left_is_ref, double* Xs,
                    int32_t* xy, int32_t* lookup, double centroid[3])
{
    int n = 0;
    v3 C = V(0, 0, 0);
    for (int y = 0; y < h; ++y) {
        const float* row = (const float*)((const char*)disp + (size_t)y * step);
        for (int x = 0; x < w; ++x) {
            lookup[(size_t)y * w + x] = -1;
            float dx = row[x];
            if (!isfinite(dx) || dx == 0.0f) continue;
            v3 X = left_is_ref ? solve3d(b, (double)x, (double)x - (double)dx, (double)y)
                               : solve3d(b, (double)x + (double)dx, (double)x, (double)y);
            C = add(C, X);
            lookup[(size_t)y * w + x] = n;
            xy[2 * n] = x; xy[2 * n + 1] = y;
            Xs[3 * n] = X.x; Xs[3 * n + 1] = X.y; Xs[3 * n + 2] = X.z;
            ++n;
        }
    }
    centroid[0] = C.x / (double)n; centroid[1] = C.y / (double)n; centroid[2] = C.z / (double)n;
    return n;
}

typedef struct { double x, y, z, w; } quat;
/* quaternion.hpp:205-213 */
static inline quat qmul(quat a, quat b)
{
    quat c;
    c.x = a.y * b.z - a.z * b.y + a.x * b.w + a.w * b.x;
    c.y = a.z * b.x - a.x * b.z + a.y * b.w + a.w * b.y;
    c.z = a.x * b.y - a.y * b.x + a.z * b.w + a.w * b.z;
    c.w = a.w * b.w - a.x * b.x - a.y * b.y - a.z * b.z;
    return c;
}
/* euclidean-transform.hpp:121-124, quaternion.hpp:126-156,274-277.  et = {tx,ty,tz, qx,qy,qz,qw, scale} */
static v3 et_apply(const double et[8], v3 X)
{
    quat q = {et[3], et[4], et[5], et[6]};
    quat inv = {-q.x, -q.y, -q.z, q.w};
    double mag2 = inv.x * inv.x + inv.y * inv.y + inv.z * inv.z + inv.w * inv.w;
    if (fabs(mag2 - 1.0) > 1e-9) {
        double s = 1.0 / sqrt(mag2);
        inv.x *= s; inv.y *= s; inv.z *= s; inv.w *= s;
    }
    v3 xs = mul(X, et[7]);
    quat v = {xs.x, xs.y, xs.z, 0.0};
    quat r = qmul(qmul(q, v), inv);
    return V(r.x + et[0], r.y + et[1], r.z + et[2]);
}
void orc_et_apply(const double et[8], const double X[3], double Y[3])
{
    v3 r = et_apply(et, V(X[0], X[1], X[2]));
    Y[0] = r.x; Y[1] = r.y; Y[2] = r.z;
}

typedef struct { float hist_sz, min_z, max_z; int weighting; int use_still; } orc_hist_params;

/* floor-histogram.cpp:19-35.  aabb = {left, top, right, bottom} */
void orc_hist_dims(const double aabb[4], float hist_sz, int* w, int* h)
{
    double inv = 1.0 / (double)hist_sz;
    *h = (int)(ceil((aabb[3] - aabb[1]) * inv) + 1e-9);
    *w = (int)(ceil((aabb[2] - aabb[0]) * inv) + 1e-9);
}

/* floor-histogram.cpp:197-286: per-camera histogram and (cells != NULL) the super-pixel table slic_to_hist_xy
 * (:222-225, :269-286; cells[l] = {x, y, count} as SparseHistCell, geometry/sparse-hist-cell.hpp:8-17).
 * cam_center = translations[cam] (floor_hist.cpp:122). labels may be NULL (lw=lh=0). */
typedef struct { int32_t x, y; float count; } orc_sparse_hist_cell;
void orc_floor_hist_camera_cells(const orc_hist_params* p, const double et[8], const double aabb[4], const double cam_center[3],
                                 const double* Xs, const int32_t* xy, int n, const int32_t* labels, int lw, int lh,
                                 const double* still_score, int n_labels, float* cam_hist, int hw, int hh,
                                 float super_pixel_min_z, orc_sparse_hist_cell* cells)
{
    /* vector<Vector4f> spixel_Cs(n_slic_labels, {0,0,0,0}), floor-histogram.cpp:222-225 */
    float* sC = cells ? (float*)calloc((size_t)(n_labels > 0 ? n_labels : 1) * 4, sizeof(float)) : NULL;
    memset(cam_hist, 0, sizeof(float) * (size_t)hw * hh);
    double hist_sz = (double)p->hist_sz;
    v3 Cc = V(cam_center[0], cam_center[1], cam_center[2]);
    for (int i = 0; i < n; ++i) {
        v3 X = V(Xs[3 * i], Xs[3 * i + 1], Xs[3 * i + 2]);
        v3 Y = et_apply(et, X);
        double zx = (Y.x - aabb[0]) / hist_sz + 1e-9, zy = (Y.y - aabb[1]) / hist_sz + 1e-9;
        int bx = (int)round(zx), by = (int)round(zy);
        if (Y.z < (double)p->min_z || Y.z > (double)p->max_z) continue;
        if (bx < 0 || bx >= hw || by < 0 || by >= hh) continue;
        v3 dv = sub(X, Cc);
        double wgt = p->weighting == 0 ? 1.0 : p->weighting == 1 ? sqrt(dot(dv, dv)) : dot(dv, dv);
        int px = xy[2 * i], py = xy[2 * i + 1];
        int label = (labels && px >= 0 && px < lw && py >= 0 && py < lh) ? labels[(size_t)py * lw + px] : -1;
        if (p->use_still && label >= 0 && label < n_labels) wgt *= 1.0 + 10.0 * still_score[label];
        const float weight = (float)wgt;
        cam_hist[(size_t)by * hw + bx] += weight;
        /* :269-273  spixel_Cs[label] += Vector4f(to_vec3f(Y) * float(weight), spixel_Cs[label].w + weight)
         * (labels >= n_labels are skipped: the reference would index past its vector) */
        if (sC && label >= 0 && label < n_labels && Y.z >= (double)super_pixel_min_z) {
            float* c = sC + 4 * (size_t)label;
            const float ax = (float)Y.x * weight, ay = (float)Y.y * weight, az = (float)Y.z * weight;
            const float aw = c[3] + weight;
            c[0] += ax; c[1] += ay; c[2] += az; c[3] += aw;
        }
    }
    if (cells) { /* :276-286 */
        for (int l = 0; l < n_labels; ++l) {
            const float* c = sC + 4 * (size_t)l;
            cells[l].x = -1; cells[l].y = -1; cells[l].count = 0.0f;
            if (c[3] == 0.0f) continue;
            const double cx = (double)c[0] / (double)c[3], cy = (double)c[1] / (double)c[3];
            const double zx = (cx - aabb[0]) / hist_sz + 1e-9, zy = (cy - aabb[1]) / hist_sz + 1e-9;
            const double rx = round(zx), ry = round(zy);
            cells[l].x = rx >= -2147483648.0 && rx <= 2147483647.0 ? (int32_t)rx : -1;   /* int(NaN / huge) is undefined there */
            cells[l].y = ry >= -2147483648.0 && ry <= 2147483647.0 ? (int32_t)ry : -1;
            cells[l].count = roundf(c[3]);
        }
        free(sC);
    }
}

void orc_floor_hist_camera(const orc_hist_params* p, const double et[8], const double aabb[4], const double cam_center[3],
                           const double* Xs, const int32_t* xy, int n, const int32_t* labels, int lw, int lh,
                           const double* still_score, int n_labels, float* cam_hist, int hw, int hh)
{
    orc_floor_hist_camera_cells(p, et, aabb, cam_center, Xs, xy, n, labels, lw, lh, still_score, n_labels, cam_hist, hw, hh, 0.25f, NULL);
}

/* floor-histogram.cpp:296-300: hist = sum over cameras in camera order (float adds). */
void orc_hist_sum(const float* cam_hists, int n_cams, int hw, int hh, float* hist)
{
    memset(hist, 0, sizeof(float) * (size_t)hw * hh);
    for (int c = 0; c < n_cams; ++c)
        for (size_t i = 0; i < (size_t)hw * hh; ++i) hist[i] += cam_hists[(size_t)c * hw * hh + i];
}

/* pipeline/nodes/disp_map_update.cpp:98-113: fill from the reference disparity, in place. */
void orc_disp_merge_ref(float* disp, const float* ref, const int32_t* labels, const uint8_t* still_cutoff, int n_labels, int w,
                        int h)
{
    for (size_t i = 0; i < (size_t)w * h; ++i) {
        int label = labels[i];
        if (!isfinite(ref[i])) continue;
        if (label < 0 || label >= n_labels) continue; /* >= n_labels: the reference would read past its vector */
        if (!still_cutoff[label] || !isfinite(disp[i]) || disp[i] <= 0.0f) disp[i] = ref[i];
    }
}

static int orc_reflect101(int i, int n)
{
    if (n == 1) return 0;
    while (i < 0 || i >= n) {
        if (i < 0) i = -i;
        if (i >= n) i = 2 * (n - 1) - i;
    }
    return i;
}
/* floor-histogram.cpp:411-434: GaussianBlur (CV_32F, ksize ceil(sigma)*4|1, BORDER_DEFAULT) + clip, in place.
 * Separable symmetric form in float without FMA; OpenCV's float path agrees to 1-2 ulp (SIMD-dispatch dependent). */
int orc_hist_post(float* hist, int w, int h, float sigma, float minc, float maxc)
{
    size_t n = (size_t)w * h;
    float* tmp = (float*)malloc(sizeof(float) * n);
    if (!tmp) return -2;
    if (isfinite(sigma) && sigma > 0.0f) {
        int ks = (int)(ceilf(sigma) * 4.0f);
        if (ks % 2 == 0) ks += 1;
        int r = ks / 2;
        double kd[65], sum = 0, s2 = -0.5 / ((double)sigma * sigma);
        float k[33];
        if (ks > 33) { free(tmp); return -1; }
        for (int i = 0; i < ks; ++i) { double x = i - (ks - 1) * 0.5; kd[i] = exp(s2 * x * x); sum += kd[i]; }
        for (int i = 0; i <= r; ++i) k[i] = (float)(kd[r + i] / sum);
        for (int y = 0; y < h; ++y)
            for (int x = 0; x < w; ++x) {
                float s = hist[(size_t)y * w + x] * k[0];
                for (int i = 1; i <= r; ++i)
                    s = s + (hist[(size_t)y * w + orc_reflect101(x - i, w)] + hist[(size_t)y * w + orc_reflect101(x + i, w)]) * k[i];
                tmp[(size_t)y * w + x] = s;
            }
        for (int y = 0; y < h; ++y)
            for (int x = 0; x < w; ++x) {
                float s = tmp[(size_t)y * w + x] * k[0];
                for (int i = 1; i <= r; ++i)
                    s = s + (tmp[(size_t)orc_reflect101(y - i, h) * w + x] + tmp[(size_t)orc_reflect101(y + i, h) * w + x]) * k[i];
                hist[(size_t)y * w + x] = s;
            }
    }
    for (size_t i = 0; i < n; ++i) {
        float v = hist[i];
        if (v < minc) v = 0.0f;
        if (v > maxc) v = maxc;
        hist[i] = v;
    }
    free(tmp);
    return 0;
}

/* ---- localization front half: cost-functions/localization/localization-data.cpp ---------------------
 * make_motion_histogram :201-233, make_radius_filtered_hist :147-187, make_bg_histogram_ :237-253,
 * make_fg_3d_histogram_ :264-285, make_fg_histogram :298-315, as called by LocalizationData::calculate :845-850
 * (pose-less branch :855-858: fg1 = fg0).  phi_function: utils/math.hpp:374-382 (0.5 * erfc(-z / sqrt 2), double). */
typedef struct {
    double n_deviations, person_diameter, background_median, background_stddev, track_median, track_stddev, prob_bg_max;
} orc_loc_params;

static double orc_phi(double z) { return 0.5 * erfc((-1.0 / sqrt(2.0)) * z); }

void orc_localization_hists(const orc_loc_params* p, const float* hist, int w, int h, double hist_sz, float* motion,
                            float* rhist, float* bg, float* fg0, float* fg)
{
    for (int i = 0; i < w * h; ++i) motion[i] = ((double)hist[i] > p->n_deviations) ? hist[i] : 0.0f;
    const double in_radius = 0.5 * p->person_diameter;
    const float radius_inv = 1.0f / (0.5f * (float)in_radius);
    const int radius = (int)ceil(in_radius / hist_sz);
    for (int y0 = 0; y0 < h; ++y0)
        for (int x0 = 0; x0 < w; ++x0) {
            double sum = 0.0;
            long n = 0;
            for (int dy = -radius; dy <= radius; ++dy)
                for (int dx = -radius; dx <= radius; ++dx)
                    if (dx * dx + dy * dy <= radius * radius) {
                        const int x = x0 + dx, y = y0 + dy;
                        if (x < 0 || y < 0 || x >= w || y >= h) continue;
                        const double cell = (double)motion[y * w + x];
                        const double dist = sqrt((double)(dx * dx + dy * dy)) * hist_sz;
                        const double wght = 1.0 - dist * (double)radius_inv;
                        if (wght >= 0.0) { sum += (double)(float)(wght * wght * cell); ++n; }
                    }
            rhist[y0 * w + x0] = (float)(n == 0 ? 0.0 : sum / (double)n);
        }
    for (int i = 0; i < w * h; ++i) {
        const double zb = ((double)rhist[i] - p->background_median) / p->background_stddev;
        const double pb = orc_phi(-zb);
        bg[i] = (float)(p->prob_bg_max < pb ? p->prob_bg_max : pb);
        const double zf = ((double)rhist[i] - p->track_median) / p->track_stddev;
        fg0[i] = (float)orc_phi(zf);
        const float a = fg0[i] > 0.10f ? fg0[i] : 0.10f;
        fg[i] = a * a;
    }
}
