/*
 * oracle/sgbm_oracle.c -- TEST INFRASTRUCTURE ONLY (never linked into the product).
 *
 * Plain-C CPU restatement of the arithmetic the reference obtains from its OpenCV
 * dependency on the disparity path:
 *   multiview_cpp/src/perceive/cost-functions/disparity/strategies/stereosgbm_disp.cpp:121-125
 *     (cv::StereoSGBM::create + compute) and :141-150 (float conversion, max_disparity = 150).
 * OpenCV itself is a third-party dependency that is NOT under /root/reference (pinned there at
 * 4.4.0, build/install-files/build-scripts/build-opencv.sh:29); this file restates its published
 * StereoSGBM algorithm (Hirschmuller SGM with Birchfield-Tomasi cost) as probed in SURVEY.md
 * Appendix A.1-A.4.  It is PINNED against the same API importable in this image as Python cv2
 * 4.13.0 (tests/test_oracle_vs_cv2.py, the tests/golden fixtures made by tests/golden/make_golden.py).
 * Version skew 4.4.0 -> 4.13.0 is recorded in DESIGN.md.
 *
 * Layouts: images are 8-bit row-major with a byte step; cost volumes are int16 [H][W1][D] with
 * D fastest; disparity is int16 Q4 [H][W].
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

typedef struct {
    int block, disp12_max_diff, min_disp, num_disp, speckle_win, speckle_range, prefilter_cap,
        uniqueness, mode, p1, p2;
} orc_sgbm_params;

enum { ORC_MODE_SGBM = 0, ORC_MODE_HH = 1, ORC_MODE_SGBM_3WAY = 2, ORC_MODE_HH4 = 3 };
#define ORC_MAX_COST 32767
#define ORC_DISP_SHIFT 4
#define ORC_DISP_SCALE 16

static inline int imin(int a, int b) { return a < b ? a : b; }
static inline int imax(int a, int b) { return a > b ? a : b; }
static inline int iclamp(int v, int lo, int hi) { return v < lo ? lo : (v > hi ? hi : v); }
static inline int16_t sat16(int v) { return (int16_t)(v < -32768 ? -32768 : (v > 32767 ? 32767 : v)); }

/* Valid-column geometry. Returns 0 if the image is too narrow. */
int orc_sgbm_dims(const orc_sgbm_params* p, int W, int* minX1, int* W1)
{
    int minD = p->min_disp, maxD = minD + p->num_disp;
    int mn = imax(maxD, 0), mx = W + imin(minD, 0);
    *minX1 = mn;
    *W1 = mx - mn;
    return *W1 > 0;
}

/* A.1: x-Sobel prefilter (clipped) + raw channel for one row y of image I. */
static void prefilter_row(const uint8_t* I, int W, int H, size_t step, int y, int ftzero, int* p,
                          int* q)
{
    const uint8_t* r = I + (size_t)y * step;
    const uint8_t* n = y > 0 ? r - step : r;
    const uint8_t* s = y < H - 1 ? r + step : r;
    for (int x = 1; x < W - 1; ++x) {
        int v = 2 * (r[x + 1] - r[x - 1]) + (n[x + 1] - n[x - 1]) + (s[x + 1] - s[x - 1]);
        p[x] = iclamp(v, -ftzero, ftzero) + ftzero;
        q[x] = r[x];
    }
    p[0] = p[W - 1] = ftzero;
    q[0] = q[W - 1] = ftzero;
}

static void lohi(const int* v, int W, int* lo, int* hi)
{
    for (int x = 0; x < W; ++x) {
        int c = v[x];
        int l = x > 0 ? (c + v[x - 1]) / 2 : c;
        int r = x < W - 1 ? (c + v[x + 1]) / 2 : c;
        lo[x] = imin(c, imin(l, r));
        hi[x] = imax(c, imax(l, r));
    }
}

/* A.1: per-pixel Birchfield-Tomasi cost for row y -> pix[W1][D] (int). */
static void pixel_cost_row(const orc_sgbm_params* p, const uint8_t* L, const uint8_t* R, int W, int H,
                           size_t step, int y, int* pix, int* scratch)
{
    int minD = p->min_disp, D = p->num_disp, minX1, W1;
    orc_sgbm_dims(p, W, &minX1, &W1);
    int ftzero = imax(p->prefilter_cap, 15) | 1;
    int* a[2] = {scratch, scratch + W};
    int* b[2] = {scratch + 2 * W, scratch + 3 * W};
    int *loa = scratch + 4 * W, *hia = scratch + 5 * W, *lob = scratch + 6 * W, *hib = scratch + 7 * W;
    prefilter_row(L, W, H, step, y, ftzero, a[0], a[1]);
    prefilter_row(R, W, H, step, y, ftzero, b[0], b[1]);
    memset(pix, 0, sizeof(int) * (size_t)W1 * D);
    for (int ch = 0; ch < 2; ++ch) {
        int shift = ch ? 2 : 0;
        lohi(a[ch], W, loa, hia);
        lohi(b[ch], W, lob, hib);
        for (int x = minX1; x < minX1 + W1; ++x) {
            int u = a[ch][x], u0 = loa[x], u1 = hia[x];
            for (int d = minD; d < minD + D; ++d) {
                int x2 = x - d;
                int v = b[ch][x2], v0 = lob[x2], v1 = hib[x2];
                int c0 = imax(0, imax(u - v1, v0 - u));
                int c1 = imax(0, imax(v - u1, u0 - v));
                pix[(size_t)(x - minX1) * D + (d - minD)] += imin(c0, c1) >> shift;
            }
        }
    }
}

/* A.1: block-summed matching cost C[H][W1][D] (wrap16). MODE_HH4 leaves rows y >= H-SH2 zero (A.3). */
int orc_sgbm_cost_volume(const orc_sgbm_params* p, const uint8_t* L, const uint8_t* R, int W, int H,
                         size_t step, int16_t* C)
{
    int D = p->num_disp, minX1, W1;
    if (!orc_sgbm_dims(p, W, &minX1, &W1)) return -1;
    int SW2 = p->block / 2, SH2 = p->block / 2;
    size_t row = (size_t)W1 * D;
    int* pix = (int*)malloc(sizeof(int) * row);
    int* scratch = (int*)malloc(sizeof(int) * 8 * (size_t)W);
    int* hsum = (int*)malloc(sizeof(int) * row * (size_t)H);
    if (!pix || !scratch || !hsum) return -2;
    for (int y = 0; y < H; ++y) {
        pixel_cost_row(p, L, R, W, H, step, y, pix, scratch);
        int* hs = hsum + (size_t)y * row;
        for (int x = 0; x < W1; ++x)
            for (int d = 0; d < D; ++d) {
                int s = 0;
                for (int dx = -SW2; dx <= SW2; ++dx) s += pix[(size_t)iclamp(x + dx, 0, W1 - 1) * D + d];
                hs[(size_t)x * D + d] = s;
            }
    }
    for (int y = 0; y < H; ++y) {
        int16_t* Cy = C + (size_t)y * row;
        if (p->mode == ORC_MODE_HH4 && y >= H - SH2) {
            memset(Cy, 0, sizeof(int16_t) * row);
            continue;
        }
        for (size_t i = 0; i < row; ++i) {
            int s = 0;
            for (int dy = -SH2; dy <= SH2; ++dy) s += hsum[(size_t)iclamp(y + dy, 0, H - 1) * row + i];
            Cy[i] = (int16_t)s;
        }
    }
    free(pix);
    free(scratch);
    free(hsum);
    return 0;
}

/* One SGM path step (A.2). Lp has sentinels at [-1] and [D]; returns min over d of the new L. */
static int path_step(const int16_t* Cp, const int16_t* Lp, int minLp, int D, int P1, int P2, int16_t* Lo)
{
    int delta = minLp + P2, mn = ORC_MAX_COST;
    for (int d = 0; d < D; ++d) {
        int v = imin(imin((int)Lp[d], Lp[d - 1] + P1), imin(Lp[d + 1] + P1, delta));
#ifdef ORC_SGBM_SUBTRACT_P2
        int L = Cp[d] + v - delta;
#else
        int L = Cp[d] + v - minLp;
#endif
        Lo[d] = (int16_t)L;
        mn = imin(mn, (int)(int16_t)L);
    }
    return mn;
}

/*
 * A.2/A.3: aggregate C -> S over the directions selected by `mode`.
 * S1 (optional) receives the volume after the first pass (forward directions only).
 * Directions are accumulated per pass with sat16, matching the per-pass saturation order.
 */
int orc_sgbm_aggregate(const orc_sgbm_params* p, const int16_t* C, int W1, int H, int16_t* S, int16_t* S1)
{
    int D = p->num_disp;
    int P1 = p->p1 > 0 ? p->p1 : 2;
    int P2 = imax(p->p2 > 0 ? p->p2 : 5, P1 + 1);
    size_t row = (size_t)W1 * D;
    int Dp = D + 2;
    /* Lr rows with one border pixel on each side: index (x+1), 4 directions, 2 rows */
    size_t lrsz = (size_t)(W1 + 2) * Dp;
    int16_t* Lr = (int16_t*)calloc(lrsz * 4 * 2, sizeof(int16_t));
    int* minLr = (int*)calloc((size_t)(W1 + 2) * 4 * 2, sizeof(int));
    if (!Lr || !minLr) return -2;
    memset(S, 0, sizeof(int16_t) * row * (size_t)H);
    int npasses = (p->mode == ORC_MODE_SGBM) ? 1 : 2;
    int hh4 = p->mode == ORC_MODE_HH4;
#define LR(buf, dir, x) (Lr + ((size_t)((buf)*4 + (dir)) * (W1 + 2) + ((x) + 1)) * Dp + 1)
#define MINLR(buf, dir, x) (minLr[(size_t)((buf)*4 + (dir)) * (W1 + 2) + ((x) + 1)])
    for (int pass = 1; pass <= npasses; ++pass) {
        int y1 = pass == 1 ? 0 : H - 1, y2 = pass == 1 ? H : -1, dy = pass == 1 ? 1 : -1;
        int x1 = pass == 1 ? 0 : W1 - 1, x2 = pass == 1 ? W1 : -1, dx = pass == 1 ? 1 : -1;
        /* clear both Lr buffers (zeros, sentinels rewritten below) */
        memset(Lr, 0, sizeof(int16_t) * lrsz * 4 * 2);
        memset(minLr, 0, sizeof(int) * (size_t)(W1 + 2) * 4 * 2);
        int cur = 0;
        for (int y = y1; y != y2; y += dy) {
            const int16_t* Cy = C + (size_t)y * row;
            int16_t* Sy = S + (size_t)y * row;
            int prev = 1 - cur;
            /* border pixels of the current row buffer are zero */
            for (int dir = 0; dir < 4; ++dir) {
                memset(LR(cur, dir, -1) - 1, 0, sizeof(int16_t) * Dp);
                memset(LR(cur, dir, W1) - 1, 0, sizeof(int16_t) * Dp);
                MINLR(cur, dir, -1) = 0;
                MINLR(cur, dir, W1) = 0;
            }
            for (int x = x1; x != x2; x += dx) {
                const int16_t* Cp = Cy + (size_t)x * D;
                int16_t* Sp = Sy + (size_t)x * D;
                /* predecessor pixels: dir0 (x-dx, y); dir1 (x-1, y-dy); dir2 (x, y-dy); dir3 (x+1, y-dy) */
                int16_t* Lp[4] = {LR(cur, 0, x - dx), LR(prev, 1, x - 1), LR(prev, 2, x), LR(prev, 3, x + 1)};
                int mp[4] = {MINLR(cur, 0, x - dx), MINLR(prev, 1, x - 1), MINLR(prev, 2, x), MINLR(prev, 3, x + 1)};
                int16_t Ln[4][512 + 2];
                int use[4] = {1, !hh4, 1, !hh4};
                for (int dir = 0; dir < 4; ++dir) {
                    if (!use[dir]) continue;
                    Lp[dir][-1] = Lp[dir][D] = ORC_MAX_COST;
                    int m = path_step(Cp, Lp[dir], mp[dir], D, P1, P2, Ln[dir]);
                    memcpy(LR(cur, dir, x), Ln[dir], sizeof(int16_t) * D);
                    MINLR(cur, dir, x) = m;
                }
                for (int d = 0; d < D; ++d) {
                    int s = Sp[d];
                    for (int dir = 0; dir < 4; ++dir)
                        if (use[dir]) s += Ln[dir][d];
                    Sp[d] = sat16(s);
                }
            }
            if (p->mode == ORC_MODE_SGBM) {
                /* fifth direction from (x+1, y), x descending, S = sat16(S + L) (A.2) */
                int16_t Lprev[512 + 2], Lcur[512 + 2];
                memset(Lprev, 0, sizeof(Lprev));
                int mprev = 0;
                for (int x = W1 - 1; x >= 0; --x) {
                    const int16_t* Cp = Cy + (size_t)x * D;
                    int16_t* Sp = Sy + (size_t)x * D;
                    Lprev[0] = Lprev[D + 1] = ORC_MAX_COST;
                    int m = path_step(Cp, Lprev + 1, mprev, D, P1, P2, Lcur + 1);
                    for (int d = 0; d < D; ++d) Sp[d] = sat16(Sp[d] + Lcur[d + 1]);
                    memcpy(Lprev, Lcur, sizeof(Lcur));
                    mprev = m;
                }
            }
            cur = prev;
        }
        if (pass == 1 && S1) memcpy(S1, S, sizeof(int16_t) * row * (size_t)H);
    }
#undef LR
#undef MINLR
    free(Lr);
    free(minLr);
    return 0;
}

/* A.2: WTA + uniqueness + subpixel + right-view consistency for all rows -> disp Q4 [H][W] */
int orc_sgbm_wta(const orc_sgbm_params* p, const int16_t* S, int W, int H, int16_t* disp, size_t dstep)
{
    int D = p->num_disp, minD = p->min_disp, minX1, W1;
    if (!orc_sgbm_dims(p, W, &minX1, &W1)) return -1;
    int ur = p->uniqueness >= 0 ? p->uniqueness : 10;
    int d12 = p->disp12_max_diff > 0 ? p->disp12_max_diff : 1;
    int INV = (minD - 1) * ORC_DISP_SCALE;
    size_t row = (size_t)W1 * D;
    int* disp2 = (int*)malloc(sizeof(int) * (size_t)W);
    int* disp2cost = (int*)malloc(sizeof(int) * (size_t)W);
    for (int y = 0; y < H; ++y) {
        int16_t* d1 = (int16_t*)((char*)disp + (size_t)y * dstep);
        for (int x = 0; x < W; ++x) {
            d1[x] = (int16_t)INV;
            disp2[x] = INV;
            disp2cost[x] = ORC_MAX_COST;
        }
        for (int x = W1 - 1; x >= 0; --x) {
            const int16_t* Sp = S + (size_t)y * row + (size_t)x * D;
            int minS = ORC_MAX_COST, best = -1, d;
            for (d = 0; d < D; ++d)
                if (Sp[d] < minS) {
                    minS = Sp[d];
                    best = d;
                }
            for (d = 0; d < D; ++d)
                if (Sp[d] * (100 - ur) < minS * 100 && abs(best - d) > 1) break;
            if (d < D) continue;
            d = best;
            int x2 = x + minX1 - d - minD;
            if (disp2cost[x2] > minS) {
                disp2cost[x2] = minS;
                disp2[x2] = d + minD;
            }
            if (0 < d && d < D - 1) {
                int den = imax(Sp[d - 1] + Sp[d + 1] - 2 * Sp[d], 1);
                d = d * ORC_DISP_SCALE + ((Sp[d - 1] - Sp[d + 1]) * ORC_DISP_SCALE + den) / (den * 2);
            } else
                d *= ORC_DISP_SCALE;
            d1[x + minX1] = (int16_t)(d + minD * ORC_DISP_SCALE);
        }
        for (int x = minX1; x < minX1 + W1; ++x) {
            int v = d1[x];
            if (v == INV) continue;
            int dl = v >> ORC_DISP_SHIFT, dh = (v + ORC_DISP_SCALE - 1) >> ORC_DISP_SHIFT;
            int xl = x - dl, xh = x - dh;
            if (0 <= xl && xl < W && disp2[xl] >= minD && abs(disp2[xl] - dl) > d12 && 0 <= xh && xh < W &&
                disp2[xh] >= minD && abs(disp2[xh] - dh) > d12)
                d1[x] = (int16_t)INV;
        }
    }
    free(disp2);
    free(disp2cost);
    return 0;
}

static inline void sort2(int* a, int* b)
{
    if (*a > *b) {
        int t = *a;
        *a = *b;
        *b = t;
    }
}

/* A.4: cv::medianBlur(ksize 3) on CV_16S, replicate border. src != dst. */
void orc_median3_s16(const int16_t* src, size_t sstep, int W, int H, int16_t* dst, size_t dstep)
{
    for (int y = 0; y < H; ++y)
        for (int x = 0; x < W; ++x) {
            int v[9], k = 0;
            for (int dy = -1; dy <= 1; ++dy)
                for (int dx = -1; dx <= 1; ++dx) {
                    int yy = iclamp(y + dy, 0, H - 1), xx = iclamp(x + dx, 0, W - 1);
                    v[k++] = ((const int16_t*)((const char*)src + (size_t)yy * sstep))[xx];
                }
            for (int i = 0; i < 9; ++i)
                for (int j = i + 1; j < 9; ++j) sort2(&v[i], &v[j]);
            ((int16_t*)((char*)dst + (size_t)y * dstep))[x] = (int16_t)v[4];
        }
}

/* A.4: cv::filterSpeckles: 4-connected components over pixels != newVal, edge iff |a-b| <= maxDiff;
 * components with size <= maxSpeckleSize are set to newVal. In place. */
int orc_filter_speckles_s16(int16_t* img, size_t step, int W, int H, int newVal, int maxSpeckleSize, int maxDiff)
{
    size_t n = (size_t)W * H;
    int* label = (int*)calloc(n, sizeof(int));
    int* stack = (int*)malloc(sizeof(int) * n);
    uint8_t* small = (uint8_t*)malloc(n + 1);
    if (!label || !stack || !small) return -2;
#define PX(i) (((int16_t*)((char*)img + (size_t)((i) / W) * step))[(i) % W])
    int cur = 0;
    for (int i = 0; i < (int)n; ++i) {
        if (PX(i) == newVal) continue;
        if (label[i]) {
            if (small[label[i]]) PX(i) = (int16_t)newVal;
            continue;
        }
        int sp = 0, count = 0;
        ++cur;
        stack[sp++] = i;
        label[i] = cur;
        while (sp) {
            int q = stack[--sp];
            ++count;
            int qy = q / W, qx = q % W, v = PX(q);
            int nb[4] = {qx < W - 1 ? q + 1 : -1, qx > 0 ? q - 1 : -1, qy < H - 1 ? q + W : -1, qy > 0 ? q - W : -1};
            for (int k = 0; k < 4; ++k) {
                int r = nb[k];
                if (r < 0 || label[r]) continue;
                int rv = PX(r);
                if (rv == newVal || abs(rv - v) > maxDiff) continue;
                label[r] = cur;
                stack[sp++] = r;
            }
        }
        small[cur] = count <= maxSpeckleSize;
        if (small[cur]) PX(i) = (int16_t)newVal;
    }
#undef PX
    free(label);
    free(stack);
    free(small);
    return 0;
}

/* Whole StereoSGBM::compute restatement. Optional taps: C, S1 (after pass 1), S, raw (pre-median). */
int orc_sgbm_compute(const orc_sgbm_params* p, const uint8_t* L, const uint8_t* R, int W, int H, size_t step,
                     int16_t* disp, int16_t* C_out, int16_t* S1_out, int16_t* S_out, int16_t* raw_out)
{
    int minX1, W1, D = p->num_disp;
    if (D <= 0 || D % 16 || D > 512 || !orc_sgbm_dims(p, W, &minX1, &W1)) return -1;
    size_t vol = (size_t)H * W1 * D;
    int16_t* C = C_out ? C_out : (int16_t*)malloc(sizeof(int16_t) * vol);
    int16_t* S = S_out ? S_out : (int16_t*)malloc(sizeof(int16_t) * vol);
    int16_t* raw = raw_out ? raw_out : (int16_t*)malloc(sizeof(int16_t) * (size_t)W * H);
    if (!C || !S || !raw) return -2;
    int rc = orc_sgbm_cost_volume(p, L, R, W, H, step, C);
    if (!rc) rc = orc_sgbm_aggregate(p, C, W1, H, S, S1_out);
    if (!rc) rc = orc_sgbm_wta(p, S, W, H, raw, sizeof(int16_t) * (size_t)W);
    if (!rc) {
        orc_median3_s16(raw, sizeof(int16_t) * (size_t)W, W, H, disp, sizeof(int16_t) * (size_t)W);
        if (p->speckle_win > 0)
            rc = orc_filter_speckles_s16(disp, sizeof(int16_t) * (size_t)W, W, H, (p->min_disp - 1) * ORC_DISP_SCALE,
                                         p->speckle_win, ORC_DISP_SCALE * p->speckle_range);
    }
    if (!C_out) free(C);
    if (!S_out) free(S);
    if (!raw_out) free(raw);
    return rc;
}

/* stereosgbm_disp.cpp:141-150: f = |d16| / 16, f > 150 -> 0 */
void orc_disp_to_float(const int16_t* disp, int W, int H, float* out)
{
    for (size_t i = 0; i < (size_t)W * H; ++i) {
        float v = (float)(disp[i] < 0 ? -(int)disp[i] : (int)disp[i]) / 16.0f;
        out[i] = v > 150.0f ? 0.0f : v;
    }
}
