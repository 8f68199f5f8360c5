/*
 * oracle/image_oracle.c -- TEST INFRASTRUCTURE ONLY (never linked into the product).
 *
 * Plain-C CPU restatement of the OpenCV image arithmetic the reference calls before StereoSGBM:
 *   - cv::remap(..., INTER_CUBIC, BORDER_CONSTANT, white)
 *       multiview_cpp/src/perceive/pipeline/nodes/calc_rectified.cpp:109-115
 *   - cv::cvtColor(BGR2GRAY) and cv::GaussianBlur(31x31, sigma) applied twice
 *       multiview_cpp/src/perceive/cost-functions/disparity/disparity.cpp:106-155
 * OpenCV is a third-party dependency absent from /root/reference (pinned 4.4.0); the algorithm is
 * restated from SURVEY.md Appendix A.5/A.6 and PINNED against cv2 4.13.0 in this image
 * (tests/test_oracle_vs_cv2.py and the tests/golden fixtures).
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

static inline int iclamp(int v, int lo, int hi) { return v < lo ? lo : (v > hi ? hi : v); }
static inline int sat16i(int v) { return v < -32768 ? -32768 : (v > 32767 ? 32767 : v); }

/* A.5: the 32x32 table of 4x4 Q15 bicubic weights, tab[(fy*32+fx)*16 + k1*4 + k2]. */
void orc_cubic_tab(int16_t* tab)
{
    float c[32][4];
    const float A = -0.75f;
    for (int i = 0; i < 32; ++i) {
        float x = (float)i * (1.0f / 32.0f);
        c[i][0] = ((A * (x + 1) - 5 * A) * (x + 1) + 8 * A) * (x + 1) - 4 * A;
        c[i][1] = ((A + 2) * x - (A + 3)) * x * x + 1;
        c[i][2] = ((A + 2) * (1 - x) - (A + 3)) * (1 - x) * (1 - x) + 1;
        c[i][3] = 1.f - c[i][0] - c[i][1] - c[i][2];
    }
    for (int fy = 0; fy < 32; ++fy)
        for (int fx = 0; fx < 32; ++fx) {
            int16_t* w = tab + (size_t)(fy * 32 + fx) * 16;
            int isum = 0;
            for (int k1 = 0; k1 < 4; ++k1)
                for (int k2 = 0; k2 < 4; ++k2) {
                    float v = c[fy][k1] * c[fx][k2];
                    int q = sat16i((int)lrintf(v * 32768.0f));
                    w[k1 * 4 + k2] = (int16_t)q;
                    isum += q;
                }
            if (isum != 32768) {
                int diff = isum - 32768, M = 2 * 4 + 2, m = 2 * 4 + 2;
                for (int k1 = 2; k1 < 4; ++k1)
                    for (int k2 = 2; k2 < 4; ++k2) {
                        int i = k1 * 4 + k2;
                        if (w[i] < w[m])
                            m = i;
                        else if (w[i] > w[M])
                            M = i;
                    }
                if (diff < 0)
                    w[M] = (int16_t)(w[M] - diff);
                else
                    w[m] = (int16_t)(w[m] - diff);
            }
        }
}

/* A.5: cv::remap INTER_CUBIC / CV_32FC1 maps / 8UC3 / BORDER_CONSTANT(border). */
void orc_remap_cubic_bgr8(const uint8_t* src, int sw, int sh, size_t sstep, const float* mapx, const float* mapy,
                          size_t mstep, int w, int h, const uint8_t border[3], uint8_t* dst, size_t dstep)
{
    int16_t* tab = (int16_t*)malloc(sizeof(int16_t) * 1024 * 16);
    orc_cubic_tab(tab);
    for (int y = 0; y < h; ++y) {
        const float* mx = (const float*)((const char*)mapx + (size_t)y * mstep);
        const float* my = (const float*)((const char*)mapy + (size_t)y * mstep);
        uint8_t* d = dst + (size_t)y * dstep;
        for (int x = 0; x < w; ++x) {
            int sx = (int)lrintf(mx[x] * 32.0f), sy = (int)lrintf(my[x] * 32.0f);
            int X = sat16i(sx >> 5) - 1, Y = sat16i(sy >> 5) - 1;
            const int16_t* wt = tab + (size_t)((sy & 31) * 32 + (sx & 31)) * 16;
            for (int c = 0; c < 3; ++c) {
                int sum = 0;
                for (int k1 = 0; k1 < 4; ++k1)
                    for (int k2 = 0; k2 < 4; ++k2) {
                        int yy = Y + k1, xx = X + k2;
                        int v = (yy >= 0 && yy < sh && xx >= 0 && xx < sw) ? src[(size_t)yy * sstep + (size_t)xx * 3 + c]
                                                                          : border[c];
                        sum += wt[k1 * 4 + k2] * v;
                    }
                d[x * 3 + c] = (uint8_t)iclamp((sum + (1 << 14)) >> 15, 0, 255);
            }
        }
    }
    free(tab);
}

/* ---- the other interpolation modes of cv::remap (calc_rectified.hpp:17-22: NEAREST, LINEAR, CUBIC, LANCZOS4) ----------
 * OpenCV builds one table of K x K Q15 weights per (fy, fx) in 1/32 pixel steps (initInterTab2D): K = 2 / 4 / 8 for
 * LINEAR / CUBIC / LANCZOS4.  Entries that do not sum to 32768 are corrected on one of the four taps k1, k2 in
 * {K/2, K/2 + 1}; for K = 2 that window reaches into the next table entry, which is still zero when OpenCV fills the
 * table in order -- reproduced here by filling a zero-initialised flat array in the same order. */
static void orc_interp_coeffs(int interp, float x, float* c)
{
    if (interp == 1) { c[0] = 1.f - x; c[1] = x; return; }
    if (interp == 2) {
        const float A = -0.75f;
        c[0] = ((A * (x + 1) - 5 * A) * (x + 1) + 8 * A) * (x + 1) - 4 * A;
        c[1] = ((A + 2) * x - (A + 3)) * x * x + 1;
        c[2] = ((A + 2) * (1 - x) - (A + 3)) * (1 - x) * (1 - x) + 1;
        c[3] = 1.f - c[0] - c[1] - c[2];
        return;
    }
    /* interpolateLanczos4 */
    static const double s45 = 0.70710678118654752440084436210485;
    static const double cs[8][2] = {{1, 0}, {-s45, -s45}, {0, 1}, {s45, -s45}, {-1, 0}, {s45, s45}, {0, -1}, {-s45, s45}};
    if (x < 1.1920928955078125e-07f) { for (int i = 0; i < 8; ++i) c[i] = 0; c[3] = 1; return; }
    float sum = 0;
    double y0 = -(x + 3) * 3.1415926535897932384626433832795 * 0.25, s0 = sin(y0), c0 = cos(y0);
    for (int i = 0; i < 8; ++i) {
        double y = -(x + 3 - i) * 3.1415926535897932384626433832795 * 0.25;
        c[i] = (float)((cs[i][0] * s0 + cs[i][1] * c0) / (y * y));
        sum += c[i];
    }
    sum = 1.f / sum;
    for (int i = 0; i < 8; ++i) c[i] *= sum;
}

int orc_interp_ksize(int interp) { return interp == 1 ? 2 : interp == 2 ? 4 : interp == 4 ? 8 : 0; }

/* tab[(fy*32+fx)*K*K + k1*K + k2]; tab must hold 1024*K*K entries */
int orc_interp_tab(int interp, int16_t* tab)
{
    const int K = orc_interp_ksize(interp);
    if (!K) return -1;
    float c[32][8];
    for (int i = 0; i < 32; ++i) orc_interp_coeffs(interp, (float)i * (1.0f / 32.0f), c[i]);
    const size_t n = (size_t)1024 * K * K;
    int16_t* flat = (int16_t*)calloc(n + 4 * K * K, sizeof(int16_t));
    if (!flat) return -2;
    for (int fy = 0; fy < 32; ++fy)
        for (int fx = 0; fx < 32; ++fx) {
            int16_t* w = flat + (size_t)(fy * 32 + fx) * K * K;
            int isum = 0;
            for (int k1 = 0; k1 < K; ++k1)
                for (int k2 = 0; k2 < K; ++k2) {
                    float v = c[fy][k1] * c[fx][k2];
                    int q = sat16i((int)lrintf(v * 32768.0f));
                    w[k1 * K + k2] = (int16_t)q;
                    isum += q;
                }
            if (isum != 32768) {
                const int diff = isum - 32768, k0 = K / 2;
                int M = k0 * K + k0, m = k0 * K + k0;
                for (int k1 = k0; k1 < k0 + 2; ++k1)
                    for (int k2 = k0; k2 < k0 + 2; ++k2) {
                        int i = k1 * K + k2;
                        if (w[i] < w[m])
                            m = i;
                        else if (w[i] > w[M])
                            M = i;
                    }
                if (diff < 0)
                    w[M] = (int16_t)(w[M] - diff);
                else
                    w[m] = (int16_t)(w[m] - diff);
            }
        }
    memcpy(tab, flat, n * sizeof(int16_t));
    free(flat);
    return 0;
}

/* cv::remap, CV_32FC1 maps, 8UC3, BORDER_CONSTANT(border); interp = cv::INTER_NEAREST 0 | LINEAR 1 | CUBIC 2 | LANCZOS4 4 */
int orc_remap_bgr8(int interp, const uint8_t* src, int sw, int sh, size_t sstep, const float* mapx, const float* mapy,
                   size_t mstep, int w, int h, const uint8_t border[3], uint8_t* dst, size_t dstep)
{
    if (interp == 0) {
        for (int y = 0; y < h; ++y) {
            const float* mx = (const float*)((const char*)mapx + (size_t)y * mstep);
            const float* my = (const float*)((const char*)mapy + (size_t)y * mstep);
            uint8_t* d = dst + (size_t)y * dstep;
            for (int x = 0; x < w; ++x) {
                /* cvRound + saturate_cast<short>; NaN / out of range give INT_MIN -> -32768 */
                int X = sat16i((fabsf(mx[x]) < 2147483648.0f) ? (int)lrintf(mx[x]) : (int)0x80000000);
                int Y = sat16i((fabsf(my[x]) < 2147483648.0f) ? (int)lrintf(my[x]) : (int)0x80000000);
                const int in = X >= 0 && X < sw && Y >= 0 && Y < sh;
                for (int c = 0; c < 3; ++c) d[x * 3 + c] = in ? src[(size_t)Y * sstep + (size_t)X * 3 + c] : border[c];
            }
        }
        return 0;
    }
    const int K = orc_interp_ksize(interp);
    if (!K) return -1;
    int16_t* tab = (int16_t*)malloc(sizeof(int16_t) * 1024 * K * K);
    if (!tab || orc_interp_tab(interp, tab)) { free(tab); return -2; }
    for (int y = 0; y < h; ++y) {
        const float* mx = (const float*)((const char*)mapx + (size_t)y * mstep);
        const float* my = (const float*)((const char*)mapy + (size_t)y * mstep);
        uint8_t* d = dst + (size_t)y * dstep;
        for (int x = 0; x < w; ++x) {
            float fx = mx[x] * 32.0f, fy = my[x] * 32.0f;
            int sx = (fabsf(fx) < 2147483648.0f) ? (int)lrintf(fx) : (int)0x80000000;
            int sy = (fabsf(fy) < 2147483648.0f) ? (int)lrintf(fy) : (int)0x80000000;
            int X = sat16i(sx >> 5) - (K / 2 - 1), Y = sat16i(sy >> 5) - (K / 2 - 1);
            const int16_t* wt = tab + (size_t)((sy & 31) * 32 + (sx & 31)) * K * K;
            for (int c = 0; c < 3; ++c) {
                int sum = 0;
                for (int k1 = 0; k1 < K; ++k1)
                    for (int k2 = 0; k2 < K; ++k2) {
                        int yy = Y + k1, xx = X + k2;
                        int v = (yy >= 0 && yy < sh && xx >= 0 && xx < sw) ? src[(size_t)yy * sstep + (size_t)xx * 3 + c]
                                                                          : border[c];
                        sum += wt[k1 * K + k2] * v;
                    }
                d[x * 3 + c] = (uint8_t)iclamp((sum + (1 << 14)) >> 15, 0, 255);
            }
        }
    }
    free(tab);
    return 0;
}

/* A.6: cvtColor BGR2GRAY 8U */
void orc_bgr2gray(const uint8_t* src, size_t sstep, int w, int h, uint8_t* dst, size_t dstep)
{
    for (int y = 0; y < h; ++y)
        for (int x = 0; x < w; ++x) {
            const uint8_t* p = src + (size_t)y * sstep + (size_t)x * 3;
            dst[(size_t)y * dstep + x] = (uint8_t)((3735 * p[0] + 19235 * p[1] + 9798 * p[2] + 16384) >> 15);
        }
}

/* A.6: Q8 Gaussian taps by error diffusion from the edge; k has n entries; returns 0 on success. */
int orc_gauss_taps_q8(int n, double sigma, int* k)
{
    if (n <= 0 || !(n & 1) || n > 63) return -1;
    double kd[64], sum = 0;
    if (sigma <= 0) sigma = ((n - 1) * 0.5 - 1) * 0.3 + 0.8;
    double scale2x = -0.5 / (sigma * sigma);
    for (int i = 0; i < n; ++i) {
        double x = i - (n - 1) * 0.5;
        kd[i] = exp(scale2x * x * x);
        sum += kd[i];
    }
    sum = 1.0 / sum;
    double err = 0;
    int acc = 0;
    for (int i = 0; i < n / 2; ++i) {
        double adj = kd[i] * sum * 256.0 + err;
        int v = (int)lrint(adj);
        err = adj - v;
        k[i] = k[n - 1 - i] = v;
        acc += 2 * v;
    }
    k[n / 2] = 256 - acc;
    return 0;
}

static inline int reflect101(int i, int n)
{
    if (n == 1) return 0;
    while (i < 0 || i >= n) {
        if (i < 0) i = -i;
        if (i >= n) i = 2 * (n - 1) - i;
    }
    return i;
}

/* A.6: GaussianBlur CV_8U (n x n, sigma), BORDER_REFLECT_101. src != dst. */
int orc_gauss_blur_u8(const uint8_t* src, size_t sstep, int w, int h, int n, double sigma, uint8_t* dst, size_t dstep)
{
    int k[64];
    if (orc_gauss_taps_q8(n, sigma, k)) return -1;
    int r = n / 2;
    uint32_t* tmp = (uint32_t*)malloc(sizeof(uint32_t) * (size_t)w * h);
    if (!tmp) return -2;
    for (int y = 0; y < h; ++y)
        for (int x = 0; x < w; ++x) {
            uint32_t s = 0;
            for (int j = -r; j <= r; ++j) s += (uint32_t)k[j + r] * src[(size_t)y * sstep + reflect101(x + j, w)];
            tmp[(size_t)y * w + x] = s;
        }
    for (int y = 0; y < h; ++y)
        for (int x = 0; x < w; ++x) {
            uint32_t s = 0;
            for (int i = -r; i <= r; ++i) s += (uint32_t)k[i + r] * tmp[(size_t)reflect101(y + i, h) * w + x];
            dst[(size_t)y * dstep + x] = (uint8_t)((s + (1u << 15)) >> 16);
        }
    free(tmp);
    return 0;
}

/* ---- input side: input_images::Task::execute, pipeline/nodes/input_images.cpp:65-76 ---------------------
 * equal = raw_image_equalize ? hist_equalize(raw) : raw  (graphics/hist-equalize.cpp:9-30: BGR -> YCrCb, equalizeHist on
 * Y, YCrCb -> BGR), grey = cvtColor(equal, BGR2GRAY).  OpenCV 8-bit fixed point (14 fractional bits), pinned against cv2. */
static inline uint8_t sat8i(long v) { return (uint8_t)(v < 0 ? 0 : (v > 255 ? 255 : v)); }
static inline void orc_bgr2ycrcb(const uint8_t* p, int* Y, int* Cr, int* Cb)
{
    const int B = p[0], G = p[1], Rr = p[2];
    const int y = (Rr * 4899 + G * 9617 + B * 1868 + (1 << 13)) >> 14;
    *Y = y;   /* <= 255 by construction */
    *Cr = sat8i(((Rr - y) * 11682 + 128 * (1 << 14) + (1 << 13)) >> 14);
    *Cb = sat8i(((B - y) * 9241 + 128 * (1 << 14) + (1 << 13)) >> 14);
}
/* cv::equalizeHist look-up table from the 256-bin histogram of `total` pixels */
void orc_equalize_lut(const int* hist, long total, uint8_t* lut)
{
    int i = 0;
    while (i < 256 && !hist[i]) ++i;
    if (i == 256 || hist[i] == total) {   /* one grey level: the image is filled with it */
        for (int k = 0; k < 256; ++k) lut[k] = (uint8_t)(i < 256 ? i : 0);
        return;
    }
    const float scale = 255.f / (float)(total - hist[i]);
    long sum = 0;
    for (int k = 0; k <= i; ++k) lut[k] = 0;
    for (++i; i < 256; ++i) {
        sum += hist[i];
        lut[i] = sat8i(lrintf((float)sum * scale));
    }
}
void orc_input_images(const uint8_t* raw, int w, int h, size_t sstep, int equalize, uint8_t* equal, size_t estep, uint8_t* grey,
                      size_t gstep)
{
    uint8_t lut[256];
    if (equalize) {
        int hist[256] = {0};
        for (int y = 0; y < h; ++y)
            for (int x = 0; x < w; ++x) {
                int Y, Cr, Cb;
                orc_bgr2ycrcb(raw + (size_t)y * sstep + (size_t)x * 3, &Y, &Cr, &Cb);
                hist[Y]++;
            }
        orc_equalize_lut(hist, (long)w * h, lut);
    }
    for (int y = 0; y < h; ++y)
        for (int x = 0; x < w; ++x) {
            const uint8_t* p = raw + (size_t)y * sstep + (size_t)x * 3;
            uint8_t o[3] = {p[0], p[1], p[2]};
            if (equalize) {
                int Y, Cr, Cb;
                orc_bgr2ycrcb(p, &Y, &Cr, &Cb);
                Y = lut[Y];
                o[0] = sat8i(Y + (((Cb - 128) * 29049 + (1 << 13)) >> 14));
                o[1] = sat8i(Y + (((Cb - 128) * -5636 + (Cr - 128) * -11698 + (1 << 13)) >> 14));
                o[2] = sat8i(Y + (((Cr - 128) * 22987 + (1 << 13)) >> 14));
            }
            if (equal) { uint8_t* e = equal + (size_t)y * estep + (size_t)x * 3; e[0] = o[0]; e[1] = o[1]; e[2] = o[2]; }
            if (grey) grey[(size_t)y * gstep + x] = (uint8_t)((3735 * o[0] + 19235 * o[1] + 9798 * o[2] + 16384) >> 15);
        }
}
