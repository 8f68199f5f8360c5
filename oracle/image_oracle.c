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
