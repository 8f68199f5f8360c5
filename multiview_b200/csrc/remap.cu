// a1: rectification remap, bit-exact with cv::remap(INTER_CUBIC, BORDER_CONSTANT) on 8UC3
// (reference call: pipeline/nodes/calc_rectified.cpp:109-115).
//
// OpenCV's fixed-point scheme: map coordinates are rounded to 1/32 pixel, the 4x4 tap weights come
// from a 32x32 table of Q15 integers whose entries are forced to sum to 32768, and the result is
// (sum + 2^14) >> 15.  One thread produces one output pixel (3 channels); the 32 KB weight table and
// the source image are read through L1/L2 (the source footprint of a tile is data dependent).
#include "common.cuh"

// Host: build the 32x32x16 weight table exactly as OpenCV's initInterTab2D(INTER_CUBIC, fixpt).
// Compiled with -ffp-contract=off so that the float expressions round like the x86 baseline build.
static void build_cubic_tab(int16_t* tab)
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
                    long q = lrintf(v * 32768.0f);
                    q = q < -32768 ? -32768 : (q > 32767 ? 32767 : q);
                    w[k1 * 4 + k2] = (int16_t)q;
                    isum += (int)q;
                }
            if (isum != 32768) {
                // the correction goes to the largest / smallest of the four central taps
                int diff = isum - 32768, M = 10, m = 10;
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

// The other fixed-point modes of cv::remap: K x K taps, K = 2 (LINEAR) / 8 (LANCZOS4), same table construction
// (initInterTab2D).  The correction window {K/2, K/2+1}^2 reaches into the next entry for K = 2, which is still zero
// while OpenCV fills the table in order: the table is built in a zero-initialised flat array in the same order.
static int interp_ksize(int interp) { return interp == MVS_INTER_LINEAR ? 2 : interp == MVS_INTER_CUBIC ? 4 : interp == MVS_INTER_LANCZOS4 ? 8 : 0; }
static void interp_coeffs(int interp, float x, float* c)
{
    if (interp == MVS_INTER_LINEAR) { c[0] = 1.f - x; c[1] = x; return; }
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
static void build_interp_tab(int interp, std::vector<int16_t>& flat)
{
    const int K = interp_ksize(interp);
    float c[32][8];
    for (int i = 0; i < 32; ++i) interp_coeffs(interp, (float)i * (1.0f / 32.0f), c[i]);
    flat.assign((size_t)1024 * K * K + 4 * K * K, 0);
    for (int fy = 0; fy < 32; ++fy)
        for (int fx = 0; fx < 32; ++fx) {
            int16_t* w = flat.data() + (size_t)(fy * 32 + fx) * K * K;
            int isum = 0;
            for (int k1 = 0; k1 < K; ++k1)
                for (int k2 = 0; k2 < K; ++k2) {
                    float v = c[fy][k1] * c[fx][k2];
                    long q = lrintf(v * 32768.0f);
                    q = q < -32768 ? -32768 : (q > 32767 ? 32767 : q);
                    w[k1 * K + k2] = (int16_t)q;
                    isum += (int)q;
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
}
static int ensure_interp_tab(mvs_ctx* ctx, int interp)
{
    if (ctx->interp_tab_dev[interp]) return MVS_OK;
    const int K = interp_ksize(interp);
    char nm[32];
    snprintf(nm, sizeof(nm), "remap.tab.%d", interp);
    MVS_ARENA(ctx, int16_t, tab, nm, (size_t)1024 * K * K);
    std::vector<int16_t> host;
    build_interp_tab(interp, host);
    MVS_CUDA(ctx, cudaMemcpyAsync(tab, host.data(), (size_t)1024 * K * K * sizeof(int16_t), cudaMemcpyHostToDevice, ctx->stream));
    MVS_CUDA(ctx, cudaStreamSynchronize(ctx->stream));  // host vector goes out of scope
    ctx->interp_tab_dev[interp] = tab;
    return MVS_OK;
}

static int ensure_cubic_tab(mvs_ctx* ctx)
{
    if (ctx->cubic_tab_dev) return MVS_OK;
    MVS_ARENA(ctx, int16_t, tab, "remap.cubic_tab", 1024 * 16);
    std::vector<int16_t> host(1024 * 16);
    build_cubic_tab(host.data());
    MVS_CUDA(ctx, cudaMemcpyAsync(tab, host.data(), host.size() * sizeof(int16_t), cudaMemcpyHostToDevice, ctx->stream));
    MVS_CUDA(ctx, cudaStreamSynchronize(ctx->stream));  // host vector goes out of scope
    ctx->cubic_tab_dev = tab;
    return MVS_OK;
}

// cvRound(float) on x86 (cvtss2si): round-half-even; NaN and out-of-range give INT_MIN.
__device__ __forceinline__ int cv_round_f(float v)
{
    return (fabsf(v) < 2147483648.0f) ? __float2int_rn(v) : (int)0x80000000;
}
__device__ __forceinline__ int sat16_i(int v) { return max(-32768, min(32767, v)); }

// d + a.lo16 * b.byte0 + a.hi16 * b.byte1 (signed weights, unsigned pixels) and the same with bytes 2, 3
__device__ __forceinline__ int dp2a_lo_su(uint32_t a, uint32_t b, int c)
{
    int d;
    asm("dp2a.lo.s32.u32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
    return d;
}
__device__ __forceinline__ int dp2a_hi_su(uint32_t a, uint32_t b, int c)
{
    int d;
    asm("dp2a.hi.s32.u32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
    return d;
}

__device__ __forceinline__ void
remap_cubic_bgr8_body(const uint8_t* __restrict__ src, int sw, int sh, size_t sstep, const float* __restrict__ mapx,
                        const float* __restrict__ mapy, size_t mstep, int w, int h, const int16_t* __restrict__ tab,
                        int b0, int b1, int b2, uint8_t* __restrict__ dst, size_t dstep, bool aligned4)
{
    int x = blockIdx.x * 32 + (threadIdx.x & 31), y = blockIdx.y * 8 + (threadIdx.x >> 5);
    if (x >= w || y >= h) return;
    int X, Y, widx;
    if (mapy) {
        float mx = *(const float*)((const char*)mapx + (size_t)y * mstep + (size_t)x * 4);
        float my = *(const float*)((const char*)mapy + (size_t)y * mstep + (size_t)x * 4);
        int sx = cv_round_f(mx * 32.0f), sy = cv_round_f(my * 32.0f);
        X = sat16_i(sx >> 5) - 1; Y = sat16_i(sy >> 5) - 1;
        widx = ((sy & 31) << 5) + (sx & 31);
    } else {
        // registered maps (mvs_maps_register): mapx points at the packed form, [h][w] uint32 {X | Y << 16} followed by
        // [h][w] uint16 weight-table indices: 6 bytes per pixel instead of 8, and no float -> fixed conversion per frame
        const uint32_t* pxy = (const uint32_t*)mapx;
        const uint16_t* pfr = (const uint16_t*)(pxy + (size_t)w * h);
        const uint32_t v = __ldg(pxy + (size_t)y * w + x);
        X = (int)(int16_t)(v & 0xffffu) - 1; Y = (int)(int16_t)(v >> 16) - 1;
        widx = __ldg(pfr + (size_t)y * w + x);
    }
    const int4* wp = (const int4*)(tab + (size_t)widx * 16);
    int4 wa = __ldg(wp), wb = __ldg(wp + 1);
    int wraw[8] = {wa.x, wa.y, wa.z, wa.w, wb.x, wb.y, wb.z, wb.w};
    int s0 = 0, s1 = 0, s2 = 0;
    uint8_t* d = dst + (size_t)y * dstep + (size_t)x * 3;
    if (X >= sw || X + 4 <= 0 || Y >= sh || Y + 4 <= 0) {
        d[0] = (uint8_t)b0; d[1] = (uint8_t)b1; d[2] = (uint8_t)b2;
        return;
    }
    bool inside = X >= 0 && X + 3 < sw && Y >= 0 && Y + 3 < sh;
    // fast path: the 12 bytes of a tap row come from four aligned words (the kernel is bound by L1 requests, not by
    // arithmetic), the weighted sums are 16-bit x 8-bit dot products.  Not taken where the last word could reach
    // past the end of the image buffer.
    if (inside && aligned4 && (Y + 3 < sh - 1 || X + 4 < sw)) {
        const uint8_t* p0 = src + (size_t)Y * sstep + (size_t)X * 3;
#pragma unroll
        for (int k1 = 0; k1 < 4; ++k1) {
            const uintptr_t addr = (uintptr_t)(p0 + (size_t)k1 * sstep);
            const int sh8 = ((int)addr & 3) * 8;   // per row: the row pitch need not be a multiple of four
            const uint32_t* a = (const uint32_t*)(addr & ~(uintptr_t)3);
            const uint32_t A0 = __ldg(a), A1 = __ldg(a + 1), A2 = __ldg(a + 2), A3 = __ldg(a + 3);
            // w0 = B0 G0 R0 B1, w1 = G1 R1 B2 G2, w2 = R2 B3 G3 R3
            const uint32_t w0 = __funnelshift_r(A0, A1, sh8), w1 = __funnelshift_r(A1, A2, sh8), w2 = __funnelshift_r(A2, A3, sh8);
            const uint32_t wl = (uint32_t)wraw[k1 * 2], wh = (uint32_t)wraw[k1 * 2 + 1];  // taps (0,1) and (2,3) as s16 pairs
            const uint32_t cb = __byte_perm(__byte_perm(w0, w1, 0x0630), w2, 0x5210);   // B0 B1 B2 B3
            const uint32_t cg = __byte_perm(__byte_perm(w0, w1, 0x0741), w2, 0x6210);   // G0 G1 G2 G3
            const uint32_t cr = __byte_perm(__byte_perm(w0, w1, 0x0052), w2, 0x7410);   // R0 R1 R2 R3
            s0 = dp2a_hi_su(wh, cb, dp2a_lo_su(wl, cb, s0));
            s1 = dp2a_hi_su(wh, cg, dp2a_lo_su(wl, cg, s1));
            s2 = dp2a_hi_su(wh, cr, dp2a_lo_su(wl, cr, s2));
        }
        d[0] = (uint8_t)max(0, min(255, (s0 + (1 << 14)) >> 15));
        d[1] = (uint8_t)max(0, min(255, (s1 + (1 << 14)) >> 15));
        d[2] = (uint8_t)max(0, min(255, (s2 + (1 << 14)) >> 15));
        return;
    }
#pragma unroll
    for (int k1 = 0; k1 < 4; ++k1) {
        int yy = Y + k1;
        const uint8_t* row = src + (size_t)yy * sstep;
#pragma unroll
        for (int k2 = 0; k2 < 4; ++k2) {
            int wr = wraw[k1 * 2 + (k2 >> 1)];
            int wt = (k2 & 1) ? (wr >> 16) : (int)(short)(wr & 0xffff);
            int xx = X + k2;
            int v0 = b0, v1 = b1, v2 = b2;
            if (inside || (yy >= 0 && yy < sh && xx >= 0 && xx < sw)) {
                const uint8_t* p = row + (size_t)xx * 3;
                v0 = __ldg(p); v1 = __ldg(p + 1); v2 = __ldg(p + 2);
            }
            s0 += wt * v0; s1 += wt * v1; s2 += wt * v2;
        }
    }
    d[0] = (uint8_t)max(0, min(255, (s0 + (1 << 14)) >> 15));
    d[1] = (uint8_t)max(0, min(255, (s1 + (1 << 14)) >> 15));
    d[2] = (uint8_t)max(0, min(255, (s2 + (1 << 14)) >> 15));
}

__global__ void __launch_bounds__(256)
remap_cubic_bgr8_kernel(const uint8_t* __restrict__ src, int sw, int sh, size_t sstep, const float* __restrict__ mapx,
                        const float* __restrict__ mapy, size_t mstep, int w, int h, const int16_t* __restrict__ tab,
                        int b0, int b1, int b2, uint8_t* __restrict__ dst, size_t dstep, bool aligned4)
{
    remap_cubic_bgr8_body(src, sw, sh, sstep, mapx, mapy, mstep, w, h, tab, b0, b1, b2, dst, dstep, aligned4);
}
// batched: image blockIdx.z of up to 32 equally sized sensor images (one launch for all sensors of a frame batch)
struct RemapBatch {
    const uint8_t* src[32];
    const float* mapx[32];
    const float* mapy[32];
    uint8_t* dst[32];
};
__global__ void __launch_bounds__(256)
remap_cubic_bgr8_batch_kernel(RemapBatch b, int sw, int sh, size_t sstep, size_t mstep, int w, int h, const int16_t* __restrict__ tab,
                              int b0, int b1, int b2, size_t dstep, bool aligned4)
{
    const int z = blockIdx.z;
    remap_cubic_bgr8_body(b.src[z], sw, sh, sstep, b.mapx[z], b.mapy[z], mstep, w, h, tab, b0, b1, b2, b.dst[z], dstep, aligned4);
}

// INTER_NEAREST: cvRound + saturate_cast<short> of the coordinates, copy or border
__global__ void __launch_bounds__(256)
remap_nearest_bgr8_kernel(const uint8_t* __restrict__ src, int sw, int sh, size_t sstep, const float* __restrict__ mapx,
                          const float* __restrict__ mapy, size_t mstep, int w, int h, int b0, int b1, int b2,
                          uint8_t* __restrict__ dst, size_t dstep)
{
    int x = blockIdx.x * 32 + (threadIdx.x & 31), y = blockIdx.y * 8 + (threadIdx.x >> 5);
    if (x >= w || y >= h) return;
    float mx = *(const float*)((const char*)mapx + (size_t)y * mstep + (size_t)x * 4);
    float my = *(const float*)((const char*)mapy + (size_t)y * mstep + (size_t)x * 4);
    int X = sat16_i(cv_round_f(mx)), Y = sat16_i(cv_round_f(my));
    uint8_t* d = dst + (size_t)y * dstep + (size_t)x * 3;
    if (X >= 0 && X < sw && Y >= 0 && Y < sh) {
        const uint8_t* p = src + (size_t)Y * sstep + (size_t)X * 3;
        d[0] = p[0]; d[1] = p[1]; d[2] = p[2];
    } else {
        d[0] = (uint8_t)b0; d[1] = (uint8_t)b1; d[2] = (uint8_t)b2;
    }
}

// INTER_LINEAR (K = 2) and INTER_LANCZOS4 (K = 8): K x K Q15 taps, one thread per output pixel
template <int K>
__global__ void __launch_bounds__(256)
remap_taps_bgr8_kernel(const uint8_t* __restrict__ src, int sw, int sh, size_t sstep, const float* __restrict__ mapx,
                       const float* __restrict__ mapy, size_t mstep, int w, int h, const int16_t* __restrict__ tab,
                       int b0, int b1, int b2, uint8_t* __restrict__ dst, size_t dstep)
{
    int x = blockIdx.x * 32 + (threadIdx.x & 31), y = blockIdx.y * 8 + (threadIdx.x >> 5);
    if (x >= w || y >= h) return;
    float mx = *(const float*)((const char*)mapx + (size_t)y * mstep + (size_t)x * 4);
    float my = *(const float*)((const char*)mapy + (size_t)y * mstep + (size_t)x * 4);
    int sx = cv_round_f(mx * 32.0f), sy = cv_round_f(my * 32.0f);
    int X = sat16_i(sx >> 5) - (K / 2 - 1), Y = sat16_i(sy >> 5) - (K / 2 - 1);
    const int16_t* wt = tab + (size_t)(((sy & 31) << 5) + (sx & 31)) * (K * K);
    uint8_t* d = dst + (size_t)y * dstep + (size_t)x * 3;
    if (X >= sw || X + K <= 0 || Y >= sh || Y + K <= 0) {
        d[0] = (uint8_t)b0; d[1] = (uint8_t)b1; d[2] = (uint8_t)b2;
        return;
    }
    const bool inside = X >= 0 && X + K - 1 < sw && Y >= 0 && Y + K - 1 < sh;
    int s0 = 0, s1 = 0, s2 = 0;
#pragma unroll
    for (int k1 = 0; k1 < K; ++k1) {
        const int yy = Y + k1;
        const uint8_t* row = src + (size_t)yy * sstep;
#pragma unroll
        for (int k2 = 0; k2 < K; ++k2) {
            const int wv = __ldg(wt + k1 * K + k2), xx = X + k2;
            int v0 = b0, v1 = b1, v2 = b2;
            if (inside || (yy >= 0 && yy < sh && xx >= 0 && xx < sw)) {
                const uint8_t* p = row + (size_t)xx * 3;
                v0 = __ldg(p); v1 = __ldg(p + 1); v2 = __ldg(p + 2);
            }
            s0 += wv * v0; s1 += wv * v1; s2 += wv * v2;
        }
    }
    d[0] = (uint8_t)max(0, min(255, (s0 + (1 << 14)) >> 15));
    d[1] = (uint8_t)max(0, min(255, (s1 + (1 << 14)) >> 15));
    d[2] = (uint8_t)max(0, min(255, (s2 + (1 << 14)) >> 15));
}

int remap_bgr8_dev(mvs_ctx* ctx, const uint8_t* src, int sw, int sh, size_t sstep, const float* mapx,
                   const float* mapy, size_t mstep, int w, int h, int interp, const uint8_t border[3], uint8_t* dst,
                   size_t dstep)
{
    dim3 grid(cdiv(w, 32), cdiv(h, 8));
    if (interp == MVS_INTER_NEAREST) {
        remap_nearest_bgr8_kernel<<<grid, 256, 0, ctx->stream>>>(src, sw, sh, sstep, mapx, mapy, mstep, w, h, border[0], border[1],
                                                                 border[2], dst, dstep);
        MVS_LAUNCH_CHECK(ctx, "remap_nearest_bgr8_kernel");
        return MVS_OK;
    }
    if (interp == MVS_INTER_LINEAR || interp == MVS_INTER_LANCZOS4) {
        int rc = ensure_interp_tab(ctx, interp);
        if (rc) return rc;
        if (interp == MVS_INTER_LINEAR)
            remap_taps_bgr8_kernel<2><<<grid, 256, 0, ctx->stream>>>(src, sw, sh, sstep, mapx, mapy, mstep, w, h, ctx->interp_tab_dev[interp],
                                                                     border[0], border[1], border[2], dst, dstep);
        else
            remap_taps_bgr8_kernel<8><<<grid, 256, 0, ctx->stream>>>(src, sw, sh, sstep, mapx, mapy, mstep, w, h, ctx->interp_tab_dev[interp],
                                                                     border[0], border[1], border[2], dst, dstep);
        MVS_LAUNCH_CHECK(ctx, "remap_taps_bgr8_kernel");
        return MVS_OK;
    }
    if (interp != MVS_INTER_CUBIC)
        return mvs_fail(ctx, MVS_ERR_UNSUPPORTED, "remap: interpolation %d (cv::INTER_NEAREST, LINEAR, CUBIC and LANCZOS4 are implemented)", interp);
    int rc = ensure_cubic_tab(ctx);
    if (rc) return rc;
    remap_cubic_bgr8_kernel<<<grid, 256, 0, ctx->stream>>>(src, sw, sh, sstep, mapx, mapy, mstep, w, h,
                                                           ctx->cubic_tab_dev, border[0], border[1], border[2], dst, dstep,
                                                           ((uintptr_t)src & 3) == 0);
    MVS_LAUNCH_CHECK(ctx, "remap_cubic_bgr8_kernel");
    return MVS_OK;
}

// n (<= 32) sensor images of equal geometry, INTER_CUBIC, one launch
int remap_cubic_bgr8_batch_dev(mvs_ctx* ctx, int n, const uint8_t* const* src, int sw, int sh, size_t sstep, const float* const* mapx,
                               const float* const* mapy, size_t mstep, int w, int h, const uint8_t border[3], uint8_t* const* dst,
                               size_t dstep)
{
    if (n < 1) return mvs_fail(ctx, MVS_ERR_ARG, "remap batch of %d images", n);
    if (n > 32) {  // larger batches go through in chunks
        for (int i0 = 0; i0 < n; i0 += 32) {
            int rc = remap_cubic_bgr8_batch_dev(ctx, n - i0 < 32 ? n - i0 : 32, src + i0, sw, sh, sstep, mapx + i0, mapy + i0, mstep, w, h, border,
                                                dst + i0, dstep);
            if (rc) return rc;
        }
        return MVS_OK;
    }
    int rc = ensure_cubic_tab(ctx);
    if (rc) return rc;
    RemapBatch b;
    bool aligned = true;
    for (int i = 0; i < n; ++i) {
        b.src[i] = src[i]; b.mapx[i] = mapx[i]; b.mapy[i] = mapy[i]; b.dst[i] = dst[i];
        aligned = aligned && ((uintptr_t)src[i] & 3) == 0;
    }
    remap_cubic_bgr8_batch_kernel<<<dim3(cdiv(w, 32), cdiv(h, 8), n), 256, 0, ctx->stream>>>(b, sw, sh, sstep, mstep, w, h, ctx->cubic_tab_dev,
                                                                                        border[0], border[1], border[2], dstep, aligned);
    MVS_LAUNCH_CHECK(ctx, "remap_cubic_bgr8_batch_kernel");
    return MVS_OK;
}

extern "C" int mvs_remap_bgr8(mvs_ctx* ctx, const uint8_t* src, int sw, int sh, size_t sstep, const float* mapx,
                              const float* mapy, size_t mstep, int w, int h, int interp, const uint8_t border_bgr[3],
                              uint8_t* dst, size_t dstep)
{
    if (!ctx) return MVS_ERR_ARG;
    if (!src || !mapx || !mapy || !dst || !border_bgr || sw <= 0 || sh <= 0 || w <= 0 || h <= 0 ||
        sstep < (size_t)sw * 3 || mstep < (size_t)w * 4 || dstep < (size_t)w * 3)
        return mvs_fail(ctx, MVS_ERR_ARG, "mvs_remap_bgr8: bad argument");
    MVS_CUDA(ctx, cudaSetDevice(ctx->device));
    MVS_ARENA(ctx, uint8_t, dsrc, "remap.src", (size_t)sh * sw * 3);
    MVS_ARENA(ctx, float, dmx, "remap.mapx", (size_t)h * w);
    MVS_ARENA(ctx, float, dmy, "remap.mapy", (size_t)h * w);
    MVS_ARENA(ctx, uint8_t, ddst, "remap.dst", (size_t)h * w * 3);
    cudaStream_t s = ctx->stream;
    MVS_CUDA(ctx, cudaMemcpy2DAsync(dsrc, (size_t)sw * 3, src, sstep, (size_t)sw * 3, sh, cudaMemcpyHostToDevice, s));
    MVS_CUDA(ctx, cudaMemcpy2DAsync(dmx, (size_t)w * 4, mapx, mstep, (size_t)w * 4, h, cudaMemcpyHostToDevice, s));
    MVS_CUDA(ctx, cudaMemcpy2DAsync(dmy, (size_t)w * 4, mapy, mstep, (size_t)w * 4, h, cudaMemcpyHostToDevice, s));
    int rc = remap_bgr8_dev(ctx, dsrc, sw, sh, (size_t)sw * 3, dmx, dmy, (size_t)w * 4, w, h, interp, border_bgr, ddst,
                            (size_t)w * 3);
    if (rc) return rc;
    MVS_CUDA(ctx, cudaMemcpy2DAsync(dst, dstep, ddst, (size_t)w * 3, (size_t)w * 3, h, cudaMemcpyDeviceToHost, s));
    MVS_CUDA(ctx, cudaStreamSynchronize(s));
    return MVS_OK;
}

// ---- f1: the per-video half of get_xy_mapping / create_cv_remap that needs no distortion model --------
// The maps a camera's get_xy_mapping task produces once per video (pipeline/nodes/get_xy_mapping.cpp:59,
// utils/create-cv-remap.cpp:10-68) are converted once into what the cubic remap derives from them for every pixel of
// every frame: the integer source corner saturate_cast<short>(cvRound(map * 32) >> 5) per axis and the index of the Q15
// weight table (the two 5-bit fractions).  Bit-exact by construction (the same operations, done earlier).
__global__ void maps_pack_kernel(const float* __restrict__ mapx, const float* __restrict__ mapy, size_t mstep, int w, int h,
                                 uint32_t* __restrict__ pxy, uint16_t* __restrict__ pfr)
{
    int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y;
    if (x >= w) return;
    float mx = *(const float*)((const char*)mapx + (size_t)y * mstep + (size_t)x * 4);
    float my = *(const float*)((const char*)mapy + (size_t)y * mstep + (size_t)x * 4);
    int sx = cv_round_f(mx * 32.0f), sy = cv_round_f(my * 32.0f);
    pxy[(size_t)y * w + x] = ((uint32_t)sat16_i(sx >> 5) & 0xffffu) | ((uint32_t)sat16_i(sy >> 5) << 16);
    pfr[(size_t)y * w + x] = (uint16_t)(((sy & 31) << 5) + (sx & 31));
}

extern "C" int mvs_maps_register(mvs_ctx* ctx, const float* mapx, const float* mapy, size_t mstep, int w, int h, int on_device,
                                 int* handle)
{
    if (!ctx) return MVS_ERR_ARG;
    if (!mapx || !mapy || !handle || w <= 0 || h <= 0 || mstep < (size_t)w * 4) return mvs_fail(ctx, MVS_ERR_ARG, "mvs_maps_register: bad argument");
    MVS_CUDA(ctx, cudaSetDevice(ctx->device));
    cudaStream_t s = ctx->stream;
    const float *dx = mapx, *dy = mapy;
    size_t dstep = mstep;
    if (!on_device) {
        float* tx = (float*)arena_get(ctx, "maps.reg.x", (size_t)w * h * 4);
        float* ty = (float*)arena_get(ctx, "maps.reg.y", (size_t)w * h * 4);
        if (!tx || !ty) return MVS_ERR_NOMEM;
        MVS_CUDA(ctx, cudaMemcpy2DAsync(tx, (size_t)w * 4, mapx, mstep, (size_t)w * 4, h, cudaMemcpyHostToDevice, s));
        MVS_CUDA(ctx, cudaMemcpy2DAsync(ty, (size_t)w * 4, mapy, mstep, (size_t)w * 4, h, cudaMemcpyHostToDevice, s));
        dx = tx; dy = ty; dstep = (size_t)w * 4;
    }
    int id = 0;
    for (; id < (int)ctx->maps.size(); ++id)
        if (!ctx->maps[id].packed) break;
    if (id == (int)ctx->maps.size()) ctx->maps.push_back(RegisteredMaps());
    void* p = nullptr;
    if (cudaMalloc(&p, (size_t)w * h * 6) != cudaSuccess) return mvs_fail(ctx, MVS_ERR_NOMEM, "mvs_maps_register: cudaMalloc(%zu) failed", (size_t)w * h * 6);
    ctx->maps[id].packed = p; ctx->maps[id].w = w; ctx->maps[id].h = h;
    maps_pack_kernel<<<dim3(cdiv(w, 256), h), 256, 0, s>>>(dx, dy, dstep, w, h, (uint32_t*)p, (uint16_t*)((uint32_t*)p + (size_t)w * h));
    MVS_LAUNCH_CHECK(ctx, "maps_pack_kernel");
    MVS_CUDA(ctx, cudaStreamSynchronize(s));   // the caller may free its maps on return
    *handle = id + 1;
    return MVS_OK;
}

extern "C" int mvs_maps_release(mvs_ctx* ctx, int handle)
{
    if (!ctx) return MVS_ERR_ARG;
    if (handle < 1 || handle > (int)ctx->maps.size() || !ctx->maps[handle - 1].packed) return mvs_fail(ctx, MVS_ERR_ARG, "mvs_maps_release: unknown handle");
    MVS_CUDA(ctx, cudaSetDevice(ctx->device));
    mvs_sync(ctx);
    cudaFree(ctx->maps[handle - 1].packed);
    ctx->maps[handle - 1] = RegisteredMaps();
    return MVS_OK;
}
