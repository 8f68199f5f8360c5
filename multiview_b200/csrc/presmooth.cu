// a2: BGR->grey and the 31x31 Gaussian pre-smoothing applied twice
// (reference: cost-functions/disparity/disparity.cpp:106-155), bit-exact with OpenCV's 8-bit path:
//   grey = (3735 B + 19235 G + 9798 R + 2^14) >> 15
//   blur = (sum_i sum_j k_i k_j src + 2^15) >> 16 with Q8 integer taps k (error-diffused rounding of
//          the normalised Gaussian), BORDER_REFLECT_101; separable, one rounding at the very end.
// The images are a few MB and stay in L2; the passes are plain coalesced kernels.
#include <math.h>

#include "common.cuh"

struct BlurTaps {
    int k[31];
    int lo, hi;  // first / last non-zero tap
};

// OpenCV getGaussianKernel (double) + fixed-point conversion with error diffusion from the edge.
static int gauss_taps_q8(double sigma, BlurTaps* t)
{
    const int n = 31;
    if (!(sigma > 0.0)) return -1;
    double kd[31], sum = 0;
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
        t->k[i] = t->k[n - 1 - i] = v;
        acc += 2 * v;
    }
    t->k[n / 2] = 256 - acc;
    t->lo = 0;
    while (t->lo < n / 2 && t->k[t->lo] == 0) ++t->lo;
    t->hi = n - 1 - t->lo;
    return 0;
}

__device__ __forceinline__ int reflect101(int i, int n)
{
    if (n == 1) return 0;
    while (i < 0 || i >= n) {
        if (i < 0) i = -i;
        if (i >= n) i = 2 * (n - 1) - i;
    }
    return i;
}

__global__ void bgr2gray_kernel(const uint8_t* __restrict__ src, size_t sstep, int w, int h, uint8_t* __restrict__ dst,
                                size_t dstep)
{
    int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y;
    if (x >= w) return;
    const uint8_t* p = src + (size_t)y * sstep + (size_t)x * 3;
    dst[(size_t)y * dstep + x] = (uint8_t)((3735 * p[0] + 19235 * p[1] + 9798 * p[2] + 16384) >> 15);
}

__global__ void copy_gray_kernel(const uint8_t* __restrict__ src, size_t sstep, int w, int h, uint8_t* __restrict__ dst,
                                 size_t dstep)
{
    int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y;
    if (x >= w) return;
    dst[(size_t)y * dstep + x] = src[(size_t)y * sstep + x];
}

// horizontal pass: u8 -> u16 (sum of Q8 taps times 8-bit pixels <= 65280)
__global__ void blur_h_kernel(const uint8_t* __restrict__ src, size_t sstep, int w, int h, BlurTaps t,
                              uint16_t* __restrict__ dst)
{
    int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y;
    if (x >= w) return;
    const uint8_t* row = src + (size_t)y * sstep;
    unsigned s = 0;
    if (x >= 15 && x + 15 < w) {
        for (int j = t.lo; j <= t.hi; ++j) s += (unsigned)t.k[j] * row[x + j - 15];
    } else {
        for (int j = t.lo; j <= t.hi; ++j) s += (unsigned)t.k[j] * row[reflect101(x + j - 15, w)];
    }
    dst[(size_t)y * w + x] = (uint16_t)s;
}

// vertical pass: u16 -> u8 with the single final rounding
__global__ void blur_v_kernel(const uint16_t* __restrict__ src, int w, int h, BlurTaps t, uint8_t* __restrict__ dst,
                              size_t dstep)
{
    int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y;
    if (x >= w) return;
    unsigned s = 0;
    if (y >= 15 && y + 15 < h) {
        for (int i = t.lo; i <= t.hi; ++i) s += (unsigned)t.k[i] * src[(size_t)(y + i - 15) * w + x];
    } else {
        for (int i = t.lo; i <= t.hi; ++i) s += (unsigned)t.k[i] * src[(size_t)reflect101(y + i - 15, h) * w + x];
    }
    dst[(size_t)y * dstep + x] = (uint8_t)((s + (1u << 15)) >> 16);
}

// ---- packed fast paths: four pixels per thread, byte / halfword dot products (IDP.4A / IDP.2A) ----------------------
// R = tap radius rounded up to 4k+3 (window 2R+2 taps, the last one zero).  Same integer sums as the scalar kernels.
struct BlurPacked {
    uint32_t t4[8];    // horizontal: taps lo'..lo'+2R+1 as bytes, four per word (lo' = 15 - R)
    uint32_t t2[16];   // vertical: the same taps, two per word (low two bytes)
};
template <int R>
__device__ __forceinline__ void
blur_h4_body(const uint8_t* __restrict__ src, size_t sstep, int w, int h, BlurTaps t, BlurPacked pk, uint16_t* __restrict__ dst)
{
    constexpr int NT = (2 * R + 2) / 4;   // dot products per output
    const int x0 = (blockIdx.x * blockDim.x + threadIdx.x) * 4, y = blockIdx.y;
    if (x0 >= w) return;
    const uint8_t* row = src + (size_t)y * sstep;
    uint16_t* out = dst + (size_t)y * w + x0;
    if (x0 - R - 1 >= 0 && x0 + R + 5 <= w) {
        // words from the aligned byte x0 - R - 1; the window of output q starts q + 1 bytes into them
        const uint32_t* wp = (const uint32_t*)(row + x0 - R - 1);
        uint32_t W[NT + 1];   // bytes x0 - R - 1 .. x0 + R + 4
#pragma unroll
        for (int n = 0; n <= NT; ++n) W[n] = __ldg(wp + n);
        uint32_t s[4] = {0, 0, 0, 0};
#pragma unroll
        for (int q = 0; q < 4; ++q)
#pragma unroll
            for (int n = 0; n < NT; ++n) {
                const uint32_t v = q == 3 ? W[n + 1] : __funnelshift_r(W[n], W[n + 1], 8 * (q + 1));
                s[q] = __dp4a(v, pk.t4[n], s[q]);
            }
        *(uint2*)out = make_uint2(s[0] | (s[1] << 16), s[2] | (s[3] << 16));
    } else {
        for (int q = 0; q < 4 && x0 + q < w; ++q) {
            unsigned a = 0;
            for (int j = t.lo; j <= t.hi; ++j) a += (unsigned)t.k[j] * row[reflect101(x0 + q + j - 15, w)];
            out[q] = (uint16_t)a;
        }
    }
}

template <int R>
__device__ __forceinline__ void
blur_v4_body(const uint16_t* __restrict__ src, int w, int h, BlurTaps t, BlurPacked pk, uint8_t* __restrict__ dst, size_t dstep)
{
    const int x0 = (blockIdx.x * blockDim.x + threadIdx.x) * 4, y = blockIdx.y;
    if (x0 >= w) return;
    if (y - R >= 0 && y + R + 1 < h) {
        const uint16_t* col = src + (size_t)(y - R) * w + x0;
        uint32_t s[4] = {0, 0, 0, 0};
#pragma unroll
        for (int n = 0; n <= R; ++n) {  // rows y - R + 2n and y - R + 2n + 1 (the last one has a zero tap)
            const uint2 a = __ldg((const uint2*)(col + (size_t)(2 * n) * w)), b = __ldg((const uint2*)(col + (size_t)(2 * n + 1) * w));
            s[0] = __dp2a_lo(__byte_perm(a.x, b.x, 0x5410), pk.t2[n], s[0]);
            s[1] = __dp2a_lo(__byte_perm(a.x, b.x, 0x7632), pk.t2[n], s[1]);
            s[2] = __dp2a_lo(__byte_perm(a.y, b.y, 0x5410), pk.t2[n], s[2]);
            s[3] = __dp2a_lo(__byte_perm(a.y, b.y, 0x7632), pk.t2[n], s[3]);
        }
        uint32_t o = 0;
#pragma unroll
        for (int q = 0; q < 4; ++q) o |= ((s[q] + (1u << 15)) >> 16) << (8 * q);
        *(uint32_t*)(dst + (size_t)y * dstep + x0) = o;
    } else {
        for (int q = 0; q < 4; ++q) {
            unsigned a = 0;
            for (int i = t.lo; i <= t.hi; ++i) a += (unsigned)t.k[i] * src[(size_t)reflect101(y + i - 15, h) * w + x0 + q];
            dst[(size_t)y * dstep + x0 + q] = (uint8_t)((a + (1u << 15)) >> 16);
        }
    }
}

template <int R>
__global__ void __launch_bounds__(128)
blur_h4_kernel(const uint8_t* __restrict__ src, size_t sstep, int w, int h, BlurTaps t, BlurPacked pk, uint16_t* __restrict__ dst)
{
    blur_h4_body<R>(src, sstep, w, h, t, pk, dst);
}
template <int R>
__global__ void __launch_bounds__(128)
blur_v4_kernel(const uint16_t* __restrict__ src, int w, int h, BlurTaps t, BlurPacked pk, uint8_t* __restrict__ dst, size_t dstep)
{
    blur_v4_body<R>(src, w, h, t, pk, dst, dstep);
}
// batched: image blockIdx.z; the u16 intermediate of image z lives at tmp + z * w * h
struct BlurBatch {
    const uint8_t* src[32];
    uint8_t* dst[32];
};
template <int R>
__global__ void __launch_bounds__(128)
blur_h4_batch_kernel(BlurBatch b, int w, int h, BlurTaps t, BlurPacked pk, uint16_t* __restrict__ tmp)
{
    blur_h4_body<R>(b.src[blockIdx.z], (size_t)w, w, h, t, pk, tmp + (size_t)blockIdx.z * w * h);
}
template <int R>
__global__ void __launch_bounds__(128)
blur_v4_batch_kernel(BlurBatch b, int w, int h, BlurTaps t, BlurPacked pk, const uint16_t* __restrict__ tmp, size_t dstep)
{
    blur_v4_body<R>(tmp + (size_t)blockIdx.z * w * h, w, h, t, pk, b.dst[blockIdx.z], dstep);
}

template <int R>
static void blur_packed_launch(cudaStream_t s, const uint8_t* src, int w, int h, const BlurTaps& t, uint16_t* tmp, uint8_t* dst,
                               size_t dstep)
{
    BlurPacked pk;
    memset(&pk, 0, sizeof(pk));
    for (int i = 0; i < 2 * R + 1; ++i) {
        const uint32_t k = (uint32_t)t.k[15 - R + i];
        pk.t4[i / 4] |= k << (8 * (i % 4));
        pk.t2[i / 2] |= k << (8 * (i % 2));
    }
    dim3 grid(cdiv(w, 512), h);
    blur_h4_kernel<R><<<grid, 128, 0, s>>>(src, (size_t)w, w, h, t, pk, tmp);
    blur_v4_kernel<R><<<grid, 128, 0, s>>>(tmp, w, h, t, pk, dst, dstep);
}

// one blur pass (horizontal then vertical): packed kernels when the rows are 4-byte aligned and every tap fits a byte
static void blur_pass(cudaStream_t s, const uint8_t* src, int w, int h, const BlurTaps& t, uint16_t* tmp, uint8_t* dst, size_t dstep)
{
    bool packed = (w % 4) == 0 && (dstep % 4) == 0 && ((uintptr_t)src % 4) == 0 && ((uintptr_t)dst % 4) == 0 && ((uintptr_t)tmp % 8) == 0;
    for (int i = t.lo; i <= t.hi; ++i) packed = packed && t.k[i] >= 0 && t.k[i] <= 255;
    const int radius = 15 - t.lo;
    if (packed && radius <= 3) blur_packed_launch<3>(s, src, w, h, t, tmp, dst, dstep);
    else if (packed && radius <= 7) blur_packed_launch<7>(s, src, w, h, t, tmp, dst, dstep);
    else if (packed && radius <= 11) blur_packed_launch<11>(s, src, w, h, t, tmp, dst, dstep);
    else if (packed) blur_packed_launch<15>(s, src, w, h, t, tmp, dst, dstep);
    else {
        dim3 grid(cdiv(w, 128), h);
        blur_h_kernel<<<grid, 128, 0, s>>>(src, (size_t)w, w, h, t, tmp);
        blur_v_kernel<<<grid, 128, 0, s>>>(tmp, w, h, t, dst, dstep);
    }
}

// ---- batched form: the grey images of a whole batch of sensors in ONE launch per pass (blockIdx.z = image) -------------
// A 2-megapixel pass is a ~10 us kernel, mostly ramp and tail; 32 images per launch fill the machine.
#define PRESMOOTH_MAX_IMAGES 32
struct ImgBatch {
    const uint8_t* src[PRESMOOTH_MAX_IMAGES];
    uint8_t* dst[PRESMOOTH_MAX_IMAGES];
};
// BGR -> grey, four pixels per thread (three aligned words in, one word out); w % 4 == 0, rows 4-byte aligned
__global__ void __launch_bounds__(128)
bgr2gray4_batch_kernel(ImgBatch b, size_t sstep, int w, int h, size_t dstep)
{
    const int x0 = (blockIdx.x * blockDim.x + threadIdx.x) * 4, y = blockIdx.y;
    if (x0 >= w) return;
    const uint32_t* p = (const uint32_t*)(b.src[blockIdx.z] + (size_t)y * sstep + (size_t)x0 * 3);
    const uint32_t w0 = __ldg(p), w1 = __ldg(p + 1), w2 = __ldg(p + 2);   // B0 G0 R0 B1 | G1 R1 B2 G2 | R2 B3 G3 R3
    auto grey = [](uint32_t bb, uint32_t gg, uint32_t rr) { return (3735u * bb + 19235u * gg + 9798u * rr + 16384u) >> 15; };
    const uint32_t g0 = grey(w0 & 0xff, (w0 >> 8) & 0xff, (w0 >> 16) & 0xff);
    const uint32_t g1 = grey(w0 >> 24, w1 & 0xff, (w1 >> 8) & 0xff);
    const uint32_t g2 = grey((w1 >> 16) & 0xff, w1 >> 24, w2 & 0xff);
    const uint32_t g3 = grey((w2 >> 8) & 0xff, (w2 >> 16) & 0xff, w2 >> 24);
    *(uint32_t*)(b.dst[blockIdx.z] + (size_t)y * dstep + x0) = g0 | (g1 << 8) | (g2 << 16) | (g3 << 24);
}

int presmooth_dev(mvs_ctx* ctx, const char* tag, const uint8_t* src, int channels, int w, int h, size_t step,
                  double sigma, uint8_t* dst, size_t dstep)
{
    if (channels != 1 && channels != 3) return mvs_fail(ctx, MVS_ERR_ARG, "presmooth: channels must be 1 or 3");
    dim3 grid(cdiv(w, 128), h);
    cudaStream_t s = ctx->stream;
    bool blur = sigma > 0.0;
    std::string nm = std::string("presmooth.") + tag;
    MVS_ARENA(ctx, uint8_t, grey, (nm + ".grey").c_str(), (size_t)w * h);
    MVS_ARENA(ctx, uint16_t, tmp, (nm + ".tmp").c_str(), (size_t)w * h);
    uint8_t* g0 = blur ? grey : dst;
    size_t g0step = blur ? (size_t)w : dstep;
    if (channels == 3)
        bgr2gray_kernel<<<grid, 128, 0, s>>>(src, step, w, h, g0, g0step);
    else
        copy_gray_kernel<<<grid, 128, 0, s>>>(src, step, w, h, g0, g0step);
    MVS_LAUNCH_CHECK(ctx, "bgr2gray_kernel");
    if (!blur) return MVS_OK;
    BlurTaps t;
    if (gauss_taps_q8(sigma, &t)) return mvs_fail(ctx, MVS_ERR_ARG, "presmooth: bad sigma");
    // pass 1: grey -> grey (through tmp); pass 2: grey -> dst
    blur_pass(s, grey, w, h, t, tmp, grey, (size_t)w);
    blur_pass(s, grey, w, h, t, tmp, dst, dstep);
    MVS_LAUNCH_CHECK_N(ctx, "blur kernels", 4);
    return MVS_OK;
}

template <int R>
static void blur_batch_launch(cudaStream_t s, const BlurBatch& b, int n, int w, int h, const BlurTaps& t, uint16_t* tmp, size_t dstep)
{
    BlurPacked pk;
    memset(&pk, 0, sizeof(pk));
    for (int i = 0; i < 2 * R + 1; ++i) {
        const uint32_t k = (uint32_t)t.k[15 - R + i];
        pk.t4[i / 4] |= k << (8 * (i % 4));
        pk.t2[i / 2] |= k << (8 * (i % 2));
    }
    dim3 grid(cdiv(w, 512), h, n);
    blur_h4_batch_kernel<R><<<grid, 128, 0, s>>>(b, w, h, t, pk, tmp);
    blur_v4_batch_kernel<R><<<grid, 128, 0, s>>>(b, w, h, t, pk, tmp, dstep);
}

// n images of equal size (<= 32) through grey + double blur with one launch per pass; falls back to the per-image
// path when the packed kernels do not apply (grey input, odd widths, unaligned rows, no blur)
int presmooth_batch_dev(mvs_ctx* ctx, int n, const uint8_t* const* src, int channels, int w, int h, size_t step, double sigma,
                        uint8_t* const* dst, size_t dstep)
{
    BlurTaps t;
    memset(&t, 0, sizeof(t));
    if (n > PRESMOOTH_MAX_IMAGES) {  // larger batches go through in chunks
        for (int i0 = 0; i0 < n; i0 += PRESMOOTH_MAX_IMAGES) {
            int rc = presmooth_batch_dev(ctx, n - i0 < PRESMOOTH_MAX_IMAGES ? n - i0 : PRESMOOTH_MAX_IMAGES, src + i0, channels, w, h, step, sigma,
                                         dst + i0, dstep);
            if (rc) return rc;
        }
        return MVS_OK;
    }
    bool packed = channels == 3 && sigma > 0.0 && n >= 1 && n <= PRESMOOTH_MAX_IMAGES && (w % 4) == 0 && (step % 4) == 0 &&
                  (dstep % 4) == 0 && gauss_taps_q8(sigma, &t) == 0;
    for (int i = 0; packed && i < n; ++i) packed = ((uintptr_t)src[i] % 4) == 0 && ((uintptr_t)dst[i] % 4) == 0;
    for (int i = t.lo; packed && i <= t.hi; ++i) packed = t.k[i] >= 0 && t.k[i] <= 255;
    if (!packed) {
        for (int i = 0; i < n; ++i) {
            int rc = presmooth_dev(ctx, (i & 1) ? "r" : "l", src[i], channels, w, h, step, sigma, dst[i], dstep);
            if (rc) return rc;
        }
        return MVS_OK;
    }
    cudaStream_t s = ctx->stream;
    const size_t px = (size_t)w * h;
    MVS_ARENA(ctx, uint8_t, grey, "presmooth.batch.grey", px * n);
    MVS_ARENA(ctx, uint16_t, tmp, "presmooth.batch.tmp", px * n);
    ImgBatch g;
    BlurBatch p1, p2;
    for (int i = 0; i < n; ++i) {
        g.src[i] = src[i]; g.dst[i] = grey + px * i;
        p1.src[i] = grey + px * i; p1.dst[i] = grey + px * i;   // pass 1: grey -> grey (through tmp)
        p2.src[i] = grey + px * i; p2.dst[i] = dst[i];          // pass 2: grey -> dst
    }
    bgr2gray4_batch_kernel<<<dim3(cdiv(w, 512), h, n), 128, 0, s>>>(g, step, w, h, (size_t)w);
    const int radius = 15 - t.lo;
    for (int pass = 0; pass < 2; ++pass) {
        const BlurBatch& b = pass ? p2 : p1;
        const size_t ds = pass ? dstep : (size_t)w;
        if (radius <= 3) blur_batch_launch<3>(s, b, n, w, h, t, tmp, ds);
        else if (radius <= 7) blur_batch_launch<7>(s, b, n, w, h, t, tmp, ds);
        else if (radius <= 11) blur_batch_launch<11>(s, b, n, w, h, t, tmp, ds);
        else blur_batch_launch<15>(s, b, n, w, h, t, tmp, ds);
    }
    MVS_LAUNCH_CHECK_N(ctx, "presmooth batch kernels", 5);
    return MVS_OK;
}

extern "C" int mvs_presmooth(mvs_ctx* ctx, const uint8_t* src, int channels, int w, int h, size_t step, double sigma,
                             uint8_t* dst, size_t dstep)
{
    if (!ctx) return MVS_ERR_ARG;
    if (!src || !dst || w <= 0 || h <= 0 || (channels != 1 && channels != 3) || step < (size_t)w * channels ||
        dstep < (size_t)w)
        return mvs_fail(ctx, MVS_ERR_ARG, "mvs_presmooth: bad argument");
    MVS_CUDA(ctx, cudaSetDevice(ctx->device));
    MVS_ARENA(ctx, uint8_t, dsrc, "presmooth.api.src", (size_t)h * w * channels);
    MVS_ARENA(ctx, uint8_t, ddst, "presmooth.api.dst", (size_t)h * w);
    cudaStream_t s = ctx->stream;
    MVS_CUDA(ctx, cudaMemcpy2DAsync(dsrc, (size_t)w * channels, src, step, (size_t)w * channels, h,
                                    cudaMemcpyHostToDevice, s));
    int rc = presmooth_dev(ctx, "api", dsrc, channels, w, h, (size_t)w * channels, sigma, ddst, (size_t)w);
    if (rc) return rc;
    MVS_CUDA(ctx, cudaMemcpy2DAsync(dst, dstep, ddst, (size_t)w, (size_t)w, h, cudaMemcpyDeviceToHost, s));
    MVS_CUDA(ctx, cudaStreamSynchronize(s));
    return MVS_OK;
}

// ---- input side: input_images::Task::execute, pipeline/nodes/input_images.cpp:65-76 ----------------------
// equal = raw_image_equalize ? hist_equalize(raw) : raw (graphics/hist-equalize.cpp:9-30: BGR -> YCrCb, equalizeHist on Y,
// YCrCb -> BGR), grey = cvtColor(equal, BGR2GRAY); OpenCV's 8-bit fixed-point arithmetic, bit-exact.
__device__ __forceinline__ int sat8(int v) { return max(0, min(255, v)); }
__device__ __forceinline__ void bgr2ycrcb(int B, int G, int Rr, int& Y, int& Cr, int& Cb)
{
    Y = (Rr * 4899 + G * 9617 + B * 1868 + (1 << 13)) >> 14;
    Cr = sat8(((Rr - Y) * 11682 + 128 * (1 << 14) + (1 << 13)) >> 14);
    Cb = sat8(((B - Y) * 9241 + 128 * (1 << 14) + (1 << 13)) >> 14);
}
__global__ void __launch_bounds__(256)
luma_hist_kernel(const uint8_t* __restrict__ src, size_t sstep, int w, int h, unsigned* __restrict__ hist)
{
    __shared__ unsigned sh[256];
    sh[threadIdx.x] = 0;
    __syncthreads();
    for (int y = blockIdx.x; y < h; y += gridDim.x) {
        const uint8_t* row = src + (size_t)y * sstep;
        for (int x = threadIdx.x; x < w; x += 256) {
            int Y, Cr, Cb;
            bgr2ycrcb(row[3 * x], row[3 * x + 1], row[3 * x + 2], Y, Cr, Cb);
            atomicAdd(&sh[Y], 1u);
        }
    }
    __syncthreads();
    if (sh[threadIdx.x]) atomicAdd(&hist[threadIdx.x], sh[threadIdx.x]);
}
// cv::equalizeHist look-up table; one thread (256 sequential steps on integers)
__global__ void equalize_lut_kernel(const unsigned* __restrict__ hist, unsigned total, uint8_t* __restrict__ lut)
{
    if (threadIdx.x || blockIdx.x) return;
    int i = 0;
    while (i < 256 && !hist[i]) ++i;
    if (i == 256 || hist[i] == total) {
        for (int k = 0; k < 256; ++k) lut[k] = (uint8_t)(i < 256 ? i : 0);
        return;
    }
    const float scale = 255.f / (float)(total - hist[i]);
    unsigned sum = 0;
    for (int k = 0; k <= i; ++k) lut[k] = 0;
    for (++i; i < 256; ++i) {
        sum += hist[i];
        lut[i] = (uint8_t)sat8(__float2int_rn((float)sum * scale));
    }
}
__global__ void __launch_bounds__(128)
equalize_apply_kernel(const uint8_t* __restrict__ src, size_t sstep, int w, int h, const uint8_t* __restrict__ lut, int equalize,
                      uint8_t* __restrict__ equal, size_t estep, uint8_t* __restrict__ grey, size_t gstep)
{
    const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y;
    if (x >= w) return;
    const uint8_t* p = src + (size_t)y * sstep + (size_t)x * 3;
    int b = p[0], g = p[1], r = p[2];
    if (equalize) {
        int Y, Cr, Cb;
        bgr2ycrcb(b, g, r, Y, Cr, Cb);
        Y = lut[Y];
        b = sat8(Y + (((Cb - 128) * 29049 + (1 << 13)) >> 14));
        g = sat8(Y + (((Cb - 128) * -5636 + (Cr - 128) * -11698 + (1 << 13)) >> 14));
        r = sat8(Y + (((Cr - 128) * 22987 + (1 << 13)) >> 14));
    }
    if (equal) {
        uint8_t* e = equal + (size_t)y * estep + (size_t)x * 3;
        e[0] = (uint8_t)b; e[1] = (uint8_t)g; e[2] = (uint8_t)r;
    }
    if (grey) grey[(size_t)y * gstep + x] = (uint8_t)((3735 * b + 19235 * g + 9798 * r + 16384) >> 15);
}

int input_images_dev(mvs_ctx* ctx, const uint8_t* raw, int w, int h, size_t sstep, int equalize, uint8_t* equal, size_t estep,
                     uint8_t* grey, size_t gstep)
{
    cudaStream_t s = ctx->stream;
    MVS_ARENA(ctx, unsigned, hist, "input.hist", 256);
    MVS_ARENA(ctx, uint8_t, lut, "input.lut", 256);
    if (equalize) {
        MVS_CUDA(ctx, cudaMemsetAsync(hist, 0, 256 * sizeof(unsigned), s));
        luma_hist_kernel<<<min(h, 592), 256, 0, s>>>(raw, sstep, w, h, hist);
        equalize_lut_kernel<<<1, 32, 0, s>>>(hist, (unsigned)((size_t)w * h), lut);
        MVS_LAUNCH_CHECK_N(ctx, "equalize kernels", 2);
    }
    equalize_apply_kernel<<<dim3(cdiv(w, 128), h), 128, 0, s>>>(raw, sstep, w, h, lut, equalize, equal, estep, grey, gstep);
    MVS_LAUNCH_CHECK(ctx, "equalize_apply_kernel");
    return MVS_OK;
}

extern "C" int mvs_input_images(mvs_ctx* ctx, const uint8_t* raw, int w, int h, size_t step, int raw_image_equalize,
                                uint8_t* equal, size_t estep, uint8_t* grey, size_t gstep)
{
    if (!ctx) return MVS_ERR_ARG;
    if (!raw || w <= 0 || h <= 0 || step < (size_t)w * 3 || (equal && estep < (size_t)w * 3) || (grey && gstep < (size_t)w) ||
        (size_t)w * h > 0xffffffffull)
        return mvs_fail(ctx, MVS_ERR_ARG, "mvs_input_images: bad argument");
    MVS_CUDA(ctx, cudaSetDevice(ctx->device));
    cudaStream_t s = ctx->stream;
    MVS_ARENA(ctx, uint8_t, dsrc, "input.api.src", (size_t)h * w * 3);
    MVS_ARENA(ctx, uint8_t, deq, "input.api.equal", (size_t)h * w * 3);
    MVS_ARENA(ctx, uint8_t, dgr, "input.api.grey", (size_t)h * w);
    MVS_CUDA(ctx, cudaMemcpy2DAsync(dsrc, (size_t)w * 3, raw, step, (size_t)w * 3, h, cudaMemcpyHostToDevice, s));
    int rc = input_images_dev(ctx, dsrc, w, h, (size_t)w * 3, raw_image_equalize, equal ? deq : nullptr, (size_t)w * 3,
                              grey ? dgr : nullptr, (size_t)w);
    if (rc) return rc;
    if (equal) MVS_CUDA(ctx, cudaMemcpy2DAsync(equal, estep, deq, (size_t)w * 3, (size_t)w * 3, h, cudaMemcpyDeviceToHost, s));
    if (grey) MVS_CUDA(ctx, cudaMemcpy2DAsync(grey, gstep, dgr, (size_t)w, (size_t)w, h, cudaMemcpyDeviceToHost, s));
    MVS_CUDA(ctx, cudaStreamSynchronize(s));
    return MVS_OK;
}

extern "C" int mvs_input_images_dev(mvs_ctx* ctx, const uint8_t* raw_dev, int w, int h, size_t step, int raw_image_equalize,
                                    uint8_t* equal_dev, size_t estep, uint8_t* grey_dev, size_t gstep)
{
    if (!ctx) return MVS_ERR_ARG;
    if (!raw_dev || w <= 0 || h <= 0 || step < (size_t)w * 3 || (equal_dev && estep < (size_t)w * 3) ||
        (grey_dev && gstep < (size_t)w) || (size_t)w * h > 0xffffffffull)
        return mvs_fail(ctx, MVS_ERR_ARG, "mvs_input_images_dev: bad argument");
    MVS_CUDA(ctx, cudaSetDevice(ctx->device));
    return input_images_dev(ctx, raw_dev, w, h, step, raw_image_equalize, equal_dev, estep, grey_dev, gstep);
}
