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
    blur_h_kernel<<<grid, 128, 0, s>>>(grey, (size_t)w, w, h, t, tmp);
    blur_v_kernel<<<grid, 128, 0, s>>>(tmp, w, h, t, grey, (size_t)w);
    blur_h_kernel<<<grid, 128, 0, s>>>(grey, (size_t)w, w, h, t, tmp);
    blur_v_kernel<<<grid, 128, 0, s>>>(tmp, w, h, t, dst, dstep);
    MVS_LAUNCH_CHECK_N(ctx, "blur kernels", 4);
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
