// a5: the reference's CUDA_SGM strategy -- SGMDisparity::calculate, cost-functions/disparity/strategies/sgm_disp.cpp:102-211,
// which runs libSGM: census 9x7 -> Hamming cost -> 8-path aggregation (P1 = 20, P2 = 100) -> left / right winner with a
// 0.95 uniqueness test -> 3x3 median -> left-right check.  The algorithm follows the vendored sources
// (multiview/contrib/libSGM/src: census.cu:27-152, matching_cost.cu:32-169, scan_scost.cu:19-20,63-330,
// winner_takes_all.cu:32-233, median_filter.cu:24-32, check_consistency.cu:20-35); nothing of their structure is kept:
// the cost volume is written in this library's 16-bit split layout and aggregated by the same wavefront sweeps as the
// StereoSGBM path (aggregate_wave.cu).  Parity is UNPINNED (oracle/census_oracle.c says why and lists the four places
// where the vendored kernels' behaviour is an artefact and the plain reading is taken instead).
#include "common.cuh"

// 9x7 census: bit = (centre > neighbour), row-major over the window without the centre; border pixels get 0
template <typename T>
__global__ void __launch_bounds__(256)
census_kernel(const T* __restrict__ img, int w, int h, unsigned long long* __restrict__ out)
{
    const int x = blockIdx.x * 32 + (threadIdx.x & 31), y = blockIdx.y * 8 + (threadIdx.x >> 5);
    if (x >= w || y >= h) return;
    unsigned long long v = 0;
    if (y >= 3 && y < h - 3 && x >= 4 && x < w - 4) {
        const unsigned c = img[(size_t)y * w + x];
#pragma unroll
        for (int dy = -3; dy <= 3; ++dy)
#pragma unroll
            for (int dx = -4; dx <= 4; ++dx) {
                if (dy == 0 && dx == 0) continue;
                v = (v << 1) | (unsigned long long)(c > (unsigned)img[(size_t)(y + dy) * w + x + dx]);
            }
    }
    out[(size_t)y * w + x] = v;
}

// Hamming cost into the split-layout volume [h][w][Dp] (u16): a warp per pixel, lane l owns words l*NW .. l*NW+NW-1,
// word k holds disparity k (low half) and k + Dp/2 (high half); cost(x, d) = popc(L(x) ^ (x < d ? 0 : R(x - d)))
template <int NW>
__global__ void __launch_bounds__(256)
census_cost_kernel(const unsigned long long* __restrict__ cl, const unsigned long long* __restrict__ cr, int w, int h,
                   uint32_t* __restrict__ C)
{
    const int lane = threadIdx.x & 31;
    const long long pix = (long long)blockIdx.x * 8 + (threadIdx.x >> 5);
    if (pix >= (long long)w * h) return;
    const int y = (int)(pix / w), x = (int)(pix - (long long)y * w);
    const unsigned long long L = cl[pix];
    const unsigned long long* R = cr + (size_t)y * w;
    uint32_t word[NW];
#pragma unroll
    for (int j = 0; j < NW; ++j) {
        const int d0 = lane * NW + j, d1 = d0 + NW * 32;
        const uint32_t c0 = __popcll(L ^ (x < d0 ? 0ull : R[x - d0])), c1 = __popcll(L ^ (x < d1 ? 0ull : R[x - d1]));
        word[j] = c0 | (c1 << 16);
    }
    uint32_t* o = C + (size_t)pix * (NW * 32) + lane * NW;
    if (NW == 4) *(uint4*)o = make_uint4(word[0], word[1 % NW], word[2 % NW], word[3 % NW]);
    else if (NW == 2) *(uint2*)o = make_uint2(word[0], word[1 % NW]);
    else o[0] = word[0];
}

// Left and right winner of a pixel (a warp per pixel).  Left: the first minimum of S(x, .) over all d; its rival is the
// first minimum over d != winner with x + d < w.  Right (the pixel seen from the other image): candidates S(x + d, d)
// with x + d < w, winner and rival likewise.  A winner is dropped (0) when rival * 0.95 < winner and the two are more
// than one disparity apart; otherwise the output is d + 1.
template <int NW>
__global__ void __launch_bounds__(256)
census_wta_kernel(const uint16_t* __restrict__ S, int w, int h, uint16_t* __restrict__ dl, uint16_t* __restrict__ dr)
{
    constexpr int HALF = NW * 32, DP = 2 * HALF;
    const int lane = threadIdx.x & 31;
    const long long pix = (long long)blockIdx.x * 8 + (threadIdx.x >> 5);
    if (pix >= (long long)w * h) return;
    const int y = (int)(pix / w), x = (int)(pix - (long long)y * w);
    const uint32_t* Sw = (const uint32_t*)(S + (size_t)pix * DP) + lane * NW;
    uint32_t kl[2 * NW], kr[2 * NW];   // (cost << 16 | d) of this lane's disparities, left and right view
#pragma unroll
    for (int j = 0; j < NW; ++j) {
        const uint32_t v = __ldg(Sw + j);
        const int d0 = lane * NW + j, d1 = d0 + HALF;
        kl[2 * j] = (v << 16) | (uint32_t)d0;
        kl[2 * j + 1] = (v & 0xffff0000u) | (uint32_t)d1;
        // right view: S(x + d, d); halfword of d in the split layout: 2 * (d mod HALF) + (d >= HALF)
        kr[2 * j] = x + d0 < w ? ((uint32_t)__ldg(S + (size_t)(pix + d0) * DP + 2 * d0) << 16) | (uint32_t)d0 : 0x7fffffffu;
        kr[2 * j + 1] = x + d1 < w ? ((uint32_t)__ldg(S + (size_t)(pix + d1) * DP + 2 * d0 + 1) << 16) | (uint32_t)d1 : 0x7fffffffu;
    }
    uint32_t m = 0xffffffffu;
#pragma unroll
    for (int k = 0; k < 2 * NW; ++k) m = min(m, kl[k]);
    const uint32_t L1 = __reduce_min_sync(0xffffffffu, m);
    const int dL1 = (int)(L1 & 0xffffu), cL1 = (int)(L1 >> 16);
    m = 0x7fffffffu;
#pragma unroll
    for (int k = 0; k < 2 * NW; ++k) {
        const int d = (int)(kl[k] & 0xffffu);
        if (d != dL1 && x + d < w) m = min(m, kl[k]);
    }
    const uint32_t L2 = __reduce_min_sync(0xffffffffu, m);
    const int cL2 = (int)(L2 >> 16), dL2 = (L2 & 0xffffu) == 0xffffu ? -1 : (int)(L2 & 0xffffu);
    m = 0x7fffffffu;
#pragma unroll
    for (int k = 0; k < 2 * NW; ++k) m = min(m, kr[k]);
    const uint32_t R1 = __reduce_min_sync(0xffffffffu, m);
    const int cR1 = (int)(R1 >> 16), dR1 = (R1 & 0xffffu) == 0xffffu ? -1 : (int)(R1 & 0xffffu);
    m = 0x7fffffffu;
#pragma unroll
    for (int k = 0; k < 2 * NW; ++k)
        if (kr[k] != 0x7fffffffu && (int)(kr[k] & 0xffffu) != dR1) m = min(m, kr[k]);
    const uint32_t R2 = __reduce_min_sync(0xffffffffu, m);
    const int cR2 = (int)(R2 >> 16), dR2 = (R2 & 0xffffu) == 0xffffu ? -1 : (int)(R2 & 0xffffu);
    if (lane == 0) {
        const float uniq = 0.95f;
        dl[pix] = ((float)cL2 * uniq < (float)cL1 && abs(dL1 - dL2) > 1) ? 0 : (uint16_t)(dL1 + 1);
        dr[pix] = ((float)cR2 * uniq < (float)cR1 && abs(dR1 - dR2) > 1) ? 0 : (uint16_t)(dR1 + 1);
    }
}

// 3x3 median with the window anchored at its top-left pixel (NPP anchor (0, 0)), right / bottom edge replicated;
// z = 0: left disparities, z = 1: right disparities (consecutive images)
__global__ void census_median_kernel(const uint16_t* __restrict__ src, int w, int h, uint16_t* __restrict__ dst)
{
    const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y;
    if (x >= w) return;
    src += (size_t)blockIdx.z * w * h;
    dst += (size_t)blockIdx.z * w * h;
    int v[9];
#pragma unroll
    for (int j = 0; j < 3; ++j)
#pragma unroll
        for (int i = 0; i < 3; ++i) v[j * 3 + i] = src[(size_t)min(y + j, h - 1) * w + min(x + i, w - 1)];
#define CSW(a, b) { int t_ = min(v[a], v[b]); v[b] = max(v[a], v[b]); v[a] = t_; }
    CSW(1, 2) CSW(4, 5) CSW(7, 8) CSW(0, 1) CSW(3, 4) CSW(6, 7) CSW(1, 2) CSW(4, 5) CSW(7, 8) CSW(0, 3) CSW(5, 8) CSW(4, 7)
    CSW(3, 6) CSW(1, 4) CSW(2, 5) CSW(4, 7) CSW(4, 2) CSW(6, 4) CSW(4, 2)
#undef CSW
    dst[(size_t)y * w + x] = (uint16_t)v[4];
}

// left-right check on the filtered images (the disparities still carry their +1, as in the vendored kernel), masking by
// source pixels that are 0, then the float conversion of sgm_disp.cpp:195-207
template <typename T>
__global__ void census_check_kernel(const uint16_t* __restrict__ ml, const uint16_t* __restrict__ mr, const T* __restrict__ left, int w,
                                    int h, int D, float* __restrict__ out, uint16_t* __restrict__ disp_tap)
{
    const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y;
    if (x >= w) return;
    const size_t i = (size_t)y * w + x;
    int d = ml[i];
    const int k = x - d;
    if (left[i] == 0 || d <= 0 || (k >= 0 && k < w && abs((int)mr[(size_t)y * w + k] - d) > 1)) d = 0;
    if (disp_tap) disp_tap[i] = (uint16_t)d;
    const float v = (float)(D == 256 ? d : (d & 0xff));   // 8-bit output for 64 / 128 disparities
    out[i] = (D == 256 && v > 512.0f) ? 0.0f : v;
}

template <typename T>
static int census_chain(mvs_ctx* ctx, const T* dl_img, const T* dr_img, int w, int h, int D, float* out_dev, uint16_t* disp_tap_dev)
{
    cudaStream_t s = ctx->stream;
    const size_t n = (size_t)w * h;
    SgbmGeom g;
    memset(&g, 0, sizeof(g));
    // every column has a cost for every disparity (an absent right pixel counts as descriptor 0): the volume is w wide
    g.W = w; g.H = h; g.D = D; g.Dp = D; g.W1 = w; g.minD = 0; g.minX1 = 0;
    g.P1 = 20; g.P2 = 100;   // scan_scost.cu:19-20
    g.mode = MVS_MODE_HH;    // all eight directions
    MVS_ARENA(ctx, unsigned long long, cen, "census.desc", n * 2);
    int16_t* C = sgbm_volume(ctx, "census.C", g);
    int16_t* S = sgbm_volume(ctx, "census.S", g);
    MVS_ARENA(ctx, uint16_t, disp, "census.disp", n * 4);   // left, right, filtered left, filtered right
    if (!C || !S) return MVS_ERR_NOMEM;
    const dim3 g2(cdiv(w, 32), cdiv(h, 8));
    census_kernel<T><<<g2, 256, 0, s>>>(dl_img, w, h, cen);
    census_kernel<T><<<g2, 256, 0, s>>>(dr_img, w, h, cen + n);
    const int pb = cdiv((long long)n, 8);
    switch (D / 64) {
    case 1: census_cost_kernel<1><<<pb, 256, 0, s>>>(cen, cen + n, w, h, (uint32_t*)C); break;
    case 2: census_cost_kernel<2><<<pb, 256, 0, s>>>(cen, cen + n, w, h, (uint32_t*)C); break;
    default: census_cost_kernel<4><<<pb, 256, 0, s>>>(cen, cen + n, w, h, (uint32_t*)C); break;
    }
    MVS_LAUNCH_CHECK_N(ctx, "census kernels", 3);
    const int16_t* Cc = C;
    int rc = sgbm_aggregate_batch_dev(ctx, g, 1, &Cc, &S, nullptr);
    if (rc) return rc;
    switch (D / 64) {
    case 1: census_wta_kernel<1><<<pb, 256, 0, s>>>((const uint16_t*)S, w, h, disp, disp + n); break;
    case 2: census_wta_kernel<2><<<pb, 256, 0, s>>>((const uint16_t*)S, w, h, disp, disp + n); break;
    default: census_wta_kernel<4><<<pb, 256, 0, s>>>((const uint16_t*)S, w, h, disp, disp + n); break;
    }
    census_median_kernel<<<dim3(cdiv(w, 128), h, 2), 128, 0, s>>>(disp, w, h, disp + 2 * n);
    census_check_kernel<T><<<dim3(cdiv(w, 128), h), 128, 0, s>>>(disp + 2 * n, disp + 3 * n, dl_img, w, h, D, out_dev, disp_tap_dev);
    MVS_LAUNCH_CHECK_N(ctx, "census winner / median / check kernels", 3);
    return MVS_OK;
}

extern "C" int mvs_disparity_census(mvs_ctx* ctx, const void* l, const void* r, int depth_bits, int w, int h, size_t step,
                                    int disparity_size, float* out, size_t ostep, uint16_t* disp_tap)
{
    if (!ctx) return MVS_ERR_ARG;
    const size_t bpp = depth_bits == 16 ? 2 : 1;
    if (!l || !r || !out || (depth_bits != 8 && depth_bits != 16) || w <= 0 || h <= 0 || step < (size_t)w * bpp || ostep < (size_t)w * 4)
        return mvs_fail(ctx, MVS_ERR_ARG, "mvs_disparity_census: bad argument");
    // sgm_disp.cpp:80-93: disparity_size must be 64, 128 or 256; libSGM wants even sizes (stereo_sgm.cpp:130-133)
    if ((disparity_size != 64 && disparity_size != 128 && disparity_size != 256) || (w & 1) || (h & 1) || w <= 8 || h <= 6)
        return mvs_fail(ctx, MVS_ERR_UNSUPPORTED, "mvs_disparity_census: disparity_size must be 64, 128 or 256 and the image size even");
    MVS_CUDA(ctx, cudaSetDevice(ctx->device));
    int rc = main_stream_waits_for_tail(ctx);
    if (rc) return rc;
    cudaStream_t s = ctx->stream;
    const size_t n = (size_t)w * h;
    MVS_ARENA(ctx, uint8_t, dimg, "census.api.img", n * bpp * 2);
    MVS_ARENA(ctx, float, dout, "census.api.out", n);
    MVS_ARENA(ctx, uint16_t, dtap, "census.api.tap", n);
    MVS_CUDA(ctx, cudaMemcpy2DAsync(dimg, (size_t)w * bpp, l, step, (size_t)w * bpp, h, cudaMemcpyHostToDevice, s));
    MVS_CUDA(ctx, cudaMemcpy2DAsync(dimg + n * bpp, (size_t)w * bpp, r, step, (size_t)w * bpp, h, cudaMemcpyHostToDevice, s));
    if (depth_bits == 16) rc = census_chain<uint16_t>(ctx, (const uint16_t*)dimg, (const uint16_t*)dimg + n, w, h, disparity_size, dout, dtap);
    else rc = census_chain<uint8_t>(ctx, dimg, dimg + n, w, h, disparity_size, dout, dtap);
    if (rc) return rc;
    MVS_CUDA(ctx, cudaMemcpy2DAsync(out, ostep, dout, (size_t)w * 4, (size_t)w * 4, h, cudaMemcpyDeviceToHost, s));
    if (disp_tap) MVS_CUDA(ctx, cudaMemcpyAsync(disp_tap, dtap, n * 2, cudaMemcpyDeviceToHost, s));
    MVS_CUDA(ctx, cudaStreamSynchronize(s));
    return sgbm_aggregate_wave_status(ctx);
}
