// a3 (i)-(iii): matching cost volume C[y][x][d] of cv::StereoSGBM (reference call site
// cost-functions/disparity/strategies/stereosgbm_disp.cpp:121-125; algorithm: SURVEY.md A.1):
//   x-Sobel prefilter clipped to [0, 2*ftzero] + raw intensity, Birchfield-Tomasi sampling-insensitive
//   cost on both channels, then a blockSize x blockSize box sum with replicate clamping in the
//   valid-column (W1) domain.  16-bit, d fastest, Dp = D rounded up to 64 per pixel (pad = 0).
#include "common.cuh"

extern "C" int mvs_sgbm_check(const mvs_sgbm_params* p, int w, int h)
{
    if (!p || w <= 0 || h <= 0) return MVS_ERR_ARG;
    if (p->num_disp <= 0 || p->num_disp % 16 || p->num_disp > 256) return MVS_ERR_UNSUPPORTED;
    if (p->min_disp < 0) return MVS_ERR_UNSUPPORTED;
    if (p->block < 1 || !(p->block & 1) || p->block > 13) return MVS_ERR_UNSUPPORTED;
    if (p->mode != MVS_MODE_SGBM && p->mode != MVS_MODE_HH && p->mode != MVS_MODE_HH4) return MVS_ERR_UNSUPPORTED;
    int ftzero = (p->prefilter_cap > 15 ? p->prefilter_cap : 15) | 1;
    if (ftzero > 127) return MVS_ERR_UNSUPPORTED;  // prefiltered values are kept in 8 bits
    int P1 = p->p1 > 0 ? p->p1 : 2;
    int P2 = p->p2 > 0 ? p->p2 : 5;
    if (P2 < P1 + 1) P2 = P1 + 1;
    long long cmax = (long long)p->block * p->block * (2 * ftzero + 63);
    if (cmax + P2 > 32767) return MVS_ERR_UNSUPPORTED;  // 16-bit path costs could wrap
    if (w - (p->min_disp + p->num_disp) <= 0) return MVS_ERR_ARG;  // no valid column
    if (w > 65534) return MVS_ERR_UNSUPPORTED;
    return MVS_OK;
}

int sgbm_geom(mvs_ctx* ctx, const mvs_sgbm_params* p, int w, int h, SgbmGeom* g)
{
    int rc = mvs_sgbm_check(p, w, h);
    if (rc)
        return mvs_fail(ctx, rc,
                        "SGBM parameters outside the supported bit-exact domain (num_disp multiple of 16 <= 256, "
                        "min_disp >= 0, odd block <= 13 with block^2*(2*ftzero+63)+P2 <= 32767, mode 0|1|3, "
                        "width > min_disp+num_disp)");
    g->W = w; g->H = h; g->D = p->num_disp; g->minD = p->min_disp;
    g->minX1 = p->min_disp + p->num_disp;
    g->W1 = w - g->minX1;
    g->Dp = (g->D + 63) / 64 * 64;
    g->SW2 = g->SH2 = p->block / 2;
    g->ftzero = (p->prefilter_cap > 15 ? p->prefilter_cap : 15) | 1;
    g->P1 = p->p1 > 0 ? p->p1 : 2;
    g->P2 = p->p2 > 0 ? p->p2 : 5;
    if (g->P2 < g->P1 + 1) g->P2 = g->P1 + 1;
    g->ur = p->uniqueness >= 0 ? p->uniqueness : 10;
    g->d12 = p->disp12_max_diff > 0 ? p->disp12_max_diff : 1;
    g->INV = (p->min_disp - 1) * 16;
    g->mode = p->mode;
    return MVS_OK;
}

// ---- prefilter: 6 planes per image: p, lo(p), hi(p), q, lo(q), hi(q) ------------------------------------
__device__ __forceinline__ int pre_p(const uint8_t* __restrict__ r, const uint8_t* __restrict__ n,
                                     const uint8_t* __restrict__ s, int x, int W, int ftzero)
{
    if (x <= 0 || x >= W - 1) return ftzero;
    int v = 2 * ((int)r[x + 1] - (int)r[x - 1]) + ((int)n[x + 1] - (int)n[x - 1]) + ((int)s[x + 1] - (int)s[x - 1]);
    return max(-ftzero, min(ftzero, v)) + ftzero;
}
__device__ __forceinline__ int pre_q(const uint8_t* __restrict__ r, int x, int W, int ftzero)
{
    return (x <= 0 || x >= W - 1) ? ftzero : (int)r[x];
}

__global__ void prefilter_kernel(const uint8_t* __restrict__ img, int W, int H, int ftzero, uint8_t* __restrict__ planes)
{
    int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y;
    if (x >= W) return;
    const uint8_t* r = img + (size_t)y * W;
    const uint8_t* n = y > 0 ? r - W : r;
    const uint8_t* s = y < H - 1 ? r + W : r;
    size_t plane = (size_t)W * H, o = (size_t)y * W + x;
    int pc = pre_p(r, n, s, x, W, ftzero), qc = pre_q(r, x, W, ftzero);
    int pl = pc, pr = pc, ql = qc, qr = qc;
    if (x > 0) { pl = (pc + pre_p(r, n, s, x - 1, W, ftzero)) >> 1; ql = (qc + pre_q(r, x - 1, W, ftzero)) >> 1; }
    if (x < W - 1) { pr = (pc + pre_p(r, n, s, x + 1, W, ftzero)) >> 1; qr = (qc + pre_q(r, x + 1, W, ftzero)) >> 1; }
    planes[o] = (uint8_t)pc;
    planes[plane + o] = (uint8_t)min(pc, min(pl, pr));
    planes[2 * plane + o] = (uint8_t)max(pc, max(pl, pr));
    planes[3 * plane + o] = (uint8_t)qc;
    planes[4 * plane + o] = (uint8_t)min(qc, min(ql, qr));
    planes[5 * plane + o] = (uint8_t)max(qc, max(ql, qr));
}

// ---- per-pixel BT cost + horizontal box sum ------------------------------------------------------------------
// One block: row y, TX output columns (W1 coordinates), all disparities.
#define COST_TX 32
__global__ void __launch_bounds__(256)
pixcost_hsum_kernel(const uint8_t* __restrict__ PL, const uint8_t* __restrict__ PR, SgbmGeom g,
                    uint16_t* __restrict__ hsum)
{
    extern __shared__ uint8_t pix[];  // [(TX + 2*SW2)][Dp]
    const int y = blockIdx.y, x0 = blockIdx.x * COST_TX;
    const int nx = COST_TX + 2 * g.SW2;
    const size_t plane = (size_t)g.W * g.H;
    const uint8_t* L = PL + (size_t)y * g.W;
    const uint8_t* R = PR + (size_t)y * g.W;
    for (int idx = threadIdx.x; idx < nx * g.Dp; idx += blockDim.x) {
        int xi = idx / g.Dp, d = idx - xi * g.Dp;
        int v = 0;
        if (d < g.D) {
            int x1 = min(max(x0 + xi - g.SW2, 0), g.W1 - 1);
            int x = x1 + g.minX1, x2 = x - (d + g.minD);
            int u = L[x], u0 = L[plane + x], u1 = L[2 * plane + x];
            int vv = R[x2], v0 = R[plane + x2], v1 = R[2 * plane + x2];
            int c0 = max(0, max(u - v1, v0 - u)), c1 = max(0, max(vv - u1, u0 - vv));
            v = min(c0, c1);
            u = L[3 * plane + x]; u0 = L[4 * plane + x]; u1 = L[5 * plane + x];
            vv = R[3 * plane + x2]; v0 = R[4 * plane + x2]; v1 = R[5 * plane + x2];
            c0 = max(0, max(u - v1, v0 - u)); c1 = max(0, max(vv - u1, u0 - vv));
            v += min(c0, c1) >> 2;
        }
        pix[idx] = (uint8_t)v;
    }
    __syncthreads();
    const int half = g.Dp >> 1;
    const int win = 2 * g.SW2 + 1;
    for (int idx = threadIdx.x; idx < COST_TX * half; idx += blockDim.x) {
        int xo = idx / half, dp = idx - xo * half;
        if (x0 + xo >= g.W1) continue;
        const uint8_t* p = pix + (size_t)xo * g.Dp + 2 * dp;
        unsigned lo = 0, hi = 0;
        for (int k = 0; k < win; ++k) {
            lo += p[(size_t)k * g.Dp];
            hi += p[(size_t)k * g.Dp + 1];
        }
        *(uint32_t*)(hsum + ((size_t)y * g.W1 + x0 + xo) * g.Dp + 2 * dp) = lo | (hi << 16);
    }
}

// ---- vertical box sum: C[y] = sum_{dy} hsum[clamp(y+dy)] ; packed u16x2, sliding along y -------------------
#define VSUM_ROWS 24
__global__ void __launch_bounds__(256)
vsum_kernel(const uint32_t* __restrict__ hsum, SgbmGeom g, uint32_t* __restrict__ C)
{
    const size_t rowwords = (size_t)g.W1 * (g.Dp >> 1);
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= rowwords) return;
    int y0 = blockIdx.y * VSUM_ROWS, y1 = min(y0 + VSUM_ROWS, g.H);
    int zero_from = (g.mode == MVS_MODE_HH4) ? g.H - g.SH2 : g.H;  // OpenCV 4.13 leaves these C rows zero in HH4
    uint32_t acc = 0;
    for (int dy = -g.SH2; dy <= g.SH2; ++dy) acc += __ldg(hsum + (size_t)min(max(y0 + dy, 0), g.H - 1) * rowwords + i);
    for (int y = y0; y < y1; ++y) {
        C[(size_t)y * rowwords + i] = (y >= zero_from) ? 0u : acc;
        uint32_t sub = __ldg(hsum + (size_t)min(max(y - g.SH2, 0), g.H - 1) * rowwords + i);
        uint32_t add = __ldg(hsum + (size_t)min(max(y + 1 + g.SH2, 0), g.H - 1) * rowwords + i);
        acc = acc - sub + add;  // component-wise: no borrow (acc >= sub) and no carry (sums <= 29237)
    }
}

int sgbm_cost_volume_dev(mvs_ctx* ctx, const SgbmGeom& g, const uint8_t* l, const uint8_t* r, int16_t* C)
{
    cudaStream_t s = ctx->stream;
    size_t plane = (size_t)g.W * g.H;
    MVS_ARENA(ctx, uint8_t, planes, "sgbm.planes", plane * 12);
    MVS_ARENA(ctx, uint16_t, hsum, "sgbm.hsum", g.cells());
    dim3 pg(cdiv(g.W, 128), g.H);
    prefilter_kernel<<<pg, 128, 0, s>>>(l, g.W, g.H, g.ftzero, planes);
    prefilter_kernel<<<pg, 128, 0, s>>>(r, g.W, g.H, g.ftzero, planes + 6 * plane);
    MVS_LAUNCH_CHECK_N(ctx, "prefilter_kernel", 2);
    dim3 cg(cdiv(g.W1, COST_TX), g.H);
    size_t smem = (size_t)(COST_TX + 2 * g.SW2) * g.Dp;
    pixcost_hsum_kernel<<<cg, 256, smem, s>>>(planes, planes + 6 * plane, g, hsum);
    MVS_LAUNCH_CHECK(ctx, "pixcost_hsum_kernel");
    size_t rowwords = (size_t)g.W1 * (g.Dp >> 1);
    dim3 vg(cdiv(rowwords, 256), cdiv(g.H, VSUM_ROWS));
    vsum_kernel<<<vg, 256, 0, s>>>((const uint32_t*)hsum, g, (uint32_t*)C);
    MVS_LAUNCH_CHECK(ctx, "vsum_kernel");
    return MVS_OK;
}
