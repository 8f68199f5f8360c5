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
    if (p->uniqueness > 99) return MVS_ERR_UNSUPPORTED;  // the WTA threshold divides by (100 - uniquenessRatio)
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
// One block: NY rows x one tile of TX output columns (W1 coordinates) x all disparities.  A thread owns a pair of
// disparities (u16x2) and walks along x keeping the last WIN pixel costs in a register ring, so the row box sum is
// a sliding window (one add, one subtract per step).  The right image's six prefilter planes are staged reversed
// in shared memory so that consecutive disparities are consecutive halfwords; a second copy shifted by one
// halfword keeps the packed loads aligned for odd offsets.  All arithmetic is packed 16-bit with a +1024 bias so
// that the Birchfield-Tomasi differences never go negative (plain 32-bit adds/subs then act per halfword).
#define COST_K2 0x04000400u
template <int WIN>
__global__ void __launch_bounds__(128)
pixcost_hsum_kernel(const uint8_t* __restrict__ PL, const uint8_t* __restrict__ PR, SgbmGeom g, uint32_t* __restrict__ hsum)
{
    constexpr int NS = 10 * WIN;            // steps per tile (a multiple of the ring length)
    constexpr int TX = NS - (WIN - 1);      // output columns per tile
    constexpr int SW = WIN / 2;
    extern __shared__ __align__(16) uint16_t sm[];
    const int half = g.Dp >> 1;             // disparity pairs = threads per row
    const int t = threadIdx.x % half, ry = threadIdx.x / half;
    const int y = blockIdx.y * (blockDim.x / half) + ry;
    const int x0 = blockIdx.x * TX;
    const size_t plane = (size_t)g.W * g.H;
    // right-image planes, reversed: Rr[k] = R[xmax - k], xmax = largest image x2 any cell of the tile touches
    const int xmax = min(x0 + TX - 1 + SW, g.W1 - 1) + g.minX1 - g.minD;
    const int span = NS + g.Dp + 2;
    const int rows = blockDim.x / half;
    uint16_t* myR = sm + (size_t)ry * (12 * span);          // [plane 0..5][copy E/O][span]
    uint32_t* myL = (uint32_t*)(sm + (((size_t)rows * 12 * span + 7) & ~(size_t)7)) + (size_t)ry * (8 * NS);  // [step][8] pre-biased left values, 16-byte aligned
    if (y < g.H) {
        const uint8_t* R = PR + (size_t)y * g.W;
        for (int idx = t; idx < 6 * span; idx += half) {
            int pl = idx / span, k = idx - pl * span;
            int xr = min(max(xmax - k, 0), g.W - 1);
            uint16_t v = R[pl * plane + xr];
            myR[(pl * 2) * span + k] = v;                    // copy E
            if (k > 0) myR[(pl * 2 + 1) * span + k - 1] = v;  // copy O: O[k] = E[k + 1]
        }
        const uint8_t* L = PL + (size_t)y * g.W;
        for (int k = t; k < NS; k += half) {
            int x = min(max(x0 - SW + k, 0), g.W1 - 1) + g.minX1;
            for (int ch = 0; ch < 2; ++ch) {
                uint32_t u = L[(3 * ch) * plane + x], u0 = L[(3 * ch + 1) * plane + x], u1 = L[(3 * ch + 2) * plane + x];
                myL[k * 8 + ch * 4 + 0] = u * 0x10001u + COST_K2;   // u + K
                myL[k * 8 + ch * 4 + 1] = COST_K2 - u * 0x10001u;   // K - u
                myL[k * 8 + ch * 4 + 2] = COST_K2 - u1 * 0x10001u;  // K - hi(u)
                myL[k * 8 + ch * 4 + 3] = u0 * 0x10001u + COST_K2;  // lo(u) + K
            }
        }
    }
    __syncthreads();
    if (y >= g.H) return;
    uint32_t ring[WIN];
#pragma unroll
    for (int i = 0; i < WIN; ++i) ring[i] = 0;
    uint32_t hs = 0;
    const bool padlo = 2 * t >= g.D, padhi = 2 * t + 1 >= g.D;
    uint32_t* out = hsum + ((size_t)y * g.W1) * half + t;
#pragma unroll 1
    for (int kb = 0; kb < NS; kb += WIN) {
#pragma unroll
        for (int i = 0; i < WIN; ++i) {
            const int k = kb + i;
            const int x = min(max(x0 - SW + k, 0), g.W1 - 1) + g.minX1;
            // halfword index of the pair (d = 2t -> x2, d = 2t+1 -> x2 - 1) in the reversed row
            const int k0 = xmax - (x - g.minD) + 2 * t;
            const uint16_t* base = myR + ((k0 & 1) ? span : 0);
            const uint32_t* pr = (const uint32_t*)(base) + (k0 >> 1);
            const uint4 lva = *(const uint4*)(myL + k * 8), lvb = *(const uint4*)(myL + k * 8 + 4);  // two broadcast 16-byte reads
            const uint32_t lv[8] = {lva.x, lva.y, lva.z, lva.w, lvb.x, lvb.y, lvb.z, lvb.w};
            uint32_t pix = 0;
#pragma unroll
            for (int ch = 0; ch < 2; ++ch) {
                const uint32_t v = pr[(ch * 3 + 0) * span], v0 = pr[(ch * 3 + 1) * span], v1 = pr[(ch * 3 + 2) * span];
                const uint32_t uK = lv[ch * 4 + 0], Ku = lv[ch * 4 + 1], Ku1 = lv[ch * 4 + 2], u0K = lv[ch * 4 + 3];
                // c0 + K = max(K, u - hi(v) + K, lo(v) - u + K); c1 + K = max(K, v - hi(u) + K, lo(u) - v + K)
                uint32_t c0 = __vimax3_u16x2(uK - v1, v0 + Ku, COST_K2);
                uint32_t c1 = __vimax3_u16x2(v + Ku1, u0K - v, COST_K2);
                uint32_t c = __vminu2(c0, c1) - COST_K2;
                pix += ch ? ((c >> 2) & 0x3fff3fffu) : c;
            }
            if (padlo) pix &= 0xffff0000u;
            if (padhi) pix &= 0x0000ffffu;
            hs = hs + pix - ring[i];
            ring[i] = pix;
            const int xo = x0 + k - (WIN - 1);
            if (k >= WIN - 1 && xo < g.W1) out[(size_t)xo * half] = hs;
        }
    }
}

// ---- vertical box sum: C[y] = sum_{dy} hsum[clamp(y+dy)] ; packed u16x2, sliding along y ------------------------
// A thread owns one packed word (x, disparity pair) of a strip of rows and keeps the WIN row sums of its window
// in a register ring, so every row sum is read once per strip (+ WIN-1 at the strip start).
#define VSUM_ROWS 96
template <int WIN>
__global__ void __launch_bounds__(256)
vsum_kernel(const uint32_t* __restrict__ hsum, SgbmGeom g, uint32_t* __restrict__ C)
{
    constexpr int SH = WIN / 2;
    const size_t rowwords = (size_t)g.W1 * (g.Dp >> 1);
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= rowwords) return;
    const int y0 = blockIdx.y * VSUM_ROWS, y1 = min(y0 + VSUM_ROWS, g.H);
    const int zero_from = (g.mode == MVS_MODE_HH4) ? g.H - g.SH2 : g.H;  // OpenCV 4.13 leaves these C rows zero in HH4
    uint32_t ring[WIN];
    uint32_t acc = 0;
#pragma unroll
    for (int k = 0; k < WIN; ++k) {  // ring[k] holds row clamp(y0 - SH + k)
        ring[k] = __ldg(hsum + (size_t)min(max(y0 - SH + k, 0), g.H - 1) * rowwords + i);
        acc += ring[k];
    }
    for (int yb = y0; yb < y1; yb += WIN) {
        uint32_t add[WIN];
#pragma unroll
        for (int k = 0; k < WIN; ++k)  // the WIN rows entering the window during this block of steps: all loads in flight at once
            add[k] = __ldg(hsum + (size_t)min(yb + k + 1 + SH, g.H - 1) * rowwords + i);
#pragma unroll
        for (int k = 0; k < WIN; ++k) {
            const int y = yb + k;
            if (y < y1) C[(size_t)y * rowwords + i] = (y >= zero_from) ? 0u : acc;
            acc = acc - ring[k] + add[k];  // per half: no borrow (acc >= ring[k]) and no carry (sums <= 29237)
            ring[k] = add[k];              // slot k held row clamp(y - SH), the one leaving the window
        }
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
    {
        const int half = g.Dp >> 1, rows = 128 / half > 0 ? 128 / half : 1, win = 2 * g.SW2 + 1;
        const int ns = 10 * win, tx = ns - (win - 1), span = ns + g.Dp + 2;
        dim3 cg(cdiv(g.W1, tx), cdiv(g.H, rows));
        size_t smem = (size_t)rows * (12 * span * sizeof(uint16_t) + 8 * ns * sizeof(uint32_t)) + 16;
        uint32_t* hs32 = (uint32_t*)hsum;
        const uint8_t* PR = planes + 6 * plane;
        switch (win) {
        case 1: pixcost_hsum_kernel<1><<<cg, rows * half, smem, s>>>(planes, PR, g, hs32); break;
        case 3: pixcost_hsum_kernel<3><<<cg, rows * half, smem, s>>>(planes, PR, g, hs32); break;
        case 5: pixcost_hsum_kernel<5><<<cg, rows * half, smem, s>>>(planes, PR, g, hs32); break;
        case 7: pixcost_hsum_kernel<7><<<cg, rows * half, smem, s>>>(planes, PR, g, hs32); break;
        case 9: pixcost_hsum_kernel<9><<<cg, rows * half, smem, s>>>(planes, PR, g, hs32); break;
        case 11: pixcost_hsum_kernel<11><<<cg, rows * half, smem, s>>>(planes, PR, g, hs32); break;
        case 13: pixcost_hsum_kernel<13><<<cg, rows * half, smem, s>>>(planes, PR, g, hs32); break;
        default: return mvs_fail(ctx, MVS_ERR_UNSUPPORTED, "cost: block size %d", win);
        }
        MVS_LAUNCH_CHECK(ctx, "pixcost_hsum_kernel");
    }
    size_t rowwords = (size_t)g.W1 * (g.Dp >> 1);
    dim3 vg(cdiv(rowwords, 256), cdiv(g.H, VSUM_ROWS));
    switch (2 * g.SH2 + 1) {
    case 1: vsum_kernel<1><<<vg, 256, 0, s>>>((const uint32_t*)hsum, g, (uint32_t*)C); break;
    case 3: vsum_kernel<3><<<vg, 256, 0, s>>>((const uint32_t*)hsum, g, (uint32_t*)C); break;
    case 5: vsum_kernel<5><<<vg, 256, 0, s>>>((const uint32_t*)hsum, g, (uint32_t*)C); break;
    case 7: vsum_kernel<7><<<vg, 256, 0, s>>>((const uint32_t*)hsum, g, (uint32_t*)C); break;
    case 9: vsum_kernel<9><<<vg, 256, 0, s>>>((const uint32_t*)hsum, g, (uint32_t*)C); break;
    case 11: vsum_kernel<11><<<vg, 256, 0, s>>>((const uint32_t*)hsum, g, (uint32_t*)C); break;
    default: vsum_kernel<13><<<vg, 256, 0, s>>>((const uint32_t*)hsum, g, (uint32_t*)C); break;
    }
    MVS_LAUNCH_CHECK(ctx, "vsum_kernel");
    return MVS_OK;
}
