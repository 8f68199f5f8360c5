// a3 (i)-(iii): matching cost volume C[y][x][d] of cv::StereoSGBM (reference call site
// cost-functions/disparity/strategies/stereosgbm_disp.cpp:121-125; algorithm: SURVEY.md A.1):
//   x-Sobel prefilter clipped to [0, 2*ftzero] + raw intensity, Birchfield-Tomasi sampling-insensitive
//   cost on both channels, then a blockSize x blockSize box sum with replicate clamping in the
//   valid-column (W1) domain.  16-bit, Dp = D rounded up to 64 disparities per pixel (pad = 0), split layout (common.cuh).
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
    g->K44 = 0;
    if (ctx && ctx->sgbm_subtract_p2) {
        const int dirs = p->mode == MVS_MODE_HH ? 8 : p->mode == MVS_MODE_SGBM ? 5 : 4;
        g->K44 = dirs * g->P2;
        // S' = S - K44 must stay above the lower end of a 16-bit cost
        if (g->K44 > 32767)
            return mvs_fail(ctx, MVS_ERR_UNSUPPORTED, "SGBM: the (min Lp + P2) form of the recurrence supports directions * P2 <= 32767 (here %d)", g->K44);
    }
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

// Four pixels per thread: the six p / q values it needs (x0 - 1 .. x0 + 4) come from eight columns of three rows read once,
// and every plane gets one 4-byte store (when the image width is a multiple of four; byte stores otherwise).
__device__ __forceinline__ void prefilter_body(const uint8_t* __restrict__ img, int W, int H, int ftzero, uint8_t* __restrict__ planes)
{
    const int x0 = (blockIdx.x * blockDim.x + threadIdx.x) * 4, y = blockIdx.y;
    if (x0 >= W) return;
    const uint8_t* r = img + (size_t)y * W;
    const uint8_t* n = y > 0 ? r - W : r;
    const uint8_t* s = y < H - 1 ? r + W : r;
    // Sobel column sums h(c) = 2 r[c] + n[c] + s[c] for c = x0 - 2 .. x0 + 5 (clamped reads: values outside are never used)
    int h[8], q[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) {
        const int c = min(max(x0 - 2 + k, 0), W - 1);
        q[k] = r[c];
        h[k] = 2 * q[k] + (int)n[c] + (int)s[c];
    }
    // p / q at x0 - 1 .. x0 + 4 (index k = x - x0 + 1); border columns hold ftzero
    int pv[6], qv[6];
#pragma unroll
    for (int k = 0; k < 6; ++k) {
        const int x = x0 - 1 + k;
        const bool border = x <= 0 || x >= W - 1;
        pv[k] = border ? ftzero : max(-ftzero, min(ftzero, h[k + 2] - h[k])) + ftzero;
        qv[k] = border ? ftzero : q[k + 1];
    }
    uint32_t out[6] = {0, 0, 0, 0, 0, 0};
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        const int x = x0 + k;
        const int pc = pv[k + 1], qc = qv[k + 1];
        int pl = pc, pr = pc, ql = qc, qr = qc;
        if (x > 0) { pl = (pc + pv[k]) >> 1; ql = (qc + qv[k]) >> 1; }
        if (x < W - 1) { pr = (pc + pv[k + 2]) >> 1; qr = (qc + qv[k + 2]) >> 1; }
        out[0] |= (uint32_t)pc << (8 * k);
        out[1] |= (uint32_t)min(pc, min(pl, pr)) << (8 * k);
        out[2] |= (uint32_t)max(pc, max(pl, pr)) << (8 * k);
        out[3] |= (uint32_t)qc << (8 * k);
        out[4] |= (uint32_t)min(qc, min(ql, qr)) << (8 * k);
        out[5] |= (uint32_t)max(qc, max(ql, qr)) << (8 * k);
    }
    const size_t plane = (size_t)W * H, o = (size_t)y * W + x0;
    if ((W & 3) == 0) {
#pragma unroll
        for (int pl_ = 0; pl_ < 6; ++pl_) *(uint32_t*)(planes + pl_ * plane + o) = out[pl_];
    } else {
        const int cnt = min(4, W - x0);
#pragma unroll
        for (int pl_ = 0; pl_ < 6; ++pl_)
            for (int k = 0; k < cnt; ++k) planes[pl_ * plane + o + k] = (uint8_t)(out[pl_] >> (8 * k));
    }
}

__global__ void prefilter_kernel(const uint8_t* __restrict__ img, int W, int H, int ftzero, uint8_t* __restrict__ planes)
{
    prefilter_body(img, W, H, ftzero, planes);
}
// all images of a batch in one launch: image z (left images 0..F-1, right images F..2F-1) -> planes + z * 6 * W * H
struct PrefilterBatch { const uint8_t* img[2 * MVS_MAX_BATCH]; };
__global__ void prefilter_batch_kernel(PrefilterBatch b, int W, int H, int ftzero, uint8_t* __restrict__ planes)
{
    prefilter_body(b.img[blockIdx.z], W, H, ftzero, planes + (size_t)blockIdx.z * 6 * W * H);
}

// ---- per-pixel BT cost + horizontal box sum ------------------------------------------------------------------
// One block: a few rows x one tile of output columns (W1 coordinates) x all disparities.  A thread owns a pair of
// disparities (u16x2) and walks along x keeping the last WIN pixel costs in a register ring, so the row box sum is
// a sliding window (one add, one subtract per step).  Step k of a tile is the virtual column xs + k; columns left
// of the image are folded into the ring's initial state, columns right of it repeat the last pixel cost
// (replicate clamping).  A pair is the two disparities that share a 32-bit word of the volume: t and t + Dp/2 (the
// split layout, common.cuh).  The right image's prefilter planes are staged reversed in shared memory, one 16-byte
// element {2 pixels Dp/2 apart x 3 planes x 2 channels, as bytes} per position, so that the element of pair t at step
// k sits at NS + t - k: every address of
// the unrolled loop is a running pointer plus an immediate, and a step costs one 16-byte load per thread (the
// kernel is bound by shared-memory wavefronts, so the bytes are widened to u16x2 with PRMT in registers).  All arithmetic is packed 16-bit with a +1024 bias so that the
// Birchfield-Tomasi differences never go negative (plain 32-bit adds/subs then act per halfword).
#define COST_K2 0x04000400u
template <int WIN, int NW>
struct CostTile {
    static constexpr int HALF = 32 * NW;                       // disparity pairs per row
#ifndef COST_PPT
#define COST_PPT 1   // measured at c2: 2 pairs per thread (shared left-image loads, 80 registers) 0.61 ms vs 0.53 ms with 1
#endif
    static constexpr int PPT = (NW % 2 == 0) ? COST_PPT : 1;   // pairs per thread (t and t + TPR): the left-image loads are shared
    static constexpr int TPR = HALF / PPT;                     // threads per row
    static constexpr int ROWS = NW == 3 ? 1 : 128 / TPR;       // rows per block
    static constexpr int THREADS = ROWS * TPR;
    static constexpr int BODY = 2 * WIN;                       // steps per unrolled body (even, a multiple of the ring length)
    static constexpr int NS = BODY * ((128 + BODY - 1) / BODY);  // steps per tile
    static constexpr int SW = WIN / 2;
    static constexpr int TX = NS - (WIN - 1);                  // output columns of an inner tile
    static constexpr int TX0 = NS - SW;                        // output columns of the first tile (its left window half is folded)
    static constexpr int R4 = NS + HALF;                       // staged right-image elements (uint4) per row
    static constexpr int L4 = 2 * NS;                          // uint4 per row: [step][channel]
    static constexpr size_t SMEM = (size_t)ROWS * (R4 + L4) * 16;
};

__device__ __forceinline__ uint32_t bt_channel_cost(uint32_t w0, uint32_t w1, const uint4& l)
{
    // right image: w0 = bytes {v, v', lo(v), lo(v')}, w1 = bytes {hi(v), hi(v'), -, -} (v' = the pair's second disparity);
    // left image: l = {u + K, K - u, K - hi(u), lo(u) + K}
    // c0 + K = max(K, u - hi(v) + K, lo(v) - u + K); c1 + K = max(K, v - hi(u) + K, lo(u) - v + K)
    const uint32_t v = __byte_perm(w0, 0u, 0x4140), vlo = __byte_perm(w0, 0u, 0x4342), vhi = __byte_perm(w1, 0u, 0x4140);
    uint32_t c0 = __vimax3_u16x2(l.x - vhi, vlo + l.y, COST_K2);
    uint32_t c1 = __vimax3_u16x2(v + l.z, l.w - v, COST_K2);
    return __vminu2(c0, c1);   // cost + K per half
}
__device__ __forceinline__ uint32_t bt_pair_cost(const uint4& r, const uint4& l0, const uint4& l1)
{
    // cost(channel 0) + cost(channel 1) / 4 per half; K is a multiple of 4, so ((c + K) >> 2) = (c >> 2) + K / 4 and
    // both biases come off in one subtraction
    return bt_channel_cost(r.x, r.y, l0) + ((bt_channel_cost(r.z, r.w, l1) >> 2) & 0x3fff3fffu) - (COST_K2 + (COST_K2 >> 2));
}

template <int WIN, int NW>
__device__ __forceinline__ void pixcost_hsum_block(const uint8_t* __restrict__ PL, const uint8_t* __restrict__ PR, const SgbmGeom& g,
                                                   uint32_t* __restrict__ hsum, int tile, int yblock)
{
    using T = CostTile<WIN, NW>;
    constexpr int NS = T::NS, SW = T::SW, HALF = T::HALF, PPT = T::PPT, TPR = T::TPR;
    extern __shared__ __align__(16) uint4 sm4[];
    const int t = threadIdx.x % TPR, ry = threadIdx.x / TPR;
    const int y = yblock * T::ROWS + ry;
    const bool first = tile == 0;
    const int xs = first ? 0 : T::TX0 + (tile - 1) * T::TX - SW;   // virtual column of step 0
    const size_t plane = (size_t)g.W * g.H;
    uint4* myR = sm4 + (size_t)ry * (T::R4 + T::L4);
    uint4* myL = myR + T::R4;
    // reversed right row: E[i] = R[xmax - i]; pair t (disparities t and t + HALF, the split layout) at step k needs
    // E[i] and E[i + HALF] with i = NS + t - k: staged element i = {E[i], E[i + HALF]}
    const int xmax = xs + NS + g.minX1 - g.minD;
    if (y < g.H) {
        const uint8_t* R = PR + (size_t)y * g.W;
        for (int idx = t; idx < T::R4; idx += TPR) {
            const int xa = min(max(xmax - idx, 0), g.W - 1), xb = min(max(xmax - idx - HALF, 0), g.W - 1);
            uint32_t w[4];
#pragma unroll
            for (int ch = 0; ch < 2; ++ch) {
                const uint8_t* q = R + (size_t)(3 * ch) * plane;
                w[2 * ch] = (uint32_t)q[xa] | ((uint32_t)q[xb] << 8) | ((uint32_t)q[plane + xa] << 16) | ((uint32_t)q[plane + xb] << 24);
                w[2 * ch + 1] = (uint32_t)q[2 * plane + xa] | ((uint32_t)q[2 * plane + xb] << 8);
            }
            myR[idx] = make_uint4(w[0], w[1], w[2], w[3]);
        }
        const uint8_t* L = PL + (size_t)y * g.W;
        for (int idx = t; idx < T::L4; idx += TPR) {
            const int k = idx >> 1, ch = idx & 1;
            const int x = min(max(xs + k, 0), g.W1 - 1) + g.minX1;
            const uint32_t u = L[(3 * ch) * plane + x], u0 = L[(3 * ch + 1) * plane + x], u1 = L[(3 * ch + 2) * plane + x];
            myL[idx] = make_uint4(u * 0x10001u + COST_K2, COST_K2 - u * 0x10001u, COST_K2 - u1 * 0x10001u, u0 * 0x10001u + COST_K2);
        }
    }
    __syncthreads();
    if (y >= g.H) return;
    uint32_t padmask[PPT];
#pragma unroll
    for (int q = 0; q < PPT; ++q) {
        const int tq = t + q * TPR;   // this thread's q-th disparity pair
        padmask[q] = (tq >= g.D ? 0u : 0xffffu) | (tq + HALF >= g.D ? 0u : 0xffff0000u);
    }
    const uint4* pR = myR + NS + t;       // element of step 0 of pair t; pair t + TPR sits TPR elements further
    const uint4* pL = myL;
    const int nvalid = min(NS, g.W1 - xs);            // steps whose column lies inside the image
    const int nout = min(NS, g.W1 - xs + SW);         // steps that still produce an output column
    const int kmin = first ? SW : WIN - 1;            // first step with a complete window
    uint32_t ring[PPT][WIN];
    uint32_t hs[PPT], pix[PPT];
#pragma unroll
    for (int q = 0; q < PPT; ++q) {
        hs[q] = 0; pix[q] = 0;
#pragma unroll
        for (int i = 0; i < WIN; ++i) ring[q][i] = 0;
    }
    if (first) {
        // columns -SW..-1 repeat column 0: they sit in the ring slots that leave the window at steps SW+1..WIN-1
#pragma unroll
        for (int q = 0; q < PPT; ++q) {
            const uint32_t pb = bt_pair_cost(pR[q * TPR], pL[0], pL[1]) & padmask[q];
#pragma unroll
            for (int i = SW + 1; i < WIN; ++i) ring[q][i] = pb;
            hs[q] = SW * pb;
        }
    }
    uint32_t* out = hsum + ((ptrdiff_t)y * g.W1 + (xs - SW)) * HALF + t;   // output column of step 0 (never stored when negative)
#pragma unroll 1
    for (int kb = 0; kb < NS; kb += T::BODY) {
        if (kb >= T::BODY && kb + T::BODY <= nvalid) {
            // interior body: every step is inside the image and produces an output (nvalid <= nout)
#pragma unroll
            for (int i = 0; i < T::BODY; ++i) {
                const uint4* r = pR - i;
                const uint4 l0 = pL[2 * i], l1 = pL[2 * i + 1];
#pragma unroll
                for (int q = 0; q < PPT; ++q) {
                    pix[q] = bt_pair_cost(r[q * TPR], l0, l1) & padmask[q];
                    hs[q] = hs[q] + pix[q] - ring[q][i % WIN];
                    ring[q][i % WIN] = pix[q];
                    out[(size_t)i * HALF + q * TPR] = hs[q];
                }
            }
            pR -= T::BODY;
            pL += 2 * T::BODY;
            out += (size_t)T::BODY * HALF;
            continue;
        }
#pragma unroll
        for (int i = 0; i < T::BODY; ++i) {
            const int k = kb + i;
            const uint4* r = pR - i;
            const uint4 l0 = pL[2 * i], l1 = pL[2 * i + 1];
#pragma unroll
            for (int q = 0; q < PPT; ++q) {
                const uint32_t c = bt_pair_cost(r[q * TPR], l0, l1) & padmask[q];
                pix[q] = k < nvalid ? c : pix[q];
                hs[q] = hs[q] + pix[q] - ring[q][i % WIN];
                ring[q][i % WIN] = pix[q];
                if (k >= kmin && k < nout) out[(size_t)i * HALF + q * TPR] = hs[q];
            }
        }
        pR -= T::BODY;
        pL += 2 * T::BODY;
        out += (size_t)T::BODY * HALF;
    }
}

// ---- vertical box sum: C[y] = sum_{dy} hsum[clamp(y+dy)] ; packed u16x2, sliding along y ------------------------
// A thread owns four packed words (16 bytes) of a strip of rows: C(y+1) = C(y) + hsum(y+SH+1) - hsum(y-SH).  The row
// leaving the window was read WIN rows earlier by the same thread and is still in L2, so HBM sees every row sum once
// per strip (+ WIN-1 at the strip start) while the instruction count stays at ~4 per word and row.
#ifndef VSUM_ROWS
#define VSUM_ROWS 96
#endif
__device__ __forceinline__ uint4 add4(uint4 a, uint4 b) { return make_uint4(a.x + b.x, a.y + b.y, a.z + b.z, a.w + b.w); }
__device__ __forceinline__ uint4 sub4(uint4 a, uint4 b) { return make_uint4(a.x - b.x, a.y - b.y, a.z - b.z, a.w - b.w); }
template <int WIN, int THREADS>
__device__ __forceinline__ void vsum_block(const uint32_t* __restrict__ hsum, const SgbmGeom& g, uint32_t* __restrict__ C, int chunk, int strip)
{
    constexpr int SH = WIN / 2;
    const size_t row4 = (size_t)g.W1 * (g.Dp >> 3);   // 16-byte elements per row (Dp/2 words per pixel, a multiple of 4)
    const size_t i = (size_t)chunk * THREADS + threadIdx.x;
    if (i >= row4) return;
    const uint4* __restrict__ h4 = (const uint4*)hsum + i;
    uint4* __restrict__ c4 = (uint4*)C + i;
    const int y0 = strip * VSUM_ROWS, y1 = min(y0 + VSUM_ROWS, g.H);
    const int zero_from = (g.mode == MVS_MODE_HH4) ? g.H - g.SH2 : g.H;  // OpenCV 4.13 leaves these C rows zero in HH4
    uint4 acc = make_uint4(0, 0, 0, 0);
#pragma unroll
    for (int k = 0; k < WIN; ++k) acc = add4(acc, __ldg(h4 + (size_t)min(max(y0 - SH + k, 0), g.H - 1) * row4));
#pragma unroll 4
    for (int y = y0; y < y1; ++y) {
        const uint4 in = __ldg(h4 + (size_t)min(y + SH + 1, g.H - 1) * row4);    // first touch: stays in L2 for the second
        const uint4 out = __ldcs(h4 + (size_t)max(y - SH, 0) * row4);            // read WIN rows ago (an L2 hit), last use
        __stcs(c4 + (size_t)y * row4, y >= zero_from ? make_uint4(0, 0, 0, 0) : acc);
        acc = sub4(add4(acc, in), out);  // per half: no borrow (acc >= out) and no carry (sums <= 29237)
    }
}

// One launch = the pixel-cost + row-sum blocks of one frame ("pix") interleaved with the column-sum blocks of the
// previous frame of the batch ("vsum").  The first kind is bound by shared-memory wavefronts and issue slots, the
// second by HBM: co-resident on the same SMs they overlap almost completely.  Either kind may be absent (n = 0).
struct CostJob {
    const uint8_t *PL, *PR;   // pix: prefilter planes of the left / right image
    uint32_t* hsum_out;       // pix: row sums written
    const uint32_t* hsum_in;  // vsum: row sums read (the other half of the double buffer)
    uint32_t* C;              // vsum: cost volume written
    int n_pix, n_vsum;        // blocks of each kind
    int ntiles, nchunks;      // pix block = (tile, row block), vsum block = (column chunk, row strip)
};
template <int WIN, int NW>
__global__ void __launch_bounds__(CostTile<WIN, NW>::THREADS)
cost_kernel(CostJob j, SgbmGeom g)
{
    using T = CostTile<WIN, NW>;
    // proportional interleave: block i is the v-th vsum block iff v = floor(i * n_vsum / n) steps up at i + 1
    const unsigned long long n = (unsigned long long)j.n_pix + j.n_vsum;
    const int v0 = (int)((unsigned long long)blockIdx.x * j.n_vsum / n), v1 = (int)(((unsigned long long)blockIdx.x + 1) * j.n_vsum / n);
    const bool pix = v1 == v0;
    const int idx = pix ? (int)blockIdx.x - v0 : v0;
    if (pix) pixcost_hsum_block<WIN, NW>(j.PL, j.PR, g, j.hsum_out, idx % j.ntiles, idx / j.ntiles);
    else vsum_block<WIN, T::THREADS>(j.hsum_in, g, j.C, idx % j.nchunks, idx / j.nchunks);
}

template <int WIN, int NW>
static int launch_cost(mvs_ctx* ctx, const SgbmGeom& g, CostJob j, bool pix, bool vsum)
{
    using T = CostTile<WIN, NW>;
    static_assert(T::SMEM <= 48 * 1024, "cost tile exceeds the default shared-memory window");
    j.ntiles = 1 + (g.W1 > T::TX0 ? cdiv(g.W1 - T::TX0, T::TX) : 0);
    j.nchunks = cdiv((long long)g.W1 * (g.Dp >> 3), T::THREADS);
    j.n_pix = pix ? j.ntiles * cdiv(g.H, T::ROWS) : 0;
    j.n_vsum = vsum ? j.nchunks * cdiv(g.H, VSUM_ROWS) : 0;
    cost_kernel<WIN, NW><<<j.n_pix + j.n_vsum, T::THREADS, T::SMEM, ctx->stream>>>(j, g);
    MVS_LAUNCH_CHECK(ctx, "cost_kernel");
    return MVS_OK;
}
template <int WIN>
static int launch_cost_nw(mvs_ctx* ctx, const SgbmGeom& g, const CostJob& j, bool pix, bool vsum)
{
    switch (g.Dp / 64) {
    case 1: return launch_cost<WIN, 1>(ctx, g, j, pix, vsum);
    case 2: return launch_cost<WIN, 2>(ctx, g, j, pix, vsum);
    case 3: return launch_cost<WIN, 3>(ctx, g, j, pix, vsum);
    case 4: return launch_cost<WIN, 4>(ctx, g, j, pix, vsum);
    }
    return mvs_fail(ctx, MVS_ERR_UNSUPPORTED, "cost: Dp %d", g.Dp);
}
static int launch_cost_win(mvs_ctx* ctx, const SgbmGeom& g, const CostJob& j, bool pix, bool vsum)
{
    switch (2 * g.SW2 + 1) {
    case 1: return launch_cost_nw<1>(ctx, g, j, pix, vsum);
    case 3: return launch_cost_nw<3>(ctx, g, j, pix, vsum);
    case 5: return launch_cost_nw<5>(ctx, g, j, pix, vsum);
    case 7: return launch_cost_nw<7>(ctx, g, j, pix, vsum);
    case 9: return launch_cost_nw<9>(ctx, g, j, pix, vsum);
    case 11: return launch_cost_nw<11>(ctx, g, j, pix, vsum);
    case 13: return launch_cost_nw<13>(ctx, g, j, pix, vsum);
    }
    return mvs_fail(ctx, MVS_ERR_UNSUPPORTED, "cost: block size %d", 2 * g.SW2 + 1);
}

// Cost volumes of a batch of F pairs: launch k builds the row sums of frame k and, from the other half of a
// double buffer, the column sums (= the cost volume) of frame k - 1.
int sgbm_cost_volume_batch_dev(mvs_ctx* ctx, const SgbmGeom& g, int F, const uint8_t* const* l, const uint8_t* const* r,
                               int16_t* const* C)
{
    cudaStream_t s = ctx->stream;
    const size_t plane = (size_t)g.W * g.H;
    MVS_ARENA(ctx, uint8_t, planes, "sgbm.planes", plane * 12 * F);   // [left images][right images] x 6 planes
    uint16_t* hsum[2];
    hsum[0] = (uint16_t*)arena_get(ctx, "sgbm.hsum", g.cells() * 2);
    hsum[1] = F > 1 ? (uint16_t*)arena_get(ctx, "sgbm.hsum.1", g.cells() * 2) : hsum[0];
    if (!hsum[0] || !hsum[1]) return MVS_ERR_NOMEM;
    {
        PrefilterBatch pb;
        for (int f = 0; f < F; ++f) { pb.img[f] = l[f]; pb.img[F + f] = r[f]; }
        prefilter_batch_kernel<<<dim3(cdiv(g.W, 4 * 128), g.H, 2 * F), 128, 0, s>>>(pb, g.W, g.H, g.ftzero, planes);
        MVS_LAUNCH_CHECK(ctx, "prefilter_batch_kernel");
    }
    for (int k = 0; k <= F; ++k) {
        CostJob j;
        memset(&j, 0, sizeof(j));
        if (k < F) {
            j.PL = planes + (size_t)k * 6 * plane; j.PR = planes + (size_t)(F + k) * 6 * plane; j.hsum_out = (uint32_t*)hsum[k & 1];
        }
        if (k > 0) { j.hsum_in = (const uint32_t*)hsum[(k - 1) & 1]; j.C = (uint32_t*)C[k - 1]; }
        int rc = launch_cost_win(ctx, g, j, k < F, k > 0);
        if (rc) return rc;
    }
    return MVS_OK;
}

int sgbm_cost_volume_dev(mvs_ctx* ctx, const SgbmGeom& g, const uint8_t* l, const uint8_t* r, int16_t* C)
{
    return sgbm_cost_volume_batch_dev(ctx, g, 1, &l, &r, &C);
}
