// a3 (iv): SGM path aggregation of cv::StereoSGBM:
//   L(p,d) = C(p,d) + min(Lp[d], Lp[d-1]+P1, Lp[d+1]+P1, min(Lp)+P2) - min(Lp),   S = sat16(sum of L)
// over 8 (MODE_HH), 5 (MODE_SGBM) or 4 (MODE_HH4) directions.
//
// Inside the domain accepted by mvs_sgbm_check all L and C lie in [0, 32767], so the reference's
// per-pass saturating accumulation equals min(sum, 32767) in ANY order; this file relies on that.
//
// Data layout: 16-bit volumes [y][x][Dp], Dp = D rounded up to 64; one warp owns one pixel's
// disparity vector, each lane 2*NW consecutive disparities packed two per 32-bit register (u16x2).
// Neighbouring disparities come from byte permutes inside the lane and one shuffle per side; the
// minimum over d is one packed min per register plus a warp REDUX.
#include <stdlib.h>

#include "common.cuh"

#define SENT 0x7FFFu
#define SENT2 0x7FFF7FFFu

template <int NW>
struct PathState {
    uint32_t L[NW];  // previous pixel's path costs (u16x2)
    uint32_t minL;   // min over d, replicated in both halves
};

__device__ __forceinline__ uint32_t dup16(uint32_t v) { return v | (v << 16); }

// One step of one path for the pixel whose costs are in Cw[]; updates st and returns L in Lout.
// padmask[j]: 0xFFFF in the halves that are padding (d >= D), else 0.
template <int NW, bool PAD>
__device__ __forceinline__ void path_step(PathState<NW>& st, const uint32_t (&Cw)[NW], uint32_t P1x2, uint32_t P2,
                                          const uint32_t (&padmask)[NW], uint32_t (&Lout)[NW], int lane)
{
    // neighbours across lanes: the last word of lane-1 (for d-1) and the first word of lane+1 (for d+1)
    uint32_t left = __shfl_up_sync(0xffffffffu, st.L[NW - 1], 1);
    uint32_t right = __shfl_down_sync(0xffffffffu, st.L[0], 1);
    if (lane == 0) left = SENT2;
    if (lane == 31) right = SENT2;
    uint32_t delta = dup16((st.minL & 0xffffu) + P2);
    uint32_t m = SENT2;
#pragma unroll
    for (int j = 0; j < NW; ++j) {
        uint32_t a = st.L[j];
        uint32_t lo = __byte_perm(j == 0 ? left : st.L[j - 1], a, 0x5432);       // Lp[d-1]
        uint32_t hi = __byte_perm(a, j == NW - 1 ? right : st.L[j + 1], 0x5432);  // Lp[d+1]
        uint32_t t = __vadd2(__vminu2(lo, hi), P1x2);
        t = __vminu2(__vminu2(t, a), delta);
        uint32_t l = Cw[j] + t - st.minL;  // per half: t >= minL, C + (t - minL) <= 32767: no borrow/carry
        if (PAD) l = (l & ~padmask[j]) | (SENT2 & padmask[j]);
        Lout[j] = l;
        m = __vminu2(m, l);
    }
    uint32_t mm = min(m & 0xffffu, m >> 16);
    mm = __reduce_min_sync(0xffffffffu, mm);
#pragma unroll
    for (int j = 0; j < NW; ++j) st.L[j] = Lout[j];
    st.minL = dup16(mm);
}

template <int NW>
__device__ __forceinline__ void load_words(const uint32_t* p, uint32_t (&w)[NW])
{
    if (NW == 4) {
        uint4 v = *(const uint4*)p;
        w[0] = v.x; w[1] = v.y; w[2] = v.z; w[3 % NW] = v.w;
    } else if (NW == 2) {
        uint2 v = *(const uint2*)p;
        w[0] = v.x; w[1 % NW] = v.y;
    } else {
#pragma unroll
        for (int j = 0; j < NW; ++j) w[j] = p[j];
    }
}
template <int NW>
__device__ __forceinline__ void store_words(uint32_t* p, const uint32_t (&w)[NW])
{
    if (NW == 4) {
        *(uint4*)p = make_uint4(w[0], w[1], w[2], w[3 % NW]);
    } else if (NW == 2) {
        *(uint2*)p = make_uint2(w[0], w[1 % NW]);
    } else {
#pragma unroll
        for (int j = 0; j < NW; ++j) p[j] = w[j];
    }
}

// v1 aggregation: one warp walks one path of one direction; S is read-modify-written per direction.
template <int NW, bool PAD>
__global__ void __launch_bounds__(256)
path_kernel(const uint32_t* __restrict__ C, uint32_t* __restrict__ S, int W1, int H, int D, int P1, int P2, int dx,
            int dy, int first)
{
    const int lane = threadIdx.x & 31;
    const int path = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    int npaths = (dy == 0) ? H : (dx == 0 ? W1 : W1 + H - 1);
    if (path >= npaths) return;
    int x, y, len;
    if (dy == 0) {
        y = path; x = dx > 0 ? 0 : W1 - 1; len = W1;
    } else {
        int ys = dy > 0 ? 0 : H - 1;
        if (dx == 0) { x = path; y = ys; len = H; }
        else if (path < W1) {
            x = path; y = ys;
            len = min(H, dx > 0 ? W1 - x : x + 1);
        } else {
            int k = path - W1 + 1;  // rows away from the starting row
            x = dx > 0 ? 0 : W1 - 1;
            y = dy > 0 ? k : H - 1 - k;
            len = min(H - k, W1);
        }
    }
    uint32_t padmask[NW];
#pragma unroll
    for (int j = 0; j < NW; ++j) {
        int d0 = (lane * NW + j) * 2;
        padmask[j] = (d0 >= D ? 0xffffu : 0u) | (d0 + 1 >= D ? 0xffff0000u : 0u);
    }
    PathState<NW> st;
#pragma unroll
    for (int j = 0; j < NW; ++j) st.L[j] = 0;
    st.minL = 0;
    const size_t wpp = (size_t)NW * 32;  // words per pixel
    long long off = ((long long)y * W1 + x) * (long long)wpp + lane * NW;
    const long long stride = ((long long)dy * W1 + dx) * (long long)wpp;
    const uint32_t P1x2 = dup16((uint32_t)P1);
    uint32_t Cw[NW], Sw[NW], Cn[NW], Sn[NW], Lw[NW];
    load_words<NW>(C + off, Cw);
    if (!first) load_words<NW>(S + off, Sw);
    for (int i = 0; i < len; ++i) {
        if (i + 1 < len) {  // prefetch the next pixel before the dependent chain
            load_words<NW>(C + off + stride, Cn);
            if (!first) load_words<NW>(S + off + stride, Sn);
        }
        path_step<NW, PAD>(st, Cw, P1x2, (uint32_t)P2, padmask, Lw, lane);
        if (first) {
#pragma unroll
            for (int j = 0; j < NW; ++j) Sw[j] = Lw[j];
        } else {
#pragma unroll
            for (int j = 0; j < NW; ++j) Sw[j] = __vminu2(Sw[j] + Lw[j], SENT2);
        }
        store_words<NW>(S + off, Sw);
        off += stride;
#pragma unroll
        for (int j = 0; j < NW; ++j) { Cw[j] = Cn[j]; Sw[j] = Sn[j]; }
    }
}

template <int NW, bool PAD>
static void launch_path(cudaStream_t s, const SgbmGeom& g, const int16_t* C, int16_t* S, int dx, int dy, int first)
{
    int npaths = (dy == 0) ? g.H : (dx == 0 ? g.W1 : g.W1 + g.H - 1);
    path_kernel<NW, PAD><<<cdiv(npaths, 8), 256, 0, s>>>((const uint32_t*)C, (uint32_t*)S, g.W1, g.H, g.D, g.P1, g.P2, dx,
                                                          dy, first);
}

int sgbm_aggregate_paths_dev(mvs_ctx* ctx, const SgbmGeom& g, const int16_t* C, int16_t* S)
{
    // direction sets (A.2/A.3); forward directions first, as the reference does
    static const int dirs_hh[8][2] = {{1, 0}, {1, 1}, {0, 1}, {-1, 1}, {-1, 0}, {1, -1}, {0, -1}, {-1, -1}};
    static const int dirs_sgbm[5][2] = {{1, 0}, {1, 1}, {0, 1}, {-1, 1}, {-1, 0}};
    static const int dirs_hh4[4][2] = {{1, 0}, {0, 1}, {-1, 0}, {0, -1}};
    const int(*dirs)[2] = g.mode == MVS_MODE_HH ? dirs_hh : (g.mode == MVS_MODE_SGBM ? dirs_sgbm : dirs_hh4);
    int nd = g.mode == MVS_MODE_HH ? 8 : (g.mode == MVS_MODE_SGBM ? 5 : 4);
    bool pad = g.Dp != g.D;
    for (int i = 0; i < nd; ++i) {
        int dx = dirs[i][0], dy = dirs[i][1], first = i == 0;
        switch (g.Dp / 64) {
        case 1: pad ? launch_path<1, true>(ctx->stream, g, C, S, dx, dy, first) : launch_path<1, false>(ctx->stream, g, C, S, dx, dy, first); break;
        case 2: pad ? launch_path<2, true>(ctx->stream, g, C, S, dx, dy, first) : launch_path<2, false>(ctx->stream, g, C, S, dx, dy, first); break;
        case 3: pad ? launch_path<3, true>(ctx->stream, g, C, S, dx, dy, first) : launch_path<3, false>(ctx->stream, g, C, S, dx, dy, first); break;
        case 4: pad ? launch_path<4, true>(ctx->stream, g, C, S, dx, dy, first) : launch_path<4, false>(ctx->stream, g, C, S, dx, dy, first); break;
        default: return mvs_fail(ctx, MVS_ERR_UNSUPPORTED, "aggregate: Dp=%d", g.Dp);
        }
        MVS_LAUNCH_CHECK(ctx, "path_kernel");
    }
    return MVS_OK;
}

// MVS_AGG=paths selects the per-direction kernels above (kept for A/B checks of the fused sweeps).
int sgbm_aggregate_batch_dev(mvs_ctx* ctx, const SgbmGeom& g, int F, const int16_t* const* C, int16_t* const* S)
{
    static const bool use_paths = [] { const char* e = getenv("MVS_AGG"); return e && !strcmp(e, "paths"); }();
    if (!use_paths) return sgbm_aggregate_wave_batch_dev(ctx, g, F, C, S);
    for (int f = 0; f < F; ++f) {
        int rc = sgbm_aggregate_paths_dev(ctx, g, C[f], S[f]);
        if (rc) return rc;
    }
    return MVS_OK;
}

int sgbm_aggregate_dev(mvs_ctx* ctx, const SgbmGeom& g, const int16_t* C, int16_t* S)
{
    return sgbm_aggregate_batch_dev(ctx, g, 1, &C, &S);
}
