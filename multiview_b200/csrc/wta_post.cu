// a3 (v)-(vii) and a4: winner-take-all with uniqueness and parabola sub-pixel (the form that reads S from memory),
// right-view consistency check, 3x3 median, speckle filter and float conversion (reference call sites
// cost-functions/disparity/strategies/stereosgbm_disp.cpp:121-125 and :141-150).
#include "common.cuh"

// The winner-take-all step (first minimum, uniqueness, sub-pixel fit, right-view key) exists twice: from 192 disparities it
// runs inside the backward aggregation sweep on the sums the sweep holds in registers (aggregate_wave.cu: wta_stage_a/b,
// wta_flush), which saves writing S and reading it back; at narrower disparity ranges its per-pixel reductions would weigh
// more than the sweep's own arithmetic (measured at 128 disparities: +0.08 ms per frame in the sweep against 0.06 ms here),
// and the kernel below reads S instead.  Both write the same buffers (sgbm_wta_buffers_dev).
// disp2 (the right-view disparity) is a per-row argmin with "first writer in descending x wins ties": realised as atomicMin
// on key = minS << 16 | (0xFFFF - x), from which the winning d follows as x + minX1 - x2 - minD (x2 = x + minX1 - d -
// minD), so no second array is needed.
// ---- WTA from S in memory: one warp per pixel ------------------------------------------------------------------------
#define WTA_PPW 8  // pixels per warp: their loads are issued together (memory-level parallelism), then processed in turn
// ur_magic = ceil(2^32 / (100 - ur)) (0 when the divisor is 1): n / (100 - ur) == umulhi(n, ur_magic) for n < 2^25
template <int NW>
__global__ void __launch_bounds__(256)
wta_kernel(const uint32_t* __restrict__ S, SgbmGeom g, uint32_t ur_magic, int16_t* __restrict__ raw, uint32_t* __restrict__ disp2key,
           int* __restrict__ sat_flag)
{
    __shared__ __align__(16) uint32_t sh[8][NW * 32];
    const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
    const long long npix = (long long)g.W1 * g.H;
    const long long pix0 = ((long long)blockIdx.x * 8 + wib) * WTA_PPW;
    if (pix0 >= npix) return;
    uint32_t wall[WTA_PPW][NW];
#pragma unroll
    for (int q = 0; q < WTA_PPW; ++q) {
        const long long pix = min(pix0 + q, npix - 1);
        const uint32_t* Sp = S + (size_t)pix * (NW * 32) + lane * NW;
        if (NW == 4) { uint4 v = __ldcs((const uint4*)Sp); wall[q][0] = v.x; wall[q][1 % NW] = v.y; wall[q][2 % NW] = v.z; wall[q][3 % NW] = v.w; }
        else if (NW == 2) { uint2 v = __ldcs((const uint2*)Sp); wall[q][0] = v.x; wall[q][1 % NW] = v.y; }
        else {
#pragma unroll
            for (int j = 0; j < NW; ++j) wall[q][j] = __ldcs(Sp + j);
        }
    }
    // split layout (common.cuh): word w of a pixel holds disparity w (low half) and w + Dp/2 (high half)
    constexpr uint32_t HALF = NW * 32;
    uint32_t padbits[NW];
#pragma unroll
    for (int j = 0; j < NW; ++j) {
        const uint32_t d0 = (uint32_t)(lane * NW + j);
        padbits[j] = (d0 >= (uint32_t)g.D ? 0xffffu : 0u) | (d0 + HALF >= (uint32_t)g.D ? 0xffff0000u : 0u);
    }
    auto hw_index = [](int d) -> int { return d >= (int)HALF ? 2 * (d - (int)HALF) + 1 : 2 * d; };  // halfword of disparity d in the staged copy
    uint32_t* mysh = sh[wib];
    uint16_t* hw = (uint16_t*)mysh;
    // lane q keeps the outcome of pixel q; the divisions, the atomic and the store happen once per warp at the end
    int my_best = -1, my_minS = 0, my_sm = 0, my_sp = 0;
#pragma unroll
    for (int q = 0; q < WTA_PPW; ++q) {
        if (pix0 + q >= npix) break;
        const uint32_t (&w)[NW] = wall[q];
        // first argmin: min over (S << 16 | d); padding (d >= D) holds 32767 and never beats a real minimum below it
        uint32_t key = 0xffffffffu;
        __syncwarp();  // the previous pixel's staged copy has been read by every lane
#pragma unroll
        for (int j = 0; j < NW; ++j) {
            const uint32_t d0 = (uint32_t)(lane * NW + j);
            key = min(key, min(w[j] * 65536u + d0, (w[j] & 0xffff0000u) | (d0 + HALF)));
            // staged copy for the uniqueness test: padding is taken out of it (0xFFFF never counts)
            mysh[lane * NW + j] = w[j] | padbits[j];
        }
        key = __reduce_min_sync(0xffffffffu, key);
        const int minS = (int)(key >> 16), best = (int)(key & 0xffffu);
        // "minus (min Lp + P2)" form of the recurrence (g.K44 = directions * P2): S here is that form's S plus K44 wherever
        // no sum saturated.  A saturated sum (32767 here, at least 32767 - K44 there) only matters where it takes part in a
        // decision: as the minimum, as a neighbour of the minimum in the sub-pixel fit, or when the uniqueness threshold
        // reaches it.  Those cases are reported (sat_flag), never guessed around; everywhere else the result is exact.
        bool satp = false;
        if (g.K44) {
            uint32_t mx = 0;
#pragma unroll
            for (int j = 0; j < NW; ++j) mx = __vmaxu2(mx, w[j] & ~padbits[j]);
            satp = __reduce_max_sync(0xffffffffu, max(mx & 0xffffu, mx >> 16)) >= 32767u;
        }
        if (minS >= 32767) {   // every cost saturated: the reference leaves / writes INV here
            if (g.K44 && lane == 0) *sat_flag = 1;
            continue;
        }
        __syncwarp();
        // uniqueness: reject if some d outside best-1..best+1 has S*(100-ur) < minS*100, i.e. S <= thr.
        // Lane q reads the two neighbours it needs for the sub-pixel fit, blanks best-1..best+1 in the staged copy,
        // and the warp takes one packed minimum over what is left.
        int sm = 0, sp = 0;
        if (lane == q) {
            if (best > 0) { sm = hw[hw_index(best - 1)]; hw[hw_index(best - 1)] = 0xffffu; }
            if (best < g.D - 1) { sp = hw[hw_index(best + 1)]; hw[hw_index(best + 1)] = 0xffffu; }
            hw[hw_index(best)] = 0xffffu;
        }
        __syncwarp();
        uint32_t m = 0xffffffffu;
#pragma unroll
        for (int j = 0; j < NW; ++j) m = __vminu2(m, mysh[lane * NW + j]);
        const uint32_t mn = __reduce_min_sync(0xffffffffu, min(m & 0xffffu, m >> 16));
        // S(d) * (100 - ur) < minS * 100 in the reference's own S; with S' = S - K44 (the other form of the recurrence,
        // K44 = 0 for OpenCV 4.13's form) that is S(d) * (100 - ur) < minS * 100 - K44 * ur on the S held here
        const int rhs = minS * 100 - g.K44 * g.ur;
        const uint32_t num = (uint32_t)(rhs - 1);
        const uint32_t thr = rhs > 0 ? (ur_magic ? __umulhi(num, ur_magic) : num) : 0xffffffffu;  // ur <= 99 (mvs_sgbm_check)
        if (satp && lane == q && ((rhs > 0 && thr >= 32767u) || (best > 0 && sm >= 32767) || (best < g.D - 1 && sp >= 32767))) *sat_flag = 1;
        if (rhs > 0 && mn != 0xffffu && mn <= thr) continue;
        if (lane == q) { my_best = best; my_minS = minS; my_sm = sm; my_sp = sp; }
    }
    if (lane < WTA_PPW && my_best >= 0) {
        int y = (int)(pix0 / g.W1), x = (int)(pix0 - (long long)y * g.W1) + lane;
        while (x >= g.W1) { x -= g.W1; ++y; }
        const int x2 = x + g.minX1 - my_best - g.minD;
        atomicMin(disp2key + (size_t)y * g.W + x2, ((uint32_t)my_minS << 16) | (uint32_t)(0xffff - x));
        int d = my_best;
        if (0 < d && d < g.D - 1) {
            int den = max(my_sm + my_sp - 2 * my_minS, 1);
            d = d * 16 + ((my_sm - my_sp) * 16 + den) / (den * 2);
        } else
            d *= 16;
        raw[(size_t)y * g.W + x + g.minX1] = (int16_t)(d + g.minD * 16);
    }
}

__device__ __forceinline__ void wta_init_body(int16_t* raw, uint32_t* disp2key, size_t n, int inv)
{
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    raw[i] = (int16_t)inv;
    disp2key[i] = 0x7fffffffu;
}

// left-right check (A.2), in place on raw
__device__ __forceinline__ void lr_check_body(int16_t* __restrict__ raw, const uint32_t* __restrict__ disp2key, const SgbmGeom& g)
{
    int x = blockIdx.x * blockDim.x + threadIdx.x + g.minX1, y = blockIdx.y;
    if (x >= g.minX1 + g.W1) return;
    int d1 = raw[(size_t)y * g.W + x];
    if (d1 == g.INV) return;
    int dl = d1 >> 4, dh = (d1 + 15) >> 4;
    int xl = x - dl, xh = x - dh;
    const uint32_t* row = disp2key + (size_t)y * g.W;
    auto disp2 = [&](int xx) -> int {
        uint32_t k = row[xx];
        if (k == 0x7fffffffu) return g.INV;  // never written: keeps INVALID_DISP_SCALED like the reference
        int xw = 0xffff - (int)(k & 0xffffu);
        return xw + g.minX1 - xx;  // = d + minD of the winning left pixel
    };
    bool fl = false, fh = false;
    if (0 <= xl && xl < g.W) { int v = disp2(xl); fl = v >= g.minD && abs(v - dl) > g.d12; }
    if (0 <= xh && xh < g.W) { int v = disp2(xh); fh = v >= g.minD && abs(v - dh) > g.d12; }
    if (fl && fh) raw[(size_t)y * g.W + x] = (int16_t)g.INV;
}

// ---- 3x3 median on s16, replicate border (cv::medianBlur) ---------------------------------------------
__device__ __forceinline__ void cswap(int& a, int& b) { int t = min(a, b); b = max(a, b); a = t; }
__device__ __forceinline__ void median3_body(const int16_t* __restrict__ src, int W, int H, int16_t* __restrict__ dst)
{
    int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y;
    if (x >= W) return;
    int v[9];
#pragma unroll
    for (int dy = -1; dy <= 1; ++dy) {
        const int16_t* r = src + (size_t)min(max(y + dy, 0), H - 1) * W;
#pragma unroll
        for (int dx = -1; dx <= 1; ++dx) v[(dy + 1) * 3 + dx + 1] = r[min(max(x + dx, 0), W - 1)];
    }
    // 19-exchange median-of-9 network
    cswap(v[1], v[2]); cswap(v[4], v[5]); cswap(v[7], v[8]); cswap(v[0], v[1]); cswap(v[3], v[4]); cswap(v[6], v[7]);
    cswap(v[1], v[2]); cswap(v[4], v[5]); cswap(v[7], v[8]); cswap(v[0], v[3]); cswap(v[5], v[8]); cswap(v[4], v[7]);
    cswap(v[3], v[6]); cswap(v[1], v[4]); cswap(v[2], v[5]); cswap(v[4], v[7]); cswap(v[4], v[2]); cswap(v[6], v[4]);
    cswap(v[4], v[2]);
    dst[(size_t)y * W + x] = (int16_t)v[4];
}

// ---- speckle filter (cv::filterSpeckles): 4-connected components, sizes <= maxSpeckleSize removed -----------
// Run-based union-find: (1) every horizontal run of connected pixels is labelled with its first pixel and its
// length is counted there (one warp per row, ballots); (2) vertically connected runs are merged with lock-free
// unions (atomicMin hooking, path halving), skipping pixels whose left neighbour already made the same merge;
// (3) every run adds its length to its component's root; (4) small components are erased.  Component sizes do
// not depend on traversal order, so the result is bit-exact with the reference's flood fill.
__device__ __forceinline__ bool sp_conn(int a, int b, int newVal, int maxDiff)
{
    return a != newVal && b != newVal && abs(a - b) <= maxDiff;
}
__device__ __forceinline__ int uf_find(int* L, int i)
{
    int p = L[i];
    while (p != i) {
        int gp = L[p];
        if (gp != p) L[i] = gp;  // path halving: racy, but only ever re-points a node to one of its ancestors
        i = p;
        p = gp;
    }
    return i;
}
__device__ __forceinline__ void uf_union(int* L, int a, int b)
{
    for (int guard = 0; guard < (1 << 22); ++guard) {
        a = uf_find(L, a);
        b = uf_find(L, b);
        if (a == b) return;
        if (a < b) { int t = a; a = b; b = t; }  // hook the larger root under the smaller
        int old = atomicMin(L + a, b);
        if (old == a) return;
        a = old;
    }
}
// one warp per row: L[i] = index of the first pixel of i's run, cnt[run start] = run length
__device__ __forceinline__ void
speckle_runs_body(const int16_t* __restrict__ img, int W, int H, int newVal, int maxDiff, int* __restrict__ L,
                    int* __restrict__ cnt)
{
    const int lane = threadIdx.x & 31;
    const int y = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (y >= H) return;
    const int16_t* row = img + (size_t)y * W;
    int carry = -1;          // run start (x) of the previous chunk's last pixel, -1 if that pixel is invalid
    int prev_last = newVal;  // value of the previous chunk's last pixel
    for (int x0 = 0; x0 < W; x0 += 32) {
        const int x = x0 + lane;
        const int v = x < W ? row[x] : newVal;
        int left = __shfl_up_sync(0xffffffffu, v, 1);
        if (lane == 0) left = prev_last;
        const bool valid = v != newVal;
        const bool link = valid && x > 0 && sp_conn(v, left, newVal, maxDiff);
        const unsigned starts = __ballot_sync(0xffffffffu, valid && !link);  // lanes where a run begins
        const unsigned vmask = __ballot_sync(0xffffffffu, valid);
        const unsigned below = starts & (0xffffffffu >> (31 - lane));        // run starts at or before this lane
        const int start = below ? x0 + (31 - __clz(below)) : carry;
        const int i = y * W + x;
        if (x < W) L[i] = valid ? y * W + start : i;
        // length of the part of each run inside this chunk, counted by the run's first lane in the chunk
        if (valid && (!link || lane == 0)) {
            unsigned stop = ~vmask | starts;                                  // a run ends before the next start / hole
            stop &= (lane == 31) ? 0u : (0xffffffffu << (lane + 1));
            const int end = stop ? __ffs(stop) - 1 : 32;
            const int len = end - lane;
            atomicAdd(cnt + y * W + start, len);  // cnt is zeroed before the kernel; a run may span several chunks
        }
        carry = __shfl_sync(0xffffffffu, valid ? start : -1, 31);
        prev_last = __shfl_sync(0xffffffffu, v, 31);
    }
}
__device__ __forceinline__ void speckle_merge_body(const int16_t* __restrict__ img, int W, int H, int newVal, int maxDiff, int* L)
{
    const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y + 1;
    if (x >= W || y >= H) return;
    const int i = y * W + x;
    const int v = img[i], u = img[i - W];
    if (!sp_conn(v, u, newVal, maxDiff)) return;
    if (x > 0) {  // the left neighbour already merges the same two runs
        const int vl = img[i - 1], ul = img[i - W - 1];
        if (sp_conn(v, vl, newVal, maxDiff) && sp_conn(u, ul, newVal, maxDiff) && sp_conn(vl, ul, newVal, maxDiff)) return;
    }
    uf_union(L, L[i], L[i - W]);
}
// every run start adds its run length to its component's root and then points straight at it
__device__ __forceinline__ void speckle_resolve_body(const int16_t* __restrict__ img, int W, int H, int newVal, int maxDiff, int* L, int* cnt)
{
    const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y;
    if (x >= W) return;
    const int i = y * W + x;
    const int v = img[i];
    if (v == newVal) return;
    if (x > 0 && sp_conn(v, img[i - 1], newVal, maxDiff)) return;  // not a run start
    int root = i;  // read-only walk: this kernel's only writes to L are final root pointers
    for (int q = L[root]; q != root; q = L[root]) root = q;
    if (root != i) {
        atomicAdd(cnt + root, cnt[i]);
        L[i] = root;
    }
}
__device__ __forceinline__ void speckle_apply_body(int16_t* img, size_t n, int newVal, int maxSize, const int* __restrict__ L,
                                                   const int* __restrict__ cnt)
{
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n || img[i] == newVal) return;
    const int root = L[L[i]];  // pixel -> run start -> component root (flattened by the resolve kernel)
    if (cnt[root] <= maxSize) img[i] = (int16_t)newVal;
}

__device__ __forceinline__ void disp_to_float_body(const int16_t* __restrict__ d, size_t n, float* __restrict__ out)
{
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float v = fabsf((float)d[i]) / 16.0f;
    out[i] = v > 150.0f ? 0.0f : v;
}

// ---- launchable forms: one image, or the images of a whole batch (blockIdx.z = frame) ----------------------------------
__global__ void disp_to_float_kernel(const int16_t* __restrict__ d, size_t n, float* __restrict__ out) { disp_to_float_body(d, n, out); }

// Batched: raw / disp2key / L / cnt of frame z are slices of one buffer (stride n = W*H); the disparity images are the
// callers' own buffers, handed over as a pointer table.
struct PostBatch {
    int16_t* disp16[MVS_MAX_BATCH];
    float* dispf[MVS_MAX_BATCH];
};
__global__ void wta_init_batch_kernel(int16_t* raw, uint32_t* disp2key, size_t n, int inv)
{
    wta_init_body(raw + blockIdx.z * n, disp2key + blockIdx.z * n, n, inv);
}
__global__ void lr_check_batch_kernel(int16_t* __restrict__ raw, const uint32_t* __restrict__ disp2key, SgbmGeom g)
{
    const size_t n = (size_t)g.W * g.H;
    lr_check_body(raw + blockIdx.z * n, disp2key + blockIdx.z * n, g);
}
__global__ void median3_batch_kernel(const int16_t* __restrict__ raw, int W, int H, PostBatch b)
{
    median3_body(raw + (size_t)blockIdx.z * W * H, W, H, b.disp16[blockIdx.z]);
}
__global__ void __launch_bounds__(256)
speckle_runs_batch_kernel(PostBatch b, int W, int H, int newVal, int maxDiff, int* __restrict__ L, int* __restrict__ cnt)
{
    const size_t n = (size_t)W * H;
    speckle_runs_body(b.disp16[blockIdx.z], W, H, newVal, maxDiff, L + blockIdx.z * n, cnt + blockIdx.z * n);
}
__global__ void speckle_merge_batch_kernel(PostBatch b, int W, int H, int newVal, int maxDiff, int* L)
{
    speckle_merge_body(b.disp16[blockIdx.z], W, H, newVal, maxDiff, L + (size_t)blockIdx.z * W * H);
}
__global__ void speckle_resolve_batch_kernel(PostBatch b, int W, int H, int newVal, int maxDiff, int* L, int* cnt)
{
    const size_t n = (size_t)W * H;
    speckle_resolve_body(b.disp16[blockIdx.z], W, H, newVal, maxDiff, L + blockIdx.z * n, cnt + blockIdx.z * n);
}
__global__ void speckle_apply_batch_kernel(PostBatch b, size_t n, int newVal, int maxSize, const int* __restrict__ L,
                                           const int* __restrict__ cnt)
{
    speckle_apply_body(b.disp16[blockIdx.z], n, newVal, maxSize, L + blockIdx.z * n, cnt + blockIdx.z * n);
}
__global__ void disp_to_float_batch_kernel(PostBatch b, size_t n) { disp_to_float_body(b.disp16[blockIdx.z], n, b.dispf[blockIdx.z]); }

// Buffers of the winner-take-all step for a batch of F frames, set `sset` (the frame path alternates two sets: the second
// half of batch k on the tail stream reads one while the sweeps of batch k + 1 fill the other), initialised.
int sgbm_wta_buffers_dev(mvs_ctx* ctx, const SgbmGeom& g, int F, int sset, SgbmWtaOut* out)
{
    const size_t n = (size_t)g.W * g.H;
    int16_t* raw = (int16_t*)arena_get(ctx, sset ? "sgbm.batch.raw.1" : "sgbm.batch.raw", sizeof(int16_t) * n * MVS_MAX_BATCH);
    uint32_t* d2 = (uint32_t*)arena_get(ctx, sset ? "sgbm.batch.disp2key.1" : "sgbm.batch.disp2key", sizeof(uint32_t) * n * MVS_MAX_BATCH);
    if (!raw || !d2) return MVS_ERR_NOMEM;
    memset(out, 0, sizeof(*out));
    for (int f = 0; f < F; ++f) { out->raw[f] = raw + n * f; out->disp2key[f] = d2 + n * f; }
    wta_init_batch_kernel<<<dim3(cdiv(n, 256), 1, F), 256, 0, ctx->stream>>>(raw, d2, n, g.INV);
    MVS_LAUNCH_CHECK(ctx, "wta_init_batch_kernel");
    return MVS_OK;
}

// The winner-take-all step from S volumes in memory (the form used up to 128 disparities), one launch per frame.
int sgbm_wta_from_s_batch_dev(mvs_ctx* ctx, const SgbmGeom& g, int F, const SgbmWtaOut& wta)
{
    cudaStream_t s = ctx->stream;
    MVS_ARENA(ctx, int, sat, "sgbm.sat44", 1);
    const int blocks = cdiv((long long)g.W1 * g.H, 8 * WTA_PPW);
    const uint32_t udiv = (uint32_t)(100 - g.ur);
    const uint32_t ur_magic = udiv <= 1 ? 0u : (uint32_t)(((1ull << 32) + udiv - 1) / udiv);
    for (int f = 0; f < F; ++f) {
        const uint32_t* S = (const uint32_t*)wta.S[f];
        switch (g.Dp / 64) {
        case 1: wta_kernel<1><<<blocks, 256, 0, s>>>(S, g, ur_magic, wta.raw[f], wta.disp2key[f], sat); break;
        case 2: wta_kernel<2><<<blocks, 256, 0, s>>>(S, g, ur_magic, wta.raw[f], wta.disp2key[f], sat); break;
        case 3: wta_kernel<3><<<blocks, 256, 0, s>>>(S, g, ur_magic, wta.raw[f], wta.disp2key[f], sat); break;
        case 4: wta_kernel<4><<<blocks, 256, 0, s>>>(S, g, ur_magic, wta.raw[f], wta.disp2key[f], sat); break;
        default: return mvs_fail(ctx, MVS_ERR_UNSUPPORTED, "wta: Dp=%d", g.Dp);
        }
        MVS_LAUNCH_CHECK(ctx, "wta_kernel");
    }
    return MVS_OK;
}

// Left-right check + median + speckle filter for the F frames of a batch whose raw disparities and right-view keys the
// backward sweep has written (buffers of sgbm_wta_buffers_dev): the small per-image kernels run once per batch
// (blockIdx.z = frame).  raw16_tap (F == 1): copy of the raw disparities after the left-right check.  The caller has opened
// the stage timer "wta"; with F == 1 it is closed behind the left-right check and the rest is timed as "post".
int sgbm_lr_post_batch_dev(mvs_ctx* ctx, const SgbmGeom& g, const mvs_sgbm_params* p, int F, const SgbmWtaOut& wta, int16_t* const* disp16,
                           int16_t* raw16_tap)
{
    cudaStream_t s = ctx->stream;
    const size_t n = (size_t)g.W * g.H;
    int16_t* raw = wta.raw[0];          // frames are consecutive slices
    uint32_t* d2 = wta.disp2key[0];
    PostBatch b;
    memset(&b, 0, sizeof(b));
    for (int f = 0; f < F; ++f) b.disp16[f] = disp16[f];
    lr_check_batch_kernel<<<dim3(cdiv(g.W1, 128), g.H, F), 128, 0, s>>>(raw, d2, g);
    MVS_LAUNCH_CHECK(ctx, "lr_check_batch_kernel");
    if (raw16_tap) MVS_CUDA(ctx, cudaMemcpyAsync(raw16_tap, raw, n * sizeof(int16_t), cudaMemcpyDeviceToDevice, s));
    if (F == 1) { stage_end(ctx, "wta"); stage_begin(ctx, "post"); }
    median3_batch_kernel<<<dim3(cdiv(g.W, 128), g.H, F), 128, 0, s>>>(raw, g.W, g.H, b);
    MVS_LAUNCH_CHECK(ctx, "median3_batch_kernel");
    if (p->speckle_win > 0) {
        MVS_ARENA(ctx, int, L, "sgbm.batch.speckle.label", n * F);
        MVS_ARENA(ctx, int, cnt, "sgbm.batch.speckle.count", n * F);
        const int maxDiff = 16 * p->speckle_range;
        MVS_CUDA(ctx, cudaMemsetAsync(cnt, 0, sizeof(int) * n * F, s));
        speckle_runs_batch_kernel<<<dim3(cdiv(g.H, 8), 1, F), 256, 0, s>>>(b, g.W, g.H, g.INV, maxDiff, L, cnt);
        if (g.H > 1) speckle_merge_batch_kernel<<<dim3(cdiv(g.W, 128), g.H - 1, F), 128, 0, s>>>(b, g.W, g.H, g.INV, maxDiff, L);
        speckle_resolve_batch_kernel<<<dim3(cdiv(g.W, 128), g.H, F), 128, 0, s>>>(b, g.W, g.H, g.INV, maxDiff, L, cnt);
        speckle_apply_batch_kernel<<<dim3(cdiv(n, 256), 1, F), 256, 0, s>>>(b, n, g.INV, p->speckle_win, L, cnt);
        MVS_LAUNCH_CHECK_N(ctx, "speckle batch kernels", g.H > 1 ? 4 : 3);
    }
    if (F == 1) stage_end(ctx, "post");
    return MVS_OK;
}

int disp_to_float_batch_dev(mvs_ctx* ctx, int F, int16_t* const* d16, int w, int h, float* const* out)
{
    PostBatch b;
    memset(&b, 0, sizeof(b));
    for (int f = 0; f < F; ++f) { b.disp16[f] = d16[f]; b.dispf[f] = out[f]; }
    const size_t n = (size_t)w * h;
    disp_to_float_batch_kernel<<<dim3(cdiv(n, 256), 1, F), 256, 0, ctx->stream>>>(b, n);
    MVS_LAUNCH_CHECK(ctx, "disp_to_float_batch_kernel");
    return MVS_OK;
}

int disp_to_float_dev(mvs_ctx* ctx, const int16_t* d16, int w, int h, float* out)
{
    size_t n = (size_t)w * h;
    disp_to_float_kernel<<<cdiv(n, 256), 256, 0, ctx->stream>>>(d16, n, out);
    MVS_LAUNCH_CHECK(ctx, "disp_to_float_kernel");
    return MVS_OK;
}

// ---- a8: disp_map_update merge with the scene's reference disparity (pipeline/nodes/disp_map_update.cpp:98-113) ------
// dst[x] = ref[x] where ref is finite, the pixel has a super-pixel label, and (the super-pixel is not flagged
// still_cuttoff, or dst is not finite, or dst <= 0).  Labels beyond the table are skipped (the reference would
// index past its vector there).
__global__ void disp_merge_ref_kernel(float* __restrict__ dst, const float* __restrict__ ref, const int32_t* __restrict__ labels,
                                      const uint8_t* __restrict__ still_cutoff, int n_labels, size_t n)
{
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float s = ref[i];
    const int label = labels[i];
    if (!isfinite(s) || label < 0 || label >= n_labels) return;
    const float d = dst[i];
    if (!still_cutoff[label] || !isfinite(d) || d <= 0.0f) dst[i] = s;
}

int disp_merge_ref_dev(mvs_ctx* ctx, float* disp, const float* ref, const int32_t* labels, const uint8_t* still_cutoff,
                       int n_labels, int w, int h)
{
    size_t n = (size_t)w * h;
    disp_merge_ref_kernel<<<cdiv(n, 256), 256, 0, ctx->stream>>>(disp, ref, labels, still_cutoff, n_labels, n);
    MVS_LAUNCH_CHECK(ctx, "disp_merge_ref_kernel");
    return MVS_OK;
}

extern "C" int mvs_disp_merge_ref(mvs_ctx* ctx, float* disp, size_t dstep, const float* ref, size_t rstep, const int32_t* labels,
                                  size_t lstep, const uint8_t* still_cutoff, int n_labels, int w, int h)
{
    if (!ctx) return MVS_ERR_ARG;
    if (!disp || !ref || !labels || (n_labels > 0 && !still_cutoff) || n_labels < 0 || w <= 0 || h <= 0 || dstep < (size_t)w * 4 ||
        rstep < (size_t)w * 4 || lstep < (size_t)w * 4)
        return mvs_fail(ctx, MVS_ERR_ARG, "mvs_disp_merge_ref: bad argument");
    MVS_CUDA(ctx, cudaSetDevice(ctx->device));
    cudaStream_t s = ctx->stream;
    size_t n = (size_t)w * h;
    MVS_ARENA(ctx, float, dd, "merge.api.disp", n);
    MVS_ARENA(ctx, float, dr, "merge.api.ref", n);
    MVS_ARENA(ctx, int32_t, dl, "merge.api.labels", n);
    MVS_ARENA(ctx, uint8_t, dc, "merge.api.cutoff", (size_t)n_labels + 1);
    MVS_CUDA(ctx, cudaMemcpy2DAsync(dd, (size_t)w * 4, disp, dstep, (size_t)w * 4, h, cudaMemcpyHostToDevice, s));
    MVS_CUDA(ctx, cudaMemcpy2DAsync(dr, (size_t)w * 4, ref, rstep, (size_t)w * 4, h, cudaMemcpyHostToDevice, s));
    MVS_CUDA(ctx, cudaMemcpy2DAsync(dl, (size_t)w * 4, labels, lstep, (size_t)w * 4, h, cudaMemcpyHostToDevice, s));
    if (n_labels > 0) MVS_CUDA(ctx, cudaMemcpyAsync(dc, still_cutoff, (size_t)n_labels, cudaMemcpyHostToDevice, s));
    int rc = disp_merge_ref_dev(ctx, dd, dr, dl, dc, n_labels, w, h);
    if (rc) return rc;
    MVS_CUDA(ctx, cudaMemcpy2DAsync(disp, dstep, dd, (size_t)w * 4, (size_t)w * 4, h, cudaMemcpyDeviceToHost, s));
    MVS_CUDA(ctx, cudaStreamSynchronize(s));
    return MVS_OK;
}
