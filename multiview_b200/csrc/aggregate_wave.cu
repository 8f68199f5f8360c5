// a3 (iv), fused form: one sweep of the SGM aggregation carries ALL directions of that sweep at once.
//
// Forward sweep (y up, x up): L0 from (x-1,y), L1 from (x-1,y-1), L2 from (x,y-1), L3 from (x+1,y-1);
// the backward sweep is the same recurrence on the image flipped in x and y.  Reading C once
// and writing S once per sweep is the algorithmic minimum (8 B/cell for both sweeps).
//
// Parallelisation: pixel (x,y) depends on (x-1,y) and on three pixels of row y-1, so all pixels with the same
// t = x + 2y are independent (a wavefront).  A CTA owns a band of R consecutive rows; warp r walks row r of the
// band, two pixels behind warp r-1, one pixel per step, all warps meeting at one __syncthreads per step.
//   * the disparity vector of a pixel is spread over the warp's 32 lanes (2*NW disparities per lane, u16x2);
//   * L0 stays in registers; L1/L2/L3 of the row above come from a shared-memory ring written by warp r-1
//     (sentinel words at both ends give the d-1 / d+1 neighbours of lanes 0 and 31 without branches);
//   * bands are chained through global memory: the last warp of a band streams its L1/L2/L3 rows to a
//     boundary buffer and publishes its progress; a helper warp of the next band waits for that progress and
//     copies the boundary row into the ring with cp.async a few pixels ahead of its consumers;
//   * bands are handed out by an atomic ticket in dependency order, so the band a CTA waits for is always
//     running or finished (no co-residency assumption); every wait is bounded and raises an error flag.
#include "common.cuh"

#define SENT2 0x7FFF7FFFu
#define WAVE_R 16            // rows (compute warps) per band
#define WAVE_THREADS ((WAVE_R + 1) * 32)
#define WAVE_PUBLISH 8       // progress is published every this many pixels
#define WAVE_SPIN_LIMIT (1 << 20)
#define WAVE_CTL_PAD 16

// A launch sweeps a batch of F camera-frames of equal geometry: tickets interleave the frames (frame = ticket % F,
// band = ticket / F) so that all frames advance together and their wavefronts fill the machine.
struct WaveFrame {
    const uint32_t* C;
    const uint32_t* Sin;   // backward sweep: S after the forward sweep (nullptr in the forward sweep)
    uint32_t* Sout;
    uint32_t* bndL;        // [nbands][W1][3][wpp] boundary path costs
    uint32_t* bndMin;      // [nbands][W1][2]      boundary minima (3 x u16 + pad)
};
struct WaveParams {
    WaveFrame f[MVS_MAX_BATCH];
    int* progress;         // [F][nbands] pixels finished by each band's last row
    int* ticket;           // work counter
    int* error;            // set when a bounded wait expires
    int W1, H, D, P1, P2, nbands, flip, F;
};

__device__ __forceinline__ uint32_t wdup16(uint32_t v) { return v | (v << 16); }
__device__ __forceinline__ int ld_acquire(const int* p)
{
    int v;
    asm volatile("ld.acquire.gpu.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_release(int* p, int v)
{
    asm volatile("st.release.gpu.global.s32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
template <int BYTES>
__device__ __forceinline__ void cp_async(void* smem, const void* gmem)
{
    unsigned s = (unsigned)__cvta_generic_to_shared(smem);
    if (BYTES == 16) asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(s), "l"(gmem) : "memory");
    else if (BYTES == 8) asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(s), "l"(gmem) : "memory");
    else asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(s), "l"(gmem) : "memory");
}
#define WAVE_L2_AHEAD 8
__device__ __forceinline__ void prefetch_l2(const void* p) { asm volatile("prefetch.global.L2 [%0];" ::"l"(p)); }
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

template <int NW>
__device__ __forceinline__ void ldw(const uint32_t* p, uint32_t (&w)[NW])
{
    if (NW == 4) { uint4 v = *(const uint4*)p; w[0] = v.x; w[1 % NW] = v.y; w[2 % NW] = v.z; w[3 % NW] = v.w; }
    else if (NW == 2) { uint2 v = *(const uint2*)p; w[0] = v.x; w[1 % NW] = v.y; }
    else {
#pragma unroll
        for (int j = 0; j < NW; ++j) w[j] = p[j];
    }
}
template <int NW>
__device__ __forceinline__ void stw(uint32_t* p, const uint32_t (&w)[NW])
{
    if (NW == 4) *(uint4*)p = make_uint4(w[0], w[1 % NW], w[2 % NW], w[3 % NW]);
    else if (NW == 2) *(uint2*)p = make_uint2(w[0], w[1 % NW]);
    else {
#pragma unroll
        for (int j = 0; j < NW; ++j) p[j] = w[j];
    }
}

// One path step.  a[] = Lp (this lane's words), left/right = neighbouring words (they supply d-1 of the lane's
// first and d+1 of its last disparity), minLp2 = min(Lp) in both halves.  Writes L to out[], returns min_d L.
template <int NW, bool PAD>
__device__ __forceinline__ uint32_t step_dir(const uint32_t (&a)[NW], uint32_t left, uint32_t right, uint32_t minLp2,
                                             const uint32_t (&Cw)[NW], uint32_t P1x2, uint32_t P2x2,
                                             const uint32_t (&padmask)[NW], uint32_t (&out)[NW])
{
    const uint32_t delta = minLp2 + P2x2;
    uint32_t m = SENT2;
#pragma unroll
    for (int j = 0; j < NW; ++j) {
        uint32_t lo = __byte_perm(j == 0 ? left : a[j == 0 ? 0 : j - 1], a[j], 0x5432);            // Lp[d-1]
        uint32_t hi = __byte_perm(a[j], j == NW - 1 ? right : a[j == NW - 1 ? j : j + 1], 0x5432);  // Lp[d+1]
        uint32_t t = __viaddmin_u16x2(__vminu2(lo, hi), P1x2, a[j]);
        t = __vminu2(t, delta);
        uint32_t l = Cw[j] + t - minLp2;  // per half: t >= minLp and C + (t - minLp) <= 32767, no borrow / carry
        if (PAD) l = (l & ~padmask[j]) | (SENT2 & padmask[j]);
        out[j] = l;
        m = __vminu2(m, l);
    }
    uint32_t mm = min(m & 0xffffu, m >> 16);
    return __reduce_min_sync(0xffffffffu, mm);
}

// Shared-memory ring geometry (32-bit words).  A slot holds one pixel's path costs with 4 pad words on each
// side; pad word 3 and pad word SLOT-4 are the d = -1 / d = D sentinels.
template <int NW>
struct Ring {
    static constexpr int WPP = NW * 32;        // words per pixel
    static constexpr int SLOT = WPP + 8;
    static constexpr int ROW_SLOTS = 10;       // compute rows: L1 x4, L2 x4, L3 x2
    static constexpr int HELP_SLOTS = 24;      // helper row: 3 directions x 8
    static constexpr int NSLOTS = WAVE_R * ROW_SLOTS + HELP_SLOTS;
    static constexpr int WORDS = NSLOTS * SLOT;
    // ring row rr in [0, R] (0 = row above the band, filled by the helper), direction 1..3, pixel x
    __device__ static __forceinline__ int slot(int rr, int dir, int x)
    {
        if (rr == 0) return (dir - 1) * 8 + (x & 7);
        int base = HELP_SLOTS + (rr - 1) * ROW_SLOTS;
        return dir == 1 ? base + (x & 3) : (dir == 2 ? base + 4 + (x & 3) : base + 8 + (x & 1));
    }
};

template <int NW, int DIRMASK, bool PAD>
__global__ void __launch_bounds__(WAVE_THREADS, 1)
wave_kernel(WaveParams p)
{
    using RG = Ring<NW>;
    constexpr bool VERT = (DIRMASK & 0xE) != 0;
    extern __shared__ __align__(16) uint32_t ring[];
    __shared__ uint16_t mins[WAVE_R + 1][3][8];          // [ring row][dir-1][x & 7]; row 0 unused (see hmins)
    __shared__ __align__(8) uint16_t hmins[8][4];        // helper row minima: [x & 7][dir-1]
    __shared__ int s_band;
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const bool helper = wid == WAVE_R;
    const int W1 = p.W1;
    const uint32_t P1x2 = wdup16((uint32_t)p.P1), P2x2 = wdup16((uint32_t)p.P2);

    for (int s = threadIdx.x; s < RG::NSLOTS; s += blockDim.x) {  // sentinels: written once
        ring[s * RG::SLOT + 3] = SENT2;
        ring[s * RG::SLOT + RG::SLOT - 4] = SENT2;
    }
    uint32_t padmask[NW];
#pragma unroll
    for (int j = 0; j < NW; ++j) {
        int d0 = (lane * NW + j) * 2;
        padmask[j] = (d0 >= p.D ? 0xffffu : 0u) | (d0 + 1 >= p.D ? 0xffff0000u : 0u);
    }

    for (;;) {
        __syncthreads();
        if (threadIdx.x == 0) s_band = atomicAdd(p.ticket, 1);
        __syncthreads();
        if (s_band >= p.nbands * p.F) break;
        const int band = s_band / p.F, fi = s_band - band * p.F;
        const WaveFrame& fr = p.f[fi];
        const int ys0 = band * WAVE_R;  // first sweep row of the band
        const int rows = min(WAVE_R, p.H - ys0);
        const bool first_band = band == 0, last_band = band == p.nbands - 1;
        const uint32_t* bL = fr.bndL + (size_t)(first_band ? 0 : band - 1) * W1 * 3 * RG::WPP;  // written by the previous band
        const uint32_t* bM = fr.bndMin + (size_t)(first_band ? 0 : band - 1) * W1 * 2;
        uint32_t* oL = fr.bndL + (size_t)band * W1 * 3 * RG::WPP;
        uint32_t* oM = fr.bndMin + (size_t)band * W1 * 2;
        int* my_progress = p.progress + (size_t)fi * p.nbands + band;
        const int* prev_progress = my_progress - (first_band ? 0 : 1);
        const bool fetch = helper && VERT && !first_band;
        int seen = 0;  // helper: cached progress of the previous band
        bool failed = false;

        // helper: request boundary pixel xp (one cp.async group per call, possibly empty)
        auto helper_issue = [&](int xp) {
            if (xp < W1) {
                const int need = xp + 1;
                if (seen < need && !failed) {
                    int spins = 0;
                    while ((seen = ld_acquire(prev_progress)) < need) {
                        // bounded: a timeout here or anywhere else in the grid ends all waiting (results are then invalid
                        // and the host reports an error), so a logic fault can never hang the GPU
                        if (++spins > WAVE_SPIN_LIMIT) { atomicExch(p.error, 1); failed = true; break; }
                        if ((spins & 1023) == 0 && *(volatile int*)p.error) { failed = true; break; }
                        __nanosleep(32);
                    }
                }
#pragma unroll
                for (int dir = 1; dir <= 3; ++dir) {
                    if (!(DIRMASK & (1 << dir))) continue;
                    uint32_t* dst = ring + RG::slot(0, dir, xp) * RG::SLOT + 4 + lane * NW;
                    const uint32_t* src = bL + ((size_t)xp * 3 + (dir - 1)) * RG::WPP + lane * NW;
                    cp_async<NW == 4 ? 16 : (NW == 2 ? 8 : 4)>(dst, src);
                    if (NW == 3) { cp_async<4>(dst + 1, src + 1); cp_async<4>(dst + 2, src + 2); }
                }
                if (lane == 0) cp_async<8>(&hmins[xp & 7][0], bM + (size_t)xp * 2);
            }
            cp_async_commit();
        };

        if (helper && VERT && first_band) {  // the row above the image: Lr = 0, minLr = 0
            for (int s = 0; s < RG::HELP_SLOTS; ++s)
                for (int k = lane; k < RG::WPP; k += 32) ring[s * RG::SLOT + 4 + k] = 0;
            if (lane < 32) ((uint16_t*)hmins)[lane] = 0;
        }
        if (fetch) {
            for (int xp = 0; xp < 4; ++xp) helper_issue(xp);
            cp_async_wait<2>();  // pixels 0 and 1 have landed
        }

        // per-warp state of a compute row
        const int r = wid;
        const bool active_row = !helper && r < rows;
        uint32_t L0[NW], Cw[NW], Sw[NW];
#pragma unroll
        for (int j = 0; j < NW; ++j) { L0[j] = 0; Sw[j] = 0; Cw[j] = 0; }
        uint32_t minL0 = 0;
        long long off = 0, stride = 0;
        if (active_row) {
            const int ys = ys0 + r;
            const int y = p.flip ? p.H - 1 - ys : ys;
            const int x0 = p.flip ? W1 - 1 : 0;
            off = ((long long)y * W1 + x0) * RG::WPP + lane * NW;
            stride = p.flip ? -(long long)RG::WPP : (long long)RG::WPP;
            ldw<NW>(fr.C + off, Cw);
            if (fr.Sin) ldw<NW>(fr.Sin + off, Sw);
        }
        const int nsteps = W1 + 2 * (rows - 1);
        __syncthreads();

        for (int t = 0; t < nsteps; ++t) {
            if (fetch) {
                helper_issue(t + 4);
                cp_async_wait<2>();  // pixel t+2 (requested two steps ago) is visible after the barrier below
            } else if (active_row) {
                const int x = t - 2 * r;
                if (x >= 0 && x < W1) {
                    uint32_t Cn[NW], Sn[NW];
                    const bool more = x + 1 < W1;
                    if (x + WAVE_L2_AHEAD < W1) {  // pull the volumes into L2 well ahead of the register prefetch
                        prefetch_l2(fr.C + off + WAVE_L2_AHEAD * stride);
                        if (fr.Sin) prefetch_l2(fr.Sin + off + WAVE_L2_AHEAD * stride);
                    }
                    if (more) {  // next pixel's C (and S) before the dependent chain
                        ldw<NW>(fr.C + off + stride, Cn);
                        if (fr.Sin) ldw<NW>(fr.Sin + off + stride, Sn);
                    }
                    uint32_t acc[NW], Lo[NW];
                    if (DIRMASK & 1) {  // direction 0: predecessor (x-1, y), kept in registers
                        uint32_t left = __shfl_up_sync(0xffffffffu, L0[NW - 1], 1);
                        uint32_t right = __shfl_down_sync(0xffffffffu, L0[0], 1);
                        if (lane == 0) left = SENT2;
                        if (lane == 31) right = SENT2;
                        uint32_t mn = step_dir<NW, PAD>(L0, left, right, minL0, Cw, P1x2, P2x2, padmask, Lo);
#pragma unroll
                        for (int j = 0; j < NW; ++j) { L0[j] = Lo[j]; acc[j] = Lo[j]; }
                        minL0 = wdup16(mn);
                    } else {
#pragma unroll
                        for (int j = 0; j < NW; ++j) acc[j] = 0;
                    }
#pragma unroll
                    for (int dir = 1; dir <= 3; ++dir) {  // predecessors in the row above, from the ring
                        if (!(DIRMASK & (1 << dir))) continue;
                        const int xp = dir == 1 ? x - 1 : (dir == 2 ? x : x + 1);
                        uint32_t a[NW], left, right, mp;
                        if (xp >= 0 && xp < W1) {
                            const uint32_t* sl = ring + RG::slot(r, dir, xp) * RG::SLOT + 4 + lane * NW;
                            ldw<NW>(sl, a);
                            left = sl[-1];
                            right = sl[NW];
                            mp = wdup16(r == 0 ? (uint32_t)hmins[xp & 7][dir - 1] : (uint32_t)mins[r][dir - 1][xp & 7]);
                        } else {  // predecessor outside the image: Lr = 0, minLr = 0
#pragma unroll
                            for (int j = 0; j < NW; ++j) a[j] = 0;
                            left = lane == 0 ? SENT2 : 0u;
                            right = lane == 31 ? SENT2 : 0u;
                            mp = 0;
                        }
                        uint32_t mn = step_dir<NW, PAD>(a, left, right, mp, Cw, P1x2, P2x2, padmask, Lo);
                        if (r + 1 < rows) {  // for the row below
                            stw<NW>(ring + RG::slot(r + 1, dir, x) * RG::SLOT + 4 + lane * NW, Lo);
                            if (lane == 0) mins[r + 1][dir - 1][x & 7] = (uint16_t)mn;
                        } else if (!last_band) {  // for the next band
                            stw<NW>(oL + ((size_t)x * 3 + (dir - 1)) * RG::WPP + lane * NW, Lo);
                            if (lane == 0) ((uint16_t*)(oM + (size_t)x * 2))[dir - 1] = (uint16_t)mn;
                        }
#pragma unroll
                        for (int j = 0; j < NW; ++j) acc[j] = __viaddmin_u16x2(acc[j], Lo[j], SENT2);
                    }
                    uint32_t So[NW];
#pragma unroll
                    for (int j = 0; j < NW; ++j) So[j] = fr.Sin ? __viaddmin_u16x2(Sw[j], acc[j], SENT2) : acc[j];
                    stw<NW>(fr.Sout + off, So);
                    if (VERT && r + 1 == rows && !last_band && (((x + 1) % WAVE_PUBLISH) == 0 || x + 1 == W1)) {
                        __threadfence();
                        __syncwarp();
                        if (lane == 0) st_release(my_progress, x + 1);
                    }
                    off += stride;
                    if (more) {
#pragma unroll
                        for (int j = 0; j < NW; ++j) { Cw[j] = Cn[j]; Sw[j] = Sn[j]; }
                    }
                }
            }
            __syncthreads();
        }
        if (fetch) cp_async_wait<0>();
    }
}

template <int NW, int DIRMASK, bool PAD>
static int launch_wave(mvs_ctx* ctx, const WaveParams& p)
{
    const int smem = Ring<NW>::WORDS * 4;
    MVS_CUDA(ctx, cudaFuncSetAttribute(wave_kernel<NW, DIRMASK, PAD>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    int per_sm = 1;
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, wave_kernel<NW, DIRMASK, PAD>, WAVE_THREADS, smem);
    if (per_sm < 1) per_sm = 1;
    int grid = min(p.nbands * p.F, ctx->sm_count * per_sm);
    wave_kernel<NW, DIRMASK, PAD><<<grid, WAVE_THREADS, smem, ctx->stream>>>(p);
    MVS_LAUNCH_CHECK(ctx, "wave_kernel");
    return MVS_OK;
}

template <int DIRMASK>
static int launch_wave_nw(mvs_ctx* ctx, const SgbmGeom& g, const WaveParams& p)
{
    bool pad = g.Dp != g.D;
    switch (g.Dp / 64) {
    case 1: return pad ? launch_wave<1, DIRMASK, true>(ctx, p) : launch_wave<1, DIRMASK, false>(ctx, p);
    case 2: return pad ? launch_wave<2, DIRMASK, true>(ctx, p) : launch_wave<2, DIRMASK, false>(ctx, p);
    case 3: return pad ? launch_wave<3, DIRMASK, true>(ctx, p) : launch_wave<3, DIRMASK, false>(ctx, p);
    case 4: return pad ? launch_wave<4, DIRMASK, true>(ctx, p) : launch_wave<4, DIRMASK, false>(ctx, p);
    }
    return mvs_fail(ctx, MVS_ERR_UNSUPPORTED, "aggregate: Dp=%d", g.Dp);
}

int sgbm_aggregate_wave_batch_dev(mvs_ctx* ctx, const SgbmGeom& g, int F, const int16_t* const* C, int16_t* const* S)
{
    if (F < 1 || F > MVS_MAX_BATCH) return mvs_fail(ctx, MVS_ERR_ARG, "aggregate: batch of %d frames (1..%d)", F, MVS_MAX_BATCH);
    const int nbands = cdiv(g.H, WAVE_R);
    const size_t wpp = (size_t)g.Dp / 2;
    const int ctl_pass = F * nbands + 2 * WAVE_CTL_PAD;
    const size_t bndL_sz = (size_t)nbands * g.W1 * 3 * wpp, bndM_sz = (size_t)nbands * g.W1 * 2;
    MVS_ARENA(ctx, uint32_t, bndL, "sgbm.wave.bndL", bndL_sz * F);
    MVS_ARENA(ctx, uint32_t, bndMin, "sgbm.wave.bndMin", bndM_sz * F);
    MVS_ARENA(ctx, int, ctl, "sgbm.wave.ctl", (size_t)2 * ctl_pass);
    MVS_ARENA(ctx, int, err, "sgbm.wave.err", 1);
    if (!ctx->wave_err_cleared) {
        MVS_CUDA(ctx, cudaMemsetAsync(err, 0, sizeof(int), ctx->stream));
        ctx->wave_err_cleared = true;
    }
    WaveParams p;
    memset(&p, 0, sizeof(p));
    p.error = err; p.F = F;
    p.W1 = g.W1; p.H = g.H; p.D = g.D; p.P1 = g.P1; p.P2 = g.P2; p.nbands = nbands;
    MVS_CUDA(ctx, cudaMemsetAsync(ctl, 0, sizeof(int) * 2 * ctl_pass, ctx->stream));
    int rc;
    // forward sweep
    for (int f = 0; f < F; ++f) {
        p.f[f].C = (const uint32_t*)C[f]; p.f[f].Sin = nullptr; p.f[f].Sout = (uint32_t*)S[f];
        p.f[f].bndL = bndL + bndL_sz * f; p.f[f].bndMin = bndMin + bndM_sz * f;
    }
    p.flip = 0; p.ticket = ctl; p.progress = ctl + WAVE_CTL_PAD;
    rc = g.mode == MVS_MODE_HH4 ? launch_wave_nw<0x5>(ctx, g, p) : launch_wave_nw<0xF>(ctx, g, p);
    if (rc) return rc;
    // backward sweep: the same recurrence on the image flipped in x and y
    for (int f = 0; f < F; ++f) p.f[f].Sin = (const uint32_t*)S[f];
    p.flip = 1; p.ticket = ctl + ctl_pass; p.progress = ctl + ctl_pass + WAVE_CTL_PAD;
    if (g.mode == MVS_MODE_HH) rc = launch_wave_nw<0xF>(ctx, g, p);
    else if (g.mode == MVS_MODE_HH4) rc = launch_wave_nw<0x5>(ctx, g, p);
    else rc = launch_wave_nw<0x1>(ctx, g, p);  // MODE_SGBM: only the (x+1, y) direction
    return rc;
}

int sgbm_aggregate_wave_dev(mvs_ctx* ctx, const SgbmGeom& g, const int16_t* C, int16_t* S)
{
    return sgbm_aggregate_wave_batch_dev(ctx, g, 1, &C, &S);
}

// Host check of the bounded-wait flag (synchronises the stream); called by the entry points that sync anyway.
int sgbm_aggregate_wave_status(mvs_ctx* ctx)
{
    auto it = ctx->arena.bufs.find("sgbm.wave.err");
    if (it == ctx->arena.bufs.end() || !it->second.p) return MVS_OK;
    int h = 0;
    MVS_CUDA(ctx, cudaMemcpyAsync(&h, it->second.p, sizeof(int), cudaMemcpyDeviceToHost, ctx->stream));
    MVS_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    if (h) {
        cudaMemsetAsync(it->second.p, 0, sizeof(int), ctx->stream);
        return mvs_fail(ctx, MVS_ERR_CUDA, "aggregation: a bounded inter-band wait expired (results invalid)");
    }
    return MVS_OK;
}
