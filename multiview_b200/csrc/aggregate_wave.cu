// a3 (iv), fused form: one sweep of the SGM aggregation carries ALL directions of that sweep at once.
//
// Forward sweep (y up, x up): L0 from (x-1,y), L1 from (x-1,y-1), L2 from (x,y-1), L3 from (x+1,y-1);
// the backward sweep is the same recurrence on the image flipped in x and y.  Reading C once
// and writing S once per sweep is the algorithmic minimum (8 B/cell for both sweeps).
//
// Parallelisation: pixel (x,y) depends on (x-1,y) and on three pixels of row y-1, so all pixels with the same
// t = x + 2y are independent (a wavefront).  A CTA owns a band of R consecutive rows; warp r walks row r of the
// band between two and four pixels behind warp r-1.
//   * the disparity vector of a pixel is spread over the warp's 32 lanes (2*NW disparities per lane, u16x2);
//   * L0 stays in registers; L1/L2/L3 of the row above come from a shared-memory ring written by warp r-1.
//     A ring slot is {min(L) x2, -, -, SENT | L[0..D) | SENT, -, -, -}: the sentinel words give the d-1 / d+1
//     neighbours of lanes 0 and 31 without branches and the minimum travels with the data;
//   * neighbouring warps are coupled only by mbarriers: per ring row eight "pixel p written" and eight "pixel p
//     consumed" barriers (index p mod 8, phase parity (p / 8) & 1; the ring depths keep producer and consumer less
//     than eight pixels apart, so a parity is never ambiguous).  arrive has release and try_wait acquire semantics
//     at CTA scope, so no memory fence is issued and a waiting warp is suspended instead of burning issue slots.
//     There is no CTA-wide barrier inside a band, and the rings are two slots deeper than lock-step needs, so a
//     warp that stalls does not stall the others at once;
//   * image borders need no branches: the slots "x = -1" and "x = W1" are filled with zeros (Lr = 0, minLr = 0
//     outside the image, as the reference initialises its Lr buffers);
//   * C (and S in the backward sweep) stream through a per-warp cp.async pipeline three pixels ahead;
//   * bands are chained through global memory: the last warp of a band streams its slots to a boundary buffer and
//     publishes its progress; a helper warp of the next band waits for that progress and copies the boundary row
//     into the ring with cp.async a few pixels ahead of its consumers;
//   * a launch covers a batch of frames; (frame, band) items are handed out by an atomic ticket in dependency
//     order, so the band a CTA waits for is always running or finished (no co-residency assumption); every wait
//     is bounded and raises an error flag instead of hanging.
#include "common.cuh"

#define SENT2 0x7FFF7FFFu
// rows (compute warps) per band by disparity words per lane: with the helper a multiple of four warps (four
// schedulers); narrower disparity ranges have smaller ring slots, so more rows fit and hide the per-pixel latency
#ifndef WAVE_R4
#define WAVE_R4 15           // Dp = 256: 16 warps
#endif
#ifndef WAVE_R3
#define WAVE_R3 19           // Dp = 192: 20 warps
#endif
#ifndef WAVE_R2
#define WAVE_R2 27           // Dp = 128: 28 warps
#endif
#ifndef WAVE_R1
#define WAVE_R1 15           // Dp = 64: 16 warps (31 rows measured slightly slower at 640x480)
#endif
template <int NW> struct WaveRows { static constexpr int R = NW == 4 ? WAVE_R4 : NW == 3 ? WAVE_R3 : NW == 2 ? WAVE_R2 : WAVE_R1; };
static int wave_rows(int nw) { return nw == 4 ? WAVE_R4 : nw == 3 ? WAVE_R3 : nw == 2 ? WAVE_R2 : WAVE_R1; }
#ifndef WAVE_PUBLISH
#define WAVE_PUBLISH 8       // inter-band progress is published every this many pixels
#endif
#ifndef WAVE_D12
#define WAVE_D12 7           // ring depth for L1 / L2 (lock-step needs 4 / 3); measured: 15 rows x 7/7/4 beats 16 x 6/6/4
#endif
#ifndef WAVE_D3
#define WAVE_D3 4            // ring depth for L3 (lock-step needs 2)
#endif
#ifndef WAVE_FULL_SLEEP
#define WAVE_FULL_SLEEP 0
#endif
#ifndef WAVE_CONS_SLEEP
#define WAVE_CONS_SLEEP 100
#endif
#define WAVE_SPIN_LIMIT (1 << 20)
#define WAVE_NB 8            // mbarriers per ring row and kind (pixel index mod WAVE_NB)
#define WAVE_CTL_PAD 16
// a producer at pixel x may overwrite its slots once the consumer has finished pixel x - WAVE_WAR - 1:
// L1 slot x % D12 held pixel x - D12 (read at x - D12 + 1), L3 slot x % D3 held pixel x - D3 (read at x - D3 - 1)
#define WAVE_WAR ((WAVE_D12 - 2) < WAVE_D3 ? (WAVE_D12 - 2) : WAVE_D3)

struct WaveFrame {
    const uint32_t* C;
    const uint32_t* Sin;   // backward sweep: S after the forward sweep (nullptr in the forward sweep)
    uint32_t* Sout;
    uint32_t* bnd;         // [nbands][W1][3][4 + wpp] boundary slots (header + path costs)
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
__device__ __forceinline__ void mbar_init(uint32_t saddr, int count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(saddr), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_inval(uint32_t saddr) { asm volatile("mbarrier.inval.shared::cta.b64 [%0];" ::"r"(saddr) : "memory"); }
__device__ __forceinline__ void mbar_arrive(uint32_t saddr)  // release at CTA scope
{
    asm volatile("{ .reg .b64 st; mbarrier.arrive.shared::cta.b64 st, [%0]; }" ::"r"(saddr) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint32_t saddr, uint32_t parity)  // acquire at CTA scope; may suspend the warp
{
    uint32_t ok;
    asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0, 1, 0, p; }"
                 : "=r"(ok) : "r"(saddr), "r"(parity) : "memory");
    return ok != 0;
}
__device__ __forceinline__ void cp_async16(uint32_t smem_addr, const void* gmem)
{
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_addr), "l"(gmem) : "memory");
}
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
// Written so that the packed mins go to the integer pipe and the plain adds may go to the FMA pipe (IMAD).
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
        uint32_t x = __vminu2(lo, hi) + P1x2;           // halves stay below 2^16: a plain add
        uint32_t t = __vimin3_u16x2(x, a[j], delta);
        uint32_t l = t + (Cw[j] - minLp2);              // per half: t >= minLp and C + (t - minLp) <= 32767
        if (PAD) l = (l & ~padmask[j]) | (SENT2 & padmask[j]);
        out[j] = l;
        m = __vminu2(m, l);
    }
    uint32_t mm = min(m & 0xffffu, m >> 16);
    return __reduce_min_sync(0xffffffffu, mm);
}

// Shared-memory geometry (32-bit words).
template <int NW>
struct Ring {
    static constexpr int WPP = NW * 32;        // words per pixel
    static constexpr int SLOT = WPP + 8;       // header(4) + data + tail(4)
    static constexpr int BSLOT = WPP + 4;      // boundary-buffer slot in global memory: header + data
    // rows written by compute warps: lock-step needs 4 / 3 / 2 slots for L1 / L2 / L3; two more give two pixels of slack
    static constexpr int D12 = WAVE_D12, D3 = WAVE_D3;
    static constexpr int ROW_SLOTS = 2 * D12 + D3;
    static constexpr int HD = 8;               // row above the band (helper): 8 slots per direction
    static constexpr int HELP_SLOTS = 3 * HD;
    static constexpr int R = WaveRows<NW>::R;  // rows per band
    static constexpr int THREADS = (R + 1) * 32;
    static constexpr int NSLOTS = HELP_SLOTS + R * ROW_SLOTS;
    static constexpr int WORDS = NSLOTS * SLOT;
    static constexpr int STAGE = 4;            // cp.async stages of the per-warp C (and S) stream
    static constexpr int STREAM_WORDS = R * STAGE * WPP;
};

template <int NW>
__device__ __forceinline__ void cp_async_lane(uint32_t smem_addr, const uint32_t* g)
{
    if (NW == 4) asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_addr), "l"(g) : "memory");
    else if (NW == 2) asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(smem_addr), "l"(g) : "memory");
    else {
#pragma unroll
        for (int j = 0; j < NW; ++j)
            asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(smem_addr + 4 * j), "l"(g + j) : "memory");
    }
}

template <int NW, int DIRMASK, bool PAD, bool BWD>
__global__ void __launch_bounds__(Ring<NW>::THREADS, 1)
wave_kernel(WaveParams p)
{
    using RG = Ring<NW>;
    constexpr bool VERT = (DIRMASK & 0xE) != 0;
    extern __shared__ __align__(16) uint32_t ring[];
    uint32_t* cstream = ring + RG::WORDS;                  // [warp][stage][WPP] cost stream
    uint32_t* sstream = cstream + RG::STREAM_WORDS;        // [warp][stage][WPP] S-in stream (backward sweep only)
    __shared__ int s_item;
    // [ring row k][p mod NB]: "pixel p of ring row k written" (k = 0: by the helper, else by warp k-1) and
    // "pixel p of ring row k consumed" (by warp k); the pseudo-pixel W1 is the zero slot behind the row
    constexpr int WAVE_R = RG::R;
    __shared__ __align__(8) uint64_t s_full[WAVE_R + 1][WAVE_NB];
    __shared__ __align__(8) uint64_t s_cons[WAVE_R + 1][WAVE_NB];
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const bool helper = wid == WAVE_R;
    // The warp scheduler favours higher warp ids among eligible warps: upstream rows (whose progress everything below
    // waits for) get the higher ids, so a warp spinning on its predecessor never takes an issue slot from it.
    const int r = helper ? WAVE_R : WAVE_R - 1 - wid;
    const int W1 = p.W1;
    const uint32_t P1x2 = wdup16((uint32_t)p.P1), P2x2 = wdup16((uint32_t)p.P2);
    const uint32_t ring_s = (uint32_t)__cvta_generic_to_shared(ring);
    const uint32_t full_s = (uint32_t)__cvta_generic_to_shared(s_full), cons_s = (uint32_t)__cvta_generic_to_shared(s_cons);
    auto bar_of = [](uint32_t base, int row, int pixel) -> uint32_t { return base + (uint32_t)(row * WAVE_NB + (pixel & (WAVE_NB - 1))) * 8u; };
    auto par_of = [](int pixel) -> uint32_t { return (uint32_t)(pixel / WAVE_NB) & 1u; };
    bool bars_valid = false;

    for (int s = threadIdx.x; s < RG::NSLOTS; s += blockDim.x) {  // sentinels: written once, never overwritten
        ring[s * RG::SLOT + 3] = SENT2;
        ring[s * RG::SLOT + RG::SLOT - 4] = SENT2;
    }
    uint32_t padmask[NW];
#pragma unroll
    for (int j = 0; j < NW; ++j) {
        int d0 = (lane * NW + j) * 2;
        padmask[j] = (d0 >= p.D ? 0xffffu : 0u) | (d0 + 1 >= p.D ? 0xffff0000u : 0u);
    }
    // ring rows: what this warp reads (the row above: ring row r) and what it writes (ring row r + 1)
    const int lw = 4 + lane * NW;  // this lane's first data word inside a slot
    const int rd12 = r == 0 ? RG::HD : RG::D12, rd3 = r == 0 ? RG::HD : RG::D3;  // depths of the rings this warp reads
    const int rd_slot0 = r == 0 ? 0 : RG::HELP_SLOTS + (r - 1) * RG::ROW_SLOTS;
    const int rb1 = rd_slot0 * RG::SLOT + lw, rb2 = (rd_slot0 + rd12) * RG::SLOT + lw, rb3 = (rd_slot0 + 2 * rd12) * RG::SLOT + lw;
    const int wr_slot0 = RG::HELP_SLOTS + r * RG::ROW_SLOTS;
    const int wb1 = wr_slot0 * RG::SLOT + lw, wb2 = (wr_slot0 + RG::D12) * RG::SLOT + lw, wb3 = (wr_slot0 + 2 * RG::D12) * RG::SLOT + lw;
    const int my_stream = (r * RG::STAGE) * RG::WPP + lane * NW;  // this lane's words in stage 0 of the warp's streams

    for (;;) {
        __syncthreads();  // every warp has left the previous band
        if (threadIdx.x == 0) s_item = atomicAdd(p.ticket, 1);
        // re-arm the mbarriers: invalidate all of them, then (after a CTA barrier: an invalidation may drop barrier state
        // another thread has just initialised) initialise them for one arrival per phase
        if (bars_valid) {
            for (int i = threadIdx.x; i < (WAVE_R + 1) * WAVE_NB; i += blockDim.x) {
                mbar_inval(full_s + (uint32_t)i * 8u);
                mbar_inval(cons_s + (uint32_t)i * 8u);
            }
            __syncthreads();
        }
        for (int i = threadIdx.x; i < (WAVE_R + 1) * WAVE_NB; i += blockDim.x) {
            mbar_init(full_s + (uint32_t)i * 8u, 1);
            mbar_init(cons_s + (uint32_t)i * 8u, 1);
        }
        bars_valid = true;
        __syncthreads();
        if (s_item >= p.nbands * p.F) break;
        const int band = s_item / p.F, fi = s_item - band * p.F;
        const WaveFrame& fr = p.f[fi];
        const int ys0 = band * WAVE_R;  // first sweep row of the band
        const int rows = min(WAVE_R, p.H - ys0);
        const bool first_band = band == 0, last_band = band == p.nbands - 1;
        const uint32_t* bin = fr.bnd + (size_t)(first_band ? 0 : band - 1) * W1 * 3 * RG::BSLOT;  // written by the previous band
        uint32_t* bout = fr.bnd + (size_t)band * W1 * 3 * RG::BSLOT;
        int* my_progress = p.progress + (size_t)fi * p.nbands + band;
        const int* prev_progress = my_progress - (first_band ? 0 : 1);
        bool failed = false;

        // bounded wait for "pixel `pixel` of ring row `row` written / consumed"
        auto wait_bar = [&](uint32_t base, int row, int pixel) {
            if (failed) return;
            const uint32_t b = bar_of(base, row, pixel), par = par_of(pixel);
            int spins = 0;
            while (!mbar_try_wait(b, par)) {
                if (++spins > WAVE_SPIN_LIMIT) {
                    // where it happened: kind (1 written / 2 consumed), ring row, band, pixel
                    atomicCAS(p.error, 0, ((base == full_s ? 1 : 2) << 28) | (row << 23) | ((band & 127) << 16) | (pixel & 0xffff));
                    failed = true;
                    break;
                }
                if ((spins & 63) == 0 && *(volatile int*)p.error) { failed = true; break; }
#if WAVE_CONS_SLEEP
                // back-pressure waits (this warp is AHEAD of its consumer): sleep instead of polling, the issue slots
                // go to the slower warps on the same scheduler
                if (base != full_s) __nanosleep(WAVE_CONS_SLEEP);
#endif
#if WAVE_FULL_SLEEP
                if (base == full_s) __nanosleep(WAVE_FULL_SLEEP);
#endif
            }
        };

        // zero slots prepared before anybody runs: the row above the image, and "x = -1" of the row each warp reads
        if (helper && VERT && first_band) {
            for (int s = 0; s < RG::HELP_SLOTS; ++s) {
                for (int k = lane; k < RG::WPP; k += 32) ring[s * RG::SLOT + 4 + k] = 0;
                if (lane == 0) ring[s * RG::SLOT] = 0;
            }
        }
        if (!helper && VERT) {
            uint32_t* z = ring + (rb1 - lw) + (rd12 - 1) * RG::SLOT;  // slot (-1 mod depth)
            for (int k = lane; k < RG::WPP; k += 32) z[4 + k] = 0;
            if (lane == 0) z[0] = 0;
        }
        __syncthreads();

        if (helper) {
            if (VERT) {
                if (!first_band) {  // (first band: everything above the image is zero, warp 0 does not wait)
                    // boundary row of the previous band -> ring row 0, four pixels ahead of publication
                    int seen = 0;
                    for (int xp = 0; xp <= W1 + 2; ++xp) {
                        if (xp <= W1) {
                            // slot xp % 8 last held pixel xp - 8, read by warp 0 (direction 1) at its pixel xp - 7
                            if (xp >= 7) wait_bar(cons_s, 0, xp - 7);  // acquire before overwriting the slot
                            if (xp < W1) {
                                const int need = xp + 1;
                                if (seen < need && !failed) {
                                    int spins = 0;
                                    while ((seen = ld_acquire(prev_progress)) < need) {
                                        // bounded: a timeout here or anywhere else in the grid ends all waiting (results are
                                        // then invalid and the host reports an error): a logic fault cannot hang the GPU
                                        if (++spins > WAVE_SPIN_LIMIT) { atomicCAS(p.error, 0, (3 << 28) | ((band & 127) << 16) | (xp & 0xffff)); failed = true; break; }
                                        if ((spins & 1023) == 0 && *(volatile int*)p.error) { failed = true; break; }
                                        __nanosleep(32);
                                    }
                                }
                                const uint32_t* src = bin + (size_t)xp * 3 * RG::BSLOT;
#pragma unroll
                                for (int dir = 1; dir <= 3; ++dir) {
                                    if (!(DIRMASK & (1 << dir))) continue;
                                    const uint32_t dst = ring_s + (((dir - 1) * RG::HD + (xp & (RG::HD - 1))) * RG::SLOT) * 4;
                                    const uint32_t* sp = src + (dir - 1) * RG::BSLOT;
                                    if (lane < RG::WPP / 4) cp_async16(dst + (lane + 1) * 16, sp + (lane + 1) * 4);
                                    if (lane == 0) cp_async16(dst, sp);  // header: {min, -, -, SENT}
                                }
                            } else {  // xp == W1: the zero slot behind the last pixel
#pragma unroll
                                for (int dir = 1; dir <= 3; ++dir) {
                                    uint32_t* dst = ring + ((dir - 1) * RG::HD + (xp & (RG::HD - 1))) * RG::SLOT;
                                    for (int k = lane; k < RG::WPP; k += 32) dst[4 + k] = 0;
                                    if (lane == 0) dst[0] = 0;
                                }
                            }
                        }
                        cp_async_commit();
                        cp_async_wait<2>();  // the group of pixel xp - 2 has landed
                        if (xp >= 2) {
                            __syncwarp();  // all lanes' copies are in; lane 0 publishes pixel xp - 2
                            if (lane == 0) mbar_arrive(bar_of(full_s, 0, xp - 2));
                        }
                    }
                }
            }
        } else if (r < rows) {
            const bool to_ring = r + 1 < rows, to_global = VERT && r + 1 == rows && !last_band;
            const bool wait_above = VERT && !(first_band && r == 0);
            uint32_t L0[NW];
#pragma unroll
            for (int j = 0; j < NW; ++j) L0[j] = 0;
            uint32_t minL0 = 0;
            const int ys = ys0 + r;
            const int y = BWD ? p.H - 1 - ys : ys;
            const uint32_t gbase = (uint32_t)(((size_t)y * W1 + (BWD ? W1 - 1 : 0)) * RG::WPP + lane * NW);  // < 2^32 words
            const int gstride = BWD ? -RG::WPP : RG::WPP;
            const uint32_t cs_s = ring_s + (RG::WORDS + my_stream) * 4, ss_s = cs_s + RG::STREAM_WORDS * 4;
            auto stream_issue = [&](int xq) {  // request pixel xq of this row into stage xq % STAGE (one group, maybe empty)
                if (xq < W1) {
                    const uint32_t o = gbase + (uint32_t)(xq * gstride);
                    cp_async_lane<NW>(cs_s + (xq & (RG::STAGE - 1)) * (RG::WPP * 4), fr.C + o);
                    if (BWD) cp_async_lane<NW>(ss_s + (xq & (RG::STAGE - 1)) * (RG::WPP * 4), fr.Sin + o);
                }
                cp_async_commit();
            };
            for (int xq = 0; xq < RG::STAGE - 1; ++xq) stream_issue(xq);
            uint32_t* gout = bout + lw;  // this lane's data words of boundary pixel 0, direction 1
            // running slot indices: reads (x-1) % rd12, x % rd12, (x+1) % rd3; writes x % D12, x % D3
            int i1 = rd12 - 1, i2 = 0, i3 = 1 % rd3, j12 = 0, j3 = 0;

            for (int x = 0; x < W1; ++x) {
                stream_issue(x + RG::STAGE - 1);
                if (wait_above) wait_bar(full_s, r, x + 1);                                       // the row above is written through pixel x + 1
                if (VERT && to_ring && x > WAVE_WAR) wait_bar(cons_s, r + 1, x - WAVE_WAR - 1);  // the slots about to be overwritten were read
                cp_async_wait<RG::STAGE - 1>();  // this lane's words of pixel x have landed (own copies only)
                uint32_t Cw[NW], Sw[NW];
                ldw<NW>(cstream + my_stream + (x & (RG::STAGE - 1)) * RG::WPP, Cw);
                if (BWD) ldw<NW>(sstream + my_stream + (x & (RG::STAGE - 1)) * RG::WPP, Sw);
                uint32_t acc[NW], Lo1[NW], Lo2[NW], Lo3[NW];
                uint32_t mn1 = 0, mn2 = 0, mn3 = 0;
                if (DIRMASK & 1) {  // direction 0: predecessor (x-1, y), kept in registers
                    uint32_t left = __shfl_up_sync(0xffffffffu, L0[NW - 1], 1);
                    uint32_t right = __shfl_down_sync(0xffffffffu, L0[0], 1);
                    if (lane == 0) left = SENT2;
                    if (lane == 31) right = SENT2;
                    uint32_t Lo[NW];
                    uint32_t mn = step_dir<NW, PAD>(L0, left, right, minL0, Cw, P1x2, P2x2, padmask, Lo);
#pragma unroll
                    for (int j = 0; j < NW; ++j) { L0[j] = Lo[j]; acc[j] = Lo[j]; }
                    minL0 = wdup16(mn);
                } else {
#pragma unroll
                    for (int j = 0; j < NW; ++j) acc[j] = 0;
                }
                if (DIRMASK & 2) {  // direction 1: predecessor (x-1, y-1)
                    const uint32_t* sl = ring + rb1 + i1 * RG::SLOT;
                    uint32_t a[NW];
                    ldw<NW>(sl, a);
                    mn1 = step_dir<NW, PAD>(a, sl[-1], sl[NW], sl[-lw], Cw, P1x2, P2x2, padmask, Lo1);
#pragma unroll
                    for (int j = 0; j < NW; ++j) acc[j] = __viaddmin_u16x2(acc[j], Lo1[j], SENT2);
                }
                if (DIRMASK & 4) {  // direction 2: predecessor (x, y-1)
                    const uint32_t* sl = ring + rb2 + i2 * RG::SLOT;
                    uint32_t a[NW];
                    ldw<NW>(sl, a);
                    mn2 = step_dir<NW, PAD>(a, sl[-1], sl[NW], sl[-lw], Cw, P1x2, P2x2, padmask, Lo2);
#pragma unroll
                    for (int j = 0; j < NW; ++j) acc[j] = __viaddmin_u16x2(acc[j], Lo2[j], SENT2);
                }
                if (DIRMASK & 8) {  // direction 3: predecessor (x+1, y-1)
                    const uint32_t* sl = ring + rb3 + i3 * RG::SLOT;
                    uint32_t a[NW];
                    ldw<NW>(sl, a);
                    mn3 = step_dir<NW, PAD>(a, sl[-1], sl[NW], sl[-lw], Cw, P1x2, P2x2, padmask, Lo3);
#pragma unroll
                    for (int j = 0; j < NW; ++j) acc[j] = __viaddmin_u16x2(acc[j], Lo3[j], SENT2);
                }
                // S = min(S_in + sum L, 32767)
                if (BWD) {
#pragma unroll
                    for (int j = 0; j < NW; ++j) acc[j] = __viaddmin_u16x2(Sw[j], acc[j], SENT2);
                }
                stw<NW>(fr.Sout + (gbase + (uint32_t)(x * gstride)), acc);
                if (to_ring) {  // for the row below
                    if (DIRMASK & 2) {
                        uint32_t* d = ring + wb1 + j12 * RG::SLOT;
                        stw<NW>(d, Lo1);
                        if (lane == 0) d[-4] = wdup16(mn1);
                    }
                    if (DIRMASK & 4) {
                        uint32_t* d = ring + wb2 + j12 * RG::SLOT;
                        stw<NW>(d, Lo2);
                        if (lane == 0) d[-4] = wdup16(mn2);
                    }
                    if (DIRMASK & 8) {
                        uint32_t* d = ring + wb3 + j3 * RG::SLOT;
                        stw<NW>(d, Lo3);
                        if (lane == 0) d[-4] = wdup16(mn3);
                    }
                } else if (to_global) {  // for the next band
                    if (DIRMASK & 2) {
                        stw<NW>(gout, Lo1);
                        if (lane == 0) *(uint4*)(gout - lw) = make_uint4(wdup16(mn1), 0u, 0u, SENT2);
                    }
                    if (DIRMASK & 4) {
                        stw<NW>(gout + RG::BSLOT, Lo2);
                        if (lane == 0) *(uint4*)(gout - lw + RG::BSLOT) = make_uint4(wdup16(mn2), 0u, 0u, SENT2);
                    }
                    if (DIRMASK & 8) {
                        stw<NW>(gout + 2 * RG::BSLOT, Lo3);
                        if (lane == 0) *(uint4*)(gout - lw + 2 * RG::BSLOT) = make_uint4(wdup16(mn3), 0u, 0u, SENT2);
                    }
                    gout += 3 * RG::BSLOT;
                    if (((x + 1) % WAVE_PUBLISH) == 0 || x + 1 == W1) {
                        // the warp barrier orders every lane's boundary stores before lane 0's release store
                        // (release is cumulative over what the warp barrier made visible to lane 0)
                        __syncwarp();
                        if (lane == 0) st_release(my_progress, x + 1);
                    }
                }
                if (VERT) {
                    // publish: ring row r+1 holds pixel x; ring row r has been read for pixel x.  The warp barrier
                    // puts every lane's ring reads and writes before lane 0's arrivals (release).
                    __syncwarp();
                    if (lane == 0) {
                        if (to_ring) mbar_arrive(bar_of(full_s, r + 1, x));
                        mbar_arrive(bar_of(cons_s, r, x));
                    }
                }
                i1 = i1 + 1 == rd12 ? 0 : i1 + 1;
                i2 = i2 + 1 == rd12 ? 0 : i2 + 1;
                i3 = i3 + 1 == rd3 ? 0 : i3 + 1;
                j12 = j12 + 1 == RG::D12 ? 0 : j12 + 1;
                j3 = j3 + 1 == RG::D3 ? 0 : j3 + 1;
            }
            if (to_ring && VERT) {
                // zero slot "x = W1" for direction 3 of the row below (its predecessor is outside the image);
                // the pseudo-pixel W1 is what the row below waits for at its last pixel
                if (DIRMASK & 8) {
                    if (W1 > RG::D3) wait_bar(cons_s, r + 1, W1 - RG::D3 - 1);  // that L3 slot held pixel W1 - D3
                    uint32_t* d = ring + wb3 + j3 * RG::SLOT;
                    for (int k = lane; k < RG::WPP; k += 32) d[k - lw + 4] = 0;
                    if (lane == 0) d[-4] = 0;
                }
                __syncwarp();
                if (lane == 0) mbar_arrive(bar_of(full_s, r + 1, W1));
            }
        }
        cp_async_wait<0>();
    }
}

template <int NW, int DIRMASK, bool PAD, bool BWD>
static int launch_wave(mvs_ctx* ctx, const WaveParams& p)
{
    const int smem = (Ring<NW>::WORDS + (BWD ? 2 : 1) * Ring<NW>::STREAM_WORDS) * 4;
    // per instantiation and device (function attributes are per device; doing this twice from two threads is harmless)
    static int per_sm_dev[MVS_MAX_DEVICES] = {};
    const int dv = ctx->device >= 0 && ctx->device < MVS_MAX_DEVICES ? ctx->device : 0;
    int per_sm = dv == ctx->device ? per_sm_dev[dv] : 0;
    if (!per_sm) {
        MVS_CUDA(ctx, cudaFuncSetAttribute(wave_kernel<NW, DIRMASK, PAD, BWD>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
        int n = 1;
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, wave_kernel<NW, DIRMASK, PAD, BWD>, Ring<NW>::THREADS, smem);
        per_sm = n < 1 ? 1 : n;
        per_sm_dev[dv] = per_sm;
    }
    int grid = min(p.nbands * p.F, ctx->sm_count * per_sm);
    wave_kernel<NW, DIRMASK, PAD, BWD><<<grid, Ring<NW>::THREADS, smem, ctx->stream>>>(p);
    MVS_LAUNCH_CHECK(ctx, "wave_kernel");
    return MVS_OK;
}

template <int DIRMASK, bool BWD>
static int launch_wave_nw(mvs_ctx* ctx, const SgbmGeom& g, const WaveParams& p)
{
    bool pad = g.Dp != g.D;
    switch (g.Dp / 64) {
    case 1: return pad ? launch_wave<1, DIRMASK, true, BWD>(ctx, p) : launch_wave<1, DIRMASK, false, BWD>(ctx, p);
    case 2: return pad ? launch_wave<2, DIRMASK, true, BWD>(ctx, p) : launch_wave<2, DIRMASK, false, BWD>(ctx, p);
    case 3: return pad ? launch_wave<3, DIRMASK, true, BWD>(ctx, p) : launch_wave<3, DIRMASK, false, BWD>(ctx, p);
    case 4: return pad ? launch_wave<4, DIRMASK, true, BWD>(ctx, p) : launch_wave<4, DIRMASK, false, BWD>(ctx, p);
    }
    return mvs_fail(ctx, MVS_ERR_UNSUPPORTED, "aggregate: Dp=%d", g.Dp);
}

int sgbm_aggregate_wave_batch_dev(mvs_ctx* ctx, const SgbmGeom& g, int F, const int16_t* const* C, int16_t* const* S)
{
    if (F < 1 || F > MVS_MAX_BATCH) return mvs_fail(ctx, MVS_ERR_ARG, "aggregate: batch of %d frames (1..%d)", F, MVS_MAX_BATCH);
    const int nbands = cdiv(g.H, wave_rows(g.Dp / 64));
    const size_t wpp = (size_t)g.Dp / 2;
    const int ctl_pass = F * nbands + 2 * WAVE_CTL_PAD;
    const size_t bnd_sz = (size_t)nbands * g.W1 * 3 * (wpp + 4);
    MVS_ARENA(ctx, uint32_t, bnd, "sgbm.wave.bnd", bnd_sz * F);
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
        p.f[f].bnd = bnd + bnd_sz * f;
    }
    p.flip = 0; p.ticket = ctl; p.progress = ctl + WAVE_CTL_PAD;
    rc = g.mode == MVS_MODE_HH4 ? launch_wave_nw<0x5, false>(ctx, g, p) : launch_wave_nw<0xF, false>(ctx, g, p);
    if (rc) return rc;
    // backward sweep: the same recurrence on the image flipped in x and y
    for (int f = 0; f < F; ++f) p.f[f].Sin = (const uint32_t*)S[f];
    p.flip = 1; p.ticket = ctl + ctl_pass; p.progress = ctl + ctl_pass + WAVE_CTL_PAD;
    if (g.mode == MVS_MODE_HH) rc = launch_wave_nw<0xF, true>(ctx, g, p);
    else if (g.mode == MVS_MODE_HH4) rc = launch_wave_nw<0x5, true>(ctx, g, p);
    else rc = launch_wave_nw<0x1, true>(ctx, g, p);  // MODE_SGBM: only the (x+1, y) direction
    return rc;
}

int sgbm_aggregate_wave_dev(mvs_ctx* ctx, const SgbmGeom& g, const int16_t* C, int16_t* S)
{
    return sgbm_aggregate_wave_batch_dev(ctx, g, 1, &C, &S);
}

// Host check of the bounded-wait flag (synchronises the stream); called by the entry points that sync anyway.
int sgbm_aggregate_wave_status(mvs_ctx* ctx, cudaStream_t stream)
{
    auto it = ctx->arena.bufs.find("sgbm.wave.err");
    if (it == ctx->arena.bufs.end() || !it->second.p) return MVS_OK;
    if (!stream) stream = ctx->stream;
    int h = 0;
    MVS_CUDA(ctx, cudaMemcpyAsync(&h, it->second.p, sizeof(int), cudaMemcpyDeviceToHost, stream));
    MVS_CUDA(ctx, cudaStreamSynchronize(stream));
    if (h) {
        cudaMemsetAsync(it->second.p, 0, sizeof(int), stream);
        return mvs_fail(ctx, MVS_ERR_CUDA, "aggregation: a bounded wait expired (results invalid; kind %d, ring row %d, band %d, pixel %d)",
                        (h >> 28) & 3, (h >> 23) & 31, (h >> 16) & 127, h & 0xffff);
    }
    return MVS_OK;
}
