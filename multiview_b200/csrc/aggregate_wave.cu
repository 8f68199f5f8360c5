// a3 (iv), fused form: one sweep of the SGM aggregation carries ALL directions of that sweep at once.
//
// Forward sweep (y up, x up): L0 from (x-1,y), L1 from (x-1,y-1), L2 from (x,y-1), L3 from (x+1,y-1);
// the backward sweep is the same recurrence on the image flipped in x and y.  Reading C once
// and writing S once per sweep is the algorithmic minimum (8 B/cell for both sweeps).
//
// Parallelisation: pixel (x,y) depends on (x-1,y) and on three pixels of row y-1, so all pixels with the same
// t = x + 2y are independent (a wavefront).  A CTA owns a band of R consecutive rows; warp r walks row r of the
// band between two and six pixels behind warp r-1.
//
// Arithmetic (every operation below was counted):
//   * a pixel's disparity vector is spread over the warp's 32 lanes, NW packed u16x2 words per lane.  The volumes
//     use the SPLIT layout (common.cuh): word w of a pixel holds disparity w in its low half and w + Dp/2 in its
//     high half, so the d-1 / d+1 neighbours of a word are simply the previous / next word (register renaming,
//     no byte permutes); only the words at the lane edges come from the neighbouring lane (one shuffle + one
//     permute that also inserts the d = -1 / d = D sentinels);
//   * the recurrence is invariant to a constant shift of Lp, so what travels from pixel to pixel is the
//     NORMALISED and CLAMPED vector  Ln(d) = min(L(d) - min L, P2)  (one fused add-min per word).  With it
//         L(p,d) = C(p,d) + min(Ln(d), Ln(d-1) + P1, Ln(d+1) + P1)
//     because min(Ln(d), .) <= P2 already carries the "min Lp + P2" term: one three-input minimum and one add per
//     word and direction, no per-direction minimum to load and no subtraction of min Lp;
//   * plain adds are written a * one + b with `one` a kernel parameter, so they issue on the FMA pipe (IMAD) and
//     leave the ALU pipe to the packed minima;
//   * the row minimum is one packed min, one permute and one CREDUX (the minimum comes back duplicated in both
//     halves, which is the form the normalisation needs).
//
// Data movement:
//   * L0 stays in registers; L1/L2/L3 of the row above come from shared-memory rings (8 / 8 / WAVE_D3 pixels deep,
//     plain vectors, no headers) written by warp r-1;
//   * C (and S in the backward sweep) are loaded straight into registers a few pixels ahead (one 16-byte load per lane and
//     pixel), S is stored from registers: no shared-memory staging, so the rings get all of it.  Every second trip a warp
//     also asks L2 for the lines eight pixels ahead: all loads of a warp complete through one hardware scoreboard, so a
//     prefetched register is only usable once the YOUNGEST load has landed, and that age is what the L2 hit has to fit in
//     (see the refill in the sweep loop);
//   * the pixel loop is unrolled by four: slot offsets, barrier addresses and global offsets are immediates (plus a
//     per-trip base that selects the half of the eight-deep rings);
//   * neighbouring warps are coupled only by mbarriers: per ring row eight "pixel p written" and eight "pixel p
//     consumed" barriers (index p mod 8, phase parity (p / 8) & 1; producer and consumer are never eight pixels
//     apart, so a parity is never ambiguous).  arrive has release and test_wait / try_wait acquire semantics at CTA
//     scope: no memory fence is issued.  A pixel's barriers are probed one pixel early (test_wait, in a predicate); a
//     failed probe is retried with try_wait, which lets the hardware suspend the warp.  No CTA-wide barrier inside a band;
//   * image borders need no branches: the slots "x = -1" and "x = W1" are zero vectors (Lr = 0, minLr = 0 outside
//     the image, as the reference initialises its Lr buffers; a zero vector is its own normalised form);
//   * bands are chained through global memory: the last warp of a band streams its vectors to a boundary buffer and
//     publishes its progress every WAVE_PUBLISH pixels; a helper warp of the next band waits for that progress and copies
//     the boundary row into ring row 0 with bulk asynchronous copies (cp.async.bulk, completing on the pixel's "written"
//     mbarrier) as far ahead of its consumer as the ring allows;
//   * a launch covers a batch of frames; (frame, band) items are handed out by an atomic ticket in dependency
//     order, so the band a CTA waits for is always running or finished (no co-residency assumption); every wait
//     is bounded in TIME and raises an error flag instead of hanging;
//   * from 192 disparities (sgbm_wta_in_sweep) the backward sweep also does the winner-take-all step of cv::StereoSGBM on the sums it holds in
//     registers (wta_stage_a / wta_stage_b / wta_flush below): S is then neither written by it nor read back.
#include "common.cuh"

#define SENT2 0x7FFF7FFFu
// rows (compute warps) per band by disparity words per lane; with the helper warp a multiple of four (four schedulers)
#ifndef WAVE_R4
#define WAVE_R4 15           // Dp = 256: 16 warps
#endif
#ifndef WAVE_R3
#define WAVE_R3 15           // Dp = 192: 16 warps (measured at 1920x1080x192, ms per frame: 11 rows 1.155, 15 1.023, 19 1.042, 23 1.151)
#endif
#ifndef WAVE_R2
#define WAVE_R2 19           // Dp = 128: 20 warps (measured at 896x672x128, ms per frame: 11 rows 0.287, 15 0.243, 17 0.251, 19 0.234, 23 0.251, 27 0.287, 31 0.340)
#endif
#ifndef WAVE_R1
#define WAVE_R1 19           // Dp = 64: 20 warps (measured at 640x480x64, ms per frame: 15 rows 0.132, 19 0.126, 23 0.126, 27 0.138, 31 0.152)
#endif
template <int NW> struct WaveRows { static constexpr int R = NW == 4 ? WAVE_R4 : NW == 3 ? WAVE_R3 : NW == 2 ? WAVE_R2 : WAVE_R1; };
static int wave_rows(int nw) { return nw == 4 ? WAVE_R4 : nw == 3 ? WAVE_R3 : nw == 2 ? WAVE_R2 : WAVE_R1; }
#ifndef WAVE_D3
#define WAVE_D3 8            // ring depth for L3 (4 or 8; lock-step needs 2); L1 and L2 are 8 deep
#endif
#ifndef WAVE_PF
#define WAVE_PF 4            // pixels of C prefetched into registers (1, 2 or 4 = WAVE_UNROLL)
#endif
#ifndef WAVE_PF_BWD
#define WAVE_PF_BWD 2        // the same in the backward sweep, whose winner-take-all step needs the registers (measured: 2 + 1
#endif                       // costs 0.8 % against 4 + 4)
#ifndef WAVE_PFS
#define WAVE_PFS 1           // pixels of S prefetched in the backward sweep
#endif
#ifndef WAVE_CONS_SLEEP
#define WAVE_CONS_SLEEP 100
#endif
#ifndef WAVE_PUBLISH
#define WAVE_PUBLISH 16      // a band's last row publishes its progress every this many pixels (a power of two >= 4;
#endif                       // measured with the GPU full, c2 batch 16: 8 and 32 are slower)
#ifndef WAVE_PUBLISH_FEW
#define WAVE_PUBLISH_FEW 4   // the same when a launch has no more (frame, band) items than SMs: the sweep is then bound by the
#endif                       // latency of the chain of bands (c2: one frame 6.9 -> 6.1 ms, two frames 3.5 -> 3.1 ms per frame)
#ifndef WAVE_DEFER_LAST
#define WAVE_DEFER_LAST 1    // the prefetch slot of a trip's last pixel is loaded inside the same trip (see the refill in the sweep loop)
#endif
#define WAVE_UNROLL 4        // pixels per trip of the sweep loop = WAVE_PF = half the ring depth
#ifndef WAVE_L2PF
#define WAVE_L2PF 8          // every second trip a warp asks L2 for the eight pixels of C (and S) that start this many pixels ahead
#endif                       // (0 = off): the register prefetch above then finds its lines in L2 instead of waiting for HBM
#ifndef WAVE_FMA_ADDS
#define WAVE_FMA_ADDS 1      // plain adds on the FMA pipe (IMAD) instead of the ALU pipe
#endif
#define WAVE_TIMEOUT_NS 2000000000ull   // a wait that lasts two seconds is a fault: flag it and drain
#define WAVE_NB 8            // mbarriers per ring row and kind (pixel index mod WAVE_NB) = unroll factor
#define WAVE_CTL_PAD 16
// a producer at pixel x may overwrite its slots once the consumer has finished pixel x - WAVE_WAR - 1:
// L1 slot x % 8 held pixel x - 8 (read at x - 7), L3 slot x % D3 held pixel x - D3 (read at x - D3 - 1)
#define WAVE_WAR (6 < WAVE_D3 ? 6 : WAVE_D3)

struct WaveFrame {
    const uint32_t* C;
    const uint32_t* Sin;   // backward sweep: S after the forward sweep (nullptr in the forward sweep)
    uint32_t* Sout;        // the sweep's sums (nullptr in a backward sweep whose winner-take-all step is the only consumer of S)
    uint32_t* bnd;         // [nbands][W1][3][wpp] boundary vectors (normalised path costs of a band's last row)
    int16_t* raw;          // backward sweep with the winner-take-all step: raw disparity image [H][W] (initialised to INV) ..
    uint32_t* disp2key;    // .. and the right-view keys [H][W] (initialised to 0x7fffffff); nullptr: no winner-take-all step
};
// winner-take-all step of cv::StereoSGBM inside the backward sweep (see wta_pixel)
struct WaveWta {
    int W, minX1, minD, ur, K44;
    uint32_t ur_magic;     // ceil(2^32 / (100 - ur)) (0 when the divisor is 1): n / (100 - ur) == umulhi(n, ur_magic) for n < 2^25
    int* sat_flag;
};
struct WaveParams {
    WaveFrame f[MVS_MAX_BATCH];
    int* progress;         // [F][nbands] pixels finished by each band's last row
    int* ticket;           // work counter
    int* error;            // set when a bounded wait expires
    int W1, H, D, P1, P2, nbands, F;
    uint32_t one;          // 1, opaque to the compiler (see WAVE_FMA_ADDS)
    int publish_mask;      // a band's last row publishes its progress every publish_mask + 1 pixels (a power of two >= 4)
    WaveWta wta;
#ifdef WAVE_DBG_PROF       // developer build: per-row clock counts {row total, waiting for the row above, waiting for the row below, slow paths}
    unsigned long long* prof;
#endif
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
__device__ __forceinline__ unsigned long long global_ns()
{
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
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
// the same, executed by the lanes with on != 0 only, without a branch around it
__device__ __forceinline__ void mbar_arrive_if(uint32_t saddr, uint32_t on)
{
    asm volatile("{ .reg .b64 st; .reg .pred q; setp.ne.u32 q, %1, 0; @q mbarrier.arrive.shared::cta.b64 st, [%0]; }" ::"r"(saddr), "r"(on) : "memory");
}
__device__ __forceinline__ void st_release_if(int* p, int v, uint32_t on)
{
    asm volatile("{ .reg .pred q; setp.ne.u32 q, %2, 0; @q st.release.gpu.global.s32 [%0], %1; }" ::"l"(p), "r"(v), "r"(on) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint32_t saddr, uint32_t parity)  // acquire at CTA scope; may suspend the warp
{
    uint32_t ok;
    asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0, 1, 0, p; }"
                 : "=r"(ok) : "r"(saddr), "r"(parity) : "memory");
    return ok != 0;
}
// Early barrier tests: the test is issued in the middle of one pixel and its outcome is read when the next pixel starts,
// so the ~90 cycles a try_wait takes (even on a completed phase) overlap the arithmetic.  The outcome stays in a PTX
// predicate in between (reading it into a register right away would wait for it): wave_kernel declares the predicates
// wave_pf / wave_pc once at function scope and these helpers name them.
// A row that has nothing to wait for in a direction (the first row of the first band has no row above, the last row of a
// band no row below) keeps that in the predicates wave_na / wave_nr (set once per row): the tests need no branch.
// test_wait only probes: it never suspends the warp.
__device__ __forceinline__ void early_test_full(uint32_t saddr, uint32_t parity)
{
    asm volatile("mbarrier.test_wait.parity.shared::cta.b64 wave_pf, [%0], %1; or.pred wave_pf, wave_pf, wave_na;" ::"r"(saddr), "r"(parity) : "memory");
}
__device__ __forceinline__ void early_test_cons(uint32_t saddr, uint32_t parity)
{
    asm volatile("mbarrier.test_wait.parity.shared::cta.b64 wave_pc, [%0], %1; or.pred wave_pc, wave_pc, wave_nr;" ::"r"(saddr), "r"(parity) : "memory");
}
__device__ __forceinline__ void early_row(uint32_t no_above, uint32_t no_below)
{
    asm volatile("setp.ne.u32 wave_na, %0, 0; setp.ne.u32 wave_nr, %1, 0;" ::"r"(no_above), "r"(no_below) : "memory");
}
__device__ __forceinline__ void early_set(uint32_t full_passed, uint32_t cons_passed)
{
    asm volatile("setp.ne.u32 wave_pf, %0, 0; setp.ne.u32 wave_pc, %1, 0;" ::"r"(full_passed), "r"(cons_passed) : "memory");
}
__device__ __forceinline__ void early_read(uint32_t& full_passed, uint32_t& cons_passed)
{
    asm volatile("selp.u32 %0, 1, 0, wave_pf; selp.u32 %1, 1, 0, wave_pc;" : "=r"(full_passed), "=r"(cons_passed)::"memory");
}
// bulk asynchronous copy global -> shared (this GPU's memory only, no multicast) that reports its bytes to an mbarrier
__device__ __forceinline__ void bulk_g2s(uint32_t smem_addr, const void* gmem, uint32_t bytes, uint32_t bar)
{
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_addr), "l"(gmem), "r"(bytes),
                 "r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint32_t saddr, uint32_t bytes)
{
    asm volatile("{ .reg .b64 st; mbarrier.arrive.expect_tx.shared::cta.b64 st, [%0], %1; }" ::"r"(saddr), "r"(bytes) : "memory");
}
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
// streaming global accesses: the volumes are far larger than L2 and every word is touched once per sweep
template <int NW>
__device__ __forceinline__ void ldg_stream(const uint32_t* p, uint32_t (&w)[NW])
{
    if (NW == 4) { uint4 v = __ldcs((const uint4*)p); w[0] = v.x; w[1 % NW] = v.y; w[2 % NW] = v.z; w[3 % NW] = v.w; }
    else if (NW == 2) { uint2 v = __ldcs((const uint2*)p); w[0] = v.x; w[1 % NW] = v.y; }
    else {
#pragma unroll
        for (int j = 0; j < NW; ++j) w[j] = __ldcs(p + j);
    }
}
template <int NW>
__device__ __forceinline__ void stg_stream(uint32_t* p, const uint32_t (&w)[NW])
{
    if (NW == 4) __stcs((uint4*)p, make_uint4(w[0], w[1 % NW], w[2 % NW], w[3 % NW]));
    else if (NW == 2) __stcs((uint2*)p, make_uint2(w[0], w[1 % NW]));
    else {
#pragma unroll
        for (int j = 0; j < NW; ++j) __stcs(p + j, w[j]);
    }
}

// Shared-memory geometry (32-bit words): ring row k (k = 0..R-1) is read by compute row k and written by compute
// row k-1 (row 0: by the helper warp, from the previous band's boundary vectors).
template <int NW>
struct Ring {
    static constexpr int WPP = NW * 32;        // words per pixel vector
    static constexpr int D12 = 8, D3 = WAVE_D3;
    static constexpr int O2 = D12 * WPP, O3 = 2 * D12 * WPP;   // word offsets of the L2 / L3 rings inside a ring row
    static constexpr int ROW_WORDS = (2 * D12 + D3) * WPP;
    static constexpr int R = WaveRows<NW>::R;  // rows per band
    static constexpr int THREADS = (R + 1) * 32;
    static constexpr int WORDS = R * ROW_WORDS;
};

// Constants of a warp's lane for the sweep arithmetic.
struct LaneConst {
    uint32_t P1x2, P2x2, one;
    uint32_t selL, selR;   // permute selectors that splice the sentinel into the words shuffled in at the lane edges
    int lm1, lp1;          // source lanes of those shuffles (rotation)
};

__device__ __forceinline__ uint32_t addf(uint32_t a, uint32_t b, uint32_t one)
{
#if WAVE_FMA_ADDS
    return a * one + b;   // IMAD: FMA pipe
#else
    return a + b;
#endif
}

template <int K> struct IC { static constexpr int value = K; };

// ---- winner-take-all step of cv::StereoSGBM inside the backward sweep --------------------------------------------------
// (reference call site strategies/stereosgbm_disp.cpp:121-125.)  The backward sweep is the last writer of
// S, and the winner-take-all step its only reader: done here, on the sums a warp holds in registers, S never goes back to
// memory (2 B per cell written and 2 B read saved).  One pixel: first minimum, uniqueness test over the disparities
// outside best-1..best+1, the two neighbours of the minimum for the sub-pixel fit; the divisions, the right-view key
// (atomicMin on minS << 16 | (0xFFFF - x): the reference's "first writer in descending x wins") and the store are done
// for 32 pixels at a time, one pixel per lane (wta_flush).
// S: this lane's words of the pixel (split layout: word j holds disparity lane * NW + j in its low half and that
// + Dp/2 in its high half; padding disparities hold 32767).
// Per pixel only what needs the whole vector is done, in two stages placed between the phases of the sweep's own
// arithmetic (the compiler keeps warp-wide operations in program order, and each stage's reduction then has the next phase
// of the sweep to complete in; no branch around a stage: it must share a basic block with that arithmetic):
//   A: the first minimum, one min-reduction over (S << 16 | d);
//   B: the two neighbours of the minimum, read from a copy of the vector the warp staged in shared memory, and one
//      sum-reduction that counts the disparities below the uniqueness threshold.
// Lane q keeps these raw outcomes of pixel q mod 32; everything scalar (threshold test, sub-pixel division, right-view key,
// store) is done for 32 pixels at a time, one pixel per lane (wta_flush).
template <int NW>
__device__ __forceinline__ uint32_t wta_stage_a(const uint32_t (&S)[NW], int lane)
{
    constexpr uint32_t HALF = NW * 32;
    const uint32_t d0 = (uint32_t)lane * NW;
    uint32_t kl = 0xffffffffu, kh = 0xffffffffu;   // padding never beats a real minimum below 32767
#pragma unroll
    for (int j = 0; j < NW; ++j) {
        kl = min(kl, S[j] * 65536u + (d0 + j));
        kh = min(kh, (S[j] & 0xffff0000u) | (d0 + j));
    }
    return __reduce_min_sync(0xffffffffu, min(kl, kh + HALF));   // minS << 16 | best
}
// Uniqueness: a pixel is rejected if some d outside best-1..best+1 has S(d) * (100 - ur) < minS * 100, i.e. S(d) <= thr;
// with the "minus (min Lp + P2)" form of the recurrence (K44 = directions * P2, see mvs_set_sgbm_recurrence) the S held
// here is that form's S plus K44 and the test reads S(d) * (100 - ur) < minS * 100 - K44 * ur = rhs.
__device__ __forceinline__ uint32_t wta_threshold(int minS, const WaveWta& w, int& rhs)
{
    rhs = minS * 100 - w.K44 * w.ur;
    const uint32_t num = (uint32_t)(rhs - 1);
    return w.ur_magic ? __umulhi(num, w.ur_magic) : num;   // meaningful when rhs > 0
}
// staged: the warp's copy of S as 16-bit values in the order of the split layout.  nb = S(best-1) | S(best+1) << 16 (the
// clamped neighbours: the flush ignores the one that does not exist); cnt = 4 * (real disparities with S <= min(thr, 32767):
// per half, bit 15 of (thr | 0x8000) - S) + 4096 * (lanes that hold a saturated sum; only the other form of the recurrence
// looks at it, but a branch would cut the basic block).
template <int NW, bool PAD>
__device__ __forceinline__ void wta_stage_b(const uint32_t (&S)[NW], const uint32_t (&padsent)[NW], int D, const WaveWta& w, uint32_t key,
                                            const uint16_t* staged, uint32_t& nb, int& cnt)
{
    constexpr int HALF = NW * 32;
    const int minS = (int)(key >> 16), best = (int)(key & 0xffffu);
    const int dm = max(best - 1, 0), dp = min(best + 1, D - 1);
    const uint32_t sm = staged[dm >= HALF ? 2 * (dm - HALF) + 1 : 2 * dm], sp = staged[dp >= HALF ? 2 * (dp - HALF) + 1 : 2 * dp];
    nb = sm | (sp << 16);
    int rhs;
    const uint32_t tc = min(wta_threshold(minS, w, rhs), 32767u);
    const uint32_t T2 = (tc | 0x8000u) * 0x10001u;
    uint32_t u[2] = {0u, 0u};
#pragma unroll
    for (int j = 0; j < NW; j += 2) {
        uint32_t t0 = T2 - S[j], t1 = j + 1 < NW ? T2 - S[j + 1 < NW ? j + 1 : j] : 0u;
        if (PAD) {
            t0 &= ~((padsent[j] + 0x00010001u) & 0x80008000u);
            if (j + 1 < NW) t1 &= ~((padsent[j + 1 < NW ? j + 1 : j] + 0x00010001u) & 0x80008000u);
        }
        // bytes of u: 0xFF where the flag (bit 15 / 31 of t0, t1) is set -- prmt's sign-replicating selectors (the
        // __byte_perm intrinsic only passes the three index bits of a selector)
        asm("prmt.b32 %0, %1, %2, 0xFDB9;" : "=r"(u[j / 2]) : "r"(t0), "r"(t1));
    }
    uint32_t mx = 0;
#pragma unroll
    for (int j = 0; j < NW; ++j) mx = __vmaxu2(mx, PAD ? S[j] & ~(padsent[j] | (padsent[j] << 1)) : S[j]);   // padding halves -> 0
    const uint32_t sat = ((mx + 0x00010001u) & 0x80008000u) != 0u ? 4096u : 0u;
    cnt = __reduce_add_sync(0xffffffffu, __popc((u[0] & 0x0f0f0f0fu) | (u[1] & 0xf0f0f0f0u)) + sat);
}

// Lane q holds the raw outcome of sweep pixel xs_first + q of image row y: uniqueness decision, sub-pixel fit, right-view
// key, store.
template <bool BWD>
__device__ __forceinline__ void wta_flush(const WaveFrame& fr, const WaveWta& w, int W1, int D, int y, int xs_first, int n, int lane,
                                          uint32_t st_key, uint32_t st_nb, int st_cnt)
{
    if (lane >= n) return;
    const int xs = xs_first + lane, x = BWD ? W1 - 1 - xs : xs;
    const int best = (int)(st_key & 0xffffu), minS = (int)(st_key >> 16);
    const bool has_m = best > 0, has_p = best < D - 1;
    const int sm = has_m ? (int)(st_nb & 0xffffu) : 0, sp = has_p ? (int)(st_nb >> 16) : 0;
    int rhs;
    const uint32_t thr = wta_threshold(minS, w, rhs), tc = min(thr, 32767u);
    const int cnt4 = st_cnt & 4095;
    const bool satp = st_cnt >= 4096;
    // the window's own members among the counted disparities
    const int nwin = ((uint32_t)minS <= tc ? 1 : 0) + (has_m && (uint32_t)sm <= tc ? 1 : 0) + (has_p && (uint32_t)sp <= tc ? 1 : 0);
    const bool not_unique = rhs > 0 && cnt4 != 4 * nwin;
    // A saturated sum (32767 here) only matters where it takes part in a decision: as the minimum, as a neighbour of the
    // minimum, or when the uniqueness threshold reaches it.  Those cases are reported, never guessed around.
    if (w.K44 && (minS >= 32767 || (satp && ((rhs > 0 && thr >= 32767u) || (has_m && sm >= 32767) || (has_p && sp >= 32767))))) *w.sat_flag = 1;
    if (minS >= 32767 || not_unique) return;   // the pixel keeps the invalid value
    const int x2 = x + w.minX1 - best - w.minD;
    atomicMin(fr.disp2key + (size_t)y * w.W + x2, ((uint32_t)minS << 16) | (uint32_t)(0xffff - x));
    int d = best;
    if (0 < d && d < D - 1) {
        const int den = max(sm + sp - 2 * minS, 1);
        d = d * 16 + ((sm - sp) * 16 + den) / (den * 2);
    } else
        d *= 16;
    fr.raw[(size_t)y * w.W + x + w.minX1] = (int16_t)(d + w.minD * 16);
}

// The long form of a pixel's two waits ("pixel written" by the row above, then "pixel consumed" by the row below), out of
// line: the sweep loop first retries each test once with try_wait, which lets the hardware suspend the warp, and comes
// here only when that was not enough.  Bounded in TIME: a wait that expires, here or anywhere else in the grid, raises the
// error flag and ends all waiting (results are then invalid and the host reports an error), so a logic fault cannot
// hang the GPU.  Returns true when waiting was abandoned.  `where`: ring row, band and pixel for the report.
__device__ __noinline__ bool wave_wait_long(uint32_t bf, uint32_t pf, uint32_t okf, uint32_t bc, uint32_t pc, uint32_t okc, int* error, int where)
{
    unsigned long long t0 = 0;
    int spins = 0;
    for (;;) {
        if (!okf) okf = mbar_try_wait(bf, pf);
        if (okf && !okc) okc = mbar_try_wait(bc, pc);
        if (okf && okc) return false;
        if ((++spins & 127) == 0) {
            if (*(volatile int*)error) return true;
            const unsigned long long t = global_ns();
            if (!t0) t0 = t;
            else if (t - t0 > WAVE_TIMEOUT_NS) {
                atomicCAS(error, 0, ((okf ? 2 : 1) << 28) | where);   // kind 1: "written", 2: "consumed"
                return true;
            }
        }
#if WAVE_CONS_SLEEP
        if (okf) __nanosleep(WAVE_CONS_SLEEP);   // only the back-pressure wait is left: this warp is AHEAD of its consumer, so it
#endif                                           // sleeps between polls and leaves the issue slots to the slower warps
    }
}

template <int NW, int DIRMASK, bool PAD, bool BWD, bool WTA>
__global__ void __launch_bounds__(Ring<NW>::THREADS, 1)
wave_kernel(WaveParams p)
{
    using RG = Ring<NW>;
    asm volatile(".reg .pred wave_pf, wave_pc, wave_na, wave_nr;");   // outcomes of the early barrier tests (see early_test_full)
    constexpr bool VERT = (DIRMASK & 0xE) != 0;
    constexpr int WAVE_R = RG::R, WPP = RG::WPP, PF = BWD ? WAVE_PF_BWD : WAVE_PF, PFS = WAVE_PFS;
    constexpr int GS = BWD ? -WPP : WPP;   // word step from pixel to pixel along the sweep
    extern __shared__ __align__(16) uint32_t ring[];
    __shared__ int s_item;
    // [ring row k][p mod NB]: "pixel p of ring row k written" (k = 0: by the helper, else by warp k-1) and
    // "pixel p of ring row k consumed" (by warp k); the pseudo-pixel W1 is the zero slot behind the row
    __shared__ __align__(8) uint64_t s_full[WAVE_R][WAVE_NB];
    __shared__ __align__(8) uint64_t s_cons[WAVE_R][WAVE_NB];
    __shared__ __align__(16) uint32_t s_wta[WTA ? WAVE_R : 1][WTA ? WPP : 4];   // winner-take-all step: a warp's copy of its last sums
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const bool helper = wid == WAVE_R;
    // The warp scheduler favours higher warp ids among eligible warps: upstream rows (whose progress everything below
    // waits for) get the higher ids, so a warp polling its predecessor never takes an issue slot from it -- except the
    // band's last two rows, which get the highest ids of all: the last row also stores the boundary vectors and publishes
    // the band's progress behind a fence, and the whole band moves at its pace (measured: -1.1 %).
    const int r = helper ? WAVE_R : (wid >= WAVE_R - 2 ? wid : WAVE_R - 3 - wid);
    const int W1 = p.W1;
    LaneConst kc;
    kc.P1x2 = wdup16((uint32_t)p.P1); kc.P2x2 = wdup16((uint32_t)p.P2); kc.one = p.one;
    kc.selL = lane == 0 ? 0x1054u : 0x3210u;    // lane 0:  {low: sentinel, high: low half of lane 31's word}
    kc.selR = lane == 31 ? 0x5432u : 0x3210u;   // lane 31: {low: high half of lane 0's word, high: sentinel}
    kc.lm1 = (lane + 31) & 31; kc.lp1 = (lane + 1) & 31;
    const uint32_t ring_s = (uint32_t)__cvta_generic_to_shared(ring);
    const uint32_t full_s = (uint32_t)__cvta_generic_to_shared(s_full), cons_s = (uint32_t)__cvta_generic_to_shared(s_cons);
    bool bars_valid = false;

    uint32_t padsent[NW];
#pragma unroll
    for (int j = 0; j < NW; ++j) {
        const int d0 = lane * NW + j, d1 = d0 + WPP;   // split layout: low half d0, high half d0 + Dp/2
        padsent[j] = (d0 >= p.D ? 0x7fffu : 0u) | (d1 >= p.D ? 0x7fff0000u : 0u);
    }
    // what this warp reads (ring row r) and writes (ring row r + 1): this lane's words of slot 0 of the L1 ring
    const uint32_t* rd = ring + (helper ? 0 : r) * RG::ROW_WORDS + lane * NW;
    uint32_t* wr = ring + (helper || r + 1 >= WAVE_R ? 0 : r + 1) * RG::ROW_WORDS + lane * NW;
    const uint32_t my_full = full_s + (uint32_t)(helper ? 0 : r) * (WAVE_NB * 8u), my_cons = cons_s + (uint32_t)(helper ? 0 : r) * (WAVE_NB * 8u);
    // ring row r + 1; the band's last warp has none: its (unused) tests then look at its own barriers, which are valid objects
    const uint32_t dn_step = helper || r + 1 >= WAVE_R ? 0u : WAVE_NB * 8u;
    const uint32_t dn_full = my_full + dn_step, dn_cons = my_cons + dn_step;

    for (;;) {
        __syncthreads();  // every warp has left the previous band
        if (threadIdx.x == 0) s_item = atomicAdd(p.ticket, 1);
        // re-arm the mbarriers: invalidate all of them, then (after a CTA barrier: an invalidation may drop barrier state
        // another thread has just initialised) initialise them for one arrival per phase
        if (bars_valid) {
            for (int i = threadIdx.x; i < WAVE_R * WAVE_NB; i += blockDim.x) {
                mbar_inval(full_s + (uint32_t)i * 8u);
                mbar_inval(cons_s + (uint32_t)i * 8u);
            }
            __syncthreads();
        }
        for (int i = threadIdx.x; i < WAVE_R * WAVE_NB; i += blockDim.x) {
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
#if defined(WAVE_DBG_NOCHAIN)           // timing experiment (wrong results): every band behaves like the first and the last one
        const bool first_band = true, last_band = true;
#elif defined(WAVE_DBG_NOCHAIN_STORE)   // timing experiment (wrong results): nobody reads the boundary rows, but they are written
        const bool first_band = true, last_band = band == p.nbands - 1;
#else
        const bool first_band = band == 0, last_band = band == p.nbands - 1;
#endif
        const uint32_t* bin = fr.bnd + (size_t)(first_band ? 0 : band - 1) * W1 * 3 * WPP;  // written by the previous band
        uint32_t* bout = fr.bnd + (size_t)band * W1 * 3 * WPP;
        int* my_progress = p.progress + (size_t)fi * p.nbands + band;
        const int* prev_progress = my_progress - (first_band ? 0 : 1);
        bool failed = false;

        // bounded wait for an mbarrier phase; kind 1: "written", 2: "consumed" (back-pressure: this warp is AHEAD of its
        // consumer, so it sleeps between polls and leaves the issue slots to the slower warps of its scheduler).  Bounded in
        // time: a wait that expires, here or anywhere else in the grid, raises the error flag and ends all waiting (results
        // are then invalid and the host reports an error), so a logic fault cannot hang the GPU.
        auto wait_bar = [&](uint32_t bar, uint32_t par, int kind, int pixel) {
            if (mbar_try_wait(bar, par)) return;
            if (failed) return;
            unsigned long long t0 = 0;
            int spins = 0;
            while (!mbar_try_wait(bar, par)) {
                if ((++spins & 127) == 0) {
                    if (*(volatile int*)p.error) { failed = true; break; }
                    const unsigned long long t = global_ns();
                    if (!t0) t0 = t;
                    else if (t - t0 > WAVE_TIMEOUT_NS) {
                        // where it happened: kind, ring row, band, pixel
                        atomicCAS(p.error, 0, (kind << 28) | ((r & 31) << 23) | ((band & 127) << 16) | (pixel & 0xffff));
                        failed = true;
                        break;
                    }
                }
#if WAVE_CONS_SLEEP
                if (kind == 2) __nanosleep(WAVE_CONS_SLEEP);
#endif
            }
        };
        // the two waits of a pixel when the early tests have not both passed: one retry each with try_wait (the hardware may
        // suspend the warp until the phase completes), then the long form.  okf / okc hold what the tests saw.
        auto wait_both = [&](uint32_t bf, uint32_t pf, uint32_t okf, uint32_t bc, uint32_t pc, uint32_t okc, int pixel) {
            if (!okf) okf = mbar_try_wait(bf, pf);
            if (okf && !okc) okc = mbar_try_wait(bc, pc);
            if (__builtin_expect((okf & okc) == 0u, 0) && !failed)
                failed = wave_wait_long(bf, pf, okf, bc, pc, okc, p.error, ((r & 31) << 23) | ((band & 127) << 16) | (pixel & 0xffff));
        };

        // zero slots prepared before anybody runs: the row above the image, and "x = -1" of the row each warp reads
        if (helper && VERT && first_band) {
            for (int k = lane; k < RG::ROW_WORDS; k += 32) ring[k] = 0;
        }
        if (!helper && VERT) {
            uint32_t* z = ring + r * RG::ROW_WORDS + (RG::D12 - 1) * WPP;  // L1 slot (-1 mod 8) of ring row r
            for (int k = lane; k < WPP; k += 32) z[k] = 0;
        }
        __syncthreads();

        if (helper) {
            if (VERT && !first_band) {  // (first band: everything above the image is zero, warp 0 does not wait)
                // boundary row of the previous band -> ring row 0, as far ahead of warp 0 as the ring allows.  Lane 0 moves a
                // pixel's vectors with bulk asynchronous copies that report their bytes to the pixel's "written" barrier: the
                // barrier's phase completes when the data has landed, nobody has to wait for the copies and arrive for them.
                int seen = 0;
                // (the zero slots above were written with ordinary stores; the copies write the same ring through the
                // asynchronous proxy)
                asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
                constexpr uint32_t VEC = WPP * 4u;
                constexpr uint32_t NVEC = ((DIRMASK >> 1) & 1) + ((DIRMASK >> 2) & 1) + ((DIRMASK >> 3) & 1);
                for (int xp = 0; xp <= W1; ++xp) {
                    // the slots of pixel xp last held pixels xp - 8 (L1: read by warp 0 at its pixel xp - 7) and xp - D3
                    // (L3: read at xp - D3 - 1)
                    if (xp > WAVE_WAR) wait_bar(cons_s + (uint32_t)((xp - WAVE_WAR - 1) & 7) * 8u, (uint32_t)((xp - WAVE_WAR - 1) >> 3) & 1u, 2, xp);
                    const uint32_t bar = full_s + (uint32_t)(xp & 7) * 8u;
                    if (xp < W1) {
                        const int need = xp + 1;
                        if (seen < need && !failed) {
                            unsigned long long t0 = 0;
                            int spins = 0;
                            while ((seen = ld_acquire(prev_progress)) < need) {
                                if ((++spins & 255) == 0) {
                                    if (*(volatile int*)p.error) { failed = true; break; }
                                    const unsigned long long t = global_ns();
                                    if (!t0) t0 = t;
                                    else if (t - t0 > WAVE_TIMEOUT_NS) { atomicCAS(p.error, 0, (3 << 28) | ((band & 127) << 16) | (xp & 0xffff)); failed = true; break; }
                                }
                                __nanosleep(32);
                            }
                            // the other band wrote these vectors with ordinary stores; the copies below read them through the
                            // asynchronous proxy
                            asm volatile("fence.proxy.async.global;" ::: "memory");
                        }
                        if (lane == 0) {
                            const uint32_t* src = bin + (size_t)xp * 3 * WPP;
                            mbar_arrive_expect_tx(bar, NVEC * VEC);
#pragma unroll
                            for (int dir = 1; dir <= 3; ++dir) {
                                if (!(DIRMASK & (1 << dir))) continue;
                                const int depth = dir == 3 ? RG::D3 : RG::D12;
                                bulk_g2s(ring_s + (uint32_t)(((dir - 1) * RG::D12 + (xp & (depth - 1))) * WPP) * 4u, src + (dir - 1) * WPP, VEC, bar);
                            }
                        }
                        // Bulk copies complete in any order, and warp 0 waits for the NEXT pixel's barrier only (for the rows
                        // written by a warp that implies the pixel itself): pixel 0 has landed before pixel 1 is asked for.
                        if (xp == 0) wait_bar(bar, 0u, 1, 0);
                    } else {   // xp == W1: the zero vector behind the last pixel (direction 3 of warp 0's last pixel)
                        if (DIRMASK & 8) {
                            uint32_t* dst = ring + RG::O3 + (xp & (RG::D3 - 1)) * WPP;
                            for (int k = lane; k < WPP; k += 32) dst[k] = 0;
                        }
                        __syncwarp();
                        if (lane == 0) mbar_arrive(bar);
                    }
                }
            }
        } else if (r < rows) {
            const bool to_ring = VERT && r + 1 < rows, to_global = VERT && r + 1 == rows && !last_band;
            const bool wait_above = VERT && !(first_band && r == 0);
            uint32_t L0[NW];
#pragma unroll
            for (int j = 0; j < NW; ++j) L0[j] = 0;
            const int ys = ys0 + r;
            const int y = BWD ? p.H - 1 - ys : ys;
            const size_t gbase = ((size_t)y * W1 + (BWD ? W1 - 1 : 0)) * WPP + lane * NW;
            const uint32_t* cptr = fr.C + gbase;       // pixel x4 of this row (this lane's words)
            const uint32_t* sptr = BWD ? fr.Sin + gbase : nullptr;
            uint32_t* optr = fr.Sout + gbase;
            uint32_t* gout = bout + lane * NW;         // this lane's words of boundary pixel x4, direction 1
            // register prefetch ring: pixel q lives in slot q % PF
            uint32_t Cq[PF][NW], Sq[PFS][NW];
            // All global loads of a warp complete through ONE hardware scoreboard, so the first use of a prefetched register
            // waits for every load issued before it, the youngest included: what counts is the age of the YOUNGEST load at
            // each first use, not the depth of the ring.  A slot is refilled right after its pixel (the load targets the same
            // registers); with a ring as deep as the trip (four pixels: one wait per trip) the slot of the trip's LAST pixel
            // is instead loaded in the middle of the same trip's first pixel, so that no load is younger than 0.7 pixels
            // at a first use.  The lines come from L2 (see WAVE_L2PF), which that age covers.
            auto refill = [&](auto kk, auto& Q, const uint32_t* ptr, auto depth) {
                constexpr int k = decltype(kk)::value, DEPTH = decltype(depth)::value;
                if (DEPTH == WAVE_UNROLL && WAVE_DEFER_LAST) {
                    if (k == 0) ldg_stream<NW>(ptr + (DEPTH - 1) * GS, Q[DEPTH - 1]);
                    if (k != DEPTH - 1) ldg_stream<NW>(ptr + (k + DEPTH) * GS, Q[k % DEPTH]);
                } else {
                    // beyond the row's end this reads up to DEPTH pixels of the neighbouring row or of the guard band the
                    // volumes are allocated with (MVS_VOLUME_GUARD); those values are never used
                    ldg_stream<NW>(ptr + (k + DEPTH) * GS, Q[k % DEPTH]);
                }
            };
#pragma unroll
            for (int q = 0; q < PF; ++q) {
#pragma unroll
                for (int j = 0; j < NW; ++j) Cq[q][j] = 0;
                if (!(PF == WAVE_UNROLL && WAVE_DEFER_LAST && q == PF - 1)) ldg_stream<NW>(cptr + q * GS, Cq[q]);
            }
#pragma unroll
            for (int q = 0; q < PFS; ++q) {
#pragma unroll
                for (int j = 0; j < NW; ++j) Sq[q][j] = 0;
                if (BWD && !(PFS == WAVE_UNROLL && WAVE_DEFER_LAST && q == PFS - 1)) ldg_stream<NW>(sptr + q * GS, Sq[q]);
            }
            const uint32_t lane0 = lane == 0 ? 1u : 0u, lane0_ring = lane == 0 && to_ring ? 1u : 0u;
            // backward sweep: the winner-take-all step runs one pixel behind the sums (its reductions and shuffles then
            // overlap the next pixel's arithmetic); lane q keeps the outcome of pixel q mod 32 until 32 are complete
            const bool store_s = !BWD || fr.Sout != nullptr;
            uint32_t accp[NW], st_key = 0xffffffffu, st_nb = 0u;
            int st_cnt = 0;
            uint32_t* wta_mine = &s_wta[WTA ? r : 0][0];
#pragma unroll
            for (int j = 0; j < NW; ++j) accp[j] = SENT2;
#if WAVE_L2PF
            // byte offset from this lane's pointer to the 128-byte line it prefetches: line `lane` of the eight pixels
            // (4 KB at Dp = 256) that start WAVE_L2PF pixels ahead in sweep order
            const ptrdiff_t l2pf_off = (ptrdiff_t)(BWD ? -(WAVE_L2PF + 7) : WAVE_L2PF) * (WPP * 4) + lane * 128 - lane * NW * 4;
            constexpr int L2PF_LINES = 8 * WPP * 4 / 128;   // lines in eight pixels (32 at Dp = 256)
#endif
            // Barrier tests for the coming pixel are issued one pixel early (their ~90 cycles of latency overlap the
            // arithmetic); a direction that has nothing to wait for counts as passed.  Pixel 0 starts untested.
            early_set(wait_above ? 0u : 1u, to_ring ? 0u : 1u);
            early_row(wait_above ? 0u : 1u, to_ring ? 0u : 1u);
#ifdef WAVE_DBG_PROF
            long long prof_t0 = clock64(), prof_full = 0, prof_cons = 0, prof_n = 0;
#endif

            // The pixel loop is unrolled by UN = 4 (the prefetch depth; half the ring depth: the instruction footprint of
            // an unroll by eight does not fit the instruction cache).  Trip t covers pixels 4t .. 4t+3, which live in
            // half h = t & 1 of the eight-deep rings and barrier sets: "c" pointers address the current half, "o"
            // pointers the other one; par is the phase parity of the current block of eight pixels.
            constexpr int UN = WAVE_UNROLL, WAR = WAVE_WAR;
            for (int x4 = 0; x4 < W1; x4 += UN) {
                const int rem = W1 - x4;                          // pixels left in the row, this trip included
                const uint32_t h = (uint32_t)(x4 >> 2) & 1u, par = (uint32_t)(x4 >> 3) & 1u;
                const int hw = h ? 4 * WPP : 0;                   // word offset of the current half of a ring
                const uint32_t hb = h ? 32u : 0u;                 // byte offset of the current half of a barrier set
                const uint32_t* rdc = rd + hw;
                const uint32_t* rdo = rd + (4 * WPP - hw);
                uint32_t* wrc = wr + hw;
                const uint32_t fullc = my_full + hb, fullo = my_full + (32u - hb);
                const uint32_t consc = dn_cons + hb, conso = dn_cons + (32u - hb);
                const uint32_t arr_full = dn_full + hb, arr_cons = my_cons + hb;
#if WAVE_L2PF
                if (h == 0 && rem >= WAVE_L2PF + 8 && lane < L2PF_LINES) {
                    asm volatile("prefetch.global.L2 [%0];" ::"l"((const char*)cptr + l2pf_off));
                    if (BWD) asm volatile("prefetch.global.L2 [%0];" ::"l"((const char*)sptr + l2pf_off));
                }
#endif
                auto body = [&](auto kk) {
                    constexpr int k = decltype(kk)::value;
                    if (k >= rem) return;
                    const int x = x4 + k;
                    uint32_t (&Cw)[NW] = Cq[k % PF];   // operands of this pixel: loaded PF pixels ago
                    uint32_t (&Sw)[NW] = Sq[k % PFS];
#ifndef WAVE_DBG_NOWAIT   // (timing experiment: wrong results)
                    // (1) the row above is written through pixel x + 1; (2) the slots about to be overwritten were read:
                    // pixel x - WAR - 1 of the row below is finished.  For x <= WAR that barrier is still in its first
                    // phase and the wait on the opposite parity passes at once (the row below cannot finish pixel
                    // x - WAR + 7 before this row has written x - WAR + 8 > x).
                    uint32_t ok_full, ok_cons;
#ifdef WAVE_DBG_PROF   // how long the outcome of each early test takes to arrive (a test suspends the warp while its phase is incomplete)
                    {
                        const long long q0 = clock64();
                        asm volatile("selp.u32 %0, 1, 0, wave_pf;" : "=r"(ok_full)::"memory");
                        const long long q1 = clock64();
                        asm volatile("selp.u32 %0, 1, 0, wave_pc;" : "=r"(ok_cons)::"memory");
                        const long long q2 = clock64();
                        prof_full += q1 - q0;
                        prof_cons += q2 - q1;
                    }
#else
                    early_read(ok_full, ok_cons);
#endif
                    if (__builtin_expect((ok_full & ok_cons) == 0u, 0)) {
                        const uint32_t bf = k == 3 ? fullo : fullc + (k + 1) * 8u, pf = k == 3 ? par ^ h : par;
                        const uint32_t bc = WAR == 6 ? (k == 3 ? conso : consc + (k + 1) * 8u) : (k == 0 ? consc + 24u : conso + (k - 1) * 8u);
                        const uint32_t pc = WAR == 6 ? pf ^ 1u : (k == 0 ? par ^ 1u : par ^ h ^ 1u);
#ifdef WAVE_DBG_PROF
                        const long long w0 = clock64();
                        const bool was_full = ok_full == 0u;
#endif
                        wait_both(bf, pf, ok_full, bc, pc, ok_cons, x);
#ifdef WAVE_DBG_PROF
                        const long long w1 = clock64() - w0;
                        if (was_full) prof_full += w1; else prof_cons += w1;
                        ++prof_n;
#endif
                    }
#endif
                    // All directions advance together, phase by phase, so that the warp always has four independent
                    // dependency chains in flight (a direction alone is one chain of ~15 dependent operations with a
                    // shuffle and a warp reduction in it; the sweep is bound by that latency, not by issue slots).
                    constexpr bool on[4] = {(DIRMASK & 1) != 0, (DIRMASK & 2) != 0, (DIRMASK & 4) != 0, (DIRMASK & 8) != 0};
                    uint32_t a[4][NW], ap[4][NW], el[4], er[4], l[4][NW], m[4], acc[NW];
#pragma unroll
                    for (int j = 0; j < NW; ++j) a[0][j] = L0[j];
                    if (on[1]) ldw<NW>(k == 0 ? rdo + 3 * WPP : rdc + (k - 1) * WPP, a[1]);                   // L1 of pixel x - 1
                    if (on[2]) ldw<NW>(rdc + RG::O2 + k * WPP, a[2]);                                         // L2 of pixel x
                    if (on[3]) ldw<NW>(RG::D3 == 8 ? (k == 3 ? rdo + RG::O3 : rdc + RG::O3 + (k + 1) * WPP)   // L3 of pixel x + 1
                                                   : rd + RG::O3 + ((k + 1) & 3) * WPP, a[3]);
#ifndef WAVE_DBG_NOWAIT
                    // test the NEXT pixel's two barriers now (looked at when the next pixel starts; behind the row's last
                    // pixel the outcome is not used: the next row starts untested)
                    if (VERT) {
                        const uint32_t bf = k >= 2 ? fullo + (k - 2) * 8u : fullc + (k + 2) * 8u, pf = k >= 2 ? par ^ h : par;
                        const uint32_t bc = WAR == 6 ? (k >= 2 ? conso + (k - 2) * 8u : consc + (k + 2) * 8u) : conso + k * 8u;
                        const uint32_t pc = WAR == 6 ? pf ^ 1u : par ^ h ^ 1u;
                        early_test_full(bf, pf);
                        early_test_cons(bc, pc);
                    }
#endif
                    // winner-take-all step of the row's previous pixel, stage A (B follows the sweep's own reductions, C the
                    // barrier arrivals)
                    // (no branch around the stages: they must share a basic block with the sweep's arithmetic to be
                    // interleaved with it.  The row's first pixel runs them on a vector of saturated sums, which gives
                    // "no disparity" for a lane slot that has just been stored.)
                    uint32_t wkey = 0, wnb = 0;
                    int wcnt = 0;
                    if (WTA) wkey = wta_stage_a<NW>(accp, lane);
#pragma unroll
                    for (int d = 0; d < 4; ++d)
                        if (on[d]) {
#pragma unroll
                            for (int j = 0; j < NW; ++j) ap[d][j] = addf(a[d][j], kc.P1x2, kc.one);
                        }
                    // neighbours across the lane edge: word NW-1 of lane-1 holds d-1 of this lane's word 0 (both halves), except
                    // in lane 0, where the low half's neighbour is d = -1 (sentinel) and the high half's is lane 31's LOW half
#pragma unroll
                    for (int d = 0; d < 4; ++d)
                        if (on[d]) {
                            el[d] = __shfl_sync(0xffffffffu, ap[d][NW - 1], kc.lm1);
                            er[d] = __shfl_sync(0xffffffffu, ap[d][0], kc.lp1);
                        }
#pragma unroll
                    for (int d = 0; d < 4; ++d)
                        if (on[d]) {
                            el[d] = __byte_perm(el[d], SENT2, kc.selL);
                            er[d] = __byte_perm(er[d], SENT2, kc.selR);
                        }
#pragma unroll
                    for (int d = 0; d < 4; ++d)
                        if (on[d]) {
#pragma unroll
                            for (int j = 0; j < NW; ++j) {
                                const uint32_t lo = j == 0 ? el[d] : ap[d][j == 0 ? 0 : j - 1];
                                const uint32_t hi = j == NW - 1 ? er[d] : ap[d][j == NW - 1 ? j : j + 1];
                                uint32_t v = addf(__vimin3_u16x2(lo, hi, a[d][j]), Cw[j], kc.one);   // per half <= P2 + Cmax <= 32767
                                if (PAD) v |= padsent[j];   // padding disparities (d >= D) act as the reference's d = D sentinel
                                l[d][j] = v;
                            }
                        }
#pragma unroll
                    for (int d = 0; d < 4; ++d)
                        if (on[d]) {
                            uint32_t mm;
                            if (NW == 4) mm = __vminu2(__vimin3_u16x2(l[d][0], l[d][1 % NW], l[d][2 % NW]), l[d][3 % NW]);
                            else if (NW == 3) mm = __vimin3_u16x2(l[d][0], l[d][1 % NW], l[d][2 % NW]);
                            else if (NW == 2) mm = __vminu2(l[d][0], l[d][1 % NW]);
                            else mm = l[d][0];
                            m[d] = __vminu2(mm, __byte_perm(mm, mm, 0x1032));   // min of the two halves, in both halves
                        }
#pragma unroll
                    for (int d = 0; d < 4; ++d)
                        if (on[d]) m[d] = __reduce_min_sync(0xffffffffu, m[d]);   // (v << 16 | v) is monotone in v: still duplicated
                    if (WTA) {
                        if (!VERT) __syncwarp();   // (with vertical directions the warp barrier before the arrivals has ordered the staged copy)
                        wta_stage_b<NW, PAD>(accp, padsent, p.D, p.wta, wkey, (const uint16_t*)wta_mine, wnb, wcnt);
                    }
                    // S accumulation (the reference saturates once per pass; all terms are non-negative, so clamping as we go is the same)
#pragma unroll
                    for (int j = 0; j < NW; ++j) {
                        uint32_t s = on[1] ? __viaddmin_u16x2(l[0][j], l[1][j], SENT2) : l[0][j];
                        if (on[2] && on[3]) s = __viaddmin_u16x2(s, __viaddmin_u16x2(l[2][j], l[3][j], SENT2), SENT2);
                        else if (on[2]) s = __viaddmin_u16x2(s, l[2][j], SENT2);
                        if (BWD) s = __viaddmin_u16x2(Sw[j], s, SENT2);   // S = min(S_in + sum L, 32767)
                        acc[j] = s;
                    }
                    if (store_s) stg_stream<NW>(optr + k * GS, acc);
                    if (WTA) {
                        // (the stage-B reads of the previous vector precede this store in program order)
                        if (lane == ((x - 1) & 31)) { st_key = wkey; st_nb = wnb; st_cnt = wcnt; }
                        stw<NW>(wta_mine + lane * NW, acc);
#pragma unroll
                        for (int j = 0; j < NW; ++j) accp[j] = acc[j];
                    }
                    // refill the prefetch slots AFTER their last use, so that the loads can target the same registers (issued
                    // earlier they get temporaries, and the move into place waits for the load one pixel later)
                    refill(kk, Cq, cptr, IC<PF>());
                    if (BWD) refill(kk, Sq, sptr, IC<PFS>());
#pragma unroll
                    for (int d = 0; d < 4; ++d)
                        if (on[d]) {
                            const uint32_t negm = __byte_perm(0u - m[d], 0u, 0x1010);   // -min per half (low half of the 32-bit negation)
#pragma unroll
                            for (int j = 0; j < NW; ++j) l[d][j] = __viaddmin_s16x2(l[d][j], negm, kc.P2x2);   // l now holds Ln
                        }
#pragma unroll
                    for (int j = 0; j < NW; ++j) L0[j] = l[0][j];
                    if (on[1]) { if (to_ring) stw<NW>(wrc + k * WPP, l[1]); if (to_global) stg_stream<NW>(gout + k * 3 * WPP, l[1]); }
                    if (on[2]) { if (to_ring) stw<NW>(wrc + RG::O2 + k * WPP, l[2]); if (to_global) stg_stream<NW>(gout + (k * 3 + 1) * WPP, l[2]); }
                    if (on[3]) {
                        if (to_ring) stw<NW>((RG::D3 == 8 ? wrc : wr) + RG::O3 + k * WPP, l[3]);
                        if (to_global) stg_stream<NW>(gout + (k * 3 + 2) * WPP, l[3]);
                    }
                    if (to_global && ((k == UN - 1 && ((x + 1) & p.publish_mask) == 0) || k + 1 == rem)) {
                        // the warp barrier orders every lane's boundary stores before lane 0's release store
                        // (release is cumulative over what the warp barrier made visible to lane 0)
                        __syncwarp();
                        st_release_if(my_progress, x + 1, lane0);
                    }
                    if (VERT) {
                        // publish: ring row r+1 holds pixel x; ring row r has been read for pixel x.  The warp barrier
                        // puts every lane's ring reads and writes before lane 0's arrivals (release).
                        __syncwarp();
                        mbar_arrive_if(arr_full + (uint32_t)k * 8u, lane0_ring);
                        mbar_arrive_if(arr_cons + (uint32_t)k * 8u, lane0);
                    }
                    if (WTA && k == 0 && (x4 & 31) == 0 && x4 > 0) wta_flush<BWD>(fr, p.wta, W1, p.D, y, x4 - 32, 32, lane, st_key, st_nb, st_cnt);
                };
                body(IC<0>()); body(IC<1>()); body(IC<2>()); body(IC<3>());
                cptr += UN * GS;
                if (BWD) sptr += UN * GS;
                optr += UN * GS;
                gout += UN * 3 * WPP;
            }
            if (WTA) {   // the row's last pixel, and the outcomes not stored yet
                uint32_t wnb;
                int wcnt;
                const uint32_t wkey = wta_stage_a<NW>(accp, lane);
                __syncwarp();
                wta_stage_b<NW, PAD>(accp, padsent, p.D, p.wta, wkey, (const uint16_t*)wta_mine, wnb, wcnt);
                if (lane == ((W1 - 1) & 31)) { st_key = wkey; st_nb = wnb; st_cnt = wcnt; }
                wta_flush<BWD>(fr, p.wta, W1, p.D, y, (W1 - 1) & ~31, ((W1 - 1) & 31) + 1, lane, st_key, st_nb, st_cnt);
                __syncwarp();   // the staged copy is overwritten by the next band's first pixel
            }
#ifdef WAVE_DBG_PROF
            if (lane == 0) {
                const int slot = (to_global ? 16 : r) * 4;
                atomicAdd(p.prof + slot, (unsigned long long)(clock64() - prof_t0));
                atomicAdd(p.prof + slot + 1, (unsigned long long)prof_full);
                atomicAdd(p.prof + slot + 2, (unsigned long long)prof_cons);
                atomicAdd(p.prof + slot + 3, (unsigned long long)prof_n);
            }
#endif
            if (to_ring) {
                // zero vector "x = W1" for direction 3 of the row below (its predecessor is outside the image);
                // the pseudo-pixel W1 is what the row below waits for at its last pixel
                if (DIRMASK & 8) {
                    // that L3 slot held pixel W1 - D3, read by the row below at its pixel W1 - D3 - 1
                    if (W1 > RG::D3) wait_bar(dn_cons + (uint32_t)((W1 - RG::D3 - 1) & 7) * 8u, (uint32_t)((W1 - RG::D3 - 1) >> 3) & 1u, 2, W1);
                    uint32_t* d = wr - lane * NW + RG::O3 + (W1 & (RG::D3 - 1)) * WPP;
                    for (int k = lane; k < WPP; k += 32) d[k] = 0;
                }
                __syncwarp();
                if (lane == 0) mbar_arrive(dn_full + (uint32_t)(W1 & 7) * 8u);
            }
        }
    }
}

template <int NW, int DIRMASK, bool PAD, bool BWD, bool WTA>
static int launch_wave(mvs_ctx* ctx, const WaveParams& p)
{
    const int smem = Ring<NW>::WORDS * 4;
    // per instantiation and device (function attributes are per device; doing this twice from two threads is harmless)
    static bool attr_set[MVS_MAX_DEVICES] = {};
    const int dv = ctx->device >= 0 && ctx->device < MVS_MAX_DEVICES ? ctx->device : -1;
    if (dv < 0 || !attr_set[dv]) {
        MVS_CUDA(ctx, cudaFuncSetAttribute(wave_kernel<NW, DIRMASK, PAD, BWD, WTA>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
        if (dv >= 0) attr_set[dv] = true;
    }
    const int grid = min(p.nbands * p.F, ctx->sm_count);   // one CTA per SM (__launch_bounds__(.., 1), > half of the shared memory)
    wave_kernel<NW, DIRMASK, PAD, BWD, WTA><<<grid, Ring<NW>::THREADS, smem, ctx->stream>>>(p);
    MVS_LAUNCH_CHECK(ctx, "wave_kernel");
    return MVS_OK;
}

template <int DIRMASK, bool BWD, bool WTA>
static int launch_wave_nw(mvs_ctx* ctx, const SgbmGeom& g, const WaveParams& p)
{
    bool pad = g.Dp != g.D;
#if defined(WAVE_ONLY_NW4)   // quick experimental builds
    if (g.Dp == 256 && !pad) return launch_wave<4, DIRMASK, false, BWD, WTA>(ctx, p);
    return mvs_fail(ctx, MVS_ERR_UNSUPPORTED, "aggregate: experimental build (Dp = 256 only)");
#elif defined(WAVE_ONLY_NW)
    if (g.Dp == 64 * WAVE_ONLY_NW && !pad && !WTA) return launch_wave<WAVE_ONLY_NW, DIRMASK, false, BWD, false>(ctx, p);
    return mvs_fail(ctx, MVS_ERR_UNSUPPORTED, "aggregate: experimental build (one Dp only)");
#else
    if constexpr (WTA) {   // the winner-take-all step is in the sweep for the wide disparity ranges only (sgbm_wta_in_sweep)
        if (g.Dp == 256) return pad ? launch_wave<4, DIRMASK, true, BWD, true>(ctx, p) : launch_wave<4, DIRMASK, false, BWD, true>(ctx, p);
        if (g.Dp == 192) return pad ? launch_wave<3, DIRMASK, true, BWD, true>(ctx, p) : launch_wave<3, DIRMASK, false, BWD, true>(ctx, p);
    } else {
        switch (g.Dp / 64) {
        case 1: return pad ? launch_wave<1, DIRMASK, true, BWD, false>(ctx, p) : launch_wave<1, DIRMASK, false, BWD, false>(ctx, p);
        case 2: return pad ? launch_wave<2, DIRMASK, true, BWD, false>(ctx, p) : launch_wave<2, DIRMASK, false, BWD, false>(ctx, p);
        case 3: return pad ? launch_wave<3, DIRMASK, true, BWD, false>(ctx, p) : launch_wave<3, DIRMASK, false, BWD, false>(ctx, p);
        case 4: return pad ? launch_wave<4, DIRMASK, true, BWD, false>(ctx, p) : launch_wave<4, DIRMASK, false, BWD, false>(ctx, p);
        }
    }
    return mvs_fail(ctx, MVS_ERR_UNSUPPORTED, "aggregate: Dp=%d", g.Dp);
#endif
}

int sgbm_aggregate_batch_dev(mvs_ctx* ctx, const SgbmGeom& g, int F, const int16_t* const* C, int16_t* const* S, const SgbmWtaOut* wta)
{
    if (F < 1 || F > MVS_MAX_BATCH) return mvs_fail(ctx, MVS_ERR_ARG, "aggregate: batch of %d frames (1..%d)", F, MVS_MAX_BATCH);
    const int nbands = cdiv(g.H, wave_rows(g.Dp / 64));
    const size_t wpp = (size_t)g.Dp / 2;
    const int ctl_pass = F * nbands + 2 * WAVE_CTL_PAD;
    const size_t bnd_sz = (size_t)nbands * g.W1 * 3 * wpp;
    MVS_ARENA(ctx, uint32_t, bnd, "sgbm.wave.bnd", bnd_sz * F);
    MVS_ARENA(ctx, int, ctl, "sgbm.wave.ctl", (size_t)2 * ctl_pass);
    MVS_ARENA(ctx, int, err, "sgbm.wave.err", 1);
    if (!ctx->wave_err_cleared) {
        MVS_CUDA(ctx, cudaMemsetAsync(err, 0, sizeof(int), ctx->stream));
        ctx->wave_err_cleared = true;
    }
    WaveParams p;
    memset(&p, 0, sizeof(p));
    p.error = err; p.F = F; p.one = 1u;
    p.publish_mask = (F * nbands <= ctx->sm_count ? WAVE_PUBLISH_FEW : WAVE_PUBLISH) - 1;
    p.W1 = g.W1; p.H = g.H; p.D = g.D; p.P1 = g.P1; p.P2 = g.P2; p.nbands = nbands;
    MVS_CUDA(ctx, cudaMemsetAsync(ctl, 0, sizeof(int) * 2 * ctl_pass, ctx->stream));
    int rc;
#ifdef WAVE_DBG_PROF
    MVS_ARENA(ctx, unsigned long long, prof, "sgbm.wave.prof", 2 * 32 * 4);
    MVS_CUDA(ctx, cudaMemsetAsync(prof, 0, sizeof(unsigned long long) * 2 * 32 * 4, ctx->stream));
    p.prof = prof;
#endif
    // forward sweep
    for (int f = 0; f < F; ++f) {
        p.f[f].C = (const uint32_t*)C[f]; p.f[f].Sin = nullptr; p.f[f].Sout = (uint32_t*)S[f];
        p.f[f].bnd = bnd + bnd_sz * f;
    }
    p.ticket = ctl; p.progress = ctl + WAVE_CTL_PAD;
    stage_begin(ctx, "aggregate_fwd");
    rc = g.mode == MVS_MODE_HH4 ? launch_wave_nw<0x5, false, false>(ctx, g, p) : launch_wave_nw<0xF, false, false>(ctx, g, p);
    stage_end(ctx, "aggregate_fwd");
    if (rc) return rc;
    // backward sweep: the same recurrence on the image flipped in x and y; with the winner-take-all step inside, S is
    // not written again unless the caller wants to look at it
    for (int f = 0; f < F; ++f) {
        p.f[f].Sin = (const uint32_t*)S[f];
        if (wta) {
            p.f[f].raw = wta->raw[f]; p.f[f].disp2key = wta->disp2key[f];
            if (!wta->store_S) p.f[f].Sout = nullptr;
        }
    }
    if (wta) {
        MVS_ARENA(ctx, int, sat, "sgbm.sat44", 1);
        const uint32_t udiv = (uint32_t)(100 - g.ur);
        p.wta.W = g.W; p.wta.minX1 = g.minX1; p.wta.minD = g.minD; p.wta.ur = g.ur; p.wta.K44 = g.K44;
        p.wta.ur_magic = udiv <= 1 ? 0u : (uint32_t)(((1ull << 32) + udiv - 1) / udiv);
        p.wta.sat_flag = sat;
    }
    p.ticket = ctl + ctl_pass; p.progress = ctl + ctl_pass + WAVE_CTL_PAD;
#ifdef WAVE_DBG_PROF
    p.prof = prof + 32 * 4;
#endif
    stage_begin(ctx, "aggregate_bwd");
    if (g.mode == MVS_MODE_HH) rc = wta ? launch_wave_nw<0xF, true, true>(ctx, g, p) : launch_wave_nw<0xF, true, false>(ctx, g, p);
    else if (g.mode == MVS_MODE_HH4) rc = wta ? launch_wave_nw<0x5, true, true>(ctx, g, p) : launch_wave_nw<0x5, true, false>(ctx, g, p);
    else rc = wta ? launch_wave_nw<0x1, true, true>(ctx, g, p) : launch_wave_nw<0x1, true, false>(ctx, g, p);  // MODE_SGBM: only the (x+1, y) direction
    stage_end(ctx, "aggregate_bwd");
#ifdef WAVE_DBG_PROF
    {
        unsigned long long h[2 * 32 * 4];
        cudaMemcpyAsync(h, prof, sizeof(h), cudaMemcpyDeviceToHost, ctx->stream);
        cudaStreamSynchronize(ctx->stream);
        for (int s = 0; s < 2; ++s)
            for (int r = 0; r <= 16; ++r) {
                const unsigned long long* q = h + (s * 32 + r) * 4;
                if (q[0]) fprintf(stderr, "prof sweep %d row %2d: total %8.2f Mclk  wait-above %5.1f%%  wait-below %5.1f%%  slow paths %llu\n", s, r,
                                  q[0] * 1e-6, 100.0 * q[1] / q[0], 100.0 * q[2] / q[0], q[3]);
            }
    }
#endif
    return rc;
}

int sgbm_aggregate_dev(mvs_ctx* ctx, const SgbmGeom& g, const int16_t* C, int16_t* S)
{
    return sgbm_aggregate_batch_dev(ctx, g, 1, &C, &S, nullptr);
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
    // the other flag of the chain: a saturated cost sum under the (min Lp + P2) form of the recurrence (wta_post.cu)
    auto st = ctx->arena.bufs.find("sgbm.sat44");
    if (st != ctx->arena.bufs.end() && st->second.p && ctx->sgbm_subtract_p2) {
        int sat = 0;
        MVS_CUDA(ctx, cudaMemcpyAsync(&sat, st->second.p, sizeof(int), cudaMemcpyDeviceToHost, stream));
        MVS_CUDA(ctx, cudaStreamSynchronize(stream));
        if (sat) {
            cudaMemsetAsync(st->second.p, 0, sizeof(int), stream);
            return mvs_fail(ctx, MVS_ERR_UNSUPPORTED, "SGBM: a cost sum saturated; under the (min Lp + P2) form of the recurrence the result is "
                                                      "not guaranteed to equal OpenCV's there (it is exact wherever nothing saturates)");
        }
    }
    if (h) {
        cudaMemsetAsync(it->second.p, 0, sizeof(int), stream);
        return mvs_fail(ctx, MVS_ERR_CUDA, "aggregation: a bounded wait expired (results invalid; kind %d, ring row %d, band %d, pixel %d)",
                        (h >> 28) & 3, (h >> 23) & 31, (h >> 16) & 127, h & 0xffff);
    }
    return MVS_OK;
}

// Test tap: raise the flag a bounded wait would raise, so that the host-side reporting path (mvs_sync, the fetch entry
// points) can be checked without provoking a real timeout.
extern "C" int mvs_test_raise_wave_error(mvs_ctx* ctx, int code)
{
    if (!ctx) return MVS_ERR_ARG;
    MVS_CUDA(ctx, cudaSetDevice(ctx->device));
    MVS_ARENA(ctx, int, err, "sgbm.wave.err", 1);
    ctx->wave_err_cleared = true;
    MVS_CUDA(ctx, cudaMemcpyAsync(err, &code, sizeof(int), cudaMemcpyHostToDevice, ctx->stream));
    MVS_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return MVS_OK;
}

// Which form of the path recurrence the SGBM chain reproduces: 0 = L = C + min(..) - min Lp, what the
// OpenCV importable here (4.13) computes and the default; 1 = L = C + min(..) - (min Lp + P2), the form of older OpenCV
// sources.  The second equals the first shifted by P2 per direction, so the sweeps are shared and only the uniqueness
// test of the winner-take-all step differs; it is exact wherever no cost sum saturates, and a saturated sum is reported
// (MVS_ERR_UNSUPPORTED from the next synchronising call) instead of being guessed around.
extern "C" int mvs_set_sgbm_recurrence(mvs_ctx* ctx, int subtract_p2)
{
    if (!ctx) return MVS_ERR_ARG;
    MVS_CUDA(ctx, cudaSetDevice(ctx->device));
    MVS_ARENA(ctx, int, sat, "sgbm.sat44", 1);
    MVS_CUDA(ctx, cudaMemsetAsync(sat, 0, sizeof(int), ctx->stream));
    ctx->sgbm_subtract_p2 = subtract_p2 != 0;
    return MVS_OK;
}
