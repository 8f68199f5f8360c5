// Internal declarations shared by the libmvstereo translation units (not part of the C ABI).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>

#include <map>
#include <string>
#include <vector>

#include "../../include/mvstereo.h"

#define MVS_MAX_DEVICES 64   // per-device one-time state (function attributes are per device)
#define MVS_MAX_SLOTS 64
#define MVS_MAX_BATCH 32  // camera-frames per batched launch

// Grow-only named device buffers: allocated on first use, reused on every later frame, freed with the
// context.  Replaces the reference's per-call cudaMalloc/cudaFree (strategies/sgm_disp.cpp:36-60,173).
struct Arena {
    struct Buf { void* p = nullptr; size_t cap = 0; };
    std::map<std::string, Buf> bufs;
    size_t total = 0;
};

struct StageTimer {
    cudaEvent_t beg = nullptr, end = nullptr;
    bool used = false;
};

struct FrameSlot {
    int w = 0, h = 0, hw = 0, hh = 0, n_labels = 0;   // n_labels > 0: the slot holds a super-pixel table
    bool valid = false;
};

// maps of a camera sensor in the packed fixed-point form of the cubic remap (mvs_maps_register)
struct RegisteredMaps {
    void* packed = nullptr;   // [h][w] uint32 {X | Y << 16}, then [h][w] uint16 weight-table index
    int w = 0, h = 0;
};

struct mvs_ctx {
    int device = 0;
    std::vector<RegisteredMaps> maps;
    cudaStream_t stream = nullptr;
    // Host<->device copies of the frame path run on their own stream so that the upload of the next batch and the
    // download of the previous one overlap the kernels of the current one (ordered by the events below).
    cudaStream_t copy_stream = nullptr;
    cudaEvent_t ev_h2d = nullptr;                       // uploads of the batch being enqueued are complete
    // The frame path runs the second half of a batch (WTA, post-filters, cloud, histogram) on the tail stream so that it
    // overlaps the aggregation sweeps of the NEXT batch on the main stream (the sweeps are latency-bound and leave issue
    // slots, registers and thread slots free on every SM).  S volumes alternate between two sets.
    cudaStream_t tail_stream = nullptr;
    cudaEvent_t ev_front_done = nullptr;                // cost + aggregation of the batch being enqueued are complete
    cudaEvent_t ev_tail_done[2] = {nullptr, nullptr};   // last tail that read S set 0 / 1
    bool tail_done_recorded[2] = {false, false};
    int s_set = 0;                                      // S set of the next batch
    cudaEvent_t ev_slot_done[MVS_MAX_SLOTS] = {};       // last kernel that touched the slot's buffers
    bool slot_done_recorded[MVS_MAX_SLOTS] = {};
    cudaEvent_t ev_slot_fetched[MVS_MAX_SLOTS] = {};    // last download of the slot's results
    bool slot_fetch_recorded[MVS_MAX_SLOTS] = {};
    Arena arena;
    char err[512];
    std::map<std::string, StageTimer> timers;
    FrameSlot slots[MVS_MAX_SLOTS];
    int16_t* cubic_tab_dev = nullptr;  // 32*32*16 Q15 bicubic weights
    int16_t* interp_tab_dev[5] = {nullptr, nullptr, nullptr, nullptr, nullptr};  // by cv::INTER_* value: 32*32*K*K Q15 weights
    int sm_count = 148;
    bool wave_err_cleared = false;
    bool sgbm_subtract_p2 = false;   // mvs_set_sgbm_recurrence
    bool sat44_cleared = false;
    long long launches = 0;  // kernels launched through this context (bench.py reports it)
};

int mvs_fail(mvs_ctx* ctx, int code, const char* fmt, ...);
// The stage entry points work on the main stream and share arena buffers (S volumes, scratch) with the second half of
// the frame path, which runs on the tail stream: they order themselves behind it first.
static inline int main_stream_waits_for_tail(mvs_ctx* ctx)
{
    for (int i = 0; i < 2; ++i)
        if (ctx->tail_done_recorded[i] && cudaStreamWaitEvent(ctx->stream, ctx->ev_tail_done[i], 0) != cudaSuccess)
            return mvs_fail(ctx, MVS_ERR_CUDA, "cudaStreamWaitEvent failed");
    return MVS_OK;
}
// Every stage helper launches on ctx->stream; the frame path enqueues its second half on the tail stream by swapping it in.
struct StreamSwap {
    mvs_ctx* c;
    cudaStream_t saved;
    StreamSwap(mvs_ctx* ctx, cudaStream_t s) : c(ctx), saved(ctx->stream) { ctx->stream = s; }
    ~StreamSwap() { c->stream = saved; }
};
void* arena_get(mvs_ctx* ctx, const char* name, size_t bytes);  // nullptr on failure (error recorded)
void stage_begin(mvs_ctx* ctx, const char* stage);
void stage_end(mvs_ctx* ctx, const char* stage);

#define MVS_CUDA(ctx, call)                                                                         \
    do {                                                                                            \
        cudaError_t e__ = (call);                                                                   \
        if (e__ != cudaSuccess)                                                                     \
            return mvs_fail(ctx, MVS_ERR_CUDA, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e__), \
                            __FILE__, __LINE__);                                                    \
    } while (0)

// n = number of kernel launches covered by this check
#define MVS_LAUNCH_CHECK_N(ctx, name, n)                                                            \
    do {                                                                                            \
        (ctx)->launches += (n);                                                                     \
        cudaError_t e__ = cudaGetLastError();                                                       \
        if (e__ != cudaSuccess)                                                                     \
            return mvs_fail(ctx, MVS_ERR_CUDA, "launch of %s failed: %s", name, cudaGetErrorString(e__)); \
    } while (0)
#define MVS_LAUNCH_CHECK(ctx, name) MVS_LAUNCH_CHECK_N(ctx, name, 1)

#define MVS_ARENA(ctx, T, var, name, count)                                   \
    T* var = (T*)arena_get(ctx, name, sizeof(T) * (size_t)(count));          \
    if (!var) return MVS_ERR_NOMEM

static inline int cdiv(long long a, long long b) { return (int)((a + b - 1) / b); }

// ---- geometry of the SGBM volumes ------------------------------------------------------------------
// Device cost volumes C and S are uint16 [H][W1][Dp] in the SPLIT layout: the 32-bit word w (0 <= w < Dp/2) of a
// pixel holds disparity w in its low half and disparity w + Dp/2 in its high half.  In packed u16x2 arithmetic the
// d-1 / d+1 neighbours of a word are then the previous / next WORD (no byte permutes in the aggregation sweep).
__host__ __device__ static inline int split_index(int d, int Dp) { return d >= Dp / 2 ? 2 * (d - Dp / 2) + 1 : 2 * d; }  // halfword of disparity d
// The aggregation sweep prefetches a few pixels past the end of a row without a bounds test: the volumes are allocated
// with this many bytes of guard on either side (never used as data).
#define MVS_VOLUME_GUARD 4096
struct SgbmGeom {
    int W, H, D, minD, minX1, W1;
    int Dp;  // disparities per pixel in the device volumes (D rounded up to a multiple of 64)
    int SW2, SH2, ftzero, P1, P2, ur, d12, INV;
    int mode;
    int K44;  // 0: OpenCV 4.13's recurrence (minus min Lp); directions * P2: the older form (minus (min Lp + P2)), see mvs_set_sgbm_recurrence
    size_t cells() const { return (size_t)H * W1 * Dp; }
};
int sgbm_geom(mvs_ctx* ctx, const mvs_sgbm_params* p, int w, int h, SgbmGeom* g);
int16_t* sgbm_volume(mvs_ctx* ctx, const char* name, const SgbmGeom& g);   // arena volume with guard bands; nullptr on failure

// ---- device-pointer stage entry points (all asynchronous on ctx->stream) ------------------------------
int remap_cubic_bgr8_batch_dev(mvs_ctx* ctx, int n, const uint8_t* const* src, int sw, int sh, size_t sstep, const float* const* mapx,
                               const float* const* mapy, size_t mstep, int w, int h, const uint8_t border[3], uint8_t* const* dst,
                               size_t dstep);
int remap_bgr8_dev(mvs_ctx* ctx, const uint8_t* src, int sw, int sh, size_t sstep, const float* mapx,
                   const float* mapy, size_t mstep, int w, int h, int interp, const uint8_t border[3], uint8_t* dst,
                   size_t dstep);
int presmooth_batch_dev(mvs_ctx* ctx, int n, const uint8_t* const* src, int channels, int w, int h, size_t step, double sigma,
                        uint8_t* const* dst, size_t dstep);
int presmooth_dev(mvs_ctx* ctx, const char* tag, const uint8_t* src, int channels, int w, int h, size_t step,
                  double sigma, uint8_t* dst, size_t dstep);
// grey images (step == w) -> disparity16 (step == w*2); volumes stay in the arena ("sgbm.C", "sgbm.S")
int sgbm_compute16_dev(mvs_ctx* ctx, const mvs_sgbm_params* p, const uint8_t* l, const uint8_t* r, int w, int h,
                       int16_t* disp16, int16_t* raw16 /* optional pre-median tap */);
int disp_to_float_dev(mvs_ctx* ctx, const int16_t* d16, int w, int h, float* out);
int point_cloud_batch_dev(mvs_ctx* ctx, int F, const mvs_bcam* const* bcam, const float* const* disp, int w, int h, int left_is_ref,
                          double* const* Xs, int32_t* const* xy, int32_t* const* lookup, int32_t* const* n_dev, double* const* centroid_dev);
int point_cloud_dev(mvs_ctx* ctx, const mvs_bcam* bcam, const float* disp, int w, int h, int left_is_ref, double* Xs,
                    int32_t* xy, int32_t* lookup, int32_t* n_dev, double* centroid_dev);
int floor_hist_camera_dev(mvs_ctx* ctx, const mvs_hist_params* p, const double et[8], const double aabb[4],
                          const double cam_center[3], const double* Xs, const int32_t* xy, const int32_t* n_dev,
                          int n_host /* used when n_dev == nullptr */, const int32_t* labels, int lw, int lh,
                          const double* still, int n_labels, float* cam_hist, int hw, int hh,
                          mvs_sparse_hist_cell* cells = nullptr /* slic_to_hist_xy, n_labels entries, optional */);
int disp_merge_ref_dev(mvs_ctx* ctx, float* disp, const float* ref, const int32_t* labels, const uint8_t* still_cutoff,
                       int n_labels, int w, int h);
int hist_post_dev(mvs_ctx* ctx, float* hist, int hw, int hh, float sigma, float minc, float maxc, const char* scratch);
int hist_sum_dev(mvs_ctx* ctx, const float* cam_hists, int n_cams, int hw, int hh, float* hist);

// stage kernels of the SGBM chain (cost.cu, aggregate.cu, wta_post.cu)
int sgbm_cost_volume_dev(mvs_ctx* ctx, const SgbmGeom& g, const uint8_t* l, const uint8_t* r, int16_t* C);
int sgbm_cost_volume_batch_dev(mvs_ctx* ctx, const SgbmGeom& g, int F, const uint8_t* const* l, const uint8_t* const* r,
                               int16_t* const* C);
int sgbm_aggregate_dev(mvs_ctx* ctx, const SgbmGeom& g, const int16_t* C, int16_t* S);       // fused wavefront sweeps
int sgbm_aggregate_wave_status(mvs_ctx* ctx, cudaStream_t stream = nullptr);  // nullptr: the compute stream
// wta != nullptr: the backward sweep also does the winner-take-all step of cv::StereoSGBM (raw disparities into raw[f],
// [H][W], initialised to INV; right-view keys into disp2key[f], [H][W], initialised to 0x7fffffff; the left-right check
// follows in sgbm_lr_post_batch_dev) and leaves S as the forward sweep wrote it unless store_S is set
struct SgbmWtaOut {
    int16_t* raw[MVS_MAX_BATCH];
    uint32_t* disp2key[MVS_MAX_BATCH];
    int16_t* S[MVS_MAX_BATCH];   // the aggregated volumes (final only when !in_sweep or store_S)
    bool in_sweep;               // the backward sweep did the winner-take-all step (sgbm_wta_in_sweep)
    int sset;                    // buffer set (0 or 1) of raw / disp2key and, when !in_sweep, of S
    bool store_S;
};
// where the winner-take-all step runs: inside the backward sweep from 192 disparities (Dp), as its own kernel below that (wta_post.cu)
#ifndef MVS_WTA_IN_SWEEP_FROM
#define MVS_WTA_IN_SWEEP_FROM 192
#endif
static inline bool sgbm_wta_in_sweep(const SgbmGeom& g) { return g.Dp >= MVS_WTA_IN_SWEEP_FROM; }
int sgbm_wta_from_s_batch_dev(mvs_ctx* ctx, const SgbmGeom& g, int F, const SgbmWtaOut& wta);
int sgbm_aggregate_batch_dev(mvs_ctx* ctx, const SgbmGeom& g, int F, const int16_t* const* C, int16_t* const* S, const SgbmWtaOut* wta);
int sgbm_compute16_batch_dev(mvs_ctx* ctx, const mvs_sgbm_params* p, int F, const uint8_t* const* l, const uint8_t* const* r, int w, int h,
                             int16_t* const* disp16, int16_t* raw16_tap /* F == 1 only */);
int sgbm_front_batch_dev(mvs_ctx* ctx, const mvs_sgbm_params* p, int F, const uint8_t* const* l, const uint8_t* const* r, int w, int h,
                         int sset, bool keep_S, SgbmWtaOut* wta);
int sgbm_tail_batch_dev(mvs_ctx* ctx, const mvs_sgbm_params* p, int F, int w, int h, const SgbmWtaOut& wta, int16_t* const* disp16,
                        int16_t* raw16_tap);
int sgbm_wta_buffers_dev(mvs_ctx* ctx, const SgbmGeom& g, int F, int sset, SgbmWtaOut* out);
int sgbm_lr_post_batch_dev(mvs_ctx* ctx, const SgbmGeom& g, const mvs_sgbm_params* p, int F, const SgbmWtaOut& wta, int16_t* const* disp16,
                           int16_t* raw16_tap);
int disp_to_float_batch_dev(mvs_ctx* ctx, int F, int16_t* const* d16, int w, int h, float* const* out);
