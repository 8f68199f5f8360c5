// SGBM chain glue: cost volume -> aggregation -> WTA/LR -> median/speckle, and the host-buffer entry points
// that replace cv::StereoSGBM::compute / Disparity::calculate (strategies/stereosgbm_disp.cpp:84-156,
// cost-functions/disparity/disparity.cpp:68-182).
#include "common.cuh"

// unpadded copy [h][w1][D] (d fastest) <- device volume [h][w1][Dp] in the split layout (common.cuh)
__global__ void unpad_kernel(const int16_t* __restrict__ src, size_t npix, int D, int Dp, int16_t* __restrict__ dst)
{
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= npix * (size_t)D) return;
    size_t p = i / D;
    int d = (int)(i - p * D);
    dst[i] = src[p * Dp + split_index(d, Dp)];
}

int16_t* sgbm_volume(mvs_ctx* ctx, const char* name, const SgbmGeom& g)
{
    char* p = (char*)arena_get(ctx, name, g.cells() * 2 + 2 * MVS_VOLUME_GUARD);
    return p ? (int16_t*)(p + MVS_VOLUME_GUARD) : nullptr;
}

static std::string fname(const char* base, int f)
{
    char buf[64];
    snprintf(buf, sizeof(buf), "%s.%d", base, f);
    return buf;
}

// A batch of F pairs of equal size through the SGBM chain, in two halves so that the frame path can run the second
// half of one batch on another stream while the first half of the next batch occupies the main stream:
//   front: cost volumes of all frames, then ONE aggregation launch per sweep for the whole batch (its wavefronts fill
//          the GPU); the backward sweep does the winner-take-all step and writes raw disparities + right-view keys into
//          the buffers of set `sset` (0 or 1).  S is final only when keep_S is set (the test tap);
//   tail:  left-right check + median + speckle filter from those buffers.
int sgbm_front_batch_dev(mvs_ctx* ctx, const mvs_sgbm_params* p, int F, const uint8_t* const* l, const uint8_t* const* r, int w, int h,
                         int sset, bool keep_S, SgbmWtaOut* wta)
{
    if (F < 1 || F > MVS_MAX_BATCH) return mvs_fail(ctx, MVS_ERR_ARG, "sgbm: batch of %d frames (1..%d)", F, MVS_MAX_BATCH);
    SgbmGeom g;
    int rc = sgbm_geom(ctx, p, w, h, &g);
    if (rc) return rc;
    int16_t *C[MVS_MAX_BATCH], *S[MVS_MAX_BATCH];
    const bool in_sweep = sgbm_wta_in_sweep(g);
    for (int f = 0; f < F; ++f) {
        C[f] = sgbm_volume(ctx, fname("sgbm.C", f).c_str(), g);
        // S is read by the tail only when the winner-take-all step is not in the sweep: then it alternates with the set
        S[f] = sgbm_volume(ctx, fname(sset && !in_sweep ? "sgbm.S1" : "sgbm.S", f).c_str(), g);
        if (!C[f] || !S[f]) return MVS_ERR_NOMEM;
    }
    stage_begin(ctx, "cost");
    rc = sgbm_cost_volume_batch_dev(ctx, g, F, l, r, C);
    stage_end(ctx, "cost");
    if (rc) return rc;
    if (in_sweep) {
        rc = sgbm_wta_buffers_dev(ctx, g, F, sset, wta);
        if (rc) return rc;
    } else
        memset(wta, 0, sizeof(*wta));
    for (int f = 0; f < F; ++f) wta->S[f] = S[f];
    wta->in_sweep = in_sweep;
    wta->sset = sset;
    wta->store_S = keep_S;
    stage_begin(ctx, "aggregate");
    rc = sgbm_aggregate_batch_dev(ctx, g, F, C, S, in_sweep ? wta : nullptr);
    stage_end(ctx, "aggregate");
    return rc;
}

int sgbm_tail_batch_dev(mvs_ctx* ctx, const mvs_sgbm_params* p, int F, int w, int h, const SgbmWtaOut& wta, int16_t* const* disp16,
                        int16_t* raw16_tap)
{
    SgbmGeom g;
    int rc = sgbm_geom(ctx, p, w, h, &g);
    if (rc) return rc;
    // "wta": the winner-take-all step when it is not in the sweep + the left-right check; F == 1 times the rest separately
    // as "post", a batch times everything under "wta"
    stage_begin(ctx, "wta");
    SgbmWtaOut wo = wta;
    if (!wo.in_sweep) {
        rc = sgbm_wta_buffers_dev(ctx, g, F, wta.sset, &wo);
        for (int f = 0; f < F; ++f) wo.S[f] = wta.S[f];
        if (!rc) rc = sgbm_wta_from_s_batch_dev(ctx, g, F, wo);
        if (rc) return rc;
    }
    rc = sgbm_lr_post_batch_dev(ctx, g, p, F, wo, disp16, F == 1 ? raw16_tap : nullptr);
    if (F > 1) stage_end(ctx, "wta");
    return rc;
}

int sgbm_compute16_batch_dev(mvs_ctx* ctx, const mvs_sgbm_params* p, int F, const uint8_t* const* l, const uint8_t* const* r,
                             int w, int h, int16_t* const* disp16, int16_t* raw16_tap)
{
    SgbmWtaOut wta;
    int rc = main_stream_waits_for_tail(ctx);
    if (rc) return rc;
    rc = sgbm_front_batch_dev(ctx, p, F, l, r, w, h, 0, /* keep_S: the tap reads it */ raw16_tap != nullptr, &wta);
    if (rc) return rc;
    return sgbm_tail_batch_dev(ctx, p, F, w, h, wta, disp16, F == 1 ? raw16_tap : nullptr);
}

int sgbm_compute16_dev(mvs_ctx* ctx, const mvs_sgbm_params* p, const uint8_t* l, const uint8_t* r, int w, int h,
                       int16_t* disp16, int16_t* raw16_tap)
{
    return sgbm_compute16_batch_dev(ctx, p, 1, &l, &r, w, h, &disp16, raw16_tap);
}

static int upload_pair(mvs_ctx* ctx, const uint8_t* l, const uint8_t* r, int channels, int w, int h, size_t step,
                       uint8_t** dl, uint8_t** dr)
{
    size_t rowb = (size_t)w * channels;
    *dl = (uint8_t*)arena_get(ctx, "sgbm.api.l", rowb * h);
    *dr = (uint8_t*)arena_get(ctx, "sgbm.api.r", rowb * h);
    if (!*dl || !*dr) return MVS_ERR_NOMEM;
    MVS_CUDA(ctx, cudaMemcpy2DAsync(*dl, rowb, l, step, rowb, h, cudaMemcpyHostToDevice, ctx->stream));
    MVS_CUDA(ctx, cudaMemcpy2DAsync(*dr, rowb, r, step, rowb, h, cudaMemcpyHostToDevice, ctx->stream));
    return MVS_OK;
}

extern "C" int mvs_sgbm_compute16(mvs_ctx* ctx, const mvs_sgbm_params* p, const uint8_t* l, const uint8_t* r, int w,
                                  int h, size_t step, int16_t* disp16, size_t dstep)
{
    if (!ctx) return MVS_ERR_ARG;
    if (!p || !l || !r || !disp16 || w <= 0 || h <= 0 || step < (size_t)w || dstep < (size_t)w * 2)
        return mvs_fail(ctx, MVS_ERR_ARG, "mvs_sgbm_compute16: bad argument");
    MVS_CUDA(ctx, cudaSetDevice(ctx->device));
    uint8_t *dl, *dr;
    int rc = upload_pair(ctx, l, r, 1, w, h, step, &dl, &dr);
    if (rc) return rc;
    MVS_ARENA(ctx, int16_t, dd, "sgbm.api.disp16", (size_t)w * h);
    rc = sgbm_compute16_dev(ctx, p, dl, dr, w, h, dd, nullptr);
    if (rc) return rc;
    MVS_CUDA(ctx, cudaMemcpy2DAsync(disp16, dstep, dd, (size_t)w * 2, (size_t)w * 2, h, cudaMemcpyDeviceToHost, ctx->stream));
    MVS_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return sgbm_aggregate_wave_status(ctx);
}

extern "C" int mvs_disparity_sgbm(mvs_ctx* ctx, const mvs_sgbm_params* p, double presmooth_sigma, const uint8_t* l,
                                  const uint8_t* r, int channels, int w, int h, size_t step, float* out, size_t ostep)
{
    if (!ctx) return MVS_ERR_ARG;
    if (!p || !l || !r || !out || w <= 0 || h <= 0 || (channels != 1 && channels != 3) ||
        step < (size_t)w * channels || ostep < (size_t)w * 4)
        return mvs_fail(ctx, MVS_ERR_ARG, "mvs_disparity_sgbm: bad argument");
    MVS_CUDA(ctx, cudaSetDevice(ctx->device));
    uint8_t *dl, *dr;
    int rc = upload_pair(ctx, l, r, channels, w, h, step, &dl, &dr);
    if (rc) return rc;
    MVS_ARENA(ctx, uint8_t, g0, "sgbm.api.grey0", (size_t)w * h);
    MVS_ARENA(ctx, uint8_t, g1, "sgbm.api.grey1", (size_t)w * h);
    MVS_ARENA(ctx, int16_t, dd, "sgbm.api.disp16", (size_t)w * h);
    MVS_ARENA(ctx, float, df, "sgbm.api.dispf", (size_t)w * h);
    rc = presmooth_dev(ctx, "l", dl, channels, w, h, (size_t)w * channels, presmooth_sigma, g0, (size_t)w);
    if (!rc) rc = presmooth_dev(ctx, "r", dr, channels, w, h, (size_t)w * channels, presmooth_sigma, g1, (size_t)w);
    if (!rc) rc = sgbm_compute16_dev(ctx, p, g0, g1, w, h, dd, nullptr);
    if (!rc) rc = disp_to_float_dev(ctx, dd, w, h, df);
    if (rc) return rc;
    MVS_CUDA(ctx, cudaMemcpy2DAsync(out, ostep, df, (size_t)w * 4, (size_t)w * 4, h, cudaMemcpyDeviceToHost, ctx->stream));
    MVS_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return sgbm_aggregate_wave_status(ctx);
}

extern "C" int mvs_sgbm_volumes(mvs_ctx* ctx, const mvs_sgbm_params* p, const uint8_t* l, const uint8_t* r, int w, int h,
                                size_t step, int16_t* C, int16_t* S, int16_t* raw_disp16)
{
    if (!ctx) return MVS_ERR_ARG;
    if (!p || !l || !r || w <= 0 || h <= 0 || step < (size_t)w) return mvs_fail(ctx, MVS_ERR_ARG, "mvs_sgbm_volumes: bad argument");
    MVS_CUDA(ctx, cudaSetDevice(ctx->device));
    SgbmGeom g;
    int rc = sgbm_geom(ctx, p, w, h, &g);
    if (rc) return rc;
    uint8_t *dl, *dr;
    rc = upload_pair(ctx, l, r, 1, w, h, step, &dl, &dr);
    if (rc) return rc;
    MVS_ARENA(ctx, int16_t, dd, "sgbm.api.disp16", (size_t)w * h);
    MVS_ARENA(ctx, int16_t, draw, "sgbm.api.raw16", (size_t)w * h);
    rc = sgbm_compute16_dev(ctx, p, dl, dr, w, h, dd, draw);
    if (rc) return rc;
    size_t npix = (size_t)g.H * g.W1;
    MVS_ARENA(ctx, int16_t, tmp, "sgbm.api.unpad", npix * g.D);
    const char* names[2] = {"sgbm.C.0", "sgbm.S.0"};
    int16_t* outs[2] = {C, S};
    for (int k = 0; k < 2; ++k) {
        if (!outs[k]) continue;
        const int16_t* src = sgbm_volume(ctx, names[k], g);
        unpad_kernel<<<cdiv(npix * g.D, 256), 256, 0, ctx->stream>>>(src, npix, g.D, g.Dp, tmp);
        MVS_LAUNCH_CHECK(ctx, "unpad_kernel");
        MVS_CUDA(ctx, cudaMemcpyAsync(outs[k], tmp, npix * g.D * 2, cudaMemcpyDeviceToHost, ctx->stream));
        MVS_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    }
    if (raw_disp16) MVS_CUDA(ctx, cudaMemcpyAsync(raw_disp16, draw, (size_t)w * h * 2, cudaMemcpyDeviceToHost, ctx->stream));
    MVS_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return sgbm_aggregate_wave_status(ctx);
}
