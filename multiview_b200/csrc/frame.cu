// Device-resident camera-frame path: rectify x2 -> grey/blur x2 -> SGBM -> float -> cloud -> floor histogram,
// i.e. the DAG slice calc_rectified[2c],[2c+1] -> disp_init[c] -> disp_map_update[c] (pass-through) ->
// floor_hist (per-camera part) of pipeline/frame-results.cpp:206-255, enqueued on one stream.
#include "common.cuh"

static std::string slot_name(int slot, const char* what)
{
    char buf[64];
    snprintf(buf, sizeof(buf), "slot%d.%s", slot, what);
    return buf;
}

extern "C" int mvs_frame_run(mvs_ctx* ctx, int slot, const mvs_frame_desc* d, mvs_frame_result* out)
{
    if (!ctx) return MVS_ERR_ARG;
    if (!d || slot < 0 || slot >= MVS_MAX_SLOTS || d->w <= 0 || d->h <= 0 || d->sw <= 0 || d->sh <= 0 ||
        !d->sensor[0] || !d->sensor[1] || !d->mapx[0] || !d->mapx[1] || !d->mapy[0] || !d->mapy[1])
        return mvs_fail(ctx, MVS_ERR_ARG, "mvs_frame_run: bad argument");
    MVS_CUDA(ctx, cudaSetDevice(ctx->device));
    cudaStream_t s = ctx->stream;
    const int w = d->w, h = d->h;
    const size_t n = (size_t)w * h;
    int hw = 0, hh = 0;
    if (mvs_hist_dims(d->aabb, d->hist.hist_sz, &hw, &hh)) return mvs_fail(ctx, MVS_ERR_ARG, "mvs_frame_run: bad histogram bounds");
    SgbmGeom g;
    int rc = sgbm_geom(ctx, &d->sgbm, w, h, &g);
    if (rc) return rc;

    const uint8_t* sensor[2];
    const float *mapx[2], *mapy[2];
    const size_t sstep = d->inputs_on_device ? d->sstep : (size_t)d->sw * 3;
    const size_t mstep = d->maps_on_device ? d->mstep : (size_t)w * 4;
    for (int i = 0; i < 2; ++i) {
        if (d->inputs_on_device) {
            sensor[i] = d->sensor[i];
        } else {
            uint8_t* ds = (uint8_t*)arena_get(ctx, slot_name(slot, i ? "sensor1" : "sensor0").c_str(), (size_t)d->sh * d->sw * 3);
            if (!ds) return MVS_ERR_NOMEM;
            MVS_CUDA(ctx, cudaMemcpy2DAsync(ds, (size_t)d->sw * 3, d->sensor[i], d->sstep, (size_t)d->sw * 3, d->sh, cudaMemcpyHostToDevice, s));
            sensor[i] = ds;
        }
        if (d->maps_on_device) {
            mapx[i] = d->mapx[i]; mapy[i] = d->mapy[i];
        } else {
            float* dx = (float*)arena_get(ctx, slot_name(slot, i ? "mapx1" : "mapx0").c_str(), n * 4);
            float* dy = (float*)arena_get(ctx, slot_name(slot, i ? "mapy1" : "mapy0").c_str(), n * 4);
            if (!dx || !dy) return MVS_ERR_NOMEM;
            MVS_CUDA(ctx, cudaMemcpy2DAsync(dx, (size_t)w * 4, d->mapx[i], d->mstep, (size_t)w * 4, h, cudaMemcpyHostToDevice, s));
            MVS_CUDA(ctx, cudaMemcpy2DAsync(dy, (size_t)w * 4, d->mapy[i], d->mstep, (size_t)w * 4, h, cudaMemcpyHostToDevice, s));
            mapx[i] = dx; mapy[i] = dy;
        }
    }

    uint8_t* rect[2];
    uint8_t* grey[2];
    for (int i = 0; i < 2; ++i) {
        rect[i] = (uint8_t*)arena_get(ctx, slot_name(slot, i ? "rect1" : "rect0").c_str(), n * 3);
        grey[i] = (uint8_t*)arena_get(ctx, slot_name(slot, i ? "grey1" : "grey0").c_str(), n);
        if (!rect[i] || !grey[i]) return MVS_ERR_NOMEM;
    }
    int16_t* disp16 = (int16_t*)arena_get(ctx, slot_name(slot, "disp16").c_str(), n * 2);
    float* dispf = (float*)arena_get(ctx, slot_name(slot, "disp").c_str(), n * 4);
    double* Xs = (double*)arena_get(ctx, slot_name(slot, "Xs").c_str(), n * 24);
    int32_t* xy = (int32_t*)arena_get(ctx, slot_name(slot, "xy").c_str(), n * 8);
    int32_t* lookup = (int32_t*)arena_get(ctx, slot_name(slot, "lookup").c_str(), n * 4);
    int32_t* nd = (int32_t*)arena_get(ctx, slot_name(slot, "n").c_str(), 4);
    double* cen = (double*)arena_get(ctx, slot_name(slot, "centroid").c_str(), 24);
    // per-camera histograms of all slots live in one array so that the camera sum is one kernel
    float* hists = (float*)arena_get(ctx, "slots.cam_hists", sizeof(float) * (size_t)hw * hh * MVS_MAX_SLOTS);
    if (!disp16 || !dispf || !Xs || !xy || !lookup || !nd || !cen || !hists) return MVS_ERR_NOMEM;
    float* cam_hist = hists + (size_t)slot * hw * hh;

    const uint8_t white[3] = {255, 255, 255};  // calc_rectified.cpp:113
    stage_begin(ctx, "rectify");
    for (int i = 0; i < 2 && !rc; ++i)
        rc = remap_bgr8_dev(ctx, sensor[i], d->sw, d->sh, sstep, mapx[i], mapy[i], mstep, w, h, MVS_INTER_CUBIC, white, rect[i], (size_t)w * 3);
    stage_end(ctx, "rectify");
    stage_begin(ctx, "presmooth");
    for (int i = 0; i < 2 && !rc; ++i)
        rc = presmooth_dev(ctx, i ? "r" : "l", rect[i], 3, w, h, (size_t)w * 3, d->presmooth_sigma, grey[i], (size_t)w);
    stage_end(ctx, "presmooth");
    if (!rc) rc = sgbm_compute16_dev(ctx, &d->sgbm, grey[0], grey[1], w, h, disp16, nullptr);
    stage_begin(ctx, "cloud");
    if (!rc) rc = disp_to_float_dev(ctx, disp16, w, h, dispf);
    if (!rc) rc = point_cloud_dev(ctx, &d->bcam, dispf, w, h, 1, Xs, xy, lookup, nd, cen);
    stage_end(ctx, "cloud");
    stage_begin(ctx, "hist");
    if (!rc) rc = floor_hist_camera_dev(ctx, &d->hist, d->et, d->aabb, d->cam_center, Xs, xy, nd, 0, nullptr, 0, 0, nullptr, 0, cam_hist, hw, hh);
    stage_end(ctx, "hist");
    if (rc) return rc;

    FrameSlot& fs = ctx->slots[slot];
    fs.w = w; fs.h = h; fs.hw = hw; fs.hh = hh; fs.valid = true;
    if (out) {
        out->rectified_dev[0] = rect[0]; out->rectified_dev[1] = rect[1];
        out->disp16_dev = disp16; out->disp_dev = dispf; out->Xs_dev = Xs; out->xy_dev = xy; out->lookup_dev = lookup;
        out->cam_hist_dev = cam_hist; out->n_dev = nd; out->centroid_dev = cen; out->hw = hw; out->hh = hh;
    }
    return MVS_OK;
}

extern "C" int mvs_frame_fetch(mvs_ctx* ctx, int slot, float* disp, size_t dstep, double* Xs, int32_t* xy,
                               int32_t* lookup, int32_t* n_out, double centroid[3], float* cam_hist)
{
    if (!ctx) return MVS_ERR_ARG;
    if (slot < 0 || slot >= MVS_MAX_SLOTS || !ctx->slots[slot].valid) return mvs_fail(ctx, MVS_ERR_ARG, "mvs_frame_fetch: empty slot");
    MVS_CUDA(ctx, cudaSetDevice(ctx->device));
    const FrameSlot& fs = ctx->slots[slot];
    cudaStream_t s = ctx->stream;
    size_t n = (size_t)fs.w * fs.h;
    int32_t nh = 0;
    MVS_CUDA(ctx, cudaMemcpyAsync(&nh, arena_get(ctx, slot_name(slot, "n").c_str(), 4), 4, cudaMemcpyDeviceToHost, s));
    if (disp) {
        if (dstep < (size_t)fs.w * 4) return mvs_fail(ctx, MVS_ERR_ARG, "mvs_frame_fetch: dstep too small");
        MVS_CUDA(ctx, cudaMemcpy2DAsync(disp, dstep, arena_get(ctx, slot_name(slot, "disp").c_str(), n * 4), (size_t)fs.w * 4, (size_t)fs.w * 4, fs.h, cudaMemcpyDeviceToHost, s));
    }
    if (lookup) MVS_CUDA(ctx, cudaMemcpyAsync(lookup, arena_get(ctx, slot_name(slot, "lookup").c_str(), n * 4), n * 4, cudaMemcpyDeviceToHost, s));
    if (centroid) MVS_CUDA(ctx, cudaMemcpyAsync(centroid, arena_get(ctx, slot_name(slot, "centroid").c_str(), 24), 24, cudaMemcpyDeviceToHost, s));
    if (cam_hist) {
        float* hists = (float*)arena_get(ctx, "slots.cam_hists", 0);
        MVS_CUDA(ctx, cudaMemcpyAsync(cam_hist, hists + (size_t)slot * fs.hw * fs.hh, sizeof(float) * (size_t)fs.hw * fs.hh, cudaMemcpyDeviceToHost, s));
    }
    MVS_CUDA(ctx, cudaStreamSynchronize(s));
    if (nh > 0 && Xs) MVS_CUDA(ctx, cudaMemcpyAsync(Xs, arena_get(ctx, slot_name(slot, "Xs").c_str(), n * 24), (size_t)nh * 24, cudaMemcpyDeviceToHost, s));
    if (nh > 0 && xy) MVS_CUDA(ctx, cudaMemcpyAsync(xy, arena_get(ctx, slot_name(slot, "xy").c_str(), n * 8), (size_t)nh * 8, cudaMemcpyDeviceToHost, s));
    MVS_CUDA(ctx, cudaStreamSynchronize(s));
    if (n_out) *n_out = nh;
    return sgbm_aggregate_wave_status(ctx);
}

extern "C" int mvs_frame_hist_sum(mvs_ctx* ctx, int n_slots, float** hist_dev, int* hw, int* hh)
{
    if (!ctx) return MVS_ERR_ARG;
    if (n_slots <= 0 || n_slots > MVS_MAX_SLOTS || !hist_dev) return mvs_fail(ctx, MVS_ERR_ARG, "mvs_frame_hist_sum: bad argument");
    for (int i = 0; i < n_slots; ++i)
        if (!ctx->slots[i].valid || ctx->slots[i].hw != ctx->slots[0].hw || ctx->slots[i].hh != ctx->slots[0].hh)
            return mvs_fail(ctx, MVS_ERR_ARG, "mvs_frame_hist_sum: slot %d empty or of different size", i);
    MVS_CUDA(ctx, cudaSetDevice(ctx->device));
    int w = ctx->slots[0].hw, h = ctx->slots[0].hh;
    float* hists = (float*)arena_get(ctx, "slots.cam_hists", 0);
    MVS_ARENA(ctx, float, sum, "slots.hist_sum", (size_t)w * h);
    int rc = hist_sum_dev(ctx, hists, n_slots, w, h, sum);
    if (rc) return rc;
    *hist_dev = sum;
    if (hw) *hw = w;
    if (hh) *hh = h;
    return MVS_OK;
}
