// Device-resident camera-frame path: rectify x2 -> grey/blur x2 -> SGBM -> float -> cloud -> floor histogram,
// i.e. the DAG slice calc_rectified[2c],[2c+1] -> disp_init[c] -> disp_map_update[c] (pass-through) ->
// floor_hist (per-camera part) of pipeline/frame-results.cpp:206-255, enqueued on one stream.
//
// mvs_frame_run_batch takes the camera-frames of one step together (the reference's thread pool runs the nodes of
// different cameras concurrently, pipeline/pipeline-task.cpp:181-200): the stages run frame after frame, except
// the SGM aggregation, which sweeps all frames of the batch in one launch so that its wavefronts fill the GPU.
#include "common.cuh"

static std::string slot_name(int slot, const char* what)
{
    char buf[64];
    snprintf(buf, sizeof(buf), "slot%d.%s", slot, what);
    return buf;
}

struct SlotBufs {
    uint8_t *rect[2], *grey[2];
    int16_t* disp16;
    float* dispf;
    double* Xs;
    int32_t *xy, *lookup, *nd;
    double* cen;
    float* cam_hist;
    const int32_t* labels;   // device, packed w x h (nullptr: no still weights)
    const double* still;
    mvs_sparse_hist_cell* cells;   // slic_to_hist_xy (nullptr without labels)
    int n_labels;
};

extern "C" int mvs_frame_run_batch(mvs_ctx* ctx, int first_slot, const mvs_frame_desc* descs, int n_frames,
                                   mvs_frame_result* out)
{
    if (!ctx) return MVS_ERR_ARG;
    if (!descs || n_frames < 1 || n_frames > MVS_MAX_BATCH || first_slot < 0 || first_slot + n_frames > MVS_MAX_SLOTS)
        return mvs_fail(ctx, MVS_ERR_ARG, "mvs_frame_run_batch: bad argument (1..%d frames, slots < %d)", MVS_MAX_BATCH, MVS_MAX_SLOTS);
    const mvs_frame_desc& d0 = descs[0];
    for (int f = 0; f < n_frames; ++f) {
        const mvs_frame_desc* d = descs + f;
        if (d->w <= 0 || d->h <= 0 || d->sw <= 0 || d->sh <= 0 || !d->sensor[0] || !d->sensor[1])
            return mvs_fail(ctx, MVS_ERR_ARG, "mvs_frame_run_batch: frame %d: bad argument", f);
        for (int i = 0; i < 2; ++i) {
            const int mh = d->maps_handle[i];
            if (mh > 0) {   // registered maps: cubic only, same output size
                if (mh > (int)ctx->maps.size() || !ctx->maps[mh - 1].packed || ctx->maps[mh - 1].w != d->w || ctx->maps[mh - 1].h != d->h ||
                    d->interp != MVS_INTER_CUBIC)
                    return mvs_fail(ctx, MVS_ERR_ARG, "mvs_frame_run_batch: frame %d: maps_handle[%d] unknown, of another size, or interp is not cubic", f, i);
            } else if (!d->mapx[i] || !d->mapy[i])
                return mvs_fail(ctx, MVS_ERR_ARG, "mvs_frame_run_batch: frame %d: no maps for sensor %d", f, i);
        }
        if (d->w != d0.w || d->h != d0.h || memcmp(&d->sgbm, &d0.sgbm, sizeof(mvs_sgbm_params)) || d->hist.hist_sz != d0.hist.hist_sz ||
            memcmp(d->aabb, d0.aabb, sizeof(d0.aabb)))
            return mvs_fail(ctx, MVS_ERR_ARG, "mvs_frame_run_batch: frame %d differs in size, SGBM parameters or histogram bounds", f);
    }
    MVS_CUDA(ctx, cudaSetDevice(ctx->device));
    cudaStream_t s = ctx->stream;
    const int w = d0.w, h = d0.h;
    const size_t n = (size_t)w * h;
    int hw = 0, hh = 0;
    if (mvs_hist_dims(d0.aabb, d0.hist.hist_sz, &hw, &hh)) return mvs_fail(ctx, MVS_ERR_ARG, "mvs_frame_run_batch: bad histogram bounds");
    SgbmGeom g;
    int rc = sgbm_geom(ctx, &d0.sgbm, w, h, &g);
    if (rc) return rc;
    // With the winner-take-all step inside the backward sweep (from 192 disparities) the second half of a batch is 8 % of it, and
    // running it beside the next batch's cost kernels measured slower than running it alone (c2, batch 16: 31.5 against
    // 30.5 ms per step): the batch then starts behind the previous batch's second half.  Narrower disparity ranges keep
    // the overlap (their second half includes the winner-take-all kernel; 896x672x128: 8.5 against 8.8 ms per step).
    if (sgbm_wta_in_sweep(g) && ctx->tail_done_recorded[ctx->s_set ^ 1]) MVS_CUDA(ctx, cudaStreamWaitEvent(s, ctx->ev_tail_done[ctx->s_set ^ 1], 0));
    // per-camera histograms of all slots live in one array so that the camera sum is one kernel
    // (slots filled with another histogram size index this array differently: they are no longer fetchable)
    for (int i = 0; i < MVS_MAX_SLOTS; ++i)
        if (ctx->slots[i].valid && (ctx->slots[i].hw != hw || ctx->slots[i].hh != hh)) ctx->slots[i].valid = false;
    float* hists = (float*)arena_get(ctx, "slots.cam_hists", sizeof(float) * (size_t)hw * hh * MVS_MAX_SLOTS);
    if (!hists) return MVS_ERR_NOMEM;

    SlotBufs sb[MVS_MAX_BATCH];
    const uint8_t* sensor[MVS_MAX_BATCH][2];
    const float *mapx[MVS_MAX_BATCH][2], *mapy[MVS_MAX_BATCH][2];
    cudaStream_t cs = ctx->copy_stream;  // uploads go here; the compute stream waits for them once, below
    bool any_upload = false;
    for (int f = 0; f < n_frames; ++f) {
        const mvs_frame_desc* d = descs + f;
        const int slot = first_slot + f;
        // the label image feeds the still weights (when enabled) and, always, the super-pixel table
        const bool use_labels = d->labels && d->n_labels > 0 && (!d->hist.use_still || d->still_score);
        if (use_labels && d->lstep < (size_t)d->w * 4) return mvs_fail(ctx, MVS_ERR_ARG, "mvs_frame_run_batch: frame %d: lstep too small", f);
        const bool maps_up = !d->maps_on_device && (d->maps_handle[0] <= 0 || d->maps_handle[1] <= 0);
        if ((!d->inputs_on_device || maps_up || (use_labels && !d->labels_on_device))) {
            any_upload = true;
            // the slot's input buffers may still be read by the kernels of the previous frame that used this slot
            if (ctx->slot_done_recorded[slot]) MVS_CUDA(ctx, cudaStreamWaitEvent(cs, ctx->ev_slot_done[slot], 0));
        }
        for (int i = 0; i < 2; ++i) {
            if (d->inputs_on_device) {
                sensor[f][i] = d->sensor[i];
            } else {
                uint8_t* ds = (uint8_t*)arena_get(ctx, slot_name(slot, i ? "sensor1" : "sensor0").c_str(), (size_t)d->sh * d->sw * 3);
                if (!ds) return MVS_ERR_NOMEM;
                MVS_CUDA(ctx, cudaMemcpy2DAsync(ds, (size_t)d->sw * 3, d->sensor[i], d->sstep, (size_t)d->sw * 3, d->sh, cudaMemcpyHostToDevice, cs));
                sensor[f][i] = ds;
            }
            if (d->maps_handle[i] > 0) {   // packed form: the remap kernel takes it through mapx with mapy == nullptr
                mapx[f][i] = (const float*)ctx->maps[d->maps_handle[i] - 1].packed; mapy[f][i] = nullptr;
            } else if (d->maps_on_device) {
                mapx[f][i] = d->mapx[i]; mapy[f][i] = d->mapy[i];
            } else {
                float* dx = (float*)arena_get(ctx, slot_name(slot, i ? "mapx1" : "mapx0").c_str(), n * 4);
                float* dy = (float*)arena_get(ctx, slot_name(slot, i ? "mapy1" : "mapy0").c_str(), n * 4);
                if (!dx || !dy) return MVS_ERR_NOMEM;
                MVS_CUDA(ctx, cudaMemcpy2DAsync(dx, (size_t)w * 4, d->mapx[i], d->mstep, (size_t)w * 4, h, cudaMemcpyHostToDevice, cs));
                MVS_CUDA(ctx, cudaMemcpy2DAsync(dy, (size_t)w * 4, d->mapy[i], d->mstep, (size_t)w * 4, h, cudaMemcpyHostToDevice, cs));
                mapx[f][i] = dx; mapy[f][i] = dy;
            }
            sb[f].rect[i] = (uint8_t*)arena_get(ctx, slot_name(slot, i ? "rect1" : "rect0").c_str(), n * 3);
            sb[f].grey[i] = (uint8_t*)arena_get(ctx, slot_name(slot, i ? "grey1" : "grey0").c_str(), n);
            if (!sb[f].rect[i] || !sb[f].grey[i]) return MVS_ERR_NOMEM;
        }
        sb[f].disp16 = (int16_t*)arena_get(ctx, slot_name(slot, "disp16").c_str(), n * 2);
        sb[f].dispf = (float*)arena_get(ctx, slot_name(slot, "disp").c_str(), n * 4);
        sb[f].Xs = (double*)arena_get(ctx, slot_name(slot, "Xs").c_str(), n * 24);
        sb[f].xy = (int32_t*)arena_get(ctx, slot_name(slot, "xy").c_str(), n * 8);
        sb[f].lookup = (int32_t*)arena_get(ctx, slot_name(slot, "lookup").c_str(), n * 4);
        sb[f].nd = (int32_t*)arena_get(ctx, slot_name(slot, "n").c_str(), 4);
        sb[f].cen = (double*)arena_get(ctx, slot_name(slot, "centroid").c_str(), 24);
        if (!sb[f].disp16 || !sb[f].dispf || !sb[f].Xs || !sb[f].xy || !sb[f].lookup || !sb[f].nd || !sb[f].cen) return MVS_ERR_NOMEM;
        sb[f].cam_hist = hists + (size_t)slot * hw * hh;
        sb[f].labels = nullptr; sb[f].still = nullptr; sb[f].cells = nullptr; sb[f].n_labels = 0;
        if (use_labels) {
            sb[f].n_labels = d->n_labels;
            sb[f].cells = (mvs_sparse_hist_cell*)arena_get(ctx, slot_name(slot, "cells").c_str(), sizeof(mvs_sparse_hist_cell) * (size_t)d->n_labels);
            if (!sb[f].cells) return MVS_ERR_NOMEM;
            if (d->labels_on_device && d->lstep == (size_t)w * 4) {
                sb[f].labels = d->labels; sb[f].still = d->still_score;
            } else {
                int32_t* dl = (int32_t*)arena_get(ctx, slot_name(slot, "labels").c_str(), n * 4);
                double* dsc = (double*)arena_get(ctx, slot_name(slot, "still").c_str(), (size_t)d->n_labels * 8);
                if (!dl || !dsc) return MVS_ERR_NOMEM;
                const cudaMemcpyKind kind = d->labels_on_device ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice;
                MVS_CUDA(ctx, cudaMemcpy2DAsync(dl, (size_t)w * 4, d->labels, d->lstep, (size_t)w * 4, h, kind, cs));
                if (d->still_score) MVS_CUDA(ctx, cudaMemcpyAsync(dsc, d->still_score, (size_t)d->n_labels * 8, kind, cs));
                sb[f].labels = dl; sb[f].still = d->still_score ? dsc : nullptr;
                any_upload = true;
            }
        }
    }

    if (any_upload) {
        MVS_CUDA(ctx, cudaEventRecord(ctx->ev_h2d, cs));
        MVS_CUDA(ctx, cudaStreamWaitEvent(s, ctx->ev_h2d, 0));
    }
    const uint8_t white[3] = {255, 255, 255};  // calc_rectified.cpp:113
    stage_begin(ctx, "rectify");
    {
        auto sstep_of = [&](const mvs_frame_desc* d) { return d->inputs_on_device ? d->sstep : (size_t)d->sw * 3; };
        auto mstep_of = [&](const mvs_frame_desc* d) { return d->maps_on_device ? d->mstep : (size_t)w * 4; };
        bool same_geometry = true;  // then all sensors of the batch go through one launch
        for (int f = 1; f < n_frames; ++f)
            same_geometry = same_geometry && descs[f].sw == d0.sw && descs[f].sh == d0.sh && sstep_of(descs + f) == sstep_of(&d0) &&
                            mstep_of(descs + f) == mstep_of(&d0) && descs[f].interp == d0.interp;
        if (same_geometry && d0.interp == MVS_INTER_CUBIC) {
            const uint8_t* ps[2 * MVS_MAX_BATCH];
            const float *px[2 * MVS_MAX_BATCH], *py[2 * MVS_MAX_BATCH];
            uint8_t* pd[2 * MVS_MAX_BATCH];
            for (int f = 0; f < n_frames; ++f)
                for (int i = 0; i < 2; ++i) {
                    ps[2 * f + i] = sensor[f][i]; px[2 * f + i] = mapx[f][i]; py[2 * f + i] = mapy[f][i]; pd[2 * f + i] = sb[f].rect[i];
                }
            rc = remap_cubic_bgr8_batch_dev(ctx, 2 * n_frames, ps, d0.sw, d0.sh, sstep_of(&d0), px, py, mstep_of(&d0), w, h, white, pd,
                                            (size_t)w * 3);
        } else {
            for (int f = 0; f < n_frames && !rc; ++f) {
                const mvs_frame_desc* d = descs + f;
                for (int i = 0; i < 2 && !rc; ++i)
                    rc = remap_bgr8_dev(ctx, sensor[f][i], d->sw, d->sh, sstep_of(d), mapx[f][i], mapy[f][i], mstep_of(d), w, h, d->interp,
                                        white, sb[f].rect[i], (size_t)w * 3);
            }
        }
    }
    stage_end(ctx, "rectify");
    stage_begin(ctx, "presmooth");
    {
        // all sensors of the batch in one launch per pass when they share the blur strength
        bool same_sigma = true;
        for (int f = 1; f < n_frames; ++f) same_sigma = same_sigma && descs[f].presmooth_sigma == d0.presmooth_sigma;
        const uint8_t* ps[2 * MVS_MAX_BATCH];
        uint8_t* pd[2 * MVS_MAX_BATCH];
        for (int f = 0; f < n_frames; ++f)
            for (int i = 0; i < 2; ++i) { ps[2 * f + i] = sb[f].rect[i]; pd[2 * f + i] = sb[f].grey[i]; }
        if (same_sigma && !rc)
            rc = presmooth_batch_dev(ctx, 2 * n_frames, ps, 3, w, h, (size_t)w * 3, d0.presmooth_sigma, pd, (size_t)w);
        else
            for (int f = 0; f < n_frames && !rc; ++f)
                rc = presmooth_batch_dev(ctx, 2, ps + 2 * f, 3, w, h, (size_t)w * 3, descs[f].presmooth_sigma, pd + 2 * f, (size_t)w);
    }
    stage_end(ctx, "presmooth");
    SgbmWtaOut wta;
    int sset = 0;
    if (rc) return rc;
    {
        const uint8_t *l[MVS_MAX_BATCH], *r[MVS_MAX_BATCH];
        for (int f = 0; f < n_frames; ++f) { l[f] = sb[f].grey[0]; r[f] = sb[f].grey[1]; }
        // first half on the main stream: cost volumes + aggregation into buffer set `sset` (S, or the outcome of the
        // winner-take-all step when it runs inside the sweep), whose previous reader must be done
        sset = ctx->s_set;
        ctx->s_set ^= 1;
        if (ctx->tail_done_recorded[sset]) MVS_CUDA(ctx, cudaStreamWaitEvent(s, ctx->ev_tail_done[sset], 0));
        rc = sgbm_front_batch_dev(ctx, &d0.sgbm, n_frames, l, r, w, h, sset, false, &wta);
        if (rc) return rc;
        MVS_CUDA(ctx, cudaEventRecord(ctx->ev_front_done, s));
    }
    // second half on the tail stream: it overlaps the sweeps of the next batch
    cudaStream_t ts = ctx->tail_stream;
    StreamSwap on_tail(ctx, ts);
    MVS_CUDA(ctx, cudaStreamWaitEvent(ts, ctx->ev_front_done, 0));
    // results of these slots may still be on their way to the host (mvs_frame_fetch copies on the copy stream)
    for (int f = 0; f < n_frames; ++f)
        if (ctx->slot_fetch_recorded[first_slot + f]) MVS_CUDA(ctx, cudaStreamWaitEvent(ts, ctx->ev_slot_fetched[first_slot + f], 0));
    if (!rc) {
        int16_t* d16[MVS_MAX_BATCH];
        for (int f = 0; f < n_frames; ++f) d16[f] = sb[f].disp16;
        rc = sgbm_tail_batch_dev(ctx, &d0.sgbm, n_frames, w, h, wta, d16, nullptr);
    }
    stage_begin(ctx, "cloud");
    if (!rc) {
        int16_t* d16[MVS_MAX_BATCH];
        float* df[MVS_MAX_BATCH];
        for (int f = 0; f < n_frames; ++f) { d16[f] = sb[f].disp16; df[f] = sb[f].dispf; }
        rc = disp_to_float_batch_dev(ctx, n_frames, d16, w, h, df);
    }
    if (!rc) {
        const mvs_bcam* pb[MVS_MAX_BATCH];
        const float* pdisp[MVS_MAX_BATCH];
        double *pX[MVS_MAX_BATCH], *pc[MVS_MAX_BATCH];
        int32_t *pxy[MVS_MAX_BATCH], *plk[MVS_MAX_BATCH], *pn[MVS_MAX_BATCH];
        for (int f = 0; f < n_frames; ++f) {
            pb[f] = &descs[f].bcam; pdisp[f] = sb[f].dispf; pX[f] = sb[f].Xs; pxy[f] = sb[f].xy; plk[f] = sb[f].lookup;
            pn[f] = sb[f].nd; pc[f] = sb[f].cen;
        }
        bool same_ref = true;
        for (int f = 1; f < n_frames; ++f) same_ref = same_ref && (descs[f].left_is_ref != 0) == (d0.left_is_ref != 0);
        if (same_ref)
            rc = point_cloud_batch_dev(ctx, n_frames, pb, pdisp, w, h, d0.left_is_ref != 0, pX, pxy, plk, pn, pc);
        else
            for (int f = 0; f < n_frames && !rc; ++f)
                rc = point_cloud_batch_dev(ctx, 1, pb + f, pdisp + f, w, h, descs[f].left_is_ref != 0, pX + f, pxy + f, plk + f, pn + f, pc + f);
    }
    stage_end(ctx, "cloud");
    stage_begin(ctx, "hist");
    for (int f = 0; f < n_frames && !rc; ++f)
        rc = floor_hist_camera_dev(ctx, &descs[f].hist, descs[f].et, descs[f].aabb, descs[f].cam_center, sb[f].Xs, sb[f].xy, sb[f].nd, (int)n /* capacity */,
                                   sb[f].labels, w, h, sb[f].still, descs[f].n_labels, sb[f].cam_hist, hw, hh, sb[f].cells);
    stage_end(ctx, "hist");
    if (rc) return rc;

    MVS_CUDA(ctx, cudaEventRecord(ctx->ev_tail_done[sset], ts));
    ctx->tail_done_recorded[sset] = true;
    for (int f = 0; f < n_frames; ++f) {
        MVS_CUDA(ctx, cudaEventRecord(ctx->ev_slot_done[first_slot + f], ts));
        ctx->slot_done_recorded[first_slot + f] = true;
        FrameSlot& fs = ctx->slots[first_slot + f];
        fs.w = w; fs.h = h; fs.hw = hw; fs.hh = hh; fs.n_labels = sb[f].n_labels; fs.valid = true;
        if (out) {
            mvs_frame_result* o = out + f;
            o->rectified_dev[0] = sb[f].rect[0]; o->rectified_dev[1] = sb[f].rect[1];
            o->disp16_dev = sb[f].disp16; o->disp_dev = sb[f].dispf; o->Xs_dev = sb[f].Xs; o->xy_dev = sb[f].xy;
            o->lookup_dev = sb[f].lookup; o->cam_hist_dev = sb[f].cam_hist; o->n_dev = sb[f].nd; o->centroid_dev = sb[f].cen;
            o->hw = hw; o->hh = hh;
            o->slic_to_hist_xy_dev = sb[f].cells; o->n_labels = sb[f].n_labels;
        }
    }
    return MVS_OK;
}

extern "C" int mvs_frame_run(mvs_ctx* ctx, int slot, const mvs_frame_desc* d, mvs_frame_result* out)
{
    return mvs_frame_run_batch(ctx, slot, d, 1, out);
}

extern "C" int mvs_frame_fetch(mvs_ctx* ctx, int slot, float* disp, size_t dstep, double* Xs, int32_t* xy,
                               int32_t* lookup, int32_t* n_out, double centroid[3], float* cam_hist)
{
    if (!ctx) return MVS_ERR_ARG;
    if (slot < 0 || slot >= MVS_MAX_SLOTS || !ctx->slots[slot].valid) return mvs_fail(ctx, MVS_ERR_ARG, "mvs_frame_fetch: empty slot");
    MVS_CUDA(ctx, cudaSetDevice(ctx->device));
    const FrameSlot& fs = ctx->slots[slot];
    // download on the copy stream, behind the slot's last kernel: the compute stream is free to run the next batch
    cudaStream_t s = ctx->copy_stream;
    if (ctx->slot_done_recorded[slot]) MVS_CUDA(ctx, cudaStreamWaitEvent(s, ctx->ev_slot_done[slot], 0));
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
    MVS_CUDA(ctx, cudaEventRecord(ctx->ev_slot_fetched[slot], s));
    ctx->slot_fetch_recorded[slot] = true;
    MVS_CUDA(ctx, cudaStreamSynchronize(s));
    if (n_out) *n_out = nh;
    return sgbm_aggregate_wave_status(ctx, s);  // read behind the slot's kernels, without waiting for later batches
}

extern "C" int mvs_frame_fetch_batch(mvs_ctx* ctx, int first_slot, int n, const mvs_fetch_desc* d)
{
    if (!ctx) return MVS_ERR_ARG;
    if (!d || n < 1 || first_slot < 0 || first_slot + n > MVS_MAX_SLOTS) return mvs_fail(ctx, MVS_ERR_ARG, "mvs_frame_fetch_batch: bad argument");
    for (int i = 0; i < n; ++i) {
        const FrameSlot& fs = ctx->slots[first_slot + i];
        if (!fs.valid) return mvs_fail(ctx, MVS_ERR_ARG, "mvs_frame_fetch_batch: slot %d is empty", first_slot + i);
        if (d[i].disp && d[i].dstep < (size_t)fs.w * 4) return mvs_fail(ctx, MVS_ERR_ARG, "mvs_frame_fetch_batch: dstep too small");
    }
    MVS_CUDA(ctx, cudaSetDevice(ctx->device));
    cudaStream_t s = ctx->copy_stream;
    std::vector<int32_t> nh((size_t)n, 0);
    float* hists = (float*)arena_get(ctx, "slots.cam_hists", 0);
    bool want_cloud = false;
    for (int i = 0; i < n; ++i) {
        const int slot = first_slot + i;
        const FrameSlot& fs = ctx->slots[slot];
        const size_t px = (size_t)fs.w * fs.h;
        if (ctx->slot_done_recorded[slot]) MVS_CUDA(ctx, cudaStreamWaitEvent(s, ctx->ev_slot_done[slot], 0));
        MVS_CUDA(ctx, cudaMemcpyAsync(&nh[i], arena_get(ctx, slot_name(slot, "n").c_str(), 4), 4, cudaMemcpyDeviceToHost, s));
        if (d[i].disp)
            MVS_CUDA(ctx, cudaMemcpy2DAsync(d[i].disp, d[i].dstep, arena_get(ctx, slot_name(slot, "disp").c_str(), px * 4), (size_t)fs.w * 4,
                                            (size_t)fs.w * 4, fs.h, cudaMemcpyDeviceToHost, s));
        if (d[i].lookup) MVS_CUDA(ctx, cudaMemcpyAsync(d[i].lookup, arena_get(ctx, slot_name(slot, "lookup").c_str(), px * 4), px * 4, cudaMemcpyDeviceToHost, s));
        if (d[i].centroid) MVS_CUDA(ctx, cudaMemcpyAsync(d[i].centroid, arena_get(ctx, slot_name(slot, "centroid").c_str(), 24), 24, cudaMemcpyDeviceToHost, s));
        if (d[i].cam_hist)
            MVS_CUDA(ctx, cudaMemcpyAsync(d[i].cam_hist, hists + (size_t)slot * fs.hw * fs.hh, sizeof(float) * (size_t)fs.hw * fs.hh, cudaMemcpyDeviceToHost, s));
        if (d[i].slic_to_hist_xy && fs.n_labels > 0)
            MVS_CUDA(ctx, cudaMemcpyAsync(d[i].slic_to_hist_xy, arena_get(ctx, slot_name(slot, "cells").c_str(), 0),
                                          sizeof(mvs_sparse_hist_cell) * (size_t)fs.n_labels, cudaMemcpyDeviceToHost, s));
        want_cloud = want_cloud || d[i].Xs || d[i].xy;
    }
    MVS_CUDA(ctx, cudaStreamSynchronize(s));   // the one blocking point of a batch without clouds
    if (want_cloud) {
        for (int i = 0; i < n; ++i) {
            const int slot = first_slot + i;
            const size_t px = (size_t)ctx->slots[slot].w * ctx->slots[slot].h;
            if (nh[i] > 0 && d[i].Xs) MVS_CUDA(ctx, cudaMemcpyAsync(d[i].Xs, arena_get(ctx, slot_name(slot, "Xs").c_str(), px * 24), (size_t)nh[i] * 24, cudaMemcpyDeviceToHost, s));
            if (nh[i] > 0 && d[i].xy) MVS_CUDA(ctx, cudaMemcpyAsync(d[i].xy, arena_get(ctx, slot_name(slot, "xy").c_str(), px * 8), (size_t)nh[i] * 8, cudaMemcpyDeviceToHost, s));
        }
    }
    for (int i = 0; i < n; ++i) {
        MVS_CUDA(ctx, cudaEventRecord(ctx->ev_slot_fetched[first_slot + i], s));
        ctx->slot_fetch_recorded[first_slot + i] = true;
        if (d[i].n_out) *d[i].n_out = nh[i];
    }
    return sgbm_aggregate_wave_status(ctx, s);  // (synchronises the copy stream: the second blocking point when clouds were asked for)
}

extern "C" int mvs_frame_hist_sum(mvs_ctx* ctx, int first_slot, int n_slots, float** hist_dev, int* hw, int* hh)
{
    if (!ctx) return MVS_ERR_ARG;
    if (first_slot < 0 || n_slots <= 0 || first_slot + n_slots > MVS_MAX_SLOTS || !hist_dev)
        return mvs_fail(ctx, MVS_ERR_ARG, "mvs_frame_hist_sum: bad argument");
    const FrameSlot& f0 = ctx->slots[first_slot];
    for (int i = first_slot; i < first_slot + n_slots; ++i)
        if (!ctx->slots[i].valid || ctx->slots[i].hw != f0.hw || ctx->slots[i].hh != f0.hh)
            return mvs_fail(ctx, MVS_ERR_ARG, "mvs_frame_hist_sum: slot %d empty or of different size", i);
    MVS_CUDA(ctx, cudaSetDevice(ctx->device));
    int w = f0.hw, h = f0.hh;
    float* hists = (float*)arena_get(ctx, "slots.cam_hists", 0);
    char nm[48];
    snprintf(nm, sizeof(nm), "slots.hist_sum.%d", first_slot);
    MVS_ARENA(ctx, float, sum, nm, (size_t)w * h);
    StreamSwap on_tail(ctx, ctx->tail_stream);   // behind the batch's second half; callers continue on mvs_tail_stream()
    int rc = hist_sum_dev(ctx, hists + (size_t)first_slot * w * h, n_slots, w, h, sum);
    if (rc) return rc;
    *hist_dev = sum;
    if (hw) *hw = w;
    if (hh) *hh = h;
    return MVS_OK;
}
