// Context, error reporting, device arena, stage timers.
#include <stdarg.h>

#include <new>

#include "common.cuh"

static const char* kVersion = "mvstereo-b200 0.1 (sm_100a)";

extern "C" const char* mvs_version(void) { return kVersion; }

int mvs_fail(mvs_ctx* ctx, int code, const char* fmt, ...)
{
    if (ctx) {
        va_list ap;
        va_start(ap, fmt);
        vsnprintf(ctx->err, sizeof(ctx->err), fmt, ap);
        va_end(ap);
    }
    return code;
}

void* arena_get(mvs_ctx* ctx, const char* name, size_t bytes)
{
    Arena::Buf& b = ctx->arena.bufs[name];
    if (bytes == 0) bytes = 256;
    if (b.cap >= bytes) return b.p;
    if (b.p) {
        // growing: earlier work on any of the context's streams may still use the old block (ctx->stream may be the
        // swapped-in tail stream here, so name all three)
        cudaStreamSynchronize(ctx->stream);
        if (ctx->tail_stream) cudaStreamSynchronize(ctx->tail_stream);
        if (ctx->copy_stream) cudaStreamSynchronize(ctx->copy_stream);
        cudaFree(b.p);
        ctx->arena.total -= b.cap;
        b.p = nullptr;
        b.cap = 0;
    }
    size_t cap = (bytes + 255) & ~(size_t)255;
    cudaError_t e = cudaMalloc(&b.p, cap);
    if (e != cudaSuccess) {
        b.p = nullptr;
        mvs_fail(ctx, MVS_ERR_NOMEM, "cudaMalloc(%zu) for '%s' failed: %s", cap, name, cudaGetErrorString(e));
        return nullptr;
    }
    b.cap = cap;
    ctx->arena.total += cap;
    return b.p;
}

void stage_begin(mvs_ctx* ctx, const char* stage)
{
    StageTimer& t = ctx->timers[stage];
    if (!t.beg) {
        cudaEventCreate(&t.beg);
        cudaEventCreate(&t.end);
    }
    cudaEventRecord(t.beg, ctx->stream);
}

void stage_end(mvs_ctx* ctx, const char* stage)
{
    StageTimer& t = ctx->timers[stage];
    if (!t.end) return;
    cudaEventRecord(t.end, ctx->stream);
    t.used = true;
}

extern "C" float mvs_stage_ms(mvs_ctx* ctx, const char* stage)
{
    if (!ctx || !stage) return -1.f;
    auto it = ctx->timers.find(stage);
    if (it == ctx->timers.end() || !it->second.used) return -1.f;
    if (cudaEventSynchronize(it->second.end) != cudaSuccess) return -1.f;
    float ms = -1.f;
    if (cudaEventElapsedTime(&ms, it->second.beg, it->second.end) != cudaSuccess) return -1.f;
    return ms;
}

extern "C" int mvs_ctx_create(int device, mvs_ctx** out)
{
    if (!out) return MVS_ERR_ARG;
    *out = nullptr;
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess || n <= 0 || device < 0 || device >= n) return MVS_ERR_CUDA;
    if (cudaSetDevice(device) != cudaSuccess) return MVS_ERR_CUDA;
    mvs_ctx* ctx = new (std::nothrow) mvs_ctx();
    if (!ctx) return MVS_ERR_NOMEM;
    ctx->device = device;
    ctx->err[0] = 0;
    bool ok = cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking) == cudaSuccess &&
              cudaStreamCreateWithFlags(&ctx->copy_stream, cudaStreamNonBlocking) == cudaSuccess &&
              cudaEventCreateWithFlags(&ctx->ev_h2d, cudaEventDisableTiming) == cudaSuccess &&
              cudaStreamCreateWithFlags(&ctx->tail_stream, cudaStreamNonBlocking) == cudaSuccess &&
              cudaEventCreateWithFlags(&ctx->ev_front_done, cudaEventDisableTiming) == cudaSuccess &&
              cudaEventCreateWithFlags(&ctx->ev_tail_done[0], cudaEventDisableTiming) == cudaSuccess &&
              cudaEventCreateWithFlags(&ctx->ev_tail_done[1], cudaEventDisableTiming) == cudaSuccess;
    for (int i = 0; ok && i < MVS_MAX_SLOTS; ++i)
        ok = cudaEventCreateWithFlags(&ctx->ev_slot_done[i], cudaEventDisableTiming) == cudaSuccess &&
             cudaEventCreateWithFlags(&ctx->ev_slot_fetched[i], cudaEventDisableTiming) == cudaSuccess;
    if (!ok) {
        mvs_ctx_destroy(ctx);
        return MVS_ERR_CUDA;
    }
    cudaDeviceGetAttribute(&ctx->sm_count, cudaDevAttrMultiProcessorCount, device);
    *out = ctx;
    return MVS_OK;
}

extern "C" void mvs_ctx_destroy(mvs_ctx* ctx)
{
    if (!ctx) return;
    cudaSetDevice(ctx->device);
    if (ctx->stream) cudaStreamSynchronize(ctx->stream);
    if (ctx->tail_stream) cudaStreamSynchronize(ctx->tail_stream);
    if (ctx->copy_stream) cudaStreamSynchronize(ctx->copy_stream);
    for (auto& kv : ctx->arena.bufs)
        if (kv.second.p) cudaFree(kv.second.p);
    for (auto& m : ctx->maps)
        if (m.packed) cudaFree(m.packed);
    for (auto& kv : ctx->timers) {
        if (kv.second.beg) cudaEventDestroy(kv.second.beg);
        if (kv.second.end) cudaEventDestroy(kv.second.end);
    }
    if (ctx->ev_h2d) cudaEventDestroy(ctx->ev_h2d);
    for (int i = 0; i < MVS_MAX_SLOTS; ++i)
        if (ctx->ev_slot_done[i]) cudaEventDestroy(ctx->ev_slot_done[i]);
    for (int i = 0; i < MVS_MAX_SLOTS; ++i)
        if (ctx->ev_slot_fetched[i]) cudaEventDestroy(ctx->ev_slot_fetched[i]);
    if (ctx->ev_front_done) cudaEventDestroy(ctx->ev_front_done);
    for (int i = 0; i < 2; ++i)
        if (ctx->ev_tail_done[i]) cudaEventDestroy(ctx->ev_tail_done[i]);
    if (ctx->tail_stream) cudaStreamDestroy(ctx->tail_stream);
    if (ctx->copy_stream) cudaStreamDestroy(ctx->copy_stream);
    if (ctx->stream) cudaStreamDestroy(ctx->stream);
    delete ctx;
}

extern "C" const char* mvs_last_error(const mvs_ctx* ctx) { return ctx ? ctx->err : "null context"; }

extern "C" int mvs_sync(mvs_ctx* ctx)
{
    if (!ctx) return MVS_ERR_ARG;
    MVS_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    MVS_CUDA(ctx, cudaStreamSynchronize(ctx->tail_stream));
    MVS_CUDA(ctx, cudaStreamSynchronize(ctx->copy_stream));
    // the device-resident frame path never copies anything back, so this is where it learns that a bounded wait of
    // the aggregation sweeps expired (results invalid)
    return sgbm_aggregate_wave_status(ctx);
}

extern "C" long long mvs_launch_count(const mvs_ctx* ctx) { return ctx ? ctx->launches : -1; }

extern "C" void* mvs_stream(mvs_ctx* ctx) { return ctx ? (void*)ctx->stream : nullptr; }
extern "C" void* mvs_tail_stream(mvs_ctx* ctx) { return ctx ? (void*)ctx->tail_stream : nullptr; }
