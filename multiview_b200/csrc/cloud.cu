// a6/a7: disparity -> 3-D point cloud (reference: geometry/point-cloud.cpp:10-72 with
// geometry/projective/binocular-camera.cpp:908-927, geometry/projective/triangulation.cpp:55-95,
// geometry/vector-4.hpp:374-390,437-450).
//
// fp64, compiled with --fmad=false: the reference is built for baseline x86-64 (no FMA, no fast-math),
// so every product and sum below is rounded once, in the order the reference writes them.
// The row-major emission order of PointCloud::init is kept by an order-preserving compaction:
// per-block counts -> exclusive scan -> per-block ranks (ballot/popc), so xy / lookup / N are exact.
#include "common.cuh"

struct V3 { double x, y, z; };
__device__ __forceinline__ V3 v3(double x, double y, double z) { V3 r; r.x = x; r.y = y; r.z = z; return r; }
__device__ __forceinline__ V3 add(V3 a, V3 b) { return v3(a.x + b.x, a.y + b.y, a.z + b.z); }
__device__ __forceinline__ V3 sub(V3 a, V3 b) { return v3(a.x - b.x, a.y - b.y, a.z - b.z); }
__device__ __forceinline__ V3 mul(V3 a, double s) { return v3(a.x * s, a.y * s, a.z * s); }
__device__ __forceinline__ double dot(V3 a, V3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
__device__ __forceinline__ V3 cross(V3 a, V3 v) { return v3(a.y * v.z - a.z * v.y, a.z * v.x - a.x * v.z, a.x * v.y - a.y * v.x); }
__device__ __forceinline__ V3 normalise(V3 a)
{
    double mag2 = a.x * a.x + a.y * a.y + a.z * a.z;
    if (fabs(mag2 - 1.0) > 1e-9) {
        double inv = 1.0 / sqrt(mag2);
        a.x *= inv; a.y *= inv; a.z *= inv;
    }
    return a;
}
__device__ __forceinline__ V3 matvec(const double* m, V3 p)
{
    return v3(m[0] * p.x + m[1] * p.y + m[2] * p.z, m[3] * p.x + m[4] * p.y + m[5] * p.z,
              m[6] * p.x + m[7] * p.y + m[8] * p.z);
}

// triangulation.cpp:55-82
__device__ __forceinline__ V3 intersect_rays(V3 u0, V3 u1, V3 v0, V3 v1)
{
    V3 u = sub(u1, u0), v = sub(v1, v0), w = sub(u0, v0);
    double a = dot(u, u), b = dot(u, v), c = dot(v, v), d = dot(u, w), e = dot(v, w);
    double D = a * c - b * b, sc, tc;
    if (D < 1e-9) {
        sc = 0.0;
        tc = (b > c ? d / b : e / c);
    } else {
        sc = (b * e - c * d) / D;
        tc = (a * e - b * d) / D;
    }
    return mul(add(add(add(u0, mul(u, sc)), v0), mul(v, tc)), 0.5);
}

// binocular-camera.cpp:917-927 -> intersect_rays_2 (triangulation.cpp:84-95)
__device__ __forceinline__ V3 solve3d(const mvs_bcam& b, double x0, double x1, double y)
{
    V3 o0 = matvec(b.krr_inv[0], v3(x0, y, 1.0)), o1 = matvec(b.krr_inv[1], v3(x1, y, 1.0));
    V3 C0 = v3(b.c[0][0], b.c[0][1], b.c[0][2]), C1 = v3(b.c[1][0], b.c[1][1], b.c[1][2]);
    V3 u1 = add(C0, o0), v1 = add(C1, o1);
    V3 c = mul(add(u1, v1), 0.5);
    V3 n = normalise(cross(sub(C0, C1), sub(C0, c)));   // plane_from_3_points
    double d = -1.0 * dot(C0, n);
    double su = u1.x * n.x + u1.y * n.y + u1.z * n.z + d;  // Plane::side
    double sv = v1.x * n.x + v1.y * n.y + v1.z * n.z + d;
    return intersect_rays(C0, sub(u1, mul(n, su)), C1, sub(v1, mul(n, sv)));  // Plane::image
}

#define PC_BLOCK 256
#define PC_PER_THREAD 4
#define PC_TILE (PC_BLOCK * PC_PER_THREAD)

__device__ __forceinline__ bool pc_valid(float dx) { return isfinite(dx) && dx != 0.0f; }

__device__ __forceinline__ void
pc_count_body(const float* __restrict__ disp, size_t n, int* __restrict__ block_counts)
{
    __shared__ int wsum[PC_BLOCK / 32];
    size_t base = (size_t)blockIdx.x * PC_TILE;
    int c = 0;
#pragma unroll
    for (int k = 0; k < PC_PER_THREAD; ++k) {
        size_t i = base + (size_t)k * PC_BLOCK + threadIdx.x;
        c += (i < n && pc_valid(disp[i])) ? 1 : 0;
    }
    c = __reduce_add_sync(0xffffffffu, c);
    if ((threadIdx.x & 31) == 0) wsum[threadIdx.x >> 5] = c;
    __syncthreads();
    if (threadIdx.x == 0) {
        int t = 0;
        for (int i = 0; i < PC_BLOCK / 32; ++i) t += wsum[i];
        block_counts[blockIdx.x] = t;
    }
}

// single block: exclusive scan of block_counts -> block_offsets, total -> *n_out
__device__ __forceinline__ void pc_scan_body(const int* __restrict__ counts, int nb, int* __restrict__ offsets, int* n_out)
{
    __shared__ int wsum[32];
    __shared__ int carry;
    if (threadIdx.x == 0) carry = 0;
    __syncthreads();
    for (int base = 0; base < nb; base += 1024) {
        int i = base + threadIdx.x;
        int v = i < nb ? counts[i] : 0;
        int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
        int incl = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            int t = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o) incl += t;
        }
        if (lane == 31) wsum[wid] = incl;
        __syncthreads();
        if (wid == 0) {
            int w = wsum[lane], wi = w;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                int t = __shfl_up_sync(0xffffffffu, wi, o);
                if (lane >= o) wi += t;
            }
            wsum[lane] = wi - w;  // exclusive prefix of warp sums
        }
        __syncthreads();
        int excl = carry + wsum[wid] + incl - v;
        if (i < nb) offsets[i] = excl;
        __syncthreads();
        if (threadIdx.x == 1023) carry = excl + v;
        __syncthreads();
    }
    if (threadIdx.x == 0) *n_out = carry;
}

__device__ __forceinline__ void
pc_emit_body(const float* __restrict__ disp, int w, size_t n, const mvs_bcam& bcam, int left_is_ref,
               const int* __restrict__ block_offsets, double* __restrict__ Xs, int32_t* __restrict__ xy,
               int32_t* __restrict__ lookup, double* __restrict__ partial)
{
    __shared__ int wcount[PC_PER_THREAD][PC_BLOCK / 32];
    __shared__ double psum[PC_BLOCK / 32][3];
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    size_t base = (size_t)blockIdx.x * PC_TILE;
    float dv[PC_PER_THREAD];
    unsigned ballots[PC_PER_THREAD];
    // the tile is laid out as PC_PER_THREAD consecutive chunks of PC_BLOCK pixels: row-major order is
    // chunk k, then warp, then lane.
#pragma unroll
    for (int k = 0; k < PC_PER_THREAD; ++k) {
        size_t i = base + (size_t)k * PC_BLOCK + threadIdx.x;
        dv[k] = i < n ? disp[i] : 0.0f;
        ballots[k] = __ballot_sync(0xffffffffu, i < n && pc_valid(dv[k]));
        if (lane == 0) wcount[k][wid] = __popc(ballots[k]);
    }
    __syncthreads();
    double sx = 0, sy = 0, sz = 0;
    int run = block_offsets[blockIdx.x];
#pragma unroll
    for (int k = 0; k < PC_PER_THREAD; ++k) {
        int before = run;
        for (int ww = 0; ww < PC_BLOCK / 32; ++ww) {
            int c = wcount[k][ww];
            if (ww < wid) before += c;
            run += c;
        }
        size_t i = base + (size_t)k * PC_BLOCK + threadIdx.x;
        if (i >= n) continue;
        bool ok = (ballots[k] >> lane) & 1u;
        int idx = before + __popc(ballots[k] & ((1u << lane) - 1u));
        lookup[i] = ok ? idx : -1;
        if (ok) {
            int y = (int)(i / (size_t)w), x = (int)(i - (size_t)y * w);
            double dx = (double)dv[k];
            V3 X = left_is_ref ? solve3d(bcam, (double)x, (double)x - dx, (double)y)
                               : solve3d(bcam, (double)x + dx, (double)x, (double)y);
            Xs[3 * (size_t)idx] = X.x; Xs[3 * (size_t)idx + 1] = X.y; Xs[3 * (size_t)idx + 2] = X.z;
            xy[2 * (size_t)idx] = x; xy[2 * (size_t)idx + 1] = y;
            sx += X.x; sy += X.y; sz += X.z;
        }
    }
    // deterministic block partial sum for the centroid
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        sx += __shfl_down_sync(0xffffffffu, sx, o);
        sy += __shfl_down_sync(0xffffffffu, sy, o);
        sz += __shfl_down_sync(0xffffffffu, sz, o);
    }
    if (lane == 0) { psum[wid][0] = sx; psum[wid][1] = sy; psum[wid][2] = sz; }
    __syncthreads();
    if (threadIdx.x < 3) {
        double t = 0;
        for (int i = 0; i < PC_BLOCK / 32; ++i) t += psum[i][threadIdx.x];
        partial[(size_t)blockIdx.x * 3 + threadIdx.x] = t;
    }
}

__device__ __forceinline__ void pc_centroid_body(const double* __restrict__ partial, int nb, const int* n_dev, double* centroid)
{
    __shared__ double sh[256][3];
    double s[3] = {0, 0, 0};
    for (int i = threadIdx.x; i < nb; i += 256)
        for (int k = 0; k < 3; ++k) s[k] += partial[(size_t)i * 3 + k];
    for (int k = 0; k < 3; ++k) sh[threadIdx.x][k] = s[k];
    __syncthreads();
    for (int o = 128; o > 0; o >>= 1) {
        if (threadIdx.x < o)
            for (int k = 0; k < 3; ++k) sh[threadIdx.x][k] += sh[threadIdx.x + o][k];
        __syncthreads();
    }
    if (threadIdx.x < 3) centroid[threadIdx.x] = sh[0][threadIdx.x] / (double)(*n_dev);  // N == 0 -> NaN, as the reference
}

// launchable forms: one frame ...
__global__ void __launch_bounds__(PC_BLOCK)
pc_count_kernel(const float* __restrict__ disp, size_t n, int* __restrict__ block_counts) { pc_count_body(disp, n, block_counts); }
__global__ void __launch_bounds__(1024) pc_scan_kernel(const int* __restrict__ counts, int nb, int* __restrict__ offsets, int* n_out)
{
    pc_scan_body(counts, nb, offsets, n_out);
}
__global__ void __launch_bounds__(PC_BLOCK)
pc_emit_kernel(const float* __restrict__ disp, int w, size_t n, mvs_bcam bcam, int left_is_ref,
               const int* __restrict__ block_offsets, double* __restrict__ Xs, int32_t* __restrict__ xy,
               int32_t* __restrict__ lookup, double* __restrict__ partial)
{
    pc_emit_body(disp, w, n, bcam, left_is_ref, block_offsets, Xs, xy, lookup, partial);
}
__global__ void __launch_bounds__(256) pc_centroid_kernel(const double* __restrict__ partial, int nb, const int* n_dev, double* centroid)
{
    pc_centroid_body(partial, nb, n_dev, centroid);
}
// ... or the frames of a batch (frame = blockIdx.y, or blockIdx.x for the single-block steps); the per-frame
// pointers and calibration sit in a small device table
struct CloudFrame {
    const float* disp;
    double* Xs;
    int32_t* xy;
    int32_t* lookup;
    int32_t* n_dev;
    double* centroid;
    mvs_bcam bcam;
};
__global__ void __launch_bounds__(PC_BLOCK)
pc_count_batch_kernel(const CloudFrame* __restrict__ fr, size_t n, int nb, int* __restrict__ counts)
{
    pc_count_body(fr[blockIdx.y].disp, n, counts + (size_t)blockIdx.y * nb);
}
__global__ void __launch_bounds__(1024) pc_scan_batch_kernel(const CloudFrame* __restrict__ fr, int nb, const int* __restrict__ counts, int* __restrict__ offsets)
{
    pc_scan_body(counts + (size_t)blockIdx.x * nb, nb, offsets + (size_t)blockIdx.x * nb, fr[blockIdx.x].n_dev);
}
__global__ void __launch_bounds__(PC_BLOCK)
pc_emit_batch_kernel(const CloudFrame* __restrict__ fr, int w, size_t n, int nb, int left_is_ref, const int* __restrict__ offsets,
                     double* __restrict__ partial)
{
    const CloudFrame& f = fr[blockIdx.y];
    pc_emit_body(f.disp, w, n, f.bcam, left_is_ref, offsets + (size_t)blockIdx.y * nb, f.Xs, f.xy, f.lookup,
                 partial + (size_t)blockIdx.y * nb * 3);
}
__global__ void __launch_bounds__(256) pc_centroid_batch_kernel(const CloudFrame* __restrict__ fr, int nb, const double* __restrict__ partial)
{
    pc_centroid_body(partial + (size_t)blockIdx.x * nb * 3, nb, fr[blockIdx.x].n_dev, fr[blockIdx.x].centroid);
}

int point_cloud_batch_dev(mvs_ctx* ctx, int F, const mvs_bcam* const* bcam, const float* const* disp, int w, int h, int left_is_ref,
                          double* const* Xs, int32_t* const* xy, int32_t* const* lookup, int32_t* const* n_dev, double* const* centroid_dev)
{
    if (F < 1 || F > MVS_MAX_BATCH) return mvs_fail(ctx, MVS_ERR_ARG, "point cloud batch of %d frames", F);
    cudaStream_t s = ctx->stream;
    const size_t n = (size_t)w * h;
    const int nb = cdiv(n, PC_TILE);
    MVS_ARENA(ctx, int, counts, "cloud.batch.counts", (size_t)nb * F);
    MVS_ARENA(ctx, int, offsets, "cloud.batch.offsets", (size_t)nb * F);
    MVS_ARENA(ctx, double, partial, "cloud.batch.partial", (size_t)nb * 3 * F);
    MVS_ARENA(ctx, CloudFrame, table, "cloud.batch.table", MVS_MAX_BATCH);
    CloudFrame host[MVS_MAX_BATCH];
    for (int f = 0; f < F; ++f) {
        host[f].disp = disp[f]; host[f].Xs = Xs[f]; host[f].xy = xy[f]; host[f].lookup = lookup[f];
        host[f].n_dev = n_dev[f]; host[f].centroid = centroid_dev[f]; host[f].bcam = *bcam[f];
    }
    // pageable source: the runtime stages it before returning, so the stack array may go out of scope
    MVS_CUDA(ctx, cudaMemcpyAsync(table, host, sizeof(CloudFrame) * F, cudaMemcpyHostToDevice, s));
    pc_count_batch_kernel<<<dim3(nb, F), PC_BLOCK, 0, s>>>(table, n, nb, counts);
    pc_scan_batch_kernel<<<F, 1024, 0, s>>>(table, nb, counts, offsets);
    pc_emit_batch_kernel<<<dim3(nb, F), PC_BLOCK, 0, s>>>(table, w, n, nb, left_is_ref, offsets, partial);
    pc_centroid_batch_kernel<<<F, 256, 0, s>>>(table, nb, partial);
    MVS_LAUNCH_CHECK_N(ctx, "point cloud batch kernels", 4);
    return MVS_OK;
}

int point_cloud_dev(mvs_ctx* ctx, const mvs_bcam* bcam, const float* disp, int w, int h, int left_is_ref, double* Xs,
                    int32_t* xy, int32_t* lookup, int32_t* n_dev, double* centroid_dev)
{
    cudaStream_t s = ctx->stream;
    size_t n = (size_t)w * h;
    int nb = cdiv(n, PC_TILE);
    MVS_ARENA(ctx, int, counts, "cloud.counts", nb);
    MVS_ARENA(ctx, int, offsets, "cloud.offsets", nb);
    MVS_ARENA(ctx, double, partial, "cloud.partial", (size_t)nb * 3);
    pc_count_kernel<<<nb, PC_BLOCK, 0, s>>>(disp, n, counts);
    pc_scan_kernel<<<1, 1024, 0, s>>>(counts, nb, offsets, n_dev);
    pc_emit_kernel<<<nb, PC_BLOCK, 0, s>>>(disp, w, n, *bcam, left_is_ref, offsets, Xs, xy, lookup, partial);
    pc_centroid_kernel<<<1, 256, 0, s>>>(partial, nb, n_dev, centroid_dev);
    MVS_LAUNCH_CHECK_N(ctx, "point cloud kernels", 4);
    return MVS_OK;
}

extern "C" int mvs_point_cloud(mvs_ctx* ctx, const mvs_bcam* bcam, const float* disp, size_t step, int w, int h,
                               int left_is_ref, double* Xs, int32_t* xy, int32_t* lookup, int32_t* n_out,
                               double centroid[3])
{
    if (!ctx) return MVS_ERR_ARG;
    if (!bcam || !disp || !Xs || !xy || !lookup || !n_out || !centroid || w <= 0 || h <= 0 || step < (size_t)w * 4)
        return mvs_fail(ctx, MVS_ERR_ARG, "mvs_point_cloud: bad argument");
    MVS_CUDA(ctx, cudaSetDevice(ctx->device));
    size_t n = (size_t)w * h;
    MVS_ARENA(ctx, float, ddisp, "cloud.api.disp", n);
    MVS_ARENA(ctx, double, dXs, "cloud.api.Xs", n * 3);
    MVS_ARENA(ctx, int32_t, dxy, "cloud.api.xy", n * 2);
    MVS_ARENA(ctx, int32_t, dlk, "cloud.api.lookup", n);
    MVS_ARENA(ctx, int32_t, dn, "cloud.api.n", 1);
    MVS_ARENA(ctx, double, dc, "cloud.api.centroid", 3);
    cudaStream_t s = ctx->stream;
    MVS_CUDA(ctx, cudaMemcpy2DAsync(ddisp, (size_t)w * 4, disp, step, (size_t)w * 4, h, cudaMemcpyHostToDevice, s));
    int rc = point_cloud_dev(ctx, bcam, ddisp, w, h, left_is_ref, dXs, dxy, dlk, dn, dc);
    if (rc) return rc;
    int32_t nh = 0;
    MVS_CUDA(ctx, cudaMemcpyAsync(&nh, dn, sizeof(int32_t), cudaMemcpyDeviceToHost, s));
    MVS_CUDA(ctx, cudaMemcpyAsync(lookup, dlk, n * 4, cudaMemcpyDeviceToHost, s));
    MVS_CUDA(ctx, cudaMemcpyAsync(centroid, dc, 3 * sizeof(double), cudaMemcpyDeviceToHost, s));
    MVS_CUDA(ctx, cudaStreamSynchronize(s));
    if (nh > 0) {
        MVS_CUDA(ctx, cudaMemcpyAsync(Xs, dXs, (size_t)nh * 24, cudaMemcpyDeviceToHost, s));
        MVS_CUDA(ctx, cudaMemcpyAsync(xy, dxy, (size_t)nh * 8, cudaMemcpyDeviceToHost, s));
        MVS_CUDA(ctx, cudaStreamSynchronize(s));
    }
    *n_out = nh;
    return MVS_OK;
}

// ---- point cloud without any allocation -------------------------------------------------------------------
// Some hosts cannot let this library allocate device memory at frame time: a real-time camera loop where
// cudaMalloc's implicit device-wide synchronisation would stall every other stream, or an application whose
// own memory pool has already claimed the whole GPU so that a later cudaMalloc simply fails.  For them the
// compaction scratch of the point-cloud kernels (per-tile counts, their prefix sums, per-tile centroid partial
// sums) lives in the fixed device array below instead of the context arena.  Every GPU that loads the library
// carries these MVS_STATIC_CLOUD_SCRATCH_BYTES (288 KiB) until the process exits.
#define MVS_STATIC_CLOUD_MAX_TILES ((MVS_STATIC_CLOUD_MAX_PIXELS + PC_TILE - 1) / PC_TILE)
#define MVS_STATIC_CLOUD_SCRATCH_BYTES (MVS_STATIC_CLOUD_MAX_TILES * (2 * sizeof(int) + 3 * sizeof(double)))
static __device__ __align__(256) unsigned char g_cloud_scratch[MVS_STATIC_CLOUD_SCRATCH_BYTES];

static unsigned char* g_cloud_scratch_addr[MVS_MAX_DEVICES];  // device address of g_cloud_scratch, per device

extern "C" int mvs_point_cloud_static_dev(mvs_ctx* ctx, const mvs_bcam* bcam, const float* disp_dev, int w, int h,
                                          int left_is_ref, double* Xs_dev, int32_t* xy_dev, int32_t* lookup_dev,
                                          int32_t* n_dev, double* centroid_dev)
{
    if (!ctx) return MVS_ERR_ARG;
    if (!bcam || !disp_dev || !Xs_dev || !xy_dev || !lookup_dev || !n_dev || !centroid_dev || w <= 0 || h <= 0)
        return mvs_fail(ctx, MVS_ERR_ARG, "mvs_point_cloud_static_dev: bad argument");
    const size_t n = (size_t)w * h;
    if (n > (size_t)MVS_STATIC_CLOUD_MAX_PIXELS)
        return mvs_fail(ctx, MVS_ERR_UNSUPPORTED,
                        "mvs_point_cloud_static_dev: %d x %d = %zu pixels exceeds the fixed scratch (at most %d pixels); "
                        "use mvs_point_cloud, which sizes its scratch from the context arena",
                        w, h, n, (int)MVS_STATIC_CLOUD_MAX_PIXELS);
    if (ctx->device < 0 || ctx->device >= MVS_MAX_DEVICES)
        return mvs_fail(ctx, MVS_ERR_UNSUPPORTED, "mvs_point_cloud_static_dev: device index %d not supported", ctx->device);
    MVS_CUDA(ctx, cudaSetDevice(ctx->device));
    // In host code `g_cloud_scratch` is only a handle for the symbol, not a device address: ask the runtime for
    // the address once per device, with that device current (done on the first call for the device).
    unsigned char* base = g_cloud_scratch_addr[ctx->device];
    if (!base) {
        MVS_CUDA(ctx, cudaGetSymbolAddress((void**)&base, g_cloud_scratch));
        if (!base) return mvs_fail(ctx, MVS_ERR_CUDA, "mvs_point_cloud_static_dev: cudaGetSymbolAddress returned null");
        g_cloud_scratch_addr[ctx->device] = base;
    }
    const int nb = cdiv(n, PC_TILE);
    int* counts = (int*)base;
    int* offsets = counts + MVS_STATIC_CLOUD_MAX_TILES;
    double* partial = (double*)(offsets + MVS_STATIC_CLOUD_MAX_TILES);
    // Nothing in the scratch has to start at zero: each of the nb entries a call reads is written earlier in
    // the same call (counts by pc_count, offsets by pc_scan, partial by pc_emit), so leftovers are never seen.
    cudaStream_t s = ctx->stream;
    pc_count_kernel<<<nb, PC_BLOCK, 0, s>>>(disp_dev, n, counts);
    pc_scan_kernel<<<1, 1024, 0, s>>>(counts, nb, offsets, n_dev);
    pc_emit_kernel<<<nb, PC_BLOCK, 0, s>>>(disp_dev, w, n, *bcam, left_is_ref, offsets, Xs_dev, xy_dev, lookup_dev, partial);
    pc_centroid_kernel<<<1, 256, 0, s>>>(partial, nb, n_dev, centroid_dev);
    MVS_LAUNCH_CHECK_N(ctx, "point cloud kernels (static scratch)", 4);
    return MVS_OK;
}
