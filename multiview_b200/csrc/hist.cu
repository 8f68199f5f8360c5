// a9/a10: floor-histogram binning of one camera's point cloud and the camera sum
// (reference: cost-functions/floor-histogram.cpp:19-35,197-300; floor-histogram.hpp:90-96;
//  geometry/euclidean-transform.hpp:121-124; geometry/quaternion.hpp:126-156,205-213,274-277).
// fp64 geometry compiled with --fmad=false (bin = round(...) flips on last-ulp differences), fp32 weights.
#include <math.h>

#include "common.cuh"

struct HistArgs {
    double et[8];   // tx,ty,tz, qx,qy,qz,qw, scale
    double aabb[4]; // left, top, right, bottom
    double cc[3];   // camera centre used by the distance weights
    double hist_sz, min_z, max_z;
    int weighting, use_still, lw, lh, n_labels, hw, hh;
};

struct Q { double x, y, z, w; };
__device__ __forceinline__ Q qmul(Q a, Q b)  // quaternion.hpp:205-213
{
    Q c;
    c.x = a.y * b.z - a.z * b.y + a.x * b.w + a.w * b.x;
    c.y = a.z * b.x - a.x * b.z + a.y * b.w + a.w * b.y;
    c.z = a.x * b.y - a.y * b.x + a.z * b.w + a.w * b.z;
    c.w = a.w * b.w - a.x * b.x - a.y * b.y - a.z * b.z;
    return c;
}

// rotation.inverse() = conjugate().normalised()
__device__ __forceinline__ Q hist_inverse_rotation(const HistArgs& a)
{
    Q inv = {-a.et[3], -a.et[4], -a.et[5], a.et[6]};
    double mag2 = inv.x * inv.x + inv.y * inv.y + inv.z * inv.z + inv.w * inv.w;
    if (fabs(mag2 - 1.0) > 1e-9) {
        double s = 1.0 / sqrt(mag2);
        inv.x *= s; inv.y *= s; inv.z *= s; inv.w *= s;
    }
    return inv;
}
// bin and weight of point i; false when the point does not land in the histogram
__device__ __forceinline__ bool hist_point(const HistArgs& a, const Q& q, const Q& inv, const double* __restrict__ Xs,
                                           const int32_t* __restrict__ xy, const int32_t* __restrict__ labels,
                                           const double* __restrict__ still, int i, int* bin, float* weight,
                                           int* label_out = nullptr, double* Y_out = nullptr)
{
    double X0 = Xs[3 * (size_t)i], X1 = Xs[3 * (size_t)i + 1], X2 = Xs[3 * (size_t)i + 2];
    Q v = {X0 * a.et[7], X1 * a.et[7], X2 * a.et[7], 0.0};
    Q r = qmul(qmul(q, v), inv);
    double Yx = r.x + a.et[0], Yy = r.y + a.et[1], Yz = r.z + a.et[2];
    double zx = (Yx - a.aabb[0]) / a.hist_sz + 1e-9, zy = (Yy - a.aabb[1]) / a.hist_sz + 1e-9;
    // int(std::round(.)) of an out-of-int-range double is UB in the reference; clamp keeps such points out
    double rx = round(zx), ry = round(zy);
    if (!(rx >= 0.0 && rx < (double)a.hw && ry >= 0.0 && ry < (double)a.hh)) return false;
    if (Yz < a.min_z || Yz > a.max_z) return false;
    int bx = (int)rx, by = (int)ry;
    double dx = X0 - a.cc[0], dy = X1 - a.cc[1], dz = X2 - a.cc[2];
    double q2 = dx * dx + dy * dy + dz * dz;
    double wgt = a.weighting == MVS_WGT_NONE ? 1.0 : (a.weighting == MVS_WGT_NORM ? sqrt(q2) : q2);
    int label = -1;
    if (labels) {   // slic_lookup.in_bounds(x) ? slic_lookup(x) : -1 (floor-histogram.cpp:236)
        int px = xy[2 * (size_t)i], py = xy[2 * (size_t)i + 1];
        label = ((unsigned)px < (unsigned)a.lw && (unsigned)py < (unsigned)a.lh) ? labels[(size_t)py * a.lw + px] : -1;
        if (a.use_still && label >= 0 && label < a.n_labels) wgt *= 1.0 + 10.0 * still[label];
    }
    if (label_out) *label_out = label;
    if (Y_out) { Y_out[0] = Yx; Y_out[1] = Yy; Y_out[2] = Yz; }
    *bin = by * a.hw + bx;
    *weight = (float)wgt;
    return true;
}

// Per point: transform to world, bin, weight.  SMEM = true: the block accumulates into a private copy of the
// histogram in shared memory (float atomics on-chip) and adds its non-zero bins to the global one at the end.
template <bool SMEM>
__global__ void __launch_bounds__(512)
floor_hist_kernel(HistArgs a, const double* __restrict__ Xs, const int32_t* __restrict__ xy, const int32_t* n_dev,
                  int n_host, const int32_t* __restrict__ labels, const double* __restrict__ still,
                  float* __restrict__ hist)
{
    extern __shared__ float sh_hist[];
    const int nbins = a.hw * a.hh;
    if (SMEM) {
        for (int k = threadIdx.x; k < nbins; k += blockDim.x) sh_hist[k] = 0.0f;
        __syncthreads();
    }
    const int n = n_dev ? *n_dev : n_host;
    const Q q = {a.et[3], a.et[4], a.et[5], a.et[6]};
    const Q inv = hist_inverse_rotation(a);
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        int bin;
        float wgt;
        if (hist_point(a, q, inv, Xs, xy, labels, still, i, &bin, &wgt)) atomicAdd((SMEM ? sh_hist : hist) + bin, wgt);
    }
    if (SMEM) {
        __syncthreads();
        for (int k = threadIdx.x; k < nbins; k += blockDim.x) {
            float v = sh_hist[k];
            if (v != 0.0f) atomicAdd(hist + k, v);
        }
    }
}

// ---- order-preserving accumulation (mvs_hist_params.ordered; the super-pixel table) -------------------------------
// The reference adds weights to a bin, and centroids to a super-pixel, in point order with float arithmetic
// (floor-histogram.cpp:265-273), so a bit-exact result needs every key's contributions summed in that order:
//   (1) key (bin or label) and values per point, in parallel; rejected points and points past N get the key `nkeys`;
//   (2) a STABLE least-significant-digit radix sort of the point indices by key (8 bits per pass, two passes for up to
//       65535 keys): stability keeps the point order inside a key;
//   (3) the values gathered into sorted order, segment bounds per key, and one warp per key adding its segment up
//       front to back (coalesced loads, one dependent float add per element).
#define RS_THREADS 256
#define RS_ROWS 16                         // 32-item rows per warp
#define RS_TILE (RS_THREADS * RS_ROWS)     // items per block
// digit counts per block: cnt[digit * nblk + block]
__global__ void __launch_bounds__(RS_THREADS)
rs_count_kernel(const uint32_t* __restrict__ keys, int n, int shift, int nblk, int* __restrict__ cnt)
{
    __shared__ int h[256];
    h[threadIdx.x] = 0;
    __syncthreads();
    const int base = blockIdx.x * RS_TILE;
#pragma unroll
    for (int k = 0; k < RS_ROWS; ++k) {
        const int i = base + k * RS_THREADS + threadIdx.x;
        if (i < n) atomicAdd(&h[(keys[i] >> shift) & 255u], 1);
    }
    __syncthreads();
    cnt[threadIdx.x * nblk + blockIdx.x] = h[threadIdx.x];
}
// exclusive scan of cnt in two levels (cnt is digit-major: digit d of block b starts after every smaller digit and after
// digit d of the blocks before b): block d scans the counts of digit d over the sort's blocks and writes the digit's total
// to tot[d]; the scatter kernel adds the exclusive scan of the 256 totals itself
__global__ void __launch_bounds__(512)
rs_scan_digit_kernel(int* __restrict__ cnt, int nblk, int* __restrict__ tot)
{
    __shared__ int part[512];
    int* row = cnt + (size_t)blockIdx.x * nblk;
    int carry = 0;
    for (int b0 = 0; b0 < nblk; b0 += 512) {
        const int b = b0 + threadIdx.x;
        const int c = b < nblk ? row[b] : 0;
        part[threadIdx.x] = c;
        __syncthreads();
        for (int d = 1; d < 512; d <<= 1) {   // inclusive Hillis-Steele scan
            const int v = threadIdx.x >= d ? part[threadIdx.x - d] : 0;
            __syncthreads();
            part[threadIdx.x] += v;
            __syncthreads();
        }
        if (b < nblk) row[b] = carry + part[threadIdx.x] - c;
        carry += part[511];
        __syncthreads();
    }
    if (threadIdx.x == 0) tot[blockIdx.x] = carry;
}
// stable scatter of (key, value) by the digit; vals_in == nullptr means the identity (the first pass sorts indices)
__global__ void __launch_bounds__(RS_THREADS)
rs_scatter_kernel(const uint32_t* __restrict__ keys_in, const uint32_t* __restrict__ vals_in, int n, int shift, int nblk,
                  const int* __restrict__ offs, const int* __restrict__ tot, uint32_t* __restrict__ keys_out, uint32_t* __restrict__ vals_out)
{
    __shared__ int wcnt[RS_THREADS / 32][256];
    __shared__ int dbase[256];   // exclusive scan of the digit totals
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    for (int k = threadIdx.x; k < (RS_THREADS / 32) * 256; k += RS_THREADS) (&wcnt[0][0])[k] = 0;
    {   // RS_THREADS == 256: thread d holds digit d's total; warp scans, then the warp totals
        static_assert(RS_THREADS == 256, "one thread per digit");
        const int t = tot[threadIdx.x];
        int v = t;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) { const int u = __shfl_up_sync(0xffffffffu, v, d); if (lane >= d) v += u; }
        __shared__ int wtot[8];
        if (lane == 31) wtot[w] = v;
        __syncthreads();
        int before = 0;
        for (int ww = 0; ww < w; ++ww) before += wtot[ww];
        dbase[threadIdx.x] = before + v - t;
    }
    __syncthreads();
    const int base = blockIdx.x * RS_TILE + w * (RS_ROWS * 32);   // a warp owns RS_ROWS consecutive rows of 32 items
    // per-warp digit counts
    for (int r = 0; r < RS_ROWS; ++r) {
        const int i = base + r * 32 + lane;
        const uint32_t d = i < n ? (keys_in[i] >> shift) & 255u : 256u;
        const unsigned m = __match_any_sync(0xffffffffu, d);
        if (d < 256u && lane == __ffs(m) - 1) wcnt[w][d] += __popc(m);
        __syncwarp();
    }
    __syncthreads();
    {   // digit threadIdx.x: running start of every warp
        int run = dbase[threadIdx.x] + offs[threadIdx.x * nblk + blockIdx.x];
        for (int ww = 0; ww < RS_THREADS / 32; ++ww) { int c = wcnt[ww][threadIdx.x]; wcnt[ww][threadIdx.x] = run; run += c; }
    }
    __syncthreads();
    for (int r = 0; r < RS_ROWS; ++r) {
        const int i = base + r * 32 + lane;
        const uint32_t key = i < n ? keys_in[i] : 0u;
        const uint32_t d = i < n ? (key >> shift) & 255u : 256u;
        const unsigned m = __match_any_sync(0xffffffffu, d);
        int pos = 0;
        if (d < 256u) pos = wcnt[w][d] + __popc(m & ((1u << lane) - 1u));
        __syncwarp();   // every lane has read the running start before a leader advances it
        if (d < 256u) {
            if (lane == __ffs(m) - 1) wcnt[w][d] += __popc(m);
            keys_out[pos] = key;
            vals_out[pos] = vals_in ? vals_in[i] : (uint32_t)i;
        }
        __syncwarp();
    }
}

// sorted point indices by key: perm (and the sorted keys) in `keys` / `perm` on return; tmp buffers of n each
static int stable_sort_by_key(mvs_ctx* ctx, uint32_t* keys, uint32_t* perm, uint32_t* keys_tmp, uint32_t* perm_tmp, int n, int nkeys,
                              int* cnt /* 256 * nblk + 256 */)
{
    cudaStream_t s = ctx->stream;
    const int nblk = cdiv(n, RS_TILE);
    int bits = 1;
    while ((1 << bits) <= nkeys) ++bits;   // keys 0 .. nkeys
    const int passes = (bits + 7) / 8 + (((bits + 7) / 8) & 1);   // an even number of passes: the result lands in keys / perm
    const uint32_t *kin = keys, *vin = nullptr;
    uint32_t *kout = keys_tmp, *vout = perm_tmp;
    for (int p = 0; p < passes; ++p) {
        rs_count_kernel<<<nblk, RS_THREADS, 0, s>>>(kin, n, 8 * p, nblk, cnt);
        rs_scan_digit_kernel<<<256, 512, 0, s>>>(cnt, nblk, cnt + (size_t)256 * nblk);
        rs_scatter_kernel<<<nblk, RS_THREADS, 0, s>>>(kin, vin, n, 8 * p, nblk, cnt, cnt + (size_t)256 * nblk, kout, vout);
        MVS_LAUNCH_CHECK_N(ctx, "stable radix sort pass", 3);
        kin = kout; vin = vout;
        kout = kout == keys_tmp ? keys : keys_tmp;
        vout = vout == perm_tmp ? perm : perm_tmp;
    }
    return MVS_OK;
}

// segment [seg[k], seg[k + 1]) of key k in the sorted order (keys that do not occur get an empty segment)
__global__ void seg_bounds_kernel(const uint32_t* __restrict__ sorted_keys, int n, int nkeys, int* __restrict__ seg)
{
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= n) return;
    const uint32_t k = min(sorted_keys[j], (uint32_t)nkeys), kp = j ? min(sorted_keys[j - 1], (uint32_t)nkeys) : 0u;
    if (j == 0) for (uint32_t q = 0; q <= k; ++q) seg[q] = 0;
    else for (uint32_t q = kp + 1; q <= k; ++q) seg[q] = j;        // keys kp+1 .. k start here (those in between are empty)
    if (j == n - 1) for (uint32_t q = k + 1; q <= (uint32_t)nkeys + 1; ++q) seg[q] = n;
}
// bin + weight per point; with cell4 != nullptr also the super-pixel key and the four floats the reference accumulates
// per super-pixel: to_vec3f(Y) * float(weight) and the weight (floor-histogram.cpp:269-273)
__global__ void __launch_bounds__(256)
hist_points_kernel(HistArgs a, const double* __restrict__ Xs, const int32_t* __restrict__ xy, const int32_t* n_dev, int n_host,
                   int cap, const int32_t* __restrict__ labels, const double* __restrict__ still, uint32_t* __restrict__ bins,
                   float* __restrict__ weights, uint32_t* __restrict__ lkeys, float4* __restrict__ cell4, double sp_min_z)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= cap) return;
    const int n = n_dev ? *n_dev : n_host;
    const Q q = {a.et[3], a.et[4], a.et[5], a.et[6]};
    const Q inv = hist_inverse_rotation(a);
    int bin = a.hw * a.hh, label = -1;
    float wgt = 0.0f;
    double Y[3] = {0.0, 0.0, 0.0};
    bool ok = i < n && hist_point(a, q, inv, Xs, xy, labels, still, i, &bin, &wgt, &label, Y);
    if (!ok) bin = a.hw * a.hh;
    if (bins) { bins[i] = (uint32_t)bin; weights[i] = wgt; }
    if (cell4) {
        const bool in = ok && label >= 0 && label < a.n_labels && Y[2] >= sp_min_z;
        lkeys[i] = in ? (uint32_t)label : (uint32_t)a.n_labels;
        cell4[i] = make_float4((float)Y[0] * wgt, (float)Y[1] * wgt, (float)Y[2] * wgt, wgt);   // (this file is built --fmad=false)
    }
}
__global__ void gather_f1_kernel(const float* __restrict__ v, const uint32_t* __restrict__ perm, int n, float* __restrict__ out)
{
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j < n) out[j] = v[perm[j]];
}
__global__ void gather_f4_kernel(const float4* __restrict__ v, const uint32_t* __restrict__ perm, int n, float4* __restrict__ out)
{
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j < n) out[j] = v[perm[j]];
}
// cam_hist(x, y) += weight in point order: one WARP per bin.  The lanes load 32 consecutive weights of the bin's segment
// at once (the next 32 are already on their way), then every lane adds them up in order from shuffled copies: the sum is
// the reference's sequential float sum, the loads are coalesced, and a crowded bin costs its length in dependent adds, not in
// memory latencies.
__global__ void __launch_bounds__(256)
hist_ordered_sum_kernel(const float* __restrict__ w, const int* __restrict__ seg, int nbins, float* __restrict__ hist)
{
    const int b = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (b >= nbins) return;
    float s = 0.0f;   // cam_hist.fill(0.0f), floor-histogram.cpp:200
    const int j0 = seg[b], j1 = seg[b + 1];
    float next = j0 + lane < j1 ? w[j0 + lane] : 0.0f;
    for (int j = j0; j < j1; j += 32) {
        const float v = next;
        next = j + 32 + lane < j1 ? w[j + 32 + lane] : 0.0f;
        const int cnt = min(32, j1 - j);
#pragma unroll
        for (int k = 0; k < 32; ++k) {
            const float t = __shfl_sync(0xffffffffu, v, k);
            if (k < cnt) s += t;
        }
    }
    if (lane == 0) hist[b] = s;
}
// slic_to_hist_xy: per super-pixel, in point order, sC += Vector4f(Yf * wf, sC.w + wf)  -- as written in the reference,
// the fourth component doubles at every point (floor-histogram.cpp:270-272) -- then the cell of the centroid (:277-285)
__global__ void __launch_bounds__(256)
slic_cells_kernel(const float4* __restrict__ c, const int* __restrict__ seg, int n_labels, double left, double top,
                  double hist_sz, mvs_sparse_hist_cell* __restrict__ out)
{
    // one warp per super-pixel, like hist_ordered_sum_kernel: coalesced loads, the sums in point order from shuffled copies
    const int l = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (l >= n_labels) return;
    float sx = 0.0f, sy = 0.0f, sz = 0.0f, sw = 0.0f;
    const int j0 = seg[l], j1 = seg[l + 1];
    const float4 zero = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
    float4 next = j0 + lane < j1 ? c[j0 + lane] : zero;
    for (int j = j0; j < j1; j += 32) {
        const float4 v = next;
        next = j + 32 + lane < j1 ? c[j + 32 + lane] : zero;
        const int cnt = min(32, j1 - j);
#pragma unroll 8
        for (int k = 0; k < 32; ++k) {
            const float tx = __shfl_sync(0xffffffffu, v.x, k), ty = __shfl_sync(0xffffffffu, v.y, k);
            const float tz = __shfl_sync(0xffffffffu, v.z, k), tw = __shfl_sync(0xffffffffu, v.w, k);
            if (k < cnt) { sx += tx; sy += ty; sz += tz; sw += sw + tw; }
        }
    }
    if (lane != 0) return;
    mvs_sparse_hist_cell cell = {-1, -1, 0.0f};   // SparseHistCell{{-1, -1}, 0}
    if (sw != 0.0f) {
        const double cx = (double)sx / (double)sw, cy = (double)sy / (double)sw;
        const double zx = (cx - left) / hist_sz + 1e-9, zy = (cy - top) / hist_sz + 1e-9;
        // to_pt2(Z.round()): int(std::round(.)); out-of-int-range / NaN is undefined in the reference, clamped here
        const double rx = round(zx), ry = round(zy);
        cell.x = rx >= -2147483648.0 && rx <= 2147483647.0 ? (int)rx : -1;
        cell.y = ry >= -2147483648.0 && ry <= 2147483647.0 ? (int)ry : -1;
        cell.count = roundf(sw);
    }
    out[l] = cell;
}

__global__ void hist_sum_kernel(const float* __restrict__ cam_hists, int n_cams, size_t n, float* __restrict__ hist)
{
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float s = 0.0f;
    for (int c = 0; c < n_cams; ++c) s += cam_hists[(size_t)c * n + i];  // camera order, like hist += cam_hist
    hist[i] = s;
}

extern "C" int mvs_hist_dims(const double aabb[4], float hist_sz, int* hw, int* hh)
{
    if (!aabb || !hw || !hh || !(hist_sz > 0.0f)) return MVS_ERR_ARG;
    double inv = 1.0 / (double)hist_sz;
    *hh = (int)(ceil((aabb[3] - aabb[1]) * inv) + 1e-9);
    *hw = (int)(ceil((aabb[2] - aabb[0]) * inv) + 1e-9);
    return (*hw > 0 && *hh > 0) ? MVS_OK : MVS_ERR_ARG;
}

// ordered histogram and / or super-pixel table of one camera (cap = capacity of Xs / xy when N lives on the device)
static int floor_hist_ordered_dev(mvs_ctx* ctx, const HistArgs& a, const double* Xs, const int32_t* xy, const int32_t* n_dev, int n_host,
                                  int cap, const int32_t* labels, const double* still, double sp_min_z, float* cam_hist /* or null */,
                                  mvs_sparse_hist_cell* cells /* or null */)
{
    cudaStream_t s = ctx->stream;
    const int nbins = a.hw * a.hh;
    if (cap <= 0) cap = 1;   // one rejected pseudo-point (N = 0 keeps it out), so that every output is still written
    const int nblk = cdiv(cap, RS_TILE);
    MVS_ARENA(ctx, uint32_t, keys, "hist.ordered.keys", (size_t)cap * 4);       // bins | tmp | labels | tmp
    MVS_ARENA(ctx, uint32_t, perm, "hist.ordered.perm", (size_t)cap * 2);       // perm | tmp
    MVS_ARENA(ctx, float, wts, "hist.ordered.weights", (size_t)cap * 2);        // weights | sorted weights
    MVS_ARENA(ctx, int, cnt, "hist.ordered.cnt", (size_t)256 * nblk + 256);
    const int nk = max(nbins, cells ? a.n_labels : 0);
    MVS_ARENA(ctx, int, seg, "hist.ordered.seg", (size_t)nk + 2);
    float4* c4 = nullptr;
    if (cells) {
        c4 = (float4*)arena_get(ctx, "hist.ordered.cell4", sizeof(float4) * (size_t)cap * 2);
        if (!c4) return MVS_ERR_NOMEM;
    }
    uint32_t *bins = keys, *lkeys = keys + 2 * (size_t)cap;
    hist_points_kernel<<<cdiv(cap, 256), 256, 0, s>>>(a, Xs, xy, n_dev, n_host, cap, labels, still, cam_hist ? bins : nullptr, wts,
                                                       cells ? lkeys : nullptr, c4, sp_min_z);
    MVS_LAUNCH_CHECK(ctx, "hist_points_kernel");
    if (cam_hist) {
        int rc = stable_sort_by_key(ctx, bins, perm, bins + cap, perm + cap, cap, nbins, cnt);
        if (rc) return rc;
        gather_f1_kernel<<<cdiv(cap, 256), 256, 0, s>>>(wts, perm, cap, wts + cap);
        seg_bounds_kernel<<<cdiv(cap, 256), 256, 0, s>>>(bins, cap, nbins, seg);
        hist_ordered_sum_kernel<<<cdiv(nbins, 8), 256, 0, s>>>(wts + cap, seg, nbins, cam_hist);
        MVS_LAUNCH_CHECK_N(ctx, "ordered histogram kernels", 3);
    }
    if (cells && a.n_labels > 0) {
        int rc = stable_sort_by_key(ctx, lkeys, perm, lkeys + cap, perm + cap, cap, a.n_labels, cnt);
        if (rc) return rc;
        gather_f4_kernel<<<cdiv(cap, 256), 256, 0, s>>>(c4, perm, cap, c4 + cap);
        seg_bounds_kernel<<<cdiv(cap, 256), 256, 0, s>>>(lkeys, cap, a.n_labels, seg);
        slic_cells_kernel<<<cdiv(a.n_labels, 8), 256, 0, s>>>(c4 + cap, seg, a.n_labels, a.aabb[0], a.aabb[1], a.hist_sz, cells);
        MVS_LAUNCH_CHECK_N(ctx, "super-pixel table kernels", 3);
    }
    return MVS_OK;
}

int floor_hist_camera_dev(mvs_ctx* ctx, const mvs_hist_params* p, const double et[8], const double aabb[4],
                          const double cam_center[3], const double* Xs, const int32_t* xy, const int32_t* n_dev,
                          int n_host, const int32_t* labels, int lw, int lh, const double* still, int n_labels,
                          float* cam_hist, int hw, int hh, mvs_sparse_hist_cell* cells)
{
    HistArgs a;
    memcpy(a.et, et, sizeof(a.et));
    memcpy(a.aabb, aabb, sizeof(a.aabb));
    memcpy(a.cc, cam_center, sizeof(a.cc));
    a.hist_sz = (double)p->hist_sz; a.min_z = (double)p->min_z; a.max_z = (double)p->max_z;
    a.weighting = p->weighting; a.use_still = p->use_still && labels && still;
    a.lw = lw; a.lh = lh; a.n_labels = n_labels; a.hw = hw; a.hh = hh;
    if (!labels || n_labels <= 0) cells = nullptr;
    if (p->ordered || cells) {   // n_host is the capacity of Xs / xy when the count lives on the device
        int rc = floor_hist_ordered_dev(ctx, a, Xs, xy, n_dev, n_host, n_host, labels, still, (double)p->super_pixel_min_z,
                                        p->ordered ? cam_hist : nullptr, cells);
        if (rc || p->ordered) return rc;
    }
    MVS_CUDA(ctx, cudaMemsetAsync(cam_hist, 0, sizeof(float) * (size_t)hw * hh, ctx->stream));
    const size_t smem = sizeof(float) * (size_t)hw * hh;
    if (smem <= 100 * 1024) {  // two blocks per SM keep a private histogram each
        static bool configured[MVS_MAX_DEVICES] = {};  // function attributes are per device; setting one twice is harmless
        const int dv = ctx->device >= 0 && ctx->device < MVS_MAX_DEVICES ? ctx->device : 0;
        if (!configured[dv] || dv != ctx->device) {
            MVS_CUDA(ctx, cudaFuncSetAttribute(floor_hist_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024));
            configured[dv] = true;
        }
        floor_hist_kernel<true><<<ctx->sm_count * 2, 512, smem, ctx->stream>>>(a, Xs, xy, n_dev, n_host, labels, still, cam_hist);
    } else {
        floor_hist_kernel<false><<<ctx->sm_count * 4, 512, 0, ctx->stream>>>(a, Xs, xy, n_dev, n_host, labels, still, cam_hist);
    }
    MVS_LAUNCH_CHECK(ctx, "floor_hist_kernel");
    return MVS_OK;
}

int hist_sum_dev(mvs_ctx* ctx, const float* cam_hists, int n_cams, int hw, int hh, float* hist)
{
    size_t n = (size_t)hw * hh;
    hist_sum_kernel<<<cdiv(n, 256), 256, 0, ctx->stream>>>(cam_hists, n_cams, n, hist);
    MVS_LAUNCH_CHECK(ctx, "hist_sum_kernel");
    return MVS_OK;
}

extern "C" int mvs_floor_hist_camera(mvs_ctx* ctx, const mvs_hist_params* p, const double et[8], const double aabb[4],
                                     const double cam_center[3], const double* Xs, const int32_t* xy, int32_t n,
                                     const int32_t* labels, int lw, int lh, const double* still_score, int n_labels,
                                     float* cam_hist, int hw, int hh)
{
    return mvs_floor_hist_camera_cells(ctx, p, et, aabb, cam_center, Xs, xy, n, labels, lw, lh, still_score, n_labels, cam_hist, hw, hh, nullptr);
}

extern "C" int mvs_floor_hist_camera_cells(mvs_ctx* ctx, const mvs_hist_params* p, const double et[8], const double aabb[4],
                                           const double cam_center[3], const double* Xs, const int32_t* xy, int32_t n,
                                           const int32_t* labels, int lw, int lh, const double* still_score, int n_labels,
                                           float* cam_hist, int hw, int hh, mvs_sparse_hist_cell* slic_to_hist_xy)
{
    if (!ctx) return MVS_ERR_ARG;
    int ew = 0, eh = 0;
    if (!p || !et || !aabb || !cam_center || !cam_hist || n < 0 || (n > 0 && (!Xs || !xy)) ||
        mvs_hist_dims(aabb, p->hist_sz, &ew, &eh) || ew != hw || eh != hh ||
        p->weighting < MVS_WGT_NONE || p->weighting > MVS_WGT_QUADRANCE)
        return mvs_fail(ctx, MVS_ERR_ARG, "mvs_floor_hist_camera: bad argument (hist is %dx%d for these bounds)", ew, eh);
    MVS_CUDA(ctx, cudaSetDevice(ctx->device));
    cudaStream_t s = ctx->stream;
    MVS_ARENA(ctx, double, dXs, "hist.api.Xs", (size_t)n * 3);
    MVS_ARENA(ctx, int32_t, dxy, "hist.api.xy", (size_t)n * 2);
    MVS_ARENA(ctx, float, dh, "hist.api.hist", (size_t)hw * hh);
    int32_t* dl = nullptr;
    double* ds = nullptr;
    if (n > 0) {
        MVS_CUDA(ctx, cudaMemcpyAsync(dXs, Xs, (size_t)n * 24, cudaMemcpyHostToDevice, s));
        MVS_CUDA(ctx, cudaMemcpyAsync(dxy, xy, (size_t)n * 8, cudaMemcpyHostToDevice, s));
    }
    if (labels && still_score && lw > 0 && lh > 0 && n_labels > 0) {
        dl = (int32_t*)arena_get(ctx, "hist.api.labels", (size_t)lw * lh * 4);
        ds = (double*)arena_get(ctx, "hist.api.still", (size_t)n_labels * 8);
        if (!dl || !ds) return MVS_ERR_NOMEM;
        MVS_CUDA(ctx, cudaMemcpyAsync(dl, labels, (size_t)lw * lh * 4, cudaMemcpyHostToDevice, s));
        MVS_CUDA(ctx, cudaMemcpyAsync(ds, still_score, (size_t)n_labels * 8, cudaMemcpyHostToDevice, s));
    }
    mvs_sparse_hist_cell* dc = nullptr;
    if (slic_to_hist_xy && dl) {
        dc = (mvs_sparse_hist_cell*)arena_get(ctx, "hist.api.cells", sizeof(mvs_sparse_hist_cell) * (size_t)n_labels);
        if (!dc) return MVS_ERR_NOMEM;
    }
    int rc = floor_hist_camera_dev(ctx, p, et, aabb, cam_center, dXs, dxy, nullptr, n, dl, lw, lh, ds, n_labels, dh, hw, hh, dc);
    if (rc) return rc;
    if (dc) MVS_CUDA(ctx, cudaMemcpyAsync(slic_to_hist_xy, dc, sizeof(mvs_sparse_hist_cell) * (size_t)n_labels, cudaMemcpyDeviceToHost, s));
    MVS_CUDA(ctx, cudaMemcpyAsync(cam_hist, dh, sizeof(float) * (size_t)hw * hh, cudaMemcpyDeviceToHost, s));
    MVS_CUDA(ctx, cudaStreamSynchronize(s));
    return MVS_OK;
}

extern "C" int mvs_hist_sum(mvs_ctx* ctx, const float* cam_hists, int n_cams, int hw, int hh, float* hist)
{
    if (!ctx) return MVS_ERR_ARG;
    if (!cam_hists || !hist || n_cams <= 0 || hw <= 0 || hh <= 0) return mvs_fail(ctx, MVS_ERR_ARG, "mvs_hist_sum: bad argument");
    MVS_CUDA(ctx, cudaSetDevice(ctx->device));
    size_t n = (size_t)hw * hh;
    MVS_ARENA(ctx, float, din, "hist.api.cams", n * n_cams);
    MVS_ARENA(ctx, float, dout, "hist.api.sum", n);
    cudaStream_t s = ctx->stream;
    MVS_CUDA(ctx, cudaMemcpyAsync(din, cam_hists, n * n_cams * 4, cudaMemcpyHostToDevice, s));
    int rc = hist_sum_dev(ctx, din, n_cams, hw, hh, dout);
    if (rc) return rc;
    MVS_CUDA(ctx, cudaMemcpyAsync(hist, dout, n * 4, cudaMemcpyDeviceToHost, s));
    MVS_CUDA(ctx, cudaStreamSynchronize(s));
    return MVS_OK;
}

// ---- a10 tail: Gaussian smoothing and clipping of the summed histogram (cost-functions/floor-histogram.cpp:411-434) ----
// cv::GaussianBlur(CV_32F, ksize = ceil(sigma)*4 | 1, sigma, BORDER_DEFAULT): separable, symmetric taps
// s = c*k0 + sum_i (src[-i] + src[+i]) * k_i, evaluated in float without FMA (this file is built --fmad=false).
// OpenCV's own float path differs from this by 1-2 ulp depending on its SIMD dispatch, so parity is <= 1e-6 relative.
struct HistBlurTaps { float k[17]; int r; };

__device__ __forceinline__ int hreflect101(int i, int n)
{
    if (n == 1) return 0;
    while (i < 0 || i >= n) {
        if (i < 0) i = -i;
        if (i >= n) i = 2 * (n - 1) - i;
    }
    return i;
}
__global__ void hist_blur_h_kernel(const float* __restrict__ src, int w, int h, HistBlurTaps t, float* __restrict__ dst)
{
    int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y;
    if (x >= w) return;
    const float* row = src + (size_t)y * w;
    float s = row[x] * t.k[0];
    for (int i = 1; i <= t.r; ++i) s = s + (row[hreflect101(x - i, w)] + row[hreflect101(x + i, w)]) * t.k[i];
    dst[(size_t)y * w + x] = s;
}
__global__ void hist_blur_v_clip_kernel(const float* __restrict__ src, int w, int h, HistBlurTaps t, int do_blur, float minc,
                                        float maxc, float* __restrict__ dst)
{
    int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y;
    if (x >= w) return;
    float s = src[(size_t)y * w + x];
    if (do_blur) {
        s = s * t.k[0];
        for (int i = 1; i <= t.r; ++i)
            s = s + (src[(size_t)hreflect101(y - i, h) * w + x] + src[(size_t)hreflect101(y + i, h) * w + x]) * t.k[i];
    }
    if (s < minc) s = 0.0f;   // floor-histogram.cpp:429-430
    if (s > maxc) s = maxc;
    dst[(size_t)y * w + x] = s;
}

int hist_post_dev(mvs_ctx* ctx, float* hist, int hw, int hh, float sigma, float minc, float maxc, const char* scratch)
{
    HistBlurTaps t;
    memset(&t, 0, sizeof(t));
    const bool blur = isfinite(sigma) && sigma > 0.0f;
    if (blur) {
        int n = (int)(ceilf(sigma) * 4.0f);
        if (n % 2 == 0) n += 1;
        if (n > 33) return mvs_fail(ctx, MVS_ERR_UNSUPPORTED, "hist_post: gaussian_sigma %g needs a %d-tap kernel (max 33)", sigma, n);
        t.r = n / 2;
        // cv::getGaussianKernel(n, sigma, CV_32F): double weights, normalised, stored as float
        double kd[33], sum = 0, s2 = -0.5 / ((double)sigma * sigma);
        for (int i = 0; i < n; ++i) { double x = i - (n - 1) * 0.5; kd[i] = exp(s2 * x * x); sum += kd[i]; }
        for (int i = 0; i <= t.r; ++i) t.k[i] = (float)(kd[t.r + i] / sum);
    }
    MVS_ARENA(ctx, float, tmp, scratch, (size_t)hw * hh);   // one scratch image per calling stream (main / tail)
    dim3 grid(cdiv(hw, 128), hh);
    const float* src = hist;
    if (blur) {
        hist_blur_h_kernel<<<grid, 128, 0, ctx->stream>>>(hist, hw, hh, t, tmp);
        src = tmp;
    } else {
        MVS_CUDA(ctx, cudaMemcpyAsync(tmp, hist, sizeof(float) * (size_t)hw * hh, cudaMemcpyDeviceToDevice, ctx->stream));
        src = tmp;
    }
    hist_blur_v_clip_kernel<<<grid, 128, 0, ctx->stream>>>(src, hw, hh, t, blur ? 1 : 0, minc, maxc, hist);
    MVS_LAUNCH_CHECK_N(ctx, "hist post kernels", blur ? 2 : 1);
    return MVS_OK;
}

extern "C" int mvs_hist_post_dev(mvs_ctx* ctx, float* hist_dev, int hw, int hh, float gaussian_sigma, float min_hist_count,
                                 float max_hist_count)
{
    if (!ctx) return MVS_ERR_ARG;
    if (!hist_dev || hw <= 0 || hh <= 0) return mvs_fail(ctx, MVS_ERR_ARG, "mvs_hist_post_dev: bad argument");
    MVS_CUDA(ctx, cudaSetDevice(ctx->device));
    StreamSwap on_tail(ctx, ctx->tail_stream);   // consumes frame results: enqueued behind them
    return hist_post_dev(ctx, hist_dev, hw, hh, gaussian_sigma, min_hist_count, max_hist_count, "hist.post.tmp.tail");
}

extern "C" int mvs_hist_post(mvs_ctx* ctx, float* hist, int hw, int hh, float gaussian_sigma, float min_hist_count,
                             float max_hist_count)
{
    if (!ctx) return MVS_ERR_ARG;
    if (!hist || hw <= 0 || hh <= 0) return mvs_fail(ctx, MVS_ERR_ARG, "mvs_hist_post: bad argument");
    MVS_CUDA(ctx, cudaSetDevice(ctx->device));
    size_t n = (size_t)hw * hh;
    MVS_ARENA(ctx, float, d, "hist.post.api", n);
    MVS_CUDA(ctx, cudaMemcpyAsync(d, hist, n * 4, cudaMemcpyHostToDevice, ctx->stream));
    int rc = hist_post_dev(ctx, d, hw, hh, gaussian_sigma, min_hist_count, max_hist_count, "hist.post.tmp.main");
    if (rc) return rc;
    MVS_CUDA(ctx, cudaMemcpyAsync(hist, d, n * 4, cudaMemcpyDeviceToHost, ctx->stream));
    MVS_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return MVS_OK;
}

// ---- localization front half: the four small images LocalizationData::calculate derives from the floor
// histogram before anything else (cost-functions/localization/localization-data.cpp:845-850; make_motion_histogram
// :201-233, make_radius_filtered_hist :147-187, make_bg_histogram_ :237-253, make_fg_3d_histogram_ :264-285) and the
// pose-less foreground product (:298-315, :855-858).  One thread per cell; same operation order as the reference, fp64
// without FMA; erfc is CUDA's (<= 1 ulp in double, invisible after the float cast except on rounding ties).
__global__ void loc_motion_kernel(const float* __restrict__ hist, int n, double n_dev, float* __restrict__ motion)
{
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) motion[i] = ((double)hist[i] > n_dev) ? hist[i] : 0.0f;
}
__global__ void loc_hists_kernel(const float* __restrict__ motion, int w, int h, double hist_sz, int radius, float radius_inv,
                                 mvs_loc_params p, float* __restrict__ rhist, float* __restrict__ bg, float* __restrict__ fg0,
                                 float* __restrict__ fg)
{
    const int x0 = blockIdx.x * blockDim.x + threadIdx.x, y0 = blockIdx.y;
    if (x0 >= w) return;
    double sum = 0.0;
    long n = 0;
    for (int dy = -radius; dy <= radius; ++dy)
        for (int dx = -radius; dx <= radius; ++dx)
            if (dx * dx + dy * dy <= radius * radius) {
                const int x = x0 + dx, y = y0 + dy;
                if (x < 0 || y < 0 || x >= w || y >= h) continue;
                const double cell = (double)motion[y * w + x];
                const double dist = sqrt((double)(dx * dx + dy * dy)) * hist_sz;
                const double wght = 1.0 - dist * (double)radius_inv;
                if (wght >= 0.0) { sum += (double)(float)(wght * wght * cell); ++n; }
            }
    const float r = (float)(n == 0 ? 0.0 : sum / (double)n);
    const int i = y0 * w + x0;
    rhist[i] = r;
    const double k = -1.0 / sqrt(2.0);
    const double zb = ((double)r - p.background_median) / p.background_stddev;
    const double pb = 0.5 * erfc(k * (-zb));
    bg[i] = (float)(p.prob_bg_max < pb ? p.prob_bg_max : pb);
    const double zf = ((double)r - p.track_median) / p.track_stddev;
    const float f0 = (float)(0.5 * erfc(k * zf));
    fg0[i] = f0;
    const float a = f0 > 0.10f ? f0 : 0.10f;
    fg[i] = a * a;
}

int localization_hists_dev(mvs_ctx* ctx, const mvs_loc_params* p, const float* hist, int w, int h, double hist_sz, float* motion,
                           float* rhist, float* bg, float* fg0, float* fg)
{
    const double in_radius = 0.5 * p->person_diameter;
    const float radius_inv = 1.0f / (0.5f * (float)in_radius);
    const int radius = (int)ceil(in_radius / hist_sz);
    if (!(hist_sz > 0.0) || radius < 0 || radius > 64) return mvs_fail(ctx, MVS_ERR_ARG, "localization: radius of %d cells", radius);
    loc_motion_kernel<<<cdiv(w * h, 256), 256, 0, ctx->stream>>>(hist, w * h, p->n_deviations, motion);
    loc_hists_kernel<<<dim3(cdiv(w, 64), h), 64, 0, ctx->stream>>>(motion, w, h, hist_sz, radius, radius_inv, *p, rhist, bg, fg0, fg);
    MVS_LAUNCH_CHECK_N(ctx, "localization kernels", 2);
    return MVS_OK;
}

extern "C" int mvs_localization_hists_dev(mvs_ctx* ctx, const mvs_loc_params* p, const float* hist_dev, int hw, int hh, double hist_sz,
                                          float* motion_dev, float* rhist_dev, float* bg_dev, float* fg0_dev, float* fg_dev)
{
    if (!ctx) return MVS_ERR_ARG;
    if (!p || !hist_dev || !motion_dev || !rhist_dev || !bg_dev || !fg0_dev || !fg_dev || hw <= 0 || hh <= 0)
        return mvs_fail(ctx, MVS_ERR_ARG, "mvs_localization_hists_dev: bad argument");
    MVS_CUDA(ctx, cudaSetDevice(ctx->device));
    StreamSwap on_tail(ctx, ctx->tail_stream);   // consumes frame results: enqueued behind them
    return localization_hists_dev(ctx, p, hist_dev, hw, hh, hist_sz, motion_dev, rhist_dev, bg_dev, fg0_dev, fg_dev);
}

extern "C" int mvs_localization_hists(mvs_ctx* ctx, const mvs_loc_params* p, const float* hist, int hw, int hh, double hist_sz,
                                      float* motion, float* rhist, float* bg, float* fg0, float* fg)
{
    if (!ctx) return MVS_ERR_ARG;
    if (!p || !hist || hw <= 0 || hh <= 0) return mvs_fail(ctx, MVS_ERR_ARG, "mvs_localization_hists: bad argument");
    MVS_CUDA(ctx, cudaSetDevice(ctx->device));
    const size_t n = (size_t)hw * hh;
    MVS_ARENA(ctx, float, d, "loc.api", n * 6);
    MVS_CUDA(ctx, cudaMemcpyAsync(d, hist, n * 4, cudaMemcpyHostToDevice, ctx->stream));
    int rc = localization_hists_dev(ctx, p, d, hw, hh, hist_sz, d + n, d + 2 * n, d + 3 * n, d + 4 * n, d + 5 * n);
    if (rc) return rc;
    float* outs[5] = {motion, rhist, bg, fg0, fg};
    for (int k = 0; k < 5; ++k)
        if (outs[k]) MVS_CUDA(ctx, cudaMemcpyAsync(outs[k], d + (size_t)(k + 1) * n, n * 4, cudaMemcpyDeviceToHost, ctx->stream));
    MVS_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return MVS_OK;
}
