// Cell-linked-list construction fused with the cell-index particle sort.
//
// Reference behaviour reproduced:
//   meshes/base_mesh.hxx:9-26,68-99          cell index / linear index / Morton key
//   meshes/cell_linked_list.cpp:84-101        UpdateCellLists (index lists + position copies)
//   shared_ck/.../update_cell_linked_list.hpp:40-106   count → scan → scatter (CSR CellOffset)
//   configuration_dynamics/particle_sorting.cpp:16-58  sort of all evolving variables + ID maps
//
// B200 design: the cell list IS a counting sort.  One pass computes the linear cell of every particle
// and its rank inside the cell with warp-aggregated atomics (__match_any_sync groups the lanes that
// fall in the same cell, one atomicAdd per group); a scan gives CellOffset; a scatter gives the
// permutation, made deterministic (stable) by ordering each cell's few entries by their previous
// index; every registered variable is then gathered into cell order, so afterwards the particles of a
// cell - and of a run of cells along the fastest axis - are contiguous in HBM and neighbour gathers
// hit consecutive sectors.  The frozen position snapshot + Vol is packed into one aligned Real4 record
// per particle (xV) so a neighbour costs one 16/32-byte load.
#include "internal.cuh"
#include <mutex>
#include <string>

using namespace sphb;

namespace sphb
{
// ---------------------------------------------------------------------------------------------
// exclusive scan (u32): 3-phase, 4096 items per block
// ---------------------------------------------------------------------------------------------
constexpr int kScanBlock = 1024;
constexpr int kScanItems = 4;
constexpr int kScanTile = kScanBlock * kScanItems;

__device__ __forceinline__ u32 blockExclusiveScan(u32 v, u32 *warp_sums, u32 &block_total)
{
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    u32 inc = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1)
    {
        u32 t = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o)
            inc += t;
    }
    if (lane == 31)
        warp_sums[wid] = inc;
    __syncthreads();
    if (wid == 0)
    {
        u32 w = lane < (blockDim.x >> 5) ? warp_sums[lane] : 0;
        u32 winc = w;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1)
        {
            u32 t = __shfl_up_sync(0xffffffffu, winc, o);
            if (lane >= o)
                winc += t;
        }
        warp_sums[lane] = winc - w;
        if (lane == 31)
            warp_sums[32] = winc;
    }
    __syncthreads();
    block_total = warp_sums[32];
    u32 r = warp_sums[wid] + inc - v;
    __syncthreads();
    return r;
}

__global__ void __launch_bounds__(kScanBlock) k_scan_tiles(const u32 *in, u32 *out, u32 *tile_sums, int64_t n)
{
    __shared__ u32 ws[33];
    int64_t base = int64_t(blockIdx.x) * kScanTile + int64_t(threadIdx.x) * kScanItems;
    u32 v[kScanItems];
    u32 s = 0;
#pragma unroll
    for (int k = 0; k < kScanItems; ++k)
    {
        v[k] = (base + k < n) ? in[base + k] : 0;
        s += v[k];
    }
    u32 total;
    u32 ex = blockExclusiveScan(s, ws, total);
#pragma unroll
    for (int k = 0; k < kScanItems; ++k)
    {
        if (base + k < n)
            out[base + k] = ex;
        ex += v[k];
    }
    if (threadIdx.x == 0)
        tile_sums[blockIdx.x] = total;
}
// single block: exclusive scan of the tile sums in place; total → tile_sums[n_tiles]
__global__ void __launch_bounds__(kScanBlock) k_scan_tile_sums(u32 *tile_sums, int64_t n_tiles)
{
    __shared__ u32 ws[33];
    u32 carry = 0;
    for (int64_t base = 0; base < n_tiles; base += kScanBlock)
    {
        int64_t i = base + threadIdx.x;
        u32 v = i < n_tiles ? tile_sums[i] : 0;
        u32 total;
        u32 ex = blockExclusiveScan(v, ws, total);
        if (i < n_tiles)
            tile_sums[i] = carry + ex;
        carry += total;
    }
    if (threadIdx.x == 0)
        tile_sums[n_tiles] = carry;
}
__global__ void k_scan_add(u32 *out, const u32 *tile_sums, int64_t n, int64_t n_tiles, int with_total)
{
    int64_t i = blockIdx.x * int64_t(blockDim.x) + threadIdx.x;
    if (i < n)
        out[i] += tile_sums[i / kScanTile];
    if (with_total && i == 0)
        out[n] = tile_sums[n_tiles];
}

sphb200_status exclusiveScanU32(sphb200_ctx *ctx, const u32 *in, u32 *out, int64_t n, bool with_total)
{
    if (n <= 0)
    {
        if (with_total)
            SPHB_CUDA(ctx, cudaMemsetAsync(out, 0, 4, ctx->stream));
        return SPHB200_OK;
    }
    int64_t n_tiles = (n + kScanTile - 1) / kScanTile;
    SPHB_CUDA(ctx, ctx->scan_tmp.ensure(size_t(n_tiles + 1) * 4));
    u32 *ts = static_cast<u32 *>(ctx->scan_tmp.p);
    k_scan_tiles<<<int(n_tiles), kScanBlock, 0, ctx->stream>>>(in, out, ts, n);
    SPHB_LAUNCH_CHECK(ctx);
    k_scan_tile_sums<<<1, kScanBlock, 0, ctx->stream>>>(ts, n_tiles);
    SPHB_LAUNCH_CHECK(ctx);
    k_scan_add<<<gridFor(n, 256), 256, 0, ctx->stream>>>(out, ts, n, n_tiles, with_total ? 1 : 0);
    SPHB_LAUNCH_CHECK(ctx);
    return SPHB200_OK;
}

// ---------------------------------------------------------------------------------------------
// binning
// ---------------------------------------------------------------------------------------------
template <int D, class R>
__global__ void __launch_bounds__(256) k_bin(const R *__restrict__ pos, int64_t n, MeshP<D, R> mesh,
                                             u32 *__restrict__ cell_of, u32 *__restrict__ rank,
                                             u32 *__restrict__ cell_count)
{
    int64_t i = blockIdx.x * int64_t(blockDim.x) + threadIdx.x;
    bool active = i < n;
    u32 c = 0xffffffffu;
    if (active)
    {
        R x[D];
#pragma unroll
        for (int d = 0; d < D; ++d)
            x[d] = pos[i * D + d];
        int ci[D];
        mesh.cellIndex(x, ci);
        c = mesh.linear(ci);
    }
    // warp-aggregated atomic: lanes with the same cell elect a leader that adds the group size
    unsigned amask = __ballot_sync(0xffffffffu, active);
    if (active)
    {
        unsigned peers = __match_any_sync(amask, c);
        int leader = __ffs(peers) - 1;
        int lane = threadIdx.x & 31;
        u32 base = 0;
        if (lane == leader)
            base = atomicAdd(cell_count + c, u32(__popc(peers)));
        base = __shfl_sync(peers, base, leader);
        cell_of[i] = c;
        rank[i] = base + u32(__popc(peers & ((1u << lane) - 1u)));
    }
}

__global__ void __launch_bounds__(256) k_scatter_perm(const u32 *__restrict__ cell_of, const u32 *__restrict__ rank,
                                                      const u32 *__restrict__ cell_offset, u32 *__restrict__ perm,
                                                      int64_t n)
{
    int64_t i = blockIdx.x * int64_t(blockDim.x) + threadIdx.x;
    if (i < n)
        perm[cell_offset[cell_of[i]] + rank[i]] = u32(i);
}

// In-cell order.  The reference's sort is unstable, so any order inside a cell is a legal outcome; this one is
// geometric: a boustrophedon (snake) walk over 8^D sub-bins of the cell, previous index as the tie-break.
// Consecutive particles are then nearest neighbours wherever the flow has taken them, which is what the
// union lists of the register-tiled kernels (tile.cuh) want - the union of a close pair is smaller - and the
// order is a function of the positions, reproducible run to run and between a slab run and the whole-domain run.
template <int D, class R>
__global__ void __launch_bounds__(256) k_cell_subkey(const R *__restrict__ pos, const u32 *__restrict__ perm,
                                                     const u32 *__restrict__ cell_of, MeshP<D, R> mesh, int64_t n,
                                                     unsigned long long *__restrict__ keys)
{
    int64_t k = blockIdx.x * int64_t(blockDim.x) + threadIdx.x;
    if (k >= n)
        return;
    const u32 i = perm[k];
    u32 c = cell_of[i]; // linear index: last axis fastest
    int q[D];
#pragma unroll
    for (int a = D - 1; a >= 0; --a)
    {
        const u32 ca = c % u32(mesh.cells[a]);
        c /= u32(mesh.cells[a]);
        const R u = (pos[size_t(i) * D + a] - (mesh.lo[a] + R(ca) * mesh.spacing)) / mesh.spacing;
        int v = int(u * R(8));
        q[a] = v < 0 ? 0 : (v > 7 ? 7 : v); // clamped particles and rounding at the faces
    }
    u32 row = 0, key = 0;
#pragma unroll
    for (int a = 0; a < D; ++a)
    {
        const int v = (row & 1u) ? 7 - q[a] : q[a];
        key = key * 8u + u32(v);
        row = row * 8u + u32(v);
    }
    keys[k] = ((unsigned long long)key << 32) | i;
}

__global__ void __launch_bounds__(256) k_cell_order(const u32 *__restrict__ cell_offset, unsigned long long *__restrict__ keys,
                                                    u32 *__restrict__ perm, int64_t n_cells)
{
    int64_t c = blockIdx.x * int64_t(blockDim.x) + threadIdx.x;
    if (c >= n_cells)
        return;
    u32 b = cell_offset[c], e = cell_offset[c + 1];
    for (u32 k = b + 1; k < e; ++k)
    {
        unsigned long long v = keys[k];
        u32 m = k;
        while (m > b && keys[m - 1] > v)
        {
            keys[m] = keys[m - 1];
            --m;
        }
        keys[m] = v;
    }
    for (u32 k = b; k < e; ++k)
        perm[k] = u32(keys[k]);
}

// gather of one variable into the new order; T = element type, width elements per particle
template <class T>
__global__ void __launch_bounds__(256) k_permute(const T *__restrict__ src, T *__restrict__ dst,
                                                 const u32 *__restrict__ perm, int64_t n, int width)
{
    int64_t t = blockIdx.x * int64_t(blockDim.x) + threadIdx.x;
    if (t >= n * width)
        return;
    int64_t i = t / width;
    int d = int(t - i * width);
    dst[t] = src[int64_t(perm[i]) * width + d];
}

// the same gather for the common widths (1, 2, 3): a thread moves two consecutive particles, W loads in flight per
// particle and W two-element (8/16-byte) stores, which are aligned because the pair starts at an even particle
template <class T>
struct Vec2Of;
template <>
struct Vec2Of<float>
{
    typedef float2 type;
};
template <>
struct Vec2Of<double>
{
    typedef double2 type;
};
template <class T, int W>
__global__ void __launch_bounds__(256) k_permute_w(const T *__restrict__ src, T *__restrict__ dst, const u32 *__restrict__ perm, int64_t n)
{
    const int64_t t = blockIdx.x * int64_t(blockDim.x) + threadIdx.x;
    const int64_t i = 2 * t;
    if (i >= n)
        return;
    T v[2 * W];
    const int64_t p0 = int64_t(perm[i]), p1 = i + 1 < n ? int64_t(perm[i + 1]) : p0;
#pragma unroll
    for (int d = 0; d < W; ++d)
    {
        v[d] = src[p0 * W + d];
        v[W + d] = src[p1 * W + d];
    }
    if (i + 1 < n)
    {
        typename Vec2Of<T>::type *out = reinterpret_cast<typename Vec2Of<T>::type *>(dst + i * W);
#pragma unroll
        for (int d = 0; d < W; ++d)
        {
            typename Vec2Of<T>::type o;
            o.x = v[2 * d];
            o.y = v[2 * d + 1];
            out[d] = o;
        }
    }
    else
    {
#pragma unroll
        for (int d = 0; d < W; ++d)
            dst[i * W + d] = v[d];
    }
}

__global__ void __launch_bounds__(256) k_permute_ids(const u32 *__restrict__ oid_src, u32 *__restrict__ oid_dst,
                                                     u32 *__restrict__ sorted_id, const u32 *__restrict__ cell_src,
                                                     u32 *__restrict__ cell_dst, const u32 *__restrict__ perm, int64_t n)
{
    int64_t i = blockIdx.x * int64_t(blockDim.x) + threadIdx.x;
    if (i >= n)
        return;
    u32 p = perm[i];
    u32 o = oid_src[p];
    oid_dst[i] = o;
    sorted_id[o] = u32(i); // particle_sorting.cpp:55-58 UpdateSortedID
    cell_dst[i] = cell_src[p];
}

template <int D, class R>
__global__ void __launch_bounds__(256) k_pack_xV(const R *__restrict__ pos, const R *__restrict__ vol,
                                                 typename Real4<R>::type *__restrict__ xV, int64_t n, int vol_only)
{
    int64_t i = blockIdx.x * int64_t(blockDim.x) + threadIdx.x;
    if (i >= n)
        return;
    typename Real4<R>::type v;
    if (vol_only)
    {
        v = xV[i];
    }
    else
    {
        v.x = pos[i * D];
        v.y = pos[i * D + 1];
        v.z = D == 3 ? pos[i * D + (D - 1)] : R(0);
    }
    v.w = vol[i];
    xV[i] = v;
}

template <int D, class R>
__global__ void __launch_bounds__(256) k_periodic_bounding(R *__restrict__ pos, int64_t n, int axis, R lo, R hi)
{
    int64_t i = blockIdx.x * int64_t(blockDim.x) + threadIdx.x;
    if (i >= n)
        return;
    // domain_bounding.h:101-137: checkLowerBound then checkUpperBound
    R L = hi - lo;
    R x = pos[i * D + axis];
    if (x < lo)
        x += L;
    if (x > hi)
        x -= L;
    pos[i * D + axis] = x;
}

__host__ __device__ inline u64 mortonPart(u64 i)
{
    // base_mesh.hxx:80-90: 10-bit spread
    u64 x = i;
    x &= 0x3ff;
    x = (x | x << 16) & 0x30000ff;
    x = (x | x << 8) & 0x300f00f;
    x = (x | x << 4) & 0x30c30c3;
    x = (x | x << 2) & 0x9249249;
    return x;
}

template <int D, class R>
__global__ void __launch_bounds__(256) k_cell_keys(const R *__restrict__ pos, int64_t n, MeshP<D, R> mesh,
                                                   long long *__restrict__ linear, long long *__restrict__ morton)
{
    int64_t i = blockIdx.x * int64_t(blockDim.x) + threadIdx.x;
    if (i >= n)
        return;
    R x[D];
#pragma unroll
    for (int d = 0; d < D; ++d)
        x[d] = pos[i * D + d];
    int c[D];
    mesh.cellIndex(x, c);
    linear[i] = (long long)mesh.linear(c);
    u64 m = mortonPart(u64(c[0])) | (mortonPart(u64(c[1])) << 1);
    if (D == 3)
        m |= mortonPart(u64(c[D - 1])) << 2;
    morton[i] = (long long)m;
}

// min / max of Vol over the whole body as u64 bit patterns of non-negative doubles (order preserving)
template <class R>
__global__ void __launch_bounds__(256) k_vol_minmax(const typename Real4<R>::type *__restrict__ xV, int64_t n,
                                                    unsigned long long *__restrict__ mm)
{
    double lo = 1.7976931348623157e308, hi = 0.0;
    for (int64_t i = blockIdx.x * int64_t(blockDim.x) + threadIdx.x; i < n; i += int64_t(gridDim.x) * blockDim.x)
    {
        const double v = double(reinterpret_cast<const R *>(xV + i)[3]);
        lo = fmin(lo, v);
        hi = fmax(hi, v);
    }
    lo = warpMin<double>(lo);
    hi = warpMax<double>(hi);
    if ((threadIdx.x & 31) == 0)
    {
        atomicMinNonNeg(mm, lo);
        atomicMaxNonNeg(mm + 1, hi);
    }
}

template <int D, class R>
sphb200_status refreshVolRange(sphb200_body *b)
{
    sphb200_ctx *ctx = b->ctx;
    SPHB_CUDA(ctx, b->vol_mm.ensure(16));
    // min slot = all ones (above every finite pattern), max slot = 0: an empty body reads as "not uniform"
    SPHB_CUDA(ctx, cudaMemsetAsync(b->vol_mm.p, 0xff, 8, ctx->stream));
    SPHB_CUDA(ctx, cudaMemsetAsync(static_cast<char *>(b->vol_mm.p) + 8, 0, 8, ctx->stream));
    if (b->n > 0)
    {
        k_vol_minmax<R><<<148 * 2, 256, 0, ctx->stream>>>(static_cast<const typename Real4<R>::type *>(b->xV.p), b->n,
                                                       static_cast<unsigned long long *>(b->vol_mm.p));
        SPHB_LAUNCH_CHECK(ctx);
    }
    return SPHB200_OK;
}
template sphb200_status refreshVolRange<2, float>(sphb200_body *);
template sphb200_status refreshVolRange<2, double>(sphb200_body *);
template sphb200_status refreshVolRange<3, float>(sphb200_body *);
template sphb200_status refreshVolRange<3, double>(sphb200_body *);

template <int D, class R>
sphb200_status ensurePacked(sphb200_body *b)
{
    sphb200_ctx *ctx = b->ctx;
    if (!b->cell_list_valid)
        return ctx->fail(SPHB200_ERR_INVALID, "body '" + b->name + "': cell_list_update has not been called");
    if (!b->xV_vol_dirty || b->n == 0)
        return SPHB200_OK;
    R *pos, *vol;
    SPHB_TRY(getVarPtr<R>(b, "Position", D, &pos));
    SPHB_TRY(getVarPtr<R>(b, "VolumetricMeasure", 1, &vol));
    k_pack_xV<D, R><<<gridFor(b->n, 256), 256, 0, ctx->stream>>>(
        pos, vol, static_cast<typename Real4<R>::type *>(b->xV.p), b->n, 1);
    SPHB_LAUNCH_CHECK(ctx);
    b->xV_vol_dirty = false;
    b->rec_b_valid = false;
    return refreshVolRange<D, R>(b);
}
template sphb200_status ensurePacked<2, float>(sphb200_body *);
template sphb200_status ensurePacked<2, double>(sphb200_body *);
template sphb200_status ensurePacked<3, float>(sphb200_body *);
template sphb200_status ensurePacked<3, double>(sphb200_body *);

template <int D, class R>
static sphb200_status cellListUpdate(sphb200_body *b)
{
    sphb200_ctx *ctx = b->ctx;
    cudaStream_t st = ctx->stream;
    const int64_t n = b->n;
    BodyP<D, R> &P = b->P<D, R>();
    const int64_t nc = b->n_cells;
    const size_t nb = size_t(n > 0 ? n : 1);
    SPHB_CUDA(ctx, b->cell_of.ensure(nb * 4));
    SPHB_CUDA(ctx, b->rank_in_cell.ensure(nb * 4));
    SPHB_CUDA(ctx, b->perm.ensure(nb * 4));
    SPHB_CUDA(ctx, b->cell_count.ensure(size_t(nc + 1) * 4));
    SPHB_CUDA(ctx, b->cell_offset.ensure(size_t(nc + 1) * 4));
    SPHB_CUDA(ctx, b->xV.ensure(nb * 4 * sizeof(R)));
    SPHB_CUDA(ctx, ctx->scratch_a.ensure(nb * 4));
    SPHB_CUDA(ctx, ctx->scratch_b.ensure(nb * 4));
    u32 *cell_of = static_cast<u32 *>(b->cell_of.p), *rank = static_cast<u32 *>(b->rank_in_cell.p),
        *perm = static_cast<u32 *>(b->perm.p), *cnt = static_cast<u32 *>(b->cell_count.p),
        *off = static_cast<u32 *>(b->cell_offset.p);
    R *pos, *vol;
    SPHB_TRY(getVarPtr<R>(b, "Position", D, &pos));
    SPHB_CUDA(ctx, cudaMemsetAsync(cnt, 0, size_t(nc + 1) * 4, st));
    if (n > 0)
    {
        k_bin<D, R><<<gridFor(n, 256), 256, 0, st>>>(pos, n, P.mesh, cell_of, rank, cnt);
        SPHB_LAUNCH_CHECK(ctx);
    }
    SPHB_TRY(exclusiveScanU32(ctx, cnt, off, nc, true));
    if (n > 0)
    {
        k_scatter_perm<<<gridFor(n, 256), 256, 0, st>>>(cell_of, rank, off, perm, n);
        SPHB_LAUNCH_CHECK(ctx);
        SPHB_CUDA(ctx, ctx->subkey_buf.ensure(size_t(n) * 8));
        auto *subkeys = static_cast<unsigned long long *>(ctx->subkey_buf.p);
        k_cell_subkey<D, R><<<gridFor(n, 256), 256, 0, st>>>(pos, perm, cell_of, P.mesh, n, subkeys);
        SPHB_LAUNCH_CHECK(ctx);
        k_cell_order<<<gridFor(nc, 256), 256, 0, st>>>(off, subkeys, perm, nc);
        SPHB_LAUNCH_CHECK(ctx);
        // gather every registered variable into cell order (UpdateSortableVariables,
        // shared_ck/.../base_configuration_dynamics.h:74-128, one temp buffer swapped in)
        for (auto &kv : b->vars)
        {
            Var &v = *kv.second;
            SPHB_CUDA(ctx, v.alt.ensure(size_t(n) * size_t(v.width) * sizeof(R)));
            const R *vs = static_cast<const R *>(v.dev.p);
            R *vd = static_cast<R *>(v.alt.p);
            const int pair_grid = gridFor((n + 1) / 2, 256);
            if (v.width == 1)
                k_permute_w<R, 1><<<pair_grid, 256, 0, st>>>(vs, vd, perm, n);
            else if (v.width == 2)
                k_permute_w<R, 2><<<pair_grid, 256, 0, st>>>(vs, vd, perm, n);
            else if (v.width == 3)
                k_permute_w<R, 3><<<pair_grid, 256, 0, st>>>(vs, vd, perm, n);
            else
                k_permute<R><<<gridFor(n * v.width, 256), 256, 0, st>>>(vs, vd, perm, n, v.width);
            SPHB_LAUNCH_CHECK(ctx);
            std::swap(v.dev, v.alt);
        }
        u32 *oid_tmp = static_cast<u32 *>(ctx->scratch_a.p), *cell_tmp = static_cast<u32 *>(ctx->scratch_b.p);
        k_permute_ids<<<gridFor(n, 256), 256, 0, st>>>(static_cast<const u32 *>(b->original_id.p), oid_tmp,
                                                      static_cast<u32 *>(b->sorted_id.p), cell_of, cell_tmp, perm, n);
        SPHB_LAUNCH_CHECK(ctx);
        SPHB_CUDA(ctx, cudaMemcpyAsync(b->original_id.p, oid_tmp, size_t(n) * 4, cudaMemcpyDeviceToDevice, st));
        SPHB_CUDA(ctx, cudaMemcpyAsync(cell_of, cell_tmp, size_t(n) * 4, cudaMemcpyDeviceToDevice, st));
        SPHB_TRY(getVarPtr<R>(b, "Position", D, &pos));
        SPHB_TRY(getVarPtr<R>(b, "VolumetricMeasure", 1, &vol));
        k_pack_xV<D, R><<<gridFor(n, 256), 256, 0, st>>>(pos, vol, static_cast<typename Real4<R>::type *>(b->xV.p), n, 0);
        SPHB_LAUNCH_CHECK(ctx);
    }
    b->cell_list_valid = true;
    b->xV_vol_dirty = false;
    b->rec_b_valid = false;
    b->config_stamp++;
    SPHB_TRY((refreshVolRange<D, R>(b)));
    return updateOwnedRange(b);
}

template <int D, class R>
static sphb200_status cellKeys(sphb200_body *b, int64_t *linear, int64_t *morton)
{
    sphb200_ctx *ctx = b->ctx;
    const int64_t n = b->n;
    if (n == 0)
        return SPHB200_OK;
    SPHB_CUDA(ctx, ctx->scratch_a.ensure(size_t(n) * 8));
    SPHB_CUDA(ctx, ctx->scratch_b.ensure(size_t(n) * 8));
    R *pos;
    SPHB_TRY(getVarPtr<R>(b, "Position", D, &pos));
    k_cell_keys<D, R><<<gridFor(n, 256), 256, 0, ctx->stream>>>(pos, n, b->P<D, R>().mesh,
                                                              static_cast<long long *>(ctx->scratch_a.p),
                                                              static_cast<long long *>(ctx->scratch_b.p));
    SPHB_LAUNCH_CHECK(ctx);
    if (linear)
        SPHB_CUDA(ctx, cudaMemcpyAsync(linear, ctx->scratch_a.p, size_t(n) * 8, cudaMemcpyDeviceToHost, ctx->stream));
    if (morton)
        SPHB_CUDA(ctx, cudaMemcpyAsync(morton, ctx->scratch_b.p, size_t(n) * 8, cudaMemcpyDeviceToHost, ctx->stream));
    SPHB_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return SPHB200_OK;
}

template <int D, class R>
static sphb200_status periodicBounding(sphb200_body *b, int axis)
{
    sphb200_ctx *ctx = b->ctx;
    if (b->n == 0)
        return SPHB200_OK;
    R *pos;
    SPHB_TRY(getVarPtr<R>(b, "Position", D, &pos));
    BodyP<D, R> &P = b->P<D, R>();
    k_periodic_bounding<D, R><<<gridFor(b->n, 256), 256, 0, ctx->stream>>>(pos, b->n, axis, P.per.lo[axis], P.per.hi[axis]);
    SPHB_LAUNCH_CHECK(ctx);
    return SPHB200_OK;
}
} // namespace sphb

extern "C"
{
sphb200_status sphb200_cell_list_update(sphb200_body *b)
{
    if (!b)
        return SPHB200_ERR_INVALID;
    cudaSetDevice(b->ctx->device);
    return SPHB_DISPATCH(b->ctx, cellListUpdate, b);
}

sphb200_status sphb200_cell_keys(sphb200_body *b, int64_t *linear, int64_t *morton)
{
    if (!b)
        return SPHB200_ERR_INVALID;
    cudaSetDevice(b->ctx->device);
    return SPHB_DISPATCH(b->ctx, cellKeys, b, linear, morton);
}

sphb200_status sphb200_cell_offsets(sphb200_body *b, int64_t *out, int64_t *n_cells)
{
    if (!b)
        return SPHB200_ERR_INVALID;
    sphb200_ctx *ctx = b->ctx;
    if (n_cells)
        *n_cells = b->n_cells;
    if (!out)
        return SPHB200_OK;
    if (!b->cell_list_valid)
        return ctx->fail(SPHB200_ERR_INVALID, "cell_offsets: cell_list_update has not been called");
    cudaSetDevice(ctx->device);
    std::vector<u32> tmp;
    tmp.resize(size_t(b->n_cells + 1));
    SPHB_CUDA(ctx, cudaMemcpyAsync(tmp.data(), b->cell_offset.p, tmp.size() * 4, cudaMemcpyDeviceToHost, ctx->stream));
    SPHB_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    for (size_t i = 0; i < tmp.size(); ++i)
        out[i] = int64_t(tmp[i]);
    return SPHB200_OK;
}

sphb200_status sphb200_periodic_bounding(sphb200_body *b, int axis)
{
    if (!b || axis < 0 || axis >= b->ctx->ndim || !b->periodic[axis])
        return SPHB200_ERR_INVALID;
    cudaSetDevice(b->ctx->device);
    return SPHB_DISPATCH(b->ctx, periodicBounding, b, axis);
}

sphb200_status sphb200_exclusive_scan_u32(sphb200_ctx *ctx, const uint32_t *in, uint32_t *out, int64_t n, uint64_t *total)
{
    if (!ctx || n < 0 || (n > 0 && (!in || !out)))
        return SPHB200_ERR_INVALID;
    cudaSetDevice(ctx->device);
    SPHB_CUDA(ctx, ctx->scratch_a.ensure(size_t(n + 1) * 4));
    SPHB_CUDA(ctx, ctx->scratch_b.ensure(size_t(n + 1) * 4));
    if (n > 0)
        SPHB_CUDA(ctx, cudaMemcpyAsync(ctx->scratch_a.p, in, size_t(n) * 4, cudaMemcpyHostToDevice, ctx->stream));
    SPHB_TRY(exclusiveScanU32(ctx, static_cast<u32 *>(ctx->scratch_a.p), static_cast<u32 *>(ctx->scratch_b.p), n, true));
    std::vector<u32> tmp;
    tmp.resize(size_t(n + 1));
    SPHB_CUDA(ctx, cudaMemcpyAsync(tmp.data(), ctx->scratch_b.p, size_t(n + 1) * 4, cudaMemcpyDeviceToHost, ctx->stream));
    SPHB_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    if (n > 0)
        memcpy(out, tmp.data(), size_t(n) * 4);
    if (total)
        *total = tmp[size_t(n)];
    return SPHB200_OK;
}
} // extern "C"

// ---------------------------------------------------------------------------------------------
// Exclusive scan that allocates nothing (no counterpart in the reference; added
// for callers that cannot call cudaMalloc: CUDA graph capture, tight serving loops, harnesses that
// forbid it).  Same three kernels as exclusiveScanU32; the tile sums - the scan's only scratch - live in a
// file-scope device array instead of ctx->scan_tmp.  Every GPU that loads the library holds these 64 KB + 256 B
// until exit.  In host code `g_scan_static` is a host-side stand-in, not a device address: the address the
// kernels get is resolved once per device by cudaGetSymbolAddress with that device current.
// ---------------------------------------------------------------------------------------------
namespace sphb
{
constexpr int64_t kScanStaticTiles = 16384;                        // x kScanTile (4096) = 67 108 864 elements
constexpr int64_t kScanStaticMaxN = kScanStaticTiles * kScanTile;
static __device__ __align__(256) u32 g_scan_static[kScanStaticTiles + 64]; // tile sums [0, n_tiles), total at [n_tiles]
constexpr int kScanStaticMaxDevices = 64;
static u32 *g_scan_static_ptr[kScanStaticMaxDevices] = {};
static std::mutex g_scan_static_mutex;

static sphb200_status scanStaticScratch(sphb200_ctx *ctx, u32 **out)
{
    if (ctx->device < 0 || ctx->device >= kScanStaticMaxDevices)
        return ctx->fail(SPHB200_ERR_INVALID, "exclusive_scan_u32_static: device ordinal out of range");
    std::lock_guard<std::mutex> lock(g_scan_static_mutex);
    if (!g_scan_static_ptr[ctx->device])
    {
        void *p = nullptr;
        SPHB_CUDA(ctx, cudaSetDevice(ctx->device)); // the address is per device: resolve it with that device current
        SPHB_CUDA(ctx, cudaGetSymbolAddress(&p, g_scan_static));
        if (!p)
            return ctx->fail(SPHB200_ERR_CUDA, "exclusive_scan_u32_static: cudaGetSymbolAddress returned a null address");
        g_scan_static_ptr[ctx->device] = static_cast<u32 *>(p);
    }
    *out = g_scan_static_ptr[ctx->device];
    return SPHB200_OK;
}
} // namespace sphb

extern "C"
{
sphb200_status sphb200_exclusive_scan_u32_static(sphb200_ctx *ctx, const uint32_t *d_in, uint32_t *d_out, int64_t n)
{
    if (!ctx || n < 0 || !d_out || (n > 0 && !d_in))
        return SPHB200_ERR_INVALID;
    if (n > kScanStaticMaxN)
        return ctx->fail(SPHB200_ERR_INVALID, "exclusive_scan_u32_static: n = " + std::to_string(n) + " exceeds the static scratch (maximum " +
                                                  std::to_string(kScanStaticMaxN) + " elements); use sphb200_exclusive_scan_u32");
    cudaSetDevice(ctx->device);
    u32 *ts = nullptr;
    SPHB_TRY(scanStaticScratch(ctx, &ts));
    if (n == 0)
    { // total of nothing; also the way to resolve the scratch address ahead of a graph capture
        SPHB_CUDA(ctx, cudaMemsetAsync(d_out, 0, 4, ctx->stream));
        return SPHB200_OK;
    }
    // nothing in the scratch has to start at zero: k_scan_tiles writes every tile sum [0, n_tiles) and
    // k_scan_tile_sums writes [n_tiles] before anything reads them, so the last call's contents are dead
    const int64_t n_tiles = (n + kScanTile - 1) / kScanTile;
    k_scan_tiles<<<int(n_tiles), kScanBlock, 0, ctx->stream>>>(d_in, d_out, ts, n);
    SPHB_LAUNCH_CHECK(ctx);
    k_scan_tile_sums<<<1, kScanBlock, 0, ctx->stream>>>(ts, n_tiles);
    SPHB_LAUNCH_CHECK(ctx);
    k_scan_add<<<gridFor(n, 256), 256, 0, ctx->stream>>>(d_out, ts, n, n_tiles, 1);
    SPHB_LAUNCH_CHECK(ctx);
    return SPHB200_OK;
}
} // extern "C"
