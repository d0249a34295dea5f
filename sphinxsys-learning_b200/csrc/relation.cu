// InnerRelation / ContactRelation: neighbour-index lists built from the cell-sorted bodies.
//
// Reference behaviour reproduced:
//   meshes/cell_linked_list.hpp:20-46              searchNeighborsByMesh (3^d cells, clamped)
//   particle_neighborhood/neighborhood.cpp:85-98   NeighborBuilderInner  (d.d < rc^2 && i != j)
//   particle_neighborhood/neighborhood.cpp:147-160 NeighborBuilderContact (larger-h kernel, no i != j)
//   general_dynamics/domian_bouding/domain_bounding.cpp:18-65  periodic ghost entries x_j ± L
//   shared_ck/.../update_body_relation.hpp:62-164  count → scan → fill (index-only list)
//
// The accepted neighbour SET per particle equals the reference's: the cut-off test is evaluated on
// the un-normalised displacement with separately rounded products and left-to-right sums
// (sqnormExact), on the positions frozen by cell_list_update.  Only indices are stored (u32, SELL-32
// layout so the 32 lanes of a warp read 128 contiguous bytes per neighbour slot); W_ij, dW_ij, e_ij,
// r_ij are recomputed from the frozen positions by the interaction kernels, which trades the
// reference's 56 B/pair of list traffic for ~15 flops.
#include "internal.cuh"

using namespace sphb;

namespace sphb
{
template <int D, class R>
struct SearchP
{
    MeshP<D, R> mesh;      // target body's mesh
    PeriodicP<D, R> per;   // target body's periodic box
    R rc2;
    R rc;
    int depth;
    int inner;
};

// Calls f(j, xj_shifted) for every candidate entry of the 3^d (depth) cell block around xi,
// including periodic images, in a deterministic order.
template <int D, class R, class F>
__device__ __forceinline__ void forEachCandidate(const R *xi, const SearchP<D, R> &sp,
                                                 const u32 *__restrict__ cell_offset,
                                                 const typename Real4<R>::type *__restrict__ xV_t, F &&f)
{
    int c[D];
    sp.mesh.cellIndex(xi, c);
    if (sp.per.mask == 0)
    {
        // non-periodic fast path: the cells of a run along the fastest axis are contiguous
        int lo[D], hi[D];
#pragma unroll
        for (int a = 0; a < D; ++a)
        {
            lo[a] = max(c[a] - sp.depth, 0);
            hi[a] = min(c[a] + sp.depth, sp.mesh.cells[a] - 1);
        }
        int cc[D];
        for (cc[0] = lo[0]; cc[0] <= hi[0]; ++cc[0])
        {
            int y0 = D == 3 ? lo[1] : 0, y1 = D == 3 ? hi[1] : 0;
            for (int cy = y0; cy <= y1; ++cy)
            {
                if (D == 3)
                    cc[1] = cy;
                cc[D - 1] = lo[D - 1];
                u32 first = sp.mesh.linear(cc);
                u32 b = __ldg(cell_offset + first);
                u32 e = __ldg(cell_offset + first + u32(hi[D - 1] - lo[D - 1]) + 1);
                for (u32 j = b; j < e; ++j)
                {
                    typename Real4<R>::type v = ldg4(xV_t + j);
                    R xj[D];
                    xj[0] = v.x;
                    xj[1] = v.y;
                    if (D == 3)
                        xj[D - 1] = v.z;
                    f(j, xj);
                }
            }
        }
        return;
    }
    // periodic path (small bodies only: Taylor-Green class cases).  Per axis the candidate cells are
    // the 3 around c plus, when xi is within rc of a periodic face, the cells that hold the
    // particles whose ghost copy the reference would have placed next to xi.
    int cell_lo[D][2], cell_hi[D][2], shift[D][2], nr[D];
#pragma unroll
    for (int a = 0; a < D; ++a)
    {
        cell_lo[a][0] = max(c[a] - sp.depth, 0);
        cell_hi[a][0] = min(c[a] + sp.depth, sp.mesh.cells[a] - 1);
        shift[a][0] = 0;
        nr[a] = 1;
        if (sp.per.mask & (1 << a))
        {
            R p[D];
#pragma unroll
            for (int q = 0; q < D; ++q)
                p[q] = xi[q];
            int t[D];
            if (xi[a] + sp.rc > sp.per.hi[a])
            { // images x_j + L of particles with lo < x_j < lo + rc
                p[a] = sp.per.lo[a];
                sp.mesh.cellIndex(p, t);
                int l = t[a];
                p[a] = sp.per.lo[a] + sp.rc;
                sp.mesh.cellIndex(p, t);
                cell_lo[a][1] = l;
                cell_hi[a][1] = t[a];
                shift[a][1] = 1;
                nr[a] = 2;
            }
            else if (xi[a] - sp.rc < sp.per.lo[a])
            { // images x_j - L of particles with hi - rc < x_j < hi
                p[a] = sp.per.hi[a] - sp.rc;
                sp.mesh.cellIndex(p, t);
                int l = t[a];
                p[a] = sp.per.hi[a];
                sp.mesh.cellIndex(p, t);
                cell_lo[a][1] = l;
                cell_hi[a][1] = t[a];
                shift[a][1] = -1;
                nr[a] = 2;
            }
        }
    }
    const int nz = D == 3 ? nr[D - 1] : 1;
    for (int r0 = 0; r0 < nr[0]; ++r0)
        for (int r1 = 0; r1 < nr[1]; ++r1)
            for (int r2 = 0; r2 < nz; ++r2)
            {
                int rr[3] = {r0, r1, r2};
                int cc[D];
                for (cc[0] = cell_lo[0][rr[0]]; cc[0] <= cell_hi[0][rr[0]]; ++cc[0])
                    for (cc[1] = cell_lo[1][rr[1]]; cc[1] <= cell_hi[1][rr[1]]; ++cc[1])
                    {
                        int z0 = D == 3 ? cell_lo[D - 1][rr[D - 1]] : 0, z1 = D == 3 ? cell_hi[D - 1][rr[D - 1]] : 0;
                        for (int cz = z0; cz <= z1; ++cz)
                        {
                            if (D == 3)
                                cc[D - 1] = cz;
                            u32 cell = sp.mesh.linear(cc);
                            u32 b = __ldg(cell_offset + cell), e = __ldg(cell_offset + cell + 1);
                            for (u32 j = b; j < e; ++j)
                            {
                                typename Real4<R>::type v = ldg4(xV_t + j);
                                R xj[D];
                                xj[0] = v.x;
                                xj[1] = v.y;
                                if (D == 3)
                                    xj[D - 1] = v.z;
                                bool ok = true;
#pragma unroll
                                for (int a = 0; a < D; ++a)
                                {
                                    int s = shift[a][rr[a]];
                                    if (s > 0)
                                    { // ghost exists iff lo < x_j < lo + rc (domain_bounding.cpp:36-44)
                                        ok = ok && (xj[a] > sp.per.lo[a]) && (xj[a] < (sp.per.lo[a] + sp.rc));
                                        xj[a] = xj[a] + sp.per.L[a];
                                    }
                                    else if (s < 0)
                                    { // ghost exists iff hi - rc < x_j < hi (domain_bounding.cpp:54-62)
                                        ok = ok && (xj[a] < sp.per.hi[a]) && (xj[a] > (sp.per.hi[a] - sp.rc));
                                        xj[a] = xj[a] - sp.per.L[a];
                                    }
                                }
                                if (ok)
                                    f(j, xj);
                            }
                        }
                    }
            }
}

template <int D, class R, bool FILL>
__global__ void __launch_bounds__(128) k_search(int64_t n_src, const typename Real4<R>::type *__restrict__ xV_s,
                                                const typename Real4<R>::type *__restrict__ xV_t,
                                                const u32 *__restrict__ cell_offset, SearchP<D, R> sp,
                                                u32 *__restrict__ count, u32 *__restrict__ slice_w,
                                                const u32 *__restrict__ slice_off, u32 *__restrict__ list)
{
    int64_t i = blockIdx.x * int64_t(blockDim.x) + threadIdx.x;
    const int lane = threadIdx.x & 31;
    u32 cnt = 0;
    if (i < n_src)
    {
        typename Real4<R>::type v = ldg4(xV_s + i);
        R xi[D];
        xi[0] = v.x;
        xi[1] = v.y;
        if (D == 3)
            xi[D - 1] = v.z;
        u32 *slot0 = nullptr;
        if (FILL)
            slot0 = listSlot(list, slice_off, i, 0);
        forEachCandidate<D, R>(xi, sp, cell_offset, xV_t,
                               [&](u32 j, const R *xj)
                               {
                                   R d[D];
#pragma unroll
                                   for (int a = 0; a < D; ++a)
                                       d[a] = xi[a] - xj[a];
                                   R m = sqnormExact<D, R>(d);
                                   if (m < sp.rc2 && (!sp.inner || j != u32(i)))
                                   {
                                       if (FILL)
                                           slot0[size_t(cnt / kS) * 32u + (cnt % kS)] = j;
                                       ++cnt;
                                   }
                               });
    }
    if (!FILL)
    {
        if (i < n_src)
            count[i] = cnt;
        u32 w = (cnt + kS - 1) / kS; // steps the kS lanes of this particle need
#pragma unroll
        for (int o = kP / 2; o > 0; o >>= 1)
            w = max(w, __shfl_xor_sync(0xffffffffu, w, o));
        if ((lane % kP) == 0 && i < n_src)
            slice_w[i / kP] = w;
    }
}

// ---------------------------------------------------------------------------------------------
// Fast path (non-periodic target, search depth 1): ONE distance pass.
//   k_search_mask : per particle, per candidate run (the <= 3 cells along the fastest axis of each of the
//                   3^(D-1) neighbouring cell columns are contiguous particles) a hit bit per candidate,
//                   kRunWords 32-bit words per run; also count[i] and the slice widths.
//   k_expand_mask : after the scan of the slice widths, turns the hit bits into the index list.
// The second pass of the generic path (another ~474 distance tests per particle in 3-D) is replaced by
// ~80 bit extractions.  A run longer than 32*kRunWords candidates raises the overflow flag and the host
// falls back to the generic two-pass search.
// ---------------------------------------------------------------------------------------------
constexpr int kRunWords = 4;
template <int D>
struct MaskLayout
{
    static constexpr int runs = D == 3 ? 9 : 3;
    static constexpr int words = runs * kRunWords;
};

template <int D, class R>
__device__ __forceinline__ void runRange(const int *c, int r, const MeshP<D, R> &mesh, const u32 *__restrict__ cell_offset,
                                         u32 &b, u32 &e)
{
    // run r = (dx, dy) column offset in 3-D, dx in 2-D; empty when the column is outside the mesh
    int cc[D];
    bool inside = true;
    if (D == 3)
    {
        cc[0] = c[0] + r / 3 - 1;
        cc[1] = c[1] + r % 3 - 1;
        inside = cc[0] >= 0 && cc[0] < mesh.cells[0] && cc[1] >= 0 && cc[1] < mesh.cells[1];
    }
    else
    {
        cc[0] = c[0] + r - 1;
        inside = cc[0] >= 0 && cc[0] < mesh.cells[0];
    }
    if (!inside)
    {
        b = e = 0;
        return;
    }
    int lo = max(c[D - 1] - 1, 0), hi = min(c[D - 1] + 1, mesh.cells[D - 1] - 1);
    cc[D - 1] = lo;
    u32 first = mesh.linear(cc);
    b = __ldg(cell_offset + first);
    e = __ldg(cell_offset + first + u32(hi - lo) + 1);
}

template <int D, class R, bool INNER>
__global__ void __launch_bounds__(128) k_search_mask(int64_t n_src, const typename Real4<R>::type *__restrict__ xV_s,
                                                     const typename Real4<R>::type *__restrict__ xV_t,
                                                     const u32 *__restrict__ cell_offset, MeshP<D, R> mesh, R rc2,
                                                     u32 *__restrict__ count, u32 *__restrict__ slice_w,
                                                     u32 *__restrict__ mask, int *__restrict__ overflow)
{
    constexpr int W = MaskLayout<D>::words;
    int64_t i = blockIdx.x * int64_t(blockDim.x) + threadIdx.x;
    const int lane = threadIdx.x & 31;
    u32 cnt = 0;
    if (i < n_src)
    {
        typename Real4<R>::type v = ldg4(xV_s + i);
        R xi[D];
        xi[0] = v.x;
        xi[1] = v.y;
        if (D == 3)
            xi[D - 1] = v.z;
        int c[D];
        mesh.cellIndex(xi, c);
        u32 *mrow = mask + (u64(i >> 5) * W) * 32u + lane;
        for (int r = 0; r < MaskLayout<D>::runs; ++r)
        {
            u32 b, e;
            runRange<D, R>(c, r, mesh, cell_offset, b, e);
            if (e - b > 32u * kRunWords)
            {
                *overflow = 1;
                e = b + 32u * kRunWords;
            }
#pragma unroll
            for (int w = 0; w < kRunWords; ++w)
            {
                u32 word = 0;
                const u32 j0 = b + 32u * w;
                if (j0 < e)
                {
                    const u32 m = min(e - j0, 32u);
#pragma unroll 4
                    for (u32 t = 0; t < m; ++t)
                    {
                        typename Real4<R>::type q = ldg4(xV_t + (j0 + t));
                        R d[D];
                        d[0] = xi[0] - q.x;
                        d[1] = xi[1] - q.y;
                        if (D == 3)
                            d[D - 1] = xi[D - 1] - q.z;
                        R mm = sqnormExact<D, R>(d);
                        word |= (mm < rc2 ? 1u : 0u) << t;
                    }
                    if (INNER)
                    { // NeighborBuilderInner: i != j (neighborhood.cpp:85-98)
                        u32 self = u32(i) - j0;
                        if (self < 32u)
                            word &= ~(1u << self);
                    }
                }
                mrow[u64(r * kRunWords + w) * 32u] = word;
                cnt += __popc(word);
            }
        }
        count[i] = cnt;
    }
    u32 w = (cnt + kS - 1) / kS;
#pragma unroll
    for (int o = kP / 2; o > 0; o >>= 1)
        w = max(w, __shfl_xor_sync(0xffffffffu, w, o));
    if ((lane % kP) == 0 && i < n_src)
        slice_w[i / kP] = w;
}

// Rows of the sliced-ELL list are assembled in shared memory (each thread drops its particle's indices
// into its own columns) and written out as whole 128-byte rows.  kExpandRows rows per slice are staged;
// the rare particle with more than kExpandRows*kS neighbours writes its tail straight to global memory.
#ifndef SPHB_EXPAND_ROWS
#define SPHB_EXPAND_ROWS 44
#endif
constexpr int kExpandRows = SPHB_EXPAND_ROWS;
constexpr int kExpandBlock = 128;
constexpr size_t kExpandSmem = size_t(kExpandBlock / 32) * kS * kExpandRows * 32 * sizeof(u32);

template <int D, class R>
__global__ void __launch_bounds__(kExpandBlock) k_expand_mask(int64_t n_src, const typename Real4<R>::type *__restrict__ xV_s,
                                                              const u32 *__restrict__ cell_offset, MeshP<D, R> mesh,
                                                              const u32 *__restrict__ mask, const u32 *__restrict__ slice_off,
                                                              u32 *__restrict__ list)
{
    extern __shared__ u32 tile_all[];
    constexpr int W = MaskLayout<D>::words;
    int64_t i = blockIdx.x * int64_t(blockDim.x) + threadIdx.x;
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    // this warp's 32 particles = kS slices of kP particles; tile[s][row][32]
    u32 *tile = tile_all + size_t(wid) * kS * kExpandRows * 32;
    u32 *my = tile + size_t(lane / kP) * kExpandRows * 32 + (lane % kP) * kS;
    if (i < n_src)
    {
        typename Real4<R>::type v = ldg4(xV_s + i);
        R xi[D];
        xi[0] = v.x;
        xi[1] = v.y;
        if (D == 3)
            xi[D - 1] = v.z;
        int c[D];
        mesh.cellIndex(xi, c);
        const u32 *mrow = mask + (u64(i >> 5) * W) * 32u + lane;
        u32 *slot0 = listSlot(list, slice_off, i, 0);
        u32 cnt = 0;
        u32 rb[MaskLayout<D>::runs], words[W];
#pragma unroll
        for (int q = 0; q < W; ++q) // independent loads, all in flight at once
            words[q] = __ldg(mrow + u64(q) * 32u);
#pragma unroll
        for (int r = 0; r < MaskLayout<D>::runs; ++r)
        {
            u32 e_unused;
            runRange<D, R>(c, r, mesh, cell_offset, rb[r], e_unused);
        }
#pragma unroll
        for (int r = 0; r < MaskLayout<D>::runs; ++r)
        {
#pragma unroll
            for (int w = 0; w < kRunWords; ++w)
            {
                u32 word = words[r * kRunWords + w];
                const u32 j0 = rb[r] + 32u * w;
                while (word)
                {
                    u32 t = __ffs(word) - 1;
                    word &= word - 1;
                    const u32 row = cnt / kS, col = cnt % kS;
                    if (row < u32(kExpandRows))
                        my[row * 32 + col] = j0 + t;
                    else
                        slot0[size_t(row) * 32u + col] = j0 + t;
                    ++cnt;
                }
            }
        }
    }
    __syncwarp();
    // copy out: slice s of this warp covers particles [i_warp + s*kP, +kP)
    const int64_t i_warp = i - lane;
#pragma unroll
    for (int sl = 0; sl < kS; ++sl)
    {
        const int64_t first = i_warp + int64_t(sl) * kP;
        if (first >= n_src)
            break;
        const u32 so = slice_off[first / kP];
        const u32 rows = min(slice_off[first / kP + 1] - so, u32(kExpandRows));
        const u32 *src = tile + size_t(sl) * kExpandRows * 32 + lane;
        u32 *dst = list + u64(so) * 32u + lane;
        for (u32 row = 0; row < rows; ++row)
            dst[size_t(row) * 32u] = src[row * 32];
    }
}

__global__ void k_sum_u32(const u32 *__restrict__ v, int64_t n, unsigned long long *out)
{
    int64_t i = blockIdx.x * int64_t(blockDim.x) + threadIdx.x;
    unsigned long long s = 0;
    for (; i < n; i += int64_t(gridDim.x) * blockDim.x)
        s += v[i];
    s = warpSum<unsigned long long>(s);
    if ((threadIdx.x & 31) == 0 && s)
        atomicAdd(out, s);
}

template <int D, class R>
static sphb200_status relationUpdate(sphb200_relation *r)
{
    sphb200_ctx *ctx = r->ctx;
    sphb200_body *src = r->body;
    cudaStream_t st = ctx->stream;
    if (!src->cell_list_valid)
        return ctx->fail(SPHB200_ERR_INVALID, "relation_update: source body '" + src->name + "' has no cell list yet");
    const int64_t n = src->n;
    const int64_t n_slices = (n + kP - 1) / kP;
    BodyP<D, R> &Ps = src->P<D, R>();
    for (size_t k = 0; k < r->targets.size(); ++k)
    {
        sphb200_body *tgt = r->targets[k];
        if (!tgt->cell_list_valid)
            return ctx->fail(SPHB200_ERR_INVALID, "relation_update: target body '" + tgt->name + "' has no cell list yet");
        BodyP<D, R> &Pt = tgt->P<D, R>();
        NeighborList &L = r->lists[k];
        if (L.cap_src < size_t(n) || !L.count)
        {
            cudaFree(L.count);
            cudaFree(L.slice_w);
            cudaFree(L.slice_off);
            L.count = L.slice_w = L.slice_off = nullptr;
            size_t cap = size_t(n) + size_t(n) / 8 + 64;
            SPHB_CUDA(ctx, cudaMalloc(&L.count, cap * 4));
            SPHB_CUDA(ctx, cudaMalloc(&L.slice_w, (cap / kP + 2) * 4));
            SPHB_CUDA(ctx, cudaMalloc(&L.slice_off, (cap / kP + 3) * 4));
            L.cap_src = cap;
        }
        SearchP<D, R> sp;
        sp.mesh = Pt.mesh;
        sp.per = Pt.per;
        sp.inner = r->inner ? 1 : 0;
        // neighborhood.cpp:75-80 chooseKernel: the kernel with the larger smoothing length
        const KernelP<R> &kern = (r->inner || Ps.h > Pt.h) ? Ps.kern : Pt.kern;
        sp.rc = kern.rc;
        sp.rc2 = kern.rc2;
        // smallest depth whose cell block covers the cut-off sphere; the reference searches
        // 1+floor(rc/spacing) cells (base_body_relation.cpp:10-16), a superset with the same accepted set
        sp.depth = 1;
        while (R(sp.depth) * Pt.mesh.spacing < sp.rc)
            sp.depth++;
        L.total_entries = L.total_pairs = 0;
        if (n == 0)
        {
            r->target_stamp[k] = tgt->config_stamp;
            L.build_id++;
            continue;
        }
        auto *xs = static_cast<const typename Real4<R>::type *>(src->xV.p);
        auto *xt = static_cast<const typename Real4<R>::type *>(tgt->xV.p);
        const u32 *off = static_cast<const u32 *>(tgt->cell_offset.p);
        unsigned long long *acc = static_cast<unsigned long long *>(ctx->reduce_dev.p) + 8;
        int *overflow = reinterpret_cast<int *>(acc + 1);
        bool fast = Pt.per.mask == 0 && sp.depth == 1;
        u32 *mask = nullptr;
        for (int attempt = 0; attempt < 2; ++attempt)
        {
            SPHB_CUDA(ctx, cudaMemsetAsync(acc, 0, 16, st));
            if (fast)
            {
                const size_t words = size_t((n + 31) / 32) * 32u * size_t(MaskLayout<D>::words);
                SPHB_CUDA(ctx, ctx->mask_buf.ensure(words * 4));
                mask = static_cast<u32 *>(ctx->mask_buf.p);
                if (r->inner)
                    k_search_mask<D, R, true><<<gridFor(n, 128), 128, 0, st>>>(n, xs, xt, off, Pt.mesh, sp.rc2, L.count, L.slice_w, mask, overflow);
                else
                    k_search_mask<D, R, false><<<gridFor(n, 128), 128, 0, st>>>(n, xs, xt, off, Pt.mesh, sp.rc2, L.count, L.slice_w, mask, overflow);
            }
            else
                k_search<D, R, false><<<gridFor(n, 128), 128, 0, st>>>(n, xs, xt, off, sp, L.count, L.slice_w, nullptr, nullptr);
            SPHB_LAUNCH_CHECK(ctx);
            SPHB_TRY(exclusiveScanU32(ctx, L.slice_w, L.slice_off, n_slices, true));
            k_sum_u32<<<148 * 4, 256, 0, st>>>(L.count, n, acc);
            SPHB_LAUNCH_CHECK(ctx);
            u32 total_w = 0;
            unsigned long long host_acc[2] = {0, 0};
            SPHB_CUDA(ctx, cudaMemcpyAsync(&total_w, L.slice_off + n_slices, 4, cudaMemcpyDeviceToHost, st));
            SPHB_CUDA(ctx, cudaMemcpyAsync(host_acc, acc, 16, cudaMemcpyDeviceToHost, st));
            SPHB_CUDA(ctx, cudaStreamSynchronize(st));
            if (fast && host_acc[1] != 0)
            { // a candidate run did not fit the mask words: redo with the generic search
                fast = false;
                continue;
            }
            L.total_entries = u64(total_w) * 32u;
            L.total_pairs = host_acc[0];
            break;
        }
        if (L.total_entries > L.cap_entries || !L.list)
        {
            cudaFree(L.list);
            L.list = nullptr;
            size_t cap = size_t(L.total_entries) + size_t(L.total_entries) / 8 + 1024;
            SPHB_CUDA(ctx, cudaMalloc(&L.list, cap * 4));
            L.cap_entries = cap;
        }
        if (fast)
        {
            static bool attr_set = false; // per (D, R) instantiation
            if (!attr_set)
            {
                SPHB_CUDA(ctx, cudaFuncSetAttribute(k_expand_mask<D, R>, cudaFuncAttributeMaxDynamicSharedMemorySize, int(kExpandSmem)));
                attr_set = true;
            }
            k_expand_mask<D, R><<<gridFor(n, kExpandBlock), kExpandBlock, kExpandSmem, st>>>(n, xs, off, Pt.mesh, mask, L.slice_off, L.list);
        }
        else
            k_search<D, R, true><<<gridFor(n, 128), 128, 0, st>>>(n, xs, xt, off, sp, L.count, L.slice_w, L.slice_off, L.list);
        SPHB_LAUNCH_CHECK(ctx);
        r->target_stamp[k] = tgt->config_stamp;
        L.build_id++;
        L.sorted = Pt.per.mask == 0; // both non-periodic searches emit ascending j per particle
    }
    r->source_stamp = src->config_stamp;
    r->valid = true;
    return SPHB200_OK;
}
} // namespace sphb

extern "C"
{
sphb200_status sphb200_relation_inner_create(sphb200_body *b, sphb200_relation **out)
{
    if (!b || !out)
        return SPHB200_ERR_INVALID;
    auto *r = new sphb200_relation();
    r->ctx = b->ctx;
    r->body = b;
    r->inner = true;
    r->targets.push_back(b);
    r->lists.resize(1);
    r->target_stamp.assign(1, 0);
    b->ctx->relations.push_back(r);
    *out = r;
    return SPHB200_OK;
}

sphb200_status sphb200_relation_contact_create(sphb200_body *b, sphb200_body *const *targets, int n_targets,
                                               sphb200_relation **out)
{
    if (!b || !out || !targets || n_targets < 1 || n_targets > kMaxTargets)
        return SPHB200_ERR_INVALID;
    for (int k = 0; k < n_targets; ++k)
        if (!targets[k] || targets[k]->ctx != b->ctx)
            return b->ctx->fail(SPHB200_ERR_INVALID, "relation_contact_create: target from another context");
    auto *r = new sphb200_relation();
    r->ctx = b->ctx;
    r->body = b;
    r->inner = false;
    for (int k = 0; k < n_targets; ++k)
        r->targets.push_back(targets[k]);
    r->lists.resize(size_t(n_targets));
    r->target_stamp.assign(size_t(n_targets), 0);
    b->ctx->relations.push_back(r);
    *out = r;
    return SPHB200_OK;
}

sphb200_status sphb200_relation_update(sphb200_relation *r)
{
    if (!r)
        return SPHB200_ERR_INVALID;
    cudaSetDevice(r->ctx->device);
    return SPHB_DISPATCH(r->ctx, relationUpdate, r);
}

sphb200_status sphb200_relation_neighbors(sphb200_relation *r, int k, int32_t *counts, int64_t *j_out, int64_t *total)
{
    if (!r || k < 0 || k >= int(r->lists.size()))
        return SPHB200_ERR_INVALID;
    sphb200_ctx *ctx = r->ctx;
    if (!r->valid)
        return ctx->fail(SPHB200_ERR_INVALID, "relation_neighbors: relation has not been updated");
    cudaSetDevice(ctx->device);
    NeighborList &L = r->lists[size_t(k)];
    const int64_t n = r->body->n;
    if (total)
        *total = int64_t(L.total_pairs);
    if (!counts && !j_out)
        return SPHB200_OK;
    std::vector<u32> c, so, lst;
    c.resize(size_t(n));
    so.resize(size_t((n + kP - 1) / kP + 1));
    if (n > 0)
    {
        SPHB_CUDA(ctx, cudaMemcpyAsync(c.data(), L.count, size_t(n) * 4, cudaMemcpyDeviceToHost, ctx->stream));
        SPHB_CUDA(ctx, cudaMemcpyAsync(so.data(), L.slice_off, so.size() * 4, cudaMemcpyDeviceToHost, ctx->stream));
        if (j_out && L.total_entries)
        {
            lst.resize(size_t(L.total_entries));
            SPHB_CUDA(ctx, cudaMemcpyAsync(lst.data(), L.list, lst.size() * 4, cudaMemcpyDeviceToHost, ctx->stream));
        }
        SPHB_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    }
    int64_t w = 0;
    for (int64_t i = 0; i < n; ++i)
    {
        if (counts)
            counts[i] = int32_t(c[size_t(i)]);
        if (j_out)
        {
            u64 base = u64(so[size_t(i / kP)]) * 32u + u64(i % kP) * u64(kS);
            for (u32 m = 0; m < c[size_t(i)]; ++m)
                j_out[w + m] = int64_t(lst[size_t(base + u64(m / kS) * 32u + (m % kS))]);
        }
        w += c[size_t(i)];
    }
    return SPHB200_OK;
}
} // extern "C"
