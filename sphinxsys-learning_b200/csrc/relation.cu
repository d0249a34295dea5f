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
        u64 base = 0;
        if (FILL)
            base = u64(slice_off[i >> 5]) * 32u + u32(lane);
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
                                           list[base + u64(cnt) * 32u] = j;
                                       ++cnt;
                                   }
                               });
    }
    if (!FILL)
    {
        if (i < n_src)
            count[i] = cnt;
        u32 w = __reduce_max_sync(0xffffffffu, cnt);
        if (lane == 0 && (i >> 5) < ((n_src + 31) >> 5))
            slice_w[i >> 5] = w;
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
    const int64_t n_slices = (n + 31) / 32;
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
            SPHB_CUDA(ctx, cudaMalloc(&L.slice_w, (cap / 32 + 2) * 4));
            SPHB_CUDA(ctx, cudaMalloc(&L.slice_off, (cap / 32 + 3) * 4));
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
            continue;
        }
        auto *xs = static_cast<const typename Real4<R>::type *>(src->xV.p);
        auto *xt = static_cast<const typename Real4<R>::type *>(tgt->xV.p);
        const u32 *off = static_cast<const u32 *>(tgt->cell_offset.p);
        k_search<D, R, false><<<gridFor(n, 128), 128, 0, st>>>(n, xs, xt, off, sp, L.count, L.slice_w, nullptr, nullptr);
        SPHB_LAUNCH_CHECK(ctx);
        SPHB_TRY(exclusiveScanU32(ctx, L.slice_w, L.slice_off, n_slices, true));
        unsigned long long *acc = static_cast<unsigned long long *>(ctx->reduce_dev.p) + 8;
        SPHB_CUDA(ctx, cudaMemsetAsync(acc, 0, 8, st));
        k_sum_u32<<<148 * 4, 256, 0, st>>>(L.count, n, acc);
        SPHB_LAUNCH_CHECK(ctx);
        u32 total_w = 0;
        unsigned long long pairs = 0;
        SPHB_CUDA(ctx, cudaMemcpyAsync(&total_w, L.slice_off + n_slices, 4, cudaMemcpyDeviceToHost, st));
        SPHB_CUDA(ctx, cudaMemcpyAsync(&pairs, acc, 8, cudaMemcpyDeviceToHost, st));
        SPHB_CUDA(ctx, cudaStreamSynchronize(st));
        L.total_entries = u64(total_w) * 32u;
        L.total_pairs = pairs;
        if (L.total_entries > L.cap_entries || !L.list)
        {
            cudaFree(L.list);
            L.list = nullptr;
            size_t cap = size_t(L.total_entries) + size_t(L.total_entries) / 8 + 1024;
            SPHB_CUDA(ctx, cudaMalloc(&L.list, cap * 4));
            L.cap_entries = cap;
        }
        k_search<D, R, true><<<gridFor(n, 128), 128, 0, st>>>(n, xs, xt, off, sp, L.count, L.slice_w, L.slice_off, L.list);
        SPHB_LAUNCH_CHECK(ctx);
        r->target_stamp[k] = tgt->config_stamp;
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
    so.resize(size_t((n + 31) / 32 + 1));
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
            u64 base = u64(so[size_t(i >> 5)]) * 32u + u64(i & 31);
            for (u32 m = 0; m < c[size_t(i)]; ++m)
                j_out[w + m] = int64_t(lst[size_t(base + u64(m) * 32u)]);
        }
        w += c[size_t(i)];
    }
    return SPHB200_OK;
}
} // extern "C"
