// Slab-decomposition support of the C ABI (multi-GPU; the reference has no equivalent).
// The exchange itself (NCCL send/recv of migrating and halo particles) is driven by the host layer
// through torch.distributed on raw device pointers; the library only has to resize a body, know which
// contiguous run of the cell-sorted particles this rank owns, and share the caller's stream.
#include "internal.cuh"
#include <algorithm>
#include <cstdlib>
#include <vector>

using namespace sphb;

namespace sphb
{
__global__ void k_iota2(u32 *a, u32 *b, int64_t n)
{
    int64_t i = blockIdx.x * int64_t(blockDim.x) + threadIdx.x;
    if (i < n)
    {
        a[i] = u32(i);
        b[i] = u32(i);
    }
}

// ---- re-decomposition after the particles have moved (one advection step): which owned rows stay, which go to
// the left / right neighbour slab (as migrants or as that slab's fresh halo).  The page of a particle is the axis-0
// cell index of its CURRENT position, evaluated exactly as the cell-list update will (MeshP::cellIndex).
template <int D, class R>
__global__ void k_slab_flags(int64_t n, int64_t ob, int64_t oe, const R *__restrict__ pos, MeshP<D, R> mesh, int p_lo, int p_hi,
                             int has_left, int has_right, u32 *__restrict__ keep, u32 *__restrict__ left, u32 *__restrict__ right)
{
    int64_t i = blockIdx.x * int64_t(blockDim.x) + threadIdx.x;
    if (i >= n)
        return;
    u32 k = 0, l = 0, r = 0;
    if (i >= ob && i < oe)
    {
        int c[D];
        mesh.cellIndex(pos + i * D, c);
        const int px = c[0];
        // kept: my pages AND the two halo pages.  A row that leaves through a cut is sent to that neighbour as a migrant
        // and stays here as a halo copy (the neighbour only sends rows IT owned, so nothing arrives twice); with the
        // owned pages changed before this call (re-cut) the same rule moves whole pages.
        k = (px >= p_lo - 1 && px < p_hi + 1) ? 1u : 0u;
        l = (has_left && px <= p_lo) ? 1u : 0u;
        r = (has_right && px >= p_hi - 1) ? 1u : 0u;
    }
    keep[i] = k;
    left[i] = l;
    right[i] = r;
}
__global__ void k_compact_idx(int64_t n, const u32 *__restrict__ flag, const u32 *__restrict__ offset, u32 *__restrict__ out)
{
    int64_t i = blockIdx.x * int64_t(blockDim.x) + threadIdx.x;
    if (i < n && flag[i])
        out[offset[i]] = u32(i);
}
// dst[r * dst_stride + dst_col + c] = src[idx[r] * width + c]   (element = 4 or 8 bytes, moved verbatim)
template <class T>
__global__ void k_gather_cols(int64_t rows, int width, const u32 *__restrict__ idx, const T *__restrict__ src, T *__restrict__ dst,
                              int dst_stride, int dst_col)
{
    int64_t e = blockIdx.x * int64_t(blockDim.x) + threadIdx.x;
    if (e >= rows * width)
        return;
    const int64_t r = e / width;
    const int c = int(e - r * width);
    dst[r * dst_stride + dst_col + c] = src[int64_t(idx[r]) * width + c];
}
// dst[(row0 + r) * width + c] = src[r * src_stride + src_col + c]
template <class T>
__global__ void k_place_cols(int64_t rows, int width, const T *__restrict__ src, int src_stride, int src_col, T *__restrict__ dst, int64_t row0)
{
    int64_t e = blockIdx.x * int64_t(blockDim.x) + threadIdx.x;
    if (e >= rows * width)
        return;
    const int64_t r = e / width;
    const int c = int(e - r * width);
    dst[(row0 + r) * width + c] = src[r * src_stride + src_col + c];
}
// halo records: dst[r] = src[idx[r]] / dst[idx[r]] = src[r], rows of row_bytes (a multiple of 16) moved as uint4
__global__ void k_rows_gather(int64_t rows, int chunks, const int64_t *__restrict__ idx, const uint4 *__restrict__ src, uint4 *__restrict__ dst)
{
    int64_t e = blockIdx.x * int64_t(blockDim.x) + threadIdx.x;
    if (e >= rows * chunks)
        return;
    const int64_t r = e / chunks;
    const int c = int(e - r * chunks);
    dst[r * chunks + c] = src[idx[r] * chunks + c];
}
__global__ void k_rows_scatter(int64_t rows, int chunks, const int64_t *__restrict__ idx, const uint4 *__restrict__ src, uint4 *__restrict__ dst)
{
    int64_t e = blockIdx.x * int64_t(blockDim.x) + threadIdx.x;
    if (e >= rows * chunks)
        return;
    const int64_t r = e / chunks;
    const int c = int(e - r * chunks);
    dst[idx[r] * chunks + c] = src[r * chunks + c];
}

template <int D, class R>
static sphb200_status slabSelect(sphb200_body *b, int has_left, int has_right, int64_t *counts3)
{
    sphb200_ctx *ctx = b->ctx;
    cudaStream_t st = ctx->stream;
    const int64_t n = b->n;
    if (b->page_lo < 0)
        return ctx->fail(SPHB200_ERR_INVALID, "slab_select: body has no owned pages (body_set_owned_pages)");
    Var *pos = b->find("Position");
    b->slab_count[0] = b->slab_count[1] = b->slab_count[2] = 0;
    if (n > 0)
    {
        const size_t stride = size_t(n) + 1;
        SPHB_CUDA(ctx, b->slab_flags.ensure(stride * 4 * 6));
        u32 *flags = static_cast<u32 *>(b->slab_flags.p);
        u32 *f[3] = {flags, flags + stride, flags + 2 * stride};
        u32 *o[3] = {flags + 3 * stride, flags + 4 * stride, flags + 5 * stride};
        k_slab_flags<D, R><<<gridFor(n, 256), 256, 0, st>>>(n, b->own_begin, b->own_end, static_cast<const R *>(pos->dev.p), b->P<D, R>().mesh,
                                                            b->page_lo, b->page_hi, has_left, has_right, f[0], f[1], f[2]);
        SPHB_LAUNCH_CHECK(ctx);
        u32 totals[3] = {0, 0, 0};
        for (int k = 0; k < 3; ++k)
        {
            SPHB_TRY(exclusiveScanU32(ctx, f[k], o[k], n, true));
            SPHB_CUDA(ctx, cudaMemcpyAsync(&totals[k], o[k] + n, 4, cudaMemcpyDeviceToHost, st));
        }
        SPHB_CUDA(ctx, cudaStreamSynchronize(st));
        for (int k = 0; k < 3; ++k)
        {
            b->slab_count[k] = totals[k];
            SPHB_CUDA(ctx, b->slab_idx[k].ensure(size_t(totals[k] > 0 ? totals[k] : 1) * 4));
            if (totals[k] > 0)
            {
                k_compact_idx<<<gridFor(n, 256), 256, 0, st>>>(n, f[k], o[k], static_cast<u32 *>(b->slab_idx[k].p));
                SPHB_LAUNCH_CHECK(ctx);
            }
        }
    }
    b->slab_selected = true;
    if (counts3)
        for (int k = 0; k < 3; ++k)
            counts3[k] = b->slab_count[k];
    return SPHB200_OK;
}

static int rowWidth(sphb200_body *b, const char *const *names, int n_names, std::vector<Var *> &vars)
{
    int w = 0;
    for (int k = 0; k < n_names; ++k)
    {
        Var *v = b->find(names[k]);
        if (!v)
            return -1;
        vars.push_back(v);
        w += v->width;
    }
    return w;
}

template <class T>
static sphb200_status slabPackT(sphb200_body *b, int side, const std::vector<Var *> &vars, int W, void *out)
{
    sphb200_ctx *ctx = b->ctx;
    const int64_t rows = b->slab_count[side];
    int col = 0;
    for (Var *v : vars)
    {
        if (rows > 0)
        {
            k_gather_cols<T><<<gridFor(rows * v->width, 256), 256, 0, ctx->stream>>>(rows, v->width, static_cast<const u32 *>(b->slab_idx[side].p),
                                                                                static_cast<const T *>(v->dev.p), static_cast<T *>(out), W, col);
            SPHB_LAUNCH_CHECK(ctx);
        }
        col += v->width;
    }
    return SPHB200_OK;
}

template <class T>
static sphb200_status slabRebuildT(sphb200_body *b, const std::vector<Var *> &vars, int W, const void *from_left, int64_t n_left,
                                   const void *from_right, int64_t n_right)
{
    sphb200_ctx *ctx = b->ctx;
    cudaStream_t st = ctx->stream;
    const int64_t n_keep = b->slab_count[0], n_new = n_keep + n_left + n_right;
    if (n_new >= (int64_t(1) << 31))
        return ctx->fail(SPHB200_ERR_INVALID, "slab_rebuild: too many particles");
    // ---- in-place path: every owned row stays (the usual case: a row leaves only through the domain faces), so the
    // kept rows are the contiguous range [ob, oe) and only the old halo copies either side of it go.  The received rows
    // are written behind oe and the tail of the result is copied into the hole [0, ob): O(halo) bytes moved instead of
    // one gather pass over every variable of the body (2.9 -> 0.3 ms per advection step at 20 M particles per slab).
    // The next cell_list_update sorts the arrangement anyway, so the row order is free.
    const int64_t ob = b->own_begin, oe = b->own_end, n_recv = n_left + n_right;
    bool in_place = n_keep == oe - ob && n_keep > 0 && ob <= n_new;
    if (const char *e = getenv("SPHB200_SLAB_INPLACE")) // 0: always the gather path (A/B timing, tests of both paths)
        in_place = in_place && atoi(e) != 0;
    if (in_place)
        for (auto &kv : b->vars)
            if (size_t(oe + n_recv) * size_t(kv.second->width) * sizeof(T) > kv.second->dev.bytes)
                in_place = false;
    if (in_place)
    {
        for (auto &kv : b->vars)
        {
            Var &v = *kv.second;
            const size_t row = size_t(v.width) * sizeof(T);
            int col = 0;
            bool carried = false;
            for (Var *m : vars)
            {
                if (m == &v)
                {
                    carried = true;
                    if (n_left > 0)
                        k_place_cols<T><<<gridFor(n_left * v.width, 256), 256, 0, st>>>(n_left, v.width, static_cast<const T *>(from_left), W, col,
                                                                                    static_cast<T *>(v.dev.p), oe);
                    if (n_right > 0)
                        k_place_cols<T><<<gridFor(n_right * v.width, 256), 256, 0, st>>>(n_right, v.width, static_cast<const T *>(from_right), W, col,
                                                                                     static_cast<T *>(v.dev.p), oe + n_left);
                    SPHB_LAUNCH_CHECK(ctx);
                    break;
                }
                col += m->width;
            }
            if (!carried && n_recv > 0)
                SPHB_CUDA(ctx, cudaMemsetAsync(static_cast<char *>(v.dev.p) + size_t(oe) * row, 0, size_t(n_recv) * row, st));
            if (ob > 0) // rows [n_new, n_new + ob) -> [0, ob); disjoint because ob <= n_new
                SPHB_CUDA(ctx, cudaMemcpyAsync(v.dev.p, static_cast<char *>(v.dev.p) + size_t(n_new) * row, size_t(ob) * row, cudaMemcpyDeviceToDevice, st));
        }
    }
    else // generic: one gather pass per variable into the twin buffer
        for (auto &kv : b->vars)
    {
        Var &v = *kv.second;
        const size_t row = size_t(v.width) * sizeof(T);
        SPHB_CUDA(ctx, v.alt.ensure(size_t(n_new > 0 ? n_new : 1) * row + size_t(n_new / 8) * row));
        if (n_keep > 0)
        {
            k_gather_cols<T><<<gridFor(n_keep * v.width, 256), 256, 0, st>>>(n_keep, v.width, static_cast<const u32 *>(b->slab_idx[0].p),
                                                                         static_cast<const T *>(v.dev.p), static_cast<T *>(v.alt.p), v.width, 0);
            SPHB_LAUNCH_CHECK(ctx);
        }
        int col = 0;
        bool carried = false;
        for (Var *m : vars)
        {
            if (m == &v)
            {
                carried = true;
                if (n_left > 0)
                    k_place_cols<T><<<gridFor(n_left * v.width, 256), 256, 0, st>>>(n_left, v.width, static_cast<const T *>(from_left), W, col,
                                                                                static_cast<T *>(v.alt.p), n_keep);
                if (n_right > 0)
                    k_place_cols<T><<<gridFor(n_right * v.width, 256), 256, 0, st>>>(n_right, v.width, static_cast<const T *>(from_right), W, col,
                                                                                 static_cast<T *>(v.alt.p), n_keep + n_left);
                SPHB_LAUNCH_CHECK(ctx);
                break;
            }
            col += m->width;
        }
        if (!carried && n_new > n_keep)
            SPHB_CUDA(ctx, cudaMemsetAsync(static_cast<char *>(v.alt.p) + size_t(n_keep) * row, 0, size_t(n_new - n_keep) * row, st));
        std::swap(v.dev, v.alt);
    }
    const size_t idb = size_t(n_new > 0 ? n_new : 1) * 4;
    SPHB_CUDA(ctx, b->original_id.ensure(idb));
    SPHB_CUDA(ctx, b->sorted_id.ensure(idb));
    if (n_new > 0)
    {
        k_iota2<<<gridFor(n_new, 256), 256, 0, st>>>(static_cast<u32 *>(b->original_id.p), static_cast<u32 *>(b->sorted_id.p), n_new);
        SPHB_LAUNCH_CHECK(ctx);
    }
    b->n = n_new;
    b->own_begin = 0;
    b->own_end = n_new;
    b->cell_list_valid = false;
    b->xV_vol_dirty = true;
    b->rec_b_valid = false;
    b->acoustic_max_valid = false;
    b->slab_selected = false;
    b->config_stamp++;
    return SPHB200_OK;
}

// called at the end of cell_list_update: translate owned pages into a particle range
sphb200_status updateOwnedRange(sphb200_body *b)
{
    sphb200_ctx *ctx = b->ctx;
    b->own_begin = 0;
    b->own_end = b->n;
    if (b->page_lo < 0 || b->n == 0)
        return SPHB200_OK;
    // cells per page = product of the cell counts of the faster axes
    int64_t per_page = 1;
    double info[12];
    SPHB_TRY(sphb200_body_info(b, info));
    for (int d = 1; d < ctx->ndim; ++d)
        per_page *= int64_t(info[8 + d]);
    int64_t pages = int64_t(info[8]);
    int64_t lo = std::min<int64_t>(std::max<int64_t>(b->page_lo, 0), pages);
    int64_t hi = std::min<int64_t>(std::max<int64_t>(b->page_hi, lo), pages);
    u32 r[2];
    const u32 *off = static_cast<const u32 *>(b->cell_offset.p);
    SPHB_CUDA(ctx, cudaMemcpyAsync(&r[0], off + lo * per_page, 4, cudaMemcpyDeviceToHost, ctx->stream));
    SPHB_CUDA(ctx, cudaMemcpyAsync(&r[1], off + hi * per_page, 4, cudaMemcpyDeviceToHost, ctx->stream));
    SPHB_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    b->own_begin = r[0];
    b->own_end = r[1];
    return SPHB200_OK;
}
} // namespace sphb

extern "C"
{
sphb200_status sphb200_ctx_set_stream(sphb200_ctx *ctx, void *cuda_stream)
{
    if (!ctx)
        return SPHB200_ERR_INVALID;
    cudaSetDevice(ctx->device);
    SPHB_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    if (ctx->owns_stream && ctx->stream)
        cudaStreamDestroy(ctx->stream);
    ctx->stream = static_cast<cudaStream_t>(cuda_stream);
    ctx->owns_stream = false;
    return SPHB200_OK;
}

sphb200_status sphb200_body_resize(sphb200_body *b, int64_t n_new)
{
    if (!b || n_new < 0 || n_new >= (int64_t(1) << 31))
        return SPHB200_ERR_INVALID;
    sphb200_ctx *ctx = b->ctx;
    cudaSetDevice(ctx->device);
    cudaStream_t st = ctx->stream;
    const int64_t keep = std::min(b->n, n_new);
    for (auto &kv : b->vars)
    {
        Var &v = *kv.second;
        size_t row = size_t(v.width) * size_t(ctx->real_bytes);
        size_t need = size_t(n_new > 0 ? n_new : 1) * row;
        if (need > v.dev.bytes)
        {
            DevBuf nb;
            SPHB_CUDA(ctx, nb.ensure(need));
            if (keep > 0)
                SPHB_CUDA(ctx, cudaMemcpyAsync(nb.p, v.dev.p, size_t(keep) * row, cudaMemcpyDeviceToDevice, st));
            SPHB_CUDA(ctx, cudaStreamSynchronize(st));
            v.dev.release();
            v.dev = nb;
        }
        if (n_new > keep)
            SPHB_CUDA(ctx, cudaMemsetAsync(static_cast<char *>(v.dev.p) + size_t(keep) * row, 0,
                                           size_t(n_new - keep) * row, st));
    }
    size_t idb = size_t(n_new > 0 ? n_new : 1) * 4;
    SPHB_CUDA(ctx, b->original_id.ensure(idb));
    SPHB_CUDA(ctx, b->sorted_id.ensure(idb));
    if (n_new > 0)
    {
        k_iota2<<<gridFor(n_new, 256), 256, 0, st>>>(static_cast<u32 *>(b->original_id.p),
                                                    static_cast<u32 *>(b->sorted_id.p), n_new);
        SPHB_LAUNCH_CHECK(ctx);
    }
    b->n = n_new;
    b->own_begin = 0;
    b->own_end = n_new;
    b->cell_list_valid = false;
    b->xV_vol_dirty = true;
    b->rec_b_valid = false;
    b->acoustic_max_valid = false;
    b->config_stamp++;
    return SPHB200_OK;
}

sphb200_status sphb200_body_keep_range(sphb200_body *b, int64_t begin, int64_t end)
{
    if (!b || begin < 0 || end < begin || end > b->n)
        return SPHB200_ERR_INVALID;
    sphb200_ctx *ctx = b->ctx;
    cudaSetDevice(ctx->device);
    cudaStream_t st = ctx->stream;
    const int64_t keep = end - begin;
    if (begin > 0 && keep > 0)
        for (auto &kv : b->vars)
        {
            Var &v = *kv.second;
            size_t row = size_t(v.width) * size_t(ctx->real_bytes);
            // through the twin buffer: source and destination ranges may overlap
            SPHB_CUDA(ctx, v.alt.ensure(size_t(b->n) * row));
            SPHB_CUDA(ctx, cudaMemcpyAsync(v.alt.p, static_cast<char *>(v.dev.p) + size_t(begin) * row, size_t(keep) * row,
                                           cudaMemcpyDeviceToDevice, st));
            std::swap(v.dev, v.alt);
        }
    return sphb200_body_resize(b, keep);
}

sphb200_status sphb200_body_page_offsets(sphb200_body *b, int64_t *out, int64_t *n_pages)
{
    if (!b || !n_pages)
        return SPHB200_ERR_INVALID;
    sphb200_ctx *ctx = b->ctx;
    cudaSetDevice(ctx->device);
    double info[12];
    SPHB_TRY(sphb200_body_info(b, info));
    int64_t per_page = 1;
    for (int d = 1; d < ctx->ndim; ++d)
        per_page *= int64_t(info[8 + d]);
    const int64_t pages = int64_t(info[8]);
    *n_pages = pages;
    if (!out)
        return SPHB200_OK;
    if (!b->cell_list_valid)
        return ctx->fail(SPHB200_ERR_INVALID, "body_page_offsets: cell_list_update has not been called");
    // one strided copy: element p of the result = cell_offset[p * per_page]
    std::vector<u32> host(size_t(pages) + 1, 0u);
    if (b->n > 0)
    {
        SPHB_CUDA(ctx, cudaMemcpy2DAsync(host.data(), 4, b->cell_offset.p, size_t(per_page) * 4, 4, size_t(pages) + 1,
                                         cudaMemcpyDeviceToHost, ctx->stream));
        SPHB_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    }
    for (int64_t p = 0; p <= pages; ++p)
        out[p] = int64_t(host[size_t(p)]);
    return SPHB200_OK;
}

sphb200_status sphb200_body_page_begin(sphb200_body *b, int page, int64_t *out)
{
    if (!b || !out)
        return SPHB200_ERR_INVALID;
    sphb200_ctx *ctx = b->ctx;
    if (!b->cell_list_valid)
        return ctx->fail(SPHB200_ERR_INVALID, "body_page_begin: cell_list_update has not been called");
    cudaSetDevice(ctx->device);
    double info[12];
    SPHB_TRY(sphb200_body_info(b, info));
    int64_t per_page = 1;
    for (int d = 1; d < ctx->ndim; ++d)
        per_page *= int64_t(info[8 + d]);
    const int64_t pages = int64_t(info[8]);
    if (page <= 0 || b->n == 0)
    {
        *out = 0;
        return SPHB200_OK;
    }
    if (page >= pages)
    {
        *out = b->n;
        return SPHB200_OK;
    }
    u32 r = 0;
    SPHB_CUDA(ctx, cudaMemcpyAsync(&r, static_cast<const u32 *>(b->cell_offset.p) + int64_t(page) * per_page, 4,
                                   cudaMemcpyDeviceToHost, ctx->stream));
    SPHB_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    *out = int64_t(r);
    return SPHB200_OK;
}

sphb200_status sphb200_body_record_ptr(sphb200_body *b, int which, void **out)
{
    if (!b || !out || which < 0 || which > 1)
        return SPHB200_ERR_INVALID;
    sphb200_ctx *ctx = b->ctx;
    cudaSetDevice(ctx->device);
    size_t need = size_t(b->n > 0 ? b->n : 1) * 8u * size_t(ctx->real_bytes);
    SPHB_CUDA(ctx, b->rec_a.ensure(need));
    SPHB_CUDA(ctx, b->rec_b.ensure(need));
    *out = which == 0 ? b->rec_a.p : b->rec_b.p;
    return SPHB200_OK;
}

sphb200_status sphb200_body_set_owned_pages(sphb200_body *b, int page_lo, int page_hi)
{
    if (!b || (page_lo >= 0 && page_hi < page_lo))
        return SPHB200_ERR_INVALID;
    b->page_lo = page_lo < 0 ? -1 : page_lo;
    b->page_hi = page_lo < 0 ? -1 : page_hi;
    if (b->cell_list_valid)
        return updateOwnedRange(b);
    return SPHB200_OK;
}

sphb200_status sphb200_slab_recut(sphb200_body *b, int page_lo, int page_hi)
{
    if (!b || page_lo < 0 || page_hi < page_lo)
        return SPHB200_ERR_INVALID;
    if (b->page_lo < 0)
        return b->ctx->fail(SPHB200_ERR_INVALID, "slab_recut: body has no owned pages (body_set_owned_pages)");
    // the owned RANGE stays that of the current arrangement (slab_select classifies exactly those rows, by the new
    // pages); the next cell_list_update translates the new pages into the new range
    b->page_lo = page_lo;
    b->page_hi = page_hi;
    return SPHB200_OK;
}

sphb200_status sphb200_body_owned_range(sphb200_body *b, int64_t *begin, int64_t *end)
{
    if (!b)
        return SPHB200_ERR_INVALID;
    if (begin)
        *begin = b->own_begin;
    if (end)
        *end = b->own_end;
    return SPHB200_OK;
}

sphb200_status sphb200_ids_device_ptr(sphb200_body *b, void **original_id, void **sorted_id)
{
    if (!b)
        return SPHB200_ERR_INVALID;
    if (original_id)
        *original_id = b->original_id.p;
    if (sorted_id)
        *sorted_id = b->sorted_id.p;
    return SPHB200_OK;
}

sphb200_status sphb200_slab_select(sphb200_body *b, int has_left, int has_right, int64_t *counts3)
{
    if (!b || !b->find("Position"))
        return SPHB200_ERR_INVALID;
    cudaSetDevice(b->ctx->device);
    return SPHB_DISPATCH(b->ctx, slabSelect, b, has_left, has_right, counts3);
}

sphb200_status sphb200_slab_pack(sphb200_body *b, int side, const char *const *names, int n_names, void *out)
{
    if (!b || side < 1 || side > 2 || !names || n_names < 1 || (!out && b->slab_count[side] > 0))
        return SPHB200_ERR_INVALID;
    sphb200_ctx *ctx = b->ctx;
    if (!b->slab_selected)
        return ctx->fail(SPHB200_ERR_INVALID, "slab_pack: call slab_select first");
    cudaSetDevice(ctx->device);
    std::vector<Var *> vars;
    const int W = rowWidth(b, names, n_names, vars);
    if (W < 0)
        return ctx->fail(SPHB200_ERR_NOT_FOUND, "slab_pack: unknown variable");
    return ctx->real_bytes == 8 ? slabPackT<double>(b, side, vars, W, out) : slabPackT<float>(b, side, vars, W, out);
}

sphb200_status sphb200_slab_rebuild(sphb200_body *b, const char *const *names, int n_names, const void *from_left, int64_t n_left,
                                    const void *from_right, int64_t n_right)
{
    if (!b || !names || n_names < 1 || n_left < 0 || n_right < 0 || (n_left > 0 && !from_left) || (n_right > 0 && !from_right))
        return SPHB200_ERR_INVALID;
    sphb200_ctx *ctx = b->ctx;
    if (!b->slab_selected)
        return ctx->fail(SPHB200_ERR_INVALID, "slab_rebuild: call slab_select first");
    cudaSetDevice(ctx->device);
    std::vector<Var *> vars;
    const int W = rowWidth(b, names, n_names, vars);
    if (W < 0)
        return ctx->fail(SPHB200_ERR_NOT_FOUND, "slab_rebuild: unknown variable");
    return ctx->real_bytes == 8 ? slabRebuildT<double>(b, vars, W, from_left, n_left, from_right, n_right)
                                : slabRebuildT<float>(b, vars, W, from_left, n_left, from_right, n_right);
}

sphb200_status sphb200_rows_gather(sphb200_ctx *ctx, const void *src, int row_bytes, const int64_t *idx, int64_t n, void *dst)
{
    if (!ctx || row_bytes < 16 || row_bytes % 16 != 0 || n < 0 || (n > 0 && (!src || !idx || !dst)))
        return SPHB200_ERR_INVALID;
    cudaSetDevice(ctx->device);
    if (n == 0)
        return SPHB200_OK;
    const int chunks = row_bytes / 16;
    k_rows_gather<<<gridFor(n * chunks, 256), 256, 0, ctx->stream>>>(n, chunks, idx, static_cast<const uint4 *>(src), static_cast<uint4 *>(dst));
    SPHB_LAUNCH_CHECK(ctx);
    return SPHB200_OK;
}

sphb200_status sphb200_rows_scatter(sphb200_ctx *ctx, void *dst, int row_bytes, const int64_t *idx, int64_t n, const void *src)
{
    if (!ctx || row_bytes < 16 || row_bytes % 16 != 0 || n < 0 || (n > 0 && (!src || !idx || !dst)))
        return SPHB200_ERR_INVALID;
    cudaSetDevice(ctx->device);
    if (n == 0)
        return SPHB200_OK;
    const int chunks = row_bytes / 16;
    k_rows_scatter<<<gridFor(n * chunks, 256), 256, 0, ctx->stream>>>(n, chunks, idx, static_cast<const uint4 *>(src), static_cast<uint4 *>(dst));
    SPHB_LAUNCH_CHECK(ctx);
    return SPHB200_OK;
}
} // extern "C"
