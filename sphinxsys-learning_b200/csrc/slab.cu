// Slab-decomposition support of the C ABI (multi-GPU; the reference has no equivalent).
// The exchange itself (NCCL send/recv of migrating and halo particles) is driven by the host layer
// through torch.distributed on raw device pointers; the library only has to resize a body, know which
// contiguous run of the cell-sorted particles this rank owns, and share the caller's stream.
#include "internal.cuh"

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
} // extern "C"
