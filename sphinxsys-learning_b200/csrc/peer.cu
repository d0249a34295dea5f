// Halo records over NVLink peer memory (multi-GPU slabs; the reference is single-address-space).
//
// The per-sub-step halo refresh is "pack the rows the neighbour slab holds copies of, move them, unpack": through a
// library send/recv that is pack kernel -> staging buffer -> transfer kernel -> staging buffer -> unpack kernel, each
// with its own launch and stream hand-over, ~0.55 ms per refresh at 8 x 20 M particles for 28 MB that NVLink moves in
// 40 us.  Here the pack kernel IS the transfer: k_rows_push gathers the rows and stores them straight into a landing
// buffer in the neighbour GPU's memory (peer mapping: cudaIpcOpenMemHandle across processes, the plain pointer inside
// one process), and the last block to finish publishes the sub-step's epoch in the neighbour's flag word
// (st.release.sys after __threadfence_system).  The neighbour's k_rows_wait_scatter spins on its own flag
// (ld.acquire.sys, bounded) and unpacks.  No host, no collective, no staging copy on the sender.
//
// Ordering needs no second handshake: a slab's refresh of rec_a for sub-step k+1 is issued after its own K3(k), which
// consumed the neighbour's rec_b(k) push, which the neighbour issued after ITS K2(k) - the last reader of the landing
// buffer for rec_a(k).  The same chain holds for rec_b, so one landing buffer and one flag per (side, record) suffice.
#include "internal.cuh"
#include <cstring>

using namespace sphb;

namespace sphb
{
__device__ __forceinline__ void stReleaseSys(unsigned long long *p, unsigned long long v)
{
    asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ unsigned long long ldAcquireSys(const unsigned long long *p)
{
    unsigned long long v;
    asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}

// dst_peer[r] = src[idx[r]] (rows of `chunks` uint4), then the last block stores `epoch` into *flag_peer
__global__ void __launch_bounds__(256)
    k_rows_push(int64_t rows, int chunks, const int64_t *__restrict__ idx, const uint4 *__restrict__ src, uint4 *__restrict__ dst_peer,
                unsigned long long *flag_peer, unsigned long long epoch, unsigned int *__restrict__ blocks_done)
{
    const int64_t e = blockIdx.x * int64_t(blockDim.x) + threadIdx.x;
    if (e < rows * chunks)
    {
        const int64_t r = e / chunks;
        const int c = int(e - r * chunks);
        dst_peer[r * chunks + c] = src[idx[r] * chunks + c];
    }
    __threadfence_system(); // this thread's peer stores are ordered before what follows, system-wide
    __syncthreads();
    if (threadIdx.x == 0)
    {
        const unsigned int done = atomicAdd(blocks_done, 1u) + 1u;
        if (done == gridDim.x)
        {
            *blocks_done = 0u; // ready for the next launch (stream order)
            __threadfence_system();
            stReleaseSys(flag_peer, epoch);
        }
    }
}

// one thread waits for the neighbour's epoch; gives up after `limit` clocks and counts the failure
__global__ void k_wait_epoch(const unsigned long long *flag, unsigned long long epoch, long long limit, unsigned int *timeouts)
{
    const long long t0 = clock64();
    while (ldAcquireSys(flag) < epoch)
    {
        if (clock64() - t0 > limit)
        {
            atomicAdd(timeouts, 1u);
            break;
        }
        __nanosleep(200);
    }
}

__global__ void __launch_bounds__(256)
    k_rows_unpack(int64_t rows, int chunks, const int64_t *__restrict__ idx, const uint4 *src, uint4 *__restrict__ dst)
{
    const int64_t e = blockIdx.x * int64_t(blockDim.x) + threadIdx.x;
    if (e >= rows * chunks)
        return;
    const int64_t r = e / chunks;
    const int c = int(e - r * chunks);
    // plain (coherent) load: the landing buffer was written by another GPU during this stream's lifetime
    dst[idx[r] * chunks + c] = __ldcg(src + r * chunks + c);
}
} // namespace sphb

extern "C"
{
sphb200_status sphb200_peer_alloc(sphb200_ctx *ctx, int64_t bytes, void **out)
{
    if (!ctx || !out || bytes <= 0)
        return SPHB200_ERR_INVALID;
    cudaSetDevice(ctx->device);
    void *p = nullptr;
    SPHB_CUDA(ctx, cudaMalloc(&p, size_t(bytes))); // cudaMalloc, not a pool: the allocation must be exportable
    SPHB_CUDA(ctx, cudaMemset(p, 0, size_t(bytes)));
    *out = p;
    return SPHB200_OK;
}

sphb200_status sphb200_peer_free(sphb200_ctx *ctx, void *p)
{
    if (!ctx)
        return SPHB200_ERR_INVALID;
    cudaSetDevice(ctx->device);
    if (p)
        SPHB_CUDA(ctx, cudaFree(p));
    return SPHB200_OK;
}

sphb200_status sphb200_peer_export(sphb200_ctx *ctx, const void *p, unsigned char handle64[64])
{
    if (!ctx || !p || !handle64)
        return SPHB200_ERR_INVALID;
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "handle size");
    cudaSetDevice(ctx->device);
    cudaIpcMemHandle_t h;
    SPHB_CUDA(ctx, cudaIpcGetMemHandle(&h, const_cast<void *>(p)));
    std::memcpy(handle64, &h, 64);
    return SPHB200_OK;
}

sphb200_status sphb200_peer_open(sphb200_ctx *ctx, const unsigned char handle64[64], void **out)
{
    if (!ctx || !handle64 || !out)
        return SPHB200_ERR_INVALID;
    cudaSetDevice(ctx->device);
    cudaIpcMemHandle_t h;
    std::memcpy(&h, handle64, 64);
    void *p = nullptr;
    SPHB_CUDA(ctx, cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess));
    *out = p;
    return SPHB200_OK;
}

sphb200_status sphb200_peer_close(sphb200_ctx *ctx, void *p)
{
    if (!ctx)
        return SPHB200_ERR_INVALID;
    cudaSetDevice(ctx->device);
    if (p)
        SPHB_CUDA(ctx, cudaIpcCloseMemHandle(p));
    return SPHB200_OK;
}

sphb200_status sphb200_rows_push(sphb200_ctx *ctx, const void *src, int row_bytes, const int64_t *idx, int64_t n, void *peer_landing,
                                 void *peer_flag, uint64_t epoch)
{
    if (!ctx || row_bytes < 16 || row_bytes % 16 != 0 || n < 0 || !peer_flag || (n > 0 && (!src || !idx || !peer_landing)))
        return SPHB200_ERR_INVALID;
    cudaSetDevice(ctx->device);
    if (!ctx->peer_words)
    {
        SPHB_CUDA(ctx, cudaMalloc(&ctx->peer_words, 64));
        SPHB_CUDA(ctx, cudaMemsetAsync(ctx->peer_words, 0, 64, ctx->stream));
    }
    const int chunks = row_bytes / 16;
    // n == 0 still publishes the epoch (one empty block): the receiver waits for it either way
    const int grid = n > 0 ? gridFor(n * chunks, 256) : 1;
    k_rows_push<<<grid, 256, 0, ctx->stream>>>(n, chunks, idx, static_cast<const uint4 *>(src), static_cast<uint4 *>(peer_landing),
                                               static_cast<unsigned long long *>(peer_flag), (unsigned long long)epoch, ctx->peer_words);
    SPHB_LAUNCH_CHECK(ctx);
    return SPHB200_OK;
}

sphb200_status sphb200_rows_wait_scatter(sphb200_ctx *ctx, void *dst, int row_bytes, const int64_t *idx, int64_t n, const void *landing,
                                         const void *flag, uint64_t epoch)
{
    if (!ctx || row_bytes < 16 || row_bytes % 16 != 0 || n < 0 || !flag || (n > 0 && (!dst || !idx || !landing)))
        return SPHB200_ERR_INVALID;
    cudaSetDevice(ctx->device);
    if (!ctx->peer_words)
    {
        SPHB_CUDA(ctx, cudaMalloc(&ctx->peer_words, 64));
        SPHB_CUDA(ctx, cudaMemsetAsync(ctx->peer_words, 0, 64, ctx->stream));
    }
    const long long limit = 10000000000ll; // ~5 s of SM clocks: a dead neighbour must not hang the device
    k_wait_epoch<<<1, 1, 0, ctx->stream>>>(static_cast<const unsigned long long *>(flag), (unsigned long long)epoch, limit, ctx->peer_words + 1);
    SPHB_LAUNCH_CHECK(ctx);
    if (n > 0)
    {
        const int chunks = row_bytes / 16;
        k_rows_unpack<<<gridFor(n * chunks, 256), 256, 0, ctx->stream>>>(n, chunks, idx, static_cast<const uint4 *>(landing), static_cast<uint4 *>(dst));
        SPHB_LAUNCH_CHECK(ctx);
    }
    return SPHB200_OK;
}

sphb200_status sphb200_peer_timeouts(sphb200_ctx *ctx, int64_t *out)
{
    if (!ctx || !out)
        return SPHB200_ERR_INVALID;
    cudaSetDevice(ctx->device);
    unsigned int v = 0;
    if (ctx->peer_words)
    {
        SPHB_CUDA(ctx, cudaMemcpyAsync(&v, ctx->peer_words + 1, 4, cudaMemcpyDeviceToHost, ctx->stream));
        SPHB_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    }
    *out = int64_t(v);
    return SPHB200_OK;
}
} // extern "C"
