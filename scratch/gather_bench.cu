// Micro-benchmark (development tool): throughput of per-lane scattered 64-byte record gathers on sm_100a
//   A: two LDG.256 per record into registers (what k_second_half_tile does)
//   B: one cp.async.bulk (UBLKCP, TMA path) per lane and record into a per-warp shared-memory stage ring,
//      completion by mbarrier transaction bytes, then 4 conflict-free LDS.128 per lane
// Question: does the TMA path avoid the "one wavefront per lane per 32 bytes" cost of the L1 data pipe?
// nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a scratch/gather_bench.cu -o scratch/gather_bench
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>
#include <vector>

#define CK(x)                                                                               \
    do                                                                                      \
    {                                                                                       \
        cudaError_t e = (x);                                                                \
        if (e != cudaSuccess)                                                               \
        {                                                                                   \
            std::printf("%s: %s (line %d)\n", #x, cudaGetErrorString(e), __LINE__);         \
            std::exit(1);                                                                   \
        }                                                                                   \
    } while (0)

struct alignas(64) Rec
{
    double a[8];
};
constexpr int K = 128;      // records per thread
constexpr int WINDOW = 4096; // neighbours lie within this many records of the thread's own (SPH-like locality)

__device__ __forceinline__ uint32_t nextIdx(uint32_t t, int k, uint32_t n)
{
    uint32_t h = (t * 2654435761u) ^ (uint32_t(k) * 40503u + 0x9e3779b9u);
    h ^= h >> 15;
    h *= 2246822519u;
    h ^= h >> 13;
    uint32_t center = (t * 2u) % n; // like 2 particles per thread
    uint32_t off = h % WINDOW;
    uint32_t j = center + off;
    return j < n ? j : j - n;
}

__device__ __forceinline__ void ld256(const double *p, double *v)
{
    asm volatile("ld.global.nc.v4.f64 {%0,%1,%2,%3}, [%4];" : "=d"(v[0]), "=d"(v[1]), "=d"(v[2]), "=d"(v[3]) : "l"(p));
}

__global__ void __launch_bounds__(128) k_ldg(const Rec *__restrict__ rec, uint32_t n, double *__restrict__ out)
{
    const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    double s = 0;
#pragma unroll 1
    for (int k0 = 0; k0 < K; k0 += 4)
    {
        double v[4][8];
#pragma unroll
        for (int u = 0; u < 4; ++u)
        {
            const Rec *r = rec + nextIdx(t, k0 + u, n);
            ld256(r->a, v[u]);
            ld256(r->a + 4, v[u] + 4);
        }
#pragma unroll
        for (int u = 0; u < 4; ++u)
#pragma unroll
            for (int q = 0; q < 8; ++q)
                s += v[u][q];
    }
    out[t] = s;
}

constexpr int STAGES = 4;
__device__ __forceinline__ uint32_t smemAddr(const void *p) { return uint32_t(__cvta_generic_to_shared(p)); }

__global__ void __launch_bounds__(128) k_bulk(const Rec *__restrict__ rec, uint32_t n, double *__restrict__ out)
{
    __shared__ Rec stage[4][STAGES][32];                 // [warp][stage][lane]
    __shared__ alignas(8) unsigned long long bar[4][STAGES];
    const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    if (lane == 0)
        for (int s = 0; s < STAGES; ++s)
            asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smemAddr(&bar[w][s])));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    __syncwarp();
    auto issue = [&](int k)
    {
        const int s = k % STAGES;
        if (lane == 0)
            asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smemAddr(&bar[w][s])), "r"(32u * 64u) : "memory");
        __syncwarp();
        const Rec *r = rec + nextIdx(t, k, n);
        asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], 64, [%2];" ::"r"(smemAddr(&stage[w][s][lane])),
                     "l"(r), "r"(smemAddr(&bar[w][s]))
                     : "memory");
    };
    for (int k = 0; k < STAGES; ++k)
        issue(k);
    double sum = 0;
#pragma unroll 1
    for (int k = 0; k < K; ++k)
    {
        const int s = k % STAGES;
        const uint32_t parity = (k / STAGES) & 1;
        uint32_t done = 0;
        while (!done)
            asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0, 1, 0, p; }"
                         : "=r"(done)
                         : "r"(smemAddr(&bar[w][s])), "r"(parity)
                         : "memory");
        const double4 *p = reinterpret_cast<const double4 *>(&stage[w][s][lane]);
        double4 a = p[0], b = p[1];
        sum += a.x + a.y + a.z + a.w + b.x + b.y + b.z + b.w;
        __syncwarp();
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        if (k + STAGES < K)
            issue(k + STAGES);
    }
    out[t] = sum;
}

int main(int argc, char **argv)
{
    const uint32_t n = argc > 1 ? uint32_t(std::atol(argv[1])) : 20000000u;
    const uint32_t threads = n / 2;
    Rec *rec;
    double *out;
    CK(cudaMalloc(&rec, size_t(n) * sizeof(Rec)));
    CK(cudaMalloc(&out, size_t(threads) * 8));
    CK(cudaMemset(rec, 0, size_t(n) * sizeof(Rec)));
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0));
    CK(cudaEventCreate(&e1));
    int sms = 0, khz = 0;
    CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0));
    CK(cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, 0));
    const int grid = int((threads + 127) / 128);
    for (int which = 0; which < 2; ++which)
    {
        float best = 1e30f;
        for (int rep = 0; rep < 4; ++rep)
        {
            CK(cudaEventRecord(e0));
            if (which == 0)
                k_ldg<<<grid, 128>>>(rec, n, out);
            else
                k_bulk<<<grid, 128>>>(rec, n, out);
            CK(cudaEventRecord(e1));
            CK(cudaEventSynchronize(e1));
            CK(cudaGetLastError());
            float ms;
            CK(cudaEventElapsedTime(&ms, e0, e1));
            if (rep > 0 && ms < best)
                best = ms;
        }
        const double recs = double(threads) * K;
        std::printf("%s: %.3f ms, %.3g lane-records/s, %.3f lane-records per clock per SM (max clock %d MHz, %d SMs)\n",
                    which == 0 ? "LDG.256 x2 " : "cp.async.bulk", best, recs / (best * 1e-3), recs / (best * 1e-3) / (double(khz) * 1e3) / sms,
                    khz / 1000, sms);
    }
    return 0;
}
