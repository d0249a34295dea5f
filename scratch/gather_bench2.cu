// Micro-benchmark (development tool), round 2: is a candidate run staged ONCE per CTA in shared memory and read
// with scattered LDS a way around the L1 data pipe's one-wavefront-per-32-byte-sector cost of per-lane LDG
// gathers ("what's weak" 2)?  Every variant moves K 64-byte records per thread into registers,
// record indices are pseudo-random inside a window of W records around the CTA (SPH-like locality).
//   A   : 2 x LDG.256 per record straight from global (k_second_half_tile's gather)             [baseline]
//   A4  : 4 x LDG.128 per record
//   C64 : window staged per CTA (coalesced 128-bit copies), AoS stride 64 B, 4 x LDS.128 per record
//   C80 : same, AoS stride 80 B (16-byte chunk c of record j in bank group (5j + c) mod 8: skewed)
//   CS  : same, SoA (8 arrays of doubles), 8 x LDS.64 per record
//   CB  : same as C80 but staged with ONE 1-D cp.async.bulk per 16 records + mbarrier (UBLKCP)
//   F4  : LDG.256 where the 4 lanes of a group read the 4 sectors of ONE 128-byte line  (does the coalescer merge them?)
//   F1  : LDG.256 where 2 adjacent lanes read the SAME sector
//   FC  : LDG.256 fully coalesced (lane l reads record base + l)                               [upper bound of the LDG path]
// nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a scratch/gather_bench2.cu -o scratch/gather_bench2
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>

#define CK(x)                                                                       \
    do                                                                              \
    {                                                                               \
        cudaError_t e = (x);                                                        \
        if (e != cudaSuccess)                                                       \
        {                                                                           \
            std::printf("%s: %s (line %d)\n", #x, cudaGetErrorString(e), __LINE__); \
            std::exit(1);                                                           \
        }                                                                           \
    } while (0)

struct alignas(64) Rec
{
    double a[8];
};
constexpr int K = 128;   // records per thread
constexpr int W = 1024;  // window (records) a CTA's threads pick from; the staged variants copy exactly this
constexpr int TPB = 128; // threads per CTA; a CTA "owns" 2*TPB records in the middle of its window

__device__ __forceinline__ uint32_t hashIdx(uint32_t t, int k)
{
    uint32_t h = (t * 2654435761u) ^ (uint32_t(k) * 40503u + 0x9e3779b9u);
    h ^= h >> 15;
    h *= 2246822519u;
    h ^= h >> 13;
    return h % W;
}
__device__ __forceinline__ uint32_t windowBase(uint32_t cta, uint32_t n)
{
    uint64_t b = uint64_t(cta) * (2u * TPB);
    return uint32_t(b + W <= n ? b : n - W);
}
__device__ __forceinline__ void ld256(const double *p, double *v)
{
    asm volatile("ld.global.nc.v4.f64 {%0,%1,%2,%3}, [%4];" : "=d"(v[0]), "=d"(v[1]), "=d"(v[2]), "=d"(v[3]) : "l"(p));
}
__device__ __forceinline__ void ld128(const double *p, double *v)
{
    asm volatile("ld.global.nc.v2.f64 {%0,%1}, [%2];" : "=d"(v[0]), "=d"(v[1]) : "l"(p));
}

template <int MODE> // 0 A, 1 A4, 2 F4, 3 F1, 4 FC
__global__ void __launch_bounds__(TPB) k_ldg(const Rec *__restrict__ rec, uint32_t n, double *__restrict__ out)
{
    const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    const uint32_t base = windowBase(blockIdx.x, n);
    const int lane = threadIdx.x & 31;
    double s = 0;
#pragma unroll 1
    for (int k0 = 0; k0 < K; k0 += 4)
    {
        double v[4][8];
#pragma unroll
        for (int u = 0; u < 4; ++u)
        {
            uint32_t j;
            if (MODE == 2) // groups of 4 lanes: one random 256-byte pair of lines, lane picks its 64-byte record inside
                j = (hashIdx(t & ~3u, k0 + u) & ~3u) + (lane & 3);
            else if (MODE == 3) // 2 adjacent lanes read the same record
                j = hashIdx(t & ~1u, k0 + u);
            else if (MODE == 4)
                j = (hashIdx(t & ~31u, k0 + u) & ~31u) + lane;
            else
                j = hashIdx(t, k0 + u);
            const Rec *r = rec + base + j;
            if (MODE == 1)
            {
                ld128(r->a, v[u]);
                ld128(r->a + 2, v[u] + 2);
                ld128(r->a + 4, v[u] + 4);
                ld128(r->a + 6, v[u] + 6);
            }
            else
            {
                ld256(r->a, v[u]);
                ld256(r->a + 4, v[u] + 4);
            }
        }
#pragma unroll
        for (int u = 0; u < 4; ++u)
#pragma unroll
            for (int q = 0; q < 8; ++q)
                s += v[u][q];
    }
    out[t] = s;
}

__device__ __forceinline__ uint32_t smemAddr(const void *p) { return uint32_t(__cvta_generic_to_shared(p)); }

template <int MODE> // 0 C64, 1 C80, 2 CS, 3 CB (bulk-staged C64... uses stride 64: bulk copies are contiguous)
__global__ void __launch_bounds__(TPB) k_staged(const Rec *__restrict__ rec, uint32_t n, double *__restrict__ out)
{
    extern __shared__ __align__(128) unsigned char smem[];
    __shared__ alignas(8) unsigned long long bar;
    const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    const uint32_t base = windowBase(blockIdx.x, n);
    constexpr int STRIDE = MODE == 1 ? 80 : 64;
    // ---- stage the window
    if (MODE == 3)
    {
        if (threadIdx.x == 0)
        {
            asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smemAddr(&bar)));
            asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        }
        __syncthreads();
        if (threadIdx.x == 0)
            asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smemAddr(&bar)), "r"(uint32_t(W * 64)) : "memory");
        __syncthreads();
        if (threadIdx.x < W / 16) // one 1 KB bulk copy per thread (16 records)
            asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], 1024, [%2];" ::"r"(
                             smemAddr(smem + threadIdx.x * 1024)),
                         "l"(rec + base + threadIdx.x * 16), "r"(smemAddr(&bar))
                         : "memory");
        uint32_t done = 0;
        while (!done)
            asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0, 1, 0, p; }"
                         : "=r"(done)
                         : "r"(smemAddr(&bar)), "r"(0u)
                         : "memory");
    }
    else
    {
        const double2 *src = reinterpret_cast<const double2 *>(rec + base);
        for (int c = threadIdx.x; c < W * 4; c += TPB) // 16-byte chunks, coalesced
        {
            double2 v = __ldg(src + c);
            const int j = c >> 2, q = c & 3;
            if (MODE == 2)
            {
                double *soa = reinterpret_cast<double *>(smem);
                soa[(2 * q) * W + j] = v.x;
                soa[(2 * q + 1) * W + j] = v.y;
            }
            else
                *reinterpret_cast<double2 *>(smem + j * STRIDE + q * 16) = v;
        }
        __syncthreads();
    }
    double s = 0;
#pragma unroll 1
    for (int k0 = 0; k0 < K; k0 += 4)
    {
        double v[4][8];
#pragma unroll
        for (int u = 0; u < 4; ++u)
        {
            const uint32_t j = hashIdx(t, k0 + u);
            if (MODE == 2)
            {
                const double *soa = reinterpret_cast<const double *>(smem);
#pragma unroll
                for (int q = 0; q < 8; ++q)
                    v[u][q] = soa[q * W + j];
            }
            else
            {
                const double2 *p = reinterpret_cast<const double2 *>(smem + j * STRIDE);
#pragma unroll
                for (int q = 0; q < 4; ++q)
                {
                    double2 x = p[q];
                    v[u][2 * q] = x.x;
                    v[u][2 * q + 1] = x.y;
                }
            }
        }
#pragma unroll
        for (int u = 0; u < 4; ++u)
#pragma unroll
            for (int q = 0; q < 8; ++q)
                s += v[u][q];
    }
    out[t] = s;
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
    const int grid = int(threads / TPB);
    const char *names[] = {"A   2xLDG.256 scattered", "A4  4xLDG.128 scattered", "F4  LDG.256 4 lanes/line", "F1  LDG.256 2 lanes/sector",
                           "FC  LDG.256 coalesced", "C64 staged AoS-64 LDS.128", "C80 staged AoS-80 LDS.128", "CS  staged SoA LDS.64",
                           "CB  bulk-staged AoS-64 LDS.128"};
    CK(cudaFuncSetAttribute(k_staged<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, W * 80));
    CK(cudaFuncSetAttribute(k_staged<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, W * 80));
    CK(cudaFuncSetAttribute(k_staged<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, W * 80));
    CK(cudaFuncSetAttribute(k_staged<3>, cudaFuncAttributeMaxDynamicSharedMemorySize, W * 80));
    for (int which = 0; which < 9; ++which)
    {
        float best = 1e30f;
        for (int rep = 0; rep < 4; ++rep)
        {
            CK(cudaEventRecord(e0));
            switch (which)
            {
            case 0: k_ldg<0><<<grid, TPB>>>(rec, n, out); break;
            case 1: k_ldg<1><<<grid, TPB>>>(rec, n, out); break;
            case 2: k_ldg<2><<<grid, TPB>>>(rec, n, out); break;
            case 3: k_ldg<3><<<grid, TPB>>>(rec, n, out); break;
            case 4: k_ldg<4><<<grid, TPB>>>(rec, n, out); break;
            case 5: k_staged<0><<<grid, TPB, W * 64>>>(rec, n, out); break;
            case 6: k_staged<1><<<grid, TPB, W * 80>>>(rec, n, out); break;
            case 7: k_staged<2><<<grid, TPB, W * 64>>>(rec, n, out); break;
            case 8: k_staged<3><<<grid, TPB, W * 64>>>(rec, n, out); break;
            }
            CK(cudaEventRecord(e1));
            CK(cudaEventSynchronize(e1));
            CK(cudaGetLastError());
            float ms;
            CK(cudaEventElapsedTime(&ms, e0, e1));
            if (rep > 0 && ms < best)
                best = ms;
        }
        const double recs = double(grid) * TPB * K;
        std::printf("%-32s %8.3f ms  %.3f lane-records per clock per SM  (%d MHz, %d SMs)\n", names[which], best,
                    recs / (best * 1e-3) / (double(khz) * 1e3) / sms, khz / 1000, sms);
    }
    return 0;
}
