// Internal types of libsphb200: context / body / relation objects behind the opaque C handles,
// device-side parameter PODs, Wendland-C2 helpers and the (ndim, Real) dispatch.
#pragma once
#include "../../include/sphb200.h"
#include <cuda_runtime.h>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <map>
#include <memory>
#include <string>
#include <vector>

namespace sphb
{
typedef uint32_t u32;
typedef uint64_t u64;

constexpr int kMaxTargets = 4;
constexpr double kPi = 3.14159265358979323846;

// common/base_data_type.h:207 TinyReal
template <class R>
struct Tiny
{
    static constexpr R value = R(2.71051e-20);
};

template <class R>
struct Real4;
template <>
struct Real4<float>
{
    typedef float4 type;
};
template <>
struct Real4<double>
{
    typedef double4 type;
};

// Packed per-particle record of the acoustic sub-step kernels: one aligned 8-Real line
// {x, y, z, p | Vol, vx, vy, vz} (32 B in fp32 = one L2 sector, 64 B in fp64 = two), fetched per
// neighbour with 256-bit loads.  x and Vol are the values frozen at configuration time.  The first half
// is all the 1st-half pair sum needs when every particle of the body has the same Vol (the usual state of a
// free-surface WCSPH body: body.vol_mm holds min/max of Vol, the kernel folds the constant in) - the L1 data
// pipe moves one 32-byte sector per cycle per SM for scattered gathers, so halving the sectors halves its time.
template <class R>
struct alignas(8 * sizeof(R)) Rec8
{
    R a[8];
};

// ---------------------------------------------------------------------------------------------
// Device-side parameter blocks (passed by value to kernels)
// ---------------------------------------------------------------------------------------------
// meshes/base_mesh.cpp:6-16, base_mesh.hxx:9-15,68-78
template <int D, class R>
struct MeshP
{
    R lo[D];
    R spacing;
    int cells[D];

    __host__ __device__ inline void cellIndex(const R *pos, int *c) const
    {
#pragma unroll
        for (int d = 0; d < D; ++d)
        {
            int v = int(floor((pos[d] - lo[d]) / spacing));
            v = v < 0 ? 0 : v;
            v = v > cells[d] - 1 ? cells[d] - 1 : v;
            c[d] = v;
        }
    }
    // x slowest, last axis fastest
    __host__ __device__ inline u32 linear(const int *c) const
    {
        if (D == 2)
            return u32(c[0]) * u32(cells[1]) + u32(c[1]);
        return (u32(c[0]) * u32(cells[1]) + u32(c[1])) * u32(cells[D - 1]) + u32(c[D - 1]);
    }
    __host__ __device__ inline u64 total() const
    {
        u64 n = 1;
        for (int d = 0; d < D; ++d)
            n *= u64(cells[d]);
        return n;
    }
};

// kernels/kernel_wendland_c2.cpp:8-35, base_kernel.cpp:6-27
template <class R>
struct KernelP
{
    R h, inv_h, rc, rc2, factor_W, factor_dW;
};

template <int D, class R>
struct PeriodicP
{
    int mask; // bit a set: axis a periodic
    R lo[D], hi[D], L[D], half_L[D];
};

// everything (D,R)-typed that a body owns on the host side
template <int D, class R>
struct BodyP
{
    R dp, h;
    KernelP<R> kern;
    R sigma0;
    MeshP<D, R> mesh;
    PeriodicP<D, R> per;
};

// Neighbour list of one (source, target) pair in a sliced-ELL layout made for kS lanes per particle:
// a slice holds kP = 32/kS consecutive particles; entry e of particle i (slice s = i/kP) sits at
//   list[32*(slice_off[s] + e/kS) + (i%kP)*kS + e%kS],   count[i] entries are valid.
// A warp that gives kS lanes to each of the kP particles of a slice (lane = (i%kP)*kS + e%kS) reads 128
// contiguous bytes per step, and the kS lanes of one particle walk ascending, mostly adjacent neighbours,
// so the lanes of one gather instruction touch fewer distinct 128-byte lines.  Measured on B200 (8 M
// particles, ms K2/K3): fp32 kS=1 1.86/1.76, 2 1.78/1.67, 4 2.18/2.06, 8 2.83/2.53; fp64 4.95/4.79, 4.54/4.24,
// 4.91/4.44, 5.72/5.06 - the L1 data stage spends one wavefront per gathered 32-byte record whatever the
// line sharing, so only kS=2 pays for its reduction shuffles.
#ifndef SPHB_LANES
#define SPHB_LANES 2
#endif
constexpr int kS = SPHB_LANES; // lanes per particle in the pair-sum kernels
constexpr int kP = 32 / kS;    // particles per slice
static_assert(kS == 1 || kS == 2 || kS == 4 || kS == 8, "SPHB_LANES must be 1, 2, 4 or 8");

// Union list of a group of G consecutive source particles (tile.cuh): the ascending merge of the members'
// neighbour lists, SELL-32 over groups (entry k of group g at list[32*(slice_off[g/32] + k) + g%32]).
// One thread of the register-tiled pair kernels owns a group: every gathered neighbour record is used for
// all G members (masked by the cut-off test), which divides the L1 gather traffic per useful pair.
struct UnionList
{
    u32 *count = nullptr;     // [n_groups]
    u32 *slice_w = nullptr;   // [n_groups/32]
    u32 *slice_off = nullptr; // [n_groups/32 + 1]
    u32 *list = nullptr;
    u32 *queue = nullptr;     // [1 + n_groups] counter, then the groups whose members lie in different cells (k_union_mask)
    size_t cap_groups = 0, cap_entries = 0;
    int G = 0;     // particles per union group
    int lanes = 1; // threads of the pair kernel that walk one group's list together (a slice holds 32/lanes groups)
    int64_t base = 0, n_groups = 0, own_begin = -1, own_end = -1;
    u64 total_entries = 0;
    uint64_t build_id = 0; // NeighborList::build_id this union was merged from
    uint64_t active_build = 0; // build whose non-empty groups are listed in queue (tile.cuh ensureActive)
    bool active_valid = false;
};

struct NeighborList
{
    UnionList uni;      // groups of 2, one thread per group (1st half)
    UnionList uni_wide; // groups of 4 shared by 2 adjacent threads of 2 members each (2nd half)
    uint64_t build_id = 0;    // bumped by every relation_update
    bool sorted = false;      // every particle's entries ascend (non-periodic searches)
    u32 *count = nullptr;     // [n_src]
    u32 *slice_w = nullptr;   // [n_slices]  max over the slice of ceil(count/kS)
    u32 *slice_off = nullptr; // [n_slices+1] exclusive scan of slice_w (units of 32 entries)
    u32 *list = nullptr;      // index list; filled on demand when the search left hit masks (ensureList)
    size_t cap_entries = 0, cap_src = 0;
    u64 total_entries = 0; // 32 * slice_off[n_slices]
    u64 total_pairs = 0;
    u32 *mask = nullptr;   // hit bits of the masked search (relation.cu), kept until the next update
    size_t cap_mask = 0;   // in words
    bool has_mask = false; // this build's hits are in mask (fast search path)
    bool expanded = false; // list holds this build's indices
};

// Device-resident control block of the acoustic sub-step loop (fluid.cu sphb200_acoustic_relaxation): the loop
// `while (relaxation_time < Dt) { ... dt = AcousticTimeStep ... }` of the reference mains runs without the host -
// every sub-step kernel reads dt (and the loop's exit flag) from here, k_substep_ctl advances it from the max slot.
struct StepCtl
{
    double dt;     // acoustic time step the next sub-step integrates with
    double relax;  // relaxation_time of the running advection step
    double Dt;     // advection time step of the running advection step
    double time;   // physical time accumulated over all sub-steps since ctx creation / reset
    double dt_bound; // external upper bound folded into dt (slab runs: stays +inf here)
    long long count_total;
    int count;     // sub-steps done in the running advection step
    int done;      // relax >= Dt: the remaining launches of this call return at once
};

// per-target device pointers handed to the interaction kernels
template <class R>
struct TargetP
{
    const u32 *count;
    const u32 *slice_off;
    const u32 *list;
    const typename Real4<R>::type *xV; // frozen position + Vol of the target body
    const R *mass, *vel, *acc, *nrm;
    R inv_rho0;
};
template <class R>
struct ContactP
{
    int n;
    TargetP<R> t[kMaxTargets];
};

// ---------------------------------------------------------------------------------------------
// Host objects
// ---------------------------------------------------------------------------------------------
struct DevBuf
{
    void *p = nullptr;
    size_t bytes = 0;
    cudaError_t ensure(size_t need, bool keep = false, cudaStream_t st = 0);
    void release();
};

struct Var
{
    std::string name;
    int width = 1;
    bool evolving = false;
    void *host = nullptr;
    DevBuf dev;
    DevBuf alt; // destination of the next cell-order gather (swapped with dev; no allocation per sort)
    DevBuf stage;          // target of var_stage, swapped with dev by body_commit_staged
    int64_t staged_n = -1; // particles the staged copy holds, -1: nothing staged
    uint64_t staged_stamp = 0; // body.config_stamp when var_stage was called
    DevBuf fetch;          // device-side snapshot var_fetch_async copies to the host from
};
} // namespace sphb

struct sphb200_ctx
{
    int device = 0, ndim = 3, real_bytes = 8;
    double sys_lo[3] = {0, 0, 0}, sys_hi[3] = {0, 0, 0};
    cudaStream_t stream = nullptr;
    bool owns_stream = true;
    cudaStream_t copy_stream = nullptr; // var_stage copies (created on first use)
    cudaEvent_t copy_done = nullptr, mirrors_free = nullptr;
    cudaStream_t fetch_stream = nullptr; // var_fetch_async copies (created on first use)
    cudaEvent_t fetch_ready = nullptr, fetch_done = nullptr;
    std::string error;
    int64_t launches = 0;
    unsigned int *peer_words = nullptr; // peer.cu: [0] blocks-done counter of k_rows_push, [1] wait time-outs
    int pair_kernel = -2; // -2: process default (env SPHB200_TILE), -1 auto, 0 per-particle lists, 1 group-tiled
    std::vector<sphb200_body *> bodies;
    std::vector<sphb200_relation *> relations;
    sphb::DevBuf scratch_a, scratch_b, scan_tmp, reduce_dev, subkey_buf;
    sphb::DevBuf step_ctl;                      // StepCtl, device side
    const sphb::StepCtl *ctl_active = nullptr;  // non-null while sphb200_acoustic_relaxation queues its sub-steps
    void *reduce_host = nullptr; // pinned, 64 doubles

    sphb200_status fail(sphb200_status s, const std::string &msg)
    {
        error = msg;
        return s;
    }
};

struct sphb200_body
{
    sphb200_ctx *ctx = nullptr;
    std::string name;
    int64_t n = 0;
    double dp = 0, h_ratio = 1.3;
    // material
    double rho0 = 1, c0 = 1, mu = 0, youngs = 0, poisson = 0, K0 = 0, G0 = 0, lambda0 = 0, cs0 = 0;
    bool periodic[3] = {false, false, false};
    double per_lo[3] = {0, 0, 0}, per_hi[3] = {0, 0, 0};
    alignas(16) unsigned char params[512]; // BodyP<D,R>
    std::map<std::string, std::unique_ptr<sphb::Var>> vars;
    sphb::DevBuf original_id, sorted_id;            // u32[n]
    sphb::DevBuf cell_of, rank_in_cell, perm;        // u32[n]
    sphb::DevBuf cell_count, cell_offset;            // u32[cells+1]
    sphb::DevBuf xV;                                 // Real4[n]: frozen position + Vol
    sphb::DevBuf rec_a, rec_b;                       // Rec8[n]: {xV | p, v} before / after the 1st-half update
    sphb::DevBuf vol_mm;                             // u64[2]: bit patterns of min / max of xV[].w (Vol), see refreshVolRange
    sphb::DevBuf wall_sum;                           // Real[n][D+1]: wall sums of the pair kernels' pre-pass (tile.cuh)
    sphb::DevBuf xf;                                 // float4[n]: fp32 copy of (x - mesh.lo) for the search pre-filter (relation.cu)
    uint64_t xf_stamp = ~uint64_t(0);                // config_stamp xf was made from
    bool rec_b_valid = false;                        // rec_b matches xV, Pressure and Velocity
    sphb::DevBuf tmp_var;                            // permutation double buffer
    int64_t n_cells = 0;
    bool cell_list_valid = false;
    bool xV_vol_dirty = true; // Vol component of xV must be refreshed before the next interaction
    uint64_t config_stamp = 0;
    bool acoustic_max_valid = false; // reduce slot 0 holds max(c0+|v|) from the last 1st half
    // slab decomposition: particles [own_begin, own_end) (cell-sorted order) are owned by this rank,
    // the rest are halo copies that only serve as neighbours j.  Defaults to the whole body.
    int64_t own_begin = 0, own_end = 0;
    int page_lo = -1, page_hi = -1; // owned range of cell pages along axis 0 (x slowest), or -1
    sphb::DevBuf slab_flags, slab_idx[3]; // slab.cu slab_select: rows that stay / go to the left / to the right slab
    int64_t slab_count[3] = {0, 0, 0};
    bool slab_selected = false;

    template <int D, class R>
    sphb::BodyP<D, R> &P() { return *reinterpret_cast<sphb::BodyP<D, R> *>(params); }
    sphb::Var *find(const char *nm)
    {
        auto it = vars.find(nm);
        return it == vars.end() ? nullptr : it->second.get();
    }
};

struct sphb200_relation
{
    sphb200_ctx *ctx = nullptr;
    sphb200_body *body = nullptr;
    bool inner = true;
    std::vector<sphb200_body *> targets; // inner: {body}
    std::vector<sphb::NeighborList> lists;
    std::vector<uint64_t> target_stamp;
    uint64_t source_stamp = 0;
    bool valid = false;
    // neighbour builder (particle_neighborhood/neighborhood.cpp): 0 NeighborBuilderInner / Contact, 1 SelfContact
    // (:123-142: r < rc and r0 > rc on InitialPosition), 2 SurfaceContact (:162-168: kernel of the averaged smoothing length)
    int builder = 0;
    sphb::DevBuf part_flag; // u8 per ORIGINAL id: the source particles of a body part (BodySurfaceLayer); empty = all
    bool has_part = false;
};

namespace sphb
{
// ---------------------------------------------------------------------------------------------
// error helpers
// ---------------------------------------------------------------------------------------------
#define SPHB_CUDA(ctx, expr)                                                                                     \
    do                                                                                                           \
    {                                                                                                            \
        cudaError_t _e = (expr);                                                                                 \
        if (_e != cudaSuccess)                                                                                   \
            return (ctx)->fail(SPHB200_ERR_CUDA, std::string(#expr) + ": " + cudaGetErrorString(_e) + " at " +   \
                                                     __FILE__ + ":" + std::to_string(__LINE__));                 \
    } while (0)

#define SPHB_TRY(expr)               \
    do                               \
    {                                \
        sphb200_status _s = (expr);  \
        if (_s != SPHB200_OK)        \
            return _s;               \
    } while (0)

#define SPHB_LAUNCH_CHECK(ctx)                  \
    do                                          \
    {                                           \
        (ctx)->launches++;                      \
        SPHB_CUDA(ctx, cudaGetLastError());     \
    } while (0)

// dispatch a templated implementation on the context's (ndim, real_bytes)
#define SPHB_DISPATCH(ctx, FN, ...)                                         \
    ((ctx)->ndim == 3 ? ((ctx)->real_bytes == 8 ? FN<3, double>(__VA_ARGS__) \
                                                : FN<3, float>(__VA_ARGS__)) \
                      : ((ctx)->real_bytes == 8 ? FN<2, double>(__VA_ARGS__) \
                                                : FN<2, float>(__VA_ARGS__)))

inline int gridFor(int64_t n, int block) { return int((n + block - 1) / block); }

// variable lookup with auto-registration (mirrors registerStateVariable's create-if-missing)
sphb200_status getVar(sphb200_body *b, const char *name, int width, bool evolving, Var **out);
template <class R>
inline sphb200_status getVarPtr(sphb200_body *b, const char *name, int width, R **out, bool evolving = false)
{
    Var *v = nullptr;
    SPHB_TRY(getVar(b, name, width, evolving, &v));
    *out = static_cast<R *>(v->dev.p);
    return SPHB200_OK;
}

// device scan of n u32 (exclusive); out may alias in.  Total written to out[n] when with_total.
sphb200_status exclusiveScanU32(sphb200_ctx *ctx, const u32 *in, u32 *out, int64_t n, bool with_total);

// refresh the Vol component of xV (and nothing else) if a dynamics changed VolumetricMeasure
template <int D, class R>
sphb200_status ensurePacked(sphb200_body *b);

// min / max of the packed Vol into body.vol_mm (device side only; the 1st-half kernels read it)
template <int D, class R>
sphb200_status refreshVolRange(sphb200_body *b);

// slab.cu: owned pages → owned particle range, after a cell_list_update
sphb200_status updateOwnedRange(sphb200_body *b);

// reduction slots (device u64 bit patterns of non-negative doubles)
sphb200_status reduceReset(sphb200_ctx *ctx, int slot, bool for_min);
sphb200_status reduceRead(sphb200_ctx *ctx, int slot, double *out);

// ---------------------------------------------------------------------------------------------
// device helpers
// ---------------------------------------------------------------------------------------------
#ifdef __CUDACC__
__device__ __forceinline__ double mul_rn(double a, double b) { return __dmul_rn(a, b); }
__device__ __forceinline__ float mul_rn(float a, float b) { return __fmul_rn(a, b); }
__device__ __forceinline__ double add_rn(double a, double b) { return __dadd_rn(a, b); }
__device__ __forceinline__ float add_rn(float a, float b) { return __fadd_rn(a, b); }
__device__ __forceinline__ double rsqrt_r(double a) { return rsqrt(a); }
__device__ __forceinline__ float rsqrt_r(float a) { return rsqrtf(a); }
__device__ __forceinline__ double sqrt_r(double a) { return sqrt(a); }
__device__ __forceinline__ float sqrt_r(float a) { return sqrtf(a); }

// squaredNorm evaluated exactly as the CPU path does: products and left-to-right sums, no FMA
template <int D, class R>
__device__ __forceinline__ R sqnormExact(const R *d)
{
    R s = mul_rn(d[0], d[0]);
#pragma unroll
    for (int k = 1; k < D; ++k)
        s = add_rn(s, mul_rn(d[k], d[k]));
    return s;
}

// read-only 16/32-byte record load (double4 has no __ldg overload: two 16-byte halves)
__device__ __forceinline__ float4 ldg4(const float4 *p) { return __ldg(p); }
__device__ __forceinline__ double4 ldg4(const double4 *p)
{
    // one LDG.E.256 (the Real4 arrays are 32-byte aligned: cudaMalloc base, 32-byte elements) instead of two 128-bit
    // loads: half the load instructions and L1 wavefronts per candidate / neighbour
    double4 v;
    asm("ld.global.nc.v4.f64 {%0,%1,%2,%3}, [%4];" : "=d"(v.x), "=d"(v.y), "=d"(v.z), "=d"(v.w) : "l"(p));
    return v;
}

// 256-bit loads/stores of record halves (LDG.E.256 on sm_100a)
__device__ __forceinline__ void ld256(const float *p, float *v)
{
    asm volatile("ld.global.nc.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=f"(v[0]), "=f"(v[1]), "=f"(v[2]), "=f"(v[3]), "=f"(v[4]), "=f"(v[5]), "=f"(v[6]), "=f"(v[7])
                 : "l"(p));
}
__device__ __forceinline__ void ld256(const double *p, double *v)
{
    asm volatile("ld.global.nc.v4.f64 {%0,%1,%2,%3}, [%4];" : "=d"(v[0]), "=d"(v[1]), "=d"(v[2]), "=d"(v[3]) : "l"(p));
}
__device__ __forceinline__ void st256(float *p, const float *v)
{
    asm volatile("st.global.v8.f32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"l"(p), "f"(v[0]), "f"(v[1]), "f"(v[2]), "f"(v[3]),
                 "f"(v[4]), "f"(v[5]), "f"(v[6]), "f"(v[7])
                 : "memory");
}
__device__ __forceinline__ void st256(double *p, const double *v)
{
    asm volatile("st.global.v4.f64 [%0], {%1,%2,%3,%4};" ::"l"(p), "d"(v[0]), "d"(v[1]), "d"(v[2]), "d"(v[3]) : "memory");
}
// whole record (fp32: one 256-bit access, fp64: two)
__device__ __forceinline__ void ldRec(const Rec8<float> *r, u32 j, float *v) { ld256(r[j].a, v); }
__device__ __forceinline__ void ldRec(const Rec8<double> *r, u32 j, double *v)
{
    ld256(r[j].a, v);
    ld256(r[j].a + 4, v + 4);
}
// {x, y, z, p, Vol}: fp32 gets the whole line anyway, fp64 adds one 8-byte load from the second sector
__device__ __forceinline__ void ldRecXVp(const Rec8<float> *r, u32 j, float *v) { ld256(r[j].a, v); }
__device__ __forceinline__ void ldRecXVp(const Rec8<double> *r, u32 j, double *v)
{
    ld256(r[j].a, v);
    v[4] = __ldg(r[j].a + 4);
}
// {x, y, z, p} only (uniform-Vol bodies): one sector in either precision
__device__ __forceinline__ void ldRecXp(const Rec8<float> *r, u32 j, float *v) { ld256(r[j].a, v); }
__device__ __forceinline__ void ldRecXp(const Rec8<double> *r, u32 j, double *v) { ld256(r[j].a, v); }
__device__ __forceinline__ void stRec(Rec8<float> *r, int64_t i, const float *v) { st256(r[i].a, v); }
__device__ __forceinline__ void stRec(Rec8<double> *r, int64_t i, const double *v)
{
    st256(r[i].a, v);
    st256(r[i].a + 4, v + 4);
}

// lane-per-particle view of the sliced-ELL list (kernels outside the acoustic sub-step)
__device__ __forceinline__ const u32 *listBase(const u32 *list, const u32 *slice_off, int64_t i)
{
    return list + u64(__ldg(slice_off + i / kP)) * 32u + u32(i % kP) * u32(kS);
}
__device__ __forceinline__ u32 listEntry(const u32 *base, u32 e) { return __ldg(base + size_t(e / kS) * 32u + (e % kS)); }
__device__ __forceinline__ u32 *listSlot(u32 *list, const u32 *slice_off, int64_t i, u32 e)
{
    return list + u64(slice_off[i / kP]) * 32u + u32(i % kP) * u32(kS) + size_t(e / kS) * 32u + (e % kS);
}
// sum / max over the kS lanes that share one particle (all 32 lanes of the warp must call)
template <class R>
__device__ __forceinline__ R groupSum(R v)
{
#pragma unroll
    for (int o = kS / 2; o > 0; o >>= 1)
        v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

template <class R>
__device__ __forceinline__ void loadX(const typename Real4<R>::type *p, u32 i, R *x, R &w)
{
    typename Real4<R>::type v = ldg4(p + i);
    x[0] = v.x;
    x[1] = v.y;
    x[2] = v.z;
    w = v.w;
}

// Pair geometry recomputed from the frozen positions.  r = |d|, e = d/r (0 when r == 0, as
// d/(r+TinyReal) gives), dW per kernel_wendland_c2.cpp:28-35, W per :17-26.
template <int D, class R>
struct Pair
{
    R e[D];
    R r, dW;
};
template <int D, class R>
__device__ __forceinline__ void pairGeometry(const R *xi, const R *xj, const KernelP<R> &k, Pair<D, R> &g)
{
    R d[D];
    R r2 = R(0);
#pragma unroll
    for (int a = 0; a < D; ++a)
    {
        d[a] = xi[a] - xj[a];
        r2 += d[a] * d[a];
    }
    R rinv = r2 > R(0) ? rsqrt_r(r2) : R(0);
    g.r = r2 * rinv;
#pragma unroll
    for (int a = 0; a < D; ++a)
        g.e[a] = d[a] * rinv;
    R q = g.r * k.inv_h;
    R t = q - R(2.0);
    g.dW = k.factor_dW * (R(0.625) * (t * t * t) * q);
}
template <class R>
__device__ __forceinline__ R kernelW(R r, const KernelP<R> &k)
{
    R q = r * k.inv_h;
    R t = R(1.0) - R(0.5) * q;
    R t2 = t * t;
    return k.factor_W * ((t2 * t2) * (R(1.0) + R(2.0) * q));
}

// minimum-image translation of x_j for periodic axes (reproduces the ghost copy x_j ± L that the
// reference inserts into the cell lists: domain_bounding.cpp:18-65)
template <int D, class R>
__device__ __forceinline__ void periodicImage(const R *xi, R *xj, const PeriodicP<D, R> &per)
{
    if (per.mask == 0)
        return;
#pragma unroll
    for (int a = 0; a < D; ++a)
        if (per.mask & (1 << a))
        {
            R t = xi[a] - xj[a];
            if (t > per.half_L[a])
                xj[a] = xj[a] + per.L[a];
            else if (t < -per.half_L[a])
                xj[a] = xj[a] - per.L[a];
        }
}

__device__ __forceinline__ void atomicMaxNonNeg(unsigned long long *slot, double v)
{
    atomicMax(slot, (unsigned long long)__double_as_longlong(v));
}
__device__ __forceinline__ void atomicMinNonNeg(unsigned long long *slot, double v)
{
    atomicMin(slot, (unsigned long long)__double_as_longlong(v));
}
template <class R>
__device__ __forceinline__ R warpMax(R v)
{
#pragma unroll
    for (int o = 16; o > 0; o >>= 1)
        v = fmax(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}
template <class R>
__device__ __forceinline__ R warpMin(R v)
{
#pragma unroll
    for (int o = 16; o > 0; o >>= 1)
        v = fmin(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}
template <class R>
__device__ __forceinline__ R warpSum(R v)
{
#pragma unroll
    for (int o = 16; o > 0; o >>= 1)
        v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
#endif // __CUDACC__
// ---------------------------------------------------------------------------------------------
// hit masks of the masked neighbour search (relation.cu: k_search_mask; readers: k_expand_mask, tile.cuh k_union_mask)
// a particle's candidates are 3^(D-1) runs of consecutive target particles (the cells c-1..c+1 along the
// fastest axis of one neighbouring column); kRunWords 32-bit words of hit bits per run
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

// fills lists[k].list from the hit masks if this build has not done so yet (relation.cu)
sphb200_status ensureList(sphb200_relation *r, size_t k);

} // namespace sphb
