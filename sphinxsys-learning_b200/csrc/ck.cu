// CK-flavour local dynamics used by the reference's SYCL dam-break tests, as sm_100a kernels:
//   CompressionSummation<Inner / Contact> + DensityRegularization<..., FreeSurface | Internal>
//        shared_ck/particle_dynamics/fluid_dynamics/density_regularization.hpp:42-111, .h:146-180
//   LinearCorrectionMatrix<Inner<WithUpdate>> + <Contact>  (alpha-weighted Tikhonov inverse)
//        shared_ck/particle_dynamics/general_dynamics/kernel_correction_ck.hpp:36-96, common/vector_functions.h:198-203
//   FreeSurfaceIndicationCK<Inner<WithUpdate>> + <Contact>
//        shared_ck/particle_dynamics/general_dynamics/surface_indication/surface_indication_ck.hpp:55-170
// evaluated with the CK smoothing kernel: the 20-interval cubic-Lagrange TABLE of Wendland C2
// (shared_ck/smoothing_kernel/kernel_tabulated_ck.h:37-71, .cpp:6-27; neighbor_method.hpp:17-52,73-120), which differs
// from the analytic kernel of the classic path at the 1e-4 level - hence its own kernels rather than a flag on the
// classic ones.  Neighbours come from the relations' index lists (ensureList), positions are the configuration-time
// ones (the CK tests commit Position once per advection step, right before the relation update).
// The reference's int variables Indicator / PreviousSurfaceIndicator are Real-typed 0 / 1 here (the registry is Real-typed).
#include "internal.cuh"
#include "smat.cuh"

using namespace sphb;

namespace sphb
{
template <class R>
struct TabKernelP
{
    R dq, dq0, dq1, dq2, dq3, dimfac, inv_h, src_inv_h;
    R w[24], dw[24];
};
template <int D, class R>
static TabKernelP<R> makeTabKernel(R h_src, R h_tar)
{
    TabKernelP<R> k;
    k.dq = R(2.0) / R(20);
    for (int i = 0; i < 24; ++i)
    {
        const R q = R(i - 1) * k.dq;
        k.w[i] = R(std::pow(R(1.0) - R(0.5) * q, 4)) * (R(1.0) + R(2.0) * q);
        k.dw[i] = R(0.625) * R(std::pow(q - R(2.0), 3)) * q;
    }
    k.dq0 = (R(-1.0) * k.dq) * (R(-2.0) * k.dq) * (R(-3.0) * k.dq);
    k.dq1 = k.dq * (R(-1.0) * k.dq) * (R(-2.0) * k.dq);
    k.dq2 = (R(2.0) * k.dq) * k.dq * (R(-1.0) * k.dq);
    k.dq3 = (R(3.0) * k.dq) * (R(2.0) * k.dq) * k.dq;
    const R h = h_src, ih = R(1.0) / h;
    k.dimfac = D == 2 ? (ih * ih * R(7.0) / (R(4.0) * R(kPi))) * h * h : (ih * ih * ih * R(21.0) / (R(16.0) * R(kPi))) * h * h * h;
    k.src_inv_h = R(1.0) / h_src;
    k.inv_h = R(1.0) / std::max(h_src, h_tar);
    return k;
}
template <class R>
__device__ __forceinline__ R tabCubic(const TabKernelP<R> &k, const R *data, R q)
{
    const int location = int(floor(q / k.dq));
    const int i = location + 1;
    const R f1 = q - R(location) * k.dq, f0 = f1 + k.dq, f2 = f1 - k.dq, f3 = f1 - R(2) * k.dq;
    return (f1 * f2 * f3) / k.dq0 * data[i - 1] + (f0 * f2 * f3) / k.dq1 * data[i] + (f0 * f1 * f3) / k.dq2 * data[i + 1] +
           (f0 * f1 * f2) / k.dq3 * data[i + 2];
}
template <int D, class R>
__device__ __forceinline__ R ipowD(R x, int extra)
{
    R r = R(1);
#pragma unroll
    for (int k = 0; k < D; ++k)
        r *= x;
    return extra ? r * x : r;
}
template <int D, class R>
__device__ __forceinline__ R tabW0(const TabKernelP<R> &k) { return ipowD<D, R>(k.src_inv_h, 0) * k.dimfac * tabCubic<R>(k, k.w, R(0)); }
template <int D, class R>
__device__ __forceinline__ R tabW(const TabKernelP<R> &k, R r) { return ipowD<D, R>(k.inv_h, 0) * k.dimfac * tabCubic<R>(k, k.w, r * k.inv_h); }
template <int D, class R>
__device__ __forceinline__ R tabdW(const TabKernelP<R> &k, R r) { return ipowD<D, R>(k.inv_h, 1) * k.dimfac * tabCubic<R>(k, k.dw, r * k.inv_h); }

struct CkList
{
    const u32 *count, *slice_off, *list;
};
static CkList ckListOf(const NeighborList &L)
{
    CkList l;
    l.count = L.count;
    l.slice_off = L.slice_off;
    l.list = L.list;
    return l;
}
template <int D, class R>
__device__ __forceinline__ R ckDisp(const R *xi, const typename Real4<R>::type *__restrict__ xV, u32 j, R *d, R &Vj)
{
    const typename Real4<R>::type v = ldg4(xV + j);
    const R xj[3] = {v.x, v.y, v.z};
    Vj = v.w;
#pragma unroll
    for (int a = 0; a < D; ++a)
        d[a] = xi[a] - xj[a];
    return sqrt_r(sqnormExact<D, R>(d));
}
#define CK_INDEX()                                                                                       \
    const int64_t i = (i0 & ~int64_t(31)) + blockIdx.x * int64_t(blockDim.x) + threadIdx.x;              \
    if (i < i0 || i >= n)                                                                                \
        return;                                                                                          \
    R xi[D];                                                                                             \
    {                                                                                                    \
        const typename Real4<R>::type v_ = ldg4(xV + i);                                                 \
        const R x3_[3] = {v_.x, v_.y, v_.z};                                                             \
        for (int a_ = 0; a_ < D; ++a_)                                                                   \
            xi[a_] = x3_[a_];                                                                            \
    }

// ---- CompressionSummation: first = 1 starts the sum with W0 * Vol_ref_i, 0 adds a contact body's part
template <int D, class R>
__global__ void __launch_bounds__(128)
    k_ck_compression(int64_t i0, int64_t n, int first, const typename Real4<R>::type *__restrict__ xV,
                     const typename Real4<R>::type *__restrict__ xV_t, CkList in, TabKernelP<R> k, const R *__restrict__ vref_i,
                     const R *__restrict__ vref_t, R *__restrict__ csum)
{
    CK_INDEX();
    R s = first ? tabW0<D, R>(k) * vref_i[i] : csum[i];
    const u32 cnt = in.count[i];
    const u32 *lst = listBase(in.list, in.slice_off, i);
    for (u32 m = 0; m < cnt; ++m)
    {
        const u32 j = listEntry(lst, m);
        R d[D], Vj;
        const R r = ckDisp<D, R>(xi, xV_t, j, d, Vj);
        s += tabW<D, R>(k, r) * vref_t[j];
    }
    csum[i] = s;
}
template <class R>
__global__ void __launch_bounds__(256)
    k_ck_regularize(int64_t i0, int64_t n, int free_surface, R rho0, const R *__restrict__ csum, R *__restrict__ comp, R *__restrict__ rho)
{
    const int64_t i = i0 + blockIdx.x * int64_t(blockDim.x) + threadIdx.x;
    if (i >= n)
        return;
    const R c = free_surface ? fmax(csum[i], R(1)) : csum[i];
    comp[i] = c;
    rho[i] = c * rho0;
}

// ---- LinearCorrectionMatrix: B -= r_ij (nablaW_ij V_j)^T over inner (first = 1) and contact lists, then the update
template <int D, class R>
__global__ void __launch_bounds__(128)
    k_ck_correction_sum(int64_t i0, int64_t n, int first, const typename Real4<R>::type *__restrict__ xV,
                        const typename Real4<R>::type *__restrict__ xV_t, CkList in, TabKernelP<R> k, R *__restrict__ B)
{
    CK_INDEX();
    SMat<D, R> M = first ? SMat<D, R>::zero() : SMat<D, R>::load(B + i * D * D);
    const u32 cnt = in.count[i];
    const u32 *lst = listBase(in.list, in.slice_off, i);
    for (u32 m = 0; m < cnt; ++m)
    {
        const u32 j = listEntry(lst, m);
        R d[D], Vj;
        const R r = ckDisp<D, R>(xi, xV_t, j, d, Vj);
        const R dw = tabdW<D, R>(k, r);
#pragma unroll
        for (int a = 0; a < D; ++a)
#pragma unroll
            for (int c = 0; c < D; ++c)
                M.a[a][c] -= d[a] * ((dw * (d[c] / r)) * Vj);
    }
    M.store(B + i * D * D);
}
template <int D, class R>
__global__ void __launch_bounds__(128) k_ck_correction_update(int64_t i0, int64_t n, R alpha, R *__restrict__ B)
{
    const int64_t i = i0 + blockIdx.x * int64_t(blockDim.x) + threadIdx.x;
    if (i >= n)
        return;
    SMat<D, R> M = SMat<D, R>::load(B + i * D * D);
    const R det = M.det();
    const R det_sqr = fmax(alpha - det, R(0));
    SMat<D, R> Mt = M.T();
    SMat<D, R> inv = (Mt * M + SMat<D, R>::identity() * R(1.0e-8)).inverse() * Mt; // inverseTikhonov(B, SqrtEps)
    const R weight = det / (det + det_sqr);
    (inv * weight + SMat<D, R>::identity() * (R(1.0) - weight)).store(B + i * D * D);
}

// ---- FreeSurfaceIndicationCK
template <int D, class R>
__global__ void __launch_bounds__(128)
    k_ck_posdiv(int64_t i0, int64_t n, int inner, const typename Real4<R>::type *__restrict__ xV,
                const typename Real4<R>::type *__restrict__ xV_t, CkList in, TabKernelP<R> k, R threshold, const R *__restrict__ prev,
                R *__restrict__ pdiv)
{
    CK_INDEX();
    R pd = R(0);
    bool near_prev = false;
    const u32 cnt = in.count[i];
    const u32 *lst = listBase(in.list, in.slice_off, i);
    for (u32 m = 0; m < cnt; ++m)
    {
        const u32 j = listEntry(lst, m);
        R d[D], Vj;
        const R r = ckDisp<D, R>(xi, xV_t, j, d, Vj);
        pd -= tabdW<D, R>(k, r) * Vj * r;
        if (inner)
            near_prev = near_prev || prev[j] == R(1);
    }
    if (inner)
    {
        if (pd < threshold && prev[i] != R(1) && !near_prev)
            pd = R(2.0) * threshold;
        pdiv[i] = pd;
    }
    else
        pdiv[i] += pd;
}
template <int D, class R>
__global__ void __launch_bounds__(128)
    k_ck_indicator(int64_t i0, int64_t n, const typename Real4<R>::type *__restrict__ xV, CkList in, R threshold, R h,
                   const R *__restrict__ pdiv, R *__restrict__ out)
{
    CK_INDEX();
    R v = R(1);
    if (pdiv[i] > threshold)
    {
        bool very_near = false;
        const u32 cnt = in.count[i];
        const u32 *lst = listBase(in.list, in.slice_off, i);
        for (u32 m = 0; m < cnt && !very_near; ++m)
        {
            const u32 j = listEntry(lst, m);
            R d[D], Vj;
            const R r = ckDisp<D, R>(xi, xV, j, d, Vj);
            very_near = pdiv[j] < threshold && r < h;
        }
        if (!very_near)
            v = R(0);
    }
    out[i] = v;
}
template <class R>
__global__ void __launch_bounds__(256) k_ck_fill(int64_t n, R value, R *__restrict__ a)
{
    const int64_t i = blockIdx.x * int64_t(blockDim.x) + threadIdx.x;
    if (i < n)
        a[i] = value;
}

// ---------------------------------------------------------------------------------------------------------------
// CK acoustic sub-step with LinearCorrectionCK (the scheme of tests_sycl/.../test_3d_dambreak_sycl and its 2-D twin):
//   AdvectionStepSetup / UpdateParticlePosition     fluid_time_step_ck.h:145-148,173-175
//   AcousticTimeStepCK / AdvectionTimeStepCK        fluid_time_step_ck.hpp:31-35,52-56 ; .cpp:23-25 ; .h:106-108
//   AcousticStep1stHalf<Inner<OneLevel, AcousticRiemannSolverCK, LinearCorrectionCK>> + <Contact<Wall, ...>>
//                                                   acoustic_step_1st_half.hpp:66-74,89-111,122-127,157-180
//   AcousticStep2ndHalf<...>                        acoustic_step_2nd_half.hpp:33-38,53-73,83-89,117-137
// State: Compression / CompressionRate / Displacement (Position is committed once per advection step, so the pair
// geometry is the configuration's, as in the reference where the sub-steps never touch Position).
// ---------------------------------------------------------------------------------------------------------------
template <int D, class R>
struct CkRiemann
{
    R inv_rho0c0, rho0c0, inv_c0; // riemann_solver.h:55-123 with both sides the same fluid
};
template <class R>
__device__ __forceinline__ R ckPJump(R rho0c0, R inv_c0, R u) { return rho0c0 * u * fmin(R(3.0) * (inv_c0 * fmax(u, R(0))), R(1)); }

template <int D, class R>
__global__ void __launch_bounds__(256) k_ck_setup(int64_t i0, int64_t n, const R *__restrict__ mass, const R *__restrict__ rho,
                                                  R *__restrict__ Vol, R *__restrict__ dpos)
{
    const int64_t i = i0 + blockIdx.x * int64_t(blockDim.x) + threadIdx.x;
    if (i >= n)
        return;
    Vol[i] = mass[i] / rho[i];
#pragma unroll
    for (int a = 0; a < D; ++a)
        dpos[i * D + a] = R(0);
}
template <class R>
__global__ void __launch_bounds__(256) k_ck_axpy(int64_t e0, int64_t e1, R *__restrict__ y, const R *__restrict__ x)
{
    const int64_t e = e0 + blockIdx.x * int64_t(blockDim.x) + threadIdx.x;
    if (e < e1)
        y[e] += x[e];
}
// kind 0: max(c0 + |v|, sqrt(4 h |F + Fp| / m));  kind 1: max |v|^2
template <int D, class R>
__global__ void __launch_bounds__(256)
    k_ck_dt_max(int64_t i0, int64_t n, int kind, R c0, R h, const R *__restrict__ vel, const R *__restrict__ force, const R *__restrict__ fprior,
                const R *__restrict__ mass, unsigned long long *__restrict__ slot)
{
    int64_t i = i0 + blockIdx.x * int64_t(blockDim.x) + threadIdx.x;
    R s = R(0);
    for (; i < n; i += int64_t(gridDim.x) * blockDim.x)
    {
        R v[D], f[D];
#pragma unroll
        for (int a = 0; a < D; ++a)
        {
            v[a] = vel[i * D + a];
            f[a] = kind == 0 ? force[i * D + a] + fprior[i * D + a] : R(0);
        }
        if (kind == 0)
        {
            const R acc_scale = sqrt_r(R(4.0) * h * sqrt_r(sqnormExact<D, R>(f)) / mass[i]);
            s = fmax(s, fmax(c0 + sqrt_r(sqnormExact<D, R>(v)), acc_scale));
        }
        else
            s = fmax(s, sqnormExact<D, R>(v));
    }
    s = warpMax<R>(s);
    if ((threadIdx.x & 31) == 0)
        atomicMaxNonNeg(slot, double(s));
}
template <int D, class R>
__global__ void __launch_bounds__(256)
    k_ck_first_init(int64_t i0, int64_t n, R dt, R rho0, R c0, R *__restrict__ comp, const R *__restrict__ crate, R *__restrict__ rho,
                    R *__restrict__ p, const R *__restrict__ vel, R *__restrict__ dpos)
{
    const int64_t i = i0 + blockIdx.x * int64_t(blockDim.x) + threadIdx.x;
    if (i >= n)
        return;
    const R c = comp[i] + R(0.5) * dt * crate[i];
    comp[i] = c;
    const R r = c * rho0;
    rho[i] = r;
    p[i] = rho0 * c0 * c0 * (r / rho0 - R(1.0));
#pragma unroll
    for (int a = 0; a < D; ++a)
        dpos[i * D + a] += vel[i * D + a] * dt * R(0.5);
}
template <int D, class R>
__device__ __forceinline__ void ckMatVec(const R *__restrict__ B, const R *e, R *out)
{
#pragma unroll
    for (int a = 0; a < D; ++a)
    {
        out[a] = R(0);
#pragma unroll
        for (int c = 0; c < D; ++c)
            out[a] += B[a * D + c] * e[c];
    }
}
template <int D, class R>
__global__ void __launch_bounds__(128)
    k_ck_first_inner(int64_t i0, int64_t n, const typename Real4<R>::type *__restrict__ xV, CkList in, TabKernelP<R> k, CkRiemann<D, R> rs,
                     const R *__restrict__ p, const R *__restrict__ B, const R *__restrict__ Vol, const R *__restrict__ comp,
                     R *__restrict__ crate, R *__restrict__ force)
{
    CK_INDEX();
    const R pi = p[i];
    R Bi[D * D];
#pragma unroll
    for (int q = 0; q < D * D; ++q)
        Bi[q] = B[i * D * D + q];
    R fs[D] = {}, diss = R(0);
    const u32 cnt = in.count[i];
    const u32 *lst = listBase(in.list, in.slice_off, i);
    for (u32 m = 0; m < cnt; ++m)
    {
        const u32 j = listEntry(lst, m);
        R d[D], Vj, e[D], M[D * D], Me[D];
        const R r = ckDisp<D, R>(xi, xV, j, d, Vj);
        const R dWV = tabdW<D, R>(k, r) * Vj;
        const R pj = p[j];
#pragma unroll
        for (int a = 0; a < D; ++a)
            e[a] = d[a] / r;
#pragma unroll
        for (int q = 0; q < D * D; ++q)
            M[q] = R(0.5) * (B[int64_t(j) * D * D + q] * pi + Bi[q] * pj);
        ckMatVec<D, R>(M, e, Me);
#pragma unroll
        for (int a = 0; a < D; ++a)
            fs[a] -= Me[a] * R(2.0) * dWV;
        diss += (pi - pj) * rs.inv_rho0c0 * dWV;
    }
    const R Vi = Vol[i];
#pragma unroll
    for (int a = 0; a < D; ++a)
        force[i * D + a] += fs[a] * Vi;
    crate[i] = diss * comp[i];
}
template <int D, class R>
__global__ void __launch_bounds__(128)
    k_ck_first_wall(int64_t i0, int64_t n, const typename Real4<R>::type *__restrict__ xV, const typename Real4<R>::type *__restrict__ xV_t,
                    CkList in, TabKernelP<R> k, CkRiemann<D, R> rs, const R *__restrict__ p, const R *__restrict__ rho, const R *__restrict__ mass,
                    const R *__restrict__ fprior, const R *__restrict__ wall_acc, const R *__restrict__ B, const R *__restrict__ Vol,
                    const R *__restrict__ comp, R *__restrict__ crate, R *__restrict__ force)
{
    CK_INDEX();
    const u32 cnt = in.count[i];
    if (cnt == 0)
        return;
    const R pi = p[i], rhoi = rho[i];
    R Bi[D * D], acc[D];
#pragma unroll
    for (int q = 0; q < D * D; ++q)
        Bi[q] = B[i * D * D + q];
#pragma unroll
    for (int a = 0; a < D; ++a)
        acc[a] = fprior[i * D + a] / mass[i];
    R fs[D] = {}, diss = R(0);
    const u32 *lst = listBase(in.list, in.slice_off, i);
    for (u32 m = 0; m < cnt; ++m)
    {
        const u32 j = listEntry(lst, m);
        R d[D], Vj, e[D], Be[D];
        const R r = ckDisp<D, R>(xi, xV_t, j, d, Vj);
        const R dWV = tabdW<D, R>(k, r) * Vj;
        R face = R(0);
#pragma unroll
        for (int a = 0; a < D; ++a)
        {
            e[a] = d[a] / r;
            face += (acc[a] - (wall_acc ? wall_acc[int64_t(j) * D + a] : R(0))) * (-e[a]);
        }
        const R p_wall = pi + rhoi * r * fmax(R(0), face);
        ckMatVec<D, R>(Bi, e, Be);
#pragma unroll
        for (int a = 0; a < D; ++a)
            fs[a] -= (pi + p_wall) * Be[a] * dWV;
        diss += (pi - p_wall) * rs.inv_rho0c0 * dWV;
    }
    const R Vi = Vol[i];
#pragma unroll
    for (int a = 0; a < D; ++a)
        force[i * D + a] += fs[a] * Vi;
    crate[i] += diss * comp[i];
}
template <int D, class R>
__global__ void __launch_bounds__(256) k_ck_first_update(int64_t i0, int64_t n, R dt, const R *__restrict__ mass, const R *__restrict__ fprior,
                                                         const R *__restrict__ force, R *__restrict__ vel)
{
    const int64_t i = i0 + blockIdx.x * int64_t(blockDim.x) + threadIdx.x;
    if (i >= n)
        return;
    const R m = mass[i];
#pragma unroll
    for (int a = 0; a < D; ++a)
        vel[i * D + a] += (fprior[i * D + a] + force[i * D + a]) / m * dt;
}
template <int D, class R>
__global__ void __launch_bounds__(128)
    k_ck_second_inner(int64_t i0, int64_t n, R dt, const typename Real4<R>::type *__restrict__ xV, CkList in, TabKernelP<R> k, CkRiemann<D, R> rs,
                      const R *__restrict__ vel, const R *__restrict__ B, const R *__restrict__ Vol, const R *__restrict__ comp,
                      R *__restrict__ crate, R *__restrict__ force, R *__restrict__ dpos)
{
    CK_INDEX();
    R vi[D], Bi[D * D];
#pragma unroll
    for (int a = 0; a < D; ++a)
    {
        vi[a] = vel[i * D + a];
        dpos[i * D + a] += vi[a] * dt * R(0.5); // InitializeKernel
    }
#pragma unroll
    for (int q = 0; q < D * D; ++q)
        Bi[q] = B[i * D * D + q];
    R div = R(0), pd[D] = {};
    const u32 cnt = in.count[i];
    const u32 *lst = listBase(in.list, in.slice_off, i);
    for (u32 m = 0; m < cnt; ++m)
    {
        const u32 j = listEntry(lst, m);
        R d[D], Vj, e[D], Be[D];
        const R r = ckDisp<D, R>(xi, xV, j, d, Vj);
        const R dWV = tabdW<D, R>(k, r) * Vj;
#pragma unroll
        for (int a = 0; a < D; ++a)
            e[a] = d[a] / r;
        ckMatVec<D, R>(Bi, e, Be);
        R dvBe = R(0), u = R(0);
#pragma unroll
        for (int a = 0; a < D; ++a)
        {
            const R vj = vel[int64_t(j) * D + a];
            const R vave = R(0.5) * (vi[a] + vj);
            dvBe += R(2.0) * (vi[a] - vave) * Be[a];
            u += (vi[a] - vj) * e[a];
        }
        div += dvBe * dWV;
        const R pj = ckPJump<R>(rs.rho0c0, rs.inv_c0, u) * dWV;
#pragma unroll
        for (int a = 0; a < D; ++a)
            pd[a] += pj * e[a];
    }
    crate[i] += div * comp[i];
    const R Vi = Vol[i];
#pragma unroll
    for (int a = 0; a < D; ++a)
        force[i * D + a] = pd[a] * Vi;
}
template <int D, class R>
__global__ void __launch_bounds__(128)
    k_ck_second_wall(int64_t i0, int64_t n, const typename Real4<R>::type *__restrict__ xV, const typename Real4<R>::type *__restrict__ xV_t,
                     CkList in, TabKernelP<R> k, CkRiemann<D, R> rs, const R *__restrict__ vel, const R *__restrict__ wall_vel,
                     const R *__restrict__ wall_n, const R *__restrict__ B, const R *__restrict__ Vol, const R *__restrict__ comp,
                     R *__restrict__ crate, R *__restrict__ force)
{
    CK_INDEX();
    const u32 cnt = in.count[i];
    if (cnt == 0)
        return;
    R vi[D], Bi[D * D];
#pragma unroll
    for (int a = 0; a < D; ++a)
        vi[a] = vel[i * D + a];
#pragma unroll
    for (int q = 0; q < D * D; ++q)
        Bi[q] = B[i * D * D + q];
    R div = R(0), pd[D] = {};
    const u32 *lst = listBase(in.list, in.slice_off, i);
    for (u32 m = 0; m < cnt; ++m)
    {
        const u32 j = listEntry(lst, m);
        R d[D], Vj, e[D], Be[D], vd[D], nj[D];
        const R r = ckDisp<D, R>(xi, xV_t, j, d, Vj);
        const R dWV = tabdW<D, R>(k, r) * Vj;
        R en = R(0);
#pragma unroll
        for (int a = 0; a < D; ++a)
        {
            e[a] = d[a] / r;
            nj[a] = wall_n[int64_t(j) * D + a];
            vd[a] = R(2.0) * (vi[a] - (wall_vel ? wall_vel[int64_t(j) * D + a] : R(0)));
            en += e[a] * nj[a];
        }
        ckMatVec<D, R>(Bi, e, Be);
        const R sgn = en > R(0) ? R(1) : (en < R(0) ? R(-1) : R(0));
        R dvBe = R(0), u = R(0);
#pragma unroll
        for (int a = 0; a < D; ++a)
        {
            dvBe += vd[a] * Be[a];
            u += vd[a] * (sgn * nj[a]);
        }
        div += dvBe * dWV;
        const R pj = ckPJump<R>(rs.rho0c0, rs.inv_c0, u) * dWV;
#pragma unroll
        for (int a = 0; a < D; ++a)
            pd[a] += pj * (sgn * nj[a]);
    }
    crate[i] += div * comp[i];
    const R Vi = Vol[i];
#pragma unroll
    for (int a = 0; a < D; ++a)
        force[i * D + a] += pd[a] * Vi;
}
template <class R>
__global__ void __launch_bounds__(256)
    k_ck_second_update(int64_t i0, int64_t n, R dt, R rho0, R *__restrict__ comp, const R *__restrict__ crate, R *__restrict__ rho)
{
    const int64_t i = i0 + blockIdx.x * int64_t(blockDim.x) + threadIdx.x;
    if (i >= n)
        return;
    const R c = comp[i] + R(0.5) * dt * crate[i];
    comp[i] = c;
    rho[i] = c * rho0;
}

// ---------------------------------------------------------------------------------------------------------------
// Force of the fluid on a structure: source of the contact relation = the solid, targets = fluid bodies.
//   FSI::PressureForceFromFluid<Contact<WithUpdate, ...>>::InteractKernel::interact   force_on_structure.hpp:113-137
//   FSI::ViscousForceFromFluid<Contact<WithUpdate, ...>>::InteractKernel::interact    force_on_structure.hpp:61-79
//   WithUpdate = ForcePriorCK::UpdateKernel                                            force_prior_ck.h:55-59
// B_t: the fluid's LinearCorrectionMatrix (LinearCorrectionCK) or null (NoKernelCorrectionCK: identity).
// ---------------------------------------------------------------------------------------------------------------
template <int D, class R>
__global__ void __launch_bounds__(128)
    k_ck_fsi_pressure(int64_t i0, int64_t n, int first, const typename Real4<R>::type *__restrict__ xV,
                      const typename Real4<R>::type *__restrict__ xV_t, CkList in, TabKernelP<R> k, R rho0c0, R inv_c0,
                      const R *__restrict__ vel_ave, const R *__restrict__ acc_ave, const R *__restrict__ nrm, const R *__restrict__ Vol,
                      const R *__restrict__ t_vel, const R *__restrict__ t_rho, const R *__restrict__ t_mass, const R *__restrict__ t_p,
                      const R *__restrict__ t_fprior, const R *__restrict__ B_t, R *__restrict__ force)
{
    CK_INDEX();
    R vi[D], ai[D], ni[D], acc[D] = {};
#pragma unroll
    for (int a = 0; a < D; ++a)
    {
        vi[a] = vel_ave ? vel_ave[i * D + a] : R(0);
        ai[a] = acc_ave ? acc_ave[i * D + a] : R(0);
        ni[a] = nrm[i * D + a];
    }
    const u32 cnt = in.count[i];
    const u32 *lst = listBase(in.list, in.slice_off, i);
    for (u32 m = 0; m < cnt; ++m)
    {
        const u32 j = listEntry(lst, m);
        R d[D], Vj, e[D], ce[D], nf[D];
        const R r = ckDisp<D, R>(xi, xV_t, j, d, Vj);
#pragma unroll
        for (int a = 0; a < D; ++a)
            e[a] = d[a] / r;
        if (B_t)
            ckMatVec<D, R>(B_t + int64_t(j) * D * D, e, ce);
        else
        {
#pragma unroll
            for (int a = 0; a < D; ++a)
                ce[a] = e[a];
        }
        const R mj = t_mass[j], pj = t_p[j];
        R face = R(0), cen = R(0);
#pragma unroll
        for (int a = 0; a < D; ++a)
        {
            face += (t_fprior[int64_t(j) * D + a] / mj - ai[a]) * e[a];
            cen += ce[a] * ni[a];
        }
        const R p_in_wall = pj + t_rho[j] * r * fmax(R(0), face);
        const R sgn = cen < R(0) ? R(-1) : (cen > R(0) ? R(1) : R(0));
        R u = R(0);
#pragma unroll
        for (int a = 0; a < D; ++a)
        {
            nf[a] = -sgn * ni[a];
            u += R(2.0) * (t_vel[int64_t(j) * D + a] - vi[a]) * nf[a];
        }
        const R jump = ckPJump<R>(rho0c0, inv_c0, u), dWV = tabdW<D, R>(k, r) * Vj;
#pragma unroll
        for (int a = 0; a < D; ++a)
            acc[a] -= (jump * nf[a] + (p_in_wall + pj) * ce[a]) * dWV;
    }
    const R Vi = Vol[i];
#pragma unroll
    for (int a = 0; a < D; ++a)
        force[i * D + a] = (first ? R(0) : force[i * D + a]) + acc[a] * Vi;
}
template <int D, class R>
__global__ void __launch_bounds__(128)
    k_ck_fsi_viscous(int64_t i0, int64_t n, int first, const typename Real4<R>::type *__restrict__ xV,
                     const typename Real4<R>::type *__restrict__ xV_t, CkList in, TabKernelP<R> k, R mu, R h_sq, const R *__restrict__ vel_ave,
                     const R *__restrict__ Vol, const R *__restrict__ t_vel, const R *__restrict__ B_t, R *__restrict__ force)
{
    CK_INDEX();
    R vi[D], acc[D] = {};
#pragma unroll
    for (int a = 0; a < D; ++a)
        vi[a] = vel_ave ? vel_ave[i * D + a] : R(0);
    const u32 cnt = in.count[i];
    const u32 *lst = listBase(in.list, in.slice_off, i);
    for (u32 m = 0; m < cnt; ++m)
    {
        const u32 j = listEntry(lst, m);
        R d[D], Vj, e[D], ce[D];
        const R r = ckDisp<D, R>(xi, xV_t, j, d, Vj);
#pragma unroll
        for (int a = 0; a < D; ++a)
            e[a] = d[a] / r;
        if (B_t)
            ckMatVec<D, R>(B_t + int64_t(j) * D * D, e, ce);
        else
        {
#pragma unroll
            for (int a = 0; a < D; ++a)
                ce[a] = e[a];
        }
        const R denom = sqnormExact<D, R>(d) + R(0.01) * h_sq;
        R rce = R(0);
#pragma unroll
        for (int a = 0; a < D; ++a)
            rce += d[a] * ce[a];
        const R w = rce * mu * tabdW<D, R>(k, r) * Vj;
#pragma unroll
        for (int a = 0; a < D; ++a)
            acc[a] += w * ((vi[a] - t_vel[int64_t(j) * D + a]) / denom);
    }
    const R Vi = Vol[i];
#pragma unroll
    for (int a = 0; a < D; ++a)
        force[i * D + a] = (first ? R(0) : force[i * D + a]) + acc[a] * Vi;
}
template <class R>
__global__ void __launch_bounds__(256)
    k_ck_force_prior_update(int64_t e0, int64_t e1, R *__restrict__ fprior, const R *__restrict__ cur, R *__restrict__ prev)
{
    const int64_t e = e0 + blockIdx.x * int64_t(blockDim.x) + threadIdx.x;
    if (e >= e1)
        return;
    const R c = cur[e];
    fprior[e] += c - prev[e];
    prev[e] = c;
}

static sphb200_status ckCheck(sphb200_relation *r, bool want_inner, const char *who)
{
    sphb200_ctx *ctx = r->ctx;
    if (r->inner != want_inner)
        return ctx->fail(SPHB200_ERR_INVALID, std::string(who) + ": wrong relation kind");
    if (!r->valid || r->source_stamp != r->body->config_stamp)
        return ctx->fail(SPHB200_ERR_INVALID, std::string(who) + ": relation is stale (call relation_update after cell_list_update)");
    for (size_t k = 0; k < r->targets.size(); ++k)
        if (r->target_stamp[k] != r->targets[k]->config_stamp)
            return ctx->fail(SPHB200_ERR_INVALID, std::string(who) + ": relation target was re-sorted; call relation_update");
    return SPHB200_OK;
}
// registerStateVariableFrom<Real>("VolumetricMeasureRef", "VolumetricMeasure"): a copy taken when first registered
template <class R>
static sphb200_status volRef(sphb200_body *b, R **out)
{
    sphb200_ctx *ctx = b->ctx;
    const bool fresh = b->find("VolumetricMeasureRef") == nullptr;
    SPHB_TRY(getVarPtr<R>(b, "VolumetricMeasureRef", 1, out, true));
    if (fresh && b->n > 0)
    {
        R *vol;
        SPHB_TRY(getVarPtr<R>(b, "VolumetricMeasure", 1, &vol, true));
        SPHB_CUDA(ctx, cudaMemcpyAsync(*out, vol, size_t(b->n) * sizeof(R), cudaMemcpyDeviceToDevice, ctx->stream));
    }
    return SPHB200_OK;
}
#define CK_GRID(b) gridFor((b)->own_end - ((b)->own_begin & ~int64_t(31)), 128), 128, 0, ctx->stream

// stage 1: CompressionSummation (inner + contact sums), 2: DensityRegularization (the update), 3: both
template <int D, class R>
static sphb200_status ckDensityRegularization(sphb200_relation *inner, sphb200_relation *contact, int free_surface, int stage = 3)
{
    sphb200_ctx *ctx = inner->ctx;
    sphb200_body *b = inner->body;
    SPHB_TRY(ckCheck(inner, true, "ck_density_regularization"));
    if (contact)
        SPHB_TRY(ckCheck(contact, false, "ck_density_regularization"));
    if (b->own_end <= b->own_begin)
        return SPHB200_OK;
    SPHB_TRY((ensurePacked<D, R>(b)));
    R *vref, *csum, *comp, *rho;
    SPHB_TRY(volRef<R>(b, &vref));
    SPHB_TRY(getVarPtr<R>(b, "CompressionSummation", 1, &csum));
    SPHB_TRY(getVarPtr<R>(b, "Compression", 1, &comp));
    SPHB_TRY(getVarPtr<R>(b, "Density", 1, &rho, true));
    SPHB_TRY(ensureList(inner, 0));
    const R h = b->P<D, R>().h;
    auto *xV = static_cast<const typename Real4<R>::type *>(b->xV.p);
    if (stage & 1)
    {
    k_ck_compression<D, R><<<CK_GRID(b)>>>(b->own_begin, b->own_end, 1, xV, xV, ckListOf(inner->lists[0]), makeTabKernel<D, R>(h, h), vref, vref, csum);
    SPHB_LAUNCH_CHECK(ctx);
    }
    for (size_t t = 0; (stage & 1) && contact && t < contact->targets.size(); ++t)
    {
        sphb200_body *tb = contact->targets[t];
        SPHB_TRY((ensurePacked<D, R>(tb)));
        R *tref;
        SPHB_TRY(volRef<R>(tb, &tref));
        SPHB_TRY(ensureList(contact, t));
        k_ck_compression<D, R><<<CK_GRID(b)>>>(b->own_begin, b->own_end, 0, xV, static_cast<const typename Real4<R>::type *>(tb->xV.p),
                                               ckListOf(contact->lists[t]), makeTabKernel<D, R>(h, tb->P<D, R>().h), vref, tref, csum);
        SPHB_LAUNCH_CHECK(ctx);
    }
    if (stage & 2)
    {
        k_ck_regularize<R><<<gridFor(b->own_end - b->own_begin, 256), 256, 0, ctx->stream>>>(b->own_begin, b->own_end, free_surface, R(b->rho0), csum, comp, rho);
        SPHB_LAUNCH_CHECK(ctx);
    }
    b->rec_b_valid = false;
    return SPHB200_OK;
}

template <int D, class R>
static sphb200_status ckRegularizeOnly(sphb200_body *b, int free_surface)
{
    sphb200_ctx *ctx = b->ctx;
    if (b->own_end <= b->own_begin)
        return SPHB200_OK;
    if (!b->find("CompressionSummation"))
        return ctx->fail(SPHB200_ERR_NOT_FOUND, "ck_density_regularize: CompressionSummation has not been computed (ck_compression_summation)");
    R *csum, *comp, *rho;
    SPHB_TRY(getVarPtr<R>(b, "CompressionSummation", 1, &csum));
    SPHB_TRY(getVarPtr<R>(b, "Compression", 1, &comp));
    SPHB_TRY(getVarPtr<R>(b, "Density", 1, &rho, true));
    k_ck_regularize<R><<<gridFor(b->own_end - b->own_begin, 256), 256, 0, ctx->stream>>>(b->own_begin, b->own_end, free_surface, R(b->rho0), csum, comp, rho);
    SPHB_LAUNCH_CHECK(ctx);
    b->rec_b_valid = false;
    return SPHB200_OK;
}

template <int D, class R>
static sphb200_status ckCorrectionMatrix(sphb200_relation *inner, sphb200_relation *contact, double alpha)
{
    sphb200_ctx *ctx = inner->ctx;
    sphb200_body *b = inner->body;
    SPHB_TRY(ckCheck(inner, true, "ck_linear_correction_matrix"));
    if (contact)
        SPHB_TRY(ckCheck(contact, false, "ck_linear_correction_matrix"));
    if (b->own_end <= b->own_begin)
        return SPHB200_OK;
    SPHB_TRY((ensurePacked<D, R>(b)));
    R *B;
    SPHB_TRY(getVarPtr<R>(b, "LinearCorrectionMatrix", D * D, &B));
    SPHB_TRY(ensureList(inner, 0));
    const R h = b->P<D, R>().h;
    auto *xV = static_cast<const typename Real4<R>::type *>(b->xV.p);
    k_ck_correction_sum<D, R><<<CK_GRID(b)>>>(b->own_begin, b->own_end, 1, xV, xV, ckListOf(inner->lists[0]), makeTabKernel<D, R>(h, h), B);
    SPHB_LAUNCH_CHECK(ctx);
    for (size_t t = 0; contact && t < contact->targets.size(); ++t)
    {
        sphb200_body *tb = contact->targets[t];
        SPHB_TRY((ensurePacked<D, R>(tb)));
        SPHB_TRY(ensureList(contact, t));
        k_ck_correction_sum<D, R><<<CK_GRID(b)>>>(b->own_begin, b->own_end, 0, xV, static_cast<const typename Real4<R>::type *>(tb->xV.p),
                                                  ckListOf(contact->lists[t]), makeTabKernel<D, R>(h, tb->P<D, R>().h), B);
        SPHB_LAUNCH_CHECK(ctx);
    }
    k_ck_correction_update<D, R><<<gridFor(b->own_end - b->own_begin, 128), 128, 0, ctx->stream>>>(b->own_begin, b->own_end, R(alpha), B);
    SPHB_LAUNCH_CHECK(ctx);
    return SPHB200_OK;
}

template <int D, class R>
static sphb200_status ckSurfaceIndication(sphb200_relation *inner, sphb200_relation *contact)
{
    sphb200_ctx *ctx = inner->ctx;
    sphb200_body *b = inner->body;
    SPHB_TRY(ckCheck(inner, true, "ck_free_surface_indication"));
    if (contact)
        SPHB_TRY(ckCheck(contact, false, "ck_free_surface_indication"));
    if (b->own_end <= b->own_begin)
        return SPHB200_OK;
    SPHB_TRY((ensurePacked<D, R>(b)));
    const bool fresh = b->find("PreviousSurfaceIndicator") == nullptr;
    R *ind, *pdiv, *prev;
    SPHB_TRY(getVarPtr<R>(b, "Indicator", 1, &ind));
    SPHB_TRY(getVarPtr<R>(b, "PositionDivergence", 1, &pdiv));
    SPHB_TRY(getVarPtr<R>(b, "PreviousSurfaceIndicator", 1, &prev, true));
    if (fresh)
    {
        k_ck_fill<R><<<gridFor(b->n, 256), 256, 0, ctx->stream>>>(b->n, R(1), prev);
        SPHB_LAUNCH_CHECK(ctx);
    }
    SPHB_TRY(ensureList(inner, 0));
    const R h = b->P<D, R>().h, threshold = R(0.75) * R(D);
    auto *xV = static_cast<const typename Real4<R>::type *>(b->xV.p);
    k_ck_posdiv<D, R><<<CK_GRID(b)>>>(b->own_begin, b->own_end, 1, xV, xV, ckListOf(inner->lists[0]), makeTabKernel<D, R>(h, h), threshold, prev, pdiv);
    SPHB_LAUNCH_CHECK(ctx);
    for (size_t t = 0; contact && t < contact->targets.size(); ++t)
    {
        sphb200_body *tb = contact->targets[t];
        SPHB_TRY((ensurePacked<D, R>(tb)));
        SPHB_TRY(ensureList(contact, t));
        k_ck_posdiv<D, R><<<CK_GRID(b)>>>(b->own_begin, b->own_end, 0, xV, static_cast<const typename Real4<R>::type *>(tb->xV.p),
                                          ckListOf(contact->lists[t]), makeTabKernel<D, R>(h, tb->P<D, R>().h), threshold, prev, pdiv);
        SPHB_LAUNCH_CHECK(ctx);
    }
    k_ck_indicator<D, R><<<CK_GRID(b)>>>(b->own_begin, b->own_end, xV, ckListOf(inner->lists[0]), threshold, h, pdiv, ind);
    SPHB_LAUNCH_CHECK(ctx);
    SPHB_CUDA(ctx, cudaMemcpyAsync(prev + b->own_begin, ind + b->own_begin, size_t(b->own_end - b->own_begin) * sizeof(R), cudaMemcpyDeviceToDevice,
                                   ctx->stream));
    return SPHB200_OK;
}

template <int D, class R>
static CkRiemann<D, R> ckRiemannOf(sphb200_body *b)
{
    // riemann_solver.h:62-70 with rho0c0_i = rho0c0_j
    const R a = R(b->rho0) * R(b->c0), c = a;
    CkRiemann<D, R> rs;
    rs.inv_rho0c0 = (a + c) / (a * a + c * c);
    rs.rho0c0 = R(2.0) * a * c * (R(1.0) / (a + c));
    rs.inv_c0 = R(0.5) * (R(b->rho0) + R(b->rho0)) * rs.inv_rho0c0;
    return rs;
}
#define CK_FLAT(b) gridFor((b)->own_end - (b)->own_begin, 256), 256, 0, ctx->stream

// registerStateVariable<Real>("Compression", Real(1)): one, not zero, when first registered
template <class R>
static sphb200_status compressionVar(sphb200_body *b, R **out)
{
    sphb200_ctx *ctx = b->ctx;
    const bool fresh = b->find("Compression") == nullptr;
    SPHB_TRY(getVarPtr<R>(b, "Compression", 1, out, true));
    if (fresh && b->n > 0)
    {
        k_ck_fill<R><<<gridFor(b->n, 256), 256, 0, ctx->stream>>>(b->n, R(1), *out);
        SPHB_LAUNCH_CHECK(ctx);
    }
    return SPHB200_OK;
}

template <int D, class R>
static sphb200_status ckAdvectionStepSetup(sphb200_body *b)
{
    sphb200_ctx *ctx = b->ctx;
    if (b->own_end <= b->own_begin)
        return SPHB200_OK;
    R *mass, *rho, *vol, *dpos;
    SPHB_TRY(getVarPtr<R>(b, "Mass", 1, &mass, true));
    SPHB_TRY(getVarPtr<R>(b, "Density", 1, &rho, true));
    SPHB_TRY(getVarPtr<R>(b, "VolumetricMeasure", 1, &vol, true));
    SPHB_TRY(getVarPtr<R>(b, "Displacement", D, &dpos, true));
    k_ck_setup<D, R><<<CK_FLAT(b)>>>(b->own_begin, b->own_end, mass, rho, vol, dpos);
    SPHB_LAUNCH_CHECK(ctx);
    b->xV_vol_dirty = true; // the neighbours' Vol_j is read from the packed xV
    return SPHB200_OK;
}
template <int D, class R>
static sphb200_status ckUpdateParticlePosition(sphb200_body *b)
{
    sphb200_ctx *ctx = b->ctx;
    if (b->own_end <= b->own_begin)
        return SPHB200_OK;
    R *pos, *dpos;
    SPHB_TRY(getVarPtr<R>(b, "Position", D, &pos, true));
    SPHB_TRY(getVarPtr<R>(b, "Displacement", D, &dpos, true));
    k_ck_axpy<R><<<gridFor((b->own_end - b->own_begin) * D, 256), 256, 0, ctx->stream>>>(b->own_begin * D, b->own_end * D, pos, dpos);
    SPHB_LAUNCH_CHECK(ctx);
    b->rec_b_valid = false;
    return SPHB200_OK;
}
template <int D, class R>
static sphb200_status ckTimeStep(sphb200_body *b, int kind, double speed_ref, double cfl, double *out)
{
    sphb200_ctx *ctx = b->ctx;
    BodyP<D, R> &P = b->P<D, R>();
    R *vel, *force, *fprior, *mass;
    SPHB_TRY(getVarPtr<R>(b, "Velocity", D, &vel, true));
    SPHB_TRY(getVarPtr<R>(b, "Force", D, &force, true));
    SPHB_TRY(getVarPtr<R>(b, "ForcePrior", D, &fprior, true));
    SPHB_TRY(getVarPtr<R>(b, "Mass", 1, &mass, true));
    SPHB_TRY(reduceReset(ctx, kind, false));
    if (b->own_end > b->own_begin)
    {
        k_ck_dt_max<D, R><<<148 * 4, 256, 0, ctx->stream>>>(b->own_begin, b->own_end, kind, R(b->c0), P.h, vel, force, fprior, mass,
                                                           static_cast<unsigned long long *>(ctx->reduce_dev.p) + kind);
        SPHB_LAUNCH_CHECK(ctx);
    }
    double m;
    SPHB_TRY(reduceRead(ctx, kind, &m));
    if (kind == 0) // fluid_time_step_ck.hpp:31-35
        *out = double(R(cfl) * P.h / (R(m) + Tiny<R>::value));
    else // fluid_time_step_ck.cpp:23-25
        *out = double(R(cfl) * P.h / (std::max(std::sqrt(R(m)), R(speed_ref)) + Tiny<R>::value));
    return SPHB200_OK;
}

template <int D, class R>
static sphb200_status ckAcousticHalf(sphb200_relation *inner, sphb200_relation *wall, double dt_, int second)
{
    sphb200_ctx *ctx = inner->ctx;
    sphb200_body *b = inner->body;
    const char *who = second ? "ck_acoustic_step_2nd_half" : "ck_acoustic_step_1st_half";
    SPHB_TRY(ckCheck(inner, true, who));
    if (wall)
        SPHB_TRY(ckCheck(wall, false, who));
    if (b->own_end <= b->own_begin)
        return SPHB200_OK;
    if (!b->find("LinearCorrectionMatrix"))
        return ctx->fail(SPHB200_ERR_NOT_FOUND, std::string(who) + ": LinearCorrectionMatrix has not been computed (ck_linear_correction_matrix)");
    SPHB_TRY((ensurePacked<D, R>(b)));
    R *rho, *mass, *p, *comp, *crate, *vel, *dpos, *force, *fprior, *B, *vol;
    SPHB_TRY(getVarPtr<R>(b, "Density", 1, &rho, true));
    SPHB_TRY(getVarPtr<R>(b, "Mass", 1, &mass, true));
    SPHB_TRY(getVarPtr<R>(b, "Pressure", 1, &p, true));
    SPHB_TRY(compressionVar<R>(b, &comp));
    SPHB_TRY(getVarPtr<R>(b, "CompressionRate", 1, &crate, true));
    SPHB_TRY(getVarPtr<R>(b, "Velocity", D, &vel, true));
    SPHB_TRY(getVarPtr<R>(b, "Displacement", D, &dpos, true));
    SPHB_TRY(getVarPtr<R>(b, "Force", D, &force, true));
    SPHB_TRY(getVarPtr<R>(b, "ForcePrior", D, &fprior, true));
    SPHB_TRY(getVarPtr<R>(b, "LinearCorrectionMatrix", D * D, &B));
    SPHB_TRY(getVarPtr<R>(b, "VolumetricMeasure", 1, &vol, true));
    SPHB_TRY(ensureList(inner, 0));
    const R h = b->P<D, R>().h, dt = R(dt_);
    const CkRiemann<D, R> rs = ckRiemannOf<D, R>(b);
    auto *xV = static_cast<const typename Real4<R>::type *>(b->xV.p);
    const int64_t i0 = b->own_begin, i1 = b->own_end;
    if (!second)
    {
        k_ck_first_init<D, R><<<CK_FLAT(b)>>>(i0, i1, dt, R(b->rho0), R(b->c0), comp, crate, rho, p, vel, dpos);
        SPHB_LAUNCH_CHECK(ctx);
        k_ck_first_inner<D, R><<<CK_GRID(b)>>>(i0, i1, xV, ckListOf(inner->lists[0]), makeTabKernel<D, R>(h, h), rs, p, B, vol, comp, crate, force);
        SPHB_LAUNCH_CHECK(ctx);
    }
    else
    {
        k_ck_second_inner<D, R><<<CK_GRID(b)>>>(i0, i1, dt, xV, ckListOf(inner->lists[0]), makeTabKernel<D, R>(h, h), rs, vel, B, vol, comp, crate,
                                                force, dpos);
        SPHB_LAUNCH_CHECK(ctx);
    }
    for (size_t t = 0; wall && t < wall->targets.size(); ++t)
    {
        sphb200_body *tb = wall->targets[t];
        SPHB_TRY((ensurePacked<D, R>(tb)));
        SPHB_TRY(ensureList(wall, t));
        auto *xVt = static_cast<const typename Real4<R>::type *>(tb->xV.p);
        const TabKernelP<R> kc = makeTabKernel<D, R>(h, tb->P<D, R>().h);
        R *w_acc = nullptr, *w_vel = nullptr, *w_n = nullptr;
        if (tb->find("Acceleration"))
            SPHB_TRY(getVarPtr<R>(tb, "Acceleration", D, &w_acc));
        if (tb->find("Velocity"))
            SPHB_TRY(getVarPtr<R>(tb, "Velocity", D, &w_vel));
        if (!second)
            k_ck_first_wall<D, R><<<CK_GRID(b)>>>(i0, i1, xV, xVt, ckListOf(wall->lists[t]), kc, rs, p, rho, mass, fprior, w_acc, B, vol, comp, crate, force);
        else
        {
            if (!tb->find("NormalDirection"))
                return ctx->fail(SPHB200_ERR_NOT_FOUND, std::string(who) + ": wall body has no NormalDirection");
            SPHB_TRY(getVarPtr<R>(tb, "NormalDirection", D, &w_n));
            k_ck_second_wall<D, R><<<CK_GRID(b)>>>(i0, i1, xV, xVt, ckListOf(wall->lists[t]), kc, rs, vel, w_vel, w_n, B, vol, comp, crate, force);
        }
        SPHB_LAUNCH_CHECK(ctx);
    }
    if (!second)
        k_ck_first_update<D, R><<<CK_FLAT(b)>>>(i0, i1, dt, mass, fprior, force, vel);
    else
        k_ck_second_update<R><<<CK_FLAT(b)>>>(i0, i1, dt, R(b->rho0), comp, crate, rho);
    SPHB_LAUNCH_CHECK(ctx);
    b->rec_b_valid = false;
    b->acoustic_max_valid = false;
    return SPHB200_OK;
}

// kind 0: PressureForceFromFluid, 1: ViscousForceFromFluid
template <int D, class R>
static sphb200_status ckForceFromFluid(sphb200_relation *contact, int kind, int corrected)
{
    sphb200_ctx *ctx = contact->ctx;
    sphb200_body *b = contact->body;
    const char *who = kind ? "ck_viscous_force_from_fluid" : "ck_pressure_force_from_fluid";
    const char *name = kind ? "ViscousForceFromFluid" : "PressureForceFromFluid";
    const char *prev_name = kind ? "PreviousViscousForceFromFluid" : "PreviousPressureForceFromFluid";
    SPHB_TRY(ckCheck(contact, false, who));
    if (b->own_end <= b->own_begin)
        return SPHB200_OK;
    SPHB_TRY((ensurePacked<D, R>(b)));
    R *force, *prev, *fprior, *vol, *vel_ave = nullptr, *acc_ave = nullptr, *nrm = nullptr;
    SPHB_TRY(getVarPtr<R>(b, name, D, &force));
    SPHB_TRY(getVarPtr<R>(b, prev_name, D, &prev, true));
    SPHB_TRY(getVarPtr<R>(b, "ForcePrior", D, &fprior, true));
    SPHB_TRY(getVarPtr<R>(b, "VolumetricMeasure", 1, &vol, true));
    if (b->find("Velocity"))
        SPHB_TRY(getVarPtr<R>(b, "Velocity", D, &vel_ave));
    if (b->find("Acceleration"))
        SPHB_TRY(getVarPtr<R>(b, "Acceleration", D, &acc_ave));
    if (kind == 0)
    {
        if (!b->find("NormalDirection"))
            return ctx->fail(SPHB200_ERR_NOT_FOUND, std::string(who) + ": the structure has no NormalDirection");
        SPHB_TRY(getVarPtr<R>(b, "NormalDirection", D, &nrm));
    }
    const R h = b->P<D, R>().h;
    auto *xV = static_cast<const typename Real4<R>::type *>(b->xV.p);
    if (contact->targets.empty())
        SPHB_CUDA(ctx, cudaMemsetAsync(force + b->own_begin * D, 0, size_t(b->own_end - b->own_begin) * D * sizeof(R), ctx->stream));
    for (size_t t = 0; t < contact->targets.size(); ++t)
    {
        sphb200_body *f = contact->targets[t];
        SPHB_TRY((ensurePacked<D, R>(f)));
        SPHB_TRY(ensureList(contact, t));
        R *t_vel, *B_t = nullptr;
        SPHB_TRY(getVarPtr<R>(f, "Velocity", D, &t_vel, true));
        if (corrected)
        {
            if (!f->find("LinearCorrectionMatrix"))
                return ctx->fail(SPHB200_ERR_NOT_FOUND, std::string(who) + ": the fluid has no LinearCorrectionMatrix (ck_linear_correction_matrix)");
            SPHB_TRY(getVarPtr<R>(f, "LinearCorrectionMatrix", D * D, &B_t));
        }
        auto *xVt = static_cast<const typename Real4<R>::type *>(f->xV.p);
        const R hf = f->P<D, R>().h;
        const TabKernelP<R> kc = makeTabKernel<D, R>(h, hf);
        if (kind == 0)
        {
            R *t_rho, *t_mass, *t_p, *t_fp;
            SPHB_TRY(getVarPtr<R>(f, "Density", 1, &t_rho, true));
            SPHB_TRY(getVarPtr<R>(f, "Mass", 1, &t_mass, true));
            SPHB_TRY(getVarPtr<R>(f, "Pressure", 1, &t_p, true));
            SPHB_TRY(getVarPtr<R>(f, "ForcePrior", D, &t_fp, true));
            const CkRiemann<D, R> rs = ckRiemannOf<D, R>(f);
            k_ck_fsi_pressure<D, R><<<CK_GRID(b)>>>(b->own_begin, b->own_end, t == 0, xV, xVt, ckListOf(contact->lists[t]), kc, rs.rho0c0, rs.inv_c0,
                                                    vel_ave, acc_ave, nrm, vol, t_vel, t_rho, t_mass, t_p, t_fp, B_t, force);
        }
        else
            k_ck_fsi_viscous<D, R><<<CK_GRID(b)>>>(b->own_begin, b->own_end, t == 0, xV, xVt, ckListOf(contact->lists[t]), kc, R(f->mu), hf * hf, vel_ave,
                                                   vol, t_vel, B_t, force);
        SPHB_LAUNCH_CHECK(ctx);
    }
    k_ck_force_prior_update<R><<<gridFor((b->own_end - b->own_begin) * D, 256), 256, 0, ctx->stream>>>(b->own_begin * D, b->own_end * D, fprior, force, prev);
    SPHB_LAUNCH_CHECK(ctx);
    return SPHB200_OK;
}
} // namespace sphb

extern "C"
{
sphb200_status sphb200_ck_pressure_force_from_fluid(sphb200_relation *contact, int corrected)
{
    if (!contact)
        return SPHB200_ERR_INVALID;
    cudaSetDevice(contact->ctx->device);
    return SPHB_DISPATCH(contact->ctx, ckForceFromFluid, contact, 0, corrected);
}
sphb200_status sphb200_ck_viscous_force_from_fluid(sphb200_relation *contact, int corrected)
{
    if (!contact)
        return SPHB200_ERR_INVALID;
    cudaSetDevice(contact->ctx->device);
    return SPHB_DISPATCH(contact->ctx, ckForceFromFluid, contact, 1, corrected);
}
sphb200_status sphb200_ck_advection_step_setup(sphb200_body *b)
{
    if (!b)
        return SPHB200_ERR_INVALID;
    cudaSetDevice(b->ctx->device);
    return SPHB_DISPATCH(b->ctx, ckAdvectionStepSetup, b);
}
sphb200_status sphb200_ck_update_particle_position(sphb200_body *b)
{
    if (!b)
        return SPHB200_ERR_INVALID;
    cudaSetDevice(b->ctx->device);
    return SPHB_DISPATCH(b->ctx, ckUpdateParticlePosition, b);
}
sphb200_status sphb200_ck_acoustic_dt(sphb200_body *b, double cfl, double *dt_out)
{
    if (!b || !dt_out)
        return SPHB200_ERR_INVALID;
    cudaSetDevice(b->ctx->device);
    return SPHB_DISPATCH(b->ctx, ckTimeStep, b, 0, 0.0, cfl, dt_out);
}
sphb200_status sphb200_ck_advection_dt(sphb200_body *b, double U_ref, double cfl, double *dt_out)
{
    if (!b || !dt_out)
        return SPHB200_ERR_INVALID;
    cudaSetDevice(b->ctx->device);
    return SPHB_DISPATCH(b->ctx, ckTimeStep, b, 1, U_ref, cfl, dt_out);
}
sphb200_status sphb200_ck_acoustic_step_1st_half(sphb200_relation *inner, sphb200_relation *wall, double dt)
{
    if (!inner || (wall && wall->ctx != inner->ctx))
        return SPHB200_ERR_INVALID;
    cudaSetDevice(inner->ctx->device);
    return SPHB_DISPATCH(inner->ctx, ckAcousticHalf, inner, wall, dt, 0);
}
sphb200_status sphb200_ck_acoustic_step_2nd_half(sphb200_relation *inner, sphb200_relation *wall, double dt)
{
    if (!inner || (wall && wall->ctx != inner->ctx))
        return SPHB200_ERR_INVALID;
    cudaSetDevice(inner->ctx->device);
    return SPHB_DISPATCH(inner->ctx, ckAcousticHalf, inner, wall, dt, 1);
}
sphb200_status sphb200_ck_density_regularization(sphb200_relation *inner, sphb200_relation *contact, int free_surface)
{
    if (!inner || (contact && contact->ctx != inner->ctx))
        return SPHB200_ERR_INVALID;
    cudaSetDevice(inner->ctx->device);
    return SPHB_DISPATCH(inner->ctx, ckDensityRegularization, inner, contact, free_surface, 3);
}
sphb200_status sphb200_ck_compression_summation(sphb200_relation *inner, sphb200_relation *contact)
{
    if (!inner || (contact && contact->ctx != inner->ctx))
        return SPHB200_ERR_INVALID;
    cudaSetDevice(inner->ctx->device);
    return SPHB_DISPATCH(inner->ctx, ckDensityRegularization, inner, contact, 0, 1);
}
sphb200_status sphb200_ck_density_regularize(sphb200_body *b, int free_surface)
{
    if (!b)
        return SPHB200_ERR_INVALID;
    cudaSetDevice(b->ctx->device);
    return SPHB_DISPATCH(b->ctx, ckRegularizeOnly, b, free_surface);
}
sphb200_status sphb200_ck_linear_correction_matrix(sphb200_relation *inner, sphb200_relation *contact, double alpha)
{
    if (!inner || (contact && contact->ctx != inner->ctx))
        return SPHB200_ERR_INVALID;
    cudaSetDevice(inner->ctx->device);
    return SPHB_DISPATCH(inner->ctx, ckCorrectionMatrix, inner, contact, alpha);
}
sphb200_status sphb200_ck_free_surface_indication(sphb200_relation *inner, sphb200_relation *contact)
{
    if (!inner || (contact && contact->ctx != inner->ctx))
        return SPHB200_ERR_INVALID;
    cudaSetDevice(inner->ctx->device);
    return SPHB_DISPATCH(inner->ctx, ckSurfaceIndication, inner, contact);
}
} // extern "C"
