// Weakly-compressible fluid dynamics of the classic SPHinXsys path as fused sm_100a kernels.
//
// Reference local dynamics reproduced, one kernel per Dynamics algorithm instead
// of one particle_for per initialization/interaction/update stage:
//   fluid_dynamics/density_summation.cpp:8-22,58-78, density_summation.hpp:11-32
//   fluid_dynamics/fluid_integration.hpp:50-113 (1st half), :167-231 (2nd half)
//   fluid_dynamics/base_fluid_dynamics.h:54-75 (InteractionWithWall data)
//   materials/riemann_solver.h:55-123, common/common_functors.h:82-94 (acoustic Riemann solver, limiter 3)
//   materials/weakly_compressible_fluid.cpp:8-30 (EOS)
//   fluid_dynamics/fluid_time_step.cpp:12-79 (time-step reductions, ReduceMax)
//   fluid_dynamics/viscous_dynamics.hpp:35-53, general_dynamics/force_prior.hpp:10-41
//   fluid_dynamics/transport_velocity_correction.hpp:38-67
//   general_dynamics/general_reduce.cpp:18-79, general_interpolation.h:63-83
//
// Data flow per particle i (thread i): the SELL-32 index list is read coalesced (lane = i mod 32),
// each neighbour costs one aligned Real4 gather of the frozen (x, Vol) record plus the gathered field
// (p_j or v_j); W/dW/e/r are recomputed in registers.  Gather-only writes (index i only), exactly the
// reference's write discipline, so no atomics on particle data.
#include "internal.cuh"

using namespace sphb;

// occupancy / ILP knobs of the two pair-sum kernels of the acoustic sub-step (resident 128-thread
// blocks per SM the register allocation must allow, and the neighbour-loop unroll depth)
#ifndef SPHB_MINB_F64
#define SPHB_MINB_F64 8
#endif
#ifndef SPHB_MINB_F32
#define SPHB_MINB_F32 8
#endif
// measured on B200 (8 M particles, ms 1st/2nd half): fp64 minb 4: 7.7/6.1, 6: 5.5/5.3, 8: 4.7/4.7, 10: 5.2/6.8;
// fp32 minb 8 unroll 4: 1.98/1.86, unroll 8: 1.90/1.78, minb 12: 2.07/1.96
#ifndef SPHB_UNROLL_F64
#define SPHB_UNROLL_F64 4
#endif
#ifndef SPHB_UNROLL_F32
#define SPHB_UNROLL_F32 8
#endif
constexpr int kUnrollF64 = SPHB_UNROLL_F64, kUnrollF32 = SPHB_UNROLL_F32;

namespace sphb
{
template <class R>
struct ListP
{
    const u32 *count, *slice_off, *list;
};

template <class R>
struct FluidP
{
    R rho0, c0, p0, mu, h;
    // riemann_solver.h:55-92 constants for identical left/right fluids
    R inv_rho0c0_ave, rho0c0_geo_ave, inv_c0_ave, limiter_slope;
    int riemann;
    const StepCtl *ctl; // device-resident dt / exit flag of the sub-step loop, or null: dt is the kernel argument
};

template <class R>
static int quadLanesK2();
static sphb200_status ensureStepCtl(sphb200_ctx *ctx);

// dt of this launch: the kernel argument, or the device-resident one; false: the loop has ended, return at once
template <class R>
__device__ __forceinline__ bool stepDt(const StepCtl *ctl, R &dt)
{
    if (!ctl)
        return true;
    if (*reinterpret_cast<const volatile int *>(&ctl->done))
        return false;
    dt = R(*reinterpret_cast<const volatile double *>(&ctl->dt));
    return true;
}

template <class R>
static FluidP<R> makeFluidP(const sphb200_body *b, R h, bool riemann)
{
    FluidP<R> f;
    f.rho0 = R(b->rho0);
    f.c0 = R(b->c0);
    f.p0 = f.rho0 * f.c0 * f.c0;
    f.mu = R(b->mu);
    f.h = h;
    R a = f.rho0 * f.c0, c = f.rho0 * f.c0;
    R inv_sum = R(1.0) / (a + c);
    f.inv_rho0c0_ave = (a + c) / (std::pow(a, R(2)) + std::pow(c, R(2)));
    f.rho0c0_geo_ave = R(2.0) * a * c * inv_sum;
    f.inv_c0_ave = R(0.5) * (f.rho0 + f.rho0) * f.inv_rho0c0_ave;
    f.limiter_slope = R(3.0);
    f.riemann = riemann ? 1 : 0;
    f.ctl = b->ctx->ctl_active;
    return f;
}

template <class R>
__device__ __forceinline__ R dissipativePJump(const FluidP<R> &f, R u)
{
    // riemann_solver.h:109-114 with TruncatedLinear (common_functors.h:82-94)
    R lim = fmin(f.limiter_slope * (f.inv_c0_ave * fmax(u, R(0))), R(1));
    return f.riemann ? f.rho0c0_geo_ave * u * lim : R(0);
}
template <class R>
__device__ __forceinline__ R dissipativeUJump(const FluidP<R> &f, R dp)
{
    return f.riemann ? dp * f.inv_rho0c0_ave : R(0);
}

template <int D, class R>
__device__ __forceinline__ void loadVec(const R *__restrict__ a, u32 j, R *v)
{
#pragma unroll
    for (int d = 0; d < D; ++d)
        v[d] = __ldg(a + size_t(j) * D + d);
}

template <int D, class R>
__device__ __forceinline__ void loadXj(const typename Real4<R>::type *__restrict__ xV, u32 j, R *x, R &w)
{
    typename Real4<R>::type v = ldg4(xV + j);
    x[0] = v.x;
    x[1] = v.y;
    if (D == 3)
        x[D - 1] = v.z;
    w = v.w;
}

// ---------------------------------------------------------------------------------------------
// DensitySummation
// ---------------------------------------------------------------------------------------------
template <int D, class R>
__global__ void __launch_bounds__(128)
    k_density_summation(int64_t i0, int64_t n, const typename Real4<R>::type *xV, ListP<R> in, ContactP<R> ct,
                        KernelP<R> kern, KernelP<R> kern_ct, PeriodicP<D, R> per, R rho0, R inv_sigma0,
                        const R *__restrict__ mass, R *__restrict__ rho_sum, R *__restrict__ rho, R *__restrict__ vol,
                        typename Real4<R>::type *xV_w, int free_surface)
{
    int64_t i = (i0 & ~int64_t(31)) + blockIdx.x * int64_t(blockDim.x) + threadIdx.x;
    if (i < i0 || i >= n)
        return;
    R xi[D], Vi;
    loadXj<D, R>(xV, u32(i), xi, Vi);
    R sigma = kern.factor_W; // W0
    {
        const u32 cnt = in.count[i];
        const u32 *lst = listBase(in.list, in.slice_off, i);
#pragma unroll 4
        for (u32 k = 0; k < cnt; ++k)
        {
            u32 j = listEntry(lst, k);
            R xj[D], Vj;
            loadXj<D, R>(xV, j, xj, Vj);
            periodicImage<D, R>(xi, xj, per);
            R r2 = R(0);
#pragma unroll
            for (int a = 0; a < D; ++a)
            {
                R d = xi[a] - xj[a];
                r2 += d * d;
            }
            sigma += kernelW<R>(sqrt_r(r2), kern);
        }
    }
    R rs = sigma * rho0 * inv_sigma0;
    if (ct.n > 0)
    {
        R s = R(0);
        for (int t = 0; t < ct.n; ++t)
        {
            const TargetP<R> &T = ct.t[t];
            const u32 cnt = T.count[i];
            const u32 *lst = listBase(T.list, T.slice_off, i);
            for (u32 k = 0; k < cnt; ++k)
            {
                u32 j = listEntry(lst, k);
                R xj[D], Vj;
                loadXj<D, R>(T.xV, j, xj, Vj);
                R r2 = R(0);
#pragma unroll
                for (int a = 0; a < D; ++a)
                {
                    R d = xi[a] - xj[a];
                    r2 += d * d;
                }
                s += kernelW<R>(sqrt_r(r2), kern_ct) * T.inv_rho0 * __ldg(T.mass + j);
            }
        }
        rs += s * rho0 * rho0 * inv_sigma0 / mass[i];
    }
    rho_sum[i] = rs;
    if (free_surface)
    {
        rho[i] = fmax(rs, rho0); // density_summation.hpp:28-32
    }
    else
    {
        rho[i] = rs; // density_summation.cpp:18-22
        R v = mass[i] / rs;
        vol[i] = v;
        // keep the packed Vol in step; no thread reads another particle's .w in this kernel
        reinterpret_cast<R *>(xV_w + i)[3] = v;
    }
}

// ---------------------------------------------------------------------------------------------
// Acoustic sub-step.  Neighbour state is gathered from packed records (internal.cuh Rec8):
//   K1 k_first_half_init : rho, p, x update; writes rec_a[i] = {x_cfg, p_new | Vol, v}
//   K2 k_first_half      : gathers rec_a[j] (x_cfg, p, Vol); writes rec_b[i] = {x_cfg, p | Vol, v_new}
//   K3 k_second_half     : gathers rec_b[j] (x_cfg, Vol, v)
// rec_a / rec_b ping-pong so that no kernel reads a line another thread of the same launch writes.
// ---------------------------------------------------------------------------------------------
template <int D, class R>
__device__ __forceinline__ void recSplit(const R *rec, R *x, R &V, R &p, R *v)
{
    x[0] = rec[0];
    x[1] = rec[1];
    if (D == 3)
        x[D - 1] = rec[2];
    p = rec[3];
    V = rec[4];
    v[0] = rec[5];
    v[1] = rec[6];
    if (D == 3)
        v[D - 1] = rec[7];
}
template <int D, class R>
__device__ __forceinline__ void recJoin(R *rec, const R *x, R V, R p, const R *v)
{
    rec[0] = x[0];
    rec[1] = x[1];
    rec[2] = D == 3 ? x[D - 1] : R(0);
    rec[3] = p;
    rec[4] = V;
    rec[5] = v[0];
    rec[6] = v[1];
    rec[7] = D == 3 ? v[D - 1] : R(0);
}

// rec[i] = {xV[i] | p[i], vel[i]} from the named variables (after a re-sort or a host-side write)
template <int D, class R>
__global__ void __launch_bounds__(256)
    k_pack_rec(int64_t n, const typename Real4<R>::type *__restrict__ xV, const R *__restrict__ p, const R *__restrict__ vel,
               Rec8<R> *__restrict__ rec)
{
    int64_t i = blockIdx.x * int64_t(blockDim.x) + threadIdx.x;
    if (i >= n)
        return;
    R x[D], V, v[D], r8[8];
    loadXj<D, R>(xV, u32(i), x, V);
#pragma unroll
    for (int a = 0; a < D; ++a)
        v[a] = vel[i * D + a];
    recJoin<D, R>(r8, x, V, p[i], v);
    stRec(rec, i, r8);
}

template <int D, class R>
__global__ void __launch_bounds__(256)
    k_first_half_init(int64_t i0, int64_t n, R dt, R p0, R rho0, const typename Real4<R>::type *__restrict__ xV,
                      R *__restrict__ rho, const R *__restrict__ drho, R *__restrict__ p, R *__restrict__ pos,
                      const R *__restrict__ vel, Rec8<R> *__restrict__ rec_a, const StepCtl *__restrict__ ctl,
                      unsigned long long *__restrict__ max_slot)
{
    if (!stepDt(ctl, dt))
        return;
    if (ctl && blockIdx.x == 0 && threadIdx.x == 0)
        *max_slot = 0ull; // the interaction kernel of this sub-step maximises c0 + |v| into it (stream order: K1 before K2)
    int64_t i = (i0 & ~int64_t(31)) + blockIdx.x * int64_t(blockDim.x) + threadIdx.x;
    if (i < i0 || i >= n)
        return;
    R r = rho[i] + drho[i] * dt * R(0.5);
    rho[i] = r;
    R pn = p0 * (r / rho0 - R(1.0));
    p[i] = pn;
    R x[D], V, v[D], r8[8];
    loadXj<D, R>(xV, u32(i), x, V);
#pragma unroll
    for (int d = 0; d < D; ++d)
    {
        v[d] = vel[i * D + d];
        pos[i * D + d] += v[d] * dt * R(0.5);
    }
    recJoin<D, R>(r8, x, V, pn, v);
    stRec(rec_a, i, r8);
}

// Lane mapping of the kS-lanes-per-particle kernels: warp w of the grid owns the slice that starts at
// particle (i0/kP)*kP + w*kP, lane l works for particle l/kS of the slice on list entries l%kS, +kS, ...
struct GroupIdx
{
    int64_t i;  // particle
    int64_t ic; // i clamped into the valid range (safe to load from)
    int sub;    // which of the kS lanes of the particle
    bool valid; // i is an owned particle
};
__device__ __forceinline__ GroupIdx groupIndex(int64_t i0, int64_t n)
{
    const int lane = threadIdx.x & 31;
    const int64_t warp = (blockIdx.x * int64_t(blockDim.x) + threadIdx.x) >> 5;
    GroupIdx g;
    g.i = (i0 / kP) * kP + warp * kP + lane / kS;
    g.sub = lane % kS;
    g.valid = g.i >= i0 && g.i < n;
    g.ic = g.valid ? g.i : i0;
    return g;
}
inline int groupGrid(int64_t i0, int64_t n, int block)
{
    int64_t slices = (n - (i0 / kP) * kP + kP - 1) / kP;
    return int((slices * 32 + block - 1) / block);
}

template <int D, class R, bool PER>
__global__ void __launch_bounds__(128, sizeof(R) == 8 ? SPHB_MINB_F64 : SPHB_MINB_F32)
    k_first_half(int64_t i0, int64_t n, const Rec8<R> *__restrict__ rec_a, Rec8<R> *__restrict__ rec_b, ListP<R> in,
                 ContactP<R> wall, KernelP<R> kern, KernelP<R> kern_w, PeriodicP<D, R> per, FluidP<R> fl, R dt,
                 const R *__restrict__ rho, const R *__restrict__ mass, const R *__restrict__ fprior, R *__restrict__ force,
                 R *__restrict__ drho, R *__restrict__ vel, unsigned long long *__restrict__ max_slot)
{
    if (!stepDt(fl.ctl, dt))
        return;
    const GroupIdx gi = groupIndex(i0, n);
    const int lane = threadIdx.x & 31;
    const int64_t i = gi.ic;
    R ri[8], xi[D], Vi, pi, vi[D];
    ldRec(rec_a, u32(i), ri);
    recSplit<D, R>(ri, xi, Vi, pi, vi);
    R f[D];
#pragma unroll
    for (int a = 0; a < D; ++a)
        f[a] = R(0);
    R diss = R(0);
    {
        const u32 cnt = gi.valid ? in.count[i] : 0u;
        const u32 *lst = in.list + u64(in.slice_off[i / kP]) * 32u + lane;
#pragma unroll(sizeof(R) == 8 ? kUnrollF64 : kUnrollF32)
        for (u32 e = gi.sub; e < cnt; e += kS, lst += 32)
        {
            u32 j = __ldg(lst);
            R rj[8], xj[D];
            ldRecXVp(rec_a, j, rj);
            xj[0] = rj[0];
            xj[1] = rj[1];
            if (D == 3)
                xj[D - 1] = rj[2];
            const R pj = rj[3], Vj = rj[4];
            if (PER)
                periodicImage<D, R>(xi, xj, per);
            Pair<D, R> g;
            pairGeometry<D, R>(xi, xj, kern, g);
            R dWV = g.dW * Vj;
            R ps = (pi + pj) * dWV;
#pragma unroll
            for (int a = 0; a < D; ++a)
                f[a] -= ps * g.e[a];
            diss += dissipativeUJump<R>(fl, pi - pj) * dWV;
        }
    }
#pragma unroll
    for (int a = 0; a < D; ++a)
        f[a] = groupSum<R>(f[a]);
    diss = groupSum<R>(diss);
    const R rhoi = rho[i], mi = mass[i];
    R fp[D];
#pragma unroll
    for (int a = 0; a < D; ++a)
        fp[a] = fprior[i * D + a];
    R fw[D];
#pragma unroll
    for (int a = 0; a < D; ++a)
        fw[a] = R(0);
    R dw = R(0);
    if (wall.n > 0)
    {
        R acc_i[D];
#pragma unroll
        for (int a = 0; a < D; ++a)
            acc_i[a] = fp[a] / mi;
        for (int t = 0; t < wall.n; ++t)
        {
            const TargetP<R> &T = wall.t[t];
            const u32 cnt = gi.valid ? T.count[i] : 0u;
            const u32 *lst = T.list + u64(T.slice_off[i / kP]) * 32u + lane;
            for (u32 e = gi.sub; e < cnt; e += kS, lst += 32)
            {
                u32 j = __ldg(lst);
                R xj[D], Vj;
                loadXj<D, R>(T.xV, j, xj, Vj);
                R aw[D];
                loadVec<D, R>(T.acc, j, aw);
                Pair<D, R> g;
                pairGeometry<D, R>(xi, xj, kern_w, g);
                R dWV = g.dW * Vj;
                R face = R(0);
#pragma unroll
                for (int a = 0; a < D; ++a)
                    face += (acc_i[a] - aw[a]) * (-g.e[a]);
                R p_wall = pi + rhoi * g.r * fmax(R(0), face);
                R ps = (pi + p_wall) * dWV;
#pragma unroll
                for (int a = 0; a < D; ++a)
                    fw[a] -= ps * g.e[a];
                dw += dissipativeUJump<R>(fl, pi - p_wall) * dWV;
            }
        }
#pragma unroll
        for (int a = 0; a < D; ++a)
            fw[a] = groupSum<R>(fw[a]);
        dw = groupSum<R>(dw);
    }
    R signal = R(0);
    if (gi.valid && gi.sub == 0)
    {
        R fo[D];
#pragma unroll
        for (int a = 0; a < D; ++a)
            fo[a] = force[i * D + a] + f[a] * Vi;
        R dr = diss * rhoi;
        if (wall.n > 0)
        {
#pragma unroll
            for (int a = 0; a < D; ++a)
                fo[a] += fw[a] * Vi;
            dr += dw * rhoi;
        }
        drho[i] = dr;
        R v[D];
#pragma unroll
        for (int a = 0; a < D; ++a)
        {
            force[i * D + a] = fo[a];
            v[a] = vi[a] + (fp[a] + fo[a]) / mi * dt; // fluid_integration.hpp:58-61
            vel[i * D + a] = v[a];
        }
        recJoin<D, R>(ri, xi, Vi, pi, v);
        stRec(rec_b, i, ri);
        signal = fl.c0 + sqrt_r(sqnormExact<D, R>(v)); // fluid_time_step.cpp:21-24
    }
    signal = warpMax<R>(signal);
    if (lane == 0)
        atomicMaxNonNeg(max_slot, double(signal));
}

// ---------------------------------------------------------------------------------------------
// Integration2ndHalf: initialization + interaction (+wall) + update in one kernel
// ---------------------------------------------------------------------------------------------
template <int D, class R, bool PER>
__global__ void __launch_bounds__(128, sizeof(R) == 8 ? SPHB_MINB_F64 : SPHB_MINB_F32)
    k_second_half(int64_t i0, int64_t n, const Rec8<R> *__restrict__ rec_b, ListP<R> in, ContactP<R> wall, KernelP<R> kern,
                  KernelP<R> kern_w, PeriodicP<D, R> per, FluidP<R> fl, R dt, R *__restrict__ pos, R *__restrict__ rho,
                  R *__restrict__ drho, R *__restrict__ force)
{
    if (!stepDt(fl.ctl, dt))
        return;
    const GroupIdx gi = groupIndex(i0, n);
    const int lane = threadIdx.x & 31;
    const int64_t i = gi.ic;
    R ri[8], xi[D], Vi, pi_unused, vi[D];
    ldRec(rec_b, u32(i), ri);
    recSplit<D, R>(ri, xi, Vi, pi_unused, vi);
    R rate = R(0);
    R pd[D];
#pragma unroll
    for (int a = 0; a < D; ++a)
        pd[a] = R(0);
    {
        const u32 cnt = gi.valid ? in.count[i] : 0u;
        const u32 *lst = in.list + u64(in.slice_off[i / kP]) * 32u + lane;
#pragma unroll(sizeof(R) == 8 ? kUnrollF64 : kUnrollF32)
        for (u32 e = gi.sub; e < cnt; e += kS, lst += 32)
        {
            u32 j = __ldg(lst);
            R rj[8], xj[D], vj[D], Vj, pj_unused;
            ldRec(rec_b, j, rj);
            recSplit<D, R>(rj, xj, Vj, pj_unused, vj);
            if (PER)
                periodicImage<D, R>(xi, xj, per);
            Pair<D, R> g;
            pairGeometry<D, R>(xi, xj, kern, g);
            R dWV = g.dW * Vj;
            R u = R(0);
#pragma unroll
            for (int a = 0; a < D; ++a)
                u += (vi[a] - vj[a]) * g.e[a];
            rate += u * dWV;
            R pj = dissipativePJump<R>(fl, u) * dWV;
#pragma unroll
            for (int a = 0; a < D; ++a)
                pd[a] += pj * g.e[a];
        }
    }
    rate = groupSum<R>(rate);
#pragma unroll
    for (int a = 0; a < D; ++a)
        pd[a] = groupSum<R>(pd[a]);
    R rw = R(0);
    R pw[D];
#pragma unroll
    for (int a = 0; a < D; ++a)
        pw[a] = R(0);
    if (wall.n > 0)
    {
        for (int t = 0; t < wall.n; ++t)
        {
            const TargetP<R> &T = wall.t[t];
            const u32 cnt = gi.valid ? T.count[i] : 0u;
            const u32 *lst = T.list + u64(T.slice_off[i / kP]) * 32u + lane;
            for (u32 e = gi.sub; e < cnt; e += kS, lst += 32)
            {
                u32 j = __ldg(lst);
                R xj[D], Vj, vw[D], nw[D];
                loadXj<D, R>(T.xV, j, xj, Vj);
                loadVec<D, R>(T.vel, j, vw);
                loadVec<D, R>(T.nrm, j, nw);
                Pair<D, R> g;
                pairGeometry<D, R>(xi, xj, kern_w, g);
                R dWV = g.dW * Vj;
                R en = R(0);
#pragma unroll
                for (int a = 0; a < D; ++a)
                    en += g.e[a] * nw[a];
                R sgn = en < R(0) ? R(-1) : (en > R(0) ? R(1) : R(0)); // scalar_functions.h:127-131
                R u1 = R(0), u2 = R(0);
#pragma unroll
                for (int a = 0; a < D; ++a)
                {
                    R v_in_wall = R(2.0) * vw[a] - vi[a];
                    u1 += (vi[a] - v_in_wall) * g.e[a];
                    u2 += (vi[a] - vw[a]) * (sgn * nw[a]);
                }
                rw += u1 * dWV;
                R pj = dissipativePJump<R>(fl, R(2.0) * u2) * dWV;
#pragma unroll
                for (int a = 0; a < D; ++a)
                    pw[a] += pj * (sgn * nw[a]);
            }
        }
        rw = groupSum<R>(rw);
#pragma unroll
        for (int a = 0; a < D; ++a)
            pw[a] = groupSum<R>(pw[a]);
    }
    if (gi.valid && gi.sub == 0)
    {
#pragma unroll
        for (int a = 0; a < D; ++a)
            pos[i * D + a] += vi[a] * dt * R(0.5); // fluid_integration.hpp:167-170
        const R rhoi = rho[i];
        R dr = drho[i] + rate * rhoi;
        R fo[D];
#pragma unroll
        for (int a = 0; a < D; ++a)
            fo[a] = pd[a] * Vi;
        if (wall.n > 0)
        {
            dr += rw * rhoi;
#pragma unroll
            for (int a = 0; a < D; ++a)
                fo[a] += pw[a] * Vi;
        }
        drho[i] = dr;
#pragma unroll
        for (int a = 0; a < D; ++a)
            force[i * D + a] = fo[a];
        rho[i] = rhoi + dr * dt * R(0.5); // fluid_integration.hpp:173-176
    }
}

// ---------------------------------------------------------------------------------------------
// time-step reductions and observables
// ---------------------------------------------------------------------------------------------
template <int D, class R>
__global__ void __launch_bounds__(256)
    k_acoustic_max(int64_t i0, int64_t n, const R *__restrict__ vel, R c0, unsigned long long *__restrict__ slot)
{
    int64_t i = i0 + blockIdx.x * int64_t(blockDim.x) + threadIdx.x;
    R s = R(0);
    for (; i < n; i += int64_t(gridDim.x) * blockDim.x)
    {
        R v[D];
#pragma unroll
        for (int a = 0; a < D; ++a)
            v[a] = vel[i * D + a];
        s = fmax(s, c0 + sqrt_r(sqnormExact<D, R>(v)));
    }
    s = warpMax<R>(s);
    if ((threadIdx.x & 31) == 0)
        atomicMaxNonNeg(slot, double(s));
}

template <int D, class R>
__global__ void __launch_bounds__(256)
    k_advection_max(int64_t i0, int64_t n, const R *__restrict__ vel, const R *__restrict__ force, const R *__restrict__ fprior,
                    const R *__restrict__ mass, R h, unsigned long long *__restrict__ slot)
{
    int64_t i = i0 + blockIdx.x * int64_t(blockDim.x) + threadIdx.x;
    R s = R(0);
    for (; i < n; i += int64_t(gridDim.x) * blockDim.x)
    {
        R v[D], a[D];
#pragma unroll
        for (int d = 0; d < D; ++d)
        {
            v[d] = vel[i * D + d];
            a[d] = force[i * D + d] + fprior[i * D + d];
        }
        // fluid_time_step.cpp:54-59
        R acc_scale = R(4.0) * h * sqrt_r(sqnormExact<D, R>(a)) / mass[i];
        s = fmax(s, fmax(sqnormExact<D, R>(v), acc_scale));
    }
    s = warpMax<R>(s);
    if ((threadIdx.x & 31) == 0)
        atomicMaxNonNeg(slot, double(s));
}

// kind 0: total kinetic energy, 1: total mechanical energy (general_reduce.cpp:54-79), 2: max |v|
template <int D, class R>
__global__ void __launch_bounds__(256)
    k_observable(int64_t i0, int64_t n, int kind, const R *__restrict__ vel, const R *__restrict__ mass, const R *__restrict__ pos,
                 R gx, R gy, R gz, double *__restrict__ partial)
{
    __shared__ double sh[8];
    const R g[3] = {gx, gy, gz};
    double s = 0.0;
    for (int64_t i = i0 + blockIdx.x * int64_t(blockDim.x) + threadIdx.x; i < n; i += int64_t(gridDim.x) * blockDim.x)
    {
        R v[D];
#pragma unroll
        for (int a = 0; a < D; ++a)
            v[a] = vel[i * D + a];
        R v2 = sqnormExact<D, R>(v);
        if (kind == 2)
        {
            s = fmax(s, double(sqrt_r(v2)));
        }
        else
        {
            R e = R(0.5) * mass[i] * v2;
            if (kind == 1)
            {
                R pot = R(0);
#pragma unroll
                for (int a = 0; a < D; ++a)
                    pot += g[a] * (R(0) - pos[i * D + a]);
                e += mass[i] * pot;
            }
            s += double(e);
        }
    }
    s = kind == 2 ? warpMax<double>(s) : warpSum<double>(s);
    if ((threadIdx.x & 31) == 0)
        sh[threadIdx.x >> 5] = s;
    __syncthreads();
    if (threadIdx.x == 0)
    {
        double t = sh[0];
        for (int w = 1; w < 8; ++w)
            t = kind == 2 ? fmax(t, sh[w]) : t + sh[w];
        partial[blockIdx.x] = t;
    }
}
__global__ void k_observable_final(int nb, int kind, const double *__restrict__ partial, double *__restrict__ out)
{
    if (threadIdx.x == 0 && blockIdx.x == 0)
    {
        double t = partial[0];
        for (int b = 1; b < nb; ++b)
            t = kind == 2 ? fmax(t, partial[b]) : t + partial[b];
        *out = t;
    }
}

// ---------------------------------------------------------------------------------------------
// forces prior to the acoustic loop
// ---------------------------------------------------------------------------------------------
template <int D, class R>
__global__ void __launch_bounds__(256)
    k_gravity(int64_t i0, int64_t n, const R *__restrict__ mass, R gx, R gy, R gz, R *__restrict__ cur, R *__restrict__ prev,
              R *__restrict__ fprior)
{
    int64_t i = (i0 & ~int64_t(31)) + blockIdx.x * int64_t(blockDim.x) + threadIdx.x;
    if (i < i0 || i >= n)
        return;
    const R g[3] = {gx, gy, gz};
#pragma unroll
    for (int a = 0; a < D; ++a)
    {
        R c = mass[i] * g[a];
        cur[i * D + a] = c;
        fprior[i * D + a] += c - prev[i * D + a]; // force_prior.hpp:22-26
        prev[i * D + a] = c;
    }
}

template <int D, class R>
__global__ void __launch_bounds__(128)
    k_viscous(int64_t i0, int64_t n, const typename Real4<R>::type *__restrict__ xV, ListP<R> in, ContactP<R> wall, KernelP<R> kern,
              KernelP<R> kern_w, PeriodicP<D, R> per, R mu, R h, const R *__restrict__ vel, R *__restrict__ vf, R *__restrict__ prev,
              R *__restrict__ fprior)
{
    int64_t i = (i0 & ~int64_t(31)) + blockIdx.x * int64_t(blockDim.x) + threadIdx.x;
    if (i < i0 || i >= n)
        return;
    R xi[D], Vi, vi[D];
    loadXj<D, R>(xV, u32(i), xi, Vi);
#pragma unroll
    for (int a = 0; a < D; ++a)
        vi[a] = vel[i * D + a];
    const R mu_ij = R(2.0) * mu * mu * (R(1.0) / (mu + mu)); // particle_functors.h:163-177
    R f[D];
#pragma unroll
    for (int a = 0; a < D; ++a)
        f[a] = R(0);
    const u32 cnt = in.count[i];
    const u32 *lst = listBase(in.list, in.slice_off, i);
#pragma unroll 2
    for (u32 k = 0; k < cnt; ++k)
    {
        u32 j = listEntry(lst, k);
        R xj[D], Vj, vj[D];
        loadXj<D, R>(xV, j, xj, Vj);
        loadVec<D, R>(vel, j, vj);
        periodicImage<D, R>(xi, xj, per);
        Pair<D, R> g;
        pairGeometry<D, R>(xi, xj, kern, g);
        R ee = R(0);
#pragma unroll
        for (int a = 0; a < D; ++a)
            ee += g.e[a] * ((R(1.0) + R(1.0)) * g.e[a]); // e·(K_i+K_j)e with unit correction matrices
        R c = ee * mu_ij * g.dW * Vj / (g.r + R(0.01) * h);
#pragma unroll
        for (int a = 0; a < D; ++a)
            f[a] += c * (vi[a] - vj[a]);
    }
    // ViscousForce<Contact<Wall>> (viscous_dynamics.hpp:92-115): 2 (v_i - v_wall) / (r + 0.01 h), 2 e.(B_i e) mu, added with += onto
    // the inner force (two separate sums times Vol_i, as the reference forms them)
    R fw[D];
#pragma unroll
    for (int a = 0; a < D; ++a)
        fw[a] = R(0);
    for (int t = 0; t < wall.n; ++t)
    {
        const TargetP<R> &T = wall.t[t];
        const u32 cw = T.count[i];
        const u32 *lw = listBase(T.list, T.slice_off, i);
        for (u32 k = 0; k < cw; ++k)
        {
            u32 j = listEntry(lw, k);
            R xj[D], Vj, vw[D];
            loadXj<D, R>(T.xV, j, xj, Vj);
            loadVec<D, R>(T.vel, j, vw);
            Pair<D, R> g;
            pairGeometry<D, R>(xi, xj, kern_w, g);
            R ee = R(0);
#pragma unroll
            for (int a = 0; a < D; ++a)
                ee += g.e[a] * g.e[a];
            R c = R(2.0) * ee * mu_ij * g.dW * Vj / (g.r + R(0.01) * h);
#pragma unroll
            for (int a = 0; a < D; ++a)
                fw[a] += c * (R(2.0) * (vi[a] - vw[a]));
        }
    }
#pragma unroll
    for (int a = 0; a < D; ++a)
    {
        R v = f[a] * Vi + fw[a] * Vi;
        vf[i * D + a] = v;
        fprior[i * D + a] += v - prev[i * D + a];
        prev[i * D + a] = v;
    }
}

template <int D, class R>
__global__ void __launch_bounds__(128)
    k_transport_velocity(int64_t i0, int64_t n, const typename Real4<R>::type *__restrict__ xV, ListP<R> in, KernelP<R> kern,
                         PeriodicP<D, R> per, R h2, R scaling, R *__restrict__ kgi, R *__restrict__ pos)
{
    int64_t i = (i0 & ~int64_t(31)) + blockIdx.x * int64_t(blockDim.x) + threadIdx.x;
    if (i < i0 || i >= n)
        return;
    R xi[D], Vi;
    loadXj<D, R>(xV, u32(i), xi, Vi);
    R inc[D];
#pragma unroll
    for (int a = 0; a < D; ++a)
        inc[a] = R(0);
    const u32 cnt = in.count[i];
    const u32 *lst = listBase(in.list, in.slice_off, i);
#pragma unroll 2
    for (u32 k = 0; k < cnt; ++k)
    {
        u32 j = listEntry(lst, k);
        R xj[D], Vj;
        loadXj<D, R>(xV, j, xj, Vj);
        periodicImage<D, R>(xi, xj, per);
        Pair<D, R> g;
        pairGeometry<D, R>(xi, xj, kern, g);
        R c = (R(1.0) + R(1.0)) * g.dW * Vj;
#pragma unroll
        for (int a = 0; a < D; ++a)
            inc[a] -= c * g.e[a];
    }
    R sq = R(0);
#pragma unroll
    for (int a = 0; a < D; ++a)
    {
        kgi[i * D + a] = inc[a];
        sq += inc[a] * inc[a];
    }
    R lim = fmin(R(100.0) * (h2 * sq), R(1)); // TruncatedLinear, slope 100 (common_functors.h:82-94)
#pragma unroll
    for (int a = 0; a < D; ++a)
        pos[i * D + a] += scaling * lim * inc[a];
}

// ObservingAQuantity: out[i] = Σ W_ij V_j q_j / (Σ W_ij V_j + tiny)
template <int D, class R>
__global__ void __launch_bounds__(128)
    k_observe(int64_t i0, int64_t n, const typename Real4<R>::type *__restrict__ xV_obs, ContactP<R> ct, KernelP<R> kern,
              const R *q0, const R *q1, const R *q2, const R *q3, int width, double *__restrict__ out)
{
    int64_t i = (i0 & ~int64_t(31)) + blockIdx.x * int64_t(blockDim.x) + threadIdx.x;
    if (i < i0 || i >= n)
        return;
    const R *qs[kMaxTargets] = {q0, q1, q2, q3};
    R xi[D], Vi;
    loadXj<D, R>(xV_obs, u32(i), xi, Vi);
    R acc[9];
    for (int d = 0; d < 9; ++d)
        acc[d] = R(0);
    R w = R(0);
    for (int t = 0; t < ct.n; ++t)
    {
        const TargetP<R> &T = ct.t[t];
        const u32 cnt = T.count[i];
        const u32 *lst = listBase(T.list, T.slice_off, i);
        for (u32 k = 0; k < cnt; ++k)
        {
            u32 j = listEntry(lst, k);
            R xj[D], Vj;
            loadXj<D, R>(T.xV, j, xj, Vj);
            R r2 = R(0);
#pragma unroll
            for (int a = 0; a < D; ++a)
            {
                R d = xi[a] - xj[a];
                r2 += d * d;
            }
            R wj = kernelW<R>(sqrt_r(r2), kern) * Vj;
            for (int d = 0; d < width; ++d)
                acc[d] += wj * qs[t][size_t(j) * width + d];
            w += wj;
        }
    }
    for (int d = 0; d < width; ++d)
        out[i * width + d] = double(acc[d] / (w + Tiny<R>::value));
}

} // namespace sphb
#ifndef SPHB_QUAD_DEFAULT_F64
#define SPHB_QUAD_DEFAULT_F64 2
#endif
#ifndef SPHB_QUAD_DEFAULT_F32
#define SPHB_QUAD_DEFAULT_F32 0
#endif
#include "tile.cuh"
namespace sphb
{
// ---------------------------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------------------------
// The register-tiled kernels (tile.cuh) serve non-periodic bodies by default.  Measured on B200 at 20 M
// particles, ms K1+K2 / K3 / density / advection step: fp64 6.35 / 8.59 / 4.3 / 97 tiled against 11.4 / 10.6 / 6.4 / 140
// per-particle; fp32 4.32 / 3.84 / 2.6 / 59.4 against 4.38 / 4.09 / 3.9 / 62.1 (fp32 is issue-bound, the masked pairs
// cost nearly what the saved gathers give; the step gains because no index list has to be written).
// SPHB200_TILE=0 / 1 forces the choice (A/B timing and the parity tests of both paths).
static int tileMode()
{
    static int mode = -2;
    if (mode == -2)
    {
        const char *e = getenv("SPHB200_TILE");
        mode = !e ? -1 : (e[0] == '0' ? 0 : 1);
    }
    return mode;
}
template <int D, class R>
static bool tileUsable(sphb200_relation *inner, sphb200_relation *wall)
{
    const int mode = inner->ctx->pair_kernel == -2 ? tileMode() : inner->ctx->pair_kernel;
    if (mode == 0)
        return false;
    if (inner->body->P<D, R>().per.mask || !inner->lists[0].sorted)
        return false;
    if (wall)
        for (auto &L : wall->lists)
            if (!L.sorted)
                return false;
    return true;
}
// wall pre-pass of the tiled pair kernels (tile.cuh): one contact body, one thread per union list
static bool wallPrepass(sphb200_relation *wall)
{
    const char *e = getenv("SPHB200_WALL_PREPASS"); // read per call: the parity test runs both ways in one process
    const bool on = !(e && e[0] == '0');
    return on && wall && wall->lists.size() == 1 && kWideLanes == 1;
}
template <int D, class R>
static sphb200_status makeUContactP(sphb200_relation *wall, const ContactP<R> &cp, UContactP<R> &out, bool wide)
{
    out.n = cp.n;
    for (int t = 0; t < cp.n; ++t)
    {
        NeighborList &L = wall->lists[size_t(t)];
        wide = wide && kWideLanes > 1;
        if (wide)
            SPHB_TRY((ensureUnion<D, R, TileG<R>::value * kWideLanes>(wall, size_t(t), L.uni_wide, kWideLanes)));
        else
            SPHB_TRY((ensureUnion<D, R, TileG<R>::value>(wall, size_t(t), L.uni, 1)));
        out.t[t].u = ulistOf(wide ? L.uni_wide : L.uni);
        out.t[t].xV = cp.t[t].xV;
        out.t[t].vel = cp.t[t].vel;
        out.t[t].acc = cp.t[t].acc;
        out.t[t].nrm = cp.t[t].nrm;
    }
    return SPHB200_OK;
}

template <class R>
static ListP<R> listOf(const NeighborList &L)
{
    ListP<R> l;
    l.count = L.count;
    l.slice_off = L.slice_off;
    l.list = L.list;
    return l;
}

static sphb200_status checkRelation(sphb200_relation *r, bool want_inner, const char *who)
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

// need: bit 0 mass, 1 velocity, 2 acceleration, 3 normal, 4 the caller walks union lists: no per-particle index list
template <int D, class R>
static sphb200_status makeContactP(sphb200_relation *c, int need, ContactP<R> &out, KernelP<R> &kern_c)
{
    out.n = 0;
    if (!c)
        return SPHB200_OK;
    sphb200_body *src = c->body;
    out.n = int(c->targets.size());
    kern_c = src->P<D, R>().kern;
    for (int t = 0; t < out.n; ++t)
    {
        sphb200_body *tb = c->targets[size_t(t)];
        SPHB_TRY((ensurePacked<D, R>(tb)));
        TargetP<R> &T = out.t[t];
        if (!(need & 16))
            SPHB_TRY(ensureList(c, size_t(t)));
        const NeighborList &L = c->lists[size_t(t)];
        T.count = L.count;
        T.slice_off = L.slice_off;
        T.list = L.list;
        T.xV = static_cast<const typename Real4<R>::type *>(tb->xV.p);
        T.mass = T.vel = T.acc = T.nrm = nullptr;
        R *ptr;
        if (need & 1)
        {
            SPHB_TRY(getVarPtr<R>(tb, "Mass", 1, &ptr));
            T.mass = ptr;
        }
        if (need & 2)
        {
            SPHB_TRY(getVarPtr<R>(tb, "Velocity", D, &ptr));
            T.vel = ptr;
        }
        if (need & 4)
        {
            SPHB_TRY(getVarPtr<R>(tb, "Acceleration", D, &ptr));
            T.acc = ptr;
        }
        if (need & 8)
        {
            SPHB_TRY(getVarPtr<R>(tb, "NormalDirection", D, &ptr));
            T.nrm = ptr;
        }
        T.inv_rho0 = R(1.0) / R(tb->rho0);
        if (tb->P<D, R>().h > kern_c.h)
            kern_c = tb->P<D, R>().kern;
    }
    return SPHB200_OK;
}

template <int D, class R>
static sphb200_status densitySummation(sphb200_relation *inner, sphb200_relation *contact, int free_surface)
{
    sphb200_ctx *ctx = inner->ctx;
    sphb200_body *b = inner->body;
    SPHB_TRY(checkRelation(inner, true, "density_summation"));
    if (contact)
        SPHB_TRY(checkRelation(contact, false, "density_summation"));
    if (b->own_end <= b->own_begin)
        return SPHB200_OK;
    BodyP<D, R> &P = b->P<D, R>();
    SPHB_TRY((ensurePacked<D, R>(b)));
    ContactP<R> ct;
    KernelP<R> kern_ct = P.kern;
    const bool tiled = tileUsable<D, R>(inner, contact);
    SPHB_TRY((makeContactP<D, R>(contact, tiled ? 1 | 16 : 1, ct, kern_ct)));
    R *mass, *rho_sum, *rho, *vol;
    SPHB_TRY(getVarPtr<R>(b, "Mass", 1, &mass, true));
    SPHB_TRY(getVarPtr<R>(b, "DensitySummation", 1, &rho_sum));
    SPHB_TRY(getVarPtr<R>(b, "Density", 1, &rho, true));
    SPHB_TRY(getVarPtr<R>(b, "VolumetricMeasure", 1, &vol, true));
    auto *xV = static_cast<typename Real4<R>::type *>(b->xV.p);
    if (tiled)
    {
        constexpr int G = TileG<R>::value;
        SPHB_TRY((ensureUnion<D, R, G>(inner, 0, inner->lists[0].uni, 1)));
        const UnionList &U = inner->lists[0].uni;
        UContactP<R> uct;
        uct.n = 0;
        if (contact)
            SPHB_TRY((makeUContactP<D, R>(contact, ct, uct, false)));
        k_density_summation_tile<D, R, G><<<gridFor(U.n_groups, 128), 128, 0, ctx->stream>>>(
            U.base, b->own_begin, b->own_end, U.n_groups, xV, ulistOf(U), ct, uct, P.kern, kern_ct, R(b->rho0), R(1.0) / P.sigma0, mass, rho_sum,
            rho, vol, xV, free_surface);
    }
    else
    {
        SPHB_TRY(ensureList(inner, 0));
        k_density_summation<D, R><<<gridFor(b->own_end - (b->own_begin & ~int64_t(31)), 128), 128, 0, ctx->stream>>>(
        b->own_begin, b->own_end, xV, listOf<R>(inner->lists[0]), ct, P.kern, kern_ct, P.per, R(b->rho0), R(1.0) / P.sigma0, mass, rho_sum,
        rho, vol, xV, free_surface);
    }
    SPHB_LAUNCH_CHECK(ctx);
    if (!free_surface)
    {
        b->rec_b_valid = false; // Vol changed
        SPHB_TRY((refreshVolRange<D, R>(b)));
    }
    return SPHB200_OK;
}

template <int D, class R>
static sphb200_status ensureRecBuffers(sphb200_body *b)
{
    size_t need = size_t(b->n > 0 ? b->n : 1) * sizeof(Rec8<R>);
    SPHB_CUDA(b->ctx, b->rec_a.ensure(need));
    SPHB_CUDA(b->ctx, b->rec_b.ensure(need));
    return SPHB200_OK;
}

// stage bit 0: initialization (K1), bit 1: interaction + update (K2); 3 = the whole Dynamics1Level::exec
template <int D, class R>
static sphb200_status firstHalf(sphb200_relation *inner, sphb200_relation *wall, int riemann, double dt_, int stage)
{
    sphb200_ctx *ctx = inner->ctx;
    sphb200_body *b = inner->body;
    SPHB_TRY(checkRelation(inner, true, "integration_1st_half"));
    if (wall)
        SPHB_TRY(checkRelation(wall, false, "integration_1st_half"));
    if (b->own_end <= b->own_begin)
        return SPHB200_OK;
    BodyP<D, R> &P = b->P<D, R>();
    SPHB_TRY((ensurePacked<D, R>(b)));
    SPHB_TRY((ensureRecBuffers<D, R>(b)));
    FluidP<R> fl = makeFluidP<R>(b, P.h, riemann != 0);
    if (dt_ < 0.0 && !fl.ctl)
    { // dt < 0 = the device-resident dt of the sub-step loop's control block (split stages of a slab run)
        SPHB_TRY(ensureStepCtl(ctx));
        fl.ctl = static_cast<const StepCtl *>(ctx->step_ctl.p);
    }
    R *rho, *p, *drho, *mass, *pos, *vel, *force, *fprior;
    SPHB_TRY(getVarPtr<R>(b, "Density", 1, &rho, true));
    SPHB_TRY(getVarPtr<R>(b, "Pressure", 1, &p, true));
    SPHB_TRY(getVarPtr<R>(b, "DensityChangeRate", 1, &drho, true));
    SPHB_TRY(getVarPtr<R>(b, "Mass", 1, &mass, true));
    SPHB_TRY(getVarPtr<R>(b, "Position", D, &pos, true));
    SPHB_TRY(getVarPtr<R>(b, "Velocity", D, &vel, true));
    SPHB_TRY(getVarPtr<R>(b, "Force", D, &force, true));
    SPHB_TRY(getVarPtr<R>(b, "ForcePrior", D, &fprior, true));
    const R dt = R(dt_);
    auto *xV = static_cast<const typename Real4<R>::type *>(b->xV.p);
    auto *rec_a = static_cast<Rec8<R> *>(b->rec_a.p);
    auto *rec_b = static_cast<Rec8<R> *>(b->rec_b.p);
    const bool slab = b->own_begin > 0 || b->own_end < b->n;
    if (stage & 1)
    {
        const int64_t span = b->own_end - (b->own_begin & ~int64_t(31));
        if (slab && stage == 3)
        {
            // halo copies are not visited as i but serve as neighbours j: their records come from the
            // named variables here; a slab driver that uses the split stages overwrites them with the
            // owner's records between the stages instead
            k_pack_rec<D, R><<<gridFor(b->n, 256), 256, 0, ctx->stream>>>(b->n, xV, p, vel, rec_a);
            SPHB_LAUNCH_CHECK(ctx);
        }
        k_first_half_init<D, R><<<gridFor(span, 256), 256, 0, ctx->stream>>>(b->own_begin, b->own_end, dt, fl.p0, fl.rho0, xV, rho, drho,
                                                                           p, pos, vel, rec_a, fl.ctl, static_cast<unsigned long long *>(ctx->reduce_dev.p));
        SPHB_LAUNCH_CHECK(ctx);
    }
    if (stage & 2)
    {
        ContactP<R> wp;
        KernelP<R> kern_w = P.kern;
        const bool tiled = tileUsable<D, R>(inner, wall);
        SPHB_TRY((makeContactP<D, R>(wall, tiled ? 4 | 16 : 4, wp, kern_w)));
        if (!tiled)
            SPHB_TRY(ensureList(inner, 0));
        if (!fl.ctl) // the device-resident loop resets the slot in K1, and only while the loop is live
            SPHB_TRY(reduceReset(ctx, 0, false));
        auto *slot = static_cast<unsigned long long *>(ctx->reduce_dev.p);
        const int grid = groupGrid(b->own_begin, b->own_end, 128);
        if (tiled)
        {
            constexpr int G = TileG<R>::value;
            SPHB_TRY((ensureUnion<D, R, G>(inner, 0, inner->lists[0].uni, 1)));
            UContactP<R> uw;
            SPHB_TRY((makeUContactP<D, R>(wall, wp, uw, false)));
            const UnionList &U = inner->lists[0].uni;
            auto *vmm = static_cast<const unsigned long long *>(b->vol_mm.p);
            if (wallPrepass(wall))
            {
                UnionList &UW = wall->lists[0].uni;
                SPHB_TRY(ensureActive(ctx, UW));
                SPHB_CUDA(ctx, b->wall_sum.ensure(size_t(b->n) * (D + 1) * sizeof(R)));
                R *ws = static_cast<R *>(b->wall_sum.p);
                k_first_half_wall<D, R, G><<<148 * 4, 128, 0, ctx->stream>>>(UW.base, b->own_begin, b->own_end, UW.queue, rec_a, uw.t[0],
                                                                           makeTileKernelP<R>(kern_w), rho, mass, fprior, ws, fl.ctl);
                SPHB_LAUNCH_CHECK(ctx);
                const int quad = quadLanesK2<R>();
                if (quad == 2)
                    k_first_half_quad<D, R, G, 2><<<gridFor(U.n_groups * 2, 128), 128, 0, ctx->stream>>>(
                        U.base, b->own_begin, b->own_end, U.n_groups, rec_a, rec_b, ulistOf(U), uw.t[0].u.count, makeTileKernelP<R>(P.kern), fl, dt,
                        rho, mass, fprior, force, drho, vel, slot, vmm, ws);
                else if (quad == 4)
                    k_first_half_quad<D, R, G, 4><<<gridFor(U.n_groups * 4, 128), 128, 0, ctx->stream>>>(
                        U.base, b->own_begin, b->own_end, U.n_groups, rec_a, rec_b, ulistOf(U), uw.t[0].u.count, makeTileKernelP<R>(P.kern), fl, dt,
                        rho, mass, fprior, force, drho, vel, slot, vmm, ws);
                else
                k_first_half_tile<D, R, G, true><<<gridFor(U.n_groups, 128), 128, 0, ctx->stream>>>(
                    U.base, b->own_begin, b->own_end, U.n_groups, rec_a, rec_b, ulistOf(U), uw, makeTileKernelP<R>(P.kern),
                    makeTileKernelP<R>(kern_w), fl, dt, rho, mass, fprior, force, drho, vel, slot, vmm, ws);
            }
            else
                k_first_half_tile<D, R, G, false><<<gridFor(U.n_groups, 128), 128, 0, ctx->stream>>>(
                    U.base, b->own_begin, b->own_end, U.n_groups, rec_a, rec_b, ulistOf(U), uw, makeTileKernelP<R>(P.kern),
                    makeTileKernelP<R>(kern_w), fl, dt, rho, mass, fprior, force, drho, vel, slot, vmm, nullptr);
        }
        else if (P.per.mask)
            k_first_half<D, R, true><<<grid, 128, 0, ctx->stream>>>(b->own_begin, b->own_end, rec_a, rec_b, listOf<R>(inner->lists[0]), wp,
                                                                  P.kern, kern_w, P.per, fl, dt, rho, mass, fprior, force, drho, vel, slot);
        else
            k_first_half<D, R, false><<<grid, 128, 0, ctx->stream>>>(b->own_begin, b->own_end, rec_a, rec_b, listOf<R>(inner->lists[0]), wp,
                                                                   P.kern, kern_w, P.per, fl, dt, rho, mass, fprior, force, drho, vel, slot);
        SPHB_LAUNCH_CHECK(ctx);
        for (auto *ob : ctx->bodies)
            ob->acoustic_max_valid = false;
        b->acoustic_max_valid = true;
        // rec_b now holds {x_cfg, Vol | p, v_new} of every owned particle; with the split stages the
        // driver completes the halo records, otherwise the 2nd half re-packs them from the variables
        b->rec_b_valid = !slab || stage != 3;
    }
    return SPHB200_OK;
}

// lanes per union group in the 2nd-half kernel: 0 = one thread per group (k_second_half_tile), 2 / 4 = k_second_half_quad
// Measured at 20.1 M particles (B200, ms per launch): fp64 tile 8.61, 2 lanes 7.99, 4 lanes 8.36; fp32 tile 3.80, 2 lanes 4.63,
// 4 lanes 4.54 (fp32 is issue-bound: the duplicated member state and the shuffles cost more than the merged sectors give).
#ifndef SPHB_QUAD_K2_DEFAULT_F64
#define SPHB_QUAD_K2_DEFAULT_F64 0
#endif
// the same for the 1st-half kernel (needs the wall pre-pass when there is a wall): env SPHB200_QUAD_K2 = 0 | 2 | 4
template <class R>
static int quadLanesK2()
{
    const char *e = getenv("SPHB200_QUAD_K2");
    if (!e || !*e)
        return sizeof(R) == 8 ? SPHB_QUAD_K2_DEFAULT_F64 : 0;
    const int v = atoi(e);
    return (v == 2 || v == 4) ? v : 0;
}
template <class R>
static int quadLanes()
{
    const char *e = getenv("SPHB200_QUAD"); // read per call: the parity test runs the variants in one process
    if (!e)
        return sizeof(R) == 8 ? SPHB_QUAD_DEFAULT_F64 : SPHB_QUAD_DEFAULT_F32;
    return e[0] == '4' ? 4 : (e[0] == '2' ? 2 : 0);
}
template <int D, class R>
static sphb200_status secondHalf(sphb200_relation *inner, sphb200_relation *wall, int riemann, double dt_)
{
    sphb200_ctx *ctx = inner->ctx;
    sphb200_body *b = inner->body;
    SPHB_TRY(checkRelation(inner, true, "integration_2nd_half"));
    if (wall)
        SPHB_TRY(checkRelation(wall, false, "integration_2nd_half"));
    if (b->own_end <= b->own_begin)
        return SPHB200_OK;
    BodyP<D, R> &P = b->P<D, R>();
    SPHB_TRY((ensurePacked<D, R>(b)));
    SPHB_TRY((ensureRecBuffers<D, R>(b)));
    ContactP<R> wp;
    KernelP<R> kern_w = P.kern;
    const bool tiled = tileUsable<D, R>(inner, wall);
    SPHB_TRY((makeContactP<D, R>(wall, tiled ? 2 | 8 | 16 : 2 | 8, wp, kern_w)));
    if (!tiled)
        SPHB_TRY(ensureList(inner, 0));
    FluidP<R> fl = makeFluidP<R>(b, P.h, riemann != 0);
    if (dt_ < 0.0 && !fl.ctl)
    {
        SPHB_TRY(ensureStepCtl(ctx));
        fl.ctl = static_cast<const StepCtl *>(ctx->step_ctl.p);
    }
    R *rho, *drho, *pos, *vel, *force, *p;
    SPHB_TRY(getVarPtr<R>(b, "Density", 1, &rho, true));
    SPHB_TRY(getVarPtr<R>(b, "DensityChangeRate", 1, &drho, true));
    SPHB_TRY(getVarPtr<R>(b, "Position", D, &pos, true));
    SPHB_TRY(getVarPtr<R>(b, "Velocity", D, &vel, true));
    SPHB_TRY(getVarPtr<R>(b, "Force", D, &force, true));
    SPHB_TRY(getVarPtr<R>(b, "Pressure", 1, &p, true));
    auto *rec_b = static_cast<Rec8<R> *>(b->rec_b.p);
    if (!b->rec_b_valid)
    {
        k_pack_rec<D, R><<<gridFor(b->n, 256), 256, 0, ctx->stream>>>(b->n, static_cast<const typename Real4<R>::type *>(b->xV.p), p, vel,
                                                                    rec_b);
        SPHB_LAUNCH_CHECK(ctx);
        b->rec_b_valid = (b->own_begin == 0 && b->own_end == b->n);
    }
    if (b->own_begin > 0 || b->own_end < b->n)
        b->rec_b_valid = false; // one use only: the next 2nd half needs the halo records refreshed again
    if (tiled)
    {
        constexpr int G = TileG<R>::value;
        // 2nd half: kWideLanes adjacent threads share the union list of their 2 * kWideLanes particles
        UnionList &U = kWideLanes > 1 ? inner->lists[0].uni_wide : inner->lists[0].uni;
        SPHB_TRY((ensureUnion<D, R, G * kWideLanes>(inner, 0, U, kWideLanes)));
        UContactP<R> uw;
        SPHB_TRY((makeUContactP<D, R>(wall, wp, uw, true)));
        const int64_t n_threads = (b->own_end - U.base + G - 1) / G;
        const int quad = quadLanes<R>();
        if (quad && kWideLanes == 1 && (!wall || wallPrepass(wall)))
        { // QL lanes per group split the union list (tile.cuh k_second_half_quad)
            const R *ws = nullptr;
            const u32 *wcount = nullptr;
            if (wall)
            {
                UnionList &UW = wall->lists[0].uni;
                SPHB_TRY(ensureActive(ctx, UW));
                SPHB_CUDA(ctx, b->wall_sum.ensure(size_t(b->n) * (D + 1) * sizeof(R)));
                R *wsum = static_cast<R *>(b->wall_sum.p);
                k_second_half_wall<D, R, G><<<148 * 4, 128, 0, ctx->stream>>>(UW.base, b->own_begin, b->own_end, UW.queue, rec_b, uw.t[0],
                                                                            makeTileKernelP<R>(kern_w), R(fl.limiter_slope * fl.inv_c0_ave), wsum, fl.ctl);
                SPHB_LAUNCH_CHECK(ctx);
                ws = wsum;
                wcount = uw.t[0].u.count;
            }
            const int64_t lanes = n_threads * quad;
            if (quad == 4 && wall)
                k_second_half_quad<D, R, G, 4, true><<<gridFor(lanes, 128), 128, 0, ctx->stream>>>(U.base, b->own_begin, b->own_end, n_threads, rec_b, ulistOf(U), wcount, makeTileKernelP<R>(P.kern), fl, R(dt_), pos, rho, drho, force, ws);
            else if (quad == 4)
                k_second_half_quad<D, R, G, 4, false><<<gridFor(lanes, 128), 128, 0, ctx->stream>>>(U.base, b->own_begin, b->own_end, n_threads, rec_b, ulistOf(U), wcount, makeTileKernelP<R>(P.kern), fl, R(dt_), pos, rho, drho, force, ws);
            else if (wall)
                k_second_half_quad<D, R, G, 2, true><<<gridFor(lanes, 128), 128, 0, ctx->stream>>>(U.base, b->own_begin, b->own_end, n_threads, rec_b, ulistOf(U), wcount, makeTileKernelP<R>(P.kern), fl, R(dt_), pos, rho, drho, force, ws);
            else
                k_second_half_quad<D, R, G, 2, false><<<gridFor(lanes, 128), 128, 0, ctx->stream>>>(U.base, b->own_begin, b->own_end, n_threads, rec_b, ulistOf(U), wcount, makeTileKernelP<R>(P.kern), fl, R(dt_), pos, rho, drho, force, ws);
        }
        else if (wallPrepass(wall))
        {
            UnionList &UW = wall->lists[0].uni;
            SPHB_TRY(ensureActive(ctx, UW));
            SPHB_CUDA(ctx, b->wall_sum.ensure(size_t(b->n) * (D + 1) * sizeof(R)));
            R *ws = static_cast<R *>(b->wall_sum.p);
            k_second_half_wall<D, R, G><<<148 * 4, 128, 0, ctx->stream>>>(UW.base, b->own_begin, b->own_end, UW.queue, rec_b, uw.t[0],
                                                                        makeTileKernelP<R>(kern_w), R(fl.limiter_slope * fl.inv_c0_ave), ws, fl.ctl);
            SPHB_LAUNCH_CHECK(ctx);
            k_second_half_tile<D, R, G, kWideLanes, true><<<gridFor(n_threads, 128), 128, 0, ctx->stream>>>(
                U.base, b->own_begin, b->own_end, n_threads, rec_b, ulistOf(U), uw, makeTileKernelP<R>(P.kern), makeTileKernelP<R>(kern_w),
                fl, R(dt_), pos, rho, drho, force, ws);
        }
        else
            k_second_half_tile<D, R, G, kWideLanes, false><<<gridFor(n_threads, 128), 128, 0, ctx->stream>>>(
                U.base, b->own_begin, b->own_end, n_threads, rec_b, ulistOf(U), uw, makeTileKernelP<R>(P.kern), makeTileKernelP<R>(kern_w),
                fl, R(dt_), pos, rho, drho, force, nullptr);
    }
    else if (P.per.mask)
        k_second_half<D, R, true><<<groupGrid(b->own_begin, b->own_end, 128), 128, 0, ctx->stream>>>(b->own_begin, b->own_end, rec_b, listOf<R>(inner->lists[0]), wp,
                                                                             P.kern, kern_w, P.per, fl, R(dt_), pos, rho, drho, force);
    else
        k_second_half<D, R, false><<<groupGrid(b->own_begin, b->own_end, 128), 128, 0, ctx->stream>>>(b->own_begin, b->own_end, rec_b, listOf<R>(inner->lists[0]), wp,
                                                                              P.kern, kern_w, P.per, fl, R(dt_), pos, rho, drho, force);
    SPHB_LAUNCH_CHECK(ctx);
    return SPHB200_OK;
}

// ---- device-resident sub-step loop control (StepCtl, internal.cuh) ----
static sphb200_status ensureStepCtl(sphb200_ctx *ctx)
{
    if (ctx->step_ctl.p)
        return SPHB200_OK;
    SPHB_CUDA(ctx, ctx->step_ctl.ensure(sizeof(StepCtl)));
    SPHB_CUDA(ctx, cudaMemsetAsync(ctx->step_ctl.p, 0, sizeof(StepCtl), ctx->stream));
    return SPHB200_OK;
}
__global__ void k_ctl_begin(StepCtl *ctl, double Dt, double dt_carry)
{
    if (Dt >= 0.0)
        ctl->Dt = Dt; // else: left by sphb200_advection_dt(..., NULL)
    if (dt_carry >= 0.0)
        ctl->dt = dt_carry;
    ctl->relax = 0.0;
    ctl->count = 0;
    ctl->done = 0;
}
// fluid_time_step.cpp:61-79 outputResult on the device (same expression, same rounding as the host variant)
template <class R>
__global__ void k_advection_ctl(StepCtl *ctl, const unsigned long long *slot, R cfl, R h, R speed_ref)
{
    const R m = R(__longlong_as_double((long long)*slot));
    const R speed_max = sqrt_r(m);
    ctl->Dt = double(cfl * h / (fmax(speed_max, speed_ref) + Tiny<R>::value));
}
// phase 0 (before the sub-step, orders 0 and 2): dt from the max slot.  phase 1 (after it): bookkeeping, order 1: new dt
template <class R>
__global__ void k_substep_ctl(StepCtl *ctl, const unsigned long long *slot, R cfl, R h, int order, int phase)
{
    if (ctl->done)
        return;
    const R m = R(__longlong_as_double((long long)*slot));
    const double dt_new = double(cfl * h / (m + Tiny<R>::value)); // fluid_time_step.cpp:27-31
    if (phase == 0)
    {
        double dt = dt_new;
        if (order == 2)
            dt = fmin(dt, ctl->Dt); // taylor_green.cpp:170 SMIN(get_fluid_time_step_size.exec(), Dt)
        ctl->dt = dt;
        return;
    }
    if (order == 1)
        ctl->dt = dt_new;
    const double dt = ctl->dt;
    ctl->relax += dt;
    ctl->time += dt;
    ctl->count += 1;
    ctl->count_total += 1;
    if (!(ctl->relax < ctl->Dt))
        ctl->done = 1;
}

template <int D, class R>
static sphb200_status acousticDt(sphb200_body *b, double cfl, double *dt_out)
{
    sphb200_ctx *ctx = b->ctx;
    BodyP<D, R> &P = b->P<D, R>();
    if (!b->acoustic_max_valid)
    {
        R *vel;
        SPHB_TRY(getVarPtr<R>(b, "Velocity", D, &vel, true));
        SPHB_TRY(reduceReset(ctx, 0, false));
        if (b->own_end > b->own_begin)
        {
            k_acoustic_max<D, R><<<148 * 4, 256, 0, ctx->stream>>>(b->own_begin, b->own_end, vel, R(b->c0),
                                                                  static_cast<unsigned long long *>(ctx->reduce_dev.p));
            SPHB_LAUNCH_CHECK(ctx);
        }
    }
    if (!dt_out)
        return SPHB200_OK;
    double m;
    SPHB_TRY(reduceRead(ctx, 0, &m));
    if (b->n == 0)
        m = double(-std::numeric_limits<R>::max()); // reduce_functors.h:73-82 reference value MinReal
    // fluid_time_step.cpp:27-31
    *dt_out = double(R(cfl) * P.h / (R(m) + Tiny<R>::value));
    return SPHB200_OK;
}

template <int D, class R>
static sphb200_status advectionDt(sphb200_body *b, double U_ref, double cfl, int viscous, double *dt_out)
{
    sphb200_ctx *ctx = b->ctx;
    BodyP<D, R> &P = b->P<D, R>();
    R *vel, *force, *fprior, *mass;
    SPHB_TRY(getVarPtr<R>(b, "Velocity", D, &vel, true));
    SPHB_TRY(getVarPtr<R>(b, "Force", D, &force, true));
    SPHB_TRY(getVarPtr<R>(b, "ForcePrior", D, &fprior, true));
    SPHB_TRY(getVarPtr<R>(b, "Mass", 1, &mass, true));
    SPHB_TRY(reduceReset(ctx, 1, false));
    if (b->own_end > b->own_begin)
    {
        k_advection_max<D, R><<<148 * 4, 256, 0, ctx->stream>>>(b->own_begin, b->own_end, vel, force, fprior, mass, P.h,
                                                               static_cast<unsigned long long *>(ctx->reduce_dev.p) + 1);
        SPHB_LAUNCH_CHECK(ctx);
    }
    // fluid_time_step.cpp:61-79
    R speed_ref = R(U_ref);
    if (viscous)
        speed_ref = std::max(R(b->mu) / R(b->rho0) / P.h, speed_ref);
    if (!dt_out)
    { // device-resident: Dt goes to the sub-step loop's control block, the host is not involved
        SPHB_TRY(ensureStepCtl(ctx));
        k_advection_ctl<R><<<1, 1, 0, ctx->stream>>>(static_cast<StepCtl *>(ctx->step_ctl.p),
                                                     static_cast<const unsigned long long *>(ctx->reduce_dev.p) + 1, R(cfl), P.h, speed_ref);
        SPHB_LAUNCH_CHECK(ctx);
        return SPHB200_OK;
    }
    double m;
    SPHB_TRY(reduceRead(ctx, 1, &m));
    R speed_max = std::sqrt(R(m));
    *dt_out = double(R(cfl) * P.h / (std::max(speed_max, speed_ref) + Tiny<R>::value));
    return SPHB200_OK;
}

// ---------------------------------------------------------------------------------------------
// The acoustic sub-step loop of one advection step, device-resident (no host round trip per sub-step).
// order 0: dt = AcousticTimeStep; step(dt); relax += dt            (2-D dam break, Dambreak.cpp:183-192)
// order 1: step(dt carried from the previous sub-step); dt = AcousticTimeStep; relax += dt   (dambreak.cpp:168-177)
// order 2: dt = min(AcousticTimeStep, Dt); relax += dt; step(dt)   (taylor_green.cpp:167-176)
// ---------------------------------------------------------------------------------------------
template <int D, class R>
static sphb200_status acousticRelaxation(sphb200_relation *inner, sphb200_relation *wall, int riemann1, int riemann2, double Dt,
                                         double cfl, int order, double dt_carry, int max_substeps, int resume, double *out4)
{
    sphb200_ctx *ctx = inner->ctx;
    sphb200_body *b = inner->body;
    BodyP<D, R> &P = b->P<D, R>();
    SPHB_TRY(ensureStepCtl(ctx));
    auto *ctl = static_cast<StepCtl *>(ctx->step_ctl.p);
    auto *slot = static_cast<unsigned long long *>(ctx->reduce_dev.p);
    if (order != 1)
        SPHB_TRY((acousticDt<D, R>(b, cfl, nullptr))); // slot 0 = max(c0 + |v|) of the current velocities
    if (!resume)
    {
        k_ctl_begin<<<1, 1, 0, ctx->stream>>>(ctl, Dt, dt_carry);
        SPHB_LAUNCH_CHECK(ctx);
    }
    int launched = 0;
    StepCtl host;
    for (;;)
    {
        const int batch = out4 ? std::max(1, std::min(max_substeps - launched, launched == 0 ? max_substeps : 4)) : max_substeps;
        ctx->ctl_active = ctl;
        sphb200_status st = SPHB200_OK;
        for (int k = 0; k < batch && st == SPHB200_OK; ++k)
        {
            if (order != 1)
            {
                k_substep_ctl<R><<<1, 1, 0, ctx->stream>>>(ctl, slot, R(cfl), P.h, order, 0);
                ctx->launches++;
            }
            st = firstHalf<D, R>(inner, wall, riemann1, 0.0, 3);
            if (st == SPHB200_OK)
                st = secondHalf<D, R>(inner, wall, riemann2, 0.0);
            k_substep_ctl<R><<<1, 1, 0, ctx->stream>>>(ctl, slot, R(cfl), P.h, order, 1);
            ctx->launches++;
        }
        ctx->ctl_active = nullptr;
        SPHB_TRY(st);
        SPHB_CUDA(ctx, cudaGetLastError());
        launched += batch;
        if (!out4)
            return SPHB200_OK;
        SPHB_CUDA(ctx, cudaMemcpyAsync(ctx->reduce_host, ctl, sizeof(StepCtl), cudaMemcpyDeviceToHost, ctx->stream));
        SPHB_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
        memcpy(&host, ctx->reduce_host, sizeof(StepCtl));
        if (host.done || launched >= 100000)
            break;
        max_substeps = launched + 4; // keep going in small batches until the device says the loop has ended
    }
    out4[0] = double(host.count);
    out4[1] = host.dt;
    out4[2] = host.time;
    out4[3] = host.relax;
    out4[4] = double(host.done);
    return SPHB200_OK;
}

template <int D, class R>
static sphb200_status observable(sphb200_body *b, int kind, const double *g, double *out)
{
    sphb200_ctx *ctx = b->ctx;
    R *vel, *mass, *pos;
    SPHB_TRY(getVarPtr<R>(b, "Velocity", D, &vel, true));
    SPHB_TRY(getVarPtr<R>(b, "Mass", 1, &mass, true));
    SPHB_TRY(getVarPtr<R>(b, "Position", D, &pos, true));
    const int nb = 148 * 2;
    SPHB_CUDA(ctx, ctx->scratch_a.ensure(size_t(nb + 1) * 8));
    double *partial = static_cast<double *>(ctx->scratch_a.p);
    k_observable<D, R><<<nb, 256, 0, ctx->stream>>>(b->own_begin, b->own_end, kind, vel, mass, pos, g ? R(g[0]) : R(0), g ? R(g[1]) : R(0),
                                                    (g && D == 3) ? R(g[2]) : R(0), partial);
    SPHB_LAUNCH_CHECK(ctx);
    k_observable_final<<<1, 32, 0, ctx->stream>>>(nb, kind, partial, partial + nb);
    SPHB_LAUNCH_CHECK(ctx);
    SPHB_CUDA(ctx, cudaMemcpyAsync(ctx->reduce_host, partial + nb, 8, cudaMemcpyDeviceToHost, ctx->stream));
    SPHB_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    *out = *static_cast<double *>(ctx->reduce_host);
    return SPHB200_OK;
}

// PositionUpperBound / PositionLowerBound (general_dynamics/general_reduce.cpp:28-49, ReduceUpperBound / ReduceLowerBound
// reduce_functors.h): component-wise max / min of the particle positions - order-independent, hence exact.
template <int D, class R>
__global__ void __launch_bounds__(256) k_position_bound(int64_t i0, int64_t n, int upper, const R *__restrict__ pos, double *__restrict__ partial)
{
    __shared__ double sh[8][3];
    double m[3];
#pragma unroll
    for (int a = 0; a < 3; ++a)
        m[a] = upper ? -INFINITY : INFINITY;
    for (int64_t i = i0 + blockIdx.x * int64_t(blockDim.x) + threadIdx.x; i < n; i += int64_t(gridDim.x) * blockDim.x)
#pragma unroll
        for (int a = 0; a < D; ++a)
        {
            const double x = double(pos[i * D + a]);
            m[a] = upper ? fmax(m[a], x) : fmin(m[a], x);
        }
#pragma unroll
    for (int a = 0; a < 3; ++a)
        m[a] = upper ? warpMax<double>(m[a]) : warpMin<double>(m[a]);
    if ((threadIdx.x & 31) == 0)
        for (int a = 0; a < 3; ++a)
            sh[threadIdx.x >> 5][a] = m[a];
    __syncthreads();
    if (threadIdx.x < 3)
    {
        double t = sh[0][threadIdx.x];
        for (int w = 1; w < 8; ++w)
            t = upper ? fmax(t, sh[w][threadIdx.x]) : fmin(t, sh[w][threadIdx.x]);
        partial[blockIdx.x * 3 + threadIdx.x] = t;
    }
}
__global__ void k_position_bound_final(int nb, int upper, const double *__restrict__ partial, double *__restrict__ out)
{
    if (threadIdx.x < 3 && blockIdx.x == 0)
    {
        double t = partial[threadIdx.x];
        for (int k = 1; k < nb; ++k)
            t = upper ? fmax(t, partial[k * 3 + threadIdx.x]) : fmin(t, partial[k * 3 + threadIdx.x]);
        out[threadIdx.x] = t;
    }
}
template <int D, class R>
static sphb200_status positionBound(sphb200_body *b, int upper, double *out3)
{
    sphb200_ctx *ctx = b->ctx;
    R *pos;
    SPHB_TRY(getVarPtr<R>(b, "Position", D, &pos, true));
    const int nb = 148 * 2;
    SPHB_CUDA(ctx, ctx->scratch_a.ensure(size_t(nb + 1) * 3 * 8));
    double *partial = static_cast<double *>(ctx->scratch_a.p);
    k_position_bound<D, R><<<nb, 256, 0, ctx->stream>>>(b->own_begin, b->own_end, upper, pos, partial);
    SPHB_LAUNCH_CHECK(ctx);
    k_position_bound_final<<<1, 32, 0, ctx->stream>>>(nb, upper, partial, partial + nb * 3);
    SPHB_LAUNCH_CHECK(ctx);
    SPHB_CUDA(ctx, cudaMemcpyAsync(ctx->reduce_host, partial + nb * 3, 24, cudaMemcpyDeviceToHost, ctx->stream));
    SPHB_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    for (int a = 0; a < 3; ++a)
        out3[a] = a < D ? static_cast<double *>(ctx->reduce_host)[a] : 0.0;
    return SPHB200_OK;
}

template <int D, class R>
static sphb200_status gravityForce(sphb200_body *b, const double *g)
{
    sphb200_ctx *ctx = b->ctx;
    if (b->own_end <= b->own_begin)
        return SPHB200_OK;
    R *mass, *cur, *prev, *fprior;
    SPHB_TRY(getVarPtr<R>(b, "Mass", 1, &mass, true));
    SPHB_TRY(getVarPtr<R>(b, "GravityForce", D, &cur));
    SPHB_TRY(getVarPtr<R>(b, "PreviousGravityForce", D, &prev, true));
    SPHB_TRY(getVarPtr<R>(b, "ForcePrior", D, &fprior, true));
    k_gravity<D, R><<<gridFor(b->own_end - (b->own_begin & ~int64_t(31)), 256), 256, 0, ctx->stream>>>(b->own_begin, b->own_end, mass, R(g[0]), R(g[1]), D == 3 ? R(g[2]) : R(0),
                                                               cur, prev, fprior);
    SPHB_LAUNCH_CHECK(ctx);
    return SPHB200_OK;
}

template <int D, class R>
static sphb200_status viscousForce(sphb200_relation *inner, sphb200_relation *wall)
{
    sphb200_ctx *ctx = inner->ctx;
    sphb200_body *b = inner->body;
    SPHB_TRY(checkRelation(inner, true, "viscous_force"));
    if (wall)
        SPHB_TRY(checkRelation(wall, false, "viscous_force"));
    if (b->own_end <= b->own_begin)
        return SPHB200_OK;
    BodyP<D, R> &P = b->P<D, R>();
    SPHB_TRY((ensurePacked<D, R>(b)));
    R *vel, *vf, *prev, *fprior;
    SPHB_TRY(getVarPtr<R>(b, "Velocity", D, &vel, true));
    SPHB_TRY(getVarPtr<R>(b, "ViscousForce", D, &vf));
    SPHB_TRY(getVarPtr<R>(b, "PreviousViscousForce", D, &prev, true));
    SPHB_TRY(getVarPtr<R>(b, "ForcePrior", D, &fprior, true));
    SPHB_TRY(ensureList(inner, 0));
    ContactP<R> wp;
    KernelP<R> kern_w = P.kern;
    SPHB_TRY((makeContactP<D, R>(wall, 2, wp, kern_w))); // wall velocity; per-particle index lists
    k_viscous<D, R><<<gridFor(b->own_end - (b->own_begin & ~int64_t(31)), 128), 128, 0, ctx->stream>>>(
        b->own_begin, b->own_end, static_cast<const typename Real4<R>::type *>(b->xV.p), listOf<R>(inner->lists[0]), wp, P.kern, kern_w,
        P.per, R(b->mu), P.h, vel, vf, prev, fprior);
    SPHB_LAUNCH_CHECK(ctx);
    return SPHB200_OK;
}

template <int D, class R>
static sphb200_status transportVelocity(sphb200_relation *inner, double coefficient)
{
    sphb200_ctx *ctx = inner->ctx;
    sphb200_body *b = inner->body;
    SPHB_TRY(checkRelation(inner, true, "transport_velocity_correction"));
    if (b->own_end <= b->own_begin)
        return SPHB200_OK;
    BodyP<D, R> &P = b->P<D, R>();
    SPHB_TRY((ensurePacked<D, R>(b)));
    R *kgi, *pos;
    SPHB_TRY(getVarPtr<R>(b, "KernelGradientIntegral", D, &kgi));
    SPHB_TRY(getVarPtr<R>(b, "Position", D, &pos, true));
    R h2 = P.h * P.h;
    SPHB_TRY(ensureList(inner, 0));
    k_transport_velocity<D, R><<<gridFor(b->own_end - (b->own_begin & ~int64_t(31)), 128), 128, 0, ctx->stream>>>(
        b->own_begin, b->own_end, static_cast<const typename Real4<R>::type *>(b->xV.p), listOf<R>(inner->lists[0]), P.kern, P.per, h2,
        R(coefficient) * h2, kgi, pos);
    SPHB_LAUNCH_CHECK(ctx);
    return SPHB200_OK;
}

template <int D, class R>
static sphb200_status observe(sphb200_relation *contact, const char *name, int width, double *out)
{
    sphb200_ctx *ctx = contact->ctx;
    sphb200_body *b = contact->body;
    SPHB_TRY(checkRelation(contact, false, "observe"));
    if (b->own_end <= b->own_begin)
        return SPHB200_OK;
    SPHB_TRY((ensurePacked<D, R>(b)));
    ContactP<R> ct;
    KernelP<R> kern = b->P<D, R>().kern;
    SPHB_TRY((makeContactP<D, R>(contact, 0, ct, kern)));
    const R *q[kMaxTargets] = {nullptr, nullptr, nullptr, nullptr};
    for (int t = 0; t < ct.n; ++t)
    {
        R *ptr;
        Var *v = contact->targets[size_t(t)]->find(name);
        if (!v)
            return ctx->fail(SPHB200_ERR_NOT_FOUND, std::string("observe: variable '") + name + "' not registered on target");
        SPHB_TRY(getVarPtr<R>(contact->targets[size_t(t)], name, width, &ptr));
        q[t] = ptr;
    }
    SPHB_CUDA(ctx, ctx->scratch_a.ensure(size_t(b->n) * size_t(width) * 8));
    double *dev_out = static_cast<double *>(ctx->scratch_a.p);
    k_observe<D, R><<<gridFor(b->own_end - (b->own_begin & ~int64_t(31)), 128), 128, 0, ctx->stream>>>(
        b->own_begin, b->own_end, static_cast<const typename Real4<R>::type *>(b->xV.p), ct, kern, q[0], q[1], q[2], q[3], width, dev_out);
    SPHB_LAUNCH_CHECK(ctx);
    SPHB_CUDA(ctx, cudaMemcpyAsync(out, dev_out, size_t(b->n) * size_t(width) * 8, cudaMemcpyDeviceToHost, ctx->stream));
    SPHB_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return SPHB200_OK;
}
} // namespace sphb

extern "C"
{
sphb200_status sphb200_ctx_set_pair_kernel(sphb200_ctx *ctx, int mode)
{
    if (!ctx || mode < -1 || mode > 1)
        return SPHB200_ERR_INVALID;
    ctx->pair_kernel = mode;
    return SPHB200_OK;
}
sphb200_status sphb200_density_summation(sphb200_relation *inner, sphb200_relation *contact, int free_surface)
{
    if (!inner)
        return SPHB200_ERR_INVALID;
    cudaSetDevice(inner->ctx->device);
    return SPHB_DISPATCH(inner->ctx, densitySummation, inner, contact, free_surface);
}
sphb200_status sphb200_integration_1st_half(sphb200_relation *inner, sphb200_relation *wall, int riemann, double dt)
{
    if (!inner)
        return SPHB200_ERR_INVALID;
    cudaSetDevice(inner->ctx->device);
    return SPHB_DISPATCH(inner->ctx, firstHalf, inner, wall, riemann, dt, 3);
}
sphb200_status sphb200_integration_1st_half_init(sphb200_relation *inner, double dt)
{
    if (!inner)
        return SPHB200_ERR_INVALID;
    cudaSetDevice(inner->ctx->device);
    return SPHB_DISPATCH(inner->ctx, firstHalf, inner, nullptr, 1, dt, 1);
}
sphb200_status sphb200_integration_1st_half_interact(sphb200_relation *inner, sphb200_relation *wall, int riemann, double dt)
{
    if (!inner)
        return SPHB200_ERR_INVALID;
    cudaSetDevice(inner->ctx->device);
    return SPHB_DISPATCH(inner->ctx, firstHalf, inner, wall, riemann, dt, 2);
}
sphb200_status sphb200_integration_2nd_half(sphb200_relation *inner, sphb200_relation *wall, int riemann, double dt)
{
    if (!inner)
        return SPHB200_ERR_INVALID;
    cudaSetDevice(inner->ctx->device);
    return SPHB_DISPATCH(inner->ctx, secondHalf, inner, wall, riemann, dt);
}
sphb200_status sphb200_acoustic_dt(sphb200_body *b, double cfl, double *dt_out)
{
    if (!b)
        return SPHB200_ERR_INVALID;
    cudaSetDevice(b->ctx->device);
    return SPHB_DISPATCH(b->ctx, acousticDt, b, cfl, dt_out);
}
sphb200_status sphb200_acoustic_relaxation(sphb200_relation *inner, sphb200_relation *wall, int riemann_1st, int riemann_2nd, double Dt,
                                           double cfl, int order, double dt_carry, int max_substeps, int resume, double *out5)
{
    if (!inner || order < 0 || order > 2 || max_substeps < 1)
        return SPHB200_ERR_INVALID;
    if (wall && wall->ctx != inner->ctx)
        return inner->ctx->fail(SPHB200_ERR_INVALID, "acoustic_relaxation: relations from different contexts");
    cudaSetDevice(inner->ctx->device);
    return SPHB_DISPATCH(inner->ctx, acousticRelaxation, inner, wall, riemann_1st, riemann_2nd, Dt, cfl, order, dt_carry, max_substeps, resume, out5);
}
// the pieces of sphb200_acoustic_relaxation for a driver that interleaves its own work (halo exchange, all-reduce of the
// max slot over the ranks) with the stages: begin -> per sub-step { stages with dt < 0 ..., substep_ctl } -> relaxation_state
sphb200_status sphb200_relaxation_begin(sphb200_ctx *ctx, double Dt, double dt_carry)
{
    if (!ctx)
        return SPHB200_ERR_INVALID;
    cudaSetDevice(ctx->device);
    SPHB_TRY(ensureStepCtl(ctx));
    k_ctl_begin<<<1, 1, 0, ctx->stream>>>(static_cast<sphb::StepCtl *>(ctx->step_ctl.p), Dt, dt_carry);
    SPHB_LAUNCH_CHECK(ctx);
    return SPHB200_OK;
}
sphb200_status sphb200_substep_ctl(sphb200_body *b, double cfl, int order, int phase)
{
    if (!b || order < 0 || order > 2 || phase < 0 || phase > 1)
        return SPHB200_ERR_INVALID;
    sphb200_ctx *ctx = b->ctx;
    cudaSetDevice(ctx->device);
    SPHB_TRY(ensureStepCtl(ctx));
    auto *ctl = static_cast<sphb::StepCtl *>(ctx->step_ctl.p);
    auto *slot = static_cast<const unsigned long long *>(ctx->reduce_dev.p);
    if (ctx->real_bytes == 8)
        k_substep_ctl<double><<<1, 1, 0, ctx->stream>>>(ctl, slot, cfl, b->h_ratio * b->dp, order, phase);
    else
        k_substep_ctl<float><<<1, 1, 0, ctx->stream>>>(ctl, slot, float(cfl), float(b->h_ratio) * float(b->dp), order, phase);
    SPHB_LAUNCH_CHECK(ctx);
    return SPHB200_OK;
}
sphb200_status sphb200_reduce_slot_ptr(sphb200_ctx *ctx, int slot, void **out)
{
    if (!ctx || !out || slot < 0 || slot > 7)
        return SPHB200_ERR_INVALID;
    *out = static_cast<unsigned long long *>(ctx->reduce_dev.p) + slot;
    return SPHB200_OK;
}
sphb200_status sphb200_relaxation_state(sphb200_ctx *ctx, double *out5)
{
    if (!ctx || !out5)
        return SPHB200_ERR_INVALID;
    cudaSetDevice(ctx->device);
    if (!ctx->step_ctl.p)
        return ctx->fail(SPHB200_ERR_INVALID, "relaxation_state: no acoustic_relaxation has run yet");
    sphb::StepCtl host;
    SPHB_CUDA(ctx, cudaMemcpyAsync(ctx->reduce_host, ctx->step_ctl.p, sizeof(sphb::StepCtl), cudaMemcpyDeviceToHost, ctx->stream));
    SPHB_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    memcpy(&host, ctx->reduce_host, sizeof(host));
    out5[0] = double(host.count);
    out5[1] = host.dt;
    out5[2] = host.time;
    out5[3] = host.relax;
    out5[4] = double(host.done);
    return SPHB200_OK;
}
sphb200_status sphb200_advection_dt(sphb200_body *b, double U_ref, double cfl, int viscous, double *dt_out)
{
    if (!b)
        return SPHB200_ERR_INVALID;
    cudaSetDevice(b->ctx->device);
    return SPHB_DISPATCH(b->ctx, advectionDt, b, U_ref, cfl, viscous, dt_out);
}
sphb200_status sphb200_position_bound(sphb200_body *b, int upper, double *out3)
{
    if (!b || !out3)
        return SPHB200_ERR_INVALID;
    cudaSetDevice(b->ctx->device);
    return SPHB_DISPATCH(b->ctx, positionBound, b, upper, out3);
}
sphb200_status sphb200_gravity_force(sphb200_body *b, const double *g)
{
    if (!b || !g)
        return SPHB200_ERR_INVALID;
    cudaSetDevice(b->ctx->device);
    return SPHB_DISPATCH(b->ctx, gravityForce, b, g);
}
sphb200_status sphb200_viscous_force(sphb200_relation *inner, sphb200_relation *wall)
{
    if (!inner || (wall && wall->ctx != inner->ctx))
        return SPHB200_ERR_INVALID;
    cudaSetDevice(inner->ctx->device);
    return SPHB_DISPATCH(inner->ctx, viscousForce, inner, wall);
}
sphb200_status sphb200_transport_velocity_correction(sphb200_relation *inner, double coefficient)
{
    if (!inner)
        return SPHB200_ERR_INVALID;
    cudaSetDevice(inner->ctx->device);
    return SPHB_DISPATCH(inner->ctx, transportVelocity, inner, coefficient);
}
sphb200_status sphb200_total_kinetic_energy(sphb200_body *b, double *out)
{
    if (!b || !out)
        return SPHB200_ERR_INVALID;
    cudaSetDevice(b->ctx->device);
    return SPHB_DISPATCH(b->ctx, observable, b, 0, nullptr, out);
}
sphb200_status sphb200_total_mechanical_energy(sphb200_body *b, const double *g, double *out)
{
    if (!b || !out || !g)
        return SPHB200_ERR_INVALID;
    cudaSetDevice(b->ctx->device);
    return SPHB_DISPATCH(b->ctx, observable, b, 1, g, out);
}
sphb200_status sphb200_maximum_speed(sphb200_body *b, double *out)
{
    if (!b || !out)
        return SPHB200_ERR_INVALID;
    cudaSetDevice(b->ctx->device);
    return SPHB_DISPATCH(b->ctx, observable, b, 2, nullptr, out);
}
sphb200_status sphb200_observe(sphb200_relation *contact, const char *name, int width, double *out)
{
    if (!contact || !name || !out || width < 1 || width > 9)
        return SPHB200_ERR_INVALID;
    cudaSetDevice(contact->ctx->device);
    return SPHB_DISPATCH(contact->ctx, observe, contact, name, width, out);
}
} // extern "C"
