// Total-Lagrangian elastic solid dynamics (test_3d_twisting_column class cases) as sm_100a kernels.
//
// Reference local dynamics reproduced:
//   general_dynamics/kernel_correction.cpp:8-29, common/vector_functions.h:198-203  (B matrix, alpha = 0)
//   solid_dynamics/elastic_dynamics.h:171-207 (Integration1stHalfPK2 interaction), :259-287 (Decomposed)
//   solid_dynamics/elastic_dynamics.cpp:96-118,162-184 (update / initialisations), :186-194 (2nd half)
//   solid_dynamics/elastic_dynamics.h:325-351 (Integration2ndHalf interaction)
//   solid_dynamics/elastic_dynamics.cpp:12-30 (AcousticTimeStep, ReduceMin)
//   materials/elastic_solid.cpp:54-80,109-132, elastic_solid.h:90-96 (Neo-Hookean stress, damping)
//   general_dynamics/general_constraint.h:106-121 (FixBodyPartConstraint)
// The neighbour configuration is the one frozen at the reference state (built once), so W/dW/e/r come
// from the frozen xV records; the shear / damping terms use the live Position as the reference does.
#include "internal.cuh"
#include "smat.cuh"

using namespace sphb;

namespace sphb
{

template <class R>
struct SolidP
{
    R rho0, G0, K0, lambda0, c0, cs0, h, inv_W0;
};

struct SList
{
    const u32 *count, *slice_off, *list;
};

template <int D, class R>
__device__ __forceinline__ void loadXs(const typename Real4<R>::type *__restrict__ xV, u32 j, R *x, R &w)
{
    typename Real4<R>::type v = ldg4(xV + j);
    x[0] = v.x;
    x[1] = v.y;
    if (D == 3)
        x[D - 1] = v.z;
    w = v.w;
}

// LinearGradientCorrectionMatrix<Inner<>>: interaction + update
template <int D, class R>
__global__ void __launch_bounds__(128)
    k_correction_matrix(int64_t i0, int64_t n, const typename Real4<R>::type *__restrict__ xV, SList in, KernelP<R> kern,
                        R *__restrict__ B)
{
    int64_t i = (i0 & ~int64_t(31)) + blockIdx.x * int64_t(blockDim.x) + threadIdx.x;
    if (i < i0 || i >= n)
        return;
    R xi[D], Vi;
    loadXs<D, R>(xV, u32(i), xi, Vi);
    SMat<D, R> local = SMat<D, R>::zero();
    const u32 cnt = in.count[i];
    const u32 *lst = listBase(in.list, in.slice_off, i);
    for (u32 k = 0; k < cnt; ++k)
    {
        u32 j = listEntry(lst, k);
        R xj[D], Vj;
        loadXs<D, R>(xV, j, xj, Vj);
        Pair<D, R> g;
        pairGeometry<D, R>(xi, xj, kern, g);
        R c = g.dW * Vj;
#pragma unroll
        for (int r = 0; r < D; ++r)
#pragma unroll
            for (int q = 0; q < D; ++q)
                local.a[r][q] -= (g.r * g.e[r]) * (c * g.e[q]);
    }
    // kernel_correction.cpp:19-29 with alpha = 0; inverseTikhonov, vector_functions.h:198-203
    const R sqrt_eps = R(1.0e-8);
    R det = local.det();
    R det_sqr = fmax(R(0) - det, R(0));
    SMat<D, R> Bt = local.T();
    SMat<D, R> inv = (Bt * local + SMat<D, R>::identity() * sqrt_eps).inverse() * Bt;
    R w1 = det / (det + det_sqr), w2 = det_sqr / (det + det_sqr);
    (inv * w1 + SMat<D, R>::identity() * w2).store(B + i * D * D);
}

// 1st half initialisation: DecomposedIntegration1stHalf (elastic_dynamics.cpp:169-184) or
// Integration1stHalfPK2 (:109-118)
template <int D, class R>
__global__ void __launch_bounds__(128)
    k_solid_first_init(int64_t i0, int64_t n, int decomposed, SolidP<R> sp, R dt, R *__restrict__ pos, const R *__restrict__ vel,
                       R *__restrict__ F, const R *__restrict__ dF, const R *__restrict__ B, R *__restrict__ rho,
                       R *__restrict__ Jterm, R *__restrict__ stress, R *__restrict__ invFT, R *__restrict__ pk1)
{
    int64_t i = (i0 & ~int64_t(31)) + blockIdx.x * int64_t(blockDim.x) + threadIdx.x;
    if (i < i0 || i >= n)
        return;
#pragma unroll
    for (int d = 0; d < D; ++d)
        pos[i * D + d] += vel[i * D + d] * dt * R(0.5);
    SMat<D, R> dFm = SMat<D, R>::load(dF + i * D * D);
    SMat<D, R> Fm = SMat<D, R>::load(F + i * D * D);
#pragma unroll
    for (int r = 0; r < D; ++r)
#pragma unroll
        for (int q = 0; q < D; ++q)
            Fm.a[r][q] += dFm.a[r][q] * dt * R(0.5);
    Fm.store(F + i * D * D);
    const R one_over_dim = R(1.0) / R(D), corr = R(1.07);
    if (decomposed)
    {
        R J = Fm.det();
        R one_over_J = R(1.0) / J;
        rho[i] = sp.rho0 * one_over_J;
        R jt = pow(one_over_J * one_over_J, one_over_dim);
        Jterm[i] = jt;
        SMat<D, R> iFT = Fm.inverse().T();
        iFT.store(invFT + i * D * D);
        R vol_kirchhoff = R(0.5) * sp.K0 * (J * J - R(1)); // NeoHookeanSolid::VolumetricKirchhoff
        R scalar = vol_kirchhoff - corr * sp.G0 * jt * (Fm * Fm.T()).trace() * one_over_dim;
        // NumericalDampingLeftCauchy (elastic_solid.h:90-96)
        SMat<D, R> strain_rate = (dFm * Fm.T() + Fm * dFm.T()) * R(0.5);
        SMat<D, R> normal_rate = SMat<D, R>::zero();
#pragma unroll
        for (int d = 0; d < D; ++d)
            normal_rate.a[d][d] = strain_rate.a[d][d];
        SMat<D, R> damping = ((strain_rate - normal_rate) * sp.cs0 + normal_rate * sp.c0) * (R(0.5) * sp.rho0) * sp.h;
        (iFT * scalar + damping * iFT).store(stress + i * D * D);
    }
    else
    {
        R J = Fm.det();
        rho[i] = sp.rho0 / J;
        // NeoHookeanSolid::StressPK2 (elastic_solid.cpp:109-117); PK1 = F * PK2, times B^T
        SMat<D, R> rc = Fm.T() * Fm;
        SMat<D, R> pk2 = SMat<D, R>::identity() * sp.G0 + rc.inverse() * ((sp.lambda0 * (J - R(1.0)) - sp.G0) * J);
        ((Fm * pk2) * SMat<D, R>::load(B + i * D * D).T()).store(pk1 + i * D * D);
    }
}
// 1st half interaction: force_i = Σ_j …  (elastic_dynamics.h:266-281 decomposed, :179-201 PK2)
template <int D, class R>
__global__ void __launch_bounds__(128)
    k_solid_first_interaction(int64_t i0, int64_t n, int decomposed, SolidP<R> sp, const typename Real4<R>::type *__restrict__ xV,
                              SList in, KernelP<R> kern, const R *__restrict__ pos, const R *__restrict__ vel,
                              const R *__restrict__ mass, const R *__restrict__ stress, const R *__restrict__ Jterm,
                              const R *__restrict__ F, const R *__restrict__ pk1, R *__restrict__ force)
{
    int64_t i = (i0 & ~int64_t(31)) + blockIdx.x * int64_t(blockDim.x) + threadIdx.x;
    if (i < i0 || i >= n)
        return;
    const R inv_rho0 = R(1.0) / sp.rho0, corr = R(1.07);
    R xi[D], Vi, pi[D], vi[D];
    loadXs<D, R>(xV, u32(i), xi, Vi);
#pragma unroll
    for (int d = 0; d < D; ++d)
    {
        pi[d] = pos[i * D + d];
        vi[d] = vel[i * D + d];
    }
    const R mi = mass[i];
    SMat<D, R> Si = SMat<D, R>::load((decomposed ? stress : pk1) + i * D * D);
    SMat<D, R> Fi = SMat<D, R>::load(F + i * D * D);
    const R Ji = decomposed ? Jterm[i] : R(0);
    R f[D];
#pragma unroll
    for (int d = 0; d < D; ++d)
        f[d] = R(0);
    const u32 cnt = in.count[i];
    const u32 *lst = listBase(in.list, in.slice_off, i);
    for (u32 k = 0; k < cnt; ++k)
    {
        u32 j = listEntry(lst, k);
        R xj[D], Vj;
        loadXs<D, R>(xV, j, xj, Vj);
        Pair<D, R> g;
        pairGeometry<D, R>(xi, xj, kern, g);
        R c = g.dW * Vj;
        R pj[D];
#pragma unroll
        for (int d = 0; d < D; ++d)
            pj[d] = __ldg(pos + size_t(j) * D + d);
        if (decomposed)
        {
            SMat<D, R> s = Si + SMat<D, R>::load(stress + size_t(j) * D * D);
            R se[D];
            s.mulVec(g.e, se);
            R shear_c = corr * sp.G0 * (Ji + __ldg(Jterm + j));
#pragma unroll
            for (int d = 0; d < D; ++d)
            {
                R shear = shear_c * (pi[d] - pj[d]) / g.r;
                f[d] += mi * (se[d] + shear) * c * inv_rho0;
            }
        }
        else
        {
            R dim_r = R(D) / g.r;
            R pv = R(0);
#pragma unroll
            for (int d = 0; d < D; ++d)
                pv += (pi[d] - pj[d]) * (vi[d] - __ldg(vel + size_t(j) * D + d));
            R strain_rate = dim_r * dim_r * pv;
            R weight = kernelW<R>(g.r, kern) * sp.inv_W0;
            R pair_damp = R(0.5) * sp.rho0 * sp.c0 * strain_rate * sp.h; // PairNumericalDamping
            SMat<D, R> num = (Fi + SMat<D, R>::load(F + size_t(j) * D * D)) * (R(0.5) * pair_damp);
            SMat<D, R> s = Si + SMat<D, R>::load(pk1 + size_t(j) * D * D) + num * (R(0.25) * weight);
            R se[D];
            s.mulVec(g.e, se);
#pragma unroll
            for (int d = 0; d < D; ++d)
                f[d] += mi * inv_rho0 * c * se[d];
        }
    }
#pragma unroll
    for (int d = 0; d < D; ++d)
        force[i * D + d] = f[d];
}

template <int D, class R>
__global__ void __launch_bounds__(256)
    k_solid_first_update(int64_t i0, int64_t n, R dt, const R *__restrict__ force, const R *__restrict__ fprior,
                         const R *__restrict__ mass, R *__restrict__ vel)
{
    int64_t i = (i0 & ~int64_t(31)) + blockIdx.x * int64_t(blockDim.x) + threadIdx.x;
    if (i < i0 || i >= n)
        return;
#pragma unroll
    for (int d = 0; d < D; ++d)
        vel[i * D + d] += (fprior[i * D + d] + force[i * D + d]) / mass[i] * dt; // elastic_dynamics.cpp:96-99
}

// 2nd half: initialization + interaction + update fused (all writes are to index i only and the
// interaction reads only Velocity of the neighbours, which this kernel does not modify)
template <int D, class R>
__global__ void __launch_bounds__(128)
    k_solid_second_half(int64_t i0, int64_t n, R dt, const typename Real4<R>::type *__restrict__ xV, SList in, KernelP<R> kern,
                        const R *__restrict__ vel, const R *__restrict__ B, R *__restrict__ pos, R *__restrict__ F,
                        R *__restrict__ dF)
{
    int64_t i = (i0 & ~int64_t(31)) + blockIdx.x * int64_t(blockDim.x) + threadIdx.x;
    if (i < i0 || i >= n)
        return;
    R xi[D], Vi, vi[D];
    loadXs<D, R>(xV, u32(i), xi, Vi);
#pragma unroll
    for (int d = 0; d < D; ++d)
    {
        vi[d] = vel[i * D + d];
        pos[i * D + d] += vi[d] * dt * R(0.5); // elastic_dynamics.cpp:186-189
    }
    SMat<D, R> rate = SMat<D, R>::zero();
    const u32 cnt = in.count[i];
    const u32 *lst = listBase(in.list, in.slice_off, i);
    for (u32 k = 0; k < cnt; ++k)
    {
        u32 j = listEntry(lst, k);
        R xj[D], Vj;
        loadXs<D, R>(xV, j, xj, Vj);
        Pair<D, R> g;
        pairGeometry<D, R>(xi, xj, kern, g);
        R c = g.dW * Vj;
#pragma unroll
        for (int r = 0; r < D; ++r)
        {
            R dv = vi[r] - __ldg(vel + size_t(j) * D + r);
#pragma unroll
            for (int q = 0; q < D; ++q)
                rate.a[r][q] -= dv * (c * g.e[q]);
        }
    }
    SMat<D, R> dFm = rate * SMat<D, R>::load(B + i * D * D); // elastic_dynamics.h:333-348
    dFm.store(dF + i * D * D);
#pragma unroll
    for (int r = 0; r < D; ++r)
#pragma unroll
        for (int q = 0; q < D; ++q)
            F[i * D * D + r * D + q] += dFm.a[r][q] * dt * R(0.5); // elastic_dynamics.cpp:191-194
}

template <int D, class R>
__global__ void k_solid_constrain(int64_t m, const long long *__restrict__ original_ids,
                                  const u32 *__restrict__ sorted_id, R *__restrict__ pos, const R *__restrict__ pos0,
                                  R *__restrict__ vel)
{
    int64_t k = blockIdx.x * int64_t(blockDim.x) + threadIdx.x;
    if (k >= m)
        return;
    u32 i = sorted_id[original_ids[k]];
#pragma unroll
    for (int d = 0; d < D; ++d)
    {
        pos[size_t(i) * D + d] = pos0[size_t(i) * D + d];
        vel[size_t(i) * D + d] = R(0);
    }
}

template <int D, class R>
__global__ void __launch_bounds__(256)
    k_solid_dt(int64_t i0, int64_t n, R cfl, R c0, R h, const R *__restrict__ vel, const R *__restrict__ force,
               const R *__restrict__ fprior, const R *__restrict__ mass, unsigned long long *__restrict__ slot)
{
    int64_t i = i0 + blockIdx.x * int64_t(blockDim.x) + threadIdx.x;
    R s = R(INFINITY);
    for (; i < n; i += int64_t(gridDim.x) * blockDim.x)
    {
        R a[D], v[D];
#pragma unroll
        for (int d = 0; d < D; ++d)
        {
            a[d] = (force[i * D + d] + fprior[i * D + d]) / mass[i];
            v[d] = vel[i * D + d];
        }
        R an = sqrt_r(sqnormExact<D, R>(a));
        R t = cfl * fmin(sqrt_r(h / (an + Tiny<R>::value)), h / (c0 + sqrt_r(sqnormExact<D, R>(v))));
        s = fmin(s, t);
    }
    s = warpMin<R>(s);
    if ((threadIdx.x & 31) == 0)
        atomicMinNonNeg(slot, double(s));
}
} // namespace sphb

namespace sphb
{
static sphb200_status checkInner(sphb200_relation *r, const char *who)
{
    sphb200_ctx *ctx = r->ctx;
    if (!r->inner)
        return ctx->fail(SPHB200_ERR_INVALID, std::string(who) + ": needs an inner relation");
    if (!r->valid || r->source_stamp != r->body->config_stamp)
        return ctx->fail(SPHB200_ERR_INVALID, std::string(who) + ": relation is stale (call relation_update)");
    return SPHB200_OK;
}
static SList slistOf(const NeighborList &L)
{
    SList s;
    s.count = L.count;
    s.slice_off = L.slice_off;
    s.list = L.list;
    return s;
}
template <int D, class R>
static SolidP<R> makeSolidP(sphb200_body *b)
{
    BodyP<D, R> &P = b->P<D, R>();
    SolidP<R> s;
    s.rho0 = R(b->rho0);
    s.G0 = R(b->G0);
    s.K0 = R(b->K0);
    s.lambda0 = R(b->lambda0);
    s.c0 = R(b->c0);
    s.cs0 = R(b->cs0);
    s.h = P.h;
    s.inv_W0 = R(1.0) / P.kern.factor_W;
    return s;
}

template <int D, class R>
static sphb200_status correctionMatrix(sphb200_relation *inner)
{
    sphb200_ctx *ctx = inner->ctx;
    sphb200_body *b = inner->body;
    SPHB_TRY(checkInner(inner, "correction_matrix"));
    if (b->n == 0)
        return SPHB200_OK;
    SPHB_TRY((ensurePacked<D, R>(b)));
    R *B;
    SPHB_TRY(getVarPtr<R>(b, "LinearGradientCorrectionMatrix", D * D, &B));
    SPHB_TRY(ensureList(inner, 0));
    k_correction_matrix<D, R><<<gridFor(b->own_end - (b->own_begin & ~int64_t(31)), 128), 128, 0, ctx->stream>>>(
        b->own_begin, b->own_end, static_cast<const typename Real4<R>::type *>(b->xV.p), slistOf(inner->lists[0]), b->P<D, R>().kern, B);
    SPHB_LAUNCH_CHECK(ctx);
    return SPHB200_OK;
}

// registerStateVariableData<Matd>("DeformationGradient", IdentityMatrix<Matd>::value) (elastic_dynamics.cpp:60,70) and
// registerStateVariableDataFrom<Vecd>("InitialPosition", "Position") (general_constraint.h:96): the initial value a variable
// gets when it is FIRST registered - a caller that registered and filled it before keeps its values
template <int D, class R>
__global__ void __launch_bounds__(256) k_fill_identity(int64_t n, R *__restrict__ F)
{
    const int64_t i = blockIdx.x * int64_t(blockDim.x) + threadIdx.x;
    if (i >= n)
        return;
#pragma unroll
    for (int a = 0; a < D; ++a)
#pragma unroll
        for (int c = 0; c < D; ++c)
            F[i * D * D + a * D + c] = a == c ? R(1) : R(0);
}
template <int D, class R>
static sphb200_status deformationGradientVar(sphb200_body *b, R **out)
{
    sphb200_ctx *ctx = b->ctx;
    const bool fresh = b->find("DeformationGradient") == nullptr;
    SPHB_TRY(getVarPtr<R>(b, "DeformationGradient", D * D, out, true));
    if (fresh && b->n > 0)
    {
        k_fill_identity<D, R><<<gridFor(b->n, 256), 256, 0, ctx->stream>>>(b->n, *out);
        SPHB_LAUNCH_CHECK(ctx);
    }
    return SPHB200_OK;
}
template <int D, class R>
static sphb200_status initialPositionVar(sphb200_body *b, R **out)
{
    sphb200_ctx *ctx = b->ctx;
    const bool fresh = b->find("InitialPosition") == nullptr;
    SPHB_TRY(getVarPtr<R>(b, "InitialPosition", D, out, true));
    if (fresh && b->n > 0)
    {
        R *pos;
        SPHB_TRY(getVarPtr<R>(b, "Position", D, &pos, true));
        SPHB_CUDA(ctx, cudaMemcpyAsync(*out, pos, size_t(b->n) * D * sizeof(R), cudaMemcpyDeviceToDevice, ctx->stream));
    }
    return SPHB200_OK;
}

template <int D, class R>
static sphb200_status solidFirstHalf(sphb200_relation *inner, int decomposed, double dt_)
{
    sphb200_ctx *ctx = inner->ctx;
    sphb200_body *b = inner->body;
    SPHB_TRY(checkInner(inner, "solid_1st_half"));
    if (b->n == 0)
        return SPHB200_OK;
    SPHB_TRY((ensurePacked<D, R>(b)));
    R *pos, *vel, *force, *fprior, *mass, *rho, *B, *F, *dF, *Jterm, *stress, *invFT, *pk1;
    SPHB_TRY(getVarPtr<R>(b, "Position", D, &pos, true));
    SPHB_TRY(getVarPtr<R>(b, "Velocity", D, &vel, true));
    SPHB_TRY(getVarPtr<R>(b, "Force", D, &force, true));
    SPHB_TRY(getVarPtr<R>(b, "ForcePrior", D, &fprior, true));
    SPHB_TRY(getVarPtr<R>(b, "Mass", 1, &mass, true));
    SPHB_TRY(getVarPtr<R>(b, "Density", 1, &rho, true));
    SPHB_TRY(getVarPtr<R>(b, "LinearGradientCorrectionMatrix", D * D, &B));
    SPHB_TRY((deformationGradientVar<D, R>(b, &F)));
    SPHB_TRY(getVarPtr<R>(b, "DeformationRate", D * D, &dF, true));
    SPHB_TRY(getVarPtr<R>(b, "DeterminantTerm", 1, &Jterm));
    SPHB_TRY(getVarPtr<R>(b, "StressOnParticle", D * D, &stress));
    SPHB_TRY(getVarPtr<R>(b, "InverseTransposedDeformation", D * D, &invFT));
    SPHB_TRY(getVarPtr<R>(b, "StressPK1OnParticle", D * D, &pk1));
    SolidP<R> sp = makeSolidP<D, R>(b);
    const R dt = R(dt_);
    auto *xV = static_cast<const typename Real4<R>::type *>(b->xV.p);
    k_solid_first_init<D, R><<<gridFor(b->own_end - (b->own_begin & ~int64_t(31)), 128), 128, 0, ctx->stream>>>(b->own_begin, b->own_end, decomposed, sp, dt, pos, vel, F, dF, B, rho,
                                                                          Jterm, stress, invFT, pk1);
    SPHB_LAUNCH_CHECK(ctx);
    SPHB_TRY(ensureList(inner, 0));
    k_solid_first_interaction<D, R><<<gridFor(b->own_end - (b->own_begin & ~int64_t(31)), 128), 128, 0, ctx->stream>>>(
        b->own_begin, b->own_end, decomposed, sp, xV, slistOf(inner->lists[0]), b->P<D, R>().kern, pos, vel, mass, stress, Jterm, F, pk1, force);
    SPHB_LAUNCH_CHECK(ctx);
    k_solid_first_update<D, R><<<gridFor(b->own_end - (b->own_begin & ~int64_t(31)), 256), 256, 0, ctx->stream>>>(b->own_begin, b->own_end, dt, force, fprior, mass, vel);
    SPHB_LAUNCH_CHECK(ctx);
    b->acoustic_max_valid = false;
    return SPHB200_OK;
}

template <int D, class R>
static sphb200_status solidSecondHalf(sphb200_relation *inner, double dt_)
{
    sphb200_ctx *ctx = inner->ctx;
    sphb200_body *b = inner->body;
    SPHB_TRY(checkInner(inner, "solid_2nd_half"));
    if (b->n == 0)
        return SPHB200_OK;
    SPHB_TRY((ensurePacked<D, R>(b)));
    R *pos, *vel, *B, *F, *dF;
    SPHB_TRY(getVarPtr<R>(b, "Position", D, &pos, true));
    SPHB_TRY(getVarPtr<R>(b, "Velocity", D, &vel, true));
    SPHB_TRY(getVarPtr<R>(b, "LinearGradientCorrectionMatrix", D * D, &B));
    SPHB_TRY((deformationGradientVar<D, R>(b, &F)));
    SPHB_TRY(getVarPtr<R>(b, "DeformationRate", D * D, &dF, true));
    SPHB_TRY(ensureList(inner, 0));
    k_solid_second_half<D, R><<<gridFor(b->own_end - (b->own_begin & ~int64_t(31)), 128), 128, 0, ctx->stream>>>(
        b->own_begin, b->own_end, R(dt_), static_cast<const typename Real4<R>::type *>(b->xV.p), slistOf(inner->lists[0]), b->P<D, R>().kern,
        vel, B, pos, F, dF);
    SPHB_LAUNCH_CHECK(ctx);
    return SPHB200_OK;
}

template <int D, class R>
static sphb200_status solidConstrain(sphb200_body *b, const int64_t *ids, int64_t m)
{
    sphb200_ctx *ctx = b->ctx;
    if (m == 0)
        return SPHB200_OK;
    R *pos, *pos0, *vel;
    SPHB_TRY(getVarPtr<R>(b, "Position", D, &pos, true));
    SPHB_TRY((initialPositionVar<D, R>(b, &pos0)));
    SPHB_TRY(getVarPtr<R>(b, "Velocity", D, &vel, true));
    SPHB_CUDA(ctx, ctx->scratch_a.ensure(size_t(m) * 8));
    SPHB_CUDA(ctx, cudaMemcpyAsync(ctx->scratch_a.p, ids, size_t(m) * 8, cudaMemcpyHostToDevice, ctx->stream));
    k_solid_constrain<D, R><<<gridFor(m, 128), 128, 0, ctx->stream>>>(m, static_cast<const long long *>(ctx->scratch_a.p),
                                                                      static_cast<const u32 *>(b->sorted_id.p), pos, pos0, vel);
    SPHB_LAUNCH_CHECK(ctx);
    SPHB_CUDA(ctx, cudaStreamSynchronize(ctx->stream)); // ids is caller memory
    b->acoustic_max_valid = false;
    return SPHB200_OK;
}

template <int D, class R>
static sphb200_status solidDt(sphb200_body *b, double cfl, double *out)
{
    sphb200_ctx *ctx = b->ctx;
    R *vel, *force, *fprior, *mass;
    SPHB_TRY(getVarPtr<R>(b, "Velocity", D, &vel, true));
    SPHB_TRY(getVarPtr<R>(b, "Force", D, &force, true));
    SPHB_TRY(getVarPtr<R>(b, "ForcePrior", D, &fprior, true));
    SPHB_TRY(getVarPtr<R>(b, "Mass", 1, &mass, true));
    SPHB_TRY(reduceReset(ctx, 2, true));
    if (b->n > 0)
    {
        k_solid_dt<D, R><<<148 * 2, 256, 0, ctx->stream>>>(b->own_begin, b->own_end, R(cfl), R(b->c0), b->P<D, R>().h, vel, force, fprior, mass,
                                                          static_cast<unsigned long long *>(ctx->reduce_dev.p) + 2);
        SPHB_LAUNCH_CHECK(ctx);
    }
    double m;
    SPHB_TRY(reduceRead(ctx, 2, &m));
    if (out)
        *out = b->n > 0 ? m : double(std::numeric_limits<R>::max());
    return SPHB200_OK;
}
} // namespace sphb

extern "C"
{
sphb200_status sphb200_correction_matrix(sphb200_relation *inner)
{
    if (!inner)
        return SPHB200_ERR_INVALID;
    cudaSetDevice(inner->ctx->device);
    return SPHB_DISPATCH(inner->ctx, correctionMatrix, inner);
}
sphb200_status sphb200_solid_1st_half(sphb200_relation *inner, int decomposed, double dt)
{
    if (!inner)
        return SPHB200_ERR_INVALID;
    cudaSetDevice(inner->ctx->device);
    return SPHB_DISPATCH(inner->ctx, solidFirstHalf, inner, decomposed, dt);
}
sphb200_status sphb200_solid_2nd_half(sphb200_relation *inner, double dt)
{
    if (!inner)
        return SPHB200_ERR_INVALID;
    cudaSetDevice(inner->ctx->device);
    return SPHB_DISPATCH(inner->ctx, solidSecondHalf, inner, dt);
}
sphb200_status sphb200_solid_constrain(sphb200_body *b, const int64_t *original_ids, int64_t n)
{
    if (!b || n < 0 || (n > 0 && !original_ids))
        return SPHB200_ERR_INVALID;
    cudaSetDevice(b->ctx->device);
    return SPHB_DISPATCH(b->ctx, solidConstrain, b, original_ids, n);
}
sphb200_status sphb200_solid_dt(sphb200_body *b, double cfl, double *dt_out)
{
    if (!b)
        return SPHB200_ERR_INVALID;
    cudaSetDevice(b->ctx->device);
    return SPHB_DISPATCH(b->ctx, solidDt, b, cfl, dt_out);
}
} // extern "C"
