// ORACLE - TEST INFRASTRUCTURE ONLY (see oracle_core.hpp header).
// CK-flavour local dynamics of the SYCL dam-break tests, restated per particle:
//   smoothing kernel        shared_ck/smoothing_kernel/kernel_tabulated_ck.h:37-71, .cpp:6-27 (20-interval cubic Lagrange table of
//                           Wendland C2), shared_ck/body_relation/neighbor_method.hpp:17-52,73-120 (inv_h = 1 / max(h_src, h_tar))
//   CompressionSummation + DensityRegularization   shared_ck/particle_dynamics/fluid_dynamics/density_regularization.hpp:42-111, .h:146-180
//   LinearCorrectionMatrix<Inner<WithUpdate>> + <Contact>   shared_ck/particle_dynamics/general_dynamics/kernel_correction_ck.hpp:36-96,
//                           common/vector_functions.h:198-203 (inverseTikhonov)
//   FreeSurfaceIndicationCK<Inner<WithUpdate>> + <Contact>  .../surface_indication/surface_indication_ck.hpp:55-170
// Neighbours are the relation's index sets; W, dW are re-evaluated with the tabulated kernel from the particle positions,
// as the CK kernels do (they store indices only).
#pragma once
#include "oracle_core.hpp"
#include "oracle_solid.hpp"

namespace orc
{
template <int D, class R>
struct TabKernel
{
    static constexpr int res = 20, size = 24;
    R dq, dq0, dq1, dq2, dq3, w[size], dw[size], dimfac, inv_h, src_inv_h;
    TabKernel(R h_src, R h_tar)
    {
        const R kernel_size = R(2.0);
        dq = kernel_size / R(res);
        for (int i = 0; i < size; ++i)
        {
            const R q = R(i - 1) * dq;
            w[i] = R(std::pow(R(1.0) - R(0.5) * q, 4)) * (R(1.0) + R(2.0) * q);
            dw[i] = R(0.625) * R(std::pow(q - R(2.0), 3)) * q;
        }
        dq0 = (R(-1.0) * dq) * (R(-2.0) * dq) * (R(-3.0) * dq);
        dq1 = dq * (R(-1.0) * dq) * (R(-2.0) * dq);
        dq2 = (R(2.0) * dq) * dq * (R(-1.0) * dq);
        dq3 = (R(3.0) * dq) * (R(2.0) * dq) * dq;
        // DimensionFactor = factor_W * h^D (base_kernel.h:91-93) with kernel_wendland_c2.cpp:11-13
        const R h = h_src, ih = R(1.0) / h;
        dimfac = D == 2 ? (ih * ih * R(7.0) / (R(4.0) * R(kPi))) * h * h : (ih * ih * ih * R(21.0) / (R(16.0) * R(kPi))) * h * h * h;
        src_inv_h = R(1.0) / h_src;
        inv_h = R(1.0) / std::max(h_src, h_tar);
    }
    R cubic(const R *data, R q) const
    {
        const int location = int(std::floor(q / dq));
        const int i = location + 1;
        const R f1 = q - R(location) * dq, f0 = f1 + dq, f2 = f1 - dq, f3 = f1 - R(2) * dq;
        return (f1 * f2 * f3) / dq0 * data[i - 1] + (f0 * f2 * f3) / dq1 * data[i] + (f0 * f1 * f3) / dq2 * data[i + 1] +
               (f0 * f1 * f2) / dq3 * data[i + 2];
    }
    R ipow(R x, int n) const
    {
        R r = R(1);
        for (int k = 0; k < n; ++k)
            r *= x;
        return r;
    }
    R W0() const { return ipow(src_inv_h, D) * dimfac * cubic(w, R(0)); }
    R W(R r) const { return ipow(inv_h, D) * dimfac * cubic(w, r * inv_h); }
    R dW(R r) const { return ipow(inv_h, D + 1) * dimfac * cubic(dw, r * inv_h); }
};

template <int D, class R>
inline R dispNorm(const R *xi, const R *xj, R *d)
{
    for (int a = 0; a < D; ++a)
        d[a] = xi[a] - xj[a];
    return std::sqrt(sqnorm<D, R>(d));
}

// registerStateVariableFrom<Real>("VolumetricMeasureRef", "VolumetricMeasure"): a copy taken when first registered
template <int D, class R>
inline R *volRef(Body<D, R> &b)
{
    const bool fresh = b.vars.find("VolumetricMeasureRef") == b.vars.end();
    R *ref = b.v("VolumetricMeasureRef", 1);
    if (fresh)
    {
        b.evolving("VolumetricMeasureRef");
        const R *vol = b.v("VolumetricMeasure");
        for (size_t i = 0; i < b.n; ++i)
            ref[i] = vol[i];
    }
    return ref;
}

template <int D, class R>
void compressionRegularization(Relation<D, R> &inner, Relation<D, R> *contact, bool free_surface)
{
    Body<D, R> &b = *inner.body;
    const R *pos = b.v("Position", D);
    R *vref = volRef<D, R>(b), *csum = b.v("CompressionSummation"), *comp = b.v("Compression"), *rho = b.v("Density");
    const TabKernel<D, R> k(b.h, b.h);
    const long n = long(b.n);
#pragma omp parallel for schedule(static)
    for (long i = 0; i < n; ++i)
    {
        R s = k.W0() * vref[i];
        const auto &nb = inner.configs[0][i];
        for (size_t m = 0; m != nb.current_size; ++m)
        {
            R d[D];
            s += k.W(dispNorm<D, R>(pos + i * D, pos + nb.j[m] * D, d)) * vref[nb.j[m]];
        }
        csum[i] = s;
    }
    if (contact)
        for (size_t t = 0; t < contact->targets.size(); ++t)
        {
            Body<D, R> &tb = *contact->targets[t];
            const R *tpos = tb.v("Position", D);
            const R *tref = volRef<D, R>(tb);
            const TabKernel<D, R> kc(b.h, tb.h);
#pragma omp parallel for schedule(static)
            for (long i = 0; i < n; ++i)
            {
                const auto &cn = contact->configs[t][i];
                for (size_t m = 0; m != cn.current_size; ++m)
                {
                    R d[D];
                    csum[i] += kc.W(dispNorm<D, R>(pos + i * D, tpos + cn.j[m] * D, d)) * tref[cn.j[m]];
                }
            }
        }
    for (long i = 0; i < n; ++i)
    {
        comp[i] = free_surface ? std::max(csum[i], R(1)) : csum[i];
        rho[i] = comp[i] * R(b.mat.rho0);
    }
}

template <int D, class R>
void linearCorrectionMatrixCK(Relation<D, R> &inner, Relation<D, R> *contact, R alpha)
{
    Body<D, R> &b = *inner.body;
    const R *pos = b.v("Position", D), *vol = b.v("VolumetricMeasure");
    R *B = b.v("LinearCorrectionMatrix", D * D);
    const TabKernel<D, R> k(b.h, b.h);
    const long n = long(b.n);
    auto accumulate = [&](long i, const R *xj, R Vj, const TabKernel<D, R> &kk)
    {
        R d[D];
        const R r = dispNorm<D, R>(pos + i * D, xj, d);
        const R dw = kk.dW(r);
        for (int a = 0; a < D; ++a)
            for (int c = 0; c < D; ++c)
                B[i * D * D + a * D + c] -= d[a] * ((dw * (d[c] / r)) * Vj); // vec_r_ij * (nablaW_ij V_j)^T
    };
#pragma omp parallel for schedule(static)
    for (long i = 0; i < n; ++i)
    {
        for (int q = 0; q < D * D; ++q)
            B[i * D * D + q] = R(0);
        const auto &nb = inner.configs[0][i];
        for (size_t m = 0; m != nb.current_size; ++m)
            accumulate(i, pos + nb.j[m] * D, vol[nb.j[m]], k);
    }
    if (contact)
        for (size_t t = 0; t < contact->targets.size(); ++t)
        {
            Body<D, R> &tb = *contact->targets[t];
            const R *tpos = tb.v("Position", D), *tvol = tb.v("VolumetricMeasure");
            const TabKernel<D, R> kc(b.h, tb.h);
#pragma omp parallel for schedule(static)
            for (long i = 0; i < n; ++i)
            {
                const auto &cn = contact->configs[t][i];
                for (size_t m = 0; m != cn.current_size; ++m)
                    accumulate(i, tpos + cn.j[m] * D, tvol[cn.j[m]], kc);
            }
        }
#pragma omp parallel for schedule(static)
    for (long i = 0; i < n; ++i)
    {
        Mat<D, R> M = Mat<D, R>::load(B + i * D * D);
        const R det = M.det();
        const R det_sqr = std::max(alpha - det, R(0));
        Mat<D, R> Mt = M.T();
        Mat<D, R> inv = (Mt * M + Mat<D, R>::identity() * R(1.0e-8)).inverse() * Mt;
        const R weight = det / (det + det_sqr);
        (inv * weight + Mat<D, R>::identity() * (R(1.0) - weight)).store(B + i * D * D);
    }
}

// Indicator / PreviousSurfaceIndicator are int variables in the reference; here Real-typed 0 / 1
template <int D, class R>
void freeSurfaceIndicationCK(Relation<D, R> &inner, Relation<D, R> *contact)
{
    Body<D, R> &b = *inner.body;
    const R *pos = b.v("Position", D), *vol = b.v("VolumetricMeasure");
    const bool fresh = b.vars.find("PreviousSurfaceIndicator") == b.vars.end();
    R *ind = b.v("Indicator"), *pdiv = b.v("PositionDivergence"), *prev = b.v("PreviousSurfaceIndicator", 1);
    const long n = long(b.n);
    if (fresh)
    {
        b.evolving("PreviousSurfaceIndicator");
        for (long i = 0; i < n; ++i)
            prev[i] = R(1);
    }
    const TabKernel<D, R> k(b.h, b.h);
    const R threshold = R(0.75) * R(D), h = b.h;
#pragma omp parallel for schedule(static)
    for (long i = 0; i < n; ++i)
    {
        R pd = R(0);
        bool near_prev = false;
        const auto &nb = inner.configs[0][i];
        for (size_t m = 0; m != nb.current_size; ++m)
        {
            R d[D];
            const R r = dispNorm<D, R>(pos + i * D, pos + nb.j[m] * D, d);
            pd -= k.dW(r) * vol[nb.j[m]] * r;
            near_prev = near_prev || prev[nb.j[m]] == R(1);
        }
        if (pd < threshold && prev[i] != R(1) && !near_prev)
            pd = R(2.0) * threshold;
        pdiv[i] = pd;
    }
    if (contact)
        for (size_t t = 0; t < contact->targets.size(); ++t)
        {
            Body<D, R> &tb = *contact->targets[t];
            const R *tpos = tb.v("Position", D), *tvol = tb.v("VolumetricMeasure");
            const TabKernel<D, R> kc(b.h, tb.h);
#pragma omp parallel for schedule(static)
            for (long i = 0; i < n; ++i)
            {
                R pd = R(0);
                const auto &cn = contact->configs[t][i];
                for (size_t m = 0; m != cn.current_size; ++m)
                {
                    R d[D];
                    const R r = dispNorm<D, R>(pos + i * D, tpos + cn.j[m] * D, d);
                    pd -= kc.dW(r) * tvol[cn.j[m]] * r;
                }
                pdiv[i] += pd;
            }
        }
    std::vector<R> out(static_cast<size_t>(n));
#pragma omp parallel for schedule(static)
    for (long i = 0; i < n; ++i)
    {
        R v = R(1);
        if (pdiv[i] > threshold)
        {
            bool very_near = false;
            const auto &nb = inner.configs[0][i];
            for (size_t m = 0; m != nb.current_size && !very_near; ++m)
            {
                R d[D];
                const R r = dispNorm<D, R>(pos + i * D, pos + nb.j[m] * D, d);
                very_near = pdiv[nb.j[m]] < threshold && r < h;
            }
            if (!very_near)
                v = R(0);
        }
        out[size_t(i)] = v;
    }
    for (long i = 0; i < n; ++i)
        ind[i] = prev[i] = out[size_t(i)];
}

// ---------------------------------------------------------------------------------------------------------------
// CK acoustic sub-step with linear kernel correction, as the SYCL dam-break mains run it:
//   AdvectionStepSetup / UpdateParticlePosition   shared_ck/particle_dynamics/fluid_dynamics/fluid_time_step_ck.h:145-148,173-175
//   AcousticTimeStepCK / AdvectionTimeStepCK      fluid_time_step_ck.hpp:31-35,52-56 ; fluid_time_step_ck.cpp:23-25 ; .h:106-108
//   AcousticStep1stHalf<Inner<OneLevel, AcousticRiemannSolverCK, LinearCorrectionCK>> + <Contact<Wall, ...>>
//                                                 acoustic_step_1st_half.hpp:66-74,89-111,122-127,157-180
//   AcousticStep2ndHalf<Inner<...>> + <Contact<Wall, ...>>   acoustic_step_2nd_half.hpp:33-38,53-73,83-89,117-137
// The state is (Compression, CompressionRate, Displacement) instead of the classic (Density, DensityChangeRate, Position):
// Position is not touched inside the sub-steps, so W, dW, e are those of the configuration.
// ---------------------------------------------------------------------------------------------------------------
template <int D, class R>
void ckAdvectionStepSetup(Body<D, R> &b)
{
    R *Vol = b.v("VolumetricMeasure"), *mass = b.v("Mass"), *rho = b.v("Density"), *dpos = b.v("Displacement", D);
    for (size_t i = 0; i < b.n; ++i)
    {
        Vol[i] = mass[i] / rho[i];
        for (int a = 0; a < D; ++a)
            dpos[i * D + a] = R(0);
    }
}
template <int D, class R>
void ckUpdateParticlePosition(Body<D, R> &b)
{
    R *pos = b.v("Position", D), *dpos = b.v("Displacement", D);
    for (size_t i = 0; i < b.n * D; ++i)
        pos[i] += dpos[i];
}
template <int D, class R>
R ckAcousticDt(Body<D, R> &b, R cfl)
{
    const R *mass = b.v("Mass"), *vel = b.v("Velocity", D), *F = b.v("Force", D), *Fp = b.v("ForcePrior", D);
    R red = R(0);
    for (size_t i = 0; i < b.n; ++i)
    {
        R f[D];
        for (int a = 0; a < D; ++a)
            f[a] = F[i * D + a] + Fp[i * D + a];
        const R acc_scale = std::sqrt(R(4.0) * b.h * std::sqrt(sqnorm<D, R>(f)) / mass[i]);
        red = std::max(red, std::max(R(b.mat.c0) + std::sqrt(sqnorm<D, R>(vel + i * D)), acc_scale));
    }
    return cfl * b.h / (red + Tiny<R>::value);
}
template <int D, class R>
R ckAdvectionDt(Body<D, R> &b, R U_ref, R cfl)
{
    const R *vel = b.v("Velocity", D);
    R red = R(0);
    for (size_t i = 0; i < b.n; ++i)
        red = std::max(red, sqnorm<D, R>(vel + i * D));
    return cfl * b.h / (std::max(std::sqrt(red), U_ref) + Tiny<R>::value);
}
template <int D, class R>
inline void matVec(const R *B, const R *e, R *out)
{
    for (int a = 0; a < D; ++a)
    {
        out[a] = R(0);
        for (int c = 0; c < D; ++c)
            out[a] += B[a * D + c] * e[c];
    }
}

template <int D, class R>
void ckAcoustic1stHalf(Relation<D, R> &inner, Relation<D, R> *wall, R dt)
{
    Body<D, R> &b = *inner.body;
    const long n = long(b.n);
    const R rho0 = R(b.mat.rho0), c0 = R(b.mat.c0);
    Riemann<R> rs(rho0, c0, true);
    if (b.vars.find("Compression") == b.vars.end())
    { // registerStateVariable<Real>("Compression", Real(1)), acoustic_step_1st_half.hpp:20
        R *c = b.v("Compression");
        for (long i = 0; i < n; ++i)
            c[i] = R(1);
    }
    const R *pos = b.v("Position", D), *Vol = b.v("VolumetricMeasure"), *mass = b.v("Mass"), *Bm = b.v("LinearCorrectionMatrix", D * D),
            *Fp = b.v("ForcePrior", D);
    R *rho = b.v("Density"), *p = b.v("Pressure"), *comp = b.v("Compression"), *crate = b.v("CompressionRate"), *vel = b.v("Velocity", D),
      *dpos = b.v("Displacement", D), *F = b.v("Force", D);
    for (long i = 0; i < n; ++i) // InitializeKernel
    {
        comp[i] += R(0.5) * dt * crate[i];
        rho[i] = comp[i] * rho0;
        p[i] = rho0 * c0 * c0 * (rho[i] / rho0 - R(1.0)); // weakly_compressible_fluid.hpp getPressure
        for (int a = 0; a < D; ++a)
            dpos[i * D + a] += vel[i * D + a] * dt * R(0.5);
    }
    const TabKernel<D, R> k(b.h, b.h);
#pragma omp parallel for schedule(static)
    for (long i = 0; i < n; ++i) // InteractKernel, inner
    {
        R fs[D] = {}, diss = R(0);
        const auto &nb = inner.configs[0][i];
        for (size_t m = 0; m != nb.current_size; ++m)
        {
            const size_t j = nb.j[m];
            R d[D], e[D], M[D * D], Me[D];
            const R r = dispNorm<D, R>(pos + i * D, pos + j * D, d);
            const R dWV = k.dW(r) * Vol[j];
            for (int a = 0; a < D; ++a)
                e[a] = d[a] / r;
            // AverageP(B_j p_i, B_i p_j) * 2 = B_j p_i + B_i p_j   (riemann_solver_ck: average = half the sum)
            for (int q = 0; q < D * D; ++q)
                M[q] = R(0.5) * (Bm[j * D * D + q] * p[i] + Bm[i * D * D + q] * p[j]);
            matVec<D, R>(M, e, Me);
            for (int a = 0; a < D; ++a)
                fs[a] -= Me[a] * R(2.0) * dWV;
            diss += rs.UJump(p[i] - p[j]) * dWV;
        }
        for (int a = 0; a < D; ++a)
            F[i * D + a] += fs[a] * Vol[i];
        crate[i] = diss * comp[i];
    }
    if (wall)
        for (size_t t = 0; t < wall->targets.size(); ++t)
        {
            Body<D, R> &tb = *wall->targets[t];
            const R *tpos = tb.v("Position", D), *tvol = tb.v("VolumetricMeasure"), *tacc = tb.v("Acceleration", D);
            const TabKernel<D, R> kc(b.h, tb.h);
#pragma omp parallel for schedule(static)
            for (long i = 0; i < n; ++i)
            {
                R fs[D] = {}, diss = R(0);
                const auto &cn = wall->configs[t][i];
                for (size_t m = 0; m != cn.current_size; ++m)
                {
                    const size_t j = cn.j[m];
                    R d[D], e[D], Be[D];
                    const R r = dispNorm<D, R>(pos + i * D, tpos + j * D, d);
                    const R dWV = kc.dW(r) * tvol[j];
                    R face = R(0);
                    for (int a = 0; a < D; ++a)
                    {
                        e[a] = d[a] / r;
                        face += (Fp[i * D + a] / mass[i] - tacc[j * D + a]) * (-e[a]);
                    }
                    const R p_wall = p[i] + rho[i] * r * std::max(R(0), face);
                    matVec<D, R>(Bm + i * D * D, e, Be);
                    for (int a = 0; a < D; ++a)
                        fs[a] -= (p[i] + p_wall) * Be[a] * dWV;
                    diss += rs.UJump(p[i] - p_wall) * dWV;
                }
                for (int a = 0; a < D; ++a)
                    F[i * D + a] += fs[a] * Vol[i];
                crate[i] += diss * comp[i];
            }
        }
    for (long i = 0; i < n; ++i) // UpdateKernel
        for (int a = 0; a < D; ++a)
            vel[i * D + a] += (Fp[i * D + a] + F[i * D + a]) / mass[i] * dt;
}

template <int D, class R>
void ckAcoustic2ndHalf(Relation<D, R> &inner, Relation<D, R> *wall, R dt)
{
    Body<D, R> &b = *inner.body;
    const long n = long(b.n);
    const R rho0 = R(b.mat.rho0), c0 = R(b.mat.c0);
    Riemann<R> rs(rho0, c0, true);
    const R *pos = b.v("Position", D), *Vol = b.v("VolumetricMeasure"), *Bm = b.v("LinearCorrectionMatrix", D * D), *vel = b.v("Velocity", D);
    R *rho = b.v("Density"), *comp = b.v("Compression"), *crate = b.v("CompressionRate"), *dpos = b.v("Displacement", D), *F = b.v("Force", D);
    for (long i = 0; i < n * D; ++i) // InitializeKernel
        dpos[i] += vel[i] * dt * R(0.5);
    const TabKernel<D, R> k(b.h, b.h);
#pragma omp parallel for schedule(static)
    for (long i = 0; i < n; ++i)
    {
        R div = R(0), pd[D] = {}, Be[D];
        const auto &nb = inner.configs[0][i];
        for (size_t m = 0; m != nb.current_size; ++m)
        {
            const size_t j = nb.j[m];
            R d[D], e[D];
            const R r = dispNorm<D, R>(pos + i * D, pos + j * D, d);
            const R dWV = k.dW(r) * Vol[j];
            for (int a = 0; a < D; ++a)
                e[a] = d[a] / r;
            matVec<D, R>(Bm + i * D * D, e, Be);
            R dvBe = R(0), u = R(0);
            for (int a = 0; a < D; ++a)
            {
                const R vave = R(0.5) * (vel[i * D + a] + vel[j * D + a]); // AverageV
                dvBe += R(2.0) * (vel[i * D + a] - vave) * Be[a];
                u += (vel[i * D + a] - vel[j * D + a]) * e[a];
            }
            div += dvBe * dWV;
            const R pj = rs.PJump(u) * dWV;
            for (int a = 0; a < D; ++a)
                pd[a] += pj * e[a];
        }
        crate[i] += div * comp[i];
        for (int a = 0; a < D; ++a)
            F[i * D + a] = pd[a] * Vol[i];
    }
    if (wall)
        for (size_t t = 0; t < wall->targets.size(); ++t)
        {
            Body<D, R> &tb = *wall->targets[t];
            const R *tpos = tb.v("Position", D), *tvol = tb.v("VolumetricMeasure"), *tvel = tb.v("Velocity", D), *tn = tb.v("NormalDirection", D);
            const TabKernel<D, R> kc(b.h, tb.h);
#pragma omp parallel for schedule(static)
            for (long i = 0; i < n; ++i)
            {
                R div = R(0), pd[D] = {}, Be[D];
                const auto &cn = wall->configs[t][i];
                for (size_t m = 0; m != cn.current_size; ++m)
                {
                    const size_t j = cn.j[m];
                    R d[D], e[D], vd[D], nf[D];
                    const R r = dispNorm<D, R>(pos + i * D, tpos + j * D, d);
                    const R dWV = kc.dW(r) * tvol[j];
                    R en = R(0);
                    for (int a = 0; a < D; ++a)
                    {
                        e[a] = d[a] / r;
                        vd[a] = R(2.0) * (vel[i * D + a] - tvel[j * D + a]);
                        en += e[a] * tn[j * D + a];
                    }
                    matVec<D, R>(Bm + i * D * D, e, Be);
                    const R sgn = en > R(0) ? R(1) : (en < R(0) ? R(-1) : R(0)); // SGN: common/scalar_functions.h:127-131
                    R dvBe = R(0), u = R(0);
                    for (int a = 0; a < D; ++a)
                    {
                        nf[a] = sgn * tn[j * D + a];
                        dvBe += vd[a] * Be[a];
                        u += vd[a] * nf[a];
                    }
                    div += dvBe * dWV;
                    const R pj = rs.PJump(u) * dWV;
                    for (int a = 0; a < D; ++a)
                        pd[a] += pj * nf[a];
                }
                crate[i] += div * comp[i];
                for (int a = 0; a < D; ++a)
                    F[i * D + a] += pd[a] * Vol[i];
            }
        }
    for (long i = 0; i < n; ++i) // UpdateKernel
    {
        comp[i] += R(0.5) * dt * crate[i];
        rho[i] = comp[i] * rho0;
    }
}

// ---------------------------------------------------------------------------------------------------------------
// Force of the fluid on a structure: the source body of the contact relation is the SOLID, its targets
// are fluid bodies.   shared_ck/particle_dynamics/fluid_structure_interaction/force_on_structure.hpp:61-79 (viscous),
// :113-137 (pressure); WithUpdate = ForcePriorCK::UpdateKernel (general_dynamics/force_prior_ck.h:55-59).
// corrected: contact_correction_(j) is the fluid particle's LinearCorrectionMatrix (LinearCorrectionCK) or 1 (NoKernelCorrectionCK).
// ---------------------------------------------------------------------------------------------------------------
template <int D, class R>
inline void forcePriorUpdate(Body<D, R> &b, const char *force_name)
{
    const std::string prev_name = std::string("Previous") + force_name;
    R *fp = b.v("ForcePrior", D), *cur = b.v(force_name, D), *prev = b.v(prev_name.c_str(), D);
    for (size_t q = 0; q < b.n * D; ++q)
    {
        fp[q] += cur[q] - prev[q];
        prev[q] = cur[q];
    }
}

template <int D, class R>
void ckPressureForceFromFluid(Relation<D, R> &contact, bool corrected)
{
    Body<D, R> &b = *contact.body;
    const long n = long(b.n);
    const R *pos = b.v("Position", D), *Vol = b.v("VolumetricMeasure"), *vel_ave = b.v("Velocity", D), *acc_ave = b.v("Acceleration", D),
            *nrm = b.v("NormalDirection", D);
    R *force = b.v("PressureForceFromFluid", D);
    for (long q = 0; q < n * D; ++q)
        force[q] = R(0);
    for (size_t t = 0; t < contact.targets.size(); ++t)
    {
        Body<D, R> &f = *contact.targets[t];
        const R *fpos = f.v("Position", D), *fvol = f.v("VolumetricMeasure"), *fvel = f.v("Velocity", D), *frho = f.v("Density"),
                *fmass = f.v("Mass"), *fp = f.v("Pressure"), *ffp = f.v("ForcePrior", D);
        const R *fB = corrected ? f.v("LinearCorrectionMatrix", D * D) : nullptr;
        Riemann<R> rs(R(f.mat.rho0), R(f.mat.c0), true);
        const TabKernel<D, R> k(b.h, f.h);
#pragma omp parallel for schedule(static)
        for (long i = 0; i < n; ++i)
        {
            R acc[D] = {};
            const auto &cn = contact.configs[t][i];
            for (size_t m = 0; m != cn.current_size; ++m)
            {
                const size_t j = cn.j[m];
                R d[D], e[D], ce[D], nf[D];
                const R r = dispNorm<D, R>(pos + i * D, fpos + j * D, d);
                for (int a = 0; a < D; ++a)
                    e[a] = d[a] / r;
                if (corrected)
                    matVec<D, R>(fB + j * D * D, e, ce);
                else
                    for (int a = 0; a < D; ++a)
                        ce[a] = e[a];
                R face = R(0), cen = R(0);
                for (int a = 0; a < D; ++a)
                {
                    face += (ffp[j * D + a] / fmass[j] - acc_ave[i * D + a]) * e[a];
                    cen += ce[a] * nrm[i * D + a];
                }
                const R p_in_wall = fp[j] + frho[j] * r * std::max(R(0), face);
                const R sgn = cen < R(0) ? R(-1) : (cen > R(0) ? R(1) : R(0));
                R u = R(0);
                for (int a = 0; a < D; ++a)
                {
                    nf[a] = -sgn * nrm[i * D + a];
                    u += R(2.0) * (fvel[j * D + a] - vel_ave[i * D + a]) * nf[a];
                }
                const R pj = rs.PJump(u), dWV = k.dW(r) * fvol[j];
                for (int a = 0; a < D; ++a)
                    acc[a] -= (pj * nf[a] + (p_in_wall + fp[j]) * ce[a]) * dWV;
            }
            for (int a = 0; a < D; ++a)
                force[i * D + a] += acc[a] * Vol[i];
        }
    }
    forcePriorUpdate<D, R>(b, "PressureForceFromFluid");
}

template <int D, class R>
void ckViscousForceFromFluid(Relation<D, R> &contact, bool corrected)
{
    Body<D, R> &b = *contact.body;
    const long n = long(b.n);
    const R *pos = b.v("Position", D), *Vol = b.v("VolumetricMeasure"), *vel_ave = b.v("Velocity", D);
    R *force = b.v("ViscousForceFromFluid", D);
    for (long q = 0; q < n * D; ++q)
        force[q] = R(0);
    for (size_t t = 0; t < contact.targets.size(); ++t)
    {
        Body<D, R> &f = *contact.targets[t];
        const R *fpos = f.v("Position", D), *fvol = f.v("VolumetricMeasure"), *fvel = f.v("Velocity", D);
        const R *fB = corrected ? f.v("LinearCorrectionMatrix", D * D) : nullptr;
        const R mu = R(f.mat.mu), h_sq = f.h * f.h;
        const TabKernel<D, R> k(b.h, f.h);
#pragma omp parallel for schedule(static)
        for (long i = 0; i < n; ++i)
        {
            R acc[D] = {};
            const auto &cn = contact.configs[t][i];
            for (size_t m = 0; m != cn.current_size; ++m)
            {
                const size_t j = cn.j[m];
                R d[D], e[D], ce[D];
                const R r = dispNorm<D, R>(pos + i * D, fpos + j * D, d);
                for (int a = 0; a < D; ++a)
                    e[a] = d[a] / r;
                if (corrected)
                    matVec<D, R>(fB + j * D * D, e, ce);
                else
                    for (int a = 0; a < D; ++a)
                        ce[a] = e[a];
                const R denom = sqnorm<D, R>(d) + R(0.01) * h_sq;
                R rce = R(0);
                for (int a = 0; a < D; ++a)
                    rce += d[a] * ce[a];
                const R w = rce * mu * k.dW(r) * fvol[j];
                for (int a = 0; a < D; ++a)
                    acc[a] += w * ((vel_ave[i * D + a] - fvel[j * D + a]) / denom);
            }
            for (int a = 0; a < D; ++a)
                force[i * D + a] += acc[a] * Vol[i];
        }
    }
    forcePriorUpdate<D, R>(b, "ViscousForceFromFluid");
}
} // namespace orc
