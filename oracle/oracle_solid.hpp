// ORACLE — TEST INFRASTRUCTURE ONLY (see oracle_core.hpp header).
// Total-Lagrangian elastic solid of the classic path (test_3d_twisting_column):
//   particle_dynamics/solid_dynamics/elastic_dynamics.h:171-207,259-287,325-351
//   particle_dynamics/solid_dynamics/elastic_dynamics.cpp:12-30,62-118,162-194
//   particle_dynamics/general_dynamics/kernel_correction.cpp:8-29
//   materials/elastic_solid.cpp:13-23,54-80,109-132 ; materials/elastic_solid.h:90-96
// Matrices are stored row-major, D*D Reals per particle.
#pragma once
#include "oracle_fluid.hpp"

namespace orc
{
template <int D, class R>
struct Mat
{
    R a[D][D];
    static Mat zero()
    {
        Mat m;
        for (int i = 0; i < D; ++i)
            for (int j = 0; j < D; ++j)
                m.a[i][j] = 0;
        return m;
    }
    static Mat identity()
    {
        Mat m = zero();
        for (int i = 0; i < D; ++i)
            m.a[i][i] = 1;
        return m;
    }
    static Mat load(const R *p)
    {
        Mat m;
        for (int i = 0; i < D; ++i)
            for (int j = 0; j < D; ++j)
                m.a[i][j] = p[i * D + j];
        return m;
    }
    void store(R *p) const
    {
        for (int i = 0; i < D; ++i)
            for (int j = 0; j < D; ++j)
                p[i * D + j] = a[i][j];
    }
    Mat T() const
    {
        Mat m;
        for (int i = 0; i < D; ++i)
            for (int j = 0; j < D; ++j)
                m.a[i][j] = a[j][i];
        return m;
    }
    Mat operator*(const Mat &o) const
    {
        Mat m = zero();
        for (int i = 0; i < D; ++i)
            for (int j = 0; j < D; ++j)
                for (int k = 0; k < D; ++k)
                    m.a[i][j] += a[i][k] * o.a[k][j];
        return m;
    }
    Mat operator*(R s) const
    {
        Mat m;
        for (int i = 0; i < D; ++i)
            for (int j = 0; j < D; ++j)
                m.a[i][j] = a[i][j] * s;
        return m;
    }
    Mat operator+(const Mat &o) const
    {
        Mat m;
        for (int i = 0; i < D; ++i)
            for (int j = 0; j < D; ++j)
                m.a[i][j] = a[i][j] + o.a[i][j];
        return m;
    }
    Mat operator-(const Mat &o) const
    {
        Mat m;
        for (int i = 0; i < D; ++i)
            for (int j = 0; j < D; ++j)
                m.a[i][j] = a[i][j] - o.a[i][j];
        return m;
    }
    R trace() const
    {
        R t = 0;
        for (int i = 0; i < D; ++i)
            t += a[i][i];
        return t;
    }
    R det() const
    {
        if constexpr (D == 2)
            return a[0][0] * a[1][1] - a[0][1] * a[1][0];
        else
            return a[0][0] * (a[1][1] * a[2][2] - a[1][2] * a[2][1]) -
                   a[0][1] * (a[1][0] * a[2][2] - a[1][2] * a[2][0]) +
                   a[0][2] * (a[1][0] * a[2][1] - a[1][1] * a[2][0]);
    }
    Mat inverse() const // closed-form cofactor inverse
    {
        Mat m;
        R inv = R(1.0) / det();
        if constexpr (D == 2)
        {
            m.a[0][0] = a[1][1] * inv;
            m.a[0][1] = -a[0][1] * inv;
            m.a[1][0] = -a[1][0] * inv;
            m.a[1][1] = a[0][0] * inv;
        }
        else
        {
            for (int i = 0; i < 3; ++i)
                for (int j = 0; j < 3; ++j)
                {
                    int i1 = (i + 1) % 3, i2 = (i + 2) % 3, j1 = (j + 1) % 3, j2 = (j + 2) % 3;
                    m.a[i][j] = (a[j1][i1] * a[j2][i2] - a[j1][i2] * a[j2][i1]) * inv;
                }
        }
        return m;
    }
    void mulVec(const R *v, R *out) const
    {
        for (int i = 0; i < D; ++i)
        {
            out[i] = 0;
            for (int j = 0; j < D; ++j)
                out[i] += a[i][j] * v[j];
        }
    }
};

// materials/elastic_solid.cpp:54-80 (LinearElasticSolid ctor; NeoHookeanSolid inherits it), :13-18
template <int D, class R>
void setElastic(Body<D, R> &b, double rho0, double E, double nu)
{
    Material &m = b.mat;
    m.rho0 = rho0;
    m.E = E;
    m.nu = nu;
    m.G0 = 0.5 * E / (1.0 + nu);
    m.K0 = E / 3.0 / (1.0 - 2.0 * nu);
    m.lambda0 = nu * E / (1.0 + nu) / (1.0 - 2.0 * nu);
    m.c0 = std::sqrt(m.K0 / rho0);
    m.cs0 = std::sqrt(m.G0 / rho0);
}

// general_dynamics/kernel_correction.cpp:8-29 (alpha = 0), common/vector_functions.h:198-203
template <int D, class R>
void linearGradientCorrectionMatrix(Relation<D, R> &inner)
{
    Body<D, R> &b = *inner.body;
    R *Vol = b.v("VolumetricMeasure"), *B = b.v("LinearGradientCorrectionMatrix", D * D);
    for (size_t i = 0; i < b.n; ++i)
    {
        Mat<D, R> local = Mat<D, R>::zero();
        const auto &nb = inner.configs[0][i];
        for (size_t k = 0; k != nb.current_size; ++k)
        {
            const R *e = &nb.e[k * D];
            R c = nb.dW[k] * Vol[nb.j[k]];
            for (int r = 0; r < D; ++r)
                for (int cidx = 0; cidx < D; ++cidx)
                    local.a[r][cidx] -= (nb.r[k] * e[r]) * (c * e[cidx]);
        }
        local.store(B + i * D * D);
    }
    const R sqrt_eps = R(1.0e-8);
    for (size_t i = 0; i < b.n; ++i)
    {
        Mat<D, R> Bm = Mat<D, R>::load(B + i * D * D);
        R det = Bm.det();
        R det_sqr = std::max(R(0) - det, R(0));
        Mat<D, R> Bt = Bm.T();
        Mat<D, R> inv = (Bt * Bm + Mat<D, R>::identity() * sqrt_eps).inverse() * Bt;
        R w1 = det / (det + det_sqr), w2 = det_sqr / (det + det_sqr);
        (inv * w1 + Mat<D, R>::identity() * w2).store(B + i * D * D);
    }
}

// elastic_dynamics.cpp:169-184 (Decomposed init) / :109-118 (PK2 init); elastic_dynamics.h:266-281 /
// :179-201 (interactions); elastic_dynamics.cpp:96-99 (update)
template <int D, class R>
void solidIntegration1stHalf(Relation<D, R> &inner, bool decomposed, R dt)
{
    Body<D, R> &b = *inner.body;
    const Material &m = b.mat;
    R *Vol = b.v("VolumetricMeasure"), *pos = b.v("Position", D), *vel = b.v("Velocity", D), *force = b.v("Force", D),
      *B = b.v("LinearGradientCorrectionMatrix", D * D), *F = b.v("DeformationGradient", D * D),
      *dF = b.v("DeformationRate", D * D), *rho = b.v("Density"), *mass = b.v("Mass"), *fprior = b.v("ForcePrior", D);
    R *Jterm = b.v("DeterminantTerm"), *stress = b.v("StressOnParticle", D * D), *invFT = b.v("InverseTransposedDeformation", D * D);
    R *pk1 = b.v("StressPK1OnParticle", D * D);
    const R rho0 = R(m.rho0), inv_rho0 = R(1.0) / rho0, G0 = R(m.G0), K0 = R(m.K0), lambda0 = R(m.lambda0),
            c0 = R(m.c0), cs0 = R(m.cs0), h = b.h, one_over_dim = R(1.0) / R(D), corr = R(1.07);
    long n = long(b.n);
#pragma omp parallel for schedule(static)
    for (long i = 0; i < n; ++i)
    {
        for (int d = 0; d < D; ++d)
            pos[i * D + d] += vel[i * D + d] * dt * R(0.5);
        for (int q = 0; q < D * D; ++q)
            F[i * D * D + q] += dF[i * D * D + q] * dt * R(0.5);
        Mat<D, R> Fm = Mat<D, R>::load(F + i * D * D);
        if (decomposed)
        {
            R J = Fm.det();
            R one_over_J = R(1.0) / J;
            rho[i] = rho0 * one_over_J;
            Jterm[i] = std::pow(one_over_J * one_over_J, one_over_dim);
            Mat<D, R> iFT = Fm.inverse().T();
            iFT.store(invFT + i * D * D);
            R vol_kirchhoff = R(0.5) * K0 * (J * J - R(1)); // NeoHookeanSolid::VolumetricKirchhoff
            R scalar = vol_kirchhoff - corr * G0 * Jterm[i] * (Fm * Fm.T()).trace() * one_over_dim;
            // NumericalDampingLeftCauchy (elastic_solid.h:90-96)
            Mat<D, R> dFm = Mat<D, R>::load(dF + i * D * D);
            Mat<D, R> strain_rate = (dFm * Fm.T() + Fm * dFm.T()) * R(0.5);
            Mat<D, R> normal_rate = Mat<D, R>::zero();
            for (int d = 0; d < D; ++d)
                normal_rate.a[d][d] = strain_rate.a[d][d];
            Mat<D, R> damping = ((strain_rate - normal_rate) * cs0 + normal_rate * c0) * (R(0.5) * rho0) * h;
            (iFT * scalar + damping * iFT).store(stress + i * D * D);
        }
        else
        {
            rho[i] = rho0 / Fm.det();
            // NeoHookeanSolid::StressPK2 (elastic_solid.cpp:109-117); PK1 = F * PK2; times B^T
            Mat<D, R> rc = Fm.T() * Fm;
            R J = Fm.det();
            Mat<D, R> pk2 = Mat<D, R>::identity() * G0 + rc.inverse() * ((lambda0 * (J - R(1.0)) - G0) * J);
            ((Fm * pk2) * Mat<D, R>::load(B + i * D * D).T()).store(pk1 + i * D * D);
        }
    }
    const R inv_W0 = R(1.0) / b.kernel.W0();
#pragma omp parallel for schedule(static)
    for (long i = 0; i < n; ++i)
    {
        R f[D] = {};
        const auto &nb = inner.configs[0][i];
        for (size_t k = 0; k != nb.current_size; ++k)
        {
            size_t j = nb.j[k];
            const R *e = &nb.e[k * D];
            R c = nb.dW[k] * Vol[j];
            if (decomposed)
            {
                Mat<D, R> s = Mat<D, R>::load(stress + i * D * D) + Mat<D, R>::load(stress + j * D * D);
                R se[D];
                s.mulVec(e, se);
                R shear_c = corr * G0 * (Jterm[i] + Jterm[j]);
                for (int d = 0; d < D; ++d)
                {
                    R shear = shear_c * (pos[i * D + d] - pos[j * D + d]) / nb.r[k];
                    f[d] += mass[i] * (se[d] + shear) * c * inv_rho0;
                }
            }
            else
            {
                R dim_r = R(D) / nb.r[k];
                R pv = 0;
                for (int d = 0; d < D; ++d)
                    pv += (pos[i * D + d] - pos[j * D + d]) * (vel[i * D + d] - vel[j * D + d]);
                R strain_rate = dim_r * dim_r * pv;
                R weight = nb.W[k] * inv_W0;
                R pair_damp = R(0.5) * rho0 * c0 * strain_rate * h; // PairNumericalDamping
                Mat<D, R> num = (Mat<D, R>::load(F + i * D * D) + Mat<D, R>::load(F + j * D * D)) * (R(0.5) * pair_damp);
                Mat<D, R> s = Mat<D, R>::load(pk1 + i * D * D) + Mat<D, R>::load(pk1 + j * D * D) + num * (R(0.25) * weight);
                R se[D];
                s.mulVec(e, se);
                for (int d = 0; d < D; ++d)
                    f[d] += mass[i] * inv_rho0 * c * se[d];
            }
        }
        for (int d = 0; d < D; ++d)
            force[i * D + d] = f[d];
    }
#pragma omp parallel for schedule(static)
    for (long i = 0; i < n; ++i)
        for (int d = 0; d < D; ++d)
            vel[i * D + d] += (fprior[i * D + d] + force[i * D + d]) / mass[i] * dt;
}

// elastic_dynamics.cpp:186-194 ; elastic_dynamics.h:333-348
template <int D, class R>
void solidIntegration2ndHalf(Relation<D, R> &inner, R dt)
{
    Body<D, R> &b = *inner.body;
    R *Vol = b.v("VolumetricMeasure"), *pos = b.v("Position", D), *vel = b.v("Velocity", D),
      *B = b.v("LinearGradientCorrectionMatrix", D * D), *F = b.v("DeformationGradient", D * D),
      *dF = b.v("DeformationRate", D * D);
    long n = long(b.n);
#pragma omp parallel for schedule(static)
    for (long i = 0; i < n; ++i)
        for (int d = 0; d < D; ++d)
            pos[i * D + d] += vel[i * D + d] * dt * R(0.5);
#pragma omp parallel for schedule(static)
    for (long i = 0; i < n; ++i)
    {
        Mat<D, R> rate = Mat<D, R>::zero();
        const auto &nb = inner.configs[0][i];
        for (size_t k = 0; k != nb.current_size; ++k)
        {
            size_t j = nb.j[k];
            R c = nb.dW[k] * Vol[j];
            for (int r = 0; r < D; ++r)
                for (int q = 0; q < D; ++q)
                    rate.a[r][q] -= (vel[i * D + r] - vel[j * D + r]) * (c * nb.e[k * D + q]);
        }
        (rate * Mat<D, R>::load(B + i * D * D)).store(dF + i * D * D);
    }
#pragma omp parallel for schedule(static)
    for (long i = 0; i < n; ++i)
        for (int q = 0; q < D * D; ++q)
            F[i * D * D + q] += dF[i * D * D + q] * dt * R(0.5);
}

// general_dynamics/general_constraint.h:106-121 FixConstraint over a BodyPartByParticle id list
template <int D, class R>
void fixBodyPart(Body<D, R> &b, const long *ids, long n)
{
    R *pos = b.v("Position", D), *pos0 = b.v("InitialPosition", D), *vel = b.v("Velocity", D);
    for (long k = 0; k < n; ++k)
        for (int d = 0; d < D; ++d)
        {
            pos[ids[k] * D + d] = pos0[ids[k] * D + d];
            vel[ids[k] * D + d] = 0;
        }
}

// elastic_dynamics.cpp:23-30 (ReduceMin, reference MaxReal)
template <int D, class R>
R solidAcousticTimeStep(Body<D, R> &b, R cfl)
{
    R *vel = b.v("Velocity", D), *force = b.v("Force", D), *fprior = b.v("ForcePrior", D), *mass = b.v("Mass");
    R c0 = R(b.mat.c0), h = b.h;
    R m = std::numeric_limits<R>::max();
    for (size_t i = 0; i < b.n; ++i)
    {
        R a[D];
        for (int d = 0; d < D; ++d)
            a[d] = (force[i * D + d] + fprior[i * D + d]) / mass[i];
        R an = vnorm<D, R>(a);
        R v = cfl * std::min(R(std::sqrt(h / (an + Tiny<R>::value))), h / (c0 + vnorm<D, R>(vel + i * D)));
        m = std::min(m, v);
    }
    return m;
}

} // namespace orc
