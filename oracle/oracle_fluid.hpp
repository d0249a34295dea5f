// ORACLE — TEST INFRASTRUCTURE ONLY (see oracle_core.hpp header).
// Weakly-compressible fluid local dynamics of the classic path, restated per particle exactly in the
// reference's initialization / interaction / update split (particle_dynamics/dynamics_algorithms.h:
// 283-290 InteractionWithUpdate, :336-352 Dynamics1Level).
#pragma once
#include "oracle_core.hpp"
#include <limits>

namespace orc
{
// materials/riemann_solver.h:55-123 — single-fluid constants, evaluated with the reference's formulas.
template <class R>
struct Riemann
{
    R inv_rho0c0_ave, rho0c0_geo_ave, inv_c0_ave, slope;
    bool on;
    Riemann(R rho0, R c0, bool enabled) : slope(R(3.0)), on(enabled)
    {
        R a = rho0 * c0, b = rho0 * c0;
        R inv_sum = R(1.0) / (a + b);
        inv_rho0c0_ave = (a + b) / (std::pow(a, R(2)) + std::pow(b, R(2)));
        rho0c0_geo_ave = R(2.0) * a * b * inv_sum;
        inv_c0_ave = R(0.5) * (rho0 + rho0) * inv_rho0c0_ave;
    }
    R PJump(R u) const
    {
        if (!on)
            return R(0);
        R lim = std::min(slope * (inv_c0_ave * std::max(u, R(0))), R(1)); // common_functors.h:82-94
        return rho0c0_geo_ave * u * lim;
    }
    R UJump(R dp) const { return on ? dp * inv_rho0c0_ave : R(0); }
};

// fluid_dynamics/density_summation.cpp:8-16,58-78 ; density_summation.hpp:28-32 ; .cpp:18-22
template <int D, class R>
void densitySummation(Relation<D, R> &inner, Relation<D, R> *contact, bool free_surface)
{
    Body<D, R> &b = *inner.body;
    R *rho = b.v("Density"), *mass = b.v("Mass"), *rho_sum = b.v("DensitySummation"), *Vol = b.v("VolumetricMeasure");
    R rho0 = R(b.mat.rho0), inv_sigma0 = R(1.0) / b.sigma0, W0 = b.kernel.W0();
    long n = long(b.n);
#pragma omp parallel for schedule(static)
    for (long i = 0; i < n; ++i)
    {
        R sigma = W0;
        const auto &nb = inner.configs[0][i];
        for (size_t k = 0; k != nb.current_size; ++k)
            sigma += nb.W[k];
        rho_sum[i] = sigma * rho0 * inv_sigma0;
        if (contact)
        {
            R s = 0;
            for (size_t t = 0; t < contact->targets.size(); ++t)
            {
                R *cm = contact->targets[t]->v("Mass");
                R inv_rho0_k = R(1.0) / R(contact->targets[t]->mat.rho0);
                const auto &cn = contact->configs[t][i];
                for (size_t k = 0; k != cn.current_size; ++k)
                    s += cn.W[k] * inv_rho0_k * cm[cn.j[k]];
            }
            rho_sum[i] += s * rho0 * rho0 * inv_sigma0 / mass[i];
        }
    }
#pragma omp parallel for schedule(static)
    for (long i = 0; i < n; ++i)
    {
        if (free_surface)
            rho[i] = std::max(rho_sum[i], rho0);
        else
        {
            rho[i] = rho_sum[i];
            Vol[i] = mass[i] / rho[i];
        }
    }
}

// fluid_dynamics/fluid_integration.hpp:50-55 (init), :64-80 (inner), :89-113 (wall), :58-61 (update)
template <int D, class R>
void integration1stHalf(Relation<D, R> &inner, Relation<D, R> *wall, bool riemann, R dt)
{
    Body<D, R> &b = *inner.body;
    R *Vol = b.v("VolumetricMeasure"), *rho = b.v("Density"), *mass = b.v("Mass"), *p = b.v("Pressure"),
      *drho = b.v("DensityChangeRate"), *pos = b.v("Position", D), *vel = b.v("Velocity", D),
      *force = b.v("Force", D), *fprior = b.v("ForcePrior", D);
    R rho0 = R(b.mat.rho0), c0 = R(b.mat.c0);
    R p0 = rho0 * c0 * c0; // materials/weakly_compressible_fluid.cpp:8-22
    Riemann<R> rs(rho0, c0, riemann);
    long n = long(b.n);
#pragma omp parallel for schedule(static)
    for (long i = 0; i < n; ++i)
    {
        rho[i] += drho[i] * dt * R(0.5);
        p[i] = p0 * (rho[i] / rho0 - R(1.0));
        for (int d = 0; d < D; ++d)
            pos[i * D + d] += vel[i * D + d] * dt * R(0.5);
    }
#pragma omp parallel for schedule(static)
    for (long i = 0; i < n; ++i)
    {
        R f[D] = {};
        R diss = 0;
        const auto &nb = inner.configs[0][i];
        for (size_t k = 0; k != nb.current_size; ++k)
        {
            size_t j = nb.j[k];
            R dWV = nb.dW[k] * Vol[j];
            const R *e = &nb.e[k * D];
            R ps = (p[i] * R(1.0) + p[j] * R(1.0)) * dWV;
            for (int d = 0; d < D; ++d)
                f[d] -= ps * e[d];
            diss += rs.UJump(p[i] - p[j]) * dWV;
        }
        for (int d = 0; d < D; ++d)
            force[i * D + d] += f[d] * Vol[i];
        drho[i] = diss * rho[i];
        if (wall)
        {
            R fw[D] = {};
            R dw = 0;
            for (size_t t = 0; t < wall->targets.size(); ++t)
            {
                R *wacc = wall->targets[t]->v("Acceleration", D);
                R *wVol = wall->targets[t]->v("VolumetricMeasure");
                const auto &cn = wall->configs[t][i];
                for (size_t k = 0; k != cn.current_size; ++k)
                {
                    size_t j = cn.j[k];
                    const R *e = &cn.e[k * D];
                    R dWV = cn.dW[k] * wVol[j];
                    R r_ij = cn.r[k];
                    R face = 0;
                    for (int d = 0; d < D; ++d)
                        face += (fprior[i * D + d] / mass[i] - wacc[j * D + d]) * (-e[d]);
                    R p_wall = p[i] + rho[i] * r_ij * std::max(R(0), face);
                    R ps = (p[i] + p_wall) * R(1.0) * dWV;
                    for (int d = 0; d < D; ++d)
                        fw[d] -= ps * e[d];
                    dw += rs.UJump(p[i] - p_wall) * dWV;
                }
            }
            for (int d = 0; d < D; ++d)
                force[i * D + d] += fw[d] * Vol[i];
            drho[i] += dw * rho[i];
        }
    }
#pragma omp parallel for schedule(static)
    for (long i = 0; i < n; ++i)
        for (int d = 0; d < D; ++d)
            vel[i * D + d] += (fprior[i * D + d] + force[i * D + d]) / mass[i] * dt;
}

// fluid_dynamics/fluid_integration.hpp:167-170 (init), :179-196 (inner), :205-231 (wall), :173-176 (update)
template <int D, class R>
void integration2ndHalf(Relation<D, R> &inner, Relation<D, R> *wall, bool riemann, R dt)
{
    Body<D, R> &b = *inner.body;
    R *Vol = b.v("VolumetricMeasure"), *rho = b.v("Density"), *drho = b.v("DensityChangeRate"),
      *pos = b.v("Position", D), *vel = b.v("Velocity", D), *force = b.v("Force", D);
    Riemann<R> rs(R(b.mat.rho0), R(b.mat.c0), riemann);
    long n = long(b.n);
#pragma omp parallel for schedule(static)
    for (long i = 0; i < n; ++i)
        for (int d = 0; d < D; ++d)
            pos[i * D + d] += vel[i * D + d] * dt * R(0.5);
#pragma omp parallel for schedule(static)
    for (long i = 0; i < n; ++i)
    {
        R rate = 0;
        R pd[D] = {};
        const auto &nb = inner.configs[0][i];
        for (size_t k = 0; k != nb.current_size; ++k)
        {
            size_t j = nb.j[k];
            const R *e = &nb.e[k * D];
            R dWV = nb.dW[k] * Vol[j];
            R u = 0;
            for (int d = 0; d < D; ++d)
                u += (vel[i * D + d] - vel[j * D + d]) * e[d];
            rate += u * dWV;
            R pj = rs.PJump(u) * dWV;
            for (int d = 0; d < D; ++d)
                pd[d] += pj * e[d];
        }
        drho[i] += rate * rho[i];
        for (int d = 0; d < D; ++d)
            force[i * D + d] = pd[d] * Vol[i];
        if (wall)
        {
            R rw = 0;
            R pw[D] = {};
            for (size_t t = 0; t < wall->targets.size(); ++t)
            {
                R *wvel = wall->targets[t]->v("Velocity", D);
                R *wn = wall->targets[t]->v("NormalDirection", D);
                R *wVol = wall->targets[t]->v("VolumetricMeasure");
                const auto &cn = wall->configs[t][i];
                for (size_t k = 0; k != cn.current_size; ++k)
                {
                    size_t j = cn.j[k];
                    const R *e = &cn.e[k * D];
                    R dWV = cn.dW[k] * wVol[j];
                    R en = 0;
                    for (int d = 0; d < D; ++d)
                        en += e[d] * wn[j * D + d];
                    R sgn = en < R(0) ? R(-1) : (en > R(0) ? R(1) : R(0)); // SGN: common/scalar_functions.h:127-131
                    R u1 = 0, u2 = 0;
                    for (int d = 0; d < D; ++d)
                    {
                        R nf = sgn * wn[j * D + d];
                        R vw = R(2.0) * wvel[j * D + d] - vel[i * D + d];
                        u1 += (vel[i * D + d] - vw) * e[d];
                        u2 += (vel[i * D + d] - wvel[j * D + d]) * nf;
                    }
                    rw += u1 * dWV;
                    R pj = rs.PJump(R(2.0) * u2) * dWV;
                    for (int d = 0; d < D; ++d)
                        pw[d] += pj * (sgn * wn[j * D + d]);
                }
            }
            drho[i] += rw * rho[i];
            for (int d = 0; d < D; ++d)
                force[i * D + d] += pw[d] * Vol[i];
        }
    }
#pragma omp parallel for schedule(static)
    for (long i = 0; i < n; ++i)
        rho[i] += drho[i] * dt * R(0.5);
}

template <int D, class R>
inline R vnorm(const R *v)
{
    return std::sqrt(sqnorm<D, R>(v));
}

// fluid_dynamics/fluid_time_step.cpp:21-31 ; reduce_functors.h:73-82 (ReduceMax, reference MinReal)
template <int D, class R>
R acousticTimeStep(Body<D, R> &b, R cfl)
{
    R *vel = b.v("Velocity", D);
    R c0 = R(b.mat.c0);
    R m = std::numeric_limits<R>::lowest();
    for (size_t i = 0; i < b.n; ++i)
        m = std::max(m, c0 + vnorm<D, R>(vel + i * D));
    return cfl * b.h / (m + Tiny<R>::value);
}

// fluid_dynamics/fluid_time_step.cpp:54-79
template <int D, class R>
R advectionTimeStep(Body<D, R> &b, R U_ref, R cfl, bool viscous)
{
    R *vel = b.v("Velocity", D), *force = b.v("Force", D), *fprior = b.v("ForcePrior", D), *mass = b.v("Mass");
    R speed_ref = U_ref;
    if (viscous)
        speed_ref = std::max(R(b.mat.mu) / R(b.mat.rho0) / b.h, speed_ref);
    R m = std::numeric_limits<R>::lowest();
    for (size_t i = 0; i < b.n; ++i)
    {
        R s[D];
        for (int d = 0; d < D; ++d)
            s[d] = force[i * D + d] + fprior[i * D + d];
        R acc_scale = R(4.0) * b.h * vnorm<D, R>(s) / mass[i];
        m = std::max(m, std::max(sqnorm<D, R>(vel + i * D), acc_scale));
    }
    R speed_max = std::sqrt(m);
    return cfl * b.h / (std::max(speed_max, speed_ref) + Tiny<R>::value);
}

// general_dynamics/force_prior.hpp:22-41 GravityForce::update (ForcePrior bookkeeping)
template <int D, class R>
void gravityForce(Body<D, R> &b, const R *g)
{
    R *mass = b.v("Mass"), *fp = b.v("ForcePrior", D), *cur = b.v("GravityForce", D), *prev = b.v("PreviousGravityForce", D);
    for (size_t i = 0; i < b.n; ++i)
        for (int d = 0; d < D; ++d)
        {
            cur[i * D + d] = mass[i] * g[d];
            fp[i * D + d] += cur[i * D + d] - prev[i * D + d];
            prev[i * D + d] = cur[i * D + d];
        }
}

// fluid_dynamics/viscous_dynamics.hpp:35-53 (inner), :92-115 (Contact<Wall>, added with +=) + force_prior.hpp:22-26 ;
// particle_functors.h:163-177.  ComplexInteraction order: inner interaction, wall interaction, then update.
template <int D, class R>
void viscousForce(Relation<D, R> &inner, Relation<D, R> *wall = nullptr)
{
    Body<D, R> &b = *inner.body;
    R *Vol = b.v("VolumetricMeasure"), *vel = b.v("Velocity", D), *vf = b.v("ViscousForce", D),
      *prev = b.v("PreviousViscousForce", D), *fp = b.v("ForcePrior", D);
    R mu = R(b.mat.mu);
    R mu_ij = R(2.0) * mu * mu * (R(1.0) / (mu + mu));
    R h = b.h;
    long n = long(b.n);
#pragma omp parallel for schedule(static)
    for (long i = 0; i < n; ++i)
    {
        R f[D] = {};
        const auto &nb = inner.configs[0][i];
        for (size_t k = 0; k != nb.current_size; ++k)
        {
            size_t j = nb.j[k];
            const R *e = &nb.e[k * D];
            R ee = 0;
            for (int d = 0; d < D; ++d)
                ee += e[d] * ((R(1.0) + R(1.0)) * e[d]);
            R denom = nb.r[k] + R(0.01) * h;
            for (int d = 0; d < D; ++d)
            {
                R vd = (vel[i * D + d] - vel[j * D + d]) / denom;
                f[d] += ee * mu_ij * vd * nb.dW[k] * Vol[j];
            }
        }
        for (int d = 0; d < D; ++d)
            vf[i * D + d] = f[d] * Vol[i];
        if (wall)
        {
            R fw[D] = {};
            for (size_t t = 0; t < wall->targets.size(); ++t)
            {
                R *wvel = wall->targets[t]->v("Velocity", D);
                R *wVol = wall->targets[t]->v("VolumetricMeasure");
                const auto &cn = wall->configs[t][i];
                for (size_t k = 0; k != cn.current_size; ++k)
                {
                    size_t j = cn.j[k];
                    const R *e = &cn.e[k * D];
                    R ee = 0;
                    for (int d = 0; d < D; ++d)
                        ee += e[d] * (R(1.0) * e[d]);
                    R denom = cn.r[k] + R(0.01) * h;
                    for (int d = 0; d < D; ++d)
                    {
                        R vd = R(2.0) * (vel[i * D + d] - wvel[j * D + d]) / denom;
                        fw[d] += R(2.0) * ee * mu_ij * vd * cn.dW[k] * wVol[j];
                    }
                }
            }
            for (int d = 0; d < D; ++d)
                vf[i * D + d] += fw[d] * Vol[i];
        }
    }
    for (long i = 0; i < n; ++i)
        for (int d = 0; d < D; ++d)
        {
            fp[i * D + d] += vf[i * D + d] - prev[i * D + d];
            prev[i * D + d] = vf[i * D + d];
        }
}

// fluid_dynamics/transport_velocity_correction.hpp:38-67 (TruncatedLinear slope 100, AllParticles)
template <int D, class R>
void transportVelocityCorrection(Relation<D, R> &inner, R coefficient)
{
    Body<D, R> &b = *inner.body;
    R *Vol = b.v("VolumetricMeasure"), *pos = b.v("Position", D), *kgi = b.v("KernelGradientIntegral", D);
    R h2 = b.h * b.h;
    R scaling = coefficient * h2;
    long n = long(b.n);
#pragma omp parallel for schedule(static)
    for (long i = 0; i < n; ++i)
    {
        R inc[D] = {};
        const auto &nb = inner.configs[0][i];
        for (size_t k = 0; k != nb.current_size; ++k)
        {
            R c = (R(1.0) + R(1.0)) * nb.dW[k] * Vol[nb.j[k]];
            for (int d = 0; d < D; ++d)
                inc[d] -= c * nb.e[k * D + d];
        }
        for (int d = 0; d < D; ++d)
            kgi[i * D + d] = inc[d];
    }
    for (long i = 0; i < n; ++i)
    {
        R sq = sqnorm<D, R>(kgi + i * D);
        R lim = std::min(R(100.0) * (h2 * sq), R(1));
        for (int d = 0; d < D; ++d)
            pos[i * D + d] += scaling * lim * kgi[i * D + d] * R(1.0) * R(1.0);
    }
}

// general_dynamics/domian_bouding/domain_bounding.h:93-137 (bounding: wrap positions).
// The reference only visits particles of the tagged bound cells; a particle outside them cannot be
// out of bounds between two configuration updates, so visiting all particles is equivalent.
template <int D, class R>
void periodicBounding(Body<D, R> &b, int axis)
{
    R *pos = b.v("Position", D);
    R L = b.per_hi[axis] - b.per_lo[axis];
    for (size_t i = 0; i < b.n; ++i)
    {
        if (pos[i * D + axis] < b.per_lo[axis])
            pos[i * D + axis] += L;
    }
    for (size_t i = 0; i < b.n; ++i)
    {
        if (pos[i * D + axis] > b.per_hi[axis])
            pos[i * D + axis] -= L;
    }
}

// domain_bounding.cpp:18-65 + for_2D_build/meshes/cell_linked_list_2d.cpp:8-44 (tagBoundingCells):
// ghost list-data entries (same index, translated position) for particles within rc of a periodic face.
template <int D, class R>
void periodicCellLinkedList(Body<D, R> &b, int axis)
{
    const Mesh<D, R> &mesh = b.mesh;
    int lc[D], uc[D];
    mesh.cellIndex(b.per_lo, lc);
    mesh.cellIndex(b.per_hi, uc);
    R L = b.per_hi[axis] - b.per_lo[axis];
    R rc = b.kernel.rc;
    // enumerate tagged cells: [bound-1, bound+2) along axis, [lower-1, upper+2) along the others
    auto forTagged = [&](int centre, auto &&fn)
    {
        int lo[3] = {0, 0, 0}, hi[3] = {1, 1, 1};
        for (int d = 0; d < D; ++d)
        {
            if (d == axis)
            {
                lo[d] = std::max(centre - 1, 0);
                hi[d] = std::min(centre + 2, mesh.cells[d]);
            }
            else
            {
                lo[d] = std::max(lc[d] - 1, 0);
                hi[d] = std::min(uc[d] + 2, mesh.cells[d]);
            }
        }
        int c[3];
        for (c[0] = lo[0]; c[0] < hi[0]; ++c[0])
            for (c[1] = lo[1]; c[1] < hi[1]; ++c[1])
                for (c[2] = lo[2]; c[2] < hi[2]; ++c[2])
                    fn(mesh.linear(c));
    };
    auto insert = [&](size_t idx, std::array<R, D> p)
    {
        int c[D];
        mesh.cellIndex(p.data(), c);
        b.cell_data_lists[mesh.linear(c)].emplace_back(idx, p);
    };
    forTagged(lc[axis], [&](size_t cell)
              {
                  auto &dl = b.cell_data_lists[cell];
                  for (size_t k = 0; k < dl.size(); ++k)
                  {
                      auto p = dl[k].second;
                      if (p[axis] > b.per_lo[axis] && p[axis] < (b.per_lo[axis] + rc))
                      {
                          p[axis] = p[axis] + L;
                          insert(dl[k].first, p);
                      }
                  } });
    forTagged(uc[axis], [&](size_t cell)
              {
                  auto &dl = b.cell_data_lists[cell];
                  for (size_t k = 0; k < dl.size(); ++k)
                  {
                      auto p = dl[k].second;
                      if (p[axis] < b.per_hi[axis] && p[axis] > (b.per_hi[axis] - rc))
                      {
                          p[axis] = p[axis] - L;
                          insert(dl[k].first, p);
                      }
                  } });
}

// general_dynamics/general_reduce.cpp:54-79 ; external_force/external_force.h:52-55
template <int D, class R>
R totalKineticEnergy(Body<D, R> &b)
{
    R *mass = b.v("Mass"), *vel = b.v("Velocity", D);
    R s = 0;
    for (size_t i = 0; i < b.n; ++i)
        s += R(0.5) * mass[i] * sqnorm<D, R>(vel + i * D);
    return s;
}
template <int D, class R>
R totalMechanicalEnergy(Body<D, R> &b, const R *g)
{
    R *mass = b.v("Mass"), *vel = b.v("Velocity", D), *pos = b.v("Position", D);
    R s = 0;
    for (size_t i = 0; i < b.n; ++i)
    {
        R pot = 0;
        for (int d = 0; d < D; ++d)
            pot += g[d] * (R(0) - pos[i * D + d]);
        s += R(0.5) * mass[i] * sqnorm<D, R>(vel + i * D) + mass[i] * pot;
    }
    return s;
}
template <int D, class R>
R maximumSpeed(Body<D, R> &b)
{
    R *vel = b.v("Velocity", D);
    R m = std::numeric_limits<R>::lowest();
    for (size_t i = 0; i < b.n; ++i)
        m = std::max(m, vnorm<D, R>(vel + i * D));
    return m;
}

// general_dynamics/general_reduce.cpp:28-49 PositionLowerBound / PositionUpperBound (component-wise min / max)
template <int D, class R>
void positionBound(Body<D, R> &b, bool upper, double *out3)
{
    R *pos = b.v("Position", D);
    for (int d = 0; d < 3; ++d)
        out3[d] = 0.0;
    for (int d = 0; d < D; ++d)
    {
        R m = upper ? std::numeric_limits<R>::lowest() : std::numeric_limits<R>::max();
        for (size_t i = 0; i < b.n; ++i)
            m = upper ? std::max(m, pos[i * D + d]) : std::min(m, pos[i * D + d]);
        out3[d] = double(m);
    }
}

// general_dynamics/general_interpolation.h:63-83 — observer interpolation of a width-wide variable
template <int D, class R>
void observe(Relation<D, R> &contact, const std::string &name, int width, R *out)
{
    for (size_t i = 0; i < contact.body->n; ++i)
    {
        std::vector<R> q(width, R(0));
        R w = 0;
        for (size_t t = 0; t < contact.targets.size(); ++t)
        {
            R *Vol = contact.targets[t]->v("VolumetricMeasure");
            R *data = contact.targets[t]->v(name, width);
            const auto &cn = contact.configs[t][i];
            for (size_t k = 0; k != cn.current_size; ++k)
            {
                R wj = cn.W[k] * Vol[cn.j[k]];
                for (int d = 0; d < width; ++d)
                    q[d] += wj * data[cn.j[k] * width + d];
                w += wj;
            }
        }
        for (int d = 0; d < width; ++d)
            out[i * width + d] = q[d] / (w + Tiny<R>::value);
    }
}

// configuration_dynamics/particle_sorting.cpp:31-58 — key = Morton(cell); the reference's in-place
// unstable quicksort is replaced by a stable sort on the same key (permutation is not
// pinned, keys are).  All evolving variables + OriginalID move; SortedID is rebuilt.
template <int D, class R>
void particleSort(Body<D, R> &b, std::vector<uint64_t> *keys_out)
{
    const R *pos = b.v("Position", D);
    std::vector<uint64_t> key(b.n);
    for (size_t i = 0; i < b.n; ++i)
    {
        int c[D];
        b.mesh.cellIndex(pos + i * D, c);
        key[i] = Mesh<D, R>::morton(c);
    }
    if (keys_out)
        *keys_out = key;
    std::vector<size_t> perm(b.n);
    for (size_t i = 0; i < b.n; ++i)
        perm[i] = i;
    std::stable_sort(perm.begin(), perm.end(), [&](size_t a, size_t c) { return key[a] < key[c]; });
    for (auto &kv : b.vars)
    {
        if (!kv.second.evolving)
            continue;
        int w = kv.second.width;
        std::vector<R> tmp(kv.second.data.size());
        for (size_t i = 0; i < b.n; ++i)
            for (int d = 0; d < w; ++d)
                tmp[i * w + d] = kv.second.data[perm[i] * w + d];
        kv.second.data.swap(tmp);
    }
    std::vector<size_t> oid(b.n);
    for (size_t i = 0; i < b.n; ++i)
        oid[i] = b.original_id[perm[i]];
    b.original_id.swap(oid);
    for (size_t i = 0; i < b.n; ++i)
        b.sorted_id[b.original_id[i]] = i;
}

} // namespace orc
