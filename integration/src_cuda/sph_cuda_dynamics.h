// integration/src_cuda (see sph_cuda.h): the fused implementations of the hot local dynamics - one C-ABI call per
// exec() - plus observer interpolation and the asynchronous output path.
#pragma once
#include "sph_cuda.h"

namespace SPH
{
inline void attachExecutionBackEnd(SPHBody &body) { cuda::attach(body); }

namespace cuda
{
inline sphb200_ctx *ctxOf(LocalDynamics &l) { return context(l.getSPHBody().getSPHSystem()); }
inline sphb200_relation *wallHandle(BaseContactRelation *w) { return w ? relationHandle(*w) : nullptr; }
template <class L>
inline void prepare(L &local)
{ // variables a constructor registered after the last upload (e.g. by a later-constructed dynamics) reach the device
    ensureOnDevice(local.getSPHBody());
}
} // namespace cuda
using execution::ParallelCudaPolicy;

// SimpleDynamics<GravityForce<Gravity>> (force_prior.hpp:28-41)
template <>
struct DeviceImplementation<ParallelCudaPolicy, GravityForce<Gravity>>
{
    static constexpr bool available = true;
    static void run(GravityForce<Gravity> &l, Real)
    {
        cuda::prepare(l);
        double g[3] = {0, 0, 0};
        for (int d = 0; d < Dimensions; ++d)
            g[d] = double(l.gravity().Acceleration()[d]);
        cuda::check(sphb200_gravity_force(cuda::bodyHandle(l.getSPHBody()), g), cuda::ctxOf(l), "GravityForce");
    }
};

// Dynamics1Level<Integration{1st,2nd}Half...> (fluid_integration.hpp:50-231)
template <class L>
struct DeviceImplementation<ParallelCudaPolicy, L, std::enable_if_t<std::is_base_of<fluid_dynamics::FluidIntegrationBase, L>::value>>
{
    static constexpr bool available = true;
    static void run(L &l, Real dt)
    {
        cuda::prepare(l);
        sphb200_relation *in = cuda::relationHandle(l.inner()), *w = cuda::wallHandle(l.wall());
        if (L::half == 1)
            cuda::check(sphb200_integration_1st_half(in, w, L::riemann ? 1 : 0, double(dt)), cuda::ctxOf(l), "Integration1stHalf");
        else
            cuda::check(sphb200_integration_2nd_half(in, w, L::riemann ? 1 : 0, double(dt)), cuda::ctxOf(l), "Integration2ndHalf");
    }
};

// InteractionWithUpdate<DensitySummation...> (density_summation.cpp:8-22,58-78)
template <class L>
struct DeviceImplementation<ParallelCudaPolicy, L, std::enable_if_t<std::is_base_of<fluid_dynamics::DensitySummationBase, L>::value>>
{
    static constexpr bool available = true;
    static void run(L &l, Real)
    {
        cuda::prepare(l);
        cuda::check(sphb200_density_summation(cuda::relationHandle(l.inner()), cuda::wallHandle(l.contact()), L::free_surface ? 1 : 0), cuda::ctxOf(l),
                    "DensitySummation");
    }
};

// ReduceDynamics<AdvectionTimeStep | AdvectionViscousTimeStep | AcousticTimeStep> (fluid_time_step.cpp:12-79)
template <class L>
struct DeviceImplementation<ParallelCudaPolicy, L, std::enable_if_t<std::is_base_of<fluid_dynamics::AdvectionTimeStep, L>::value>>
{
    static constexpr bool available = true;
    static Real run(L &l, Real)
    {
        double out = 0;
        cuda::check(sphb200_advection_dt(cuda::bodyHandle(l.getSPHBody()), double(l.speed_ref_), double(l.advectionCFL_), L::viscous ? 1 : 0, &out),
                    cuda::ctxOf(l), "AdvectionTimeStep");
        return Real(out);
    }
};
template <>
struct DeviceImplementation<ParallelCudaPolicy, fluid_dynamics::AcousticTimeStep>
{
    static constexpr bool available = true;
    static Real run(fluid_dynamics::AcousticTimeStep &l, Real)
    {
        double out = 0;
        cuda::check(sphb200_acoustic_dt(cuda::bodyHandle(l.getSPHBody()), double(l.acousticCFL_), &out), cuda::ctxOf(l), "AcousticTimeStep");
        return Real(out);
    }
};

// body reductions (general_reduce.cpp:18-79)
template <>
struct DeviceImplementation<ParallelCudaPolicy, TotalMechanicalEnergy>
{
    static constexpr bool available = true;
    static Real run(TotalMechanicalEnergy &l, Real)
    {
        double g[3] = {0, 0, 0}, out = 0;
        for (int d = 0; d < Dimensions; ++d)
            g[d] = double(l.gravity_.Acceleration()[d]);
        cuda::check(sphb200_total_mechanical_energy(cuda::bodyHandle(l.getSPHBody()), g, &out), cuda::ctxOf(l), "TotalMechanicalEnergy");
        return Real(out);
    }
};
template <>
struct DeviceImplementation<ParallelCudaPolicy, TotalKineticEnergy>
{
    static constexpr bool available = true;
    static Real run(TotalKineticEnergy &l, Real)
    {
        double out = 0;
        cuda::check(sphb200_total_kinetic_energy(cuda::bodyHandle(l.getSPHBody()), &out), cuda::ctxOf(l), "TotalKineticEnergy");
        return Real(out);
    }
};
template <>
struct DeviceImplementation<ParallelCudaPolicy, MaximumSpeed>
{
    static constexpr bool available = true;
    static Real run(MaximumSpeed &l, Real)
    {
        double out = 0;
        cuda::check(sphb200_maximum_speed(cuda::bodyHandle(l.getSPHBody()), &out), cuda::ctxOf(l), "MaximumSpeed");
        return Real(out);
    }
};
template <>
struct DeviceImplementation<ParallelCudaPolicy, PositionUpperBound>
{
    static constexpr bool available = true;
    static Vecd run(PositionUpperBound &l, Real)
    {
        double out[3] = {0, 0, 0};
        cuda::check(sphb200_position_bound(cuda::bodyHandle(l.getSPHBody()), 1, out), cuda::ctxOf(l), "PositionUpperBound");
        Vecd r;
        for (int d = 0; d < Dimensions; ++d)
            r[d] = Real(out[d]);
        return r;
    }
};

// observer interpolation + output path
template <>
struct DeviceObservation<ParallelCudaPolicy>
{
    static constexpr bool available = true;
    static void observe(BaseContactRelation &contact, const std::string &name, int width, std::vector<double> &row)
    {
        SPHBody &obs = contact.getSPHBody();
        const size_t n = obs.getBaseParticles().TotalRealParticles();
        std::vector<double> out(n * size_t(width));
        cuda::check(sphb200_observe(cuda::relationHandle(contact), name.c_str(), width, out.data()), cuda::context(obs.getSPHSystem()), "ObservingAQuantity");
        row.insert(row.end(), out.begin(), out.end());
    }
    // BaseVariable::prepareForOutput(policy): asynchronous device -> host copies into the variables' own host arrays,
    // completed by one wait (the copies overlap whatever the compute stream does next)
    static void prepareForOutput(SPHBody &body)
    {
        sphb200_body *h = cuda::bodyHandle(body);
        sphb200_ctx *ctx = cuda::context(body.getSPHSystem());
        BaseParticles &p = body.getBaseParticles();
        std::vector<std::string> names = p.VariablesToWrite();
        names.push_back("Position");
        for (const std::string &nm : names)
        {
            void *dst = nullptr;
            if (auto *r = p.getVariableByName<Real>(nm))
                dst = r->Data();
            else if (auto *x = p.getVariableByName<Vecd>(nm))
                dst = x->Data();
            if (dst)
                cuda::check(sphb200_var_fetch_async(h, nm.c_str(), dst), ctx, "prepareForOutput");
        }
        cuda::check(sphb200_ctx_fetch_wait(ctx), ctx, "prepareForOutput");
    }
};
} // namespace SPH
