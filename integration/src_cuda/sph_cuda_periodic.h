// integration/src_cuda (see sph_cuda.h): device implementations for tests/2d_examples/test_2d_taylor_green/taylor_green.cpp -
// viscous force, transport-velocity correction and the periodic condition (whose ghost images the library generates inside
// the neighbour search, so the "update cell linked list" moment has nothing to insert).
#pragma once
#include "sph_periodic.h"
#include "sph_cuda_dynamics.h"

namespace SPH
{
// InteractionWithUpdate<fluid_dynamics::ViscousForceInner> (viscous_dynamics.hpp:35-53 + ForcePrior::update)
template <>
struct DeviceImplementation<ParallelCudaPolicy, fluid_dynamics::ViscousForceInner>
{
    static constexpr bool available = true;
    static void run(fluid_dynamics::ViscousForceInner &l, Real)
    {
        cuda::prepare(l);
        cuda::check(sphb200_viscous_force(cuda::relationHandle(l.inner()), nullptr), cuda::ctxOf(l), "ViscousForceInner");
    }
};
// InteractionWithUpdate<fluid_dynamics::TransportVelocityCorrectionInner<TruncatedLinear, AllParticles>>
// (transport_velocity_correction.hpp:38-67)
template <class Limiter, class Scope>
struct DeviceImplementation<ParallelCudaPolicy, fluid_dynamics::TransportVelocityCorrectionInner<Limiter, Scope>>
{
    static constexpr bool available = true;
    static void run(fluid_dynamics::TransportVelocityCorrectionInner<Limiter, Scope> &l, Real)
    {
        cuda::prepare(l);
        cuda::check(sphb200_transport_velocity_correction(cuda::relationHandle(l.inner()), double(l.coefficient())), cuda::ctxOf(l),
                    "TransportVelocityCorrectionInner");
    }
};
// PeriodicConditionUsingCellLinkedList (domain_bounding.h:93-175)
template <>
struct DevicePeriodic<ParallelCudaPolicy>
{
    static void declare(RealBody &body, int axis, Real lo, Real hi)
    {
        cuda::check(sphb200_body_set_periodic(cuda::bodyHandle(body), axis, double(lo), double(hi)), cuda::context(body.getSPHSystem()), "PeriodicAlongAxis");
    }
    static void bounding(RealBody &body, int axis)
    {
        cuda::check(sphb200_periodic_bounding(cuda::bodyHandle(body), axis), cuda::context(body.getSPHSystem()), "PeriodicCondition::bounding_");
    }
    static void updateCellLinkedList(RealBody &, int) {} // periodic images are generated inside the neighbour search
};
} // namespace SPH
