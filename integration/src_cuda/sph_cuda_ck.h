// integration/src_cuda (see sph_cuda.h): device implementations of the CK local dynamics that
// tests/tests_sycl/3d_examples/test_3d_dambreak_sycl/dambreak.cpp constructs - one C-ABI call per exec().
// The SYCL back end runs these as generic computing kernels; a C ABI cannot carry a kernel object, so each local
// dynamics type is routed to its fused sm_100a implementation (csrc/ck.cu).
#pragma once
#include "sph_ck.h"
#include "sph_cuda_dynamics.h"

namespace SPH
{
namespace cuda
{
inline sphb200_relation *contactHandle(Contact<> &c) { return relationHandle(static_cast<BaseContactRelation &>(c)); }
inline sphb200_relation *innerHandle(Inner<> &r) { return relationHandle(static_cast<BaseInnerRelation &>(r)); }
} // namespace cuda

// StateDynamics<MainExecutionPolicy, GravityForceCK<Gravity>> (force_prior_ck.h:66-90)
template <>
struct DeviceImplementation<ParallelCudaPolicy, GravityForceCK<Gravity>>
{
    static constexpr bool available = true;
    static void run(GravityForceCK<Gravity> &l, Real)
    {
        cuda::prepare(l);
        double g[3] = {0, 0, 0};
        for (int d = 0; d < Dimensions; ++d)
            g[d] = double(l.gravity().Acceleration()[d]);
        cuda::check(sphb200_gravity_force(cuda::bodyHandle(l.getSPHBody()), g), cuda::ctxOf(l), "GravityForceCK");
    }
};
// StateDynamics<..., fluid_dynamics::AdvectionStepSetup | UpdateParticlePosition> (fluid_time_step_ck.h:126-185)
template <>
struct DeviceImplementation<ParallelCudaPolicy, fluid_dynamics::AdvectionStepSetup>
{
    static constexpr bool available = true;
    static void run(fluid_dynamics::AdvectionStepSetup &l, Real)
    {
        cuda::prepare(l);
        cuda::check(sphb200_ck_advection_step_setup(cuda::bodyHandle(l.getSPHBody())), cuda::ctxOf(l), "AdvectionStepSetup");
    }
};
template <>
struct DeviceImplementation<ParallelCudaPolicy, fluid_dynamics::UpdateParticlePosition>
{
    static constexpr bool available = true;
    static void run(fluid_dynamics::UpdateParticlePosition &l, Real)
    {
        cuda::check(sphb200_ck_update_particle_position(cuda::bodyHandle(l.getSPHBody())), cuda::ctxOf(l), "UpdateParticlePosition");
    }
};
// InteractionDynamicsCK<..., LinearCorrectionMatrixComplex> (kernel_correction_ck.hpp:36-96)
template <>
struct DeviceImplementation<ParallelCudaPolicy, LinearCorrectionMatrixComplex>
{
    static constexpr bool available = true;
    static void run(LinearCorrectionMatrixComplex &l, Real)
    {
        cuda::prepare(l);
        cuda::check(sphb200_ck_linear_correction_matrix(cuda::innerHandle(l.inner()), cuda::contactHandle(l.contact()), double(l.alpha())), cuda::ctxOf(l),
                    "LinearCorrectionMatrixComplex");
    }
};
// InteractionDynamicsCK<..., AcousticStep{1st,2nd}HalfWithWallRiemannCorrectionCK> (acoustic_step_1st_half.hpp, _2nd_half.hpp)
template <>
struct DeviceImplementation<ParallelCudaPolicy, fluid_dynamics::AcousticStep1stHalfWithWallRiemannCorrectionCK>
{
    static constexpr bool available = true;
    static void run(fluid_dynamics::AcousticStep1stHalfWithWallRiemannCorrectionCK &l, Real dt)
    {
        cuda::check(sphb200_ck_acoustic_step_1st_half(cuda::innerHandle(l.inner()), cuda::contactHandle(l.contact()), double(dt)), cuda::ctxOf(l),
                    "AcousticStep1stHalf");
    }
};
template <>
struct DeviceImplementation<ParallelCudaPolicy, fluid_dynamics::AcousticStep2ndHalfWithWallRiemannCorrectionCK>
{
    static constexpr bool available = true;
    static void run(fluid_dynamics::AcousticStep2ndHalfWithWallRiemannCorrectionCK &l, Real dt)
    {
        cuda::check(sphb200_ck_acoustic_step_2nd_half(cuda::innerHandle(l.inner()), cuda::contactHandle(l.contact()), double(dt)), cuda::ctxOf(l),
                    "AcousticStep2ndHalf");
    }
};
// InteractionDynamicsCK<..., CompressionSummation<Inner<>, Contact<>>> + StateDynamics<..., DensityRegularization<...>>
// (density_regularization.hpp:42-111)
template <>
struct DeviceImplementation<ParallelCudaPolicy, fluid_dynamics::CompressionSummation<Inner<>, Contact<>>>
{
    static constexpr bool available = true;
    static void run(fluid_dynamics::CompressionSummation<Inner<>, Contact<>> &l, Real)
    {
        cuda::prepare(l);
        cuda::check(sphb200_ck_compression_summation(cuda::innerHandle(l.inner()), cuda::contactHandle(l.contact())), cuda::ctxOf(l), "CompressionSummation");
    }
};
template <class Identifier, class FluidType, class FlowType>
struct DeviceImplementation<ParallelCudaPolicy, fluid_dynamics::DensityRegularization<Identifier, FluidType, FlowType>>
{
    using L = fluid_dynamics::DensityRegularization<Identifier, FluidType, FlowType>;
    static constexpr bool available = true;
    static void run(L &l, Real)
    {
        cuda::check(sphb200_ck_density_regularize(cuda::bodyHandle(l.getSPHBody()), L::free_surface ? 1 : 0), cuda::ctxOf(l), "DensityRegularization");
    }
};
// InteractionDynamicsCK<..., FreeSurfaceIndicationComplexSpatialTemporalCK> (surface_indication_ck.hpp:55-170)
template <>
struct DeviceImplementation<ParallelCudaPolicy, fluid_dynamics::FreeSurfaceIndicationComplexSpatialTemporalCK>
{
    static constexpr bool available = true;
    static void run(fluid_dynamics::FreeSurfaceIndicationComplexSpatialTemporalCK &l, Real)
    {
        cuda::check(sphb200_ck_free_surface_indication(cuda::innerHandle(l.inner()), cuda::contactHandle(l.contact())), cuda::ctxOf(l),
                    "FreeSurfaceIndication");
    }
};
// ReduceDynamicsCK<..., AdvectionTimeStepCK | AcousticTimeStepCK<Fluid> | TotalMechanicalEnergyCK>
template <>
struct DeviceImplementation<ParallelCudaPolicy, fluid_dynamics::AdvectionTimeStepCK>
{
    static constexpr bool available = true;
    static Real run(fluid_dynamics::AdvectionTimeStepCK &l, Real)
    {
        double out = 0;
        cuda::check(sphb200_ck_advection_dt(cuda::bodyHandle(l.getSPHBody()), double(l.speed_ref_), double(l.advectionCFL_), &out), cuda::ctxOf(l),
                    "AdvectionTimeStepCK");
        return Real(out);
    }
};
template <class FluidType>
struct DeviceImplementation<ParallelCudaPolicy, fluid_dynamics::AcousticTimeStepCK<FluidType>>
{
    static constexpr bool available = true;
    static Real run(fluid_dynamics::AcousticTimeStepCK<FluidType> &l, Real)
    {
        double out = 0;
        cuda::check(sphb200_ck_acoustic_dt(cuda::bodyHandle(l.getSPHBody()), double(l.acousticCFL_), &out), cuda::ctxOf(l), "AcousticTimeStepCK");
        return Real(out);
    }
};
template <>
struct DeviceImplementation<ParallelCudaPolicy, TotalMechanicalEnergyCK>
{
    static constexpr bool available = true;
    static Real run(TotalMechanicalEnergyCK &l, Real)
    {
        double g[3] = {0, 0, 0}, out = 0;
        for (int d = 0; d < Dimensions; ++d)
            g[d] = double(l.gravity_.Acceleration()[d]);
        cuda::check(sphb200_total_mechanical_energy(cuda::bodyHandle(l.getSPHBody()), g, &out), cuda::ctxOf(l), "TotalMechanicalEnergyCK");
        return Real(out);
    }
};
} // namespace SPH
