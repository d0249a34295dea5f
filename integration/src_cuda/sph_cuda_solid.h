// integration/src_cuda (see sph_cuda.h): device implementations of the total-Lagrangian solid dynamics of
// tests/3d_examples/test_3d_twisting_column/twisting_column.cpp and the restart hand-over of RestartIO.
#pragma once
#include "sph_solid.h"
#include "sph_cuda_dynamics.h"

namespace SPH
{
// InteractionWithUpdate<LinearGradientCorrectionMatrixInner> (kernel_correction.cpp:8-29)
template <>
struct DeviceImplementation<ParallelCudaPolicy, LinearGradientCorrectionMatrixInner>
{
    static constexpr bool available = true;
    static void run(LinearGradientCorrectionMatrixInner &l, Real)
    {
        cuda::prepare(l);
        cuda::check(sphb200_correction_matrix(cuda::relationHandle(l.inner())), cuda::ctxOf(l), "LinearGradientCorrectionMatrixInner");
    }
};
// Dynamics1Level<solid_dynamics::DecomposedIntegration1stHalf | Integration1stHalfPK2 | Integration2ndHalf>
// (elastic_dynamics.h:171-351, elastic_dynamics.cpp:62-194)
template <class L>
struct DeviceImplementation<ParallelCudaPolicy, L, std::enable_if_t<std::is_base_of<solid_dynamics::SolidIntegrationBase, L>::value>>
{
    static constexpr bool available = true;
    static void run(L &l, Real dt)
    {
        cuda::prepare(l);
        if (L::half == 1)
            cuda::check(sphb200_solid_1st_half(cuda::relationHandle(l.inner()), L::decomposed ? 1 : 0, double(dt)), cuda::ctxOf(l), "solid Integration1stHalf");
        else
            cuda::check(sphb200_solid_2nd_half(cuda::relationHandle(l.inner()), double(dt)), cuda::ctxOf(l), "solid Integration2ndHalf");
    }
};
// ReduceDynamics<solid_dynamics::AcousticTimeStep> (elastic_dynamics.cpp:12-30)
template <>
struct DeviceImplementation<ParallelCudaPolicy, solid_dynamics::AcousticTimeStep>
{
    static constexpr bool available = true;
    static Real run(solid_dynamics::AcousticTimeStep &l, Real)
    {
        double out = 0;
        cuda::check(sphb200_solid_dt(cuda::bodyHandle(l.getSPHBody()), double(l.CFL_), &out), cuda::ctxOf(l), "solid AcousticTimeStep");
        return Real(out);
    }
};
// SimpleDynamics<FixBodyPartConstraint> (general_constraint.h:106-121): the part's particles by ORIGINAL id
template <>
struct DeviceImplementation<ParallelCudaPolicy, FixBodyPartConstraint>
{
    static constexpr bool available = true;
    static void run(FixBodyPartConstraint &l, Real)
    {
        const std::vector<long long> &ids = l.part().OriginalIds();
        static_assert(sizeof(long long) == sizeof(int64_t), "id width");
        cuda::check(sphb200_solid_constrain(cuda::bodyHandle(l.getSPHBody()), reinterpret_cast<const int64_t *>(ids.data()), (long long)ids.size()),
                    cuda::ctxOf(l), "FixBodyPartConstraint");
    }
};
// RestartIO: the evolving state of a body through its DiscreteVariables' synchronizeWithDevice / synchronizeToDevice
template <>
struct DeviceRestart<ParallelCudaPolicy>
{
    static void toHost(SPHBody &body) { cuda::syncAll(body, true); }
    static void toDevice(SPHBody &body) { cuda::syncAll(body, false); }
};
} // namespace SPH
