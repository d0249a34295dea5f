// integration/src_cuda (see sph_cuda.h): what the complex interactions assembled by ParticleMethodContainer's factories
// (tests/tests_sycl/2d_examples/test_2d_dambreak_sycl/dambreak_sycl.cpp:96-118) run on the CUDA policy.
#pragma once
#include "sph_solver.h"
#include "sph_cuda_ck.h"

namespace SPH
{
namespace cuda
{
template <class Chain>
inline sphb200_ctx *ctxOfChain(Chain &c) { return context(c.inner().getSPHBody().getSPHSystem()); }
template <class Chain>
inline sphb200_relation *contactOfChain(Chain &c) { return c.contact() ? contactHandle(*c.contact()) : nullptr; }
} // namespace cuda

// addInteractionDynamics<LinearCorrectionMatrix, WithUpdate>(inner, alpha).addPostContactInteraction(contact)
template <>
struct DeviceChain<ParallelCudaPolicy, LinearCorrectionMatrix>
{
    static constexpr bool available = true;
    template <class Chain>
    static void run(Chain &c, Real)
    {
        cuda::ensureOnDevice(c.inner().getSPHBody());
        cuda::check(sphb200_ck_linear_correction_matrix(cuda::innerHandle(c.inner()), cuda::contactOfChain(c), double(c.scalarArgument())),
                    cuda::ctxOfChain(c), "LinearCorrectionMatrix");
    }
};
// addInteractionDynamics<AcousticStep1stHalf | 2ndHalf, OneLevel, AcousticRiemannSolverCK, LinearCorrectionCK>(inner)
//     .addPostContactInteraction<Wall, AcousticRiemannSolverCK, LinearCorrectionCK>(contact)
template <>
struct DeviceChain<ParallelCudaPolicy, fluid_dynamics::AcousticStep1stHalf>
{
    static constexpr bool available = true;
    template <class Chain>
    static void run(Chain &c, Real dt)
    {
        cuda::ensureOnDevice(c.inner().getSPHBody());
        cuda::check(sphb200_ck_acoustic_step_1st_half(cuda::innerHandle(c.inner()), cuda::contactOfChain(c), double(dt)), cuda::ctxOfChain(c),
                    "AcousticStep1stHalf");
    }
};
template <>
struct DeviceChain<ParallelCudaPolicy, fluid_dynamics::AcousticStep2ndHalf>
{
    static constexpr bool available = true;
    template <class Chain>
    static void run(Chain &c, Real dt)
    {
        cuda::check(sphb200_ck_acoustic_step_2nd_half(cuda::innerHandle(c.inner()), cuda::contactOfChain(c), double(dt)), cuda::ctxOfChain(c),
                    "AcousticStep2ndHalf");
    }
};
// addInteractionDynamics<CompressionSummation>(inner).addPostContactInteraction(contact)
//     .addPostStateDynamics<DensityRegularization, WeaklyCompressibleFluid, FreeSurface>(body)
template <>
struct DeviceChain<ParallelCudaPolicy, fluid_dynamics::CompressionSummation>
{
    static constexpr bool available = true;
    template <class Chain>
    static void run(Chain &c, Real)
    {
        cuda::ensureOnDevice(c.inner().getSPHBody());
        cuda::check(sphb200_ck_compression_summation(cuda::innerHandle(c.inner()), cuda::contactOfChain(c)), cuda::ctxOfChain(c), "CompressionSummation");
        if (c.postStateBody())
            cuda::check(sphb200_ck_density_regularize(cuda::bodyHandle(*c.postStateBody()), c.postFreeSurface() ? 1 : 0), cuda::ctxOfChain(c),
                        "DensityRegularization");
    }
};
} // namespace SPH
