// integration/skeleton/sphinxsys.h - stands in for the reference's umbrella header (src/shared/include/sphinxsys.h)
// when a reference main is compiled against the skeleton + a device back end.  -DSPHINXSYS_USE_CUDA selects the CUDA
// execution policy as the default ExecutionPolicy of the algorithm templates, which is the ONE change a user makes
// (the reference's mains leave the policy argument defaulted).
#pragma once
#ifdef SPHINXSYS_USE_CUDA
namespace SPH
{
namespace execution
{
class ParallelCudaPolicy;
}
} // namespace SPH
#define SPHINXSYS_DEFAULT_POLICY execution::ParallelCudaPolicy
#endif
#include "sph_base.h"
#include "sph_bodies.h"
#include "sph_dynamics.h"
#include "sph_io.h"
#include "sph_ck.h"
#include "sph_solid.h"
#include "sph_periodic.h"
#ifdef SPHINXSYS_USE_CUDA
#include "sph_cuda.h"
#include "sph_cuda_dynamics.h"
#include "sph_cuda_ck.h"
#include "sph_cuda_solid.h"
#include "sph_cuda_periodic.h"
#include "sph_cuda_solver.h"
#else
#include "sph_solver.h"
namespace SPH
{
inline void attachExecutionBackEnd(SPHBody &) {}
} // namespace SPH
#endif
