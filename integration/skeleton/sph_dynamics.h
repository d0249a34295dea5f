// integration/skeleton (see sph_base.h): the Dynamics API - algorithm templates with an execution-policy parameter
// (particle_dynamics/dynamics_algorithms.h:100-353, common/base_dynamics.h:42-63, base_local_dynamics.h:55-88) and
// the local dynamics of the path as declarations with the reference's constructor signatures.
#pragma once
#include "sph_bodies.h"

namespace SPH
{
#ifndef SPHINXSYS_DEFAULT_POLICY // the reference's default is ParallelPolicy; a CUDA build passes -DSPHINXSYS_USE_CUDA
#define SPHINXSYS_DEFAULT_POLICY execution::ParallelPolicy
#endif

template <class ReturnType = void>
class BaseDynamics
{
  public:
    virtual ~BaseDynamics() = default;
    virtual ReturnType exec(Real dt = Real(0)) = 0;
};

class LocalDynamics // base_local_dynamics.h:55-88
{
  public:
    explicit LocalDynamics(SPHBody &body) : sph_body_(&body), particles_(&body.getBaseParticles()) {}
    SPHBody &getSPHBody() { return *sph_body_; }
    BaseParticles *getParticles() { return particles_; }
    void setupDynamics(Real = Real(0)) {}

  protected:
    SPHBody *sph_body_;
    BaseParticles *particles_;
};

// The customisation point of a device back end for the classic algorithm templates.  A generic lambda cannot cross a
// C ABI, so a back end states PER LOCAL-DYNAMICS TYPE that it has a fused implementation (available = true, run());
// every other local dynamics falls back to: synchronizeWithDevice -> host loop -> synchronizeToDevice.
template <class ExecutionPolicy, class LocalDynamicsType, class Enable = void>
struct DeviceImplementation
{
    static constexpr bool available = false;
};

namespace detail
{
template <class T, class = void>
struct has_update : std::false_type {};
template <class T>
struct has_update<T, std::void_t<decltype(std::declval<T &>().update(size_t(0), Real(0)))>> : std::true_type {};
template <class T, class = void>
struct has_reduce : std::false_type {};
template <class T>
struct has_reduce<T, std::void_t<decltype(std::declval<T &>().reduce(size_t(0), Real(0)))>> : std::true_type {};
[[noreturn]] inline void noHostNumerics(const char *what)
{
    std::cout << "\n Error: " << what << " has no host implementation in the integration skeleton (the reference owns it)." << std::endl;
    std::exit(1);
}
// particle_for(policy, range, f): sequential here, the reference runs TBB (particle_iterators.h)
template <class Policy, class F>
inline void particle_for(const Policy &, size_t n, const F &f)
{
    for (size_t i = 0; i < n; ++i)
        f(i);
}
template <class L>
inline void hostFallbackBegin(L &local)
{
    auto &sync = local.getParticles()->sync_all_;
    if (sync)
        sync(true);
}
template <class L>
inline void hostFallbackEnd(L &local)
{
    auto &sync = local.getParticles()->sync_all_;
    if (sync)
        sync(false);
}
} // namespace detail

// SimpleDynamics (dynamics_algorithms.h:100-125): update(i, dt) for every particle
template <class LocalDynamicsType, class ExecutionPolicy = SPHINXSYS_DEFAULT_POLICY>
class SimpleDynamics : public LocalDynamicsType, public BaseDynamics<void>
{
  public:
    template <typename... Args>
    SimpleDynamics(Args &&...args) : LocalDynamicsType(std::forward<Args>(args)...) {}
    void exec(Real dt = Real(0)) override
    {
        using Impl = DeviceImplementation<ExecutionPolicy, LocalDynamicsType>;
        this->setupDynamics(dt);
        if constexpr (Impl::available)
            Impl::run(*this, dt);
        else if constexpr (detail::has_update<LocalDynamicsType>::value)
        {
            if constexpr (is_device_policy<ExecutionPolicy>)
                detail::hostFallbackBegin(*this);
            detail::particle_for(execution::par_host, this->particles_->TotalRealParticles(), [&](size_t i) { this->update(i, dt); });
            if constexpr (is_device_policy<ExecutionPolicy>)
                detail::hostFallbackEnd(*this);
        }
        else
            detail::noHostNumerics("SimpleDynamics<...>");
    }
};

// ReduceDynamics (dynamics_algorithms.h:127-160): reduce over the particles, then outputResult
template <class LocalDynamicsType, class ExecutionPolicy = SPHINXSYS_DEFAULT_POLICY>
class ReduceDynamics : public LocalDynamicsType, public BaseDynamics<typename LocalDynamicsType::ReturnType>
{
  public:
    using ReturnType = typename LocalDynamicsType::ReturnType;
    template <typename... Args>
    ReduceDynamics(Args &&...args) : LocalDynamicsType(std::forward<Args>(args)...) {}
    std::string QuantityName() { return this->quantity_name_; }
    ReturnType exec(Real dt = Real(0)) override
    {
        using Impl = DeviceImplementation<ExecutionPolicy, LocalDynamicsType>;
        this->setupDynamics(dt);
        if constexpr (Impl::available)
            return Impl::run(*this, dt);
        else
            detail::noHostNumerics("ReduceDynamics<...>");
    }
};

// InteractionDynamics / InteractionWithUpdate / Dynamics1Level (dynamics_algorithms.h:237-353): the hot ones.
template <class LocalDynamicsType, class ExecutionPolicy = SPHINXSYS_DEFAULT_POLICY>
class InteractionWithUpdate : public LocalDynamicsType, public BaseDynamics<void>
{
  public:
    template <typename... Args>
    InteractionWithUpdate(Args &&...args) : LocalDynamicsType(std::forward<Args>(args)...) {}
    void exec(Real dt = Real(0)) override
    {
        using Impl = DeviceImplementation<ExecutionPolicy, LocalDynamicsType>;
        this->setupDynamics(dt);
        if constexpr (Impl::available)
            Impl::run(*this, dt);
        else
            detail::noHostNumerics("InteractionWithUpdate<...>");
    }
};
template <class LocalDynamicsType, class ExecutionPolicy = SPHINXSYS_DEFAULT_POLICY>
class Dynamics1Level : public LocalDynamicsType, public BaseDynamics<void>
{
  public:
    template <typename... Args>
    Dynamics1Level(Args &&...args) : LocalDynamicsType(std::forward<Args>(args)...) {}
    void exec(Real dt = Real(0)) override
    {
        using Impl = DeviceImplementation<ExecutionPolicy, LocalDynamicsType>;
        this->setupDynamics(dt);
        if constexpr (Impl::available)
            Impl::run(*this, dt);
        else
            detail::noHostNumerics("Dynamics1Level<...>");
    }
};

// ---------------------------------------------------------------------------------------------------------
// local dynamics of the path: constructor signatures and registered variables of the reference, no numerics
// ---------------------------------------------------------------------------------------------------------
class Gravity // external_force/external_force.h:38-57
{
  public:
    explicit Gravity(const Vecd &g) : g_(g) {}
    const Vecd &Acceleration() const { return g_; }

  private:
    Vecd g_;
};
template <class GravityType>
class GravityForce : public LocalDynamics // general_dynamics/force_prior.hpp:28-41
{
  public:
    GravityForce(SPHBody &body, const GravityType &gravity) : LocalDynamics(body), gravity_(gravity)
    {
        particles_->registerStateVariable<Vecd>("ForcePrior");
        particles_->registerStateVariable<Vecd>("PreviousGravityForce");
        particles_->addEvolvingVariable<Vecd>("PreviousGravityForce");
    }
    const GravityType &gravity() const { return gravity_; }

  private:
    GravityType gravity_;
};
// general_dynamics/general_geometric.cpp:18-26: a NON-hot local dynamics, it really runs on the host here and
// exercises the generic fall-back of a device policy
class NormalDirectionFromBodyShape : public LocalDynamics
{
  public:
    explicit NormalDirectionFromBodyShape(SPHBody &body)
        : LocalDynamics(body), shape_(body.getInitialShape()), pos_(particles_->getVariableDataByName<Vecd>("Position")),
          n_(particles_->registerStateVariableData<Vecd>("NormalDirection"))
    {
    }
    void update(size_t i, Real = Real(0)) { n_[i] = shape_.findNormalDirection(pos_[i]); }

  private:
    Shape &shape_;
    Vecd *pos_, *n_;
};

namespace fluid_dynamics
{
class FluidIntegrationBase : public LocalDynamics // BaseIntegration ctor, fluid_integration.hpp:9-47
{
  public:
    FluidIntegrationBase(BaseInnerRelation &inner, BaseContactRelation *wall) : LocalDynamics(inner.getSPHBody()), inner_(inner), wall_(wall)
    {
        particles_->registerStateVariable<Real>("Pressure");
        particles_->registerStateVariable<Real>("DensityChangeRate");
        particles_->registerStateVariable<Vecd>("Velocity");
        particles_->registerStateVariable<Vecd>("Force");
        particles_->registerStateVariable<Vecd>("ForcePrior");
        particles_->addEvolvingVariable<Vecd>("Velocity");
        particles_->addEvolvingVariable<Real>("Mass");
        particles_->addEvolvingVariable<Vecd>("ForcePrior");
        particles_->addEvolvingVariable<Vecd>("Force");
        particles_->addEvolvingVariable<Real>("DensityChangeRate");
        particles_->addEvolvingVariable<Real>("Density");
        particles_->addEvolvingVariable<Real>("Pressure");
        particles_->addVariableToWrite<Vecd>("Velocity");
        if (wall)
            for (RealBody *w : wall->getContactBodies())
            { // InteractionWithWall, base_fluid_dynamics.h:54-75
                w->getBaseParticles().registerStateVariable<Vecd>("Velocity");
                w->getBaseParticles().registerStateVariable<Vecd>("Acceleration");
                w->getBaseParticles().registerStateVariable<Vecd>("NormalDirection");
            }
    }
    BaseInnerRelation &inner() { return inner_; }
    BaseContactRelation *wall() { return wall_; }

  private:
    BaseInnerRelation &inner_;
    BaseContactRelation *wall_;
};
#define SPH_SKELETON_INTEGRATION(NAME, HALF, RIEMANN, WITH_WALL)                                                      \
    class NAME : public FluidIntegrationBase                                                                         \
    {                                                                                                                \
      public:                                                                                                        \
        static constexpr int half = HALF;                                                                            \
        static constexpr bool riemann = RIEMANN, with_wall = WITH_WALL;                                              \
        NAME(BaseInnerRelation &inner, BaseContactRelation &wall) : FluidIntegrationBase(inner, &wall) {}            \
        explicit NAME(BaseInnerRelation &inner) : FluidIntegrationBase(inner, nullptr) {}                            \
    };
SPH_SKELETON_INTEGRATION(Integration1stHalfWithWallRiemann, 1, true, true) // fluid_integration.h:141-146, 204-207
SPH_SKELETON_INTEGRATION(Integration2ndHalfWithWallRiemann, 2, true, true)
SPH_SKELETON_INTEGRATION(Integration1stHalfWithWallNoRiemann, 1, false, true)
SPH_SKELETON_INTEGRATION(Integration2ndHalfWithWallNoRiemann, 2, false, true)
SPH_SKELETON_INTEGRATION(Integration1stHalfInnerRiemann, 1, true, false) // fluid_integration.h:101-102, 171-172
SPH_SKELETON_INTEGRATION(Integration2ndHalfInnerRiemann, 2, true, false)
SPH_SKELETON_INTEGRATION(Integration1stHalfInnerNoRiemann, 1, false, false)
SPH_SKELETON_INTEGRATION(Integration2ndHalfInnerNoRiemann, 2, false, false)
#undef SPH_SKELETON_INTEGRATION

class DensitySummationBase : public LocalDynamics // density_summation.hpp:11-21
{
  public:
    DensitySummationBase(BaseInnerRelation &inner, BaseContactRelation *contact) : LocalDynamics(inner.getSPHBody()), inner_(inner), contact_(contact)
    {
        particles_->registerStateVariable<Real>("DensitySummation");
    }
    BaseInnerRelation &inner() { return inner_; }
    BaseContactRelation *contact() { return contact_; }

  private:
    BaseInnerRelation &inner_;
    BaseContactRelation *contact_;
};
class DensitySummationComplexFreeSurface : public DensitySummationBase // density_summation.h:180
{
  public:
    static constexpr bool free_surface = true;
    DensitySummationComplexFreeSurface(BaseInnerRelation &inner, BaseContactRelation &contact) : DensitySummationBase(inner, &contact) {}
};
class DensitySummationComplex : public DensitySummationBase // density_summation.h:178
{
  public:
    static constexpr bool free_surface = false;
    DensitySummationComplex(BaseInnerRelation &inner, BaseContactRelation &contact) : DensitySummationBase(inner, &contact) {}
};
class DensitySummationInner : public DensitySummationBase // density_summation.h:77
{
  public:
    static constexpr bool free_surface = false;
    explicit DensitySummationInner(BaseInnerRelation &inner) : DensitySummationBase(inner, nullptr) {}
};

class AdvectionTimeStep : public LocalDynamics // fluid_time_step.cpp:44-65
{
  public:
    using ReturnType = Real;
    static constexpr bool viscous = false;
    AdvectionTimeStep(SPHBody &body, Real U_ref, Real advectionCFL = Real(0.25)) : LocalDynamics(body), speed_ref_(U_ref), advectionCFL_(advectionCFL) {}
    Real speed_ref_, advectionCFL_;
    std::string quantity_name_ = "AdvectionTimeStep";
};
class AdvectionViscousTimeStep : public AdvectionTimeStep // fluid_time_step.cpp:67-79
{
  public:
    static constexpr bool viscous = true;
    using AdvectionTimeStep::AdvectionTimeStep;
};
class AcousticTimeStep : public LocalDynamics // fluid_time_step.cpp:12-31
{
  public:
    using ReturnType = Real;
    explicit AcousticTimeStep(SPHBody &body, Real acousticCFL = Real(0.6)) : LocalDynamics(body), acousticCFL_(acousticCFL) {}
    Real acousticCFL_;
    std::string quantity_name_ = "AcousticTimeStep";
};
} // namespace fluid_dynamics

class TotalMechanicalEnergy : public LocalDynamics // general_reduce.cpp:66-79
{
  public:
    using ReturnType = Real;
    TotalMechanicalEnergy(SPHBody &body, Gravity &gravity) : LocalDynamics(body), gravity_(gravity) {}
    Gravity &gravity_;
    std::string quantity_name_ = "TotalMechanicalEnergy";
};
class TotalKineticEnergy : public LocalDynamics // general_reduce.cpp:54-64
{
  public:
    using ReturnType = Real;
    explicit TotalKineticEnergy(SPHBody &body) : LocalDynamics(body) {}
    std::string quantity_name_ = "TotalKineticEnergy";
};
class MaximumSpeed : public LocalDynamics // general_reduce.cpp:18-27
{
  public:
    using ReturnType = Real;
    explicit MaximumSpeed(SPHBody &body) : LocalDynamics(body) {}
    std::string quantity_name_ = "MaximumSpeed";
};
class PositionUpperBound : public LocalDynamics // general_reduce.cpp:40-49
{
  public:
    using ReturnType = Vecd;
    explicit PositionUpperBound(SPHBody &body) : LocalDynamics(body) {}
    std::string quantity_name_ = "PositionUpperBound";
};

// configuration_dynamics/particle_sorting.h: ParticleSorting<ExecutionPolicy>::exec()
template <class ExecutionPolicy = SPHINXSYS_DEFAULT_POLICY>
class ParticleSortingT
{
  public:
    explicit ParticleSortingT(RealBody &body) : body_(body) {}
    void exec(Real = Real(0)) { body_.updateCellLinkedList(); } // the cell-list update of a device back end keeps the body sorted

  private:
    RealBody &body_;
};
using ParticleSorting = ParticleSortingT<>;
} // namespace SPH
