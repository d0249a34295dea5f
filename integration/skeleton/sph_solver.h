// integration/skeleton (see sph_base.h): the solver / method-container form of the CK API that
// tests/tests_sycl/2d_examples/test_2d_dambreak_sycl/dambreak_sycl.cpp uses (shared_ck/particle_dynamics/sph_solver.h:41-206,
// particle_method_container.h): SPHSolver, ParticleMethodContainer<ExecutionPolicy> with its add...Dynamics factories, the
// chained addPostContactInteraction / addPostStateDynamics of a complex interaction, and TimeStepper with its interval
// triggers (sph_solver.cpp:53-107).  Included after the execution back end: `par_ck` is an INSTANCE of the main policy.
#pragma once
#include "sph_ck.h"
#include "sph_solid.h"

namespace SPH
{
inline constexpr MainExecutionPolicy par_ck{};

// template tags of the complex-interaction factories
struct WithUpdate {};
struct OneLevel {};
struct Wall {};
struct AcousticRiemannSolverCK {};
struct LinearCorrectionCK {};
struct NoKernelCorrectionCK {};
template <typename...>
class LinearCorrectionMatrix; // kernel_correction_ck.h - named as a template in addInteractionDynamics<LinearCorrectionMatrix, WithUpdate>
namespace fluid_dynamics
{
template <typename...>
class AcousticStep1stHalf; // acoustic_step_1st_half.h
template <typename...>
class AcousticStep2ndHalf; // acoustic_step_2nd_half.h
} // namespace fluid_dynamics

template <template <typename...> class T>
struct TemplateTag {};
template <template <typename...> class A, template <typename...> class B>
inline constexpr bool same_template = std::is_same<TemplateTag<A>, TemplateTag<B>>::value;

// what a back end runs for a complex interaction assembled by the factories
template <class ExecutionPolicy, template <typename...> class DynamicsType, class Enable = void>
struct DeviceChain
{
    static constexpr bool available = false;
};

// InteractionDynamicsCK<ExecutionPolicy, Dyn<Inner<...>>> + post contact interaction + post state dynamics
template <class ExecutionPolicy, template <typename...> class DynamicsType>
class InteractionChain : public BaseDynamics<void>
{
  public:
    InteractionChain(Inner<> &inner, Real scalar_arg) : inner_(inner), scalar_arg_(scalar_arg)
    {
        if constexpr (same_template<DynamicsType, fluid_dynamics::AcousticStep1stHalf> || same_template<DynamicsType, fluid_dynamics::AcousticStep2ndHalf>)
        { // AcousticStep ctor, acoustic_step_1st_half.hpp:13-39
            BaseParticles &p = inner.getSPHBody().getBaseParticles();
            p.registerStateVariable<Real>("Pressure");
            const bool fresh = p.getVariableByName<Real>("Compression") == nullptr;
            Real *compression = p.registerStateVariableData<Real>("Compression");
            for (size_t i = 0; fresh && i < p.TotalRealParticles(); ++i)
                compression[i] = Real(1);
            p.registerStateVariable<Real>("CompressionRate");
            p.registerStateVariable<Vecd>("Velocity");
            p.registerStateVariable<Vecd>("Force");
            p.registerStateVariable<Vecd>("ForcePrior");
            p.addEvolvingVariable<Vecd>("Velocity");
            p.addEvolvingVariable<Real>("Mass");
            p.addEvolvingVariable<Vecd>("ForcePrior");
            p.addEvolvingVariable<Real>("Compression");
            p.addEvolvingVariable<Real>("CompressionRate");
        }
    }
    template <typename... ContactParameters>
    InteractionChain &addPostContactInteraction(Contact<> &contact)
    {
        contact_ = &contact;
        for (RealBody *w : contact.getContactBodies())
            if (dynamic_cast<Solid *>(&w->getBaseMaterial()))
            { // Interaction<Wall>, interaction_ck.hpp:80-90
                w->getBaseParticles().registerStateVariable<Vecd>("Velocity");
                w->getBaseParticles().registerStateVariable<Vecd>("Acceleration");
                w->getBaseParticles().registerStateVariable<Vecd>("NormalDirection");
            }
        return *this;
    }
    template <template <typename...> class StateDynamicsType, typename... StateParameters>
    InteractionChain &addPostStateDynamics(SPHBody &body)
    {
        post_state_body_ = &body;
        post_free_surface_ = (std::is_same<StateParameters, FreeSurface>::value || ...);
        return *this;
    }
    void exec(Real dt = Real(0)) override
    {
        using Impl = DeviceChain<ExecutionPolicy, DynamicsType>;
        if constexpr (Impl::available)
            Impl::run(*this, dt);
        else
            detail::noHostNumerics("ParticleMethodContainer::addInteractionDynamics<...>");
    }
    Inner<> &inner() { return inner_; }
    Contact<> *contact() { return contact_; }
    Real scalarArgument() const { return scalar_arg_; }
    SPHBody *postStateBody() { return post_state_body_; }
    bool postFreeSurface() const { return post_free_surface_; }

  private:
    Inner<> &inner_;
    Contact<> *contact_ = nullptr;
    Real scalar_arg_;
    SPHBody *post_state_body_ = nullptr;
    bool post_free_surface_ = false;
};

template <class ExecutionPolicy>
class RestartIOCK : public RestartIOT<ExecutionPolicy> // io_system: restart through the variables' device delegation
{
  public:
    using RestartIOT<ExecutionPolicy>::RestartIOT;
};

// particle_method_container.h: owns the dynamics it creates and hands out references
template <class ExecutionPolicy>
class ParticleMethodContainer
{
  public:
    explicit ParticleMethodContainer(const ExecutionPolicy &) {}
    auto &addCellLinkedListDynamics(RealBody &body) { return keep<UpdateCellLinkedList<ExecutionPolicy, RealBody>>(body); }
    template <typename... RelationTypes>
    auto &addRelationDynamics(RelationTypes &...relations) { return keep<UpdateRelation<ExecutionPolicy, RelationTypes...>>(relations...); }
    auto &addSortDynamics(RealBody &body) { return keep<ParticleSortCK<ExecutionPolicy>>(body); }
    template <class LocalDynamicsType, typename... Args>
    auto &addStateDynamics(Args &&...args) { return keep<StateDynamics<ExecutionPolicy, LocalDynamicsType>>(std::forward<Args>(args)...); }
    template <class LocalDynamicsType, typename... Args>
    auto &addReduceDynamics(Args &&...args) { return keep<ReduceDynamicsCK<ExecutionPolicy, LocalDynamicsType>>(std::forward<Args>(args)...); }
    template <template <typename...> class DynamicsType, typename... Parameters>
    auto &addInteractionDynamics(Inner<> &inner, Real scalar_arg = Real(0))
    {
        return keep<InteractionChain<ExecutionPolicy, DynamicsType>>(inner, scalar_arg);
    }
    template <template <typename...> class RecorderType>
    auto &addBodyStateRecorder(SPHSystem &system) { return keep<RecorderType<ExecutionPolicy>>(system); }
    template <template <typename...> class IOType>
    auto &addIODynamics(SPHSystem &system) { return keep<IOType<ExecutionPolicy>>(system); }
    template <template <typename...> class RegressionType, class LocalReduceMethodType, typename... Args>
    auto &addReduceRegression(Args &&...args)
    {
        return keep<RegressionType<ReducedQuantityRecording<ExecutionPolicy, LocalReduceMethodType>>>(std::forward<Args>(args)...);
    }
    template <template <typename...> class RegressionType, typename DataType, typename... Args>
    auto &addObserveRegression(Args &&...args)
    {
        return keep<RegressionType<ObservedQuantityRecording<ExecutionPolicy, DataType>>>(std::forward<Args>(args)...);
    }

  private:
    template <class T, typename... Args>
    T &keep(Args &&...args)
    {
        auto owned = std::make_shared<T>(std::forward<Args>(args)...);
        T &ref = *owned;
        keeper_.push_back(std::move(owned));
        return ref;
    }
    std::vector<std::shared_ptr<void>> keeper_;
};

class TimeStepper // sph_solver.h:65-173, sph_solver.cpp:53-157
{
  public:
    explicit TimeStepper(SPHSystem &system) : sv_physical_time_(system.getSystemVariableByName<Real>("PhysicalTime")) {}
    class TriggerByInterval
    {
      public:
        explicit TriggerByInterval(Real initial_interval) : present_time_(Real(0)), interval_(initial_interval) {}
        bool operator()(BaseDynamics<Real> &interval_evaluator)
        {
            if (present_time_ > interval_)
            {
                present_time_ = Real(0);
                interval_ = interval_evaluator.exec();
                return true;
            }
            return false;
        }
        bool operator()()
        {
            if (present_time_ > interval_)
            {
                present_time_ = Real(0);
                return true;
            }
            return false;
        }
        Real getInterval() const { return interval_; }
        void incrementPresentTime(Real dt) { present_time_ += dt; }

      private:
        Real present_time_, interval_;
    };
    TriggerByInterval &addTriggerByInterval(Real initial_interval)
    {
        triggers_.push_back(std::make_unique<TriggerByInterval>(initial_interval));
        return *triggers_.back();
    }
    bool isEndTime(Real end_time) { return sv_physical_time_->getValue() >= end_time; }
    void setPhysicalTime(Real t) { sv_physical_time_->setValue(t); }
    Real getPhysicalTime() { return sv_physical_time_->getValue(); }
    Real getGlobalTimeStepSize() { return global_dt_; }
    Real incrementPhysicalTime(Real global_time_step)
    {
        global_dt_ = global_time_step;
        sv_physical_time_->incrementValue(global_dt_);
        for (auto &t : triggers_)
            t->incrementPresentTime(global_dt_);
        return global_dt_;
    }
    Real incrementPhysicalTime(BaseDynamics<Real> &step_evaluator) { return incrementPhysicalTime(step_evaluator.exec()); }

  private:
    SingleVariable<Real> *sv_physical_time_;
    Real global_dt_ = Real(0);
    std::vector<std::unique_ptr<TriggerByInterval>> triggers_;
};

class SPHSolver
{
  public:
    explicit SPHSolver(SPHSystem &system) : time_stepper_(system) {}
    template <class ExecutionPolicy>
    ParticleMethodContainer<ExecutionPolicy> &addParticleMethodContainer(const ExecutionPolicy &policy)
    {
        auto owned = std::make_shared<ParticleMethodContainer<ExecutionPolicy>>(policy);
        auto &ref = *owned;
        keeper_.push_back(std::move(owned));
        return ref;
    }
    TimeStepper &getTimeStepper() { return time_stepper_; }

  private:
    TimeStepper time_stepper_;
    std::vector<std::shared_ptr<void>> keeper_;
};
} // namespace SPH
