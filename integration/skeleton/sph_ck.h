// integration/skeleton (see sph_base.h): the CK ("computing kernel") form of the API that the reference's device-ready
// mains use (tests/tests_sycl/...): relations Inner<> / Contact<>, the algorithm templates with the execution policy as
// FIRST template argument - StateDynamics, InteractionDynamicsCK, ReduceDynamicsCK, UpdateCellLinkedList, UpdateRelation,
// ParticleSortCK (shared_ck/particle_dynamics/simple_algorithms_ck.h:41-121, interaction_algorithms_ck.h:91-261,
// configuration_dynamics/*.h) - and the local dynamics test_3d_dambreak_sycl/dambreak.cpp:113-150 constructs, as
// declarations with the reference's constructor signatures.  No numerics: a device back end supplies
// DeviceImplementation<ExecutionPolicy, LocalDynamicsType> (sph_dynamics.h), the same customisation point as for the
// classic templates.
#pragma once
#include "sph_io.h"
#include <tuple>

namespace SPH
{
// the reference's sphinxsys.h aliases MainExecutionPolicy to ParallelDevicePolicy in a SYCL build, ParallelPolicy otherwise
using MainExecutionPolicy = SPHINXSYS_DEFAULT_POLICY;

struct FreeSurface {};
struct Internal {};

// ---- relations (shared_ck/body_relation/relation_ck.h): Inner<> / Contact<> are relation classes AND template tags
template <typename...>
class Inner;
template <>
class Inner<> : public InnerRelation
{
  public:
    explicit Inner(RealBody &body) : InnerRelation(body) {}
};
template <typename...>
class Contact;
template <>
class Contact<> : public ContactRelation
{
  public:
    Contact(SPHBody &body, StdVec<RealBody *> contact_bodies) : ContactRelation(body, std::move(contact_bodies)) {}
};

// DynamicsArgs(identifier, others...): constructor arguments bundled for one part of a complex interaction
template <class Identifier, typename... Others>
struct DynamicsArgs
{
    DynamicsArgs(Identifier &identifier, Others... others) : identifier_(identifier), others_(others...) {}
    Identifier &identifier_;
    std::tuple<Others...> others_;
};

namespace detail
{
template <class ExecutionPolicy, class L>
inline void execLocal(L &local, Real dt, const char *what)
{
    using Impl = DeviceImplementation<ExecutionPolicy, L>;
    if constexpr (Impl::available)
        Impl::run(local, dt);
    else if constexpr (has_update<L>::value)
    { // a host kernel (or the generic fall-back of a device policy): device -> host, loop, host -> device
        hostFallbackBegin(local);
        particle_for(execution::par_host, local.getParticles()->TotalRealParticles(), [&](size_t i) { local.update(i, dt); });
        hostFallbackEnd(local);
    }
    else
        noHostNumerics(what);
}
} // namespace detail

// ---- algorithm templates, execution policy first
template <class ExecutionPolicy, class LocalDynamicsType>
class StateDynamics : public LocalDynamicsType, public BaseDynamics<void> // simple_algorithms_ck.h:41-75
{
  public:
    template <typename... Args>
    StateDynamics(Args &&...args) : LocalDynamicsType(std::forward<Args>(args)...) {}
    void exec(Real dt = Real(0)) override { detail::execLocal<ExecutionPolicy>(static_cast<LocalDynamicsType &>(*this), dt, "StateDynamics<...>"); }
};
template <class ExecutionPolicy, class LocalDynamicsType>
class InteractionDynamicsCK : public LocalDynamicsType, public BaseDynamics<void> // interaction_algorithms_ck.h:91-261
{
  public:
    template <typename... Args>
    InteractionDynamicsCK(Args &&...args) : LocalDynamicsType(std::forward<Args>(args)...) {}
    void exec(Real dt = Real(0)) override
    {
        detail::execLocal<ExecutionPolicy>(static_cast<LocalDynamicsType &>(*this), dt, "InteractionDynamicsCK<...>");
    }
};
template <class ExecutionPolicy, class LocalDynamicsType>
class ReduceDynamicsCK : public LocalDynamicsType, public BaseDynamics<typename LocalDynamicsType::ReturnType> // simple_algorithms_ck.h:77-121
{
  public:
    using ReturnType = typename LocalDynamicsType::ReturnType;
    template <typename... Args>
    ReduceDynamicsCK(Args &&...args) : LocalDynamicsType(std::forward<Args>(args)...) {}
    std::string QuantityName() { return this->quantity_name_; }
    ReturnType exec(Real dt = Real(0)) override
    {
        using Impl = DeviceImplementation<ExecutionPolicy, LocalDynamicsType>;
        if constexpr (Impl::available)
            return Impl::run(static_cast<LocalDynamicsType &>(*this), dt);
        else
            detail::noHostNumerics("ReduceDynamicsCK<...>");
    }
};

// ---- spatial operators (configuration_dynamics/update_cell_linked_list.h, update_body_relation.h, particle_sort_ck.h)
template <class ExecutionPolicy, class BodyType>
class UpdateCellLinkedList
{
  public:
    explicit UpdateCellLinkedList(RealBody &body) : body_(body) {}
    void exec(Real = Real(0)) { body_.updateCellLinkedList(); }

  private:
    RealBody &body_;
};
template <class ExecutionPolicy, typename... RelationTypes>
class UpdateRelation
{
  public:
    explicit UpdateRelation(RelationTypes &...relations) : relations_(relations...) {}
    void exec(Real = Real(0))
    {
        std::apply([](auto &...r) { (r.updateConfiguration(), ...); }, relations_);
    }

  private:
    std::tuple<RelationTypes &...> relations_;
};
template <class ExecutionPolicy>
class ParticleSortCK
{
  public:
    explicit ParticleSortCK(RealBody &body) : body_(body) {}
    void exec(Real = Real(0)) { body_.updateCellLinkedList(); } // the cell-list update of a device back end keeps the body sorted

  private:
    RealBody &body_;
};
} // namespace SPH
#include "sph_ck_dynamics.h"
