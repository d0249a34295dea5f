// sphb200_adapter.hpp - header-only C++17 host side above the C ABI (include/sphb200.h).
//
// It mirrors the reference's user-facing Dynamics API for the hot path, name for name, so that a SPHinXsys main
// (tests/3d_examples/test_3d_dambreak/dambreak.cpp:66-195) reads the same with this header:
//
//   SPHSystem, SPHBody/FluidBody/SolidBody/ObserverBody, BaseParticles named variables
//                                   sphinxsys_system/sph_system.h:50-154, bodies/base_body.h:62-201,
//                                   particles/base_particles.h:119-162
//   InnerRelation, ContactRelation, ComplexRelation::updateConfiguration
//                                   body_relations/inner_body_relation.h, contact_body_relation.h, complex_body_relation.h
//   SimpleDynamics / ReduceDynamics / InteractionWithUpdate / Dynamics1Level <LocalDynamics>::exec(dt)
//                                   particle_dynamics/dynamics_algorithms.h:100-353, common/base_dynamics.h:42-63
//   local dynamics tags             fluid_dynamics/{density_summation,fluid_integration,fluid_time_step,
//                                   viscous_dynamics,transport_velocity_correction}.h, general_dynamics/{force_prior,
//                                   general_reduce,general_interpolation}.h, solid_dynamics/elastic_dynamics.h
//
// The reference's own headers cannot be compiled in this image (Eigen/TBB/Boost/Simbody absent), so this is a
// stand-alone mirror rather than a set of specialisations inside the reference tree; INTEGRATION.md shows the
// in-tree form.  Every exec() is one C-ABI call; there is no CPU implementation behind any of them.  Error
// behaviour follows the reference (print + exit(1)) unless SPHB200_ADAPTER_THROW is defined.
#pragma once
#include "sphb200.h"

#include <array>
#include <cstdio>
#include <cstdlib>
#include <initializer_list>
#include <stdexcept>
#include <string>
#include <utility>
#include <vector>

namespace SPH
{
#ifdef SPHINXSYS_USE_FLOAT // common/base_data_type.h:50-60
using Real = float;
#else
using Real = double;
#endif
#ifndef SPH_NDIM
#define SPH_NDIM 3
#endif
constexpr int Dimensions = SPH_NDIM;
using Vecd = std::array<Real, Dimensions>;
using BoundingBox = std::pair<Vecd, Vecd>; // first = lower, second = upper (base_data_type.h BoundingBox)

namespace detail
{
inline void check(sphb200_status s, sphb200_ctx *ctx, const char *what)
{
    if (s == SPHB200_OK)
        return;
    std::string msg = std::string(what) + ": " + (ctx ? sphb200_last_error(ctx) : "no context") +
                      (s == SPHB200_ERR_NO_DEVICE ? " (no CUDA device; there is no CPU fallback)" : "");
#ifdef SPHB200_ADAPTER_THROW
    throw std::runtime_error(msg);
#else
    std::fprintf(stderr, "\n Error: %s\n", msg.c_str()); // the reference's convention: print and exit(1)
    std::exit(1);
#endif
}
} // namespace detail

class SPHBody;
class SPHRelation;

// sph_system.h:50-154.  The user's bounds are expanded by 4 dp like sph_system.cpp:39.
class SPHSystem
{
  public:
    SPHSystem(const BoundingBox &system_domain_bounds, Real resolution_ref, int device = 0)
        : resolution_ref_(resolution_ref)
    {
        double lo[3] = {0, 0, 0}, hi[3] = {0, 0, 0};
        for (int d = 0; d < Dimensions; ++d)
        {
            lo[d] = double(system_domain_bounds.first[d]) - 4.0 * double(resolution_ref);
            hi[d] = double(system_domain_bounds.second[d]) + 4.0 * double(resolution_ref);
            bounds_.first[d] = Real(lo[d]);
            bounds_.second[d] = Real(hi[d]);
        }
        detail::check(sphb200_ctx_create(device, Dimensions, int(sizeof(Real)), lo, hi, &ctx_), ctx_, "SPHSystem");
    }
    ~SPHSystem()
    {
        if (ctx_)
            sphb200_ctx_destroy(ctx_);
    }
    SPHSystem(const SPHSystem &) = delete;
    SPHSystem &operator=(const SPHSystem &) = delete;

    Real ReferenceResolution() const { return resolution_ref_; }
    const BoundingBox &getSystemDomainBounds() const { return bounds_; }
    sphb200_ctx *handle() const { return ctx_; }
    void addBody(SPHBody *b) { bodies_.push_back(b); }
    void addRelation(SPHRelation *r) { relations_.push_back(r); }
    inline void initializeSystemCellLinkedLists(); // sph_system.cpp:58-66
    inline void initializeSystemConfigurations();  // sph_system.cpp:68-77
    void synchronize() { detail::check(sphb200_ctx_synchronize(ctx_), ctx_, "synchronize"); }
    long long kernelLaunches() const { return (long long)sphb200_ctx_launch_count(ctx_); }

  private:
    Real resolution_ref_;
    BoundingBox bounds_;
    sphb200_ctx *ctx_ = nullptr;
    std::vector<SPHBody *> bodies_;
    std::vector<SPHRelation *> relations_;
};

// materials: only the constants the path needs (weakly_compressible_fluid.h, elastic_solid.h, base_material.h)
struct WeaklyCompressibleFluid
{
    Real rho0, c0, mu;
    WeaklyCompressibleFluid(Real rho0_, Real c0_, Real mu_ = Real(0)) : rho0(rho0_), c0(c0_), mu(mu_) {}
};
struct Solid
{
    Real rho0 = Real(1);
};
struct NeoHookeanSolid
{
    Real rho0, youngs_modulus, poisson_ratio;
    NeoHookeanSolid(Real rho0_, Real E, Real nu) : rho0(rho0_), youngs_modulus(E), poisson_ratio(nu) {}
};

// SPHBody + its BaseParticles (bodies/base_body.h:62-201, particles/base_particles.h:80-263).  Variables are
// addressed by name exactly like registerStateVariable / getVariableDataByName; the device copy is the master
// between updateCellLinkedList() calls and arrays are handed over in the CURRENT particle order.
class SPHBody
{
  public:
    SPHBody(SPHSystem &system, const std::string &name, Real h_ratio = Real(1.3))
        : system_(system), name_(name), h_ratio_(h_ratio)
    {
        system.addBody(this);
    }
    virtual ~SPHBody() = default;
    SPHSystem &getSPHSystem() { return system_; }
    const std::string &getName() const { return name_; }
    sphb200_body *handle() const { return body_; }
    size_t TotalRealParticles() const { return n_; }
    Real getSPHBodyResolutionRef() const { return system_.ReferenceResolution(); }

    void defineMaterial(const WeaklyCompressibleFluid &m) { fluid_ = m, kind_ = 1; }
    void defineMaterial(const Solid &m) { fluid_ = WeaklyCompressibleFluid(m.rho0, Real(1)), kind_ = 2; }
    void defineMaterial(const NeoHookeanSolid &m) { elastic_ = m, kind_ = 3; }

    // generateParticles<BaseParticles, Lattice>() with the lattice evaluated by the caller (geometry is outside
    // the path): positions in reference order, one volume; rho = rho0, m = rho0 * Vol (base_material.cpp:35-41)
    void generateParticles(const std::vector<Vecd> &positions, Real volume)
    {
        n_ = positions.size();
        ck(sphb200_body_create(system_.handle(), name_.c_str(), (long long)n_, double(getSPHBodyResolutionRef()), double(h_ratio_), &body_),
           "generateParticles");
        Real rho0 = Real(1);
        if (kind_ == 1 || kind_ == 2)
        {
            ck(sphb200_body_set_fluid(body_, double(fluid_.rho0), double(fluid_.c0), double(fluid_.mu)), "defineMaterial");
            rho0 = fluid_.rho0;
        }
        else if (kind_ == 3)
        {
            ck(sphb200_body_set_elastic(body_, double(elastic_.rho0), double(elastic_.youngs_modulus), double(elastic_.poisson_ratio)),
               "defineMaterial");
            rho0 = elastic_.rho0;
        }
        setVariable("Position", positions);
        setVariable("VolumetricMeasure", std::vector<Real>(n_, volume));
        if (kind_ != 0)
        {
            setVariable("Mass", std::vector<Real>(n_, rho0 * volume));
            setVariable("Density", std::vector<Real>(n_, rho0));
        }
    }

    // registerStateVariable<DataType>(name): zero-initialised, carried along by the particle sort
    void registerStateVariable(const std::string &name, int width = 1, bool evolving = true)
    {
        ck(sphb200_var_register(body_, name.c_str(), width, evolving ? 1 : 0, nullptr), "registerStateVariable");
    }
    void setVariable(const std::string &name, const std::vector<Real> &values)
    {
        registerStateVariable(name, 1);
        if (n_)
            ck(sphb200_var_upload(body_, name.c_str(), values.data()), "setVariable");
    }
    void setVariable(const std::string &name, const std::vector<Vecd> &values)
    {
        registerStateVariable(name, Dimensions);
        if (n_)
            ck(sphb200_var_upload(body_, name.c_str(), values.data()), "setVariable");
    }
    // Double-buffered synchronizeToDevice: stageVariable starts the copy of a (pinned) host array on the library's
    // copy stream while the dynamics already queued keep running; commitStagedVariables swaps the copies in.
    // The host array must stay untouched in between; update the cell linked list and the relations afterwards.
    void stageVariable(const std::string &name, const Real *values)
    {
        if (n_)
            ck(sphb200_var_stage(body_, name.c_str(), values), "stageVariable");
    }
    void commitStagedVariables() { ck(sphb200_body_commit_staged(body_), "commitStagedVariables"); }
    // getVariableDataByName + synchronizeWithDevice: a host copy in the current particle order
    std::vector<Real> getVariableDataByName(const std::string &name)
    {
        int width = 1;
        ck(sphb200_var_width(body_, name.c_str(), &width), "getVariableDataByName");
        std::vector<Real> out(n_ * size_t(width));
        if (n_)
            ck(sphb200_var_download(body_, name.c_str(), out.data()), "getVariableDataByName");
        return out;
    }
    // OriginalID / SortedID (base_particles.cpp:35-38, particle_sorting.cpp:55-58)
    void getParticleIDs(std::vector<long long> &original_id, std::vector<long long> &sorted_id)
    {
        original_id.resize(n_), sorted_id.resize(n_);
        ck(sphb200_ids_to_host(body_, reinterpret_cast<int64_t *>(original_id.data()), reinterpret_cast<int64_t *>(sorted_id.data())),
           "getParticleIDs");
    }
    // RealBody::updateCellLinkedList (base_body.cpp:113-116); also keeps the body cell-sorted (ParticleSorting)
    void updateCellLinkedList() { ck(sphb200_cell_list_update(body_), "updateCellLinkedList"); }
    void ck(sphb200_status s, const char *what) { detail::check(s, system_.handle(), what); }

  protected:
    SPHSystem &system_;
    std::string name_;
    Real h_ratio_;
    sphb200_body *body_ = nullptr;
    size_t n_ = 0;
    int kind_ = 0;
    WeaklyCompressibleFluid fluid_{Real(1), Real(1)};
    NeoHookeanSolid elastic_{Real(1), Real(1), Real(0.3)};
};
using RealBody = SPHBody;
using FluidBody = SPHBody;
using SolidBody = SPHBody;
using ObserverBody = SPHBody;

// ParticleSorting<ExecutionPolicy>::exec (particle_sorting.hpp:15-21): the cell-list update already sorts
class ParticleSorting
{
  public:
    explicit ParticleSorting(RealBody &body) : body_(body) {}
    void exec() { body_.updateCellLinkedList(); }

  private:
    RealBody &body_;
};

// body_relations/base_body_relation.h: updateConfiguration() is the one virtual entry
class SPHRelation
{
  public:
    explicit SPHRelation(SPHBody &body) : body_(body) { body.getSPHSystem().addRelation(this); }
    virtual ~SPHRelation() = default;
    virtual void updateConfiguration() = 0;
    SPHBody &getSPHBody() { return body_; }

  protected:
    SPHBody &body_;
};
class InnerRelation : public SPHRelation
{
  public:
    explicit InnerRelation(RealBody &body) : SPHRelation(body)
    {
        body.ck(sphb200_relation_inner_create(body.handle(), &rel_), "InnerRelation");
    }
    void updateConfiguration() override { body_.ck(sphb200_relation_update(rel_), "InnerRelation::updateConfiguration"); }
    sphb200_relation *handle() const { return rel_; }

  private:
    sphb200_relation *rel_ = nullptr;
};
class ContactRelation : public SPHRelation
{
  public:
    ContactRelation(SPHBody &body, std::initializer_list<RealBody *> contact_bodies) : SPHRelation(body)
    {
        std::vector<sphb200_body *> t;
        for (RealBody *b : contact_bodies)
        {
            contact_bodies_.push_back(b);
            t.push_back(b->handle());
        }
        body.ck(sphb200_relation_contact_create(body.handle(), t.data(), int(t.size()), &rel_), "ContactRelation");
    }
    void updateConfiguration() override { body_.ck(sphb200_relation_update(rel_), "ContactRelation::updateConfiguration"); }
    sphb200_relation *handle() const { return rel_; }
    const std::vector<RealBody *> &contactBodies() const { return contact_bodies_; }

  private:
    sphb200_relation *rel_ = nullptr;
    std::vector<RealBody *> contact_bodies_;
};
class ComplexRelation : public SPHRelation // complex_body_relation.h: inner + contact updated together
{
  public:
    ComplexRelation(InnerRelation &inner, ContactRelation &contact) : SPHRelation(inner.getSPHBody()), inner_(inner), contact_(contact) {}
    void updateConfiguration() override
    {
        inner_.updateConfiguration();
        contact_.updateConfiguration();
    }

  private:
    InnerRelation &inner_;
    ContactRelation &contact_;
};

inline void SPHSystem::initializeSystemCellLinkedLists()
{
    for (SPHBody *b : bodies_)
        if (b->handle())
            b->updateCellLinkedList();
}
inline void SPHSystem::initializeSystemConfigurations()
{
    for (SPHRelation *r : relations_)
        if (dynamic_cast<ComplexRelation *>(r) == nullptr) // its parts are registered themselves
            r->updateConfiguration();
}

// ---------------------------------------------------------------------------------------------------------
// Local dynamics (tags) and the algorithm templates that execute them.  Like in the reference the algorithm
// template takes the local dynamics as its parameter and forwards the constructor arguments
// (dynamics_algorithms.h:100-353); exec(dt) of a hot local dynamics is ONE call into the library.
// ---------------------------------------------------------------------------------------------------------
struct Gravity // external_force/external_force.h:38-57
{
    Vecd g;
    explicit Gravity(const Vecd &g_) : g(g_) {}
};
template <class ForceType>
struct GravityForce // general_dynamics/force_prior.hpp:28-41
{
};
namespace fluid_dynamics
{
struct DensitySummationComplex {};            // density_summation.h: Inner<> + Contact<>
struct DensitySummationComplexFreeSurface {}; // ... Inner<FreeSurface> + Contact<>
struct DensitySummationInner {};
struct Integration1stHalfWithWallRiemann {};  // fluid_integration.h aliases
struct Integration2ndHalfWithWallRiemann {};
struct Integration1stHalfWithWallNoRiemann {};
struct Integration2ndHalfWithWallNoRiemann {};
struct Integration1stHalfInnerRiemann {};
struct Integration2ndHalfInnerRiemann {};
struct Integration1stHalfInnerNoRiemann {};
struct Integration2ndHalfInnerNoRiemann {};
struct AdvectionTimeStep {};        // fluid_time_step.h
struct AdvectionViscousTimeStep {};
struct AcousticTimeStep {};
struct ViscousForceInner {};        // viscous_dynamics.h
struct TransportVelocityCorrectionInner {}; // transport_velocity_correction.h
} // namespace fluid_dynamics
namespace solid_dynamics
{
struct Integration1stHalfPK2 {};          // elastic_dynamics.h:259-287
struct DecomposedIntegration1stHalf {};   // elastic_dynamics.h:325-351
struct Integration2ndHalf {};
struct AcousticTimeStep {};
} // namespace solid_dynamics
struct LinearGradientCorrectionMatrixInner {}; // general_dynamics/kernel_correction.h
struct TotalKineticEnergy {};                  // general_dynamics/general_reduce.h
struct TotalMechanicalEnergy {};
struct MaximumSpeed {};

namespace detail
{
template <class T>
struct IntegrationTraits;
#define SPHB200_INTEGRATION(NAME, HALF, WALL, RIEMANN)               \
    template <>                                                      \
    struct IntegrationTraits<fluid_dynamics::NAME>                   \
    {                                                                \
        static constexpr int half = HALF;                            \
        static constexpr bool wall = WALL, riemann = RIEMANN;        \
    };
SPHB200_INTEGRATION(Integration1stHalfWithWallRiemann, 1, true, true)
SPHB200_INTEGRATION(Integration2ndHalfWithWallRiemann, 2, true, true)
SPHB200_INTEGRATION(Integration1stHalfWithWallNoRiemann, 1, true, false)
SPHB200_INTEGRATION(Integration2ndHalfWithWallNoRiemann, 2, true, false)
SPHB200_INTEGRATION(Integration1stHalfInnerRiemann, 1, false, true)
SPHB200_INTEGRATION(Integration2ndHalfInnerRiemann, 2, false, true)
SPHB200_INTEGRATION(Integration1stHalfInnerNoRiemann, 1, false, false)
SPHB200_INTEGRATION(Integration2ndHalfInnerNoRiemann, 2, false, false)
#undef SPHB200_INTEGRATION
} // namespace detail

// common/base_dynamics.h:42-63
template <class ReturnType = void>
class BaseDynamics
{
  public:
    virtual ~BaseDynamics() = default;
    virtual ReturnType exec(Real dt = Real(0)) = 0;
};

// ---- Dynamics1Level<Integration{1st,2nd}Half...> (dynamics_algorithms.h:289-353) ----
template <class LocalDynamics>
class Dynamics1Level : public BaseDynamics<void>
{
    using T = detail::IntegrationTraits<LocalDynamics>;

  public:
    Dynamics1Level(InnerRelation &inner, ContactRelation &wall) : inner_(inner), wall_(&wall) { static_assert(T::wall, "use the inner-only ctor"); }
    explicit Dynamics1Level(InnerRelation &inner) : inner_(inner), wall_(nullptr) { static_assert(!T::wall, "wall contact relation required"); }
    void exec(Real dt = Real(0)) override
    {
        sphb200_relation *w = wall_ ? wall_->handle() : nullptr;
        if (T::half == 1)
            inner_.getSPHBody().ck(sphb200_integration_1st_half(inner_.handle(), w, T::riemann ? 1 : 0, double(dt)), "Integration1stHalf");
        else
            inner_.getSPHBody().ck(sphb200_integration_2nd_half(inner_.handle(), w, T::riemann ? 1 : 0, double(dt)), "Integration2ndHalf");
    }

  private:
    InnerRelation &inner_;
    ContactRelation *wall_;
};
template <>
class Dynamics1Level<solid_dynamics::Integration1stHalfPK2> : public BaseDynamics<void>
{
  public:
    explicit Dynamics1Level(InnerRelation &inner) : inner_(inner) {}
    void exec(Real dt = Real(0)) override { inner_.getSPHBody().ck(sphb200_solid_1st_half(inner_.handle(), 0, double(dt)), "Integration1stHalfPK2"); }

  private:
    InnerRelation &inner_;
};
template <>
class Dynamics1Level<solid_dynamics::DecomposedIntegration1stHalf> : public BaseDynamics<void>
{
  public:
    explicit Dynamics1Level(InnerRelation &inner) : inner_(inner) {}
    void exec(Real dt = Real(0)) override { inner_.getSPHBody().ck(sphb200_solid_1st_half(inner_.handle(), 1, double(dt)), "DecomposedIntegration1stHalf"); }

  private:
    InnerRelation &inner_;
};
template <>
class Dynamics1Level<solid_dynamics::Integration2ndHalf> : public BaseDynamics<void>
{
  public:
    explicit Dynamics1Level(InnerRelation &inner) : inner_(inner) {}
    void exec(Real dt = Real(0)) override { inner_.getSPHBody().ck(sphb200_solid_2nd_half(inner_.handle(), double(dt)), "solid Integration2ndHalf"); }

  private:
    InnerRelation &inner_;
};

// ---- InteractionWithUpdate<...> (dynamics_algorithms.h:254-287) ----
template <class LocalDynamics>
class InteractionWithUpdate;
template <>
class InteractionWithUpdate<fluid_dynamics::DensitySummationComplexFreeSurface> : public BaseDynamics<void>
{
  public:
    InteractionWithUpdate(InnerRelation &inner, ContactRelation &contact) : inner_(inner), contact_(contact) {}
    void exec(Real = Real(0)) override { inner_.getSPHBody().ck(sphb200_density_summation(inner_.handle(), contact_.handle(), 1), "DensitySummation"); }

  private:
    InnerRelation &inner_;
    ContactRelation &contact_;
};
template <>
class InteractionWithUpdate<fluid_dynamics::DensitySummationComplex> : public BaseDynamics<void>
{
  public:
    InteractionWithUpdate(InnerRelation &inner, ContactRelation &contact) : inner_(inner), contact_(contact) {}
    void exec(Real = Real(0)) override { inner_.getSPHBody().ck(sphb200_density_summation(inner_.handle(), contact_.handle(), 0), "DensitySummation"); }

  private:
    InnerRelation &inner_;
    ContactRelation &contact_;
};
template <>
class InteractionWithUpdate<fluid_dynamics::DensitySummationInner> : public BaseDynamics<void>
{
  public:
    explicit InteractionWithUpdate(InnerRelation &inner) : inner_(inner) {}
    void exec(Real = Real(0)) override { inner_.getSPHBody().ck(sphb200_density_summation(inner_.handle(), nullptr, 0), "DensitySummation"); }

  private:
    InnerRelation &inner_;
};
template <>
class InteractionWithUpdate<fluid_dynamics::ViscousForceInner> : public BaseDynamics<void>
{
  public:
    explicit InteractionWithUpdate(InnerRelation &inner) : inner_(inner) {}
    void exec(Real = Real(0)) override { inner_.getSPHBody().ck(sphb200_viscous_force(inner_.handle(), nullptr), "ViscousForce"); }

  private:
    InnerRelation &inner_;
};
template <>
class InteractionWithUpdate<fluid_dynamics::TransportVelocityCorrectionInner> : public BaseDynamics<void>
{
  public:
    explicit InteractionWithUpdate(InnerRelation &inner, Real coefficient = Real(0.2)) : inner_(inner), coefficient_(coefficient) {}
    void exec(Real = Real(0)) override
    {
        inner_.getSPHBody().ck(sphb200_transport_velocity_correction(inner_.handle(), double(coefficient_)), "TransportVelocityCorrection");
    }

  private:
    InnerRelation &inner_;
    Real coefficient_;
};
template <>
class InteractionWithUpdate<LinearGradientCorrectionMatrixInner> : public BaseDynamics<void>
{
  public:
    explicit InteractionWithUpdate(InnerRelation &inner) : inner_(inner) {}
    void exec(Real = Real(0)) override { inner_.getSPHBody().ck(sphb200_correction_matrix(inner_.handle()), "LinearGradientCorrectionMatrix"); }

  private:
    InnerRelation &inner_;
};

// ---- SimpleDynamics<GravityForce<Gravity>> (dynamics_algorithms.h:100-125) ----
template <class LocalDynamics>
class SimpleDynamics;
template <>
class SimpleDynamics<GravityForce<Gravity>> : public BaseDynamics<void>
{
  public:
    SimpleDynamics(SPHBody &body, const Gravity &gravity) : body_(body), gravity_(gravity) {}
    void exec(Real = Real(0)) override
    {
        double g[3] = {0, 0, 0};
        for (int d = 0; d < Dimensions; ++d)
            g[d] = double(gravity_.g[d]);
        body_.ck(sphb200_gravity_force(body_.handle(), g), "GravityForce");
    }

  private:
    SPHBody &body_;
    Gravity gravity_;
};

// ---- ReduceDynamics<...> (dynamics_algorithms.h:127-160): exec() returns outputResult(reduced value) ----
template <class LocalDynamics>
class ReduceDynamics;
template <>
class ReduceDynamics<fluid_dynamics::AdvectionTimeStep> : public BaseDynamics<Real>
{
  public:
    ReduceDynamics(SPHBody &body, Real U_ref, Real advectionCFL = Real(0.25)) : body_(body), U_ref_(U_ref), cfl_(advectionCFL) {}
    Real exec(Real = Real(0)) override
    {
        double out = 0;
        body_.ck(sphb200_advection_dt(body_.handle(), double(U_ref_), double(cfl_), 0, &out), "AdvectionTimeStep");
        return Real(out);
    }

  private:
    SPHBody &body_;
    Real U_ref_, cfl_;
};
template <>
class ReduceDynamics<fluid_dynamics::AdvectionViscousTimeStep> : public BaseDynamics<Real>
{
  public:
    ReduceDynamics(SPHBody &body, Real U_ref, Real advectionCFL = Real(0.25)) : body_(body), U_ref_(U_ref), cfl_(advectionCFL) {}
    Real exec(Real = Real(0)) override
    {
        double out = 0;
        body_.ck(sphb200_advection_dt(body_.handle(), double(U_ref_), double(cfl_), 1, &out), "AdvectionViscousTimeStep");
        return Real(out);
    }

  private:
    SPHBody &body_;
    Real U_ref_, cfl_;
};
template <>
class ReduceDynamics<fluid_dynamics::AcousticTimeStep> : public BaseDynamics<Real>
{
  public:
    explicit ReduceDynamics(SPHBody &body, Real acousticCFL = Real(0.6)) : body_(body), cfl_(acousticCFL) {}
    Real exec(Real = Real(0)) override
    {
        double out = 0;
        body_.ck(sphb200_acoustic_dt(body_.handle(), double(cfl_), &out), "AcousticTimeStep");
        return Real(out);
    }

  private:
    SPHBody &body_;
    Real cfl_;
};
template <>
class ReduceDynamics<solid_dynamics::AcousticTimeStep> : public BaseDynamics<Real>
{
  public:
    explicit ReduceDynamics(SPHBody &body, Real CFL = Real(0.6)) : body_(body), cfl_(CFL) {}
    Real exec(Real = Real(0)) override
    {
        double out = 0;
        body_.ck(sphb200_solid_dt(body_.handle(), double(cfl_), &out), "solid AcousticTimeStep");
        return Real(out);
    }

  private:
    SPHBody &body_;
    Real cfl_;
};
template <>
class ReduceDynamics<TotalMechanicalEnergy> : public BaseDynamics<Real>
{
  public:
    ReduceDynamics(SPHBody &body, const Gravity &gravity) : body_(body), gravity_(gravity) {}
    Real exec(Real = Real(0)) override
    {
        double g[3] = {0, 0, 0}, out = 0;
        for (int d = 0; d < Dimensions; ++d)
            g[d] = double(gravity_.g[d]);
        body_.ck(sphb200_total_mechanical_energy(body_.handle(), g, &out), "TotalMechanicalEnergy");
        return Real(out);
    }

  private:
    SPHBody &body_;
    Gravity gravity_;
};
template <>
class ReduceDynamics<TotalKineticEnergy> : public BaseDynamics<Real>
{
  public:
    explicit ReduceDynamics(SPHBody &body) : body_(body) {}
    Real exec(Real = Real(0)) override
    {
        double out = 0;
        body_.ck(sphb200_total_kinetic_energy(body_.handle(), &out), "TotalKineticEnergy");
        return Real(out);
    }

  private:
    SPHBody &body_;
};
template <>
class ReduceDynamics<MaximumSpeed> : public BaseDynamics<Real>
{
  public:
    explicit ReduceDynamics(SPHBody &body) : body_(body) {}
    Real exec(Real = Real(0)) override
    {
        double out = 0;
        body_.ck(sphb200_maximum_speed(body_.handle(), &out), "MaximumSpeed");
        return Real(out);
    }

  private:
    SPHBody &body_;
};

// ObservedQuantityRecording<Real>: kernel interpolation of a fluid variable at the observer particles
// (general_interpolation.h:63-83, io_observation.h:56-120) - the values, not the file output
class ObservingAQuantity
{
  public:
    ObservingAQuantity(ContactRelation &observer_contact, const std::string &variable_name, int width = 1)
        : contact_(observer_contact), name_(variable_name), width_(width) {}
    std::vector<double> exec()
    {
        std::vector<double> out(contact_.getSPHBody().TotalRealParticles() * size_t(width_));
        contact_.getSPHBody().ck(sphb200_observe(contact_.handle(), name_.c_str(), width_, out.data()), "ObservingAQuantity");
        return out;
    }

  private:
    ContactRelation &contact_;
    std::string name_;
    int width_;
};
} // namespace SPH
