// integration/skeleton (see sph_base.h): what the other two reference mains of the path need beyond the dam break -
// the total-Lagrangian solid dynamics of tests/3d_examples/test_3d_twisting_column/twisting_column.cpp (declarations with
// the reference's constructor signatures, no numerics; a user-defined initial condition runs through the generic host
// fall-back) and RestartIO of tests/2d_examples/test_2d_dambreak/Dambreak.cpp.
#pragma once
#include "sph_io.h"

namespace SPH
{
// body parts (bodies/body_part_by_particle: BodyRegionByParticle tags the particles inside a shape, by ORIGINAL id)
class BodyPartByParticle
{
  public:
    explicit BodyPartByParticle(SPHBody &body) : sph_body_(body) {}
    SPHBody &getSPHBody() { return sph_body_; }
    const std::vector<long long> &OriginalIds() const { return ids_; }

  protected:
    SPHBody &sph_body_;
    std::vector<long long> ids_;
};
class BodyRegionByParticle : public BodyPartByParticle
{
  public:
    BodyRegionByParticle(SPHBody &body, Shape &shape) : BodyPartByParticle(body)
    { // constructed before the first cell-list update: host order = original order
        BaseParticles &p = body.getBaseParticles();
        Vecd *pos = p.ParticlePositions();
        for (size_t i = 0; i < p.TotalRealParticles(); ++i)
            if (shape.checkContain(pos[i]))
                ids_.push_back((long long)i);
    }
};
class FixBodyPartConstraint : public LocalDynamics // general_dynamics/general_constraint.h:106-121
{
  public:
    explicit FixBodyPartConstraint(BodyPartByParticle &part) : LocalDynamics(part.getSPHBody()), part_(part)
    { // registerStateVariableDataFrom<Vecd>("InitialPosition", "Position"), general_constraint.h:96
        if (!particles_->getVariableByName<Vecd>("InitialPosition"))
        {
            Vecd *pos = particles_->ParticlePositions(), *pos0 = particles_->registerStateVariableData<Vecd>("InitialPosition");
            for (size_t i = 0; i < particles_->TotalRealParticles(); ++i)
                pos0[i] = pos[i];
        }
        particles_->registerStateVariable<Vecd>("Velocity");
    }
    BodyPartByParticle &part() { return part_; }

  private:
    BodyPartByParticle &part_;
};
class LinearGradientCorrectionMatrixInner : public LocalDynamics // general_dynamics/kernel_correction.cpp:8-29
{
  public:
    explicit LinearGradientCorrectionMatrixInner(BaseInnerRelation &inner) : LocalDynamics(inner.getSPHBody()), inner_(inner) {}
    BaseInnerRelation &inner() { return inner_; }

  private:
    BaseInnerRelation &inner_;
};

namespace solid_dynamics
{
// elastic_dynamics.h:57-70: the base of user-defined initial conditions - pos_ and vel_ as the reference names them
class ElasticDynamicsInitialCondition : public LocalDynamics
{
  public:
    explicit ElasticDynamicsInitialCondition(SPHBody &body)
        : LocalDynamics(body), pos_(particles_->getVariableDataByName<Vecd>("Position")), vel_(particles_->registerStateVariableData<Vecd>("Velocity"))
    {
    }

  protected:
    Vecd *pos_, *vel_;
};
class SolidIntegrationBase : public LocalDynamics
{
  public:
    explicit SolidIntegrationBase(BaseInnerRelation &inner) : LocalDynamics(inner.getSPHBody()), inner_(inner)
    {
        particles_->registerStateVariable<Vecd>("Velocity");
    }
    BaseInnerRelation &inner() { return inner_; }

  private:
    BaseInnerRelation &inner_;
};
class DecomposedIntegration1stHalf : public SolidIntegrationBase // elastic_dynamics.h:259-287
{
  public:
    static constexpr int half = 1;
    static constexpr bool decomposed = true;
    using SolidIntegrationBase::SolidIntegrationBase;
};
class Integration1stHalfPK2 : public SolidIntegrationBase // elastic_dynamics.h:217-236
{
  public:
    static constexpr int half = 1;
    static constexpr bool decomposed = false;
    using SolidIntegrationBase::SolidIntegrationBase;
};
class Integration2ndHalf : public SolidIntegrationBase // elastic_dynamics.h:325-351
{
  public:
    static constexpr int half = 2;
    static constexpr bool decomposed = false;
    using SolidIntegrationBase::SolidIntegrationBase;
};
class AcousticTimeStep : public LocalDynamics // elastic_dynamics.cpp:12-30 (ReduceMin)
{
  public:
    using ReturnType = Real;
    explicit AcousticTimeStep(SPHBody &body, Real CFL = Real(0.6)) : LocalDynamics(body), CFL_(CFL) {}
    Real CFL_;
    std::string quantity_name_ = "AcousticTimeStep";
};
} // namespace solid_dynamics

// how a back end moves a body's evolving state to and from the host for a restart (io_base.cpp:67-97 writes the evolving
// variables of every real body; the XML container is out of scope, the hand-over is the point)
template <class ExecutionPolicy, class Enable = void>
struct DeviceRestart
{
    static void toHost(SPHBody &) {}
    static void toDevice(SPHBody &) {}
};
template <class ExecutionPolicy = SPHINXSYS_DEFAULT_POLICY>
class RestartIOT
{
  public:
    explicit RestartIOT(SPHSystem &system) : system_(system) { ::mkdir("restart", 0755); }
    void writeToFile(size_t iteration_step)
    {
        std::ofstream out(path(iteration_step), std::ios::binary);
        const double t = double(*system_.template getSystemVariableDataByName<Real>("PhysicalTime"));
        out.write(reinterpret_cast<const char *>(&t), sizeof(t));
        for (RealBody *b : system_.getRealBodies())
        {
            DeviceRestart<ExecutionPolicy>::toHost(*b);
            forEachEvolving(*b, [&](void *data, size_t bytes) { out.write(static_cast<const char *>(data), std::streamsize(bytes)); });
        }
    }
    void readRestartFiles(size_t restart_step)
    {
        std::ifstream in(path(restart_step), std::ios::binary);
        if (!in)
        {
            std::cout << "\n Error: the restart file '" << path(restart_step) << "' does not exist!" << std::endl;
            std::exit(1);
        }
        double t = 0;
        in.read(reinterpret_cast<char *>(&t), sizeof(t));
        *system_.template getSystemVariableDataByName<Real>("PhysicalTime") = Real(t);
        for (RealBody *b : system_.getRealBodies())
        {
            forEachEvolving(*b, [&](void *data, size_t bytes) { in.read(static_cast<char *>(data), std::streamsize(bytes)); });
            DeviceRestart<ExecutionPolicy>::toDevice(*b);
        }
    }

  private:
    static std::string path(size_t step) { return "restart/Restart_" + std::to_string(step) + ".bin"; }
    template <class F>
    static void forEachEvolving(SPHBody &body, const F &f)
    {
        BaseParticles &p = body.getBaseParticles();
        const size_t n = p.TotalRealParticles();
        for (const std::string &name : p.VariableNames())
        {
            if (auto *r = p.getVariableByName<Real>(name))
            {
                if (r->isEvolving())
                    f(r->Data(), n * sizeof(Real));
            }
            else if (auto *v = p.getVariableByName<Vecd>(name))
            {
                if (v->isEvolving())
                    f(v->Data(), n * sizeof(Vecd));
            }
        }
    }
    SPHSystem &system_;
};
using RestartIO = RestartIOT<>;
} // namespace SPH
