// integration/skeleton (see sph_ck.h): the CK local dynamics of tests/tests_sycl/3d_examples/test_3d_dambreak_sycl/dambreak.cpp
// - constructor signatures of the reference, no numerics - and the CK forms of the recorders.
#pragma once
#include "sph_ck.h"

namespace SPH
{
// general_dynamics/force_prior_ck.h:66-90
template <class GravityType>
class GravityForceCK : public LocalDynamics
{
  public:
    GravityForceCK(SPHBody &body, const GravityType &gravity) : LocalDynamics(body), gravity_(gravity)
    {
        particles_->registerStateVariable<Vecd>("ForcePrior");
        particles_->addEvolvingVariable<Vecd>("ForcePrior");
        particles_->addEvolvingVariable<Vecd>("PreviousGravityForce");
    }
    const GravityType &gravity() const { return gravity_; }

  private:
    GravityType gravity_;
};
// general_dynamics/geometric_dynamics: a host kernel in the reference main too ("run on CPU")
class NormalFromBodyShapeCK : public LocalDynamics
{
  public:
    explicit NormalFromBodyShapeCK(SPHBody &body)
        : LocalDynamics(body), shape_(body.getInitialShape()), pos_(particles_->getVariableDataByName<Vecd>("Position")),
          n_(particles_->registerStateVariableData<Vecd>("NormalDirection"))
    {
    }
    void update(size_t i, Real = Real(0)) { n_[i] = shape_.findNormalDirection(pos_[i]); }

  private:
    Shape &shape_;
    Vecd *pos_, *n_;
};
// general_dynamics/kernel_correction_ck.h: LinearCorrectionMatrixComplex(DynamicsArgs(inner, alpha), contact)
class LinearCorrectionMatrixComplex : public LocalDynamics
{
  public:
    template <class InnerArgs>
    LinearCorrectionMatrixComplex(InnerArgs inner_args, Contact<> &contact)
        : LocalDynamics(inner_args.identifier_.getSPHBody()), inner_(inner_args.identifier_), contact_(contact), alpha_(Real(std::get<0>(inner_args.others_)))
    {
    }
    Inner<> &inner() { return inner_; }
    Contact<> &contact() { return contact_; }
    Real alpha() const { return alpha_; }

  private:
    Inner<> &inner_;
    Contact<> &contact_;
    Real alpha_;
};
// every complex interaction below is constructed as (inner, contact)
class ComplexInteractionCK : public LocalDynamics
{
  public:
    ComplexInteractionCK(Inner<> &inner, Contact<> &contact) : LocalDynamics(inner.getSPHBody()), inner_(inner), contact_(contact) {}
    Inner<> &inner() { return inner_; }
    Contact<> &contact() { return contact_; }

  private:
    Inner<> &inner_;
    Contact<> &contact_;
};
class TotalMechanicalEnergyCK : public LocalDynamics // general_reduce_ck.h
{
  public:
    using ReturnType = Real;
    TotalMechanicalEnergyCK(SPHBody &body, Gravity &gravity) : LocalDynamics(body), gravity_(gravity) {}
    Gravity &gravity_;
    std::string quantity_name_ = "TotalMechanicalEnergy";
};

namespace fluid_dynamics
{
class AdvectionStepSetup : public LocalDynamics // fluid_time_step_ck.h:126-160
{
  public:
    explicit AdvectionStepSetup(SPHBody &body) : LocalDynamics(body) { particles_->registerStateVariable<Vecd>("Displacement"); }
};
class UpdateParticlePosition : public LocalDynamics // fluid_time_step_ck.h:162-185
{
  public:
    explicit UpdateParticlePosition(SPHBody &body) : LocalDynamics(body) { particles_->registerStateVariable<Vecd>("Displacement"); }
};
class AcousticStepBaseCK : public ComplexInteractionCK // AcousticStep ctor, acoustic_step_1st_half.hpp:13-39
{
  public:
    AcousticStepBaseCK(Inner<> &inner, Contact<> &contact) : ComplexInteractionCK(inner, contact)
    {
        particles_->registerStateVariable<Real>("Pressure");
        const bool fresh = particles_->getVariableByName<Real>("Compression") == nullptr;
        Real *compression = particles_->registerStateVariableData<Real>("Compression"); // registerStateVariable<Real>("Compression", Real(1))
        for (size_t i = 0; fresh && i < particles_->TotalRealParticles(); ++i)
            compression[i] = Real(1);
        particles_->registerStateVariable<Real>("CompressionRate");
        particles_->registerStateVariable<Vecd>("Velocity");
        particles_->registerStateVariable<Vecd>("Force");
        particles_->registerStateVariable<Vecd>("ForcePrior");
        particles_->addEvolvingVariable<Vecd>("Velocity");
        particles_->addEvolvingVariable<Real>("Mass");
        particles_->addEvolvingVariable<Vecd>("ForcePrior");
        particles_->addEvolvingVariable<Real>("Compression");
        particles_->addEvolvingVariable<Real>("CompressionRate");
        particles_->addVariableToWrite<Vecd>("Velocity");
        for (RealBody *w : contact.getContactBodies())
        { // Interaction<Wall>, interaction_ck.hpp:80-90: the wall's averaged velocity / acceleration and its normal
            w->getBaseParticles().registerStateVariable<Vecd>("Velocity");
            w->getBaseParticles().registerStateVariable<Vecd>("Acceleration");
            w->getBaseParticles().registerStateVariable<Vecd>("NormalDirection");
        }
    }
};
class AcousticStep1stHalfWithWallRiemannCorrectionCK : public AcousticStepBaseCK // acoustic_step_1st_half.h aliases
{
  public:
    using AcousticStepBaseCK::AcousticStepBaseCK;
};
class AcousticStep2ndHalfWithWallRiemannCorrectionCK : public AcousticStepBaseCK
{
  public:
    using AcousticStepBaseCK::AcousticStepBaseCK;
};
template <typename...>
class CompressionSummation;
template <>
class CompressionSummation<Inner<>, Contact<>> : public ComplexInteractionCK // density_regularization.h:45-106
{
  public:
    using ComplexInteractionCK::ComplexInteractionCK;
};
template <class DynamicsIdentifier, class FluidType, class FlowType>
class DensityRegularization : public LocalDynamics // density_regularization.h:113-144
{
  public:
    static constexpr bool free_surface = std::is_same<FlowType, FreeSurface>::value;
    explicit DensityRegularization(SPHBody &body) : LocalDynamics(body) {}
};
class FreeSurfaceIndicationComplexSpatialTemporalCK : public ComplexInteractionCK // surface_indication_ck.h
{
  public:
    using ComplexInteractionCK::ComplexInteractionCK;
};
class AdvectionTimeStepCK : public LocalDynamics // fluid_time_step_ck.h:82-124
{
  public:
    using ReturnType = Real;
    AdvectionTimeStepCK(SPHBody &body, Real U_ref, Real advectionCFL = Real(0.25)) : LocalDynamics(body), speed_ref_(U_ref), advectionCFL_(advectionCFL) {}
    Real speed_ref_, advectionCFL_;
    std::string quantity_name_ = "AdvectionTimeStep";
};
template <class FluidType>
class AcousticTimeStepCK : public LocalDynamics // fluid_time_step_ck.h:44-80
{
  public:
    using ReturnType = Real;
    explicit AcousticTimeStepCK(SPHBody &body, Real acousticCFL = Real(0.6)) : LocalDynamics(body), acousticCFL_(acousticCFL) {}
    Real acousticCFL_;
    std::string quantity_name_ = "AcousticTimeStep";
};
} // namespace fluid_dynamics

// ---- CK forms of the recorders: <ExecutionPolicy, Method> and (contact, name)
template <class ExecutionPolicy, class LocalReduceMethodType>
class ReducedQuantityRecording<ExecutionPolicy, LocalReduceMethodType> : public BaseQuantityRecording
{
  public:
    template <class DynamicsIdentifier, typename... Args>
    ReducedQuantityRecording(DynamicsIdentifier &identifier, Args &&...args)
        : BaseQuantityRecording(identifier.getSPHBody().getSPHSystem(), identifier.getSPHBody().getName()),
          reduce_method_(identifier, std::forward<Args>(args)...)
    {
        quantity_name_ = reduce_method_.QuantityName();
    }
    virtual void writeToFile(size_t = 0)
    {
        const double t = double(*system_.template getSystemVariableDataByName<Real>("PhysicalTime"));
        rows_.push_back({t, double(reduce_method_.exec())});
    }

  protected:
    ReduceDynamicsCK<ExecutionPolicy, LocalReduceMethodType> reduce_method_;
    std::string quantity_name_;
    std::vector<std::vector<double>> rows_;
};
template <class ExecutionPolicy>
class ObservedQuantityRecording<ExecutionPolicy, Real> : public BaseQuantityRecording
{
  public:
    ObservedQuantityRecording(Contact<> &contact_relation, const std::string &quantity_name)
        : BaseQuantityRecording(contact_relation.getSPHBody().getSPHSystem(), contact_relation.getSPHBody().getName()), contact_(contact_relation),
          quantity_name_(quantity_name)
    {
    }
    virtual void writeToFile(size_t = 0)
    {
        using Obs = DeviceObservation<ExecutionPolicy>;
        static_assert(Obs::available, "observer interpolation needs an execution back end");
        std::vector<double> row{double(*system_.template getSystemVariableDataByName<Real>("PhysicalTime"))};
        Obs::observe(contact_, quantity_name_, 1, row);
        rows_.push_back(std::move(row));
    }

  protected:
    Contact<> &contact_;
    std::string quantity_name_;
    std::vector<std::vector<double>> rows_;
};
template <class ExecutionPolicy>
class BodyStatesRecordingToVtpCK : public BodyStatesRecordingT<ExecutionPolicy>
{
  public:
    using BodyStatesRecordingT<ExecutionPolicy>::BodyStatesRecordingT;
};
} // namespace SPH
