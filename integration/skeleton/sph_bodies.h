// integration/skeleton (see sph_base.h): execution policies, variables, particles, bodies, relations.
#pragma once
#include "sph_base.h"

namespace SPH
{
// ---------------------------------------------------------------------------------------------------------
// particle_dynamics/execution/execution_policy.h:37-83 - policy tags; a device policy derives from DeviceExecution<>
// ---------------------------------------------------------------------------------------------------------
namespace execution
{
class SequencedPolicy {};
class ParallelPolicy {};
template <typename...>
class DeviceExecution;
template <>
class DeviceExecution<> {};
template <typename PolicyType>
class DeviceExecution<PolicyType> : public DeviceExecution<>, public PolicyType {};
using ParallelDevicePolicy = DeviceExecution<ParallelPolicy>;
inline constexpr auto seq = SequencedPolicy{};
inline constexpr auto par_host = ParallelPolicy{};
inline constexpr auto par_device = ParallelDevicePolicy{};
template <class Policy>
inline constexpr bool is_device_policy = std::is_base_of<DeviceExecution<>, Policy>::value;
} // namespace execution
using namespace execution;

// ---------------------------------------------------------------------------------------------------------
// common/sphinxsys_variable.h:130-378 - SingleVariable / DiscreteVariable with the device delegation points.
// DelegatedData(policy) hands out the array a kernel of that policy works on; for a device policy that is
// DelegatedOnDevice(), whose definition (with synchronizeWithDevice / synchronizeToDevice) a device back end
// supplies - src_sycl/shared/common/sphinxsys_variable_sycl.hpp:35-87 for SYCL, integration/src_cuda for CUDA.
// ---------------------------------------------------------------------------------------------------------
class BaseVariable
{
  public:
    explicit BaseVariable(const std::string &name) : name_(name) {}
    virtual ~BaseVariable() = default;
    const std::string &Name() const { return name_; }

  protected:
    std::string name_;
};

template <typename DataType>
class SingleVariable : public BaseVariable
{
  public:
    SingleVariable(const std::string &name, const DataType &value) : BaseVariable(name), value_(value) {}
    DataType *Data() { return &value_; }
    DataType getValue() const { return value_; }
    void setValue(const DataType &v) { value_ = v; }
    void incrementValue(const DataType &v) { value_ += v; }

  private:
    DataType value_;
};

struct DeviceBinding; // what a device back end attaches to a variable (integration/src_cuda/sph_cuda.h)

template <typename DataType>
class DiscreteVariable : public BaseVariable
{
  public:
    DiscreteVariable(const std::string &name, size_t data_size) : BaseVariable(name), data_size_(data_size), data_field_(new DataType[data_size]())
    {
    }
    ~DiscreteVariable() { delete[] data_field_; }
    DataType *Data() { return data_field_; }
    size_t getDataSize() const { return data_size_; }
    template <class Policy>
    DataType *DelegatedData(const Policy &)
    {
        if constexpr (is_device_policy<Policy>)
            return DelegatedOnDevice();
        else
            return data_field_;
    }
    void reallocateData(size_t n) // sphinxsys_variable.h:368-373: the owner re-registers with the device afterwards
    {
        delete[] data_field_;
        data_field_ = new DataType[n]();
        data_size_ = n;
    }
    // ---- customisation points of a device back end (declared here, defined by the back end) ----
    DataType *DelegatedOnDevice();
    void synchronizeWithDevice(); // device -> host
    void synchronizeToDevice();   // host -> device
    void bindDevice(std::shared_ptr<DeviceBinding> b) { device_ = std::move(b); }
    DeviceBinding *deviceBinding() { return device_.get(); }
    bool isEvolving() const { return evolving_; }
    void setEvolving() { evolving_ = true; }

  private:
    size_t data_size_;
    DataType *data_field_; // host array, owned here (new T[]), borrowed by the device back end
    bool evolving_ = false;
    std::shared_ptr<DeviceBinding> device_;
};

// ---------------------------------------------------------------------------------------------------------
// particles/base_particles.h:80-263 - the named-variable registry of a body
// ---------------------------------------------------------------------------------------------------------
class SPHBody;
class BaseParticles
{
  public:
    explicit BaseParticles(SPHBody &body) : sph_body_(body) {}
    SPHBody &getSPHBody() { return sph_body_; }
    size_t TotalRealParticles() const { return total_real_particles_; }
    void initializeAllParticlesBounds(size_t n) { total_real_particles_ = n; }

    template <typename DataType>
    DiscreteVariable<DataType> *registerStateVariable(const std::string &name)
    {
        auto *existing = getVariableByName<DataType>(name);
        if (existing)
            return existing;
        auto v = std::make_unique<DiscreteVariable<DataType>>(name, std::max<size_t>(total_real_particles_, 1));
        auto *raw = v.get();
        variables_[name] = std::move(v);
        order_.push_back(name);
        onVariableRegistered(raw, name, int(sizeof(DataType) / sizeof(Real)));
        return raw;
    }
    template <typename DataType>
    DataType *registerStateVariableData(const std::string &name) { return registerStateVariable<DataType>(name)->Data(); }
    template <typename DataType>
    DiscreteVariable<DataType> *getVariableByName(const std::string &name)
    {
        auto it = variables_.find(name);
        return it == variables_.end() ? nullptr : dynamic_cast<DiscreteVariable<DataType> *>(it->second.get());
    }
    template <typename DataType>
    DataType *getVariableDataByName(const std::string &name)
    {
        auto *v = getVariableByName<DataType>(name);
        if (!v)
        {
            std::cout << "\n Error: the variable '" << name << "' is not registered!" << std::endl;
            std::exit(1);
        }
        return v->Data();
    }
    template <typename DataType>
    void addEvolvingVariable(const std::string &name)
    {
        registerStateVariable<DataType>(name)->setEvolving();
        onEvolving(name);
    }
    template <typename DataType>
    void addVariableToWrite(const std::string &name) { to_write_.push_back(name); }
    Vecd *ParticlePositions() { return getVariableDataByName<Vecd>("Position"); }
    const std::vector<std::string> &VariablesToWrite() const { return to_write_; }
    const std::vector<std::string> &VariableNames() const { return order_; }
    BaseVariable *findVariable(const std::string &name)
    {
        auto it = variables_.find(name);
        return it == variables_.end() ? nullptr : it->second.get();
    }
    // hooks a device back end installs (the reference's registerDiscreteVariable would call the back end directly)
    std::function<void(BaseVariable *, const std::string &, int)> on_register_;
    std::function<void(const std::string &)> on_evolving_;
    std::function<void(bool to_host)> sync_all_; // generic fall-back of non-hot local dynamics

  private:
    void onVariableRegistered(BaseVariable *v, const std::string &name, int width)
    {
        if (on_register_)
            on_register_(v, name, width);
    }
    void onEvolving(const std::string &name)
    {
        if (on_evolving_)
            on_evolving_(name);
    }
    SPHBody &sph_body_;
    size_t total_real_particles_ = 0;
    std::map<std::string, std::unique_ptr<BaseVariable>> variables_;
    std::vector<std::string> order_, to_write_;
};

// ---------------------------------------------------------------------------------------------------------
// materials (base_material.h, weakly_compressible_fluid.h): constants only
// ---------------------------------------------------------------------------------------------------------
class BaseMaterial
{
  public:
    explicit BaseMaterial(Real rho0 = Real(1)) : rho0_(rho0) {}
    virtual ~BaseMaterial() = default;
    Real ReferenceDensity() const { return rho0_; }

  protected:
    Real rho0_;
};
class Fluid : public BaseMaterial
{
  public:
    Fluid(Real rho0, Real c0) : BaseMaterial(rho0), c0_(c0) {}
    Real ReferenceSoundSpeed() const { return c0_; }

  protected:
    Real c0_;
};
class WeaklyCompressibleFluid : public Fluid
{
  public:
    WeaklyCompressibleFluid(Real rho0, Real c0) : Fluid(rho0, c0) {}
};
class Solid : public BaseMaterial
{
  public:
    explicit Solid(Real rho0 = Real(1)) : BaseMaterial(rho0) {}
};
class NeoHookeanSolid : public Solid // materials/elastic_solid.h:120-140 (constants only)
{
  public:
    NeoHookeanSolid(Real rho0, Real youngs_modulus, Real poisson_ratio) : Solid(rho0), E_(youngs_modulus), nu_(poisson_ratio) {}
    Real YoungsModulus() const { return E_; }
    Real PoissonRatio() const { return nu_; }

  private:
    Real E_, nu_;
};

// adaptations/adaptation.h: spacing and smoothing-length ratio
class SPHAdaptation
{
  public:
    SPHAdaptation(Real resolution_ref, Real h_spacing_ratio = Real(1.3)) : spacing_ref_(resolution_ref), h_spacing_ratio_(h_spacing_ratio) {}
    Real ReferenceSpacing() const { return spacing_ref_; }
    Real MinimumSpacing() const { return spacing_ref_; }
    Real SmoothingLengthRatio() const { return h_spacing_ratio_; }
    Real ReferenceSmoothingLength() const { return h_spacing_ratio_ * spacing_ref_; }

  private:
    Real spacing_ref_, h_spacing_ratio_;
};

// ---------------------------------------------------------------------------------------------------------
// sphinxsys_system/sph_system.h:50-154
// ---------------------------------------------------------------------------------------------------------
class RealBody;
class SPHRelation;
class SPHSystem
{
  public:
    SPHSystem(BoundingBoxd system_domain_bounds, Real global_resolution)
        : system_bounds_(system_domain_bounds.expand(global_resolution * 4)), global_resolution_(global_resolution),
          physical_time_("PhysicalTime", Real(0))
    {
    }
    void handleCommandlineOptions(int ac, char *av[])
    { // sph_system.cpp:100-190: --state_recording, --regression, --restart_step (others are accepted and ignored here)
        for (int k = 1; k < ac; ++k)
        {
            const std::string a(av[k]);
            auto value = [&](const char *key) -> std::string
            {
                const std::string p = std::string("--") + key + "=";
                return a.compare(0, p.size(), p) == 0 ? a.substr(p.size()) : std::string();
            };
            std::string v;
            if (!(v = value("state_recording")).empty())
                state_recording_ = (v == "1" || v == "true" || v == "True");
            if (!(v = value("regression")).empty())
                generate_regression_data_ = (v == "1" || v == "true" || v == "True");
            if (!(v = value("restart_step")).empty())
                restart_step_ = size_t(std::atol(v.c_str()));
        }
    }
    const BoundingBoxd &getSystemDomainBounds() const { return system_bounds_; }
    Real GlobalResolution() const { return global_resolution_; }
    bool GenerateRegressionData() const { return generate_regression_data_; }
    void setReloadParticles(bool v) { reload_particles_ = v; }
    bool ReloadParticles() const { return reload_particles_; }
    bool StateRecording() const { return state_recording_; }
    size_t RestartStep() const { return restart_step_; }
    template <typename DataType>
    DataType *getSystemVariableDataByName(const std::string &) { return physical_time_.Data(); }
    template <typename DataType>
    SingleVariable<DataType> *getSystemVariableByName(const std::string &) { return &physical_time_; }
    void addRealBody(RealBody *b) { real_bodies_.push_back(b); }
    const std::vector<RealBody *> &getRealBodies() const { return real_bodies_; }
    void addRelation(SPHRelation *r) { relations_.push_back(r); }
    inline void initializeSystemCellLinkedLists(); // sph_system.cpp:58-66
    inline void initializeSystemConfigurations();  // sph_system.cpp:68-77
    void *device_context_ = nullptr;               // owned by the device back end

  private:
    BoundingBoxd system_bounds_;
    Real global_resolution_;
    SingleVariable<Real> physical_time_;
    bool generate_regression_data_ = false, state_recording_ = true, reload_particles_ = false;
    size_t restart_step_ = 0;
    std::vector<RealBody *> real_bodies_;
    std::vector<SPHRelation *> relations_;
};

// ---------------------------------------------------------------------------------------------------------
// bodies/base_body.h:62-201, predefined_bodies.h:36-70; particle generators as tags
// ---------------------------------------------------------------------------------------------------------
struct Lattice {};
struct ObserverParticles {};
class SPHBody;
inline void attachExecutionBackEnd(SPHBody &body); // defined by the execution back end (no-op for the host policies)

class SPHBody
{
  public:
    SPHBody(SPHSystem &system, Shape &shape) : system_(system), shape_(&shape), name_(shape.getName()) { init(); }
    SPHBody(SPHSystem &system, SharedPtr<Shape> shape) : system_(system), shape_keeper_(shape), shape_(shape.get()), name_(shape->getName()) { init(); }
    SPHBody(SPHSystem &system, const std::string &name) : system_(system), shape_(nullptr), name_(name) { init(); }
    virtual ~SPHBody() = default;
    SPHSystem &getSPHSystem() { return system_; }
    SPHBody &getSPHBody() { return *this; }
    const std::string &getName() const { return name_; }
    const std::string &Name() const { return name_; }
    Shape &getInitialShape() { return *shape_; }
    SPHAdaptation &getSPHAdaptation() { return *adaptation_; }
    BaseParticles &getBaseParticles() { return *particles_; }
    BaseMaterial &getBaseMaterial() { return *material_; }
    BoundingBoxd getSPHSystemBounds() { return system_.getSystemDomainBounds(); }
    BoundingBoxd getSPHBodyBounds() { return shape_->getBounds(); }
    // addMaterialProperty<Viscosity>(mu) (base_body.h): the one property on the path is the dynamic viscosity
    template <class PropertyType, typename... Args>
    void addMaterialProperty(Args &&...args) { viscosity_ = Real(PropertyType(std::forward<Args>(args)...).ReferenceViscosity()); }
    Real ReferenceViscosity() const { return viscosity_; }
    template <class AdaptationType, typename... Args>
    void defineAdaptation(Args &&...args)
    {
        adaptation_ = std::make_unique<AdaptationType>(system_.GlobalResolution(), std::forward<Args>(args)...);
    }
    template <class MaterialType, typename... Args>
    MaterialType *defineMatterMaterial(Args &&...args)
    {
        auto m = std::make_unique<MaterialType>(std::forward<Args>(args)...);
        auto *raw = m.get();
        material_ = std::move(m);
        return raw;
    }
    // generateParticles<BaseParticles, Lattice>() (particle_generator_lattice_3d.cpp:12-25) and
    // generateParticles<ObserverParticles>(positions)
    template <class ParticleType, class Method>
    void generateParticles()
    {
        static_assert(std::is_same<Method, Lattice>::value, "only the lattice generator is on the path");
        const Real dp = adaptation_->MinimumSpacing();
        const BoundingBoxd box = getSPHSystemBounds();
        int n[3] = {1, 1, 1};
        for (int d = 0; d < Dimensions; ++d)
            n[d] = int(std::ceil((box.upper_[d] - box.lower_[d]) / dp));
        std::vector<Vecd> pos;
        int c[3];
        for (c[0] = 0; c[0] < n[0]; ++c[0])
            for (c[1] = 0; c[1] < n[1]; ++c[1])
                for (c[2] = 0; c[2] < n[2]; ++c[2])
                {
                    Vecd p;
                    for (int d = 0; d < Dimensions; ++d)
                        p[d] = (box.lower_[d] + Real(c[d]) * dp) + Real(0.5) * dp;
                    if (shape_->checkContain(p))
                        pos.push_back(p);
                }
        Real vol = dp;
        for (int d = 1; d < Dimensions; ++d)
            vol *= dp;
        createParticles(pos, vol);
    }
    template <class ParticleType>
    void generateParticles(const StdVec<Vecd> &positions) { createParticles(positions, Real(0)); }
    // generateParticles<BaseParticles, Reload>(name): relaxed positions from a reload file - out of scope here
    template <class ParticleType, class Method>
    void generateParticles(const std::string &)
    {
        std::cout << "\n Error: particle reload files are not supported by the integration skeleton." << std::endl;
        std::exit(1);
    }
    void *device_body_ = nullptr; // owned by the device back end
    bool device_cell_list_ready_ = false;

  protected:
    void init()
    {
        adaptation_ = std::make_unique<SPHAdaptation>(system_.GlobalResolution());
        material_ = std::make_unique<BaseMaterial>();
        particles_ = std::make_unique<BaseParticles>(*this);
    }
    inline void createParticles(const std::vector<Vecd> &pos, Real vol);
    SPHSystem &system_;
    SharedPtr<Shape> shape_keeper_;
    Shape *shape_;
    std::string name_;
    std::unique_ptr<SPHAdaptation> adaptation_;
    std::unique_ptr<BaseMaterial> material_;
    std::unique_ptr<BaseParticles> particles_;
    Real viscosity_ = Real(0);
};

class RealBody : public SPHBody
{
  public:
    template <typename... Args>
    RealBody(SPHSystem &system, Args &&...args) : SPHBody(system, std::forward<Args>(args)...) { system.addRealBody(this); }
    inline void updateCellLinkedList(); // base_body.cpp:113-116 - defined by the execution back end
};
class FluidBody : public RealBody
{
  public:
    template <typename... Args>
    FluidBody(Args &&...args) : RealBody(std::forward<Args>(args)...) {}
};
class SolidBody : public RealBody
{
  public:
    template <typename... Args>
    SolidBody(Args &&...args) : RealBody(std::forward<Args>(args)...) { defineAdaptation<SPHAdaptation>(Real(1.15)); } // predefined_bodies.h:44-54
};
class ObserverBody : public SPHBody
{
  public:
    template <typename... Args>
    ObserverBody(Args &&...args) : SPHBody(std::forward<Args>(args)...) {}
};

// ---------------------------------------------------------------------------------------------------------
// body_relations/*.h: updateConfiguration() is the one virtual entry; a back end fills in the work
// ---------------------------------------------------------------------------------------------------------
class SPHRelation
{
  public:
    explicit SPHRelation(SPHBody &body) : sph_body_(body) { body.getSPHSystem().addRelation(this); }
    virtual ~SPHRelation() = default;
    SPHBody &getSPHBody() { return sph_body_; }
    virtual void updateConfiguration() = 0;
    virtual bool isComposite() const { return false; }
    void *device_relation_ = nullptr; // owned by the device back end

  protected:
    SPHBody &sph_body_;
};
class BaseInnerRelation : public SPHRelation
{
  public:
    explicit BaseInnerRelation(RealBody &body) : SPHRelation(body), real_body_(body) {}
    RealBody &getRealBody() { return real_body_; }

  protected:
    RealBody &real_body_;
};
class InnerRelation : public BaseInnerRelation
{
  public:
    explicit InnerRelation(RealBody &body) : BaseInnerRelation(body) {}
    inline void updateConfiguration() override;
};
class BaseContactRelation : public SPHRelation
{
  public:
    BaseContactRelation(SPHBody &body, StdVec<RealBody *> contact_bodies) : SPHRelation(body), contact_bodies_(std::move(contact_bodies)) {}
    const StdVec<RealBody *> &getContactBodies() const { return contact_bodies_; }

  protected:
    StdVec<RealBody *> contact_bodies_;
};
class ContactRelation : public BaseContactRelation
{
  public:
    ContactRelation(SPHBody &body, StdVec<RealBody *> contact_bodies) : BaseContactRelation(body, std::move(contact_bodies)) {}
    inline void updateConfiguration() override;
};
class ComplexRelation : public SPHRelation
{
  public:
    ComplexRelation(BaseInnerRelation &inner, BaseContactRelation &contact) : SPHRelation(inner.getSPHBody()), inner_(inner), contact_(contact) {}
    void updateConfiguration() override
    {
        inner_.updateConfiguration();
        contact_.updateConfiguration();
    }
    bool isComposite() const override { return true; }

  private:
    BaseInnerRelation &inner_;
    BaseContactRelation &contact_;
};

inline void SPHSystem::initializeSystemCellLinkedLists()
{
    for (RealBody *b : real_bodies_)
        b->updateCellLinkedList();
}
inline void SPHSystem::initializeSystemConfigurations()
{
    for (SPHRelation *r : relations_)
        if (!r->isComposite())
            r->updateConfiguration();
}

inline void SPHBody::createParticles(const std::vector<Vecd> &pos, Real vol)
{
    particles_->initializeAllParticlesBounds(pos.size());
    attachExecutionBackEnd(*this); // a device back end hooks the variable registry before the first variable registers
    Vecd *x = particles_->registerStateVariableData<Vecd>("Position");
    Real *V = particles_->registerStateVariableData<Real>("VolumetricMeasure");
    for (size_t i = 0; i < pos.size(); ++i)
        x[i] = pos[i], V[i] = vol;
    particles_->addEvolvingVariable<Vecd>("Position");
    particles_->addEvolvingVariable<Real>("VolumetricMeasure");
    if (vol > Real(0))
    { // base_material.cpp:35-41: rho = rho0, m = rho0 * Vol
        Real *rho = particles_->registerStateVariableData<Real>("Density");
        Real *m = particles_->registerStateVariableData<Real>("Mass");
        for (size_t i = 0; i < pos.size(); ++i)
            rho[i] = material_->ReferenceDensity(), m[i] = material_->ReferenceDensity() * vol;
    }
}
} // namespace SPH
