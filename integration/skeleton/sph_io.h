// integration/skeleton (see sph_base.h): observation / state recording / regression recording
// (io_system/io_observation.h:56-170, io_vtk.h, regression_test/dynamic_time_warping_method.h:68-135).
// The recorders collect the series; testResult() / generateDataBase() write them to ./output/<body>_<quantity>.dat
// in the reference's layout ("run_time" then the values per line) so a test can compare them - the DTW comparator
// itself is restated in oracle/validate_golden.py, not here.
#pragma once
#include "sph_dynamics.h"
#include <sys/stat.h>

namespace SPH
{
// how a back end interpolates at observer particles (general_interpolation.h:63-83) and reads variables for output
template <class ExecutionPolicy, class Enable = void>
struct DeviceObservation
{
    static constexpr bool available = false;
};

class BaseQuantityRecording
{
  public:
    BaseQuantityRecording(SPHSystem &system, const std::string &body_name) : system_(system), body_name_(body_name) {}
    virtual ~BaseQuantityRecording() = default;

  protected:
    void dump(const std::string &quantity, const std::vector<std::vector<double>> &rows)
    {
        ::mkdir("output", 0755);
        std::ofstream out("output/" + body_name_ + "_" + quantity + ".dat");
        out.precision(17);
        out << "\"run_time\"   \"" << quantity << "\"\n";
        for (const auto &r : rows)
        {
            for (double v : r)
                out << v << "   ";
            out << "\n";
        }
    }
    SPHSystem &system_;
    std::string body_name_;
};

// The reference declares the recorders as `template <typename...> class X;` with one specialisation for the classic
// form X<Method> and one for the CK form X<ExecutionPolicy, Method> (io_observation.h, io_observation_ck.h)
template <typename...>
class ReducedQuantityRecording;
template <typename...>
class ObservedQuantityRecording;
template <class>
struct DefaultPolicyFor // keeps the policy a dependent type: the back end's specialisations are defined after this header
{
    using type = SPHINXSYS_DEFAULT_POLICY;
};

template <class LocalReduceMethodType>
class ReducedQuantityRecording<LocalReduceMethodType> : public BaseQuantityRecording // io_observation.h:123-170
{
    using ExecutionPolicy = typename DefaultPolicyFor<LocalReduceMethodType>::type;

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
        const double t = double(*system_.getSystemVariableDataByName<Real>("PhysicalTime"));
        rows_.push_back({t, double(reduce_method_.exec())});
    }

  protected:
    ReduceDynamics<LocalReduceMethodType, ExecutionPolicy> reduce_method_;
    std::string quantity_name_;
    std::vector<std::vector<double>> rows_;
};

template <typename DataType>
class ObservedQuantityRecording<DataType> : public BaseQuantityRecording // io_observation.h:56-120
{
    using ExecutionPolicy = typename DefaultPolicyFor<DataType>::type;

  public:
    ObservedQuantityRecording(const std::string &quantity_name, BaseContactRelation &contact_relation)
        : BaseQuantityRecording(contact_relation.getSPHBody().getSPHSystem(), contact_relation.getSPHBody().getName()),
          contact_(contact_relation), quantity_name_(quantity_name)
    {
    }
    virtual void writeToFile(size_t = 0)
    {
        using Obs = DeviceObservation<ExecutionPolicy>;
        static_assert(Obs::available, "observer interpolation needs an execution back end");
        std::vector<double> row{double(*system_.getSystemVariableDataByName<Real>("PhysicalTime"))};
        Obs::observe(contact_, quantity_name_, int(sizeof(DataType) / sizeof(Real)), row);
        rows_.push_back(std::move(row));
    }

  protected:
    BaseContactRelation &contact_;
    std::string quantity_name_;
    std::vector<std::vector<double>> rows_;
};

template <class ObserveMethodType>
class RegressionTestDynamicTimeWarping : public ObserveMethodType // dynamic_time_warping_method.h:68-135
{
  public:
    template <typename... Args>
    explicit RegressionTestDynamicTimeWarping(Args &&...args) : ObserveMethodType(std::forward<Args>(args)...) {}
    void generateDataBase(Real, const std::string & = "false") { this->dump(this->quantity_name_, this->rows_); }
    void testResult(const std::string & = "false") { this->dump(this->quantity_name_, this->rows_); }
};

// io_vtk.h BodyStatesRecordingToVtp: the variables to write come back through the back end's asynchronous
// output path (BaseVariable::prepareForOutput(policy), sphinxsys_variable.h:350-356); the file itself is out of scope
template <class ExecutionPolicy = SPHINXSYS_DEFAULT_POLICY>
class BodyStatesRecordingT
{
  public:
    explicit BodyStatesRecordingT(SPHSystem &system) : system_(system) {}
    template <typename DataType>
    void addToWrite(SPHBody &body, const std::string &name)
    {
        body.getBaseParticles().template addVariableToWrite<DataType>(name);
        bodies_.push_back(&body);
    }
    void addBody(SPHBody &body) { bodies_.push_back(&body); }
    void writeToFile(size_t = 0)
    {
        if (!system_.StateRecording())
            return;
        using Obs = DeviceObservation<ExecutionPolicy>;
        if constexpr (Obs::available)
            for (SPHBody *b : bodies_)
                Obs::prepareForOutput(*b); // asynchronous D2H of Position + the variables to write
        ++frames_;
    }
    size_t framesWritten() const { return frames_; }

  private:
    SPHSystem &system_;
    std::vector<SPHBody *> bodies_;
    size_t frames_ = 0;
};
using BodyStatesRecordingToVtp = BodyStatesRecordingT<>;
} // namespace SPH
