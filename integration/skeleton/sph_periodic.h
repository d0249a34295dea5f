// integration/skeleton (see sph_base.h): what tests/2d_examples/test_2d_taylor_green/taylor_green.cpp needs beyond the dam
// break - MultiPolygonShape, the Viscosity material property, FluidInitialCondition (base of the main's own host initial
// condition), ViscousForceInner, TransportVelocityCorrectionInner, the periodic condition and ReloadParticleIO.
// Declarations with the reference's constructor signatures; numerics come from the execution back end.
#pragma once
#include "sph_io.h"

namespace SPH
{
constexpr int xAxis = 0, yAxis = 1, zAxis = 2;
template <class T>
inline T SMIN(const T &a, const T &b) { return a < b ? a : b; }
template <class T>
inline T SMAX(const T &a, const T &b) { return a > b ? a : b; }

// geometries/multi_polygon_shape.h: boolean operations on closed polygons (2-D); add / sub of simple polygons here
enum class GeometricOps { add, sub };
class MultiPolygon
{
  public:
    void addPolygon(const std::vector<VecN<2>> &points, GeometricOps op) { parts_.emplace_back(points, op == GeometricOps::add); }
    bool checkContain(const VecN<2> &p) const
    {
        bool in = false;
        for (const auto &part : parts_)
        {
            const bool inside = contains(part.first, p);
            in = part.second ? (in || inside) : (in && !inside);
        }
        return in;
    }
    VecN<2> findClosestPoint(const VecN<2> &p) const
    {
        VecN<2> best = p;
        Real best_d = std::numeric_limits<Real>::max();
        for (const auto &part : parts_)
            for (size_t k = 0; k + 1 < part.first.size(); ++k)
            {
                const VecN<2> a = part.first[k], ab = part.first[k + 1] - a;
                const Real len2 = ab.dot(ab);
                const Real t = len2 > Real(0) ? std::min(std::max((p - a).dot(ab) / len2, Real(0)), Real(1)) : Real(0);
                const VecN<2> c = a + ab * t;
                const Real d = (p - c).norm();
                if (d < best_d)
                    best_d = d, best = c;
            }
        return best;
    }
    void bounds(VecN<2> &lo, VecN<2> &hi) const
    {
        bool first = true;
        for (const auto &part : parts_)
            for (const auto &q : part.first)
            {
                for (int d = 0; d < 2; ++d)
                {
                    lo[d] = first ? q[d] : std::min(lo[d], q[d]);
                    hi[d] = first ? q[d] : std::max(hi[d], q[d]);
                }
                first = false;
            }
    }

  private:
    static bool contains(const std::vector<VecN<2>> &poly, const VecN<2> &p)
    { // even-odd rule; points on the boundary count as inside (closed shape), as boost::geometry::covered_by does
        bool in = false;
        for (size_t i = 0, j = poly.size() - 1; i < poly.size(); j = i++)
        {
            const VecN<2> &a = poly[i], &b = poly[j];
            const Real cross = (b[0] - a[0]) * (p[1] - a[1]) - (b[1] - a[1]) * (p[0] - a[0]);
            if (cross == Real(0) && std::min(a[0], b[0]) <= p[0] && p[0] <= std::max(a[0], b[0]) && std::min(a[1], b[1]) <= p[1] &&
                p[1] <= std::max(a[1], b[1]))
                return true;
            if ((a[1] > p[1]) != (b[1] > p[1]) && p[0] < (b[0] - a[0]) * (p[1] - a[1]) / (b[1] - a[1]) + a[0])
                in = !in;
        }
        return in;
    }
    std::vector<std::pair<std::vector<VecN<2>>, bool>> parts_;
};
#if SPH_NDIM == 2
class MultiPolygonShape : public Shape
{
  public:
    explicit MultiPolygonShape(const std::string &name) : Shape(name) {}
    bool checkContain(const Vecd &p) override { return multi_polygon_.checkContain(p); }
    Vecd findClosestPoint(const Vecd &p) override { return multi_polygon_.findClosestPoint(p); }
    BoundingBoxd getBounds()
    {
        Vecd lo, hi;
        multi_polygon_.bounds(lo, hi);
        return BoundingBoxd(lo, hi);
    }

  protected:
    MultiPolygon multi_polygon_;
};
#endif

class Viscosity // materials/viscosity.h:40-67 (a material property: the constant only)
{
  public:
    explicit Viscosity(Real mu) : mu_(mu) {}
    Real ReferenceViscosity() const { return mu_; }

  private:
    Real mu_;
};
struct Reload {};
struct TruncatedLinear {};
struct AllParticles {};

namespace fluid_dynamics
{
class FluidInitialCondition : public LocalDynamics // base_fluid_dynamics.h:40-52
{
  public:
    explicit FluidInitialCondition(SPHBody &body)
        : LocalDynamics(body), pos_(particles_->getVariableDataByName<Vecd>("Position")), vel_(particles_->registerStateVariableData<Vecd>("Velocity"))
    {
    }

  protected:
    Vecd *pos_, *vel_;
};
class ViscousForceInner : public LocalDynamics // viscous_dynamics.h:43-53
{
  public:
    explicit ViscousForceInner(BaseInnerRelation &inner) : LocalDynamics(inner.getSPHBody()), inner_(inner) {}
    BaseInnerRelation &inner() { return inner_; }

  private:
    BaseInnerRelation &inner_;
};
template <class LimiterType, class ParticleScope>
class TransportVelocityCorrectionInner : public LocalDynamics // transport_velocity_correction.h:87-89
{
  public:
    explicit TransportVelocityCorrectionInner(BaseInnerRelation &inner, Real coefficient = Real(0.2))
        : LocalDynamics(inner.getSPHBody()), inner_(inner), coefficient_(coefficient)
    {
    }
    BaseInnerRelation &inner() { return inner_; }
    Real coefficient() const { return coefficient_; }

  private:
    BaseInnerRelation &inner_;
    Real coefficient_;
};
} // namespace fluid_dynamics

// general_dynamics/domian_bouding/domain_bounding.h:36-175
class PeriodicAlongAxis
{
  public:
    PeriodicAlongAxis(BoundingBoxd bounding_bounds, int axis) : bounds_(bounding_bounds), axis_(axis) {}
    int getAxis() const { return axis_; }
    const BoundingBoxd &getBoundingBox() const { return bounds_; }

  private:
    BoundingBoxd bounds_;
    int axis_;
};
// what a back end does for the three moments of the periodic condition
template <class ExecutionPolicy, class Enable = void>
struct DevicePeriodic
{
    static void declare(RealBody &, int, Real, Real) {}
    static void bounding(RealBody &, int) {}
    static void updateCellLinkedList(RealBody &, int) {}
};
template <class ExecutionPolicy = SPHINXSYS_DEFAULT_POLICY>
class PeriodicConditionUsingCellLinkedListT
{
    struct Step
    {
        RealBody *body;
        int axis;
        bool bounding;
        void exec(Real = Real(0))
        {
            if (bounding)
                DevicePeriodic<ExecutionPolicy>::bounding(*body, axis);
            else
                DevicePeriodic<ExecutionPolicy>::updateCellLinkedList(*body, axis);
        }
    };

  public:
    PeriodicConditionUsingCellLinkedListT(RealBody &body, PeriodicAlongAxis &periodic)
        : bounding_{&body, periodic.getAxis(), true}, update_cell_linked_list_{&body, periodic.getAxis(), false}
    {
        const int a = periodic.getAxis();
        DevicePeriodic<ExecutionPolicy>::declare(body, a, periodic.getBoundingBox().lower_[a], periodic.getBoundingBox().upper_[a]);
    }
    Step bounding_, update_cell_linked_list_;
};
using PeriodicConditionUsingCellLinkedList = PeriodicConditionUsingCellLinkedListT<>;

class ReloadParticleIO // io_base.h: writes the relaxed particle positions for a later Reload; the file is out of scope
{
  public:
    explicit ReloadParticleIO(SPHBody &) {}
    void writeToFile(size_t = 0) {}
};
} // namespace SPH
