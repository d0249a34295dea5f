// integration/skeleton - a compile target for the CUDA execution policy (integration/src_cuda).
//
// The reference cannot be built in this image (Eigen, oneTBB, Boost, Simbody absent), so its class SKELETON for the
// hot path is restated here, in our own words and only as far as the path and tests/3d_examples/test_3d_dambreak/
// dambreak.cpp need it: same class names, template signatures, constructor signatures and customisation points as
// the reference headers cited at each block, bodies kept to the host-side bookkeeping.  It is NOT a copy of the
// reference's sources and carries none of its numerics: the pair sums live in libsphb200 only.  With it the
// reference's dam-break main compiles UNMODIFIED (from /root/reference, by integration/Makefile) on top of the
// CUDA policy - the drop-in statement of SURVEY section 8b made checkable.
//
// This file: scalar / vector types, bounding box, shapes (base_data_type.h, geometries/*).
#pragma once
#include <algorithm>
#include <array>
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <fstream>
#include <functional>
#include <initializer_list>
#include <iomanip>
#include <iostream>
#include <map>
#include <memory>
#include <string>
#include <type_traits>
#include <utility>
#include <vector>

namespace SPH
{
#ifdef SPHINXSYS_USE_FLOAT // common/base_data_type.h:50-60
using Real = float;
#else
using Real = double;
#endif
using UnsignedInt = size_t;
#ifndef SPH_NDIM
#define SPH_NDIM 3
#endif
constexpr int Dimensions = SPH_NDIM;
constexpr Real Eps = std::numeric_limits<Real>::epsilon();
constexpr Real Pi = Real(M_PI);

// stand-in for the fixed-size Eigen vectors of base_data_type.h (only what the mains and the skeleton use)
template <int N>
struct VecN
{
    Real v[N];
    VecN()
    {
        for (int k = 0; k < N; ++k)
            v[k] = Real(0);
    }
    VecN(Real a, Real b)
    {
        static_assert(N == 2, "two components");
        v[0] = a, v[1] = b;
    }
    VecN(Real a, Real b, Real c)
    {
        static_assert(N == 3, "three components");
        v[0] = a, v[1] = b, v[2] = c;
    }
    static VecN Zero() { return VecN(); }
    static VecN Ones()
    {
        VecN r;
        for (int k = 0; k < N; ++k)
            r.v[k] = Real(1);
        return r;
    }
    Real &operator[](int k) { return v[k]; }
    const Real &operator[](int k) const { return v[k]; }
    VecN operator+(const VecN &o) const
    {
        VecN r;
        for (int k = 0; k < N; ++k)
            r.v[k] = v[k] + o.v[k];
        return r;
    }
    VecN operator-(const VecN &o) const
    {
        VecN r;
        for (int k = 0; k < N; ++k)
            r.v[k] = v[k] - o.v[k];
        return r;
    }
    VecN operator*(Real s) const
    {
        VecN r;
        for (int k = 0; k < N; ++k)
            r.v[k] = v[k] * s;
        return r;
    }
    VecN operator/(Real s) const
    {
        VecN r;
        for (int k = 0; k < N; ++k)
            r.v[k] = v[k] / s;
        return r;
    }
    Real dot(const VecN &o) const
    {
        Real s = Real(0);
        for (int k = 0; k < N; ++k)
            s += v[k] * o.v[k];
        return s;
    }
    Real squaredNorm() const { return dot(*this); }
    Real norm() const { return std::sqrt(squaredNorm()); }
};
template <int N>
inline VecN<N> operator*(Real s, const VecN<N> &a) { return a * s; }
using Vec2d = VecN<2>;
using Vec3d = VecN<3>;
using Vecd = VecN<Dimensions>;
using Arrayi = std::array<int, Dimensions>;
template <class T>
using StdVec = std::vector<T>;
template <class T>
using SharedPtr = std::shared_ptr<T>;
template <class T, typename... Args>
SharedPtr<T> makeShared(Args &&...args) { return std::make_shared<T>(std::forward<Args>(args)...); }

struct BoundingBoxd // base_data_type.h BaseBoundingBox: lower_, upper_, expand()
{
    Vecd lower_, upper_;
    BoundingBoxd() {}
    BoundingBoxd(const Vecd &lo, const Vecd &hi) : lower_(lo), upper_(hi) {}
    BoundingBoxd expand(Real d) const { return BoundingBoxd(lower_ - Vecd::Ones() * d, upper_ + Vecd::Ones() * d); }
};

// geometries/base_geometry.h, geometric_shape.h, complex_geometry.h: only translations and boxes
class Transform
{
  public:
    Transform() {}
    explicit Transform(const Vecd &translation) : translation_(translation) {}
    Vecd shiftBaseStationToFrame(const Vecd &p) const { return p - translation_; }
    Vecd shiftFrameStationToBase(const Vecd &p) const { return p + translation_; }
    const Vecd &Translation() const { return translation_; }

  private:
    Vecd translation_;
};

class Shape
{
  public:
    explicit Shape(const std::string &name) : name_(name) {}
    virtual ~Shape() = default;
    const std::string &getName() const { return name_; }
    bool isValid() const { return true; }
    virtual bool checkContain(const Vecd &p) = 0;
    virtual Vecd findClosestPoint(const Vecd &p) = 0;
    virtual BoundingBoxd getBounds() { return BoundingBoxd(); } // base_geometry.h: bounding box of the shape
    // base_geometry.cpp:45-59: unit vector to the closest surface point, sign by containment
    Vecd findNormalDirection(const Vecd &p)
    {
        const bool inside = checkContain(p);
        Vecd to_surface = findClosestPoint(p) - p;
        const Real len = to_surface.norm();
        Vecd dir = len > Real(0) ? to_surface / len : Vecd::Zero();
        return inside ? dir : dir * Real(-1);
    }

  protected:
    std::string name_;
};

// geometric_element.cpp:19-60 GeometricBox + the Transform wrapper of geometric_shape.h
class GeometricShapeBox : public Shape
{
  public:
    GeometricShapeBox(const Transform &transform, const Vecd &halfsize, const std::string &name = "GeometricShapeBox")
        : Shape(name), transform_(transform), halfsize_(halfsize) {}
    bool checkContain(const Vecd &p) override
    {
        const Vecd c = transform_.shiftBaseStationToFrame(p);
        for (int d = 0; d < Dimensions; ++d)
            if (std::abs(c[d]) > halfsize_[d])
                return false;
        return true;
    }
    Vecd findClosestPoint(const Vecd &p) override
    {
        const Vecd c = transform_.shiftBaseStationToFrame(p);
        Vecd out = c;
        bool inside = true;
        for (int d = 0; d < Dimensions; ++d)
        {
            const Real clamped = std::min(std::max(c[d], -halfsize_[d]), halfsize_[d]);
            inside = inside && clamped == c[d];
            out[d] = clamped;
        }
        if (inside)
        { // project onto the nearest face (first minimum wins)
            int which = 0;
            Real best = halfsize_[0] - std::abs(c[0]);
            for (int d = 1; d < Dimensions; ++d)
                if (halfsize_[d] - std::abs(c[d]) < best)
                    best = halfsize_[d] - std::abs(c[d]), which = d;
            out[which] = c[which] < Real(0) ? -halfsize_[which] : halfsize_[which];
        }
        return transform_.shiftFrameStationToBase(out);
    }

  private:
    Transform transform_;
    Vecd halfsize_;
};

// complex_geometry.h/.cpp BinaryShapes / ComplexShape: boolean add / subtract of sub-shapes
class ComplexShape : public Shape
{
  public:
    explicit ComplexShape(const std::string &name) : Shape(name) {}
    template <class ShapeType, typename... Args>
    void add(Args &&...args) { parts_.emplace_back(std::make_shared<ShapeType>(std::forward<Args>(args)...), true); }
    template <class ShapeType, typename... Args>
    void subtract(Args &&...args) { parts_.emplace_back(std::make_shared<ShapeType>(std::forward<Args>(args)...), false); }
    bool checkContain(const Vecd &p) override
    {
        bool in = false;
        for (auto &s : parts_)
            in = s.second ? (in || s.first->checkContain(p)) : (in && !s.first->checkContain(p));
        return in;
    }
    Vecd findClosestPoint(const Vecd &p) override
    { // complex_geometry.cpp:118-140: nearest over the sub-shapes, a later sub-shape wins ties
        Vecd best = p;
        Real best_d = std::numeric_limits<Real>::max();
        for (auto &s : parts_)
        {
            const Vecd c = s.first->findClosestPoint(p);
            const Real d = (p - c).norm();
            if (d <= best_d)
                best_d = d, best = c;
        }
        return best;
    }

  private:
    std::vector<std::pair<SharedPtr<Shape>, bool>> parts_;
};

// common/sphinxsys_timer (tick_count wrappers)
struct TimeInterval
{
    double s = 0.0;
    double seconds() const { return s; }
    TimeInterval &operator+=(const TimeInterval &o)
    {
        s += o.s;
        return *this;
    }
    TimeInterval operator-(const TimeInterval &o) const { return TimeInterval{s - o.s}; }
};
struct TickCount
{
    std::chrono::steady_clock::time_point t;
    static TickCount now() { return TickCount{std::chrono::steady_clock::now()}; }
    TimeInterval operator-(const TickCount &o) const { return TimeInterval{std::chrono::duration<double>(t - o.t).count()}; }
};
} // namespace SPH
