// 3-D dam break through the C++ adapter: the reference's main
// (tests/3d_examples/test_3d_dambreak/dambreak.cpp:10-208) with the same objects in the same order; only the
// particle generation is spelled out (box lattices, particle_generator_lattice_3d.cpp:12-25) because shapes and
// level sets are outside the path.  Build: make -C examples   Run: examples/dambreak_3d [dp] [advection steps]
#include "sphb200_adapter.hpp"

#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
using namespace SPH;

// dambreak.cpp:10-24
static const Real DL = 5.366, DH = 2.0, DW = 0.5, LL = 2.0, LH = 1.0, LW = 0.5;
static const Real rho0_f = 1.0, gravity_g = 1.0;

struct Box
{
    Vecd center, half;
    bool contains(const Vecd &p) const
    {
        for (int d = 0; d < 3; ++d)
            if (std::fabs(p[d] - center[d]) > half[d])
                return false;
        return true;
    }
    // GeometricBox::findClosestPoint (geometric_element.cpp:19-60)
    Vecd closest(const Vecd &p) const
    {
        Vecd c, cl;
        bool inside = true;
        for (int d = 0; d < 3; ++d)
        {
            c[d] = p[d] - center[d];
            cl[d] = std::min(std::max(c[d], -half[d]), half[d]);
            inside = inside && cl[d] == c[d];
        }
        if (inside)
        {
            int which = 0;
            Real best = half[0] - std::fabs(c[0]);
            for (int d = 1; d < 3; ++d)
                if (half[d] - std::fabs(c[d]) < best)
                    best = half[d] - std::fabs(c[d]), which = d;
            cl = c;
            cl[which] = c[which] < 0 ? -half[which] : half[which];
        }
        for (int d = 0; d < 3; ++d)
            cl[d] += center[d];
        return cl;
    }
};

// lattice points of the (expanded) system box that satisfy `keep`, x-major, z fastest
template <class Keep>
static std::vector<Vecd> lattice(const BoundingBox &b, Real dp, Keep keep)
{
    long n[3];
    for (int d = 0; d < 3; ++d)
        n[d] = long(std::ceil((b.second[d] - b.first[d]) / dp));
    std::vector<Vecd> out;
    for (long i = 0; i < n[0]; ++i)
        for (long j = 0; j < n[1]; ++j)
            for (long k = 0; k < n[2]; ++k)
            {
                Vecd p = {(b.first[0] + Real(i) * dp) + Real(0.5) * dp, (b.first[1] + Real(j) * dp) + Real(0.5) * dp,
                          (b.first[2] + Real(k) * dp) + Real(0.5) * dp};
                if (keep(p))
                    out.push_back(p);
            }
    return out;
}

int main(int argc, char **argv)
{
    const Real particle_spacing_ref = argc > 1 ? Real(std::atof(argv[1])) : Real(0.05);
    const int advection_steps = argc > 2 ? std::atoi(argv[2]) : 10;
    const Real BW = particle_spacing_ref * 4;
    const Real U_f = 2.0 * std::sqrt(gravity_g * LH), c_f = 10.0 * U_f;

    BoundingBox system_domain_bounds{Vecd{-BW, -BW, -BW}, Vecd{DL + BW, DH + BW, DW + BW}};
    SPHSystem sph_system(system_domain_bounds, particle_spacing_ref);
    const Box water{{Real(0.5) * LL, Real(0.5) * LH, Real(0.5) * LW}, {Real(0.5) * LL, Real(0.5) * LH, Real(0.5) * LW}};
    const Box inner{{Real(0.5) * DL, Real(0.5) * DH, Real(0.5) * DW}, {Real(0.5) * DL, Real(0.5) * DH, Real(0.5) * DW}};
    const Box outer{inner.center, {inner.half[0] + BW, inner.half[1] + BW, inner.half[2] + BW}};
    const Real volume = particle_spacing_ref * particle_spacing_ref * particle_spacing_ref;

    FluidBody water_block(sph_system, "WaterBody");
    water_block.defineMaterial(WeaklyCompressibleFluid(rho0_f, c_f));
    water_block.generateParticles(lattice(sph_system.getSystemDomainBounds(), particle_spacing_ref, [&](const Vecd &p) { return water.contains(p); }),
                                  volume);
    for (const char *name : {"Velocity", "Force", "ForcePrior", "PreviousGravityForce"})
        water_block.registerStateVariable(name, Dimensions);
    for (const char *name : {"Pressure", "DensityChangeRate"})
        water_block.registerStateVariable(name, 1);

    SolidBody wall_boundary(sph_system, "WallBoundary");
    wall_boundary.defineMaterial(Solid());
    std::vector<Vecd> wall_points =
        lattice(sph_system.getSystemDomainBounds(), particle_spacing_ref, [&](const Vecd &p) { return outer.contains(p) && !inner.contains(p); });
    wall_boundary.generateParticles(wall_points, volume);
    { // SimpleDynamics<NormalDirectionFromBodyShape> (general_geometric.cpp:18-26, base_geometry.cpp:45-59)
        std::vector<Vecd> normals(wall_points.size());
        for (size_t i = 0; i < wall_points.size(); ++i)
        {
            const Vecd &p = wall_points[i];
            Vecd co = outer.closest(p), ci = inner.closest(p);
            auto dist = [&](const Vecd &q) { return std::sqrt((p[0] - q[0]) * (p[0] - q[0]) + (p[1] - q[1]) * (p[1] - q[1]) + (p[2] - q[2]) * (p[2] - q[2])); };
            const Vecd &c = dist(ci) <= dist(co) ? ci : co;
            Real len = dist(c);
            normals[i] = {(c[0] - p[0]) / len, (c[1] - p[1]) / len, (c[2] - p[2]) / len};
        }
        wall_boundary.setVariable("NormalDirection", normals);
        wall_boundary.registerStateVariable("Velocity", Dimensions);
        wall_boundary.registerStateVariable("Acceleration", Dimensions);
    }

    ObserverBody fluid_observer(sph_system, "FluidObserver"); // dambreak.cpp:52-63
    {
        std::vector<Vecd> probes;
        for (Real y : {0.01, 0.1, 0.2, 0.24, 0.252, 0.266})
            probes.push_back({DL, y, Real(0.5) * DW});
        fluid_observer.generateParticles(probes, volume);
    }

    InnerRelation water_block_inner(water_block);
    ContactRelation water_wall_contact(water_block, {&wall_boundary});
    ContactRelation fluid_observer_contact(fluid_observer, {&water_block});
    ComplexRelation water_block_complex(water_block_inner, water_wall_contact);

    Gravity gravity(Vecd{0.0, -gravity_g, 0.0});
    SimpleDynamics<GravityForce<Gravity>> constant_gravity(water_block, gravity);
    Dynamics1Level<fluid_dynamics::Integration1stHalfWithWallRiemann> pressure_relaxation(water_block_inner, water_wall_contact);
    Dynamics1Level<fluid_dynamics::Integration2ndHalfWithWallRiemann> density_relaxation(water_block_inner, water_wall_contact);
    InteractionWithUpdate<fluid_dynamics::DensitySummationComplexFreeSurface> update_density_by_summation(water_block_inner, water_wall_contact);
    ReduceDynamics<fluid_dynamics::AdvectionTimeStep> get_fluid_advection_time_step_size(water_block, U_f);
    ReduceDynamics<fluid_dynamics::AcousticTimeStep> get_fluid_time_step_size(water_block);
    ReduceDynamics<TotalMechanicalEnergy> write_water_mechanical_energy(water_block, gravity);
    ObservingAQuantity write_recorded_water_pressure(fluid_observer_contact, "Pressure");

    sph_system.initializeSystemCellLinkedLists();
    sph_system.initializeSystemConfigurations();
    constant_gravity.exec();

    // dambreak.cpp:159-195
    Real physical_time = 0, dt = 0;
    long number_of_iterations = 0, acoustic_sub_steps = 0;
    sph_system.synchronize();
    auto t0 = std::chrono::steady_clock::now();
    while (number_of_iterations < advection_steps)
    {
        Real Dt = get_fluid_advection_time_step_size.exec();
        update_density_by_summation.exec();
        Real relaxation_time = 0;
        long k = 0;
        while (relaxation_time < Dt)
        {
            pressure_relaxation.exec(dt);
            density_relaxation.exec(dt);
            dt = get_fluid_time_step_size.exec();
            relaxation_time += dt;
            physical_time += dt;
            ++k;
        }
        acoustic_sub_steps += k;
        ++number_of_iterations;
        std::printf("N=%ld Dt=%.17g sub_steps=%ld dt=%.17g\n", number_of_iterations, double(Dt), k, double(dt));
        water_block.updateCellLinkedList();
        water_block_complex.updateConfiguration();
        fluid_observer_contact.updateConfiguration();
    }
    sph_system.synchronize();
    double seconds = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    std::printf("energy=%.17g\n", double(write_water_mechanical_energy.exec()));
    std::vector<double> p = write_recorded_water_pressure.exec();
    std::printf("probes=");
    for (double v : p)
        std::printf("%.17g ", v);
    std::printf("\nfluid=%zu wall=%zu acoustic_sub_steps=%ld time=%.17g launches=%lld particle_steps_per_s=%.4g\n",
                water_block.TotalRealParticles(), wall_boundary.TotalRealParticles(), acoustic_sub_steps, double(physical_time),
                sph_system.kernelLaunches(), double(water_block.TotalRealParticles()) * double(acoustic_sub_steps) / seconds);
    return 0;
}
