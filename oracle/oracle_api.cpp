// ORACLE — TEST INFRASTRUCTURE ONLY (see oracle_core.hpp header).
// C entry points (ctypes) over the templated restatement.  A "system" owns bodies and relations;
// handles are small integers.  Runtime dispatch over (ndim, real_bytes).
#include "oracle_fluid.hpp"
#include "oracle_solid.hpp"
#include "oracle_ck.hpp"
#include <cstdio>
#ifdef _OPENMP
#include <omp.h>
#endif

using namespace orc;

namespace
{
struct SysBase
{
    int ndim, real_bytes;
    virtual ~SysBase() {}
};
template <int D, class R>
struct Sys : SysBase
{
    R lo[D], hi[D];
    std::vector<std::unique_ptr<Body<D, R>>> bodies;
    std::vector<std::unique_ptr<Relation<D, R>>> rels;
};
} // namespace

#define DISPATCH(sys, ...)                                   \
    do                                                       \
    {                                                        \
        SysBase *_s = (SysBase *)(sys);                      \
        if (_s->ndim == 2 && _s->real_bytes == 8)            \
        {                                                    \
            auto &S = *static_cast<Sys<2, double> *>(_s);    \
            using R = double;                                \
            constexpr int D = 2;                             \
            (void)S; (void)D; (void)sizeof(R);               \
            __VA_ARGS__;                                            \
        }                                                    \
        else if (_s->ndim == 3 && _s->real_bytes == 8)       \
        {                                                    \
            auto &S = *static_cast<Sys<3, double> *>(_s);    \
            using R = double;                                \
            constexpr int D = 3;                             \
            (void)S; (void)D; (void)sizeof(R);               \
            __VA_ARGS__;                                            \
        }                                                    \
        else if (_s->ndim == 2 && _s->real_bytes == 4)       \
        {                                                    \
            auto &S = *static_cast<Sys<2, float> *>(_s);     \
            using R = float;                                 \
            constexpr int D = 2;                             \
            (void)S; (void)D; (void)sizeof(R);               \
            __VA_ARGS__;                                            \
        }                                                    \
        else                                                 \
        {                                                    \
            auto &S = *static_cast<Sys<3, float> *>(_s);     \
            using R = float;                                 \
            constexpr int D = 3;                             \
            (void)S; (void)D; (void)sizeof(R);               \
            __VA_ARGS__;                                            \
        }                                                    \
    } while (0)

extern "C"
{
// system bounds are the *expanded* bounds (sphinxsys_system/sph_system.cpp:39: user box ± 4 dp)
void *orc_create(int ndim, int real_bytes, const double *lo, const double *hi)
{
    SysBase *s;
    if (ndim == 2 && real_bytes == 8)
        s = new Sys<2, double>();
    else if (ndim == 3 && real_bytes == 8)
        s = new Sys<3, double>();
    else if (ndim == 2 && real_bytes == 4)
        s = new Sys<2, float>();
    else
        s = new Sys<3, float>();
    s->ndim = ndim;
    s->real_bytes = real_bytes;
    DISPATCH(s, for (int d = 0; d < D; ++d) {
        S.lo[d] = R(lo[d]);
        S.hi[d] = R(hi[d]);
    });
    return s;
}
void orc_destroy(void *sys) { delete (SysBase *)sys; }

// bodies/base_body.h + adaptations/adaptation.cpp:12-19 (h = h_ratio*dp, Wendland C2, sigma0) +
// cell linked list mesh (meshes/cell_linked_list.cpp:10-20: spacing = cut-off radius, buffer 2)
int orc_body_add(void *sys, const char *name, long n, double dp, double h_ratio)
{
    int id = -1;
    DISPATCH(sys, {
        auto b = std::make_unique<Body<D, R>>();
        b->name = name;
        b->n = size_t(n);
        b->dp = R(dp);
        b->h = R(h_ratio) * R(dp);
        b->kernel = Kernel<D, R>(b->h);
        b->sigma0 = latticeNumberDensity<D, R>(b->kernel, b->dp);
        b->mesh = Mesh<D, R>(S.lo, S.hi, b->kernel.rc, 2);
        for (int d = 0; d < D; ++d)
            b->periodic[d] = false;
        b->original_id.resize(b->n);
        b->sorted_id.resize(b->n);
        for (size_t i = 0; i < b->n; ++i)
            b->original_id[i] = b->sorted_id[i] = i;
        b->v("Position", D);
        b->v("VolumetricMeasure");
        b->evolving("Position");
        b->evolving("VolumetricMeasure");
        S.bodies.push_back(std::move(b));
        id = int(S.bodies.size()) - 1;
    });
    return id;
}
void orc_body_material(void *sys, int body, double rho0, double c0, double mu)
{
    DISPATCH(sys, {
        S.bodies[body]->mat.rho0 = rho0;
        S.bodies[body]->mat.c0 = c0;
        S.bodies[body]->mat.mu = mu;
    });
}
void orc_body_elastic(void *sys, int body, double rho0, double E, double nu)
{
    DISPATCH(sys, setElastic<D, R>(*S.bodies[body], rho0, E, nu));
}
void orc_body_periodic(void *sys, int body, int axis, double lo, double hi)
{
    DISPATCH(sys, {
        S.bodies[body]->periodic[axis] = true;
        S.bodies[body]->per_lo[axis] = R(lo);
        S.bodies[body]->per_hi[axis] = R(hi);
    });
}
// registry access: returns a pointer into the oracle's own storage (array of Vecd, width Reals each)
void *orc_var(void *sys, int body, const char *name, int width, int evolving)
{
    void *p = nullptr;
    DISPATCH(sys, {
        p = S.bodies[body]->v(name, width);
        if (evolving)
            S.bodies[body]->evolving(name);
    });
    return p;
}
int orc_var_width(void *sys, int body, const char *name)
{
    int w = 0;
    DISPATCH(sys, {
        auto it = S.bodies[body]->vars.find(name);
        w = it == S.bodies[body]->vars.end() ? 0 : it->second.width;
    });
    return w;
}
void orc_body_info(void *sys, int body, double *out /* h, rc, W0, sigma0, lo[3], spacing, cells[3] */)
{
    DISPATCH(sys, {
        auto &b = *S.bodies[body];
        out[0] = b.h;
        out[1] = b.kernel.rc;
        out[2] = b.kernel.W0();
        out[3] = b.sigma0;
        for (int d = 0; d < 3; ++d)
        {
            out[4 + d] = d < D ? double(b.mesh.lo[d]) : 0.0;
            out[8 + d] = d < D ? double(b.mesh.cells[d]) : 1.0;
        }
        out[7] = b.mesh.spacing;
    });
}
void orc_ids(void *sys, int body, long *original_id, long *sorted_id)
{
    DISPATCH(sys, {
        auto &b = *S.bodies[body];
        for (size_t i = 0; i < b.n; ++i)
        {
            original_id[i] = long(b.original_id[i]);
            sorted_id[i] = long(b.sorted_id[i]);
        }
    });
}
void orc_cell_keys(void *sys, int body, long *linear, long *morton)
{
    DISPATCH(sys, {
        auto &b = *S.bodies[body];
        const R *pos = b.v("Position", D);
        for (size_t i = 0; i < b.n; ++i)
        {
            int c[D];
            b.mesh.cellIndex(pos + i * D, c);
            linear[i] = long(b.mesh.linear(c));
            morton[i] = long(Mesh<D, R>::morton(c));
        }
    });
}
void orc_update_cell_list(void *sys, int body)
{
    DISPATCH(sys, {
        auto &b = *S.bodies[body];
        updateCellLists<D, R>(b);
    });
}
void orc_periodic_bounding(void *sys, int body, int axis) { DISPATCH(sys, periodicBounding<D, R>(*S.bodies[body], axis)); }
void orc_periodic_cell_list(void *sys, int body, int axis) { DISPATCH(sys, periodicCellLinkedList<D, R>(*S.bodies[body], axis)); }
void orc_sort(void *sys, int body) { DISPATCH(sys, particleSort<D, R>(*S.bodies[body], nullptr)); }

int orc_relation_inner(void *sys, int body)
{
    int id = -1;
    DISPATCH(sys, {
        auto r = std::make_unique<Relation<D, R>>();
        r->body = S.bodies[body].get();
        r->inner = true;
        setupRelation<D, R>(*r);
        S.rels.push_back(std::move(r));
        id = int(S.rels.size()) - 1;
    });
    return id;
}
int orc_relation_contact(void *sys, int body, const int *targets, int n_targets)
{
    int id = -1;
    DISPATCH(sys, {
        auto r = std::make_unique<Relation<D, R>>();
        r->body = S.bodies[body].get();
        r->inner = false;
        for (int k = 0; k < n_targets; ++k)
            r->targets.push_back(S.bodies[targets[k]].get());
        setupRelation<D, R>(*r);
        S.rels.push_back(std::move(r));
        id = int(S.rels.size()) - 1;
    });
    return id;
}
// SelfSurfaceContactRelation / SurfaceContactRelation over a body part by particle (original ids; n_part = 0: all)
int orc_relation_self_contact(void *sys, int body, const long *part, long n_part)
{
    int id = orc_relation_inner(sys, body);
    DISPATCH(sys, {
        auto &r = *S.rels[id];
        r.builder = 1;
        if (n_part > 0)
        {
            r.part.assign(r.body->n, 0);
            for (long k = 0; k < n_part; ++k)
                r.part[size_t(part[k])] = 1;
        }
    });
    return id;
}
int orc_relation_surface_contact(void *sys, int body, const int *targets, int n_targets, const long *part, long n_part)
{
    int id = orc_relation_contact(sys, body, targets, n_targets);
    DISPATCH(sys, {
        auto &r = *S.rels[id];
        r.builder = 2;
        setupRelation<D, R>(r);
        if (n_part > 0)
        {
            r.part.assign(r.body->n, 0);
            for (long k = 0; k < n_part; ++k)
                r.part[size_t(part[k])] = 1;
        }
    });
    return id;
}
void orc_relation_update(void *sys, int rel) { DISPATCH(sys, updateRelation<D, R>(*S.rels[rel])); }
// neighbour dump: counts[n]; if j_out != null, flat concatenation in list order (k-th target)
long orc_neighbors(void *sys, int rel, int k, int *counts, long *j_out)
{
    long total = 0;
    DISPATCH(sys, {
        auto &cfg = S.rels[rel]->configs[k];
        for (size_t i = 0; i < cfg.size(); ++i)
        {
            if (counts)
                counts[i] = int(cfg[i].current_size);
            if (j_out)
                for (size_t m = 0; m < cfg[i].current_size; ++m)
                    j_out[total + long(m)] = long(cfg[i].j[m]);
            total += long(cfg[i].current_size);
        }
    });
    return total;
}
int orc_search_depth(void *sys, int rel, int k)
{
    int d = 0;
    DISPATCH(sys, d = S.rels[rel]->depths[k]);
    return d;
}

void orc_density_summation(void *sys, int inner, int contact, int free_surface)
{
    DISPATCH(sys, densitySummation<D, R>(*S.rels[inner], contact >= 0 ? S.rels[contact].get() : nullptr, free_surface != 0));
}
void orc_integration_1st_half(void *sys, int inner, int wall, int riemann, double dt)
{
    DISPATCH(sys, integration1stHalf<D, R>(*S.rels[inner], wall >= 0 ? S.rels[wall].get() : nullptr, riemann != 0, R(dt)));
}
void orc_integration_2nd_half(void *sys, int inner, int wall, int riemann, double dt)
{
    DISPATCH(sys, integration2ndHalf<D, R>(*S.rels[inner], wall >= 0 ? S.rels[wall].get() : nullptr, riemann != 0, R(dt)));
}
double orc_acoustic_dt(void *sys, int body, double cfl)
{
    double r = 0;
    DISPATCH(sys, r = acousticTimeStep<D, R>(*S.bodies[body], R(cfl)));
    return r;
}
double orc_advection_dt(void *sys, int body, double U_ref, double cfl, int viscous)
{
    double r = 0;
    DISPATCH(sys, r = advectionTimeStep<D, R>(*S.bodies[body], R(U_ref), R(cfl), viscous != 0));
    return r;
}
void orc_gravity_force(void *sys, int body, const double *g)
{
    DISPATCH(sys, {
        R gg[D];
        for (int d = 0; d < D; ++d)
            gg[d] = R(g[d]);
        gravityForce<D, R>(*S.bodies[body], gg);
    });
}
void orc_viscous_force(void *sys, int inner, int wall)
{
    DISPATCH(sys, viscousForce<D, R>(*S.rels[inner], wall >= 0 ? S.rels[wall].get() : nullptr));
}
void orc_transport_velocity_correction(void *sys, int inner, double coeff)
{
    DISPATCH(sys, transportVelocityCorrection<D, R>(*S.rels[inner], R(coeff)));
}
void orc_ck_density_regularization(void *sys, int inner, int contact, int free_surface)
{
    DISPATCH(sys, compressionRegularization<D, R>(*S.rels[inner], contact >= 0 ? S.rels[contact].get() : nullptr, free_surface != 0));
}
void orc_ck_linear_correction_matrix(void *sys, int inner, int contact, double alpha)
{
    DISPATCH(sys, linearCorrectionMatrixCK<D, R>(*S.rels[inner], contact >= 0 ? S.rels[contact].get() : nullptr, R(alpha)));
}
void orc_ck_free_surface_indication(void *sys, int inner, int contact)
{
    DISPATCH(sys, freeSurfaceIndicationCK<D, R>(*S.rels[inner], contact >= 0 ? S.rels[contact].get() : nullptr));
}
void orc_ck_advection_step_setup(void *sys, int body) { DISPATCH(sys, ckAdvectionStepSetup<D, R>(*S.bodies[body])); }
void orc_ck_update_particle_position(void *sys, int body) { DISPATCH(sys, ckUpdateParticlePosition<D, R>(*S.bodies[body])); }
double orc_ck_acoustic_dt(void *sys, int body, double cfl)
{
    double r = 0;
    DISPATCH(sys, r = double(ckAcousticDt<D, R>(*S.bodies[body], R(cfl))));
    return r;
}
double orc_ck_advection_dt(void *sys, int body, double U_ref, double cfl)
{
    double r = 0;
    DISPATCH(sys, r = double(ckAdvectionDt<D, R>(*S.bodies[body], R(U_ref), R(cfl))));
    return r;
}
void orc_ck_acoustic_1st_half(void *sys, int inner, int wall, double dt)
{
    DISPATCH(sys, ckAcoustic1stHalf<D, R>(*S.rels[inner], wall >= 0 ? S.rels[wall].get() : nullptr, R(dt)));
}
void orc_ck_acoustic_2nd_half(void *sys, int inner, int wall, double dt)
{
    DISPATCH(sys, ckAcoustic2ndHalf<D, R>(*S.rels[inner], wall >= 0 ? S.rels[wall].get() : nullptr, R(dt)));
}
void orc_ck_pressure_force_from_fluid(void *sys, int contact, int corrected)
{
    DISPATCH(sys, ckPressureForceFromFluid<D, R>(*S.rels[contact], corrected != 0));
}
void orc_ck_viscous_force_from_fluid(void *sys, int contact, int corrected)
{
    DISPATCH(sys, ckViscousForceFromFluid<D, R>(*S.rels[contact], corrected != 0));
}
void orc_position_bound(void *sys, int body, int upper, double *out3)
{
    DISPATCH(sys, positionBound<D, R>(*S.bodies[body], upper != 0, out3));
}
double orc_total_kinetic_energy(void *sys, int body)
{
    double r = 0;
    DISPATCH(sys, r = totalKineticEnergy<D, R>(*S.bodies[body]));
    return r;
}
double orc_total_mechanical_energy(void *sys, int body, const double *g)
{
    double r = 0;
    DISPATCH(sys, {
        R gg[D];
        for (int d = 0; d < D; ++d)
            gg[d] = R(g[d]);
        r = totalMechanicalEnergy<D, R>(*S.bodies[body], gg);
    });
    return r;
}
double orc_maximum_speed(void *sys, int body)
{
    double r = 0;
    DISPATCH(sys, r = maximumSpeed<D, R>(*S.bodies[body]));
    return r;
}
void orc_observe(void *sys, int contact, const char *name, int width, double *out)
{
    DISPATCH(sys, {
        std::vector<R> tmp(S.rels[contact]->body->n * size_t(width));
        observe<D, R>(*S.rels[contact], name, width, tmp.data());
        for (size_t i = 0; i < tmp.size(); ++i)
            out[i] = double(tmp[i]);
    });
}

// ---- total-Lagrangian solid (oracle_solid.hpp) ----
void orc_correction_matrix(void *sys, int inner) { DISPATCH(sys, linearGradientCorrectionMatrix<D, R>(*S.rels[inner])); }
void orc_solid_1st_half(void *sys, int inner, int decomposed, double dt)
{
    DISPATCH(sys, solidIntegration1stHalf<D, R>(*S.rels[inner], decomposed != 0, R(dt)));
}
void orc_solid_2nd_half(void *sys, int inner, double dt) { DISPATCH(sys, solidIntegration2ndHalf<D, R>(*S.rels[inner], R(dt))); }
void orc_solid_constrain(void *sys, int body, const long *ids, long n)
{
    DISPATCH(sys, fixBodyPart<D, R>(*S.bodies[body], ids, n));
}
double orc_solid_dt(void *sys, int body, double cfl)
{
    double r = 0;
    DISPATCH(sys, r = solidAcousticTimeStep<D, R>(*S.bodies[body], R(cfl)));
    return r;
}
int orc_num_threads()
{
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}
// the bench's CPU legs run on every host core whatever OMP_NUM_THREADS the launcher exported
// (torch.distributed.run sets it to 1): returns the thread count in effect afterwards
int orc_set_num_threads(int n)
{
#ifdef _OPENMP
    if (n > 0)
        omp_set_num_threads(n);
    return omp_get_max_threads();
#else
    return 1;
#endif
}
} // extern "C"
