// ORACLE — TEST INFRASTRUCTURE ONLY. Never linked, imported or executed by the product path
// (sphinxsys-learning_b200/). Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
// --impl reference legs may use it, and there only as the checker / CPU baseline.
//
// CPU restatement of the *classic* (TBB) SPHinXsys particle-interaction path, written from the
// reference's algorithm description, with no Eigen/TBB/Boost dependency.  All citations are
// relative to /root/reference/src/shared unless noted.
//
// Parity status: the reference cannot be compiled in this container (no Eigen/TBB/Boost/Simbody),
// so this restatement is pinned against the reference's own regression fixtures
// (tests/*/regression_test_tool/*.xml, extracted to tests/golden/ by tests/golden/make_golden.py)
// and closed-form known answers.  Per-step fields are unpinned by the reference
// itself; this oracle is the pin.
#pragma once
#include <algorithm>
#include <array>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <map>
#include <memory>
#include <string>
#include <vector>

namespace orc
{
constexpr double kPi = 3.14159265358979323846;

template <class R>
struct Tiny
{
    // common/base_data_type.h:207  constexpr Real TinyReal = Real(2.71051e-20);
    static constexpr R value = R(2.71051e-20);
};

// ---------------------------------------------------------------------------------------------
// Mesh: meshes/base_mesh.cpp:6-16 (ctor), base_mesh.hxx:9-15 (CellIndexFromPosition),
//       base_mesh.hxx:68-78 (transferMeshIndexTo1D), base_mesh.hxx:80-99 (Morton).
// ---------------------------------------------------------------------------------------------
template <int D, class R>
struct Mesh
{
    R lo[D];
    R spacing;
    int grid_points[D];
    int cells[D];

    Mesh() : spacing(1) {}
    // tentative bounds + buffer_width cells of padding on both sides
    Mesh(const R *blo, const R *bhi, R grid_spacing, int buffer_width) : spacing(grid_spacing)
    {
        for (int d = 0; d < D; ++d)
        {
            R buffer = R(buffer_width) * grid_spacing * R(1);
            lo[d] = blo[d] - buffer;
            R dim = bhi[d] + buffer - lo[d];
            grid_points[d] = int(std::ceil(dim / grid_spacing)) + 1;
            cells[d] = grid_points[d] - 1;
        }
    }
    void cellIndex(const R *pos, int *out) const
    {
        for (int d = 0; d < D; ++d)
        {
            int c = int(std::floor((pos[d] - lo[d]) / spacing));
            c = std::max(c, 0);
            c = std::min(c, grid_points[d] - 2);
            out[d] = c;
        }
    }
    size_t linear(const int *c) const
    {
        if (D == 2)
            return size_t(c[0]) * cells[1] + c[1];
        return size_t(c[0]) * cells[1] * cells[D - 1] + size_t(c[1]) * cells[D - 1] + c[D - 1];
    }
    size_t totalCells() const
    {
        size_t n = 1;
        for (int d = 0; d < D; ++d)
            n *= size_t(cells[d]);
        return n;
    }
    static uint64_t mortonPart(uint64_t i)
    {
        uint64_t x = i;
        x &= 0x3ff;
        x = (x | x << 16) & 0x30000ff;
        x = (x | x << 8) & 0x300f00f;
        x = (x | x << 4) & 0x30c30c3;
        x = (x | x << 2) & 0x9249249;
        return x;
    }
    static uint64_t morton(const int *c)
    {
        uint64_t m = mortonPart(uint64_t(c[0])) | (mortonPart(uint64_t(c[1])) << 1);
        if (D == 3)
            m |= (mortonPart(uint64_t(c[D - 1])) << 2);
        return m;
    }
    // base_mesh.cpp:26-29 CellPositionFromIndex (used by the lattice generator)
    void cellCenter(const int *c, R *out) const
    {
        for (int d = 0; d < D; ++d)
            out[d] = (lo[d] + R(c[d]) * spacing) + R(0.5) * R(1) * spacing;
    }
};

// ---------------------------------------------------------------------------------------------
// Wendland C2: kernels/kernel_wendland_c2.cpp:8-35, kernels/base_kernel.cpp:6-27,50-78,
//              kernels/base_kernel.h:99-114.
// ---------------------------------------------------------------------------------------------
template <int D, class R>
struct Kernel
{
    R h, inv_h, rc, rc2, factor_W, factor_dW;
    Kernel() {}
    explicit Kernel(R h_) : h(h_), inv_h(R(1.0) / h_)
    {
        rc = R(2.0) * h;
        rc2 = rc * rc;
        if (D == 2)
            factor_W = inv_h * inv_h * R(7.0) / (R(4.0) * R(kPi));
        else
            factor_W = inv_h * inv_h * inv_h * R(21.0) / (R(16.0) * R(kPi));
        factor_dW = inv_h * factor_W;
    }
    R W(R r) const
    {
        R q = r * inv_h;
        return factor_W * (R(std::pow(R(1.0) - R(0.5) * q, 4)) * (R(1.0) + R(2.0) * q));
    }
    R dW(R r) const
    {
        R q = r * inv_h;
        return factor_dW * (R(0.625) * R(std::pow(q - R(2.0), 3)) * q);
    }
    R W0() const { return factor_W; }
};

// adaptations/adaptation.cpp:26-60 computeLatticeNumberDensity
template <int D, class R>
R latticeNumberDensity(const Kernel<D, R> &k, R dp)
{
    R sigma = 0;
    int depth = int(k.rc / dp) + 1;
    if (D == 2)
    {
        for (int j = -depth; j <= depth; ++j)
            for (int i = -depth; i <= depth; ++i)
            {
                R x = i * dp, y = j * dp;
                R dist = std::sqrt(x * x + y * y);
                if (dist < k.rc)
                    sigma += k.W(dist);
            }
    }
    else
    {
        for (int kk = -depth; kk <= depth; ++kk)
            for (int j = -depth; j <= depth; ++j)
                for (int i = -depth; i <= depth; ++i)
                {
                    R x = i * dp, y = j * dp, z = kk * dp;
                    R dist = std::sqrt(x * x + y * y + z * z);
                    if (dist < k.rc)
                        sigma += k.W(dist);
                }
    }
    return sigma;
}

// ---------------------------------------------------------------------------------------------
// Neighborhood: particle_neighborhood/neighborhood.h:49-66 (stored j, W, dW, r, e per pair).
// ---------------------------------------------------------------------------------------------
template <int D, class R>
struct Neighborhood
{
    size_t current_size = 0;
    std::vector<size_t> j;
    std::vector<R> W, dW, r;
    std::vector<R> e; // D per neighbour
    void push(size_t jj, R w, R dw, R rr, const R *ee)
    {
        if (current_size >= j.size())
        {
            j.push_back(jj);
            W.push_back(w);
            dW.push_back(dw);
            r.push_back(rr);
            for (int d = 0; d < D; ++d)
                e.push_back(ee[d]);
        }
        else
        {
            j[current_size] = jj;
            W[current_size] = w;
            dW[current_size] = dw;
            r[current_size] = rr;
            for (int d = 0; d < D; ++d)
                e[current_size * D + d] = ee[d];
        }
        current_size++;
    }
};

// A named per-particle variable (particles/base_particles.h:119-162 registry, array-of-Vecd layout
// common/sphinxsys_variable.h:204-236).
template <class R>
struct Var
{
    int width = 1;
    bool evolving = false;
    std::vector<R> data;
};

struct Material
{
    double rho0 = 1.0, c0 = 1.0, mu = 0.0;
    // elastic solid (materials/elastic_solid.cpp)
    double E = 0, nu = 0, K0 = 0, G0 = 0, lambda0 = 0, cs0 = 0;
};

template <int D, class R>
struct Body
{
    std::string name;
    size_t n = 0;
    R dp, h;
    Kernel<D, R> kernel;
    R sigma0;
    Mesh<D, R> mesh;
    Material mat;
    std::map<std::string, Var<R>> vars;
    std::vector<size_t> original_id, sorted_id;
    // cell lists: meshes/cell_linked_list.cpp:84-101 (index lists) and :36-56 (index+pos copies)
    std::vector<std::vector<size_t>> cell_index_lists;
    std::vector<std::vector<std::pair<size_t, std::array<R, D>>>> cell_data_lists;
    // periodic axes (general_dynamics/domian_bouding/domain_bounding.h:49-70)
    bool periodic[D];
    R per_lo[D], per_hi[D];

    R *v(const std::string &nm, int width = 1)
    {
        auto it = vars.find(nm);
        if (it == vars.end())
        {
            Var<R> nv;
            nv.width = width;
            nv.data.assign(n * size_t(width), R(0));
            it = vars.emplace(nm, std::move(nv)).first;
        }
        return it->second.data.data();
    }
    bool has(const std::string &nm) const { return vars.count(nm) != 0; }
    void evolving(const std::string &nm) { vars.at(nm).evolving = true; }
};

// ---------------------------------------------------------------------------------------------
// UpdateCellLists: meshes/cell_linked_list.cpp:84-101,181-186 — serial insertion (ascending
// particle index) makes the oracle deterministic; the reference's order is arbitrary.
// ---------------------------------------------------------------------------------------------
template <int D, class R>
void updateCellLists(Body<D, R> &b)
{
    size_t nc = b.mesh.totalCells();
    if (b.cell_index_lists.size() != nc)
    {
        b.cell_index_lists.assign(nc, {});
        b.cell_data_lists.assign(nc, {});
    }
    for (auto &l : b.cell_index_lists)
        l.clear();
    const R *pos = b.v("Position", D);
    for (size_t i = 0; i < b.n; ++i)
    {
        int c[D];
        b.mesh.cellIndex(pos + i * D, c);
        b.cell_index_lists[b.mesh.linear(c)].push_back(i);
    }
    for (size_t c = 0; c < nc; ++c)
    {
        auto &dl = b.cell_data_lists[c];
        dl.clear();
        for (size_t idx : b.cell_index_lists[c])
        {
            std::array<R, D> p;
            for (int d = 0; d < D; ++d)
                p[d] = pos[idx * D + d];
            dl.emplace_back(idx, p);
        }
    }
}

// squaredNorm without FMA contraction (build flag -ffp-contract=off), left-to-right sum.
template <int D, class R>
inline R sqnorm(const R *d)
{
    R s = d[0] * d[0];
    for (int k = 1; k < D; ++k)
        s = s + d[k] * d[k];
    return s;
}

// searchNeighborsByMesh: meshes/cell_linked_list.hpp:20-46 with NeighborBuilderInner
// (particle_neighborhood/neighborhood.cpp:85-98) or NeighborBuilderContact (:147-160).
// builder 1: NeighborBuilderSelfContact (neighborhood.cpp:123-142), part: the body part's particles (current indices
// flagged through original ids), others keep empty neighbourhoods (inner_body_relation.cpp:57-75)
template <int D, class R>
void searchNeighbors(Body<D, R> &src, Body<D, R> &tgt, const Kernel<D, R> &kernel, int depth, bool inner,
                     std::vector<Neighborhood<D, R>> &cfg, int builder = 0, const std::vector<unsigned char> *part = nullptr)
{
    const R *pos = src.v("Position", D);
    const R *pos0_s = builder == 1 ? src.v("InitialPosition", D) : nullptr;
    const R *pos0_t = builder == 1 ? tgt.v("InitialPosition", D) : nullptr;
    cfg.resize(src.n);
    const Mesh<D, R> &mesh = tgt.mesh;
#pragma omp parallel for schedule(dynamic, 256)
    for (long ii = 0; ii < long(src.n); ++ii)
    {
        size_t i = size_t(ii);
        Neighborhood<D, R> &nb = cfg[i];
        nb.current_size = 0;
        if (part && !(*part)[src.original_id[i]])
            continue;
        int c[D], lo[D], hi[D];
        mesh.cellIndex(pos + i * D, c);
        for (int d = 0; d < D; ++d)
        {
            lo[d] = std::max(0, c[d] - depth);
            hi[d] = std::min(mesh.cells[d], c[d] + depth + 1);
        }
        int cc[3] = {0, 0, 0};
        int l3 = (D == 3) ? lo[D - 1] : 0, h3 = (D == 3) ? hi[D - 1] : 1;
        for (cc[0] = lo[0]; cc[0] < hi[0]; ++cc[0])
            for (cc[1] = lo[1]; cc[1] < hi[1]; ++cc[1])
                for (cc[2] = l3; cc[2] < h3; ++cc[2])
                {
                    const auto &cell = tgt.cell_data_lists[mesh.linear(cc)];
                    for (const auto &entry : cell)
                    {
                        size_t j = entry.first;
                        R dsp[D];
                        for (int d = 0; d < D; ++d)
                            dsp[d] = pos[i * D + d] - entry.second[d];
                        R metric = sqnorm<D, R>(dsp);
                        bool accept = metric < kernel.rc2 && (!inner || i != j);
                        if (builder == 1)
                        {
                            R d0[D];
                            for (int d = 0; d < D; ++d)
                                d0[d] = pos0_s[i * D + d] - pos0_t[j * D + d];
                            accept = std::sqrt(metric) < kernel.rc && std::sqrt(sqnorm<D, R>(d0)) > kernel.rc;
                        }
                        if (accept)
                        {
                            R dist = std::sqrt(metric);
                            R e[D];
                            for (int d = 0; d < D; ++d)
                                e[d] = dsp[d] / (dist + Tiny<R>::value);
                            nb.push(j, kernel.W(dist), kernel.dW(dist), dist, e);
                        }
                    }
                }
    }
}

template <int D, class R>
struct Relation
{
    Body<D, R> *body = nullptr;
    bool inner = true;
    std::vector<Body<D, R> *> targets;
    // configs[k][i] : inner → k == 0; contact → one per target body
    std::vector<std::vector<Neighborhood<D, R>>> configs;
    std::vector<Kernel<D, R>> kernels;
    std::vector<int> depths;
    int builder = 0; // 0 inner / contact, 1 self contact, 2 surface contact
    std::vector<unsigned char> part; // by original id; empty: all particles
};

// body_relations/inner_body_relation.cpp:14-20; contact_body_relation.cpp:22-32;
// base_body_relation.cpp:10-16 (SearchDepthContact); neighborhood.cpp:75-80 (chooseKernel).
template <int D, class R>
void setupRelation(Relation<D, R> &rel)
{
    rel.kernels.clear();
    rel.depths.clear();
    if (rel.inner)
    {
        rel.kernels.push_back(rel.body->kernel);
        rel.depths.push_back(1);
        rel.configs.resize(1);
        return;
    }
    rel.configs.resize(rel.targets.size());
    for (auto *t : rel.targets)
    {
        if (rel.builder == 2) // NeighborBuilderSurfaceContact, neighborhood.cpp:162-168
            rel.kernels.push_back(Kernel<D, R>(R(0.5) * (rel.body->kernel.h + t->kernel.h)));
        else
            rel.kernels.push_back(rel.body->kernel.h > t->kernel.h ? rel.body->kernel : t->kernel);
        R inv_grid_spacing = R(1.0) / t->mesh.spacing;
        rel.depths.push_back(1 + int(std::floor(rel.body->kernel.rc * inv_grid_spacing)));
    }
}

template <int D, class R>
void updateRelation(Relation<D, R> &rel)
{
    const std::vector<unsigned char> *part = rel.part.empty() ? nullptr : &rel.part;
    if (rel.inner)
        searchNeighbors<D, R>(*rel.body, *rel.body, rel.kernels[0], 1, rel.builder != 1, rel.configs[0], rel.builder, part);
    else
        for (size_t k = 0; k < rel.targets.size(); ++k)
            searchNeighbors<D, R>(*rel.body, *rel.targets[k], rel.kernels[k], rel.depths[k], false, rel.configs[k], rel.builder, part);
}

} // namespace orc
