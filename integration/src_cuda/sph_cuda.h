// integration/src_cuda - the CUDA execution policy of SPHinXsys' Dynamics API over libsphb200 (include/sphb200.h):
// what a maintainer adds next to src/src_sycl/.  It supplies exactly the customisation points the SYCL back end
// supplies:
//   * a DeviceExecution-derived policy tag                         execution_policy.h:53-79
//   * DiscreteVariable<T>::DelegatedOnDevice / synchronizeWithDevice / synchronizeToDevice
//                                                                  sphinxsys_variable.h:179-190, sphinxsys_variable_sycl.hpp:35-87
//   * the spatial operators of a device build (cell linked list + sort, relations)
//                                                                  shared_ck/.../update_cell_linked_list.hpp, update_body_relation.hpp
//   * the per-local-dynamics fused implementations                 DeviceImplementation<ParallelCudaPolicy, LocalDynamics>
//   * observer interpolation and the asynchronous output path      general_interpolation.h:63-83, sphinxsys_variable.h:350-356
// Everything else (non-hot local dynamics) runs through the generic fall-back of the algorithm templates:
// synchronizeWithDevice -> host loop -> synchronizeToDevice.
// Compiled against integration/skeleton here (the reference's own headers need Eigen/TBB/Boost/Simbody); the names,
// template signatures and members used below are the reference's.
#pragma once
#include "sphb200.h"
#include "sph_io.h"

namespace SPH
{
namespace execution
{
class ParallelCudaPolicy : public DeviceExecution<ParallelPolicy> {};
inline constexpr auto par_cuda = ParallelCudaPolicy{};
} // namespace execution

namespace cuda
{
// reference convention on errors: print and exit(1)
inline void check(sphb200_status s, sphb200_ctx *ctx, const char *what)
{
    if (s == SPHB200_OK)
        return;
    std::cout << "\n Error: " << what << ": " << (ctx ? sphb200_last_error(ctx) : "no context")
              << (s == SPHB200_ERR_NO_DEVICE ? " (no CUDA device; there is no CPU fallback)" : "") << std::endl;
    std::exit(1);
}

// one context per SPHSystem (per GPU / rank) - the analogue of src_sycl's ExecutionInstance
inline sphb200_ctx *context(SPHSystem &system)
{
    if (!system.device_context_)
    {
        const BoundingBoxd &b = system.getSystemDomainBounds(); // already expanded by 4 dp (sph_system.cpp:39)
        double lo[3] = {0, 0, 0}, hi[3] = {0, 0, 0};
        for (int d = 0; d < Dimensions; ++d)
            lo[d] = double(b.lower_[d]), hi[d] = double(b.upper_[d]);
        sphb200_ctx *ctx = nullptr;
        check(sphb200_ctx_create(0, Dimensions, int(sizeof(Real)), lo, hi, &ctx), ctx, "SPHSystem");
        system.device_context_ = ctx;
    }
    return static_cast<sphb200_ctx *>(system.device_context_);
}
} // namespace cuda

// what the CUDA back end attaches to a DiscreteVariable
struct DeviceBinding
{
    SPHBody *body = nullptr;
    int width = 1;
    bool on_device = false; // registered with the library and uploaded once
};

namespace cuda
{
inline sphb200_body *bodyHandle(SPHBody &body)
{
    if (!body.device_body_)
    {
        sphb200_ctx *ctx = context(body.getSPHSystem());
        sphb200_body *h = nullptr;
        check(sphb200_body_create(ctx, body.getName().c_str(), (long long)body.getBaseParticles().TotalRealParticles(),
                                  double(body.getSPHAdaptation().ReferenceSpacing()), double(body.getSPHAdaptation().SmoothingLengthRatio()), &h),
              ctx, "SPHBody");
        body.device_body_ = h;
        if (auto *e = dynamic_cast<NeoHookeanSolid *>(&body.getBaseMaterial()))
            check(sphb200_body_set_elastic(h, double(e->ReferenceDensity()), double(e->YoungsModulus()), double(e->PoissonRatio())), ctx,
                  "defineMatterMaterial");
        else if (auto *f = dynamic_cast<Fluid *>(&body.getBaseMaterial()))
            check(sphb200_body_set_fluid(h, double(f->ReferenceDensity()), double(f->ReferenceSoundSpeed()), double(body.ReferenceViscosity())), ctx,
                  "defineMatterMaterial");
        else
            check(sphb200_body_set_fluid(h, double(body.getBaseMaterial().ReferenceDensity()), 1.0, 0.0), ctx, "defineMatterMaterial");
    }
    return static_cast<sphb200_body *>(body.device_body_);
}

// every variable of the body that is registered on the host but not yet with the library: register + upload
template <typename DataType>
inline bool pushIf(BaseVariable *v, sphb200_body *h, sphb200_ctx *ctx, bool to_host)
{
    auto *dv = dynamic_cast<DiscreteVariable<DataType> *>(v);
    if (!dv)
        return false;
    if (to_host)
        dv->synchronizeWithDevice();
    else
        dv->synchronizeToDevice();
    return true;
}
inline void syncAll(SPHBody &body, bool to_host)
{
    sphb200_body *h = bodyHandle(body);
    sphb200_ctx *ctx = context(body.getSPHSystem());
    for (const std::string &name : body.getBaseParticles().VariableNames())
    {
        BaseVariable *v = body.getBaseParticles().findVariable(name);
        if (!pushIf<Real>(v, h, ctx, to_host))
            pushIf<Vecd>(v, h, ctx, to_host);
    }
}
inline void ensureOnDevice(SPHBody &body)
{
    for (const std::string &name : body.getBaseParticles().VariableNames())
    {
        BaseVariable *v = body.getBaseParticles().findVariable(name);
        if (auto *r = dynamic_cast<DiscreteVariable<Real> *>(v))
            r->DelegatedOnDevice();
        else if (auto *x = dynamic_cast<DiscreteVariable<Vecd> *>(v))
            x->DelegatedOnDevice();
    }
}
// installs the back end's hooks on a body's particles (the in-tree form calls them from BaseParticles directly)
inline void attach(SPHBody &body)
{
    BaseParticles &p = body.getBaseParticles();
    if (p.on_register_)
        return;
    p.on_register_ = [&body](BaseVariable *v, const std::string &, int width)
    {
        auto b = std::make_shared<DeviceBinding>();
        b->body = &body, b->width = width;
        if (auto *r = dynamic_cast<DiscreteVariable<Real> *>(v))
            r->bindDevice(b);
        else if (auto *x = dynamic_cast<DiscreteVariable<Vecd> *>(v))
            x->bindDevice(b);
    };
    p.sync_all_ = [&body](bool to_host) { syncAll(body, to_host); };
}
} // namespace cuda

// ---------------------------------------------------------------------------------------------------------
// DiscreteVariable device delegation (the three customisation points)
// ---------------------------------------------------------------------------------------------------------
template <typename DataType>
DataType *DiscreteVariable<DataType>::DelegatedOnDevice()
{
    DeviceBinding *b = deviceBinding();
    if (!b)
    {
        std::cout << "\n Error: variable '" << name_ << "' is not attached to a CUDA body" << std::endl;
        std::exit(1);
    }
    sphb200_body *h = cuda::bodyHandle(*b->body);
    sphb200_ctx *ctx = cuda::context(b->body->getSPHSystem());
    if (!b->on_device)
    { // the host array stays owned by this DiscreteVariable; the library borrows the pointer and owns the device mirror
        cuda::check(sphb200_var_register(h, name_.c_str(), b->width, isEvolving() ? 1 : 0, data_field_), ctx, "DelegatedOnDevice");
        if (b->body->getBaseParticles().TotalRealParticles() > 0)
            cuda::check(sphb200_var_to_device(h, name_.c_str()), ctx, "DelegatedOnDevice");
        b->on_device = true;
    }
    void *p = nullptr;
    cuda::check(sphb200_var_device_ptr(h, name_.c_str(), &p), ctx, "DelegatedOnDevice");
    return static_cast<DataType *>(p);
}
template <typename DataType>
void DiscreteVariable<DataType>::synchronizeWithDevice()
{
    DeviceBinding *b = deviceBinding();
    if (b && b->on_device && b->body->getBaseParticles().TotalRealParticles() > 0)
        cuda::check(sphb200_var_to_host(cuda::bodyHandle(*b->body), name_.c_str()), cuda::context(b->body->getSPHSystem()), "synchronizeWithDevice");
}
template <typename DataType>
void DiscreteVariable<DataType>::synchronizeToDevice()
{
    DeviceBinding *b = deviceBinding();
    if (!b)
        return;
    if (!b->on_device)
    {
        DelegatedOnDevice(); // registers and uploads
        return;
    }
    if (b->body->getBaseParticles().TotalRealParticles() > 0)
        cuda::check(sphb200_var_to_device(cuda::bodyHandle(*b->body), name_.c_str()), cuda::context(b->body->getSPHSystem()), "synchronizeToDevice");
}

// ---------------------------------------------------------------------------------------------------------
// spatial operators: cell linked list (+ sort) and relations
// ---------------------------------------------------------------------------------------------------------
inline void RealBody::updateCellLinkedList()
{
    cuda::ensureOnDevice(*this);
    cuda::check(sphb200_cell_list_update(cuda::bodyHandle(*this)), cuda::context(getSPHSystem()), "updateCellLinkedList");
}
namespace cuda
{
inline sphb200_relation *relationHandle(BaseInnerRelation &r)
{
    if (!r.device_relation_)
    {
        sphb200_relation *h = nullptr;
        check(sphb200_relation_inner_create(bodyHandle(r.getSPHBody()), &h), context(r.getSPHBody().getSPHSystem()), "InnerRelation");
        r.device_relation_ = h;
    }
    return static_cast<sphb200_relation *>(r.device_relation_);
}
inline sphb200_relation *relationHandle(BaseContactRelation &r)
{
    if (!r.device_relation_)
    {
        std::vector<sphb200_body *> t;
        sphb200_ctx *ctx = context(r.getSPHBody().getSPHSystem());
        double src[12];
        check(sphb200_body_info(bodyHandle(r.getSPHBody()), src), ctx, "ContactRelation");
        for (RealBody *b : r.getContactBodies())
        {
            sphb200_body *tb = bodyHandle(*b);
            // a target with a smaller kernel than the source (SolidBody wall, h = 1.15 dp, under a fluid with 1.3 dp): the
            // reference searches a deeper block of its small cells (SearchDepthContact); the library widens the cells
            // instead - same neighbour set, single-pass masked search.  One re-binning of the target, once.
            double tar[12];
            check(sphb200_body_info(tb, tar), ctx, "ContactRelation");
            if (tar[7] < src[1] && dynamic_cast<RealBody *>(&r.getSPHBody()))
            {
                check(sphb200_body_set_cell_spacing(tb, src[1]), ctx, "ContactRelation: cell spacing of the target");
                check(sphb200_cell_list_update(tb), ctx, "ContactRelation: re-binning the target");
            }
            t.push_back(tb);
        }
        sphb200_relation *h = nullptr;
        check(sphb200_relation_contact_create(bodyHandle(r.getSPHBody()), t.data(), int(t.size()), &h), context(r.getSPHBody().getSPHSystem()),
              "ContactRelation");
        r.device_relation_ = h;
    }
    return static_cast<sphb200_relation *>(r.device_relation_);
}
inline void ensureCellList(SPHBody &body)
{ // an ObserverBody is not a RealBody in the reference (no cell list of its own); the library bins every relation source
    if (!body.device_body_ || !static_cast<bool>(body.device_cell_list_ready_))
    {
        ensureOnDevice(body);
        check(sphb200_cell_list_update(bodyHandle(body)), context(body.getSPHSystem()), "observer cell list");
        body.device_cell_list_ready_ = true;
    }
}
} // namespace cuda
inline void InnerRelation::updateConfiguration()
{
    cuda::check(sphb200_relation_update(cuda::relationHandle(*this)), cuda::context(getSPHBody().getSPHSystem()), "InnerRelation::updateConfiguration");
}
inline void ContactRelation::updateConfiguration()
{
    if (!dynamic_cast<RealBody *>(&getSPHBody()))
        cuda::ensureCellList(getSPHBody());
    cuda::check(sphb200_relation_update(cuda::relationHandle(*this)), cuda::context(getSPHBody().getSPHSystem()), "ContactRelation::updateConfiguration");
}
} // namespace SPH
