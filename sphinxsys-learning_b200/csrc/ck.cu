// CK-flavour local dynamics used by the reference's SYCL dam-break tests, as sm_100a kernels:
//   CompressionSummation<Inner / Contact> + DensityRegularization<..., FreeSurface | Internal>
//        shared_ck/particle_dynamics/fluid_dynamics/density_regularization.hpp:42-111, .h:146-180
//   LinearCorrectionMatrix<Inner<WithUpdate>> + <Contact>  (alpha-weighted Tikhonov inverse)
//        shared_ck/particle_dynamics/general_dynamics/kernel_correction_ck.hpp:36-96, common/vector_functions.h:198-203
//   FreeSurfaceIndicationCK<Inner<WithUpdate>> + <Contact>
//        shared_ck/particle_dynamics/general_dynamics/surface_indication/surface_indication_ck.hpp:55-170
// evaluated with the CK smoothing kernel: the 20-interval cubic-Lagrange TABLE of Wendland C2
// (shared_ck/smoothing_kernel/kernel_tabulated_ck.h:37-71, .cpp:6-27; neighbor_method.hpp:17-52,73-120), which differs
// from the analytic kernel of the classic path at the 1e-4 level - hence its own kernels rather than a flag on the
// classic ones.  Neighbours come from the relations' index lists (ensureList), positions are the configuration-time
// ones (the CK tests commit Position once per advection step, right before the relation update).
// The reference's int variables Indicator / PreviousSurfaceIndicator are Real-typed 0 / 1 here (the registry is Real-typed).
#include "internal.cuh"
#include "smat.cuh"

using namespace sphb;

namespace sphb
{
template <class R>
struct TabKernelP
{
    R dq, dq0, dq1, dq2, dq3, dimfac, inv_h, src_inv_h;
    R w[24], dw[24];
};
template <int D, class R>
static TabKernelP<R> makeTabKernel(R h_src, R h_tar)
{
    TabKernelP<R> k;
    k.dq = R(2.0) / R(20);
    for (int i = 0; i < 24; ++i)
    {
        const R q = R(i - 1) * k.dq;
        k.w[i] = R(std::pow(R(1.0) - R(0.5) * q, 4)) * (R(1.0) + R(2.0) * q);
        k.dw[i] = R(0.625) * R(std::pow(q - R(2.0), 3)) * q;
    }
    k.dq0 = (R(-1.0) * k.dq) * (R(-2.0) * k.dq) * (R(-3.0) * k.dq);
    k.dq1 = k.dq * (R(-1.0) * k.dq) * (R(-2.0) * k.dq);
    k.dq2 = (R(2.0) * k.dq) * k.dq * (R(-1.0) * k.dq);
    k.dq3 = (R(3.0) * k.dq) * (R(2.0) * k.dq) * k.dq;
    const R h = h_src, ih = R(1.0) / h;
    k.dimfac = D == 2 ? (ih * ih * R(7.0) / (R(4.0) * R(kPi))) * h * h : (ih * ih * ih * R(21.0) / (R(16.0) * R(kPi))) * h * h * h;
    k.src_inv_h = R(1.0) / h_src;
    k.inv_h = R(1.0) / std::max(h_src, h_tar);
    return k;
}
template <class R>
__device__ __forceinline__ R tabCubic(const TabKernelP<R> &k, const R *data, R q)
{
    const int location = int(floor(q / k.dq));
    const int i = location + 1;
    const R f1 = q - R(location) * k.dq, f0 = f1 + k.dq, f2 = f1 - k.dq, f3 = f1 - R(2) * k.dq;
    return (f1 * f2 * f3) / k.dq0 * data[i - 1] + (f0 * f2 * f3) / k.dq1 * data[i] + (f0 * f1 * f3) / k.dq2 * data[i + 1] +
           (f0 * f1 * f2) / k.dq3 * data[i + 2];
}
template <int D, class R>
__device__ __forceinline__ R ipowD(R x, int extra)
{
    R r = R(1);
#pragma unroll
    for (int k = 0; k < D; ++k)
        r *= x;
    return extra ? r * x : r;
}
template <int D, class R>
__device__ __forceinline__ R tabW0(const TabKernelP<R> &k) { return ipowD<D, R>(k.src_inv_h, 0) * k.dimfac * tabCubic<R>(k, k.w, R(0)); }
template <int D, class R>
__device__ __forceinline__ R tabW(const TabKernelP<R> &k, R r) { return ipowD<D, R>(k.inv_h, 0) * k.dimfac * tabCubic<R>(k, k.w, r * k.inv_h); }
template <int D, class R>
__device__ __forceinline__ R tabdW(const TabKernelP<R> &k, R r) { return ipowD<D, R>(k.inv_h, 1) * k.dimfac * tabCubic<R>(k, k.dw, r * k.inv_h); }

struct CkList
{
    const u32 *count, *slice_off, *list;
};
static CkList ckListOf(const NeighborList &L)
{
    CkList l;
    l.count = L.count;
    l.slice_off = L.slice_off;
    l.list = L.list;
    return l;
}
template <int D, class R>
__device__ __forceinline__ R ckDisp(const R *xi, const typename Real4<R>::type *__restrict__ xV, u32 j, R *d, R &Vj)
{
    const typename Real4<R>::type v = ldg4(xV + j);
    const R xj[3] = {v.x, v.y, v.z};
    Vj = v.w;
#pragma unroll
    for (int a = 0; a < D; ++a)
        d[a] = xi[a] - xj[a];
    return sqrt_r(sqnormExact<D, R>(d));
}
#define CK_INDEX()                                                                                       \
    const int64_t i = (i0 & ~int64_t(31)) + blockIdx.x * int64_t(blockDim.x) + threadIdx.x;              \
    if (i < i0 || i >= n)                                                                                \
        return;                                                                                          \
    R xi[D];                                                                                             \
    {                                                                                                    \
        const typename Real4<R>::type v_ = ldg4(xV + i);                                                 \
        const R x3_[3] = {v_.x, v_.y, v_.z};                                                             \
        for (int a_ = 0; a_ < D; ++a_)                                                                   \
            xi[a_] = x3_[a_];                                                                            \
    }

// ---- CompressionSummation: first = 1 starts the sum with W0 * Vol_ref_i, 0 adds a contact body's part
template <int D, class R>
__global__ void __launch_bounds__(128)
    k_ck_compression(int64_t i0, int64_t n, int first, const typename Real4<R>::type *__restrict__ xV,
                     const typename Real4<R>::type *__restrict__ xV_t, CkList in, TabKernelP<R> k, const R *__restrict__ vref_i,
                     const R *__restrict__ vref_t, R *__restrict__ csum)
{
    CK_INDEX();
    R s = first ? tabW0<D, R>(k) * vref_i[i] : csum[i];
    const u32 cnt = in.count[i];
    const u32 *lst = listBase(in.list, in.slice_off, i);
    for (u32 m = 0; m < cnt; ++m)
    {
        const u32 j = listEntry(lst, m);
        R d[D], Vj;
        const R r = ckDisp<D, R>(xi, xV_t, j, d, Vj);
        s += tabW<D, R>(k, r) * vref_t[j];
    }
    csum[i] = s;
}
template <class R>
__global__ void __launch_bounds__(256)
    k_ck_regularize(int64_t i0, int64_t n, int free_surface, R rho0, const R *__restrict__ csum, R *__restrict__ comp, R *__restrict__ rho)
{
    const int64_t i = i0 + blockIdx.x * int64_t(blockDim.x) + threadIdx.x;
    if (i >= n)
        return;
    const R c = free_surface ? fmax(csum[i], R(1)) : csum[i];
    comp[i] = c;
    rho[i] = c * rho0;
}

// ---- LinearCorrectionMatrix: B -= r_ij (nablaW_ij V_j)^T over inner (first = 1) and contact lists, then the update
template <int D, class R>
__global__ void __launch_bounds__(128)
    k_ck_correction_sum(int64_t i0, int64_t n, int first, const typename Real4<R>::type *__restrict__ xV,
                        const typename Real4<R>::type *__restrict__ xV_t, CkList in, TabKernelP<R> k, R *__restrict__ B)
{
    CK_INDEX();
    SMat<D, R> M = first ? SMat<D, R>::zero() : SMat<D, R>::load(B + i * D * D);
    const u32 cnt = in.count[i];
    const u32 *lst = listBase(in.list, in.slice_off, i);
    for (u32 m = 0; m < cnt; ++m)
    {
        const u32 j = listEntry(lst, m);
        R d[D], Vj;
        const R r = ckDisp<D, R>(xi, xV_t, j, d, Vj);
        const R dw = tabdW<D, R>(k, r);
#pragma unroll
        for (int a = 0; a < D; ++a)
#pragma unroll
            for (int c = 0; c < D; ++c)
                M.a[a][c] -= d[a] * ((dw * (d[c] / r)) * Vj);
    }
    M.store(B + i * D * D);
}
template <int D, class R>
__global__ void __launch_bounds__(128) k_ck_correction_update(int64_t i0, int64_t n, R alpha, R *__restrict__ B)
{
    const int64_t i = i0 + blockIdx.x * int64_t(blockDim.x) + threadIdx.x;
    if (i >= n)
        return;
    SMat<D, R> M = SMat<D, R>::load(B + i * D * D);
    const R det = M.det();
    const R det_sqr = fmax(alpha - det, R(0));
    SMat<D, R> Mt = M.T();
    SMat<D, R> inv = (Mt * M + SMat<D, R>::identity() * R(1.0e-8)).inverse() * Mt; // inverseTikhonov(B, SqrtEps)
    const R weight = det / (det + det_sqr);
    (inv * weight + SMat<D, R>::identity() * (R(1.0) - weight)).store(B + i * D * D);
}

// ---- FreeSurfaceIndicationCK
template <int D, class R>
__global__ void __launch_bounds__(128)
    k_ck_posdiv(int64_t i0, int64_t n, int inner, const typename Real4<R>::type *__restrict__ xV,
                const typename Real4<R>::type *__restrict__ xV_t, CkList in, TabKernelP<R> k, R threshold, const R *__restrict__ prev,
                R *__restrict__ pdiv)
{
    CK_INDEX();
    R pd = R(0);
    bool near_prev = false;
    const u32 cnt = in.count[i];
    const u32 *lst = listBase(in.list, in.slice_off, i);
    for (u32 m = 0; m < cnt; ++m)
    {
        const u32 j = listEntry(lst, m);
        R d[D], Vj;
        const R r = ckDisp<D, R>(xi, xV_t, j, d, Vj);
        pd -= tabdW<D, R>(k, r) * Vj * r;
        if (inner)
            near_prev = near_prev || prev[j] == R(1);
    }
    if (inner)
    {
        if (pd < threshold && prev[i] != R(1) && !near_prev)
            pd = R(2.0) * threshold;
        pdiv[i] = pd;
    }
    else
        pdiv[i] += pd;
}
template <int D, class R>
__global__ void __launch_bounds__(128)
    k_ck_indicator(int64_t i0, int64_t n, const typename Real4<R>::type *__restrict__ xV, CkList in, R threshold, R h,
                   const R *__restrict__ pdiv, R *__restrict__ out)
{
    CK_INDEX();
    R v = R(1);
    if (pdiv[i] > threshold)
    {
        bool very_near = false;
        const u32 cnt = in.count[i];
        const u32 *lst = listBase(in.list, in.slice_off, i);
        for (u32 m = 0; m < cnt && !very_near; ++m)
        {
            const u32 j = listEntry(lst, m);
            R d[D], Vj;
            const R r = ckDisp<D, R>(xi, xV, j, d, Vj);
            very_near = pdiv[j] < threshold && r < h;
        }
        if (!very_near)
            v = R(0);
    }
    out[i] = v;
}
template <class R>
__global__ void __launch_bounds__(256) k_ck_fill(int64_t n, R value, R *__restrict__ a)
{
    const int64_t i = blockIdx.x * int64_t(blockDim.x) + threadIdx.x;
    if (i < n)
        a[i] = value;
}

static sphb200_status ckCheck(sphb200_relation *r, bool want_inner, const char *who)
{
    sphb200_ctx *ctx = r->ctx;
    if (r->inner != want_inner)
        return ctx->fail(SPHB200_ERR_INVALID, std::string(who) + ": wrong relation kind");
    if (!r->valid || r->source_stamp != r->body->config_stamp)
        return ctx->fail(SPHB200_ERR_INVALID, std::string(who) + ": relation is stale (call relation_update after cell_list_update)");
    for (size_t k = 0; k < r->targets.size(); ++k)
        if (r->target_stamp[k] != r->targets[k]->config_stamp)
            return ctx->fail(SPHB200_ERR_INVALID, std::string(who) + ": relation target was re-sorted; call relation_update");
    return SPHB200_OK;
}
// registerStateVariableFrom<Real>("VolumetricMeasureRef", "VolumetricMeasure"): a copy taken when first registered
template <class R>
static sphb200_status volRef(sphb200_body *b, R **out)
{
    sphb200_ctx *ctx = b->ctx;
    const bool fresh = b->find("VolumetricMeasureRef") == nullptr;
    SPHB_TRY(getVarPtr<R>(b, "VolumetricMeasureRef", 1, out, true));
    if (fresh && b->n > 0)
    {
        R *vol;
        SPHB_TRY(getVarPtr<R>(b, "VolumetricMeasure", 1, &vol, true));
        SPHB_CUDA(ctx, cudaMemcpyAsync(*out, vol, size_t(b->n) * sizeof(R), cudaMemcpyDeviceToDevice, ctx->stream));
    }
    return SPHB200_OK;
}
#define CK_GRID(b) gridFor((b)->own_end - ((b)->own_begin & ~int64_t(31)), 128), 128, 0, ctx->stream

template <int D, class R>
static sphb200_status ckDensityRegularization(sphb200_relation *inner, sphb200_relation *contact, int free_surface)
{
    sphb200_ctx *ctx = inner->ctx;
    sphb200_body *b = inner->body;
    SPHB_TRY(ckCheck(inner, true, "ck_density_regularization"));
    if (contact)
        SPHB_TRY(ckCheck(contact, false, "ck_density_regularization"));
    if (b->own_end <= b->own_begin)
        return SPHB200_OK;
    SPHB_TRY((ensurePacked<D, R>(b)));
    R *vref, *csum, *comp, *rho;
    SPHB_TRY(volRef<R>(b, &vref));
    SPHB_TRY(getVarPtr<R>(b, "CompressionSummation", 1, &csum));
    SPHB_TRY(getVarPtr<R>(b, "Compression", 1, &comp));
    SPHB_TRY(getVarPtr<R>(b, "Density", 1, &rho, true));
    SPHB_TRY(ensureList(inner, 0));
    const R h = b->P<D, R>().h;
    auto *xV = static_cast<const typename Real4<R>::type *>(b->xV.p);
    k_ck_compression<D, R><<<CK_GRID(b)>>>(b->own_begin, b->own_end, 1, xV, xV, ckListOf(inner->lists[0]), makeTabKernel<D, R>(h, h), vref, vref, csum);
    SPHB_LAUNCH_CHECK(ctx);
    for (size_t t = 0; contact && t < contact->targets.size(); ++t)
    {
        sphb200_body *tb = contact->targets[t];
        SPHB_TRY((ensurePacked<D, R>(tb)));
        R *tref;
        SPHB_TRY(volRef<R>(tb, &tref));
        SPHB_TRY(ensureList(contact, t));
        k_ck_compression<D, R><<<CK_GRID(b)>>>(b->own_begin, b->own_end, 0, xV, static_cast<const typename Real4<R>::type *>(tb->xV.p),
                                               ckListOf(contact->lists[t]), makeTabKernel<D, R>(h, tb->P<D, R>().h), vref, tref, csum);
        SPHB_LAUNCH_CHECK(ctx);
    }
    k_ck_regularize<R><<<gridFor(b->own_end - b->own_begin, 256), 256, 0, ctx->stream>>>(b->own_begin, b->own_end, free_surface, R(b->rho0), csum, comp, rho);
    SPHB_LAUNCH_CHECK(ctx);
    b->rec_b_valid = false;
    return SPHB200_OK;
}

template <int D, class R>
static sphb200_status ckCorrectionMatrix(sphb200_relation *inner, sphb200_relation *contact, double alpha)
{
    sphb200_ctx *ctx = inner->ctx;
    sphb200_body *b = inner->body;
    SPHB_TRY(ckCheck(inner, true, "ck_linear_correction_matrix"));
    if (contact)
        SPHB_TRY(ckCheck(contact, false, "ck_linear_correction_matrix"));
    if (b->own_end <= b->own_begin)
        return SPHB200_OK;
    SPHB_TRY((ensurePacked<D, R>(b)));
    R *B;
    SPHB_TRY(getVarPtr<R>(b, "LinearCorrectionMatrix", D * D, &B));
    SPHB_TRY(ensureList(inner, 0));
    const R h = b->P<D, R>().h;
    auto *xV = static_cast<const typename Real4<R>::type *>(b->xV.p);
    k_ck_correction_sum<D, R><<<CK_GRID(b)>>>(b->own_begin, b->own_end, 1, xV, xV, ckListOf(inner->lists[0]), makeTabKernel<D, R>(h, h), B);
    SPHB_LAUNCH_CHECK(ctx);
    for (size_t t = 0; contact && t < contact->targets.size(); ++t)
    {
        sphb200_body *tb = contact->targets[t];
        SPHB_TRY((ensurePacked<D, R>(tb)));
        SPHB_TRY(ensureList(contact, t));
        k_ck_correction_sum<D, R><<<CK_GRID(b)>>>(b->own_begin, b->own_end, 0, xV, static_cast<const typename Real4<R>::type *>(tb->xV.p),
                                                  ckListOf(contact->lists[t]), makeTabKernel<D, R>(h, tb->P<D, R>().h), B);
        SPHB_LAUNCH_CHECK(ctx);
    }
    k_ck_correction_update<D, R><<<gridFor(b->own_end - b->own_begin, 128), 128, 0, ctx->stream>>>(b->own_begin, b->own_end, R(alpha), B);
    SPHB_LAUNCH_CHECK(ctx);
    return SPHB200_OK;
}

template <int D, class R>
static sphb200_status ckSurfaceIndication(sphb200_relation *inner, sphb200_relation *contact)
{
    sphb200_ctx *ctx = inner->ctx;
    sphb200_body *b = inner->body;
    SPHB_TRY(ckCheck(inner, true, "ck_free_surface_indication"));
    if (contact)
        SPHB_TRY(ckCheck(contact, false, "ck_free_surface_indication"));
    if (b->own_end <= b->own_begin)
        return SPHB200_OK;
    SPHB_TRY((ensurePacked<D, R>(b)));
    const bool fresh = b->find("PreviousSurfaceIndicator") == nullptr;
    R *ind, *pdiv, *prev;
    SPHB_TRY(getVarPtr<R>(b, "Indicator", 1, &ind));
    SPHB_TRY(getVarPtr<R>(b, "PositionDivergence", 1, &pdiv));
    SPHB_TRY(getVarPtr<R>(b, "PreviousSurfaceIndicator", 1, &prev, true));
    if (fresh)
    {
        k_ck_fill<R><<<gridFor(b->n, 256), 256, 0, ctx->stream>>>(b->n, R(1), prev);
        SPHB_LAUNCH_CHECK(ctx);
    }
    SPHB_TRY(ensureList(inner, 0));
    const R h = b->P<D, R>().h, threshold = R(0.75) * R(D);
    auto *xV = static_cast<const typename Real4<R>::type *>(b->xV.p);
    k_ck_posdiv<D, R><<<CK_GRID(b)>>>(b->own_begin, b->own_end, 1, xV, xV, ckListOf(inner->lists[0]), makeTabKernel<D, R>(h, h), threshold, prev, pdiv);
    SPHB_LAUNCH_CHECK(ctx);
    for (size_t t = 0; contact && t < contact->targets.size(); ++t)
    {
        sphb200_body *tb = contact->targets[t];
        SPHB_TRY((ensurePacked<D, R>(tb)));
        SPHB_TRY(ensureList(contact, t));
        k_ck_posdiv<D, R><<<CK_GRID(b)>>>(b->own_begin, b->own_end, 0, xV, static_cast<const typename Real4<R>::type *>(tb->xV.p),
                                          ckListOf(contact->lists[t]), makeTabKernel<D, R>(h, tb->P<D, R>().h), threshold, prev, pdiv);
        SPHB_LAUNCH_CHECK(ctx);
    }
    k_ck_indicator<D, R><<<CK_GRID(b)>>>(b->own_begin, b->own_end, xV, ckListOf(inner->lists[0]), threshold, h, pdiv, ind);
    SPHB_LAUNCH_CHECK(ctx);
    SPHB_CUDA(ctx, cudaMemcpyAsync(prev + b->own_begin, ind + b->own_begin, size_t(b->own_end - b->own_begin) * sizeof(R), cudaMemcpyDeviceToDevice,
                                   ctx->stream));
    return SPHB200_OK;
}
} // namespace sphb

extern "C"
{
sphb200_status sphb200_ck_density_regularization(sphb200_relation *inner, sphb200_relation *contact, int free_surface)
{
    if (!inner || (contact && contact->ctx != inner->ctx))
        return SPHB200_ERR_INVALID;
    cudaSetDevice(inner->ctx->device);
    return SPHB_DISPATCH(inner->ctx, ckDensityRegularization, inner, contact, free_surface);
}
sphb200_status sphb200_ck_linear_correction_matrix(sphb200_relation *inner, sphb200_relation *contact, double alpha)
{
    if (!inner || (contact && contact->ctx != inner->ctx))
        return SPHB200_ERR_INVALID;
    cudaSetDevice(inner->ctx->device);
    return SPHB_DISPATCH(inner->ctx, ckCorrectionMatrix, inner, contact, alpha);
}
sphb200_status sphb200_ck_free_surface_indication(sphb200_relation *inner, sphb200_relation *contact)
{
    if (!inner || (contact && contact->ctx != inner->ctx))
        return SPHB200_ERR_INVALID;
    cudaSetDevice(inner->ctx->device);
    return SPHB_DISPATCH(inner->ctx, ckSurfaceIndication, inner, contact);
}
} // extern "C"
