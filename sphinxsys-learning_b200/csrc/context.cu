// Context, bodies and the named-variable registry behind the C ABI (include/sphb200.h).
// Mirrors, for the hot path only: SPHSystem bounds (sphinxsys_system/sph_system.cpp:36-51),
// SPHBody + SPHAdaptation (bodies/base_body.h:62-201, adaptations/adaptation.cpp:12-60),
// BaseParticles' variable registry (particles/base_particles.h:119-162) and the device delegation of
// DiscreteVariable (common/sphinxsys_variable.h:179-190).
#include "internal.cuh"
#include <algorithm>

using namespace sphb;

namespace sphb
{
cudaError_t DevBuf::ensure(size_t need, bool keep, cudaStream_t st)
{
    if (need <= bytes && p)
        return cudaSuccess;
    size_t cap = need + need / 8 + 256;
    void *np = nullptr;
    cudaError_t e = cudaMalloc(&np, cap);
    if (e != cudaSuccess)
        return e;
    if (keep && p && bytes)
    {
        e = cudaMemcpyAsync(np, p, bytes, cudaMemcpyDeviceToDevice, st);
        if (e != cudaSuccess)
            return e;
        cudaStreamSynchronize(st);
    }
    if (p)
        cudaFree(p);
    p = np;
    bytes = cap;
    return cudaSuccess;
}
void DevBuf::release()
{
    if (p)
        cudaFree(p);
    p = nullptr;
    bytes = 0;
}

sphb200_status getVar(sphb200_body *b, const char *name, int width, bool evolving, Var **out)
{
    Var *v = b->find(name);
    if (!v)
    {
        SPHB_TRY(sphb200_var_register(b, name, width, evolving ? 1 : 0, nullptr));
        v = b->find(name);
    }
    if (v->width != width)
        return b->ctx->fail(SPHB200_ERR_INVALID, std::string("variable '") + name + "' registered with width " +
                                                     std::to_string(v->width) + ", requested " + std::to_string(width));
    // a dynamics that needs the field to follow its particle through the cell-order sort gets that, whoever registered it
    if (evolving && !v->evolving)
        v->evolving = true;
    *out = v;
    return SPHB200_OK;
}

sphb200_status reduceReset(sphb200_ctx *ctx, int slot, bool for_min)
{
    unsigned long long *dst = static_cast<unsigned long long *>(ctx->reduce_dev.p) + slot;
    if (!for_min)
    {
        SPHB_CUDA(ctx, cudaMemsetAsync(dst, 0, 8, ctx->stream));
        return SPHB200_OK;
    }
    // +inf comes from the last word of the pinned result line (a pageable source could make the copy synchronous)
    unsigned long long *inf = static_cast<unsigned long long *>(ctx->reduce_host) + 63;
    *inf = 0x7ff0000000000000ull;
    SPHB_CUDA(ctx, cudaMemcpyAsync(dst, inf, 8, cudaMemcpyHostToDevice, ctx->stream));
    return SPHB200_OK;
}
sphb200_status reduceRead(sphb200_ctx *ctx, int slot, double *out)
{
    SPHB_CUDA(ctx, cudaMemcpyAsync(ctx->reduce_host, static_cast<unsigned long long *>(ctx->reduce_dev.p) + slot, 8,
                                   cudaMemcpyDeviceToHost, ctx->stream));
    SPHB_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    *out = *static_cast<double *>(ctx->reduce_host);
    return SPHB200_OK;
}

// host-side Wendland C2 (exactly the reference's expression order; used for W0 / sigma0 only)
template <int D, class R>
static KernelP<R> makeKernel(R h)
{
    KernelP<R> k;
    k.h = h;
    k.inv_h = R(1.0) / h;
    k.rc = R(2.0) * h; // kernel_wendland_c2.cpp:8-15 kernel_size_ = 2
    k.rc2 = k.rc * k.rc;
    if (D == 2)
        k.factor_W = k.inv_h * k.inv_h * R(7.0) / (R(4.0) * R(kPi));
    else
        k.factor_W = k.inv_h * k.inv_h * k.inv_h * R(21.0) / (R(16.0) * R(kPi));
    k.factor_dW = k.inv_h * k.factor_W;
    return k;
}
template <class R>
static R hostW(const KernelP<R> &k, R r)
{
    R q = r * k.inv_h;
    return k.factor_W * (R(std::pow(R(1.0) - R(0.5) * q, 4)) * (R(1.0) + R(2.0) * q));
}
// adaptations/adaptation.cpp:26-60 computeLatticeNumberDensity
template <int D, class R>
static R latticeSigma0(const KernelP<R> &k, R dp)
{
    R sigma = 0;
    int depth = int(k.rc / dp) + 1;
    int kz0 = D == 3 ? -depth : 0, kz1 = D == 3 ? depth : 0;
    for (int kk = kz0; kk <= kz1; ++kk)
        for (int j = -depth; j <= depth; ++j)
            for (int i = -depth; i <= depth; ++i)
            {
                R x = i * dp, y = j * dp, z = kk * dp;
                R dist = D == 3 ? std::sqrt(x * x + y * y + z * z) : std::sqrt(x * x + y * y);
                if (dist < k.rc)
                    sigma += hostW(k, dist);
            }
    return sigma;
}

template <int D, class R>
static sphb200_status setMeshSpacing(sphb200_body *b, double spacing);

template <int D, class R>
static sphb200_status initBodyParams(sphb200_body *b)
{
    BodyP<D, R> &P = b->P<D, R>();
    P.dp = R(b->dp);
    P.h = R(b->h_ratio) * R(b->dp);
    P.kern = makeKernel<D, R>(P.h);
    P.sigma0 = latticeSigma0<D, R>(P.kern, P.dp);
    SPHB_TRY((setMeshSpacing<D, R>(b, double(P.kern.rc))));
    P.per.mask = 0;
    for (int d = 0; d < D; ++d)
        P.per.lo[d] = P.per.hi[d] = P.per.L[d] = P.per.half_L[d] = R(0);
    return SPHB200_OK;
}

// Mesh ctor, base_mesh.cpp:6-16 with buffer_width = 2; the reference's spacing is the body's own cut-off radius
// (cell_linked_list.cpp:84-101 via the adaptation), sphb200_body_set_cell_spacing may choose a larger one
template <int D, class R>
static sphb200_status setMeshSpacing(sphb200_body *b, double spacing)
{
    sphb200_ctx *ctx = b->ctx;
    BodyP<D, R> &P = b->P<D, R>();
    P.mesh.spacing = R(spacing);
    u64 total = 1;
    for (int d = 0; d < D; ++d)
    {
        R blo = R(ctx->sys_lo[d]), bhi = R(ctx->sys_hi[d]);
        R buffer = R(2) * P.mesh.spacing;
        P.mesh.lo[d] = blo - buffer;
        R extent = bhi + buffer - P.mesh.lo[d];
        int grid_points = int(std::ceil(extent / P.mesh.spacing)) + 1;
        P.mesh.cells[d] = grid_points - 1;
        total *= u64(P.mesh.cells[d]);
    }
    if (total >= (u64(1) << 31))
        return ctx->fail(SPHB200_ERR_INVALID, "cell-linked-list mesh has more than 2^31 cells");
    b->n_cells = int64_t(total);
    return SPHB200_OK;
}

template <int D, class R>
static sphb200_status setPeriodic(sphb200_body *b, int axis, double lo, double hi)
{
    BodyP<D, R> &P = b->P<D, R>();
    P.per.mask |= (1 << axis);
    P.per.lo[axis] = R(lo);
    P.per.hi[axis] = R(hi);
    P.per.L[axis] = P.per.hi[axis] - P.per.lo[axis];
    P.per.half_L[axis] = R(0.5) * P.per.L[axis];
    return SPHB200_OK;
}

template <int D, class R>
static sphb200_status bodyInfo(sphb200_body *b, double *out)
{
    BodyP<D, R> &P = b->P<D, R>();
    out[0] = P.h;
    out[1] = P.kern.rc;
    out[2] = P.kern.factor_W;
    out[3] = P.sigma0;
    for (int d = 0; d < 3; ++d)
    {
        out[4 + d] = d < D ? double(P.mesh.lo[d]) : 0.0;
        out[8 + d] = d < D ? double(P.mesh.cells[d]) : 1.0;
    }
    out[7] = P.mesh.spacing;
    out[11] = double(b->n);
    return SPHB200_OK;
}

__global__ void k_iota(u32 *a, u32 *b, int64_t n)
{
    int64_t i = blockIdx.x * int64_t(blockDim.x) + threadIdx.x;
    if (i < n)
    {
        a[i] = u32(i);
        b[i] = u32(i);
    }
}
} // namespace sphb

extern "C"
{
sphb200_status sphb200_ctx_create(int device, int ndim, int real_bytes, const double *sys_lo, const double *sys_hi,
                                  sphb200_ctx **out)
{
    if (!out)
        return SPHB200_ERR_INVALID;
    *out = nullptr;
    if ((ndim != 2 && ndim != 3) || (real_bytes != 4 && real_bytes != 8) || !sys_lo || !sys_hi)
        return SPHB200_ERR_INVALID;
    int count = 0;
    if (cudaGetDeviceCount(&count) != cudaSuccess || count == 0 || device < 0 || device >= count)
    {
        cudaGetLastError();
        return SPHB200_ERR_NO_DEVICE; // no CPU fallback by design
    }
    if (cudaSetDevice(device) != cudaSuccess)
        return SPHB200_ERR_CUDA;
    auto *ctx = new sphb200_ctx();
    ctx->device = device;
    ctx->ndim = ndim;
    ctx->real_bytes = real_bytes;
    for (int d = 0; d < ndim; ++d)
    {
        ctx->sys_lo[d] = sys_lo[d];
        ctx->sys_hi[d] = sys_hi[d];
    }
    if (cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking) != cudaSuccess ||
        ctx->reduce_dev.ensure(64 * 8) != cudaSuccess || cudaMallocHost(&ctx->reduce_host, 64 * 8) != cudaSuccess)
    {
        delete ctx;
        return SPHB200_ERR_CUDA;
    }
    cudaMemsetAsync(ctx->reduce_dev.p, 0, 64 * 8, ctx->stream);
    *out = ctx;
    return SPHB200_OK;
}

void sphb200_ctx_destroy(sphb200_ctx *ctx)
{
    if (!ctx)
        return;
    cudaSetDevice(ctx->device);
    cudaStreamSynchronize(ctx->stream);
    if (ctx->copy_stream)
        cudaStreamSynchronize(ctx->copy_stream);
    cudaFree(ctx->peer_words);
    for (auto *r : ctx->relations)
    {
        for (auto &l : r->lists)
        {
            cudaFree(l.count);
            cudaFree(l.slice_w);
            cudaFree(l.slice_off);
            cudaFree(l.list);
            cudaFree(l.mask);
            cudaFree(l.uni.count);
            cudaFree(l.uni.slice_w);
            cudaFree(l.uni.slice_off);
            cudaFree(l.uni.list);
            cudaFree(l.uni.queue);
            cudaFree(l.uni_wide.queue);
            cudaFree(l.uni_wide.count);
            cudaFree(l.uni_wide.slice_w);
            cudaFree(l.uni_wide.slice_off);
            cudaFree(l.uni_wide.list);
        }
        delete r;
    }
    for (auto *b : ctx->bodies)
    {
        for (auto &kv : b->vars)
        {
            kv.second->dev.release();
            kv.second->alt.release();
            kv.second->stage.release();
        }
        for (DevBuf *d : {&b->original_id, &b->sorted_id, &b->cell_of, &b->rank_in_cell, &b->perm, &b->cell_count,
                          &b->cell_offset, &b->xV, &b->tmp_var, &b->rec_a, &b->rec_b, &b->vol_mm, &b->xf, &b->wall_sum})
            d->release();
        delete b;
    }
    for (DevBuf *d : {&ctx->scratch_a, &ctx->scratch_b, &ctx->scan_tmp, &ctx->reduce_dev, &ctx->subkey_buf})
        d->release();
    if (ctx->reduce_host)
        cudaFreeHost(ctx->reduce_host);
    if (ctx->copy_stream)
    {
        cudaStreamSynchronize(ctx->copy_stream);
        cudaStreamDestroy(ctx->copy_stream);
        cudaEventDestroy(ctx->copy_done);
        cudaEventDestroy(ctx->mirrors_free);
    }
    if (ctx->owns_stream)
        cudaStreamDestroy(ctx->stream);
    delete ctx;
}

sphb200_status sphb200_ctx_synchronize(sphb200_ctx *ctx)
{
    if (!ctx)
        return SPHB200_ERR_INVALID;
    SPHB_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return SPHB200_OK;
}
const char *sphb200_last_error(sphb200_ctx *ctx) { return ctx ? ctx->error.c_str() : "null context"; }
void *sphb200_ctx_stream(sphb200_ctx *ctx) { return ctx ? (void *)ctx->stream : nullptr; }
int64_t sphb200_ctx_launch_count(sphb200_ctx *ctx) { return ctx ? ctx->launches : 0; }

sphb200_status sphb200_body_create(sphb200_ctx *ctx, const char *name, int64_t n, double dp, double h_ratio,
                                   sphb200_body **out)
{
    if (!ctx || !out || !name)
        return SPHB200_ERR_INVALID;
    if (n < 0 || n >= (int64_t(1) << 31) || !(dp > 0) || !(h_ratio > 0))
        return ctx->fail(SPHB200_ERR_INVALID, "body_create: bad n / dp / h_ratio");
    cudaSetDevice(ctx->device);
    auto *b = new sphb200_body();
    b->ctx = ctx;
    b->name = name;
    b->n = n;
    b->own_end = n;
    b->dp = dp;
    b->h_ratio = h_ratio;
    memset(b->params, 0, sizeof(b->params));
    sphb200_status s = SPHB_DISPATCH(ctx, initBodyParams, b);
    if (s != SPHB200_OK)
    {
        delete b;
        return s;
    }
    size_t idb = size_t(n > 0 ? n : 1) * 4;
    if (b->original_id.ensure(idb) != cudaSuccess || b->sorted_id.ensure(idb) != cudaSuccess)
    {
        delete b;
        return ctx->fail(SPHB200_ERR_CUDA, "body_create: out of device memory");
    }
    if (n > 0)
    {
        k_iota<<<gridFor(n, 256), 256, 0, ctx->stream>>>(static_cast<u32 *>(b->original_id.p),
                                                       static_cast<u32 *>(b->sorted_id.p), n);
        ctx->launches++;
    }
    ctx->bodies.push_back(b);
    // base_particles.cpp:33-37: Position and VolumetricMeasure always exist and evolve
    sphb200_var_register(b, "Position", ctx->ndim, 1, nullptr);
    sphb200_var_register(b, "VolumetricMeasure", 1, 1, nullptr);
    *out = b;
    return SPHB200_OK;
}

sphb200_status sphb200_body_set_cell_spacing(sphb200_body *b, double spacing)
{
    if (!b)
        return SPHB200_ERR_INVALID;
    sphb200_ctx *ctx = b->ctx;
    if (!(spacing > 0))
        return ctx->fail(SPHB200_ERR_INVALID, "body_set_cell_spacing: spacing must be positive");
    cudaSetDevice(ctx->device);
    SPHB_TRY(SPHB_DISPATCH(ctx, setMeshSpacing, b, spacing));
    b->cell_list_valid = false; // the next cell_list_update bins on the new mesh; relations on the old one are stale
    b->config_stamp++;
    return SPHB200_OK;
}

sphb200_status sphb200_body_set_fluid(sphb200_body *b, double rho0, double c0, double mu)
{
    if (!b)
        return SPHB200_ERR_INVALID;
    b->rho0 = rho0;
    b->c0 = c0;
    b->mu = mu;
    return SPHB200_OK;
}

sphb200_status sphb200_body_set_elastic(sphb200_body *b, double rho0, double E, double nu)
{
    if (!b)
        return SPHB200_ERR_INVALID;
    // materials/elastic_solid.cpp:54-80 (LinearElasticSolid ctor), :13-18 (setSoundSpeeds)
    b->rho0 = rho0;
    b->youngs = E;
    b->poisson = nu;
    b->G0 = 0.5 * E / (1.0 + nu);
    b->K0 = E / 3.0 / (1.0 - 2.0 * nu);
    b->lambda0 = nu * E / (1.0 + nu) / (1.0 - 2.0 * nu);
    b->c0 = std::sqrt(b->K0 / rho0);
    b->cs0 = std::sqrt(b->G0 / rho0);
    return SPHB200_OK;
}

sphb200_status sphb200_body_set_periodic(sphb200_body *b, int axis, double lo, double hi)
{
    if (!b || axis < 0 || axis >= b->ctx->ndim || !(hi > lo))
        return SPHB200_ERR_INVALID;
    b->periodic[axis] = true;
    b->per_lo[axis] = lo;
    b->per_hi[axis] = hi;
    return SPHB_DISPATCH(b->ctx, setPeriodic, b, axis, lo, hi);
}

sphb200_status sphb200_body_info(sphb200_body *b, double *out12)
{
    if (!b || !out12)
        return SPHB200_ERR_INVALID;
    return SPHB_DISPATCH(b->ctx, bodyInfo, b, out12);
}

sphb200_status sphb200_var_register(sphb200_body *b, const char *name, int width, int evolving, void *host_ptr)
{
    if (!b || !name || width < 1 || width > 9)
        return SPHB200_ERR_INVALID;
    sphb200_ctx *ctx = b->ctx;
    cudaSetDevice(ctx->device);
    Var *v = b->find(name);
    if (v)
    {
        if (v->width != width)
            return ctx->fail(SPHB200_ERR_INVALID, std::string("var_register: '") + name + "' exists with another width");
        if (host_ptr)
            v->host = host_ptr;
        v->evolving = v->evolving || evolving != 0;
        return SPHB200_OK;
    }
    auto nv = std::make_unique<Var>();
    nv->name = name;
    nv->width = width;
    nv->evolving = evolving != 0;
    nv->host = host_ptr;
    size_t bytes = size_t(b->n > 0 ? b->n : 1) * size_t(width) * size_t(ctx->real_bytes);
    SPHB_CUDA(ctx, nv->dev.ensure(bytes));
    SPHB_CUDA(ctx, cudaMemsetAsync(nv->dev.p, 0, nv->dev.bytes, ctx->stream));
    b->vars.emplace(name, std::move(nv));
    return SPHB200_OK;
}

static void touched(sphb200_body *b, const char *name)
{
    if (!strcmp(name, "VolumetricMeasure"))
        b->xV_vol_dirty = true;
    if (!strcmp(name, "Velocity"))
        b->acoustic_max_valid = false;
    if (!strcmp(name, "Velocity") || !strcmp(name, "VolumetricMeasure") || !strcmp(name, "Pressure"))
        b->rec_b_valid = false;
    if (!strcmp(name, "Position"))
        b->cell_list_valid = b->cell_list_valid; // the frozen snapshot stays; a new cell_list_update refreshes it
}

sphb200_status sphb200_var_upload(sphb200_body *b, const char *name, const void *src)
{
    if (!b || !name || !src)
        return SPHB200_ERR_INVALID;
    sphb200_ctx *ctx = b->ctx;
    Var *v = b->find(name);
    if (!v)
        return ctx->fail(SPHB200_ERR_NOT_FOUND, std::string("variable '") + name + "' is not registered");
    cudaSetDevice(ctx->device);
    size_t bytes = size_t(b->n) * size_t(v->width) * size_t(ctx->real_bytes);
    SPHB_CUDA(ctx, cudaMemcpyAsync(v->dev.p, src, bytes, cudaMemcpyHostToDevice, ctx->stream));
    SPHB_CUDA(ctx, cudaStreamSynchronize(ctx->stream)); // src may be pageable and reused by the caller
    touched(b, name);
    return SPHB200_OK;
}
sphb200_status sphb200_var_stage(sphb200_body *b, const char *name, const void *src)
{
    if (!b || !name || !src)
        return SPHB200_ERR_INVALID;
    sphb200_ctx *ctx = b->ctx;
    Var *v = b->find(name);
    if (!v)
        return ctx->fail(SPHB200_ERR_NOT_FOUND, std::string("variable '") + name + "' is not registered");
    cudaSetDevice(ctx->device);
    if (!ctx->copy_stream)
    {
        SPHB_CUDA(ctx, cudaStreamCreateWithFlags(&ctx->copy_stream, cudaStreamNonBlocking));
        SPHB_CUDA(ctx, cudaEventCreateWithFlags(&ctx->copy_done, cudaEventDisableTiming));
        SPHB_CUDA(ctx, cudaEventCreateWithFlags(&ctx->mirrors_free, cudaEventDisableTiming));
    }
    const size_t bytes = size_t(b->n) * size_t(v->width) * size_t(ctx->real_bytes);
    if (v->stage.bytes < bytes)
    { // first use (or the body grew): the allocation synchronises the device once
        v->stage.release();
        SPHB_CUDA(ctx, v->stage.ensure(bytes));
    }
    // the staging mirror was a live mirror until the last commit: kernels queued before it may still read it
    SPHB_CUDA(ctx, cudaStreamWaitEvent(ctx->copy_stream, ctx->mirrors_free, 0));
    SPHB_CUDA(ctx, cudaMemcpyAsync(v->stage.p, src, bytes, cudaMemcpyHostToDevice, ctx->copy_stream));
    v->staged_n = b->n;
    v->staged_stamp = b->config_stamp;
    return SPHB200_OK;
}
// blocks the host until every var_stage copy issued so far has left its host buffer
sphb200_status sphb200_ctx_stage_wait(sphb200_ctx *ctx)
{
    if (!ctx)
        return SPHB200_ERR_INVALID;
    cudaSetDevice(ctx->device);
    if (ctx->copy_stream)
        SPHB_CUDA(ctx, cudaStreamSynchronize(ctx->copy_stream));
    return SPHB200_OK;
}
__global__ void k_iota_ids(u32 *__restrict__ a, u32 *__restrict__ b, int64_t n)
{
    int64_t i = int64_t(blockIdx.x) * blockDim.x + threadIdx.x;
    if (i < n)
        a[i] = b[i] = u32(i);
}
sphb200_status sphb200_body_commit_staged(sphb200_body *b)
{
    if (!b)
        return SPHB200_ERR_INVALID;
    sphb200_ctx *ctx = b->ctx;
    cudaSetDevice(ctx->device);
    bool any = false, reordered = false;
    for (auto &kv : b->vars)
        if (kv.second->staged_n >= 0)
        {
            if (kv.second->staged_n != b->n)
                return ctx->fail(SPHB200_ERR_INVALID, "commit_staged: body '" + b->name + "' was resized after var_stage of '" + kv.first + "'");
            any = true;
            reordered = reordered || kv.second->staged_stamp != b->config_stamp;
        }
    if (!any)
        return SPHB200_OK;
    if (reordered)
    {
        // a cell_list_update ran between var_stage and this commit: the rows of the resident variables are in another
        // particle order than the staged rows.  Swapping in a subset would silently mix the two, so the staged set must
        // cover every evolving variable; the staged (host) row order then DEFINES the body's order, as a reload does
        // (io_base.cpp:67-97 readFromFile + Dambreak.cpp:132-138): ids restart from the identity (sphb200_ids_from_host
        // restores saved ones) and the cell list / configurations must be rebuilt before the next interaction.
        for (auto &kv : b->vars)
            if (kv.second->evolving && kv.second->staged_n < 0)
                return ctx->fail(SPHB200_ERR_INVALID, "commit_staged: body '" + b->name + "' was re-sorted after var_stage; every evolving "
                                                      "variable must be staged then, '" + kv.first + "' was not");
        if (b->n > 0)
        {
            k_iota_ids<<<gridFor(b->n, 256), 256, 0, ctx->stream>>>(static_cast<u32 *>(b->original_id.p), static_cast<u32 *>(b->sorted_id.p), b->n);
            SPHB_LAUNCH_CHECK(ctx);
        }
        b->cell_list_valid = false;
        for (sphb200_relation *r : ctx->relations)
            if (r->body == b || std::find(r->targets.begin(), r->targets.end(), b) != r->targets.end())
                r->valid = false;
    }
    SPHB_CUDA(ctx, cudaEventRecord(ctx->copy_done, ctx->copy_stream));
    SPHB_CUDA(ctx, cudaStreamWaitEvent(ctx->stream, ctx->copy_done, 0));
    for (auto &kv : b->vars)
    {
        Var &v = *kv.second;
        if (v.staged_n < 0)
            continue;
        std::swap(v.dev, v.stage);
        v.staged_n = -1;
        touched(b, kv.first.c_str());
    }
    // what is now the staging side may be overwritten once everything queued so far has run
    SPHB_CUDA(ctx, cudaEventRecord(ctx->mirrors_free, ctx->stream));
    return SPHB200_OK;
}
sphb200_status sphb200_var_download(sphb200_body *b, const char *name, void *dst)
{
    if (!b || !name || !dst)
        return SPHB200_ERR_INVALID;
    sphb200_ctx *ctx = b->ctx;
    Var *v = b->find(name);
    if (!v)
        return ctx->fail(SPHB200_ERR_NOT_FOUND, std::string("variable '") + name + "' is not registered");
    cudaSetDevice(ctx->device);
    size_t bytes = size_t(b->n) * size_t(v->width) * size_t(ctx->real_bytes);
    SPHB_CUDA(ctx, cudaMemcpyAsync(dst, v->dev.p, bytes, cudaMemcpyDeviceToHost, ctx->stream));
    SPHB_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return SPHB200_OK;
}
// Asynchronous, double-buffered device -> host read of a variable (the output / restart path:
// BaseVariable::prepareForOutput(ParallelDevicePolicy), sphinxsys_variable.h:350-356 -> synchronizeWithDevice):
// a device-side snapshot is taken in stream order (so later kernels and re-sorts do not disturb it), then copied to
// dst on a separate stream while the compute stream goes on.  dst (pinned for a truly asynchronous copy) holds the
// rows in the body's CURRENT order once sphb200_ctx_fetch_wait returns.
sphb200_status sphb200_var_fetch_async(sphb200_body *b, const char *name, void *dst)
{
    if (!b || !name || !dst)
        return SPHB200_ERR_INVALID;
    sphb200_ctx *ctx = b->ctx;
    Var *v = b->find(name);
    if (!v)
        return ctx->fail(SPHB200_ERR_NOT_FOUND, std::string("variable '") + name + "' is not registered");
    cudaSetDevice(ctx->device);
    if (!ctx->fetch_stream)
    {
        SPHB_CUDA(ctx, cudaStreamCreateWithFlags(&ctx->fetch_stream, cudaStreamNonBlocking));
        SPHB_CUDA(ctx, cudaEventCreateWithFlags(&ctx->fetch_ready, cudaEventDisableTiming));
        SPHB_CUDA(ctx, cudaEventCreateWithFlags(&ctx->fetch_done, cudaEventDisableTiming));
        SPHB_CUDA(ctx, cudaEventRecord(ctx->fetch_done, ctx->fetch_stream));
    }
    const size_t bytes = size_t(b->n) * size_t(v->width) * size_t(ctx->real_bytes);
    if (bytes == 0)
        return SPHB200_OK;
    if (v->fetch.bytes < bytes)
    {
        SPHB_CUDA(ctx, cudaStreamSynchronize(ctx->fetch_stream));
        v->fetch.release();
        SPHB_CUDA(ctx, v->fetch.ensure(bytes));
    }
    // the snapshot may be overwritten only after the previous copy out of it has finished
    SPHB_CUDA(ctx, cudaStreamWaitEvent(ctx->stream, ctx->fetch_done, 0));
    SPHB_CUDA(ctx, cudaMemcpyAsync(v->fetch.p, v->dev.p, bytes, cudaMemcpyDeviceToDevice, ctx->stream));
    SPHB_CUDA(ctx, cudaEventRecord(ctx->fetch_ready, ctx->stream));
    SPHB_CUDA(ctx, cudaStreamWaitEvent(ctx->fetch_stream, ctx->fetch_ready, 0));
    SPHB_CUDA(ctx, cudaMemcpyAsync(dst, v->fetch.p, bytes, cudaMemcpyDeviceToHost, ctx->fetch_stream));
    SPHB_CUDA(ctx, cudaEventRecord(ctx->fetch_done, ctx->fetch_stream));
    return SPHB200_OK;
}
sphb200_status sphb200_ctx_fetch_wait(sphb200_ctx *ctx)
{
    if (!ctx)
        return SPHB200_ERR_INVALID;
    cudaSetDevice(ctx->device);
    if (ctx->fetch_stream)
        SPHB_CUDA(ctx, cudaStreamSynchronize(ctx->fetch_stream));
    return SPHB200_OK;
}
sphb200_status sphb200_var_to_device(sphb200_body *b, const char *name)
{
    if (!b || !name)
        return SPHB200_ERR_INVALID;
    Var *v = b->find(name);
    if (!v)
        return b->ctx->fail(SPHB200_ERR_NOT_FOUND, std::string("variable '") + name + "' is not registered");
    if (!v->host)
        return b->ctx->fail(SPHB200_ERR_INVALID, std::string("variable '") + name + "' has no host array bound");
    return sphb200_var_upload(b, name, v->host);
}
sphb200_status sphb200_var_to_host(sphb200_body *b, const char *name)
{
    if (!b || !name)
        return SPHB200_ERR_INVALID;
    Var *v = b->find(name);
    if (!v)
        return b->ctx->fail(SPHB200_ERR_NOT_FOUND, std::string("variable '") + name + "' is not registered");
    if (!v->host)
        return b->ctx->fail(SPHB200_ERR_INVALID, std::string("variable '") + name + "' has no host array bound");
    return sphb200_var_download(b, name, v->host);
}
sphb200_status sphb200_var_device_ptr(sphb200_body *b, const char *name, void **out)
{
    if (!b || !name || !out)
        return SPHB200_ERR_INVALID;
    Var *v = b->find(name);
    if (!v)
        return b->ctx->fail(SPHB200_ERR_NOT_FOUND, std::string("variable '") + name + "' is not registered");
    *out = v->dev.p;
    // the caller may write through this pointer: be conservative about cached derived state
    touched(b, name);
    return SPHB200_OK;
}

sphb200_status sphb200_var_width(sphb200_body *b, const char *name, int *width)
{
    if (!b || !name || !width)
        return SPHB200_ERR_INVALID;
    Var *v = b->find(name);
    if (!v)
        return b->ctx->fail(SPHB200_ERR_NOT_FOUND, std::string("variable '") + name + "' is not registered");
    *width = v->width;
    return SPHB200_OK;
}

__global__ void k_sorted_from_original(const u32 *__restrict__ oid, u32 *__restrict__ sid, int64_t n)
{
    int64_t i = int64_t(blockIdx.x) * blockDim.x + threadIdx.x;
    if (i < n)
        sid[oid[i]] = u32(i); // particle_sorting.cpp:55-58 UpdateSortedID
}
// restores OriginalID after a reload (the reference's restart file carries it as an evolving variable,
// base_particles.cpp:33-37); SortedID is rebuilt.  original_id must be a permutation of 0..n-1.
sphb200_status sphb200_ids_from_host(sphb200_body *b, const int64_t *original_id)
{
    if (!b || !original_id)
        return SPHB200_ERR_INVALID;
    sphb200_ctx *ctx = b->ctx;
    cudaSetDevice(ctx->device);
    std::vector<u32> tmp(static_cast<size_t>(b->n));
    std::vector<bool> seen(static_cast<size_t>(b->n), false);
    for (int64_t i = 0; i < b->n; ++i)
    {
        const int64_t o = original_id[i];
        if (o < 0 || o >= b->n || seen[size_t(o)])
            return ctx->fail(SPHB200_ERR_INVALID, "ids_from_host: original_id is not a permutation of 0..n-1");
        seen[size_t(o)] = true;
        tmp[size_t(i)] = u32(o);
    }
    if (b->n > 0)
    {
        SPHB_CUDA(ctx, cudaMemcpyAsync(b->original_id.p, tmp.data(), size_t(b->n) * 4, cudaMemcpyHostToDevice, ctx->stream));
        k_sorted_from_original<<<gridFor(b->n, 256), 256, 0, ctx->stream>>>(static_cast<const u32 *>(b->original_id.p),
                                                                          static_cast<u32 *>(b->sorted_id.p), b->n);
        SPHB_LAUNCH_CHECK(ctx);
        SPHB_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    }
    return SPHB200_OK;
}

sphb200_status sphb200_ids_to_host(sphb200_body *b, int64_t *original_id, int64_t *sorted_id)
{
    if (!b)
        return SPHB200_ERR_INVALID;
    sphb200_ctx *ctx = b->ctx;
    cudaSetDevice(ctx->device);
    std::vector<u32> tmp;
    tmp.resize(size_t(b->n));
    for (int pass = 0; pass < 2; ++pass)
    {
        int64_t *dst = pass == 0 ? original_id : sorted_id;
        if (!dst)
            continue;
        DevBuf &src = pass == 0 ? b->original_id : b->sorted_id;
        SPHB_CUDA(ctx, cudaMemcpyAsync(tmp.data(), src.p, size_t(b->n) * 4, cudaMemcpyDeviceToHost, ctx->stream));
        SPHB_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
        for (int64_t i = 0; i < b->n; ++i)
            dst[i] = int64_t(tmp[size_t(i)]);
    }
    return SPHB200_OK;
}
} // extern "C"
