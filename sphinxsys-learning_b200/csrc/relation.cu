// InnerRelation / ContactRelation: neighbour-index lists built from the cell-sorted bodies.
//
// Reference behaviour reproduced:
//   meshes/cell_linked_list.hpp:20-46              searchNeighborsByMesh (3^d cells, clamped)
//   particle_neighborhood/neighborhood.cpp:85-98   NeighborBuilderInner  (d.d < rc^2 && i != j)
//   particle_neighborhood/neighborhood.cpp:147-160 NeighborBuilderContact (larger-h kernel, no i != j)
//   general_dynamics/domian_bouding/domain_bounding.cpp:18-65  periodic ghost entries x_j ± L
//   shared_ck/.../update_body_relation.hpp:62-164  count → scan → fill (index-only list)
//
// The accepted neighbour SET per particle equals the reference's: the cut-off test is evaluated on
// the un-normalised displacement with separately rounded products and left-to-right sums
// (sqnormExact), on the positions frozen by cell_list_update.  Only indices are stored (u32, SELL-32
// layout so the 32 lanes of a warp read 128 contiguous bytes per neighbour slot); W_ij, dW_ij, e_ij,
// r_ij are recomputed from the frozen positions by the interaction kernels, which trades the
// reference's 56 B/pair of list traffic for ~15 flops.
#include "internal.cuh"

using namespace sphb;

namespace sphb
{
template <int D, class R>
struct SearchP
{
    MeshP<D, R> mesh;      // target body's mesh
    PeriodicP<D, R> per;   // target body's periodic box
    R rc2;
    R rc;
    int depth;
    int inner;
    // other neighbour builders (generic search only)
    int builder;               // 0 inner / contact, 1 self contact, 2 surface contact (rc of the averaged kernel)
    const unsigned char *part; // source particles of a body part, by original id; null: every particle
    const u32 *oid_s;          // OriginalID of the source body
    const R *pos0_s, *pos0_t;  // InitialPosition (self contact)
};

// Calls f(j, xj_shifted) for every candidate entry of the 3^d (depth) cell block around xi,
// including periodic images, in a deterministic order.
template <int D, class R, class F>
__device__ __forceinline__ void forEachCandidate(const R *xi, const SearchP<D, R> &sp,
                                                 const u32 *__restrict__ cell_offset,
                                                 const typename Real4<R>::type *__restrict__ xV_t, F &&f)
{
    int c[D];
    sp.mesh.cellIndex(xi, c);
    if (sp.per.mask == 0)
    {
        // non-periodic fast path: the cells of a run along the fastest axis are contiguous
        int lo[D], hi[D];
#pragma unroll
        for (int a = 0; a < D; ++a)
        {
            lo[a] = max(c[a] - sp.depth, 0);
            hi[a] = min(c[a] + sp.depth, sp.mesh.cells[a] - 1);
        }
        int cc[D];
        for (cc[0] = lo[0]; cc[0] <= hi[0]; ++cc[0])
        {
            int y0 = D == 3 ? lo[1] : 0, y1 = D == 3 ? hi[1] : 0;
            for (int cy = y0; cy <= y1; ++cy)
            {
                if (D == 3)
                    cc[1] = cy;
                cc[D - 1] = lo[D - 1];
                u32 first = sp.mesh.linear(cc);
                u32 b = __ldg(cell_offset + first);
                u32 e = __ldg(cell_offset + first + u32(hi[D - 1] - lo[D - 1]) + 1);
                for (u32 j = b; j < e; ++j)
                {
                    typename Real4<R>::type v = ldg4(xV_t + j);
                    R xj[D];
                    xj[0] = v.x;
                    xj[1] = v.y;
                    if (D == 3)
                        xj[D - 1] = v.z;
                    f(j, xj);
                }
            }
        }
        return;
    }
    // periodic path (small bodies only: Taylor-Green class cases).  Per axis the candidate cells are
    // the 3 around c plus, when xi is within rc of a periodic face, the cells that hold the
    // particles whose ghost copy the reference would have placed next to xi.
    int cell_lo[D][2], cell_hi[D][2], shift[D][2], nr[D];
#pragma unroll
    for (int a = 0; a < D; ++a)
    {
        cell_lo[a][0] = max(c[a] - sp.depth, 0);
        cell_hi[a][0] = min(c[a] + sp.depth, sp.mesh.cells[a] - 1);
        shift[a][0] = 0;
        nr[a] = 1;
        if (sp.per.mask & (1 << a))
        {
            R p[D];
#pragma unroll
            for (int q = 0; q < D; ++q)
                p[q] = xi[q];
            int t[D];
            if (xi[a] + sp.rc > sp.per.hi[a])
            { // images x_j + L of particles with lo < x_j < lo + rc
                p[a] = sp.per.lo[a];
                sp.mesh.cellIndex(p, t);
                int l = t[a];
                p[a] = sp.per.lo[a] + sp.rc;
                sp.mesh.cellIndex(p, t);
                cell_lo[a][1] = l;
                cell_hi[a][1] = t[a];
                shift[a][1] = 1;
                nr[a] = 2;
            }
            else if (xi[a] - sp.rc < sp.per.lo[a])
            { // images x_j - L of particles with hi - rc < x_j < hi
                p[a] = sp.per.hi[a] - sp.rc;
                sp.mesh.cellIndex(p, t);
                int l = t[a];
                p[a] = sp.per.hi[a];
                sp.mesh.cellIndex(p, t);
                cell_lo[a][1] = l;
                cell_hi[a][1] = t[a];
                shift[a][1] = -1;
                nr[a] = 2;
            }
        }
    }
    const int nz = D == 3 ? nr[D - 1] : 1;
    for (int r0 = 0; r0 < nr[0]; ++r0)
        for (int r1 = 0; r1 < nr[1]; ++r1)
            for (int r2 = 0; r2 < nz; ++r2)
            {
                int rr[3] = {r0, r1, r2};
                int cc[D];
                for (cc[0] = cell_lo[0][rr[0]]; cc[0] <= cell_hi[0][rr[0]]; ++cc[0])
                    for (cc[1] = cell_lo[1][rr[1]]; cc[1] <= cell_hi[1][rr[1]]; ++cc[1])
                    {
                        int z0 = D == 3 ? cell_lo[D - 1][rr[D - 1]] : 0, z1 = D == 3 ? cell_hi[D - 1][rr[D - 1]] : 0;
                        for (int cz = z0; cz <= z1; ++cz)
                        {
                            if (D == 3)
                                cc[D - 1] = cz;
                            u32 cell = sp.mesh.linear(cc);
                            u32 b = __ldg(cell_offset + cell), e = __ldg(cell_offset + cell + 1);
                            for (u32 j = b; j < e; ++j)
                            {
                                typename Real4<R>::type v = ldg4(xV_t + j);
                                R xj[D];
                                xj[0] = v.x;
                                xj[1] = v.y;
                                if (D == 3)
                                    xj[D - 1] = v.z;
                                bool ok = true;
#pragma unroll
                                for (int a = 0; a < D; ++a)
                                {
                                    int s = shift[a][rr[a]];
                                    if (s > 0)
                                    { // ghost exists iff lo < x_j < lo + rc (domain_bounding.cpp:36-44)
                                        ok = ok && (xj[a] > sp.per.lo[a]) && (xj[a] < (sp.per.lo[a] + sp.rc));
                                        xj[a] = xj[a] + sp.per.L[a];
                                    }
                                    else if (s < 0)
                                    { // ghost exists iff hi - rc < x_j < hi (domain_bounding.cpp:54-62)
                                        ok = ok && (xj[a] < sp.per.hi[a]) && (xj[a] > (sp.per.hi[a] - sp.rc));
                                        xj[a] = xj[a] - sp.per.L[a];
                                    }
                                }
                                if (ok)
                                    f(j, xj);
                            }
                        }
                    }
            }
}

template <int D, class R, bool FILL>
__global__ void __launch_bounds__(128) k_search(int64_t n_src, const typename Real4<R>::type *__restrict__ xV_s,
                                                const typename Real4<R>::type *__restrict__ xV_t,
                                                const u32 *__restrict__ cell_offset, SearchP<D, R> sp,
                                                u32 *__restrict__ count, u32 *__restrict__ slice_w,
                                                const u32 *__restrict__ slice_off, u32 *__restrict__ list)
{
    int64_t i = blockIdx.x * int64_t(blockDim.x) + threadIdx.x;
    const int lane = threadIdx.x & 31;
    u32 cnt = 0;
    if (i < n_src && (!sp.part || sp.part[sp.oid_s[i]]))
    {
        typename Real4<R>::type v = ldg4(xV_s + i);
        R xi[D];
        xi[0] = v.x;
        xi[1] = v.y;
        if (D == 3)
            xi[D - 1] = v.z;
        u32 *slot0 = nullptr;
        if (FILL)
            slot0 = listSlot(list, slice_off, i, 0);
        forEachCandidate<D, R>(xi, sp, cell_offset, xV_t,
                               [&](u32 j, const R *xj)
                               {
                                   R d[D];
#pragma unroll
                                   for (int a = 0; a < D; ++a)
                                       d[a] = xi[a] - xj[a];
                                   R m = sqnormExact<D, R>(d);
                                   bool accept = m < sp.rc2 && (!sp.inner || j != u32(i));
                                   if (sp.builder == 1)
                                   { // NeighborBuilderSelfContact (neighborhood.cpp:123-142): distance < rc and initial distance > rc
                                       R d0[D];
#pragma unroll
                                       for (int a = 0; a < D; ++a)
                                           d0[a] = sp.pos0_s[i * D + a] - sp.pos0_t[size_t(j) * D + a];
                                       accept = sqrt_r(m) < sp.rc && sqrt_r(sqnormExact<D, R>(d0)) > sp.rc;
                                   }
                                   if (accept)
                                   {
                                       if (FILL)
                                           slot0[size_t(cnt / kS) * 32u + (cnt % kS)] = j;
                                       ++cnt;
                                   }
                               });
    }
    if (!FILL)
    {
        if (i < n_src)
            count[i] = cnt;
        u32 w = (cnt + kS - 1) / kS; // steps the kS lanes of this particle need
#pragma unroll
        for (int o = kP / 2; o > 0; o >>= 1)
            w = max(w, __shfl_xor_sync(0xffffffffu, w, o));
        if ((lane % kP) == 0 && i < n_src)
            slice_w[i / kP] = w;
    }
}

// ---------------------------------------------------------------------------------------------
// Fast path (non-periodic target, search depth 1): ONE distance pass.
//   k_search_mask : per particle, per candidate run (the <= 3 cells along the fastest axis of each of the
//                   3^(D-1) neighbouring cell columns are contiguous particles) a hit bit per candidate,
//                   kRunWords 32-bit words per run; also count[i] and the slice widths.
//   k_expand_mask : after the scan of the slice widths, turns the hit bits into the index list - only when a
//                   kernel asks for the list (ensureList); the group-tiled kernels build their union lists from
//                   the bits themselves (tile.cuh k_union_mask).
// The second pass of the generic path (another ~474 distance tests per particle in 3-D) is replaced by
// ~80 bit extractions.  A run longer than 32*kRunWords candidates raises the overflow flag and the host
// falls back to the generic two-pass search.
// ---------------------------------------------------------------------------------------------
// fp32 pre-filter of the fp64 search.  xf = float(x - mesh.lo); a candidate whose fp32 squared distance is
// below pre.lo is a hit, above pre.hi a miss, and only what falls between (or is NaN: a particle outside
// the mesh) takes the exact fp64 test, so the accepted set is bit-for-bit that of the exact test.  With
// u = 2^-24, M = the largest |x - lo| of a particle inside the mesh (+ one cell) and d the true offset:
// every fp32 coordinate is off by <= u M, a component of the fp32 offset by delta <= 2 u M + u |d_k|, and
// the fp32 sum of squares by <= 2 sqrt(D) |d| delta + D delta^2 + 4 u |d|^2.  Near |d| = rc this is
// tau0 = 2 sqrt(D) rc delta + D delta^2 + 4 u rc^2; the band is rc^2 -+ 4 tau0 (candidates lie within
// 3.5 rc, where the error grows at most linearly in |d| while the margin |d|^2 - rc^2 grows faster).
// The fp32 tests issue at twice the rate of the fp64 ones and read 16 instead of 32 bytes per candidate.
struct PrefilterP
{
    float lo, hi;    // hit below lo, miss above hi
    float limit[3];  // a coordinate outside [-spacing, limit] is stored as NaN
};
template <int D, class R>
__global__ void __launch_bounds__(256)
    k_make_xf(int64_t n, const typename Real4<R>::type *__restrict__ xV, MeshP<D, R> mesh, PrefilterP pre, float4 *__restrict__ xf)
{
    int64_t i = blockIdx.x * int64_t(blockDim.x) + threadIdx.x;
    if (i >= n)
        return;
    typename Real4<R>::type v = ldg4(xV + i);
    R x[3] = {v.x, v.y, D == 3 ? v.z : R(0)};
    float f[3] = {0.f, 0.f, 0.f};
    bool inside = true;
#pragma unroll
    for (int d = 0; d < D; ++d)
    {
        const R rel = x[d] - mesh.lo[d];
        f[d] = float(rel);
        inside = inside && rel >= -mesh.spacing && f[d] <= pre.limit[d];
    }
    const float bad = __int_as_float(0x7fc00000);
    xf[i] = inside ? make_float4(f[0], f[1], f[2], 0.f) : make_float4(bad, bad, bad, 0.f);
}

template <int D, class R, bool INNER, bool PRE>
__global__ void __launch_bounds__(128) k_search_mask(int64_t n_src, const typename Real4<R>::type *__restrict__ xV_s,
                                                     const typename Real4<R>::type *__restrict__ xV_t,
                                                     const u32 *__restrict__ cell_offset, MeshP<D, R> mesh, R rc2,
                                                     u32 *__restrict__ count, u32 *__restrict__ slice_w,
                                                     u32 *__restrict__ mask, int *__restrict__ overflow,
                                                     const float4 *__restrict__ xf_t, PrefilterP pre)
{
    constexpr int W = MaskLayout<D>::words;
    int64_t i = blockIdx.x * int64_t(blockDim.x) + threadIdx.x;
    const int lane = threadIdx.x & 31;
    u32 cnt = 0;
    if (i < n_src)
    {
        typename Real4<R>::type v = ldg4(xV_s + i);
        R xi[D];
        xi[0] = v.x;
        xi[1] = v.y;
        if (D == 3)
            xi[D - 1] = v.z;
        int c[D];
        mesh.cellIndex(xi, c);
        float xif[D];
        if (PRE)
        {
#pragma unroll
            for (int d = 0; d < D; ++d)
            {
                const R rel = xi[d] - mesh.lo[d];
                xif[d] = float(rel);
                if (!(rel >= -mesh.spacing && xif[d] <= pre.limit[d]))
                    xif[d] = __int_as_float(0x7fc00000); // outside the mesh: every candidate takes the exact test
            }
        }
        u32 *mrow = mask + (u64(i >> 5) * W) * 32u + lane;
        for (int r = 0; r < MaskLayout<D>::runs; ++r)
        {
            u32 b, e;
            runRange<D, R>(c, r, mesh, cell_offset, b, e);
            if (e - b > 32u * kRunWords)
            {
                *overflow = 1;
                e = b + 32u * kRunWords;
            }
#pragma unroll
            for (int w = 0; w < kRunWords; ++w)
            {
                u32 word = 0;
                const u32 j0 = b + 32u * w;
                if (j0 < e)
                {
                    const u32 m = min(e - j0, 32u);
                    u32 todo = 0; // candidates that need the exact test
                    if (PRE)
                    {
                        // the kernel is issue-bound on the integer pipe: the two decisions are taken as sign bits
                        // (mm - lo < 0: hit, hi - mm < 0: miss; a NaN is 0x7fffffff in fp32, sign clear: undecided)
                        // and shifted into the words by one funnel shift each; candidate t lands on bit m-1-t
                        const float4 *cand = xf_t + j0;
                        u32 hits = 0u, misses = 0u;
#pragma unroll 8
                        for (u32 t = 0; t < m; ++t)
                        {
                            const float4 q = __ldg(cand + t);
                            const float dx = xif[0] - q.x, dy = xif[1] - q.y;
                            float mm = dx * dx + dy * dy;
                            if (D == 3)
                            {
                                const float dz = xif[D - 1] - q.z;
                                mm += dz * dz;
                            }
                            const u32 a = __float_as_uint(mm - pre.lo), b = __float_as_uint(pre.hi - mm);
                            hits = __funnelshift_l(a, hits, 1);
                            misses = __funnelshift_l(b, misses, 1);
                        }
                        const u32 live = m == 32u ? 0xffffffffu : (1u << m) - 1u;
                        word = __brev(hits) >> (32u - m);
                        todo = ~(__brev(hits | misses) >> (32u - m)) & live;
                    }
                    else
                        todo = m == 32u ? 0xffffffffu : (1u << m) - 1u;
                    if (PRE)
                    {
                        while (todo)
                        {
                            const u32 t = u32(__ffs(int(todo)) - 1);
                            todo &= todo - 1u;
                            typename Real4<R>::type q = ldg4(xV_t + (j0 + t));
                            R d[D];
                            d[0] = xi[0] - q.x;
                            d[1] = xi[1] - q.y;
                            if (D == 3)
                                d[D - 1] = xi[D - 1] - q.z;
                            R mm = sqnormExact<D, R>(d);
                            word |= (mm < rc2 ? 1u : 0u) << t;
                        }
                    }
                    else
                    {
                        const typename Real4<R>::type *cand = xV_t + j0;
                        if (sizeof(R) == 4)
                        { // fp32 build: mm < rc2 as the sign of mm - rc2 (exact for finite values, clear for a NaN)
                            u32 hits = 0u;
#pragma unroll 8
                            for (u32 t = 0; t < m; ++t)
                            {
                                typename Real4<R>::type q = ldg4(cand + t);
                                R d[D];
                                d[0] = xi[0] - q.x;
                                d[1] = xi[1] - q.y;
                                if (D == 3)
                                    d[D - 1] = xi[D - 1] - q.z;
                                const float diff = float(sqnormExact<D, R>(d)) - float(rc2);
                                hits = __funnelshift_l(__float_as_uint(diff), hits, 1);
                            }
                            word = __brev(hits) >> (32u - m);
                        }
                        else
                        {
#pragma unroll 4
                            for (u32 t = 0; t < m; ++t)
                            {
                                typename Real4<R>::type q = ldg4(cand + t);
                                R d[D];
                                d[0] = xi[0] - q.x;
                                d[1] = xi[1] - q.y;
                                if (D == 3)
                                    d[D - 1] = xi[D - 1] - q.z;
                                R mm = sqnormExact<D, R>(d);
                                word |= (mm < rc2 ? 1u : 0u) << t;
                            }
                        }
                    }
                    if (INNER)
                    { // NeighborBuilderInner: i != j (neighborhood.cpp:85-98)
                        u32 self = u32(i) - j0;
                        if (self < 32u)
                            word &= ~(1u << self);
                    }
                }
                mrow[u64(r * kRunWords + w) * 32u] = word;
                cnt += __popc(word);
            }
        }
        count[i] = cnt;
    }
    u32 w = (cnt + kS - 1) / kS;
#pragma unroll
    for (int o = kP / 2; o > 0; o >>= 1)
        w = max(w, __shfl_xor_sync(0xffffffffu, w, o));
    if ((lane % kP) == 0 && i < n_src)
        slice_w[i / kP] = w;
}

// Rows of the sliced-ELL list are assembled in shared memory (each thread drops its particle's indices
// into its own columns) and written out as whole 128-byte rows.  kExpandRows rows per slice are staged;
// the rare particle with more than kExpandRows*kS neighbours writes its tail straight to global memory.
#ifndef SPHB_EXPAND_ROWS
#define SPHB_EXPAND_ROWS 44
#endif
constexpr int kExpandRows = SPHB_EXPAND_ROWS;
constexpr int kExpandBlock = 128;
constexpr size_t kExpandSmem = size_t(kExpandBlock / 32) * kS * kExpandRows * 32 * sizeof(u32);

template <int D, class R>
__global__ void __launch_bounds__(kExpandBlock) k_expand_mask(int64_t n_src, const typename Real4<R>::type *__restrict__ xV_s,
                                                              const u32 *__restrict__ cell_offset, MeshP<D, R> mesh,
                                                              const u32 *__restrict__ mask, const u32 *__restrict__ slice_off,
                                                              u32 *__restrict__ list)
{
    extern __shared__ u32 tile_all[];
    constexpr int W = MaskLayout<D>::words;
    int64_t i = blockIdx.x * int64_t(blockDim.x) + threadIdx.x;
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    // this warp's 32 particles = kS slices of kP particles; tile[s][row][32]
    u32 *tile = tile_all + size_t(wid) * kS * kExpandRows * 32;
    u32 *my = tile + size_t(lane / kP) * kExpandRows * 32 + (lane % kP) * kS;
    if (i < n_src)
    {
        typename Real4<R>::type v = ldg4(xV_s + i);
        R xi[D];
        xi[0] = v.x;
        xi[1] = v.y;
        if (D == 3)
            xi[D - 1] = v.z;
        int c[D];
        mesh.cellIndex(xi, c);
        const u32 *mrow = mask + (u64(i >> 5) * W) * 32u + lane;
        u32 *slot0 = listSlot(list, slice_off, i, 0);
        u32 cnt = 0;
        u32 rb[MaskLayout<D>::runs], words[W];
#pragma unroll
        for (int q = 0; q < W; ++q) // independent loads, all in flight at once
            words[q] = __ldg(mrow + u64(q) * 32u);
#pragma unroll
        for (int r = 0; r < MaskLayout<D>::runs; ++r)
        {
            u32 e_unused;
            runRange<D, R>(c, r, mesh, cell_offset, rb[r], e_unused);
        }
#pragma unroll
        for (int r = 0; r < MaskLayout<D>::runs; ++r)
        {
#pragma unroll
            for (int w = 0; w < kRunWords; ++w)
            {
                u32 word = words[r * kRunWords + w];
                const u32 j0 = rb[r] + 32u * w;
                while (word)
                {
                    u32 t = __ffs(word) - 1;
                    word &= word - 1;
                    const u32 row = cnt / kS, col = cnt % kS;
                    if (row < u32(kExpandRows))
                        my[row * 32 + col] = j0 + t;
                    else
                        slot0[size_t(row) * 32u + col] = j0 + t;
                    ++cnt;
                }
            }
        }
    }
    __syncwarp();
    // copy out: slice s of this warp covers particles [i_warp + s*kP, +kP)
    const int64_t i_warp = i - lane;
#pragma unroll
    for (int sl = 0; sl < kS; ++sl)
    {
        const int64_t first = i_warp + int64_t(sl) * kP;
        if (first >= n_src)
            break;
        const u32 so = slice_off[first / kP];
        const u32 rows = min(slice_off[first / kP + 1] - so, u32(kExpandRows));
        const u32 *src = tile + size_t(sl) * kExpandRows * 32 + lane;
        u32 *dst = list + u64(so) * 32u + lane;
        for (u32 row = 0; row < rows; ++row)
            dst[size_t(row) * 32u] = src[row * 32];
    }
}

__global__ void k_sum_u32(const u32 *__restrict__ v, int64_t n, unsigned long long *out)
{
    int64_t i = blockIdx.x * int64_t(blockDim.x) + threadIdx.x;
    unsigned long long s = 0;
    for (; i < n; i += int64_t(gridDim.x) * blockDim.x)
        s += v[i];
    s = warpSum<unsigned long long>(s);
    if ((threadIdx.x & 31) == 0 && s)
        atomicAdd(out, s);
}

// thresholds of the fp32 pre-filter (see PrefilterP); false when the band would be too wide to be worth it
template <int D, class R>
static bool searchPrefilter(const MeshP<D, R> &mesh, R rc_, PrefilterP &pre)
{
    const char *env = getenv("SPHB200_SEARCH_PREFILTER");
    if (env && env[0] == '0')
        return false;
    const double u = 1.0 / 16777216.0, rc = double(rc_), sp = double(mesh.spacing);
    double M = 0;
    for (int d = 0; d < 3; ++d)
        pre.limit[d] = 0.f;
    for (int d = 0; d < D; ++d)
    {
        const double ext = double(mesh.cells[d]) * sp + sp;
        M = ext > M ? ext : M;
        pre.limit[d] = float(ext);
    }
    M *= 1.0 + 4 * u; // limit[] is compared in fp32
    const double delta = 2 * u * M + u * 4 * rc;
    const double tau = 4.0 * (2.0 * sqrt(double(D)) * rc * delta + D * delta * delta + 4 * u * rc * rc) + 4 * u * rc * rc;
    const double rc2 = rc * rc;
    if (tau > 0.05 * rc2)
        return false;
    pre.lo = nextafterf(float(rc2 - tau), 0.f);
    pre.hi = nextafterf(float(rc2 + tau), INFINITY);
    return true;
}

static sphb200_status reserveList(sphb200_ctx *ctx, NeighborList &L)
{
    if (L.total_entries > L.cap_entries || !L.list)
    {
        cudaFree(L.list);
        L.list = nullptr;
        L.cap_entries = 0;
        size_t cap = size_t(L.total_entries) + size_t(L.total_entries) / 8 + 1024;
        SPHB_CUDA(ctx, cudaMalloc(&L.list, cap * 4));
        L.cap_entries = cap;
    }
    return SPHB200_OK;
}

template <int D, class R>
static sphb200_status expandList(sphb200_relation *r, size_t k)
{
    sphb200_ctx *ctx = r->ctx;
    NeighborList &L = r->lists[k];
    if (L.expanded)
        return SPHB200_OK;
    if (!L.has_mask)
        return ctx->fail(SPHB200_ERR_INVALID, "neighbour list requested before relation_update");
    const int64_t n = r->body->n;
    SPHB_TRY(reserveList(ctx, L));
    if (n > 0)
    {
        // the opt-in is a per-device attribute of the function: remember it per (D, R) instantiation AND device
        static bool attr_set[64] = {};
        if (ctx->device < 0 || ctx->device >= 64 || !attr_set[ctx->device])
        {
            SPHB_CUDA(ctx, cudaFuncSetAttribute(k_expand_mask<D, R>, cudaFuncAttributeMaxDynamicSharedMemorySize, int(kExpandSmem)));
            if (ctx->device >= 0 && ctx->device < 64)
                attr_set[ctx->device] = true;
        }
        sphb200_body *tgt = r->targets[k];
        k_expand_mask<D, R><<<gridFor(n, kExpandBlock), kExpandBlock, kExpandSmem, ctx->stream>>>(
            n, static_cast<const typename Real4<R>::type *>(r->body->xV.p), static_cast<const u32 *>(tgt->cell_offset.p),
            tgt->P<D, R>().mesh, L.mask, L.slice_off, L.list);
        SPHB_LAUNCH_CHECK(ctx);
    }
    L.expanded = true;
    return SPHB200_OK;
}

template <int D, class R>
static sphb200_status relationUpdate(sphb200_relation *r)
{
    sphb200_ctx *ctx = r->ctx;
    sphb200_body *src = r->body;
    cudaStream_t st = ctx->stream;
    if (!src->cell_list_valid)
        return ctx->fail(SPHB200_ERR_INVALID, "relation_update: source body '" + src->name + "' has no cell list yet");
    const int64_t n = src->n;
    const int64_t n_slices = (n + kP - 1) / kP;
    BodyP<D, R> &Ps = src->P<D, R>();
    for (size_t k = 0; k < r->targets.size(); ++k)
    {
        sphb200_body *tgt = r->targets[k];
        if (!tgt->cell_list_valid)
            return ctx->fail(SPHB200_ERR_INVALID, "relation_update: target body '" + tgt->name + "' has no cell list yet");
        BodyP<D, R> &Pt = tgt->P<D, R>();
        NeighborList &L = r->lists[k];
        if (L.cap_src < size_t(n) || !L.count)
        {
            cudaFree(L.count);
            cudaFree(L.slice_w);
            cudaFree(L.slice_off);
            L.count = L.slice_w = L.slice_off = nullptr;
            size_t cap = size_t(n) + size_t(n) / 8 + 64;
            SPHB_CUDA(ctx, cudaMalloc(&L.count, cap * 4));
            SPHB_CUDA(ctx, cudaMalloc(&L.slice_w, (cap / kP + 2) * 4));
            SPHB_CUDA(ctx, cudaMalloc(&L.slice_off, (cap / kP + 3) * 4));
            L.cap_src = cap;
        }
        SearchP<D, R> sp;
        sp.mesh = Pt.mesh;
        sp.per = Pt.per;
        sp.inner = r->inner ? 1 : 0;
        sp.builder = r->builder;
        sp.part = r->has_part ? static_cast<const unsigned char *>(r->part_flag.p) : nullptr;
        sp.oid_s = static_cast<const u32 *>(src->original_id.p);
        sp.pos0_s = sp.pos0_t = nullptr;
        if (r->builder == 1)
        {
            Var *p0 = src->find("InitialPosition");
            if (!p0 || p0->width != D)
                return ctx->fail(SPHB200_ERR_INVALID, "self-contact relation: the body has no InitialPosition variable");
            sp.pos0_s = sp.pos0_t = static_cast<const R *>(p0->dev.p);
            sp.inner = 0; // NeighborBuilderSelfContact has no i != j test: r0 = 0 fails r0 > rc
        }
        // neighborhood.cpp:75-80 chooseKernel: the kernel with the larger smoothing length
        const KernelP<R> &kern = (r->inner || Ps.h > Pt.h) ? Ps.kern : Pt.kern;
        sp.rc = kern.rc;
        sp.rc2 = kern.rc2;
        if (r->builder == 2)
        { // NeighborBuilderSurfaceContact: KernelWendlandC2(0.5 * (h_source + h_target)), cut-off 2 h
            const R h_avg = R(0.5) * (Ps.h + Pt.h);
            sp.rc = R(2.0) * h_avg;
            sp.rc2 = sp.rc * sp.rc;
        }
        // smallest depth whose cell block covers the cut-off sphere; the reference searches
        // 1+floor(rc/spacing) cells (base_body_relation.cpp:10-16), a superset with the same accepted set
        sp.depth = 1;
        while (R(sp.depth) * Pt.mesh.spacing < sp.rc)
            sp.depth++;
        L.total_entries = L.total_pairs = 0;
        if (n == 0)
        {
            r->target_stamp[k] = tgt->config_stamp;
            L.build_id++;
            L.has_mask = false;
            L.expanded = true;
            continue;
        }
        auto *xs = static_cast<const typename Real4<R>::type *>(src->xV.p);
        auto *xt = static_cast<const typename Real4<R>::type *>(tgt->xV.p);
        const u32 *off = static_cast<const u32 *>(tgt->cell_offset.p);
        unsigned long long *acc = static_cast<unsigned long long *>(ctx->reduce_dev.p) + 8;
        int *overflow = reinterpret_cast<int *>(acc + 1);
        bool fast = Pt.per.mask == 0 && sp.depth == 1 && r->builder == 0 && !r->has_part;
        u32 *mask = nullptr;
        for (int attempt = 0; attempt < 2; ++attempt)
        {
            SPHB_CUDA(ctx, cudaMemsetAsync(acc, 0, 16, st));
            if (fast)
            {
                const size_t words = size_t((n + 31) / 32) * 32u * size_t(MaskLayout<D>::words);
                if (L.cap_mask < words)
                {
                    cudaFree(L.mask);
                    L.mask = nullptr;
                    L.cap_mask = 0;
                    const size_t cap = words + words / 8;
                    SPHB_CUDA(ctx, cudaMalloc(&L.mask, cap * 4));
                    L.cap_mask = cap;
                }
                mask = L.mask;
                PrefilterP pre;
                const bool prefilter = sizeof(R) == 8 && searchPrefilter<D, R>(Pt.mesh, sp.rc, pre);
                const float4 *xf = nullptr;
                if (prefilter)
                {
                    if (tgt->xf_stamp != tgt->config_stamp || !tgt->xf.p)
                    {
                        SPHB_CUDA(ctx, tgt->xf.ensure(size_t(tgt->n) * sizeof(float4)));
                        k_make_xf<D, R><<<gridFor(tgt->n, 256), 256, 0, st>>>(tgt->n, xt, Pt.mesh, pre, static_cast<float4 *>(tgt->xf.p));
                        SPHB_LAUNCH_CHECK(ctx);
                        tgt->xf_stamp = tgt->config_stamp;
                    }
                    xf = static_cast<const float4 *>(tgt->xf.p);
                }
                const int grid = gridFor(n, 128);
                if (r->inner && prefilter)
                    k_search_mask<D, R, true, true><<<grid, 128, 0, st>>>(n, xs, xt, off, Pt.mesh, sp.rc2, L.count, L.slice_w, mask, overflow, xf, pre);
                else if (r->inner)
                    k_search_mask<D, R, true, false><<<grid, 128, 0, st>>>(n, xs, xt, off, Pt.mesh, sp.rc2, L.count, L.slice_w, mask, overflow, xf, pre);
                else if (prefilter)
                    k_search_mask<D, R, false, true><<<grid, 128, 0, st>>>(n, xs, xt, off, Pt.mesh, sp.rc2, L.count, L.slice_w, mask, overflow, xf, pre);
                else
                    k_search_mask<D, R, false, false><<<grid, 128, 0, st>>>(n, xs, xt, off, Pt.mesh, sp.rc2, L.count, L.slice_w, mask, overflow, xf, pre);
            }
            else
                k_search<D, R, false><<<gridFor(n, 128), 128, 0, st>>>(n, xs, xt, off, sp, L.count, L.slice_w, nullptr, nullptr);
            SPHB_LAUNCH_CHECK(ctx);
            SPHB_TRY(exclusiveScanU32(ctx, L.slice_w, L.slice_off, n_slices, true));
            k_sum_u32<<<148 * 4, 256, 0, st>>>(L.count, n, acc);
            SPHB_LAUNCH_CHECK(ctx);
            u32 total_w = 0;
            unsigned long long host_acc[2] = {0, 0};
            SPHB_CUDA(ctx, cudaMemcpyAsync(&total_w, L.slice_off + n_slices, 4, cudaMemcpyDeviceToHost, st));
            SPHB_CUDA(ctx, cudaMemcpyAsync(host_acc, acc, 16, cudaMemcpyDeviceToHost, st));
            SPHB_CUDA(ctx, cudaStreamSynchronize(st));
            if (fast && host_acc[1] != 0)
            { // a candidate run did not fit the mask words: redo with the generic search
                fast = false;
                continue;
            }
            L.total_entries = u64(total_w) * 32u;
            L.total_pairs = host_acc[0];
            break;
        }
        // the masked search leaves the hits as bits; the index list is written when a kernel asks for it
        // (ensureList) - the group-tiled kernels build their union lists from the bits directly (tile.cuh)
        L.has_mask = fast;
        L.expanded = false;
        if (!fast)
        {
            SPHB_TRY(reserveList(ctx, L));
            k_search<D, R, true><<<gridFor(n, 128), 128, 0, st>>>(n, xs, xt, off, sp, L.count, L.slice_w, L.slice_off, L.list);
            SPHB_LAUNCH_CHECK(ctx);
            L.expanded = true;
        }
        r->target_stamp[k] = tgt->config_stamp;
        L.build_id++;
        L.sorted = Pt.per.mask == 0; // both non-periodic searches emit ascending j per particle
    }
    r->source_stamp = src->config_stamp;
    r->valid = true;
    return SPHB200_OK;
}
sphb200_status ensureList(sphb200_relation *r, size_t k)
{
    if (r->lists[k].expanded)
        return SPHB200_OK;
    return SPHB_DISPATCH(r->ctx, expandList, r, k);
}
} // namespace sphb

extern "C"
{
sphb200_status sphb200_relation_inner_create(sphb200_body *b, sphb200_relation **out)
{
    if (!b || !out)
        return SPHB200_ERR_INVALID;
    auto *r = new sphb200_relation();
    r->ctx = b->ctx;
    r->body = b;
    r->inner = true;
    r->targets.push_back(b);
    r->lists.resize(1);
    r->target_stamp.assign(1, 0);
    b->ctx->relations.push_back(r);
    *out = r;
    return SPHB200_OK;
}

sphb200_status sphb200_relation_contact_create(sphb200_body *b, sphb200_body *const *targets, int n_targets,
                                               sphb200_relation **out)
{
    if (!b || !out || !targets || n_targets < 1 || n_targets > kMaxTargets)
        return SPHB200_ERR_INVALID;
    for (int k = 0; k < n_targets; ++k)
        if (!targets[k] || targets[k]->ctx != b->ctx)
            return b->ctx->fail(SPHB200_ERR_INVALID, "relation_contact_create: target from another context");
    auto *r = new sphb200_relation();
    r->ctx = b->ctx;
    r->body = b;
    r->inner = false;
    for (int k = 0; k < n_targets; ++k)
        r->targets.push_back(targets[k]);
    r->lists.resize(size_t(n_targets));
    r->target_stamp.assign(size_t(n_targets), 0);
    b->ctx->relations.push_back(r);
    *out = r;
    return SPHB200_OK;
}

static sphb200_status setPart(sphb200_relation *r, const int64_t *part_original_ids, int64_t n_part)
{
    sphb200_ctx *ctx = r->ctx;
    r->has_part = false;
    if (!part_original_ids || n_part <= 0)
        return SPHB200_OK;
    std::vector<unsigned char> flag(size_t(r->body->n), 0);
    for (int64_t k = 0; k < n_part; ++k)
    {
        if (part_original_ids[k] < 0 || part_original_ids[k] >= r->body->n)
            return ctx->fail(SPHB200_ERR_INVALID, "relation: body-part particle id out of range");
        flag[size_t(part_original_ids[k])] = 1;
    }
    SPHB_CUDA(ctx, r->part_flag.ensure(flag.size()));
    SPHB_CUDA(ctx, cudaMemcpyAsync(r->part_flag.p, flag.data(), flag.size(), cudaMemcpyHostToDevice, ctx->stream));
    SPHB_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    r->has_part = true;
    return SPHB200_OK;
}

sphb200_status sphb200_relation_self_contact_create(sphb200_body *b, const int64_t *part_original_ids, int64_t n_part, sphb200_relation **out)
{
    SPHB_TRY(sphb200_relation_inner_create(b, out));
    (*out)->builder = 1;
    cudaSetDevice(b->ctx->device);
    return setPart(*out, part_original_ids, n_part);
}

sphb200_status sphb200_relation_surface_contact_create(sphb200_body *b, sphb200_body *const *targets, int n_targets,
                                                       const int64_t *part_original_ids, int64_t n_part, sphb200_relation **out)
{
    SPHB_TRY(sphb200_relation_contact_create(b, targets, n_targets, out));
    (*out)->builder = 2;
    cudaSetDevice(b->ctx->device);
    return setPart(*out, part_original_ids, n_part);
}

sphb200_status sphb200_relation_update(sphb200_relation *r)
{
    if (!r)
        return SPHB200_ERR_INVALID;
    cudaSetDevice(r->ctx->device);
    return SPHB_DISPATCH(r->ctx, relationUpdate, r);
}

sphb200_status sphb200_relation_neighbors(sphb200_relation *r, int k, int32_t *counts, int64_t *j_out, int64_t *total)
{
    if (!r || k < 0 || k >= int(r->lists.size()))
        return SPHB200_ERR_INVALID;
    sphb200_ctx *ctx = r->ctx;
    if (!r->valid)
        return ctx->fail(SPHB200_ERR_INVALID, "relation_neighbors: relation has not been updated");
    cudaSetDevice(ctx->device);
    NeighborList &L = r->lists[size_t(k)];
    const int64_t n = r->body->n;
    if (total)
        *total = int64_t(L.total_pairs);
    if (!counts && !j_out)
        return SPHB200_OK;
    if (j_out)
        SPHB_TRY(sphb::ensureList(r, size_t(k)));
    std::vector<u32> c, so, lst;
    c.resize(size_t(n));
    so.resize(size_t((n + kP - 1) / kP + 1));
    if (n > 0)
    {
        SPHB_CUDA(ctx, cudaMemcpyAsync(c.data(), L.count, size_t(n) * 4, cudaMemcpyDeviceToHost, ctx->stream));
        SPHB_CUDA(ctx, cudaMemcpyAsync(so.data(), L.slice_off, so.size() * 4, cudaMemcpyDeviceToHost, ctx->stream));
        if (j_out && L.total_entries)
        {
            lst.resize(size_t(L.total_entries));
            SPHB_CUDA(ctx, cudaMemcpyAsync(lst.data(), L.list, lst.size() * 4, cudaMemcpyDeviceToHost, ctx->stream));
        }
        SPHB_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    }
    int64_t w = 0;
    for (int64_t i = 0; i < n; ++i)
    {
        if (counts)
            counts[i] = int32_t(c[size_t(i)]);
        if (j_out)
        {
            u64 base = u64(so[size_t(i / kP)]) * 32u + u64(i % kP) * u64(kS);
            for (u32 m = 0; m < c[size_t(i)]; ++m)
                j_out[w + m] = int64_t(lst[size_t(base + u64(m / kS) * 32u + (m % kS))]);
        }
        w += c[size_t(i)];
    }
    return SPHB200_OK;
}

sphb200_status sphb200_relation_stats(sphb200_relation *r, int k, double *out8)
{
    if (!r || !out8 || k < 0 || k >= int(r->lists.size()))
        return SPHB200_ERR_INVALID;
    if (!r->valid)
        return r->ctx->fail(SPHB200_ERR_INVALID, "relation_stats: relation has not been updated");
    const NeighborList &L = r->lists[size_t(k)];
    const bool uni = L.uni.build_id == L.build_id && L.uni.n_groups > 0;
    out8[0] = double(L.total_pairs);
    out8[1] = L.has_mask ? 1.0 : 0.0;
    out8[2] = L.expanded ? 1.0 : 0.0;
    out8[3] = uni ? double(L.uni.total_entries) : 0.0;
    out8[4] = uni ? double(L.uni.n_groups) : 0.0;
    out8[5] = uni ? double(L.uni.G) : 0.0;
    out8[6] = double(r->body->n);
    out8[7] = double(L.build_id);
    return SPHB200_OK;
}
} // extern "C"
