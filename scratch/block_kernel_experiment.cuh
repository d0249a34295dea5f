// EXPERIMENT, NOT BUILT INTO libsphb200 (round 2).  Measured on B200, 20.1 M fluid particles, fp64, wall pre-pass on:
//   k_second_half_tile (production)            8.60 ms
//   k_second_half_block (this file)           18.96 ms      results bit-identical to the tiled kernel
// The gathers themselves are cheaper from shared memory (profiles/r2a_gather_microbench.log: 0.68 vs 0.41
// lane-records/clk/SM), but a block's candidate pieces cost 6-8 staged records per owned particle, so ~100 KB of
// shared memory hold the candidates of only ~210 owned particles: 2 CTAs x 5 warps per SM at 66 % thread use against
// the tiled kernel's 16 warps per SM - too few warps to keep the FP64 pipe (34.5 instructions per evaluated pair)
// busy.  Kept for the record; DESIGN.md section 3 has the discussion.
// Cell-block variant of the 2nd-half pair kernel (included by fluid.cu after tile.cuh): the candidate records of a
// CTA are staged ONCE in shared memory and gathered from there.
//
// A CTA owns the particles of a block of cells: kBlkLy consecutive columns along y (one x, y0 <= y < y1) cut to the
// cell range [z0, z1) along the fastest axis.  Its owned particles are <= kBlkLy contiguous runs of the cell-sorted
// order, and every neighbour of every owned particle lies in one of the 3 * (kBlkLy + 2) candidate PIECES
// (column (x + dx, y), dx in -1..1, y0 - 1 <= y <= y1, cells z0 - 1 .. z1) - each piece a contiguous run of records
// (MeshP::linear, x slowest / z fastest), the same run structure the masked search uses (internal.cuh runRange).
// The pieces are copied into shared memory with coalesced 128-bit loads, {x, y, z, vx, vy, vz} at a 48-byte stride
// (fp64: 16-byte chunk c of record j falls into bank group (3 j + c) mod 8, so scattered LDS.128 spread over the
// banks), Vol beside it.  The threads then walk the same union lists as k_second_half_tile, translate each global
// index to its shared-memory slot with a running piece pointer (list entries ascend) and gather with LDS.128 instead
// of two L1 sectors per record.  Sums, their order and the epilogue are k_second_half_tile's: results are
// bit-identical (tests/test_gpu_parity.py::test_cell_block_kernel_equals_tiled_kernel).
//
// Why: profiles/r2a_gather_microbench.log - scattered 64-byte gathers: 2 x LDG.256 0.41 lane-records/clk/SM, the
// same gathers from a CTA-staged window at this stride 0.68 (1.65x); k_second_half_tile sits on the former
// (L1 91.6 %, profiles/r1i_full_K2K3_wallprepass_f64_20M.txt).
#pragma once

namespace sphb
{
constexpr int kBlkLy = 3;                       // columns along y per block
constexpr int kBlkPieces = 3 * (kBlkLy + 2);    // candidate pieces per block
constexpr int kBlkThreads = 160;                // one thread per union group (2 particles): <= 320 owned particles
constexpr int kBlkSmemBytes = 100 * 1024;       // dynamic shared memory per CTA (2 CTAs per SM with the static ring + tables)

__device__ unsigned int g_blk_err[8]; // development guard counters (SPHB200_BLOCK_CHECK=1 prints them)

struct BlockDesc
{
    int cx, cy0, z0, z1;
};

template <class R>
__host__ __device__ constexpr int blkCapacity() { return kBlkSmemBytes / int(7 * sizeof(R)); } // 6 R per record + Vol

// Greedy cut of every (x, y-strip) column group along z: extend [z0, z1) while the candidate pieces fit the
// shared-memory capacity and the owned groups fit the CTA.  One thread per strip; empty blocks are skipped.
template <class R>
__global__ void __launch_bounds__(128)
    k_block_plan(int nx, int ny, int nz, const u32 *__restrict__ off, int64_t i0, int64_t i1, u32 *__restrict__ counter,
                 BlockDesc *__restrict__ out, u32 max_blocks)
{
    const int strips_y = (ny + kBlkLy - 1) / kBlkLy;
    const int s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= nx * strips_y)
        return;
    const int cx = s / strips_y, cy0 = (s % strips_y) * kBlkLy, cy1 = min(cy0 + kBlkLy, ny);
    // the strip is empty of owned particles: nothing to do
    {
        const u32 a = off[(size_t(cx) * ny + cy0) * nz], e = off[(size_t(cx) * ny + (cy1 - 1)) * nz + nz];
        if (a == e || int64_t(e) <= i0 || int64_t(a) >= i1)
            return;
    }
    const int cap = blkCapacity<R>();
    auto staged = [&](int z0, int z1)
    {
        const int zl = max(z0 - 1, 0), zh = min(z1 + 1, nz);
        u32 t = 0;
        for (int dx = -1; dx <= 1; ++dx)
        {
            const int x = cx + dx;
            if (x < 0 || x >= nx)
                continue;
            const int ya = max(cy0 - 1, 0), yb = min(cy1, ny - 1);
            for (int y = ya; y <= yb; ++y)
            {
                const size_t col = (size_t(x) * ny + y) * nz;
                t += off[col + zh] - off[col + zl];
            }
        }
        return t;
    };
    auto owned_groups = [&](int z0, int z1)
    {
        u32 g = 0;
        for (int y = cy0; y < cy1; ++y)
        {
            const size_t col = (size_t(cx) * ny + y) * nz;
            const u32 a = off[col + z0], e = off[col + z1];
            g += (e - a + 1) / 2 + 1; // upper bound: a run of n particles touches at most n / 2 + 1 pairs
        }
        return g;
    };
    int z0 = 0;
    while (z0 < nz)
    {
        int z1 = z0 + 1;
        while (z1 < nz && staged(z0, z1 + 1) <= u32(cap) && owned_groups(z0, z1 + 1) <= u32(kBlkThreads))
            ++z1;
        u32 own = 0;
        for (int y = cy0; y < cy1; ++y)
        {
            const size_t col = (size_t(cx) * ny + y) * nz;
            own += off[col + z1] - off[col + z0];
        }
        if (own)
        {
            const u32 k = atomicAdd(counter, 1u);
            if (k < max_blocks)
                out[k] = BlockDesc{cx, cy0, z0, z1};
            // a single cell layer that does not fit (staged > cap or too many groups) is flagged for the caller
            if (staged(z0, z1) > u32(cap) || owned_groups(z0, z1) > u32(kBlkThreads))
                atomicAdd(counter + 1, 1u);
        }
        z0 = z1;
    }
}

template <class R, bool WPRE>
__global__ void __launch_bounds__(kBlkThreads, 2)
    k_second_half_block(const BlockDesc *__restrict__ blocks, int nx, int ny, int nz, const u32 *__restrict__ off, int64_t base, int64_t i0,
                        int64_t i1, const Rec8<R> *__restrict__ rec_b, UListP in, const u32 *__restrict__ wall_count, TileKernelP<R> kern,
                        FluidP<R> fl, R dt, R *__restrict__ pos, R *__restrict__ rho, R *__restrict__ drho, R *__restrict__ force,
                        const R *__restrict__ wall_sum, const unsigned long long *__restrict__ vol_mm)
{
    constexpr int D = 3, G = 2;
    if (!stepDt(fl.ctl, dt))
        return;
    extern __shared__ __align__(16) unsigned char blk_smem[];
    constexpr int CAP = blkCapacity<R>();
    R *srec = reinterpret_cast<R *>(blk_smem);          // [CAP][6]
    R *svol = srec + size_t(CAP) * 6;                   // [CAP]
    __shared__ u32 pc_beg[kBlkPieces], pc_end[kBlkPieces], pc_off[kBlkPieces + 1];
    __shared__ u32 rg_beg[kBlkLy], rg_end[kBlkLy], rg_g0[kBlkLy], rg_pre[kBlkLy + 1];
    __shared__ uint4 ring[kIdxDepth][kBlkThreads];
    const BlockDesc b = blocks[blockIdx.x];
    const int tid = threadIdx.x;
    const int cy1 = min(b.cy0 + kBlkLy, ny);
    if (tid < kBlkPieces)
    {
        const int x = b.cx + tid / (kBlkLy + 2) - 1, y = b.cy0 - 1 + tid % (kBlkLy + 2);
        u32 pb = 0, pe = 0;
        if (x >= 0 && x < nx && y >= 0 && y < ny && y <= cy1)
        {
            const size_t col = (size_t(x) * ny + y) * nz;
            pb = off[col + max(b.z0 - 1, 0)];
            pe = off[col + min(b.z1 + 1, nz)];
        }
        pc_beg[tid] = pb;
        pc_end[tid] = pe;
    }
    if (tid < kBlkLy)
    {
        u32 a = 0, e = 0;
        if (b.cy0 + tid < cy1)
        {
            const size_t col = (size_t(b.cx) * ny + b.cy0 + tid) * nz;
            a = max(off[col + b.z0], u32(i0));
            e = min(off[col + b.z1], u32(i1));
            e = max(e, a);
        }
        rg_beg[tid] = a;
        rg_end[tid] = e;
    }
    __syncthreads();
    if (tid == 0)
    {
        u32 o = 0;
        for (int p = 0; p < kBlkPieces; ++p)
        {
            pc_off[p] = o;
            o += pc_end[p] - pc_beg[p];
        }
        pc_off[kBlkPieces] = o;
        u32 t = 0;
        for (int k = 0; k < kBlkLy; ++k)
        {
            rg_pre[k] = t;
            u32 n = 0, g0 = 0;
            if (rg_end[k] > rg_beg[k])
            {
                g0 = u32((int64_t(rg_beg[k]) - base) >> 1);
                n = u32((int64_t(rg_end[k]) - 1 - base) >> 1) - g0 + 1;
            }
            rg_g0[k] = g0;
            t += n;
        }
        rg_pre[kBlkLy] = t;
    }
    __syncthreads();
    if (pc_off[kBlkPieces] > u32(CAP) || rg_pre[kBlkLy] > u32(kBlkThreads))
    { // cannot happen with a plan of this cell list; refuse instead of overrunning shared memory
        if (tid == 0)
            atomicAdd(&g_blk_err[pc_off[kBlkPieces] > u32(CAP) ? 0 : 1], 1u);
        return;
    }
    // ---- stage the candidate pieces: one thread per record, coalesced 128-bit loads
    for (int p = 0; p < kBlkPieces; ++p)
    {
        const u32 pb = pc_beg[p], len = pc_end[p] - pb, so = pc_off[p];
        for (u32 r = tid; r < len; r += kBlkThreads)
        {
            R v[8];
            ldRec(rec_b, pb + r, v);
            R *dst = srec + size_t(so + r) * 6;
            dst[0] = v[0], dst[1] = v[1], dst[2] = v[2];
            dst[3] = v[5], dst[4] = v[6], dst[5] = v[7];
            svol[so + r] = v[4];
        }
    }
    __syncthreads();
    const unsigned long long vol_lo = __ldg(vol_mm), vol_hi = __ldg(vol_mm + 1);
    const bool vol_uniform = vol_lo == vol_hi;
    const R vol_all = R(__longlong_as_double((long long)vol_lo));
    // ---- thread -> union group
    int k = 0;
#pragma unroll
    for (int q = 1; q < kBlkLy; ++q)
        k += (u32(tid) >= rg_pre[q]) ? 1 : 0;
    const bool gvalid = u32(tid) < rg_pre[kBlkLy];
    const int64_t g = gvalid ? int64_t(rg_g0[k]) + (tid - int(rg_pre[k])) : 0;
    bool ok[G];
    int64_t idx[G];
    R xi[G][D], Vi[G], vi[G][D];
#pragma unroll
    for (int m = 0; m < G; ++m)
    {
        const int64_t i = base + g * G + m;
        ok[m] = gvalid && i >= int64_t(rg_beg[k]) && i < int64_t(rg_end[k]);
        idx[m] = ok[m] ? i : int64_t(rg_beg[k]);
        R ri[8], p_unused;
        ldRec(rec_b, u32(idx[m]), ri);
        recSplit<D, R>(ri, xi[m], Vi[m], p_unused, vi[m]);
    }
    const R k3 = fl.limiter_slope * fl.inv_c0_ave;
    R rate[G], pd[G][D];
#pragma unroll
    for (int m = 0; m < G; ++m)
    {
        rate[m] = R(0);
#pragma unroll
        for (int a = 0; a < D; ++a)
            pd[m][a] = R(0);
    }
    {
        const u32 cnt = gvalid ? __ldg(in.count + g) : 0u;
        const u32 nch = (cnt + kChk - 1) / kChk;
        const uint4 *rows = reinterpret_cast<const uint4 *>(in.list) + (u64(__ldg(in.slice_off + (g >> 5))) * 32u + u32(g & 31));
        // the index ring of tile.cuh, sized for this CTA
#pragma unroll
        for (int s = 0; s < kIdxDepth; ++s)
        {
            if (u32(s) < nch)
                cpAsync16(&ring[s][tid], rows + size_t(s) * 32u);
            cpAsyncCommit();
        }
        int pp = 0; // running piece pointer: the list ascends
        for (u32 c = 0; c < nch; ++c)
        {
            cpAsyncWait<kIdxDepth - 1>();
            const uint4 nxv = ring[c % kIdxDepth][tid];
            const u32 jj[kChk] = {nxv.x, nxv.y, nxv.z, nxv.w};
            R rj[kChk][6], Vj[kChk];
            bool mine[kChk];
#pragma unroll
            for (int s = 0; s < kChk; ++s)
            {
                // A group that straddles the end of an owned run carries the neighbours of its other member too; those
                // may lie outside this block's pieces.  They cannot be neighbours of the owned member: masked like padding.
                u32 slot = 0;
                mine[s] = false;
                if (c * kChk + s < cnt)
                {
                    while (pp < kBlkPieces - 1 && jj[s] >= pc_end[pp])
                        ++pp;
                    mine[s] = jj[s] >= pc_beg[pp] && jj[s] < pc_end[pp];
                    slot = mine[s] ? pc_off[pp] + (jj[s] - pc_beg[pp]) : 0u;
                    if (slot >= pc_off[kBlkPieces])
                    {
                        atomicAdd(&g_blk_err[2], 1u);
                        slot = 0u;
                        mine[s] = false;
                    }
                }
                const R *src = srec + size_t(slot) * 6;
                if (sizeof(R) == 8)
                {
                    const double2 *s2 = reinterpret_cast<const double2 *>(src);
                    const double2 a0 = s2[0], a1 = s2[1], a2 = s2[2];
                    rj[s][0] = R(a0.x), rj[s][1] = R(a0.y), rj[s][2] = R(a1.x);
                    rj[s][3] = R(a1.y), rj[s][4] = R(a2.x), rj[s][5] = R(a2.y);
                }
                else
                {
                    const float2 *s2 = reinterpret_cast<const float2 *>(src);
                    const float2 a0 = s2[0], a1 = s2[1], a2 = s2[2];
                    rj[s][0] = R(a0.x), rj[s][1] = R(a0.y), rj[s][2] = R(a1.x);
                    rj[s][3] = R(a1.y), rj[s][4] = R(a2.x), rj[s][5] = R(a2.y);
                }
                Vj[s] = vol_uniform ? vol_all : svol[slot];
            }
            if (c + kIdxDepth < nch)
                cpAsync16(&ring[c % kIdxDepth][tid], rows + size_t(c + kIdxDepth) * 32u);
            cpAsyncCommit();
#pragma unroll
            for (int s = 0; s < kChk; ++s)
            {
                const R cWVj = mine[s] ? kern.cW * Vj[s] : R(0);
#pragma unroll
                for (int m = 0; m < G; ++m)
                {
                    R d[D], rinv, r, dWV;
                    tilePair<D, R>(xi[m], rj[s], kern, cWVj, d, rinv, r, dWV);
                    R ud = R(0);
#pragma unroll
                    for (int a = 0; a < D; ++a)
                        ud += (vi[m][a] - rj[s][3 + a]) * d[a];
                    const R u = ud * rinv;
                    rate[m] += u * dWV;
                    const R lim = minOne<R>(k3 * maxZero<R>(u));
                    const R pjs = (u * lim) * (dWV * rinv);
#pragma unroll
                    for (int a = 0; a < D; ++a)
                        pd[m][a] += pjs * d[a];
                }
            }
        }
    }
    R rw[G], pw[G][D];
#pragma unroll
    for (int m = 0; m < G; ++m)
    {
        rw[m] = R(0);
#pragma unroll
        for (int a = 0; a < D; ++a)
            pw[m][a] = R(0);
    }
    if (WPRE)
    { // left by k_second_half_wall
        const u32 cnt = gvalid ? __ldg(wall_count + g) : 0u;
        if (cnt)
        {
#pragma unroll
            for (int m = 0; m < G; ++m)
            {
#pragma unroll
                for (int a = 0; a < D; ++a)
                    pw[m][a] = wall_sum[idx[m] * (D + 1) + a];
                rw[m] = wall_sum[idx[m] * (D + 1) + D];
            }
        }
    }
    const R pjump = fl.riemann ? fl.rho0c0_geo_ave : R(0);
#pragma unroll
    for (int m = 0; m < G; ++m)
    {
        if (ok[m])
        {
            const int64_t i = idx[m];
#pragma unroll
            for (int a = 0; a < D; ++a)
                pos[i * D + a] += vi[m][a] * dt * R(0.5);
            const R rhoi = rho[i];
            const R dr = drho[i] + rate[m] * rhoi + rw[m] * rhoi;
            drho[i] = dr;
#pragma unroll
            for (int a = 0; a < D; ++a)
                force[i * D + a] = (pd[m][a] * pjump) * Vi[m] + (pw[m][a] * pjump) * Vi[m];
            rho[i] = rhoi + dr * dt * R(0.5);
        }
    }
}
} // namespace sphb
