// small dense matrices in registers (shared by solid.cu and ck.cu)
#pragma once
#include "internal.cuh"

namespace sphb
{
template <int D, class R>
struct SMat
{
    R a[D][D];
    __device__ static SMat zero()
    {
        SMat m;
#pragma unroll
        for (int i = 0; i < D; ++i)
#pragma unroll
            for (int j = 0; j < D; ++j)
                m.a[i][j] = R(0);
        return m;
    }
    __device__ static SMat identity()
    {
        SMat m = zero();
#pragma unroll
        for (int i = 0; i < D; ++i)
            m.a[i][i] = R(1);
        return m;
    }
    __device__ static SMat load(const R *p)
    {
        SMat m;
#pragma unroll
        for (int i = 0; i < D; ++i)
#pragma unroll
            for (int j = 0; j < D; ++j)
                m.a[i][j] = p[i * D + j];
        return m;
    }
    __device__ void store(R *p) const
    {
#pragma unroll
        for (int i = 0; i < D; ++i)
#pragma unroll
            for (int j = 0; j < D; ++j)
                p[i * D + j] = a[i][j];
    }
    __device__ SMat T() const
    {
        SMat m;
#pragma unroll
        for (int i = 0; i < D; ++i)
#pragma unroll
            for (int j = 0; j < D; ++j)
                m.a[i][j] = a[j][i];
        return m;
    }
    __device__ SMat operator*(const SMat &o) const
    {
        SMat m = zero();
#pragma unroll
        for (int i = 0; i < D; ++i)
#pragma unroll
            for (int j = 0; j < D; ++j)
#pragma unroll
                for (int k = 0; k < D; ++k)
                    m.a[i][j] += a[i][k] * o.a[k][j];
        return m;
    }
    __device__ SMat operator*(R s) const
    {
        SMat m;
#pragma unroll
        for (int i = 0; i < D; ++i)
#pragma unroll
            for (int j = 0; j < D; ++j)
                m.a[i][j] = a[i][j] * s;
        return m;
    }
    __device__ SMat operator+(const SMat &o) const
    {
        SMat m;
#pragma unroll
        for (int i = 0; i < D; ++i)
#pragma unroll
            for (int j = 0; j < D; ++j)
                m.a[i][j] = a[i][j] + o.a[i][j];
        return m;
    }
    __device__ SMat operator-(const SMat &o) const
    {
        SMat m;
#pragma unroll
        for (int i = 0; i < D; ++i)
#pragma unroll
            for (int j = 0; j < D; ++j)
                m.a[i][j] = a[i][j] - o.a[i][j];
        return m;
    }
    __device__ R trace() const
    {
        R t = R(0);
#pragma unroll
        for (int i = 0; i < D; ++i)
            t += a[i][i];
        return t;
    }
    __device__ R det() const
    {
        if (D == 2)
            return a[0][0] * a[1][1] - a[0][1] * a[1][0];
        return a[0][0] * (a[1][1] * a[D - 1][D - 1] - a[1][D - 1] * a[D - 1][1]) -
               a[0][1] * (a[1][0] * a[D - 1][D - 1] - a[1][D - 1] * a[D - 1][0]) +
               a[0][D - 1] * (a[1][0] * a[D - 1][1] - a[1][1] * a[D - 1][0]);
    }
    __device__ SMat inverse() const
    {
        SMat m;
        R inv = R(1.0) / det();
        if (D == 2)
        {
            m.a[0][0] = a[1][1] * inv;
            m.a[0][1] = -a[0][1] * inv;
            m.a[1][0] = -a[1][0] * inv;
            m.a[1][1] = a[0][0] * inv;
        }
        else
        {
#pragma unroll
            for (int i = 0; i < D; ++i)
#pragma unroll
                for (int j = 0; j < D; ++j)
                {
                    const int i1 = (i + 1) % D, i2 = (i + 2) % D, j1 = (j + 1) % D, j2 = (j + 2) % D;
                    m.a[i][j] = (a[j1][i1] * a[j2][i2] - a[j1][i2] * a[j2][i1]) * inv;
                }
        }
        return m;
    }
    __device__ void mulVec(const R *v, R *out) const
    {
#pragma unroll
        for (int i = 0; i < D; ++i)
        {
            out[i] = R(0);
#pragma unroll
            for (int j = 0; j < D; ++j)
                out[i] += a[i][j] * v[j];
        }
    }
};
} // namespace sphb
