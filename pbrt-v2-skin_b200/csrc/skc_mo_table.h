// skc_mo_table.h -- the compact, shared-abscissa form of the Rd profile tables that the Mo kernel keeps in
// shared memory (k_mo_lambda, skc_mo.cu), and the exact rule it falls back to.
//
// Reference: core/multipole.cpp:60-73 (sampleProfile): per wavelength its own table (119 766 entries at
// desiredLength 512) and its own rcpDsqSpacing; value = lerp of data[seg], data[seg+1] at f = d2 * rcp, 0 once
// f >= len-1.  Thirty scattered 8-byte reads per contribution.
//
// Here: ONE abscissa for all wavelengths, u = M * log2(1 + d2/a), a row of 32 floats per integer u, 4-point
// Lagrange (cubic) interpolation in u -- a few hundred rows (tens of KB) hold every wavelength's profile over the
// range where it is smooth.  The profile tables are NOT smooth everywhere at the 1e-5 level: near d2 = 0 the long
// wavelengths show the kinks of the reference's own resampling (piecewise linear between the i^2+j^2 radii; up to
// 6e-5 of the value), and the far half of every table carries the fp32 noise of the dipole sums spread by the
// layer convolution (1e-8 of the peak, i.e. percents of the local value).  The builder therefore VERIFIES the
// compact form against sampleProfile at four points of every table segment and gives each wavelength a window
// [fLo, fHi) of f inside which the compact value is within `tol` (relative to the exact value) -- outside the
// window the kernel does the exact lookup: from a copy of the table's head in shared memory (f < headLen), else
// from the full table in global memory.  So every lookup is either exact or within tol of exact by construction.
#pragma once
#include <stdint.h>
#include <math.h>
#include <vector>

#ifdef __CUDACC__
#define SKC_HD __host__ __device__ __forceinline__
#else
#define SKC_HD inline
#endif

#define MOC_ROWS_MAX 512

struct MoCompactParams {
    float invA;          // 1/a
    float cellsPerBinade;// M (integer valued)
    int K;               // integer abscissae 0..K-1; rows stored for u = -1..K+1 (K+3 rows, row index = u+1)
};

// Abscissa of d2: cell k (0 <= k) and the four Lagrange weights of rows k-1, k, k+1, k+2 (array rows k..k+3).
// lg2m = log2 of the mantissa in [1,2): the device uses MUFU.LG2 (absolute error 2^-22 there), the host log2f; the
// difference moves t by < 1e-5 of a cell, the value by < 1e-6 of itself (covered by the builder's margin).
SKC_HD void mocCell(const MoCompactParams &P, float d2, int &k, float &t) {
    float y = d2 * P.invA + 1.f;
#ifdef __CUDA_ARCH__
    int bits = __float_as_int(y);
    float m = __int_as_float((bits & 0x007fffff) | 0x3f800000);
    float lg; asm("lg2.approx.ftz.f32 %0, %1;" : "=f"(lg) : "f"(m));
#else
    union { float f; int i; } cv; cv.f = y;
    int bits = cv.i;
    cv.i = (bits & 0x007fffff) | 0x3f800000;
    float lg = log2f(cv.f);
#endif
    int e = (bits >> 23) - 127;
    float s = lg * P.cellsPerBinade;
    int ks = (int)s;
    t = s - (float)ks;
    k = e * (int)P.cellsPerBinade + ks;
}
// 4-point Lagrange weights of the abscissae -1, 0, 1, 2 at t in [0, 1)
SKC_HD void mocWeights(float t, float (&w)[4]) {
    float tm1 = t - 1.f, tm2 = t - 2.f, tp1 = t + 1.f;
    float a = t * tm1, b = tp1 * tm2;
    w[0] = (a * tm2) * (-1.f / 6.f);
    w[1] = (b * tm1) * 0.5f;
    w[2] = (b * t) * (-0.5f);
    w[3] = (a * tp1) * (1.f / 6.f);
}
SKC_HD void mocCoord(const MoCompactParams &P, float d2, int &k, float (&w)[4]) {
    float t;
    mocCell(P, d2, k, t);
    mocWeights(t, w);
}

// the interpolated value from the four rows' entries (fused multiply-adds on both sides, so host and device agree)
SKC_HD float mocCubic(const float (&w)[4], float r0, float r1, float r2, float r3) {
#ifdef __CUDA_ARCH__
    return __fmaf_rn(w[0], r0, __fmaf_rn(w[1], r1, __fmaf_rn(w[2], r2, w[3] * r3)));
#else
    return fmaf(w[0], r0, fmaf(w[1], r1, fmaf(w[2], r2, w[3] * r3)));
#endif
}

// sampleProfile (core/multipole.cpp:60-73), the reference's arithmetic: fp32 product, exact compare and fraction
SKC_HD float mocExact(const float *data, uint32_t len, float rcp, float d2) {
    float f = d2 * rcp;
    if (f >= (float)(len - 1)) return 0.f;
    uint32_t seg = (uint32_t)f;
    float t = f - (float)seg;
    return (1.f - t) * data[seg] + t * data[seg + 1];
}

struct MoCompactTable {
    MoCompactParams P;
    std::vector<float> rows;        // [(K+3)][32]
    std::vector<float> heads;       // concatenated table heads: wavelength l at headOff[l], headLen[l]+1 entries
    uint32_t headOff[32], headLen[32];
    float fLo[32], fHi[32];         // compact window per wavelength, in units of f = d2 * rcp (fLo >= headLen possible)
    float maxErr[32];               // largest verified relative error inside the window
    double coverage;                // share of table segments (log-weighted in f) the window covers, mean over wavelengths
    float tol;
};

// Builds the compact table for 30 exact tables table[l][len] with spacings rcp[l]; K rows of M cells per binade,
// window tolerance tol, at most headCapFloats floats of shared memory for the heads.  Returns false if no wavelength
// got a window (the caller then stays on the exact kernel).
bool mocBuild(const float *table, const float *rcp, uint32_t len, float tol, int K, int M, uint32_t headCapFloats, MoCompactTable &out);
// value the kernel computes for (wavelength l, d2): head / compact / exact tail
float mocEvalHost(const MoCompactTable &C, const float *table, const float *rcp, uint32_t len, int l, float d2);
