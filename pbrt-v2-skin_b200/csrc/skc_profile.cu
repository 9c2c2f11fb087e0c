// skc_profile.cu -- MPC_ComputeDiffusionProfile (+ uniform-r^2 resample) on the device, batched.
//
// Reference: multipole/MultipoleProfileCalculator/MultipoleProfileCalculator.cpp:151-230
// (ComputeLayerProfile), :233-280 (To{Frequency,Time}Domain, CombineLayerProfiles), :294-346
// (MPC_ComputeDiffusionProfile), :355-426 (resample), DipoleCalculator.cpp:38-91, numutil.h
// (ScaleAndShift, Sum); FFT = kissfft 1.3.0 kiss_fftndr in double.
//
// Per parameter set (one wavelength of one layer stack):
//   k_layer_profile   one thread per octant cell of the (2L)^2 grid: 11 dipole pairs in the
//                     reference's fp32 arithmetic, accumulated in double in its order, thin-slab
//                     lerp, centre delta, 8-fold mirroring -- writes z = R + iT
//   combine           the layer grids are real and even, so their spectra are real: ONE complex
//                     FFT of z = R + iT yields FR (real part) and FT (imaginary part).  Two
//                     forward 2-D FFTs (4L)^2 + pointwise adding equations + one inverse replace the
//                     reference's four real forward and two inverse transforms.
//   k_fft_rows        one CTA per row, row in shared memory (double2), radix-2 DIT, twiddle table
//   k_transpose       32x32 tiles through shared memory
//   k_extract/k_resample  the reference's unique-radius list (host-built index table, it is
//                     data independent) and its lower_bound + fp32 lerp
#include <algorithm>
#include <map>
#include <thread>
#include <vector>
#include <math.h>
#include <string.h>
#include "skc_internal.h"

// ---------------------------------------------------------------------------------------------
#define SKC_PROFILE_MAX_LAYERS 16
struct DipoleSet {            // DipoleCalculator members for zi = -5..5 (DipoleCalculator.cpp:51-69)
    float zpos[11], zneg[11];
    float d, sigma_tr, alphap;
    double lerp;              // thin-slab factor (MultipoleProfileCalculator.cpp:163-171)
    double normalize;         // stepSize*stepSize
    float step2;              // (float)(stepSize*stepSize)
};

static float FresnelDiffuseReflectance(float eta) {   // DipoleCalculator.cpp:38-45
    if (eta >= 1.f) return -1.4399f / (eta * eta) + 0.7099f / eta + 0.6681f + 0.0636f * eta;
    float eta2 = eta * eta;
    return -0.4399f + 0.7099f / eta - 0.3319f / eta2 + 0.0636f / (eta2 * eta);
}

static DipoleSet makeDipoles(const skc_layer_spec &spec, float iorUpper, float iorLower, float stepSize, bool lerpOnThinSlab) {
    DipoleSet ds;
    float thickness = spec.thickness;
    double mfp2 = 2. / (spec.mua + spec.musp);
    double lerp = 1.;
    if (lerpOnThinSlab) {
        lerp = (thickness < mfp2) ? (1. - exp(-thickness * 2. / mfp2)) / (1. - exp(-2.)) : 1.;
        if (thickness < 0.01 * mfp2) thickness = (float)(0.01 * mfp2);
    }
    ds.lerp = lerp;
    ds.step2 = stepSize * stepSize;
    ds.normalize = (double)ds.step2;
    float sigma_a = spec.mua, sigmap_s = spec.musp, d = thickness;
    float sigmap_t = sigma_a + sigmap_s;
    ds.d = d;
    ds.sigma_tr = sqrtf(3 * sigma_a * sigmap_t);
    ds.alphap = sigmap_s / sigmap_t;
    float Fdr0 = FresnelDiffuseReflectance(iorUpper), Fdrd = FresnelDiffuseReflectance(iorLower);
    float A_0 = (1.f + Fdr0) / (1.f - Fdr0), A_d = (1.f + Fdrd) / (1.f - Fdrd);
    float D = 1.f / (3.f * sigmap_t);
    float zb_0 = 2.f * A_0 * D, zb_d = 2.f * A_d * D;
    float l = 1.f / sigmap_t;
    if (lerpOnThinSlab && l > d * .5f) l = d * .5f;
    for (int zi = -5; zi <= 5; ++zi) {
        float zpos = 2.f * (float)zi * (d + zb_0 + zb_d) + l;
        ds.zpos[zi + 5] = zpos;
        ds.zneg[zi + 5] = zpos - 2.f * (l + zb_0);
    }
    return ds;
}

// expf as glibc >= 2.27 computes it (sysdeps/ieee754/flt-32/e_expf.c: 32-entry 2^(i/32) table, cubic in
// double, one rounding to float), so that the dipole sums -- which cancel heavily for thin slabs without the
// lerp -- see the same fp32 inputs as the reference run on a glibc host.  A correctly rounded expf differs from
// it in ~5e-4 of all arguments by one ulp, which the cancellation amplifies to ~1e-5 of the table value.
__constant__ unsigned long long c_exp2Tab[32];
static void uploadExpTable(cudaStream_t st) {
    unsigned long long tab[32];
    for (int i = 0; i < 32; ++i) {
        double v = (double)exp2l(i / 32.0L);
        unsigned long long u; memcpy(&u, &v, 8);
        tab[i] = u - (((unsigned long long)i << 52) / 32);
    }
    cudaMemcpyToSymbolAsync(c_exp2Tab, tab, sizeof(tab), 0, cudaMemcpyHostToDevice, st);
    cudaStreamSynchronize(st);
}
__device__ __forceinline__ float expf_glibc(float x, const unsigned long long *tab) {
    if (x < -0x1.9fe368p6f) return 0.f;
    const double N = 32, shift = 0x1.8p+52, inv = 0x1.71547652b82fep+0 * N;
    const double C0 = 0x1.c6af84b912394p-5 / N / N / N, C1 = 0x1.ebfce50fac4f3p-3 / N / N, C2 = 0x1.62e42ff0c52d6p-1 / N;
    // libm's FMA build (the one selected on the reference host's CPU family): the scaled argument enters both steps fused
    double z, kd = __fma_rn(inv, (double)x, shift);
    unsigned long long ki = (unsigned long long)__double_as_longlong(kd);
    kd -= shift;
    double r = __fma_rn(inv, (double)x, -kd);
    double s = __longlong_as_double((long long)(tab[ki & 31] + (ki << 47)));
    z = C0 * r + C1;
    double r2 = r * r, y = C2 * r + 1.;
    y = z * r2 + y;
    return (float)(y * s);
}

#define INV_FOURPI_F (0.25f / 3.141592654f)   /* numutil.h:45-46 */

// z[G][G] (double2 = {R, T}); one thread per (row<=col) cell of the [0,extent]^2 quadrant
// blockIdx.z = set of the batch (its DipoleSet and its grid)
__global__ void k_layer_profile(const DipoleSet *__restrict__ dsAll, uint32_t G, double2 *__restrict__ z, uint32_t *__restrict__ tNonZero) {
    const DipoleSet ds = dsAll[blockIdx.z];
    z += (size_t)blockIdx.z * G * G;
    uint32_t center = (G - 1) / 2, extent = center;
    __shared__ unsigned long long sTab[32];
    if (threadIdx.y == 0) sTab[threadIdx.x] = c_exp2Tab[threadIdx.x];
    __syncthreads();
    uint32_t col = blockIdx.x * blockDim.x + threadIdx.x, row = blockIdx.y * blockDim.y + threadIdx.y;
    if (row > extent || col > extent || col < row) return;
    double drow2 = (double)(row * row), dcol2 = (double)(col * col);
    double r2 = (drow2 + dcol2) * (double)ds.step2;
    float dsq = (float)r2;
    double R = 0., T = 0.;
    for (int k = 0; k < 11; ++k) {
        float zpos = ds.zpos[k], zneg = ds.zneg[k], d = ds.d, str = ds.sigma_tr;
        {   // DipoleCalculator::Rd (DipoleCalculator.cpp:71-80)
            float dpos = sqrtf(dsq + zpos * zpos), dneg = sqrtf(dsq + zneg * zneg);
            float dpos3 = dpos * dpos * dpos, dneg3 = dneg * dneg * dneg;
            float rd = ds.alphap * INV_FOURPI_F * (zpos * (1 + str * dpos) * expf_glibc(-str * dpos, sTab) / dpos3 -
                                                   zneg * (1 + str * dneg) * expf_glibc(-str * dneg, sTab) / dneg3);
            R += (double)rd * ds.normalize;
        }
        {   // DipoleCalculator::Td (:82-91)
            float dpos = sqrtf(dsq + (d - zpos) * (d - zpos)), dneg = sqrtf(dsq + (d - zneg) * (d - zneg));
            float dpos3 = dpos * dpos * dpos, dneg3 = dneg * dneg * dneg;
            float td = ds.alphap * INV_FOURPI_F * ((d - zpos) * (1 + str * dpos) * expf_glibc(-str * dpos, sTab) / dpos3 -
                                                   (d - zneg) * (1 + str * dneg) * expf_glibc(-str * dneg, sTab) / dneg3);
            T += (double)td * ds.normalize;
        }
    }
    if (ds.lerp < 1.) {
        R *= ds.lerp; T *= ds.lerp;
        if (row == 0 && col == 0) T += 1. - ds.lerp;
    }
    // see k_unpack; 1e-36: below it the reference's fp32 dipole terms are denormal or zero
    if (fabs(T) >= 1e-36 * ds.normalize && ((volatile uint32_t *)tNonZero)[blockIdx.z] == 0u) tNonZero[blockIdx.z] = 1u;
    double2 v = make_double2(R, T);
    // symmetric fill (MultipoleProfileCalculator.cpp:210-229)
    uint32_t rr[2] = {center + row, center - row}, cc[2] = {center + col, center - col};
    for (int a = 0; a < 2; ++a)
        for (int b = 0; b < 2; ++b) {
            z[(size_t)rr[a] * G + cc[b]] = v;
            z[(size_t)cc[b] * G + rr[a]] = v;
        }
}

// Z[N][N] <- ScaleAndShift(grid, center) (numutil.h:192-204): grid cell (i,j) -> ((i-c) mod N, (j-c) mod N)
// Only the G rows that hold grid cells are written (blockIdx.y counts from the first of them, (N - c) mod N): the other rows
// are zero by construction and nothing reads them -- the first row pass skips them and the transpose substitutes zeros.
__global__ void k_pack(const double2 *__restrict__ z, uint32_t G, uint32_t N, double2 *__restrict__ Z) {
    uint32_t jj = blockIdx.x * blockDim.x + threadIdx.x, ii = (N - (G - 1) / 2 + blockIdx.y) & (N - 1);
    if (jj >= N) return;
    z += (size_t)blockIdx.z * G * G; Z += (size_t)blockIdx.z * N * N;
    uint32_t c = (G - 1) / 2;
    uint32_t i = (ii + c) % N, j = (jj + c) % N;
    Z[(size_t)ii * N + jj] = (i < G && j < G) ? z[(size_t)i * G + j] : make_double2(0., 0.);
}

// In-place FFT of every row of Z[rows][N], N a power of two <= 4096, one CTA per row, the row in shared memory.
// Radix-2 decimation in time on bit-reversed input; up to three consecutive radix-2 stages are done per pass on 8
// values held in registers (same butterflies, twiddles and operation order as stage-by-stage radix-2, so the result is
// bit-identical to it) -- 4 shared-memory round trips and barriers for N = 2048 instead of 11.
// shared-memory index padding: one 16 B slot after every 8 and every 64 elements, so that the stride-8 (first pass),
// 8-consecutive-then-stride-64 (second pass) and bit-reversed (load) patterns spread over the banks
__device__ __forceinline__ uint32_t fftSlot(uint32_t i) { return i + (i >> 3) + (i >> 6); }

template <int R>
__device__ __forceinline__ void fftPass(double2 *s, uint32_t N, uint32_t st, const double2 *__restrict__ tw, int inverse) {
    const uint32_t half = 1u << (st - 1), items = N >> R;
    for (uint32_t item = threadIdx.x; item < items; item += blockDim.x) {
        uint32_t pos = item & (half - 1), base = ((item >> (st - 1)) << (st - 1 + R)) + pos;
        double2 v[1 << R];
#pragma unroll
        for (int k = 0; k < (1 << R); ++k) v[k] = s[fftSlot(base + k * half)];
#pragma unroll
        for (int q = 0; q < R; ++q) {
            const int hs = 1 << q;
            const uint32_t stride = N >> (st + q);
#pragma unroll
            for (int k = 0; k < (1 << R); ++k) {
                if (k & hs) continue;
                double2 w = __ldg(&tw[(pos + (k & (hs - 1)) * half) * stride]);
                if (inverse) w.y = -w.y;
                double2 a = v[k], c = v[k + hs];
                double2 t = make_double2(w.x * c.x - w.y * c.y, w.x * c.y + w.y * c.x);
                v[k] = make_double2(a.x + t.x, a.y + t.y);
                v[k + hs] = make_double2(a.x - t.x, a.y - t.y);
            }
        }
#pragma unroll
        for (int k = 0; k < (1 << R); ++k) s[fftSlot(base + k * half)] = v[k];
    }
}

// rows (rowStart + blockIdx.x) mod N of every set (blockIdx.y): the whole matrix, or the band of rows that is not all zero
// (first pass of a forward transform of the zero-padded grid) / that anything reads afterwards (last pass of the inverse)
__global__ void __launch_bounds__(512) k_fft_rows(double2 *__restrict__ Z, uint32_t N, uint32_t logN, const double2 *__restrict__ tw, int inverse,
                                                  uint32_t rowStart) {
    extern __shared__ double2 s[];
    double2 *row = Z + ((size_t)blockIdx.y * N + ((rowStart + blockIdx.x) & (N - 1))) * N;
    for (uint32_t i = threadIdx.x; i < N; i += blockDim.x) s[fftSlot(__brev(i) >> (32 - logN))] = row[i];
    __syncthreads();
    uint32_t st = 1;
    while (st <= logN) {
        uint32_t left = logN - st + 1;
        if (left >= 3) { fftPass<3>(s, N, st, tw, inverse); st += 3; }
        else if (left == 2) { fftPass<2>(s, N, st, tw, inverse); st += 2; }
        else { fftPass<1>(s, N, st, tw, inverse); st += 1; }
        __syncthreads();
    }
    for (uint32_t i = threadIdx.x; i < N; i += blockDim.x) row[i] = s[fftSlot(i)];
}

// ---- N = 2048 (desiredLength 512, the size BASELINE.json configs[4] and skin.pbrt use): the same radix-2 DIT butterflies in
// the same order (bit-identical results), re-tiled so that a row crosses shared memory twice instead of five times:
//   pass 1 = stages 1..4 on 16 values that come STRAIGHT FROM GLOBAL MEMORY: position i of the bit-reversed sequence holds
//            x[brev(i)], so the 16 inputs of item brev7(j) are x[j + brev4(k) * 128] -- for a fixed k the 128 threads of a row read
//            128 consecutive elements (coalesced), and no bit-reversed scatter through shared memory is needed;
//   pass 2 = stages 5..8, 16 values per thread, shared memory in place;
//   pass 3 = stages 9..11 on 8 values, whose results are positions pos + k * 256 of the natural-order output: written straight
//            to global memory, coalesced.
// 128 threads per row, two rows per CTA (independent: named barriers), 33 KB of shared memory per row with the index padding
// i + (i >> 7) + (i >> 10), which makes every 16-byte access pattern of the three passes conflict-free per quarter-warp.
// Shared-memory traffic per row: 128 KB + 58 KB of twiddles (k_fft_rows: 320 KB + 110 KB).
#define FFT2_N 2048u
#define FFT2_SLOTS 2080u      // >= slot(2047) + 1 = 2064, kept a multiple of 16
__device__ __forceinline__ uint32_t fft2Slot(uint32_t i) { return i + (i >> 7) + (i >> 10); }

template <int R>
__device__ __forceinline__ void fftButterflies(double2 (&v)[1 << R], uint32_t pos, uint32_t half, uint32_t st, uint32_t N,
                                               const double2 *__restrict__ tw, int inverse) {
#pragma unroll
    for (int q = 0; q < R; ++q) {
        const int hs = 1 << q;
        const double2 *stage = tw + (N >> 1) + (1u << (st + q - 1)) - 1u;       // this stage's own contiguous copy of its twiddles
#pragma unroll
        for (int k = 0; k < (1 << R); ++k) {
            if (k & hs) continue;
            double2 w = __ldg(&stage[pos + (k & (hs - 1)) * half]);
            if (inverse) w.y = -w.y;
            double2 a = v[k], c = v[k + hs];
            double2 t = make_double2(w.x * c.x - w.y * c.y, w.x * c.y + w.y * c.x);
            v[k] = make_double2(a.x + t.x, a.y + t.y);
            v[k + hs] = make_double2(a.x - t.x, a.y - t.y);
        }
    }
}

__global__ void __launch_bounds__(256) k_fft_rows2048(double2 *__restrict__ Z, const double2 *__restrict__ tw, int inverse,
                                                      uint32_t rowStart, uint32_t nRows) {
    constexpr uint32_t N = FFT2_N;
    extern __shared__ double2 s2[];
    const uint32_t sub = threadIdx.x >> 7, t = threadIdx.x & 127;
    const uint32_t r = blockIdx.x * 2 + sub;
    if (r >= nRows) return;                  // the whole 128-thread half leaves; the barriers below are per half
    double2 *s = s2 + sub * FFT2_SLOTS;
    double2 *row = Z + ((size_t)blockIdx.y * N + ((rowStart + r) & (N - 1))) * N;
    const uint32_t barId = 1 + sub;
    {
        double2 v[16];
#pragma unroll
        for (int k = 0; k < 16; ++k) v[k] = row[t + (uint32_t)(((k & 1) << 3) | ((k & 2) << 1) | ((k & 4) >> 1) | ((k & 8) >> 3)) * 128u];
        fftButterflies<4>(v, 0u, 1u, 1u, N, tw, inverse);
        const uint32_t base = (__brev(t) >> 25) * 16u;
#pragma unroll
        for (int k = 0; k < 16; ++k) s[fft2Slot(base + k)] = v[k];
    }
    asm volatile("bar.sync %0, 128;" ::"r"(barId) : "memory");
    {
        double2 v[16];
        const uint32_t pos = t & 15u, base = (t >> 4) * 256u + pos;
#pragma unroll
        for (int k = 0; k < 16; ++k) v[k] = s[fft2Slot(base + k * 16u)];
        fftButterflies<4>(v, pos, 16u, 5u, N, tw, inverse);
#pragma unroll
        for (int k = 0; k < 16; ++k) s[fft2Slot(base + k * 16u)] = v[k];
    }
    asm volatile("bar.sync %0, 128;" ::"r"(barId) : "memory");
#pragma unroll 1
    for (uint32_t pos = t; pos < 256u; pos += 128u) {
        double2 u[8];
#pragma unroll
        for (int k = 0; k < 8; ++k) u[k] = s[fft2Slot(pos + k * 256u)];
        fftButterflies<3>(u, pos, 256u, 9u, N, tw, inverse);
#pragma unroll
        for (int k = 0; k < 8; ++k) row[pos + k * 256u] = u[k];
    }
}

// mode 0: all of it.  mode 1 (forward transform): input rows outside the band [bandStart, bandStart + G) mod N are zero and are
// not read.  mode 2 (inverse): only the output rows inside the band are needed (the last row pass and k_unpack touch no others):
// the other input columns are neither read nor written.
__global__ void k_transpose(const double2 *__restrict__ in, uint32_t N, double2 *__restrict__ out, int mode, uint32_t bandStart, uint32_t G) {
    __shared__ double2 tile[32][33];
    in += (size_t)blockIdx.z * N * N; out += (size_t)blockIdx.z * N * N;
    uint32_t x = blockIdx.x * 32 + threadIdx.x, y0 = blockIdx.y * 32;
    const bool xIn = mode != 2 || ((x - bandStart) & (N - 1)) < G;
    for (uint32_t k = threadIdx.y; k < 32; k += blockDim.y) {
        const bool yIn = mode != 1 || ((y0 + k - bandStart) & (N - 1)) < G;
        tile[k][threadIdx.x] = (xIn && yIn) ? in[(size_t)(y0 + k) * N + x] : make_double2(0., 0.);
    }
    __syncthreads();
    uint32_t xo = blockIdx.y * 32 + threadIdx.x, yo0 = blockIdx.x * 32;
    for (uint32_t k = threadIdx.y; k < 32; k += blockDim.y) {
        if (mode == 2 && ((yo0 + k - bandStart) & (N - 1)) >= G) continue;
        out[(size_t)(yo0 + k) * N + xo] = tile[threadIdx.x][k];
    }
}

// adding equations in the frequency domain (MultipoleProfileCalculator.cpp:253-280).  R and T are real, so the
// transform of z = R + iT separates into the two Hermitian spectra FR(k) = (Z(k) + conj Z(-k))/2 and
// FT(k) = (Z(k) - conj Z(-k))/(2i); a fresh layer grid is even (real spectra) but a combined one is not quite
// (its last row/column has no mirror partner), so the general form is used throughout.  The result is packed
// the same way, W = FR12 + i FT12, and one inverse transform returns R12 in the real and T12 in the imaginary part.
__device__ __forceinline__ double2 cmul(double2 a, double2 b) { return make_double2(a.x * b.x - a.y * b.y, a.x * b.y + a.y * b.x); }
__device__ __forceinline__ double2 cdiv(double2 a, double2 b) {
    double d = b.x * b.x + b.y * b.y;
    return make_double2((a.x * b.x + a.y * b.y) / d, (a.y * b.x - a.x * b.y) / d);
}
__device__ __forceinline__ void splitSpectra(double2 zk, double2 zm, double2 &FR, double2 &FT) {
    // zm = Z(-k); conj(zm) = (zm.x, -zm.y)
    FR = make_double2(0.5 * (zk.x + zm.x), 0.5 * (zk.y - zm.y));
    double dx = zk.x - zm.x, dy = zk.y + zm.y;              // Z(k) - conj Z(-k)
    FT = make_double2(0.5 * dy, -0.5 * dx);                // divided by 2i
}
__device__ __forceinline__ double2 addingAt(double2 z1k, double2 z1m, double2 z2k, double2 z2m) {
    double2 R1, T1, R2, T2;
    splitSpectra(z1k, z1m, R1, T1);
    splitSpectra(z2k, z2m, R2, T2);
    double2 one = cmul(R2, R1);
    one = make_double2(1. - one.x, -one.y);
    double2 R12 = cdiv(cmul(cmul(T1, R2), T1), one);
    R12.x += R1.x; R12.y += R1.y;
    double2 T12 = cdiv(cmul(T1, T2), one);
    return make_double2(R12.x - T12.y, R12.y + T12.x);
}
// One thread per PAIR of mirror frequencies {k, -k}: W(k) needs Z(k) and Z(-k) of both layers and W(-k) needs the same four
// values, so each element is read once instead of twice.  blockIdx.y = row a in [0, N/2]; rows 0 and N/2 are their own mirrors
// (there the pair partner is the mirrored column of the same row and only columns <= N/2 start a pair).
__global__ void k_adding(const double2 *__restrict__ Z1, const double2 *__restrict__ Z2, uint32_t N, double2 *__restrict__ W) {
    uint32_t b = blockIdx.x * blockDim.x + threadIdx.x, a = blockIdx.y;
    if (b >= N) return;
    const uint32_t am = (N - a) & (N - 1), bm = (N - b) & (N - 1);
    if (am == a && b > bm) return;
    Z1 += (size_t)blockIdx.z * N * N; Z2 += (size_t)blockIdx.z * N * N; W += (size_t)blockIdx.z * N * N;
    const size_t i = (size_t)a * N + b, j = (size_t)am * N + bm;
    const double2 z1i = Z1[i], z1j = Z1[j], z2i = Z2[i], z2j = Z2[j];
    W[i] = addingAt(z1i, z1j, z2i, z2j);
    if (j != i) W[j] = addingAt(z1j, z1i, z2j, z2i);
}

// ScaleAndShiftReversed (numutil.h:205-217) of the inverse transform, scaled by 1/N^2 (:132-139)
// A layer that transmits nothing (every dipole term underflows or is denormal in fp32: an optically thick slab) has T == 0, or
// ~1e-41, in every cell, and the reference, which transforms R and T separately, then gets T12 == 0 exactly (or denormals that
// its float tables round to 0).  Here T travels as the imaginary part of R + iT and would come back as rounding noise (1e-24 of
// R); tNonZero[layer][set] (k_layer_profile) says which layers had any representable T, and a stack with a layer that had
// none gets the reference's zero.
__global__ void k_unpack(const double2 *__restrict__ W, uint32_t N, uint32_t G, double2 *__restrict__ z, const uint32_t *__restrict__ tNonZero,
                         uint32_t nLayersSoFar, uint32_t setStride) {
    uint32_t j = blockIdx.x * blockDim.x + threadIdx.x, i = blockIdx.y;
    if (j >= G) return;
    W += (size_t)blockIdx.z * N * N; z += (size_t)blockIdx.z * G * G;
    bool opaque = false;
    for (uint32_t l = 0; l < nLayersSoFar; ++l) opaque |= tNonZero[(size_t)l * setStride + blockIdx.z] == 0u;
    uint32_t c = (G - 1) / 2;
    uint32_t ii = (N - c + i) % N, jj = (N - c + j) % N;
    double scale = (double)1 / ((double)N * (double)N);
    double2 v = W[(size_t)ii * N + jj];
    z[(size_t)i * G + j] = make_double2(v.x * scale, opaque ? 0. : v.y * scale);
}

// block sums of R and T over the grid (totalReflectance / totalTransmittance, :338-339)
__global__ void k_grid_sum(const double2 *__restrict__ z, size_t n, double2 *__restrict__ partial) {
    __shared__ double2 s[256];
    z += (size_t)blockIdx.y * n; partial += (size_t)blockIdx.y * gridDim.x;      // blockIdx.y = set
    double r = 0., t = 0.;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) { r += z[i].x; t += z[i].y; }
    s[threadIdx.x] = make_double2(r, t);
    __syncthreads();
    for (int st = 128; st > 0; st >>= 1) {
        if (threadIdx.x < st) { s[threadIdx.x].x += s[threadIdx.x + st].x; s[threadIdx.x].y += s[threadIdx.x + st].y; }
        __syncthreads();
    }
    if (threadIdx.x == 0) partial[blockIdx.x] = s[0];
}

// unique radii (:316-330): entry e <- grid cell idx[e] = (i, j); d2 = (float)(i*i+j*j)*step*step
// blockIdx.y = set: its grid, its step size, its [d2 | R | T] rows of nEntries floats each
__global__ void k_extract(const double2 *__restrict__ z, uint32_t G, const uint2 *__restrict__ idx, uint32_t nEntries, const float *__restrict__ stepSizes,
                          float *__restrict__ raw) {
    uint32_t e = blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= nEntries) return;
    z += (size_t)blockIdx.y * G * G;
    const float stepSize = stepSizes[blockIdx.y];
    float *d2 = raw + (size_t)blockIdx.y * 3 * nEntries, *R = d2 + nEntries, *T = R + nEntries;
    uint32_t c = G / 2 - 1;       // center = length - 1
    uint2 ij = idx[e];
    float denorm = 1.f / (stepSize * stepSize);
    double2 v = z[(size_t)(c + ij.x) * G + (c + ij.y)];
    d2[e] = (float)(ij.x * ij.x + ij.y * ij.y) * stepSize * stepSize;
    R[e] = (float)v.x * denorm;
    T[e] = (float)v.y * denorm;
}

// MPC_ResampleForUniformDistanceSquaredDistribution (:404-426) with resample (:355-402)
__global__ void k_resample(const float *__restrict__ raw, uint32_t len, uint32_t target, float *__restrict__ res) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= target) return;
    const float *d2 = raw + (size_t)blockIdx.y * 3 * len, *R = d2 + len, *T = R + len;      // blockIdx.y = set
    float *od2 = res + (size_t)blockIdx.y * 3 * target, *oR = od2 + target, *oT = oR + target;
    float extent = d2[len - 1];
    float dsq = (float)i * extent / (float)(target - 1);
    od2[i] = dsq;
    float r, t;
    if (dsq > extent) { r = t = 0.f; }
    else {
        uint32_t lo = 0, hi = len - 1;          // first index with d2[lo] >= dsq, capped at len-1
        while (lo < hi) { uint32_t mid = (lo + hi) / 2; if (dsq > d2[mid]) lo = mid + 1; else hi = mid; }
        if (lo) {
            float la = (dsq - d2[lo - 1]) / (d2[lo] - d2[lo - 1]);
            la = la <= 0.f ? 0.f : (la >= 1.f ? 1.f : la);   // Clamp (template above :348-353)
            if (la != la) la = 0.5f;
            r = (1.f - la) * R[lo - 1] + la * R[lo];
            t = (1.f - la) * T[lo - 1] + la * T[lo];
        } else { r = R[0]; t = T[0]; }
    }
    oR[i] = r; oT[i] = t;
}

// ---------------------------------------------------------------------------------------------
static uint32_t roundUpPow2(uint32_t v) { v--; v |= v >> 1; v |= v >> 2; v |= v >> 4; v |= v >> 8; v |= v >> 16; return v + 1; }

// the reference's std::set<OutEntry> keyed by distanceSquared: unique i*i+j*j <= extent^2 in
// ascending order, first inserted (smallest i) wins (:316-330)
static const std::vector<uint2> &uniqueRadii(uint32_t length) {
    static std::map<uint32_t, std::vector<uint2>> cache;      // the list depends on the length only
    auto it = cache.find(length);
    if (it != cache.end()) return it->second;
    uint32_t extent = length - 1;
    std::map<uint32_t, uint2> m;
    for (uint32_t i = 0; i <= extent; i++)
        for (uint32_t j = i; i * i + j * j <= extent * extent; j++) m.insert({i * i + j * j, make_uint2(i, j)});
    std::vector<uint2> v;
    v.reserve(m.size());
    for (auto &kv : m) v.push_back(kv.second);
    return cache[length] = std::move(v);
}

// Work areas of one GROUP of K sets: every array holds K consecutive per-set blocks and every kernel takes the set from
// its grid (blockIdx.z, or .y where noted), so a group's profiles go through each stage in one launch.
struct ProfileWork {
    double2 *grid[2] = {nullptr, nullptr}, *Z[3] = {nullptr, nullptr, nullptr}, *tw = nullptr, *partial = nullptr;
    uint2 *idx = nullptr;
    float *raw = nullptr, *res = nullptr, *steps = nullptr;
    float *out[2] = {nullptr, nullptr}, *hostOut[2] = {nullptr, nullptr};      // a group's [set][d2 | R | T][len] rows: device, and pinned host
    double2 *hostPart[2] = {nullptr, nullptr};
    cudaEvent_t computed[2] = {nullptr, nullptr}, copied[2] = {nullptr, nullptr};
    DipoleSet *ds = nullptr;
    uint32_t *tNonZero = nullptr;      // [SKC_PROFILE_MAX_LAYERS][K]
    uint32_t G = 0, N = 0, nEntries = 0, K = 1;
    void release() {
        void *p[] = {grid[0], grid[1], Z[0], Z[1], Z[2], tw, partial, idx, raw, res, steps, ds, tNonZero, out[0], out[1]};
        for (void *q : p) if (q) cudaFree(q);
        for (int b = 0; b < 2; ++b) {
            if (hostOut[b]) cudaFreeHost(hostOut[b]);
            if (hostPart[b]) cudaFreeHost(hostPart[b]);
            if (computed[b]) cudaEventDestroy(computed[b]);
            if (copied[b]) cudaEventDestroy(copied[b]);
        }
        *this = ProfileWork();
    }
};

static size_t fftSmemBytes(uint32_t N) { return (size_t)(N + (N >> 3) + (N >> 6) + 2) * sizeof(double2); }

static int fft2d(skc_ctx *ctx, ProfileWork &w, double2 *A, double2 *B, uint32_t logN, int inverse, uint32_t k) {
    // rows of A in place, transpose into B, rows of B in place: B holds the 2-D transform (transposed layout)
    // The G x G grid sits zero-padded in the N x N matrix, on the rows and columns (i - c) mod N, i < G, c = (G-1)/2 (k_pack):
    // a forward transform's first pass meets N - G rows of zeros (whose transform is zeros, already in place), and of the
    // inverse's last pass only that band of rows is ever read (k_unpack) -- both passes run on the G-row band alone.
    uint32_t N = w.N, G = w.G, threads = std::max(32u, std::min(N / 8, 512u)), band = N - (G - 1) / 2;
    const bool oldRows = getenv("SKC_FFT_ROWS") && !strcmp(getenv("SKC_FFT_ROWS"), "generic");
    auto rows = [&](double2 *M, uint32_t nRows, uint32_t rowStart) {
        if (N == FFT2_N && !oldRows) {
            const size_t smem = 2 * FFT2_SLOTS * sizeof(double2);
            static bool attr = false;
            if (!attr) { cudaFuncSetAttribute(k_fft_rows2048, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem); attr = true; }
            k_fft_rows2048<<<dim3((nRows + 1) / 2, k), 256, smem, ctx->stream>>>(M, w.tw, inverse, rowStart, nRows);
            SKC_LAUNCHED("k_fft_rows2048");
        } else {
            k_fft_rows<<<dim3(nRows, k), threads, fftSmemBytes(N), ctx->stream>>>(M, N, logN, w.tw, inverse, rowStart);
            SKC_LAUNCHED("k_fft_rows");
        }
    };
    rows(A, inverse ? N : G, inverse ? 0u : band);
    k_transpose<<<dim3(N / 32, N / 32, k), dim3(32, 8), 0, ctx->stream>>>(A, N, B, inverse ? 2 : 1, band, G);
    SKC_LAUNCHED("k_transpose");
    rows(B, inverse ? G : N, inverse ? band : 0u);
    return 0;
}

extern "C" int skc_profile_batch(skc_ctx *ctx, const skc_layer_spec *specs, const uint32_t *layersPerSet,
                                 const skc_mpc_options *opts, uint32_t nSets, int resample,
                                 uint32_t *lenOut, float *d2, float *R, float *T, float *totals) {
    if (!ctx) return 1;
    cudaSetDevice(ctx->device);
    if (!nSets) { if (lenOut) *lenOut = 0; return 0; }
    uint32_t length = roundUpPow2(opts[0].desiredLength);
    for (uint32_t s = 0; s < nSets; ++s) {
        if (roundUpPow2(opts[s].desiredLength) != length) return skc_fail(ctx, "skc_profile_batch: all sets of a batch must share desiredLength");
        if (!layersPerSet[s]) return skc_fail(ctx, "skc_profile_batch: set %u has no layers", s);
    }
    if (length < 2 || length > 1024) return skc_fail(ctx, "skc_profile_batch: desiredLength %u out of range (2..1024)", length);
    const std::vector<uint2> &radii = uniqueRadii(length);
    uint32_t nEntries = (uint32_t)radii.size(), outLen = resample ? 2 * nEntries : nEntries;
    if (lenOut) *lenOut = outLen;
    if (!d2 && !R && !T && !totals) return 0;

    ProfileWork w;
    w.G = 2 * length; w.N = 2 * w.G; w.nEntries = nEntries;
    uint32_t G = w.G, N = w.N, logN = 0;
    while ((1u << logN) < N) ++logN;
    if (fftSmemBytes(N) > 48 * 1024)
        SKC_CUDA(cudaFuncSetAttribute(k_fft_rows, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)fftSmemBytes(N)));
    // group size: as many sets per launch as fit a quarter of the free memory, at most 32
    bool multi = false;
    for (uint32_t s = 0; s < nSets; ++s) multi |= layersPerSet[s] > 1;
    size_t perSet = 2 * (size_t)G * G * sizeof(double2) + (multi ? 3 * (size_t)N * N * sizeof(double2) : 0) + 3 * (size_t)(nEntries + outLen) * sizeof(float);
    size_t freeB = 0, totalB = 0;
    cudaMemGetInfo(&freeB, &totalB);
    uint32_t K = (uint32_t)std::max<size_t>(1, std::min<size_t>(std::min<size_t>(nSets, 32), freeB / 4 / perSet));
    if (const char *e = getenv("SKC_PROFILE_GROUP")) K = (uint32_t)std::max(1, std::min(atoi(e), (int)nSets));
    w.K = K;
    // more than one group: the groups' outputs are staged through pinned memory and delivered while the next group computes;
    // a single group (a material's 30 wavelengths) copies straight into the caller's arrays
    const bool pipelined = nSets > K;
    int rc = 1;
    do {
        if (cudaMalloc((void **)&w.grid[0], (size_t)K * G * G * sizeof(double2)) != cudaSuccess) break;
        if (cudaMalloc((void **)&w.grid[1], (size_t)K * G * G * sizeof(double2)) != cudaSuccess) break;
        if (multi) {
            bool ok = true;
            for (int k = 0; k < 3; ++k) ok &= cudaMalloc((void **)&w.Z[k], (size_t)K * N * N * sizeof(double2)) == cudaSuccess;
            if (!ok) break;
            std::vector<double2> tw(N / 2);
            tw.reserve(N / 2 + N);
            for (uint32_t k = 0; k < N / 2; ++k) { long double a = -2.0L * 3.14159265358979323846264338327950288L * k / N; tw[k] = make_double2((double)cosl(a), (double)sinl(a)); }
            // the same entries again, stage by stage (stage st reads tw[i * (N >> st)], i < 2^(st-1)): contiguous per stage, so that
            // k_fft_rows2048's twiddle reads are coalesced; stage st starts at N/2 + 2^(st-1) - 1
            for (uint32_t st = 1; (1u << st) <= N; ++st)
                for (uint32_t i = 0; i < (1u << (st - 1)); ++i) tw.push_back(tw[(size_t)i * (N >> st)]);
            if (cudaMalloc((void **)&w.tw, tw.size() * sizeof(double2)) != cudaSuccess) break;
            if (cudaMemcpyAsync(w.tw, tw.data(), tw.size() * sizeof(double2), cudaMemcpyHostToDevice, ctx->stream) != cudaSuccess) break;
            if (cudaStreamSynchronize(ctx->stream) != cudaSuccess) break;
        }
        if (cudaMalloc((void **)&w.partial, 2 * (size_t)K * 256 * sizeof(double2)) != cudaSuccess) break;
        {
            bool ok = true;
            for (int b = 0; b < (pipelined ? 2 : 1); ++b) {
                ok &= cudaMalloc((void **)&w.out[b], (size_t)K * 3 * outLen * sizeof(float)) == cudaSuccess;
                if (pipelined) ok &= cudaMallocHost((void **)&w.hostOut[b], (size_t)K * 3 * outLen * sizeof(float)) == cudaSuccess;
                if (pipelined) ok &= cudaMallocHost((void **)&w.hostPart[b], (size_t)K * 256 * sizeof(double2)) == cudaSuccess;
                ok &= cudaEventCreateWithFlags(&w.computed[b], cudaEventDisableTiming) == cudaSuccess;
                ok &= cudaEventCreateWithFlags(&w.copied[b], cudaEventDisableTiming) == cudaSuccess;
            }
            if (!ok) break;
        }
        if (cudaMalloc((void **)&w.idx, nEntries * sizeof(uint2)) != cudaSuccess) break;
        if (cudaMalloc((void **)&w.raw, (size_t)K * 3 * nEntries * sizeof(float)) != cudaSuccess) break;
        if (cudaMalloc((void **)&w.steps, (size_t)K * sizeof(float)) != cudaSuccess) break;
        if (cudaMalloc((void **)&w.ds, (size_t)K * sizeof(DipoleSet)) != cudaSuccess) break;
        if (cudaMalloc((void **)&w.tNonZero, (size_t)SKC_PROFILE_MAX_LAYERS * K * sizeof(uint32_t)) != cudaSuccess) break;
        if (cudaMemcpyAsync(w.idx, radii.data(), nEntries * sizeof(uint2), cudaMemcpyHostToDevice, ctx->stream) != cudaSuccess) break;
        if (cudaStreamSynchronize(ctx->stream) != cudaSuccess) break;
        rc = 0;
    } while (0);
    if (rc) { w.release(); return skc_fail(ctx, "skc_profile_batch: device allocation failed (%s)", cudaGetErrorString(cudaGetLastError())); }

    uploadExpTable(ctx->stream);
    skc_stage_begin(ctx);
    std::vector<size_t> specOffset(nSets + 1, 0);
    for (uint32_t s = 0; s < nSets; ++s) specOffset[s + 1] = specOffset[s] + layersPerSet[s];
    std::vector<double2> hpart((size_t)K * 256);
    std::vector<DipoleSet> hds(K);
    std::vector<float> hsteps(K);
    uint32_t ext = (G - 1) / 2 + 1;
    dim3 lpBlock(32, 8);
    // a group = up to K consecutive sets with the same number of layers: layer l of all of them goes through each stage at once
    uint32_t groupIndex = 0;
    struct Joining { std::thread t; ~Joining() { if (t.joinable()) t.join(); } } bg;      // joins on every way out of the function
    std::thread &worker = bg.t;
    int workerStatus = 0;
    const int device = ctx->device;
    for (uint32_t s0 = 0; s0 < nSets && !rc;) {
        const uint32_t nL = layersPerSet[s0];
        uint32_t k = 1;
        while (k < K && s0 + k < nSets && layersPerSet[s0 + k] == nL) ++k;
        auto layer = [&](uint32_t l, double2 *dst) -> int {
            for (uint32_t q = 0; q < k; ++q) {
                const skc_layer_spec *sp = specs + specOffset[s0 + q];
                float iorUpper = l == 0 ? sp[0].ior : sp[l].ior / sp[l - 1].ior;                     // :303-309
                float iorLower = nL > l + 1 ? sp[l].ior / sp[l + 1].ior : sp[l].ior;
                hds[q] = makeDipoles(sp[l], iorUpper, iorLower, opts[s0 + q].desiredStepSize, opts[s0 + q].lerpOnThinSlab != 0);
            }
            SKC_CUDA(cudaMemcpyAsync(w.ds, hds.data(), k * sizeof(DipoleSet), cudaMemcpyHostToDevice, ctx->stream));      // pageable source: staged before the call returns
            SKC_CUDA(cudaMemsetAsync(dst, 0, (size_t)k * G * G * sizeof(double2), ctx->stream));   // profile.Clear()
            k_layer_profile<<<dim3((ext + 31) / 32, (ext + 7) / 8, k), lpBlock, 0, ctx->stream>>>(w.ds, G, dst, w.tNonZero + (size_t)l * K);
            SKC_LAUNCHED("k_layer_profile");
            return 0;
        };
        if (nL > SKC_PROFILE_MAX_LAYERS) { rc = skc_fail(ctx, "skc_profile_batch: %u layers in one set (at most %d)", nL, SKC_PROFILE_MAX_LAYERS); break; }
        SKC_CUDA(cudaMemsetAsync(w.tNonZero, 0, (size_t)SKC_PROFILE_MAX_LAYERS * K * sizeof(uint32_t), ctx->stream));
        rc = layer(0, w.grid[0]);
        for (uint32_t l = 1; l < nL && !rc; ++l) {
            if ((rc = layer(l, w.grid[1]))) break;
            dim3 pg((N + 255) / 256, G, k);
            k_pack<<<pg, 256, 0, ctx->stream>>>(w.grid[0], G, N, w.Z[0]);
            k_pack<<<pg, 256, 0, ctx->stream>>>(w.grid[1], G, N, w.Z[1]);
            ctx->launches += 2;
            // forward: Z0 -> Z2 (transposed spectrum of layer 1); Z1 -> Z0 (layer 2)
            if ((rc = fft2d(ctx, w, w.Z[0], w.Z[2], logN, 0, k))) break;
            if ((rc = fft2d(ctx, w, w.Z[1], w.Z[0], logN, 0, k))) break;
            k_adding<<<dim3((N + 255) / 256, N / 2 + 1, k), 256, 0, ctx->stream>>>(w.Z[2], w.Z[0], N, w.Z[1]);
            ctx->launches++;
            if ((rc = fft2d(ctx, w, w.Z[1], w.Z[2], logN, 1, k))) break;      // back to the natural layout
            k_unpack<<<dim3((G + 255) / 256, G, k), 256, 0, ctx->stream>>>(w.Z[2], N, G, w.grid[0], w.tNonZero, l + 1, K);
            ctx->launches++;
        }
        if (rc) break;
        for (uint32_t q = 0; q < k; ++q) hsteps[q] = opts[s0 + q].desiredStepSize;
        if (cudaMemcpyAsync(w.steps, hsteps.data(), k * sizeof(float), cudaMemcpyHostToDevice, ctx->stream) != cudaSuccess) { rc = 1; break; }
        k_extract<<<dim3((nEntries + 255) / 256, k), 256, 0, ctx->stream>>>(w.grid[0], G, w.idx, nEntries, w.steps, w.raw);
        ctx->launches++;
        // Output of the group: double-buffered on the device (out[gi & 1]) and in pinned host memory; the copy runs on the copy
        // stream and the scatter into the caller's (pageable) arrays on the host while the next group is being computed.
        const int bsel = pipelined ? (int)(groupIndex & 1) : 0;
        float *o = w.out[bsel];
        double2 *part = w.partial + (size_t)bsel * K * 256;
        if (pipelined && groupIndex >= 2) SKC_CUDA(cudaStreamWaitEvent(ctx->stream, w.copied[bsel], 0));      // the copy that last read this buffer
        if (resample) {
            k_resample<<<dim3((outLen + 255) / 256, k), 256, 0, ctx->stream>>>(w.raw, nEntries, outLen, o);
            ctx->launches++;
        } else SKC_CUDA(cudaMemcpyAsync(o, w.raw, (size_t)k * 3 * nEntries * sizeof(float), cudaMemcpyDeviceToDevice, ctx->stream));
        if (totals) {
            k_grid_sum<<<dim3(256, k), 256, 0, ctx->stream>>>(w.grid[0], (size_t)G * G, part);
            ctx->launches++;
        }
        if (!pipelined) {
            if (totals && cudaMemcpyAsync(hpart.data(), part, (size_t)k * 256 * sizeof(double2), cudaMemcpyDeviceToHost, ctx->stream) != cudaSuccess) { rc = 1; break; }
            const size_t rowB = (size_t)outLen * sizeof(float);
            float *dsts[3] = {d2, R, T};
            for (int a = 0; a < 3 && !rc; ++a)
                if (dsts[a] && cudaMemcpy2DAsync(dsts[a] + (size_t)s0 * outLen, rowB, o + (size_t)a * outLen, 3 * rowB, rowB, k, cudaMemcpyDeviceToHost, ctx->stream) != cudaSuccess) rc = 1;
            if (rc) break;
            if (cudaStreamSynchronize(ctx->stream) != cudaSuccess) { rc = 1; break; }
            if (totals)
                for (uint32_t q = 0; q < k; ++q) {
                    double r = 0., t = 0.;
                    for (int j = 0; j < 256; ++j) { r += hpart[(size_t)q * 256 + j].x; t += hpart[(size_t)q * 256 + j].y; }
                    totals[2 * (s0 + q)] = (float)r; totals[2 * (s0 + q) + 1] = (float)t;
                }
            s0 += k;
            ++groupIndex;
            continue;
        }
        SKC_CUDA(cudaEventRecord(w.computed[bsel], ctx->stream));
        SKC_CUDA(cudaStreamWaitEvent(ctx->copyStream, w.computed[bsel], 0));
        SKC_CUDA(cudaMemcpyAsync(w.hostOut[bsel], o, (size_t)k * 3 * outLen * sizeof(float), cudaMemcpyDeviceToHost, ctx->copyStream));
        if (totals) SKC_CUDA(cudaMemcpyAsync(w.hostPart[bsel], part, (size_t)k * 256 * sizeof(double2), cudaMemcpyDeviceToHost, ctx->copyStream));
        SKC_CUDA(cudaEventRecord(w.copied[bsel], ctx->copyStream));
        // rows [set][d2 | R | T][outLen] of a finished group -> the caller's three [set][outLen] arrays: on a worker thread (the
        // scatter itself on up to four), so that the launching thread goes on to the next group -- a 46 MB pageable scatter
        // per group on the launching thread kept the GPU waiting for its next launches
        auto deliver = [&w, d2, R, T, totals, outLen, device](int bb, uint32_t gs0, uint32_t gk, int *status) {
            if (cudaSetDevice(device) != cudaSuccess || cudaEventSynchronize(w.copied[bb]) != cudaSuccess) { *status = 1; return; }
            const float *h = w.hostOut[bb];
            float *dsts[3] = {d2, R, T};
            auto part = [&](uint32_t t, uint32_t nT) {
                for (uint32_t q = t; q < gk; q += nT)
                    for (int a = 0; a < 3; ++a)
                        if (dsts[a]) memcpy(dsts[a] + (size_t)(gs0 + q) * outLen, h + ((size_t)q * 3 + a) * outLen, (size_t)outLen * sizeof(float));
            };
            const uint32_t nT = std::min<uint32_t>(4u, gk);
            std::vector<std::thread> helpers;
            for (uint32_t t = 1; t < nT; ++t) helpers.emplace_back(part, t, nT);
            part(0, nT);
            for (auto &th : helpers) th.join();
            if (totals)
                for (uint32_t q = 0; q < gk; ++q) {
                    double r = 0., t = 0.;
                    for (int j = 0; j < 256; ++j) { r += w.hostPart[bb][(size_t)q * 256 + j].x; t += w.hostPart[bb][(size_t)q * 256 + j].y; }
                    totals[2 * (gs0 + q)] = (float)r; totals[2 * (gs0 + q) + 1] = (float)t;
                }
        };
        // the worker of the group before the previous one read hostOut[bsel ^ 1] ... it must be done before that buffer's next copy
        // is enqueued, which is the copy of the NEXT group: joining here, one group of GPU work after it started, rarely waits
        if (worker.joinable()) { worker.join(); if (workerStatus) { rc = 1; break; } }
        worker = std::thread(deliver, bsel, s0, k, &workerStatus);
        s0 += k;
        ++groupIndex;
    }
    if (worker.joinable()) { worker.join(); if (workerStatus) rc = 1; }
    if (!rc) skc_stage_end(ctx, 6);
    cudaError_t e = cudaGetLastError();
    cudaStreamSynchronize(ctx->stream);
    w.release();
    if (rc || e != cudaSuccess) return skc_fail(ctx, "skc_profile_batch: %s", ctx->err.empty() ? cudaGetErrorString(e) : ctx->err.c_str());
    return 0;
}

// ---------------------------------------------------------------------------------------------
// resample (MultipoleProfileCalculator.cpp:355-402) at arbitrary query points: what both MPC_ResampleForUniformDistanceSquared-
// Distribution (:404-426, any target length) and MPC_ResampleDistribution (:428-448) reduce to.  One thread per query: first
// entry with d2 >= query (the reference's interpolation + linear search finds the same entry; d2 is strictly increasing), fp32
// lerp with the reference's clamp and NaN rule, 0 beyond the table.
__global__ void k_resample_at(const float *__restrict__ d2, const float *__restrict__ R, const float *__restrict__ T, uint32_t len,
                              const float *__restrict__ q, uint32_t nq, float *__restrict__ oR, float *__restrict__ oT) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nq) return;
    const float dsq = q[i], extent = d2[len - 1];
    float r, t;
    if (dsq > extent) { r = t = 0.f; }
    else {
        uint32_t lo = 0, hi = len - 1;
        while (lo < hi) { uint32_t mid = (lo + hi) / 2; if (dsq > d2[mid]) lo = mid + 1; else hi = mid; }
        if (lo) {
            float la = (dsq - d2[lo - 1]) / (d2[lo] - d2[lo - 1]);
            la = la < 0.f ? 0.f : (la > 1.f ? 1.f : la);
            if (la != la) la = 0.5f;
            r = (1.f - la) * R[lo - 1] + la * R[lo];
            t = (1.f - la) * T[lo - 1] + la * T[lo];
        } else { r = R[0]; t = T[0]; }
    }
    oR[i] = r; oT[i] = t;
}

extern "C" int skc_profile_resample(skc_ctx *ctx, const float *d2, const float *R, const float *T, uint32_t len,
                                    const float *queryDsq, uint32_t nQueries, float *Rout, float *Tout) {
    if (!ctx) return 1;
    cudaSetDevice(ctx->device);
    if (!nQueries) return 0;
    if (!len || !d2 || !R || !T || !queryDsq || !Rout || !Tout) return skc_fail(ctx, "skc_profile_resample: empty table or null argument");
    float *buf = (float *)skc_scratch(ctx, 0, ((size_t)3 * len + (size_t)3 * nQueries) * sizeof(float));
    if (!buf) return 1;
    float *dD = buf, *dR = dD + len, *dT = dR + len, *dQ = dT + len, *dOR = dQ + nQueries, *dOT = dOR + nQueries;
    SKC_CUDA(cudaMemcpyAsync(dD, d2, len * sizeof(float), cudaMemcpyHostToDevice, ctx->stream));
    SKC_CUDA(cudaMemcpyAsync(dR, R, len * sizeof(float), cudaMemcpyHostToDevice, ctx->stream));
    SKC_CUDA(cudaMemcpyAsync(dT, T, len * sizeof(float), cudaMemcpyHostToDevice, ctx->stream));
    SKC_CUDA(cudaMemcpyAsync(dQ, queryDsq, nQueries * sizeof(float), cudaMemcpyHostToDevice, ctx->stream));
    k_resample_at<<<(nQueries + 255) / 256, 256, 0, ctx->stream>>>(dD, dR, dT, len, dQ, nQueries, dOR, dOT);
    SKC_LAUNCHED("k_resample_at");
    SKC_CUDA(cudaMemcpyAsync(Rout, dOR, nQueries * sizeof(float), cudaMemcpyDeviceToHost, ctx->stream));
    SKC_CUDA(cudaMemcpyAsync(Tout, dOT, nQueries * sizeof(float), cudaMemcpyDeviceToHost, ctx->stream));
    SKC_CUDA(cudaStreamSynchronize(ctx->stream));
    return 0;
}
