// Weight-gradient tile reductions of the fp32 (CUDA-core) backward path.  Every CTA owns a private row of the
// partial-gradient buffer, walks its tiles in a fixed order and adds into that row; a final kernel sums the rows in
// CTA order.  No atomics: reruns are bit-identical.  add = false: the CTA's first tile stores (the row needs no zero-fill).
#pragma once
#include "common.cuh"

// dst[m*ldd + c] += sum_{r<rows} GZ[r*ldg + m] * A[r*lda + c]      m < M, c < Kc
// 256 threads = 16 (m groups of 8) x 16 (column lanes owning float4 at 4*lc and 64 + 4*lc of a 128-column block).
// GZ / A are shared-memory tiles whose row strides are >= round4(width) + 4 (over-reads stay inside the row).
__device__ __forceinline__ void tile_wgrad(const float* __restrict__ GZ, int ldg, const float* __restrict__ A, int lda,
                                           int rows, int M, int Kc, float* __restrict__ dst, int ldd, bool add = true) {
    const int lc = threadIdx.x & 15, lm = threadIdx.x >> 4;
    if (M <= 8) {
        // a few output rows (mlp_h's second Linear, M = H): one thread per column instead of 16 busy threads
        for (int c = threadIdx.x; c < Kc; c += 256) {
            float acc[8];
#pragma unroll
            for (int i = 0; i < 8; ++i) acc[i] = 0.f;
            for (int r = 0; r < rows; ++r) {
                const float av = A[r * lda + c];
                const float4 g0 = *reinterpret_cast<const float4*>(GZ + r * ldg);
                const float4 g1 = *reinterpret_cast<const float4*>(GZ + r * ldg + 4);
                const float g[8] = {g0.x, g0.y, g0.z, g0.w, g1.x, g1.y, g1.z, g1.w};
#pragma unroll
                for (int i = 0; i < 8; ++i) acc[i] = fmaf(g[i], av, acc[i]);
            }
#pragma unroll
            for (int i = 0; i < 8; ++i)
                if (i < M) dst[(size_t)i * ldd + c] = add ? dst[(size_t)i * ldd + c] + acc[i] : acc[i];
        }
        return;
    }
    // a few trailing columns (Kc = d + H): one thread per output row m instead of a 128-column block with one lane busy
    const int Kmain = (Kc > 128 && Kc % 128 != 0 && Kc % 128 <= 8) ? Kc - Kc % 128 : Kc;
    if (Kmain < Kc) {
        const int nt = Kc - Kmain;
        for (int m = threadIdx.x; m < M; m += 256) {
            float acc[8];
#pragma unroll
            for (int j = 0; j < 8; ++j) acc[j] = 0.f;
            for (int r = 0; r < rows; ++r) {
                const float gv = GZ[r * ldg + m];
#pragma unroll
                for (int j = 0; j < 8; ++j)
                    if (j < nt) acc[j] = fmaf(gv, A[r * lda + Kmain + j], acc[j]);
            }
#pragma unroll
            for (int j = 0; j < 8; ++j)
                if (j < nt) dst[(size_t)m * ldd + Kmain + j] = add ? dst[(size_t)m * ldd + Kmain + j] + acc[j] : acc[j];
        }
    }
    for (int mb = 0; mb < M; mb += 128) {
        const int m0 = mb + lm * 8;
        if (m0 >= M) continue;
        for (int cb = 0; cb < Kmain; cb += 128) {
            const int c0 = cb + lc * 4, c1 = c0 + 64;
            const bool v0 = c0 < Kmain, v1 = c1 < Kmain;
            if (!v0) continue;
            float acc[8][8];
#pragma unroll
            for (int i = 0; i < 8; ++i)
#pragma unroll
                for (int j = 0; j < 8; ++j) acc[i][j] = 0.f;
            for (int r = 0; r < rows; ++r) {
                const float4 g0 = *reinterpret_cast<const float4*>(GZ + r * ldg + m0);
                const float4 g1 = *reinterpret_cast<const float4*>(GZ + r * ldg + m0 + 4);
                const float4 a0 = *reinterpret_cast<const float4*>(A + r * lda + c0);
                float4 a1 = make_float4(0.f, 0.f, 0.f, 0.f);
                if (v1) a1 = *reinterpret_cast<const float4*>(A + r * lda + c1);
                const float g[8] = {g0.x, g0.y, g0.z, g0.w, g1.x, g1.y, g1.z, g1.w};
                const float a[8] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w};
#pragma unroll
                for (int i = 0; i < 8; ++i)
#pragma unroll
                    for (int j = 0; j < 8; ++j) acc[i][j] = fmaf(g[i], a[j], acc[i][j]);
            }
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                const int m = m0 + i;
                if (m >= M) break;
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                    const int c = j < 4 ? c0 + j : c1 + j - 4;
                    if (c < Kmain) dst[(size_t)m * ldd + c] = add ? dst[(size_t)m * ldd + c] + acc[i][j] : acc[i][j];
                }
            }
        }
    }
}

// dst[c] += sum_{r<rows} coef[r] * buf[r*ld + c]   (coef == nullptr -> 1)
__device__ __forceinline__ void tile_colsum(const float* __restrict__ buf, int ld, int rows, int N,
                                            const float* __restrict__ coef, float* __restrict__ dst, bool add = true) {
    for (int c = threadIdx.x; c < N; c += blockDim.x) {
        float s = 0.f;
        for (int r = 0; r < rows; ++r) s = fmaf(coef ? coef[r] : 1.f, buf[r * ld + c], s);
        dst[c] = add ? dst[c] + s : s;
    }
}
