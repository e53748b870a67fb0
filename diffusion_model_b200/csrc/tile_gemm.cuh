// CUDA-core (FFMA) tile GEMM used by the EGNN_PREC_FP32 path: true fp32 products and accumulation, like the
// reference's nn.Linear on a device with TF32 off (EquivariantGraphNeuralNetwork.py:13-34).
//
//   Z[r][n] = sum_k A[r][k] * Bt[k][n]      r in [0, 4*RT),  n in [0, N),  k in [0, K)
//
// A lives in shared memory (row stride lda, lda % 4 == 0, 16-byte aligned rows, columns K..round4(K)-1 finite),
// Bt in global memory (row stride ldb) and is read through the read-only path: a warp reads 32 consecutive
// floats of one Bt row, every A read is a 128-bit broadcast.  256 threads: 4 row groups x 64 column lanes,
// each thread owns RT rows x 2 columns (n, n+64) of a 128-column pass.
#pragma once
#include "common.cuh"

// k-steps of 4 whose weight loads are issued together in the edge kernels: wide layers force small register tiles (d = 1024:
// 16-edge tiles, RT = 4), where 8 loads per 32 FMAs leave the L2 latency exposed (parameters.yaml network: 196 ms per step);
// small tiles have the registers to keep 32 loads in flight.  The accumulation order does not change: results are identical.
#define EGNN_EDGE_KU(RT) ((RT) <= 8 ? 4 : (RT) <= 16 ? 2 : 1)

template <int RT, int KU = 1, class Epi>
__device__ __forceinline__ void tile_gemm(const float* __restrict__ As, int lda, const float* __restrict__ Bt, int ldb,
                                          int K, int N, Epi epi) {
    const int tid = threadIdx.x;
    const int r0 = (tid >> 6) * RT;
    const int cg = tid & 63;
    for (int nb = 0; nb < N; nb += 128) {
        if (nb > 0 && N - nb <= 8) {
            // a few trailing columns (e.g. N = d + H): a full 128-column pass would run with 3 of 64 lanes active.
            // 64/RT threads per row, each one whole dot product; the Bt tail is a handful of sectors shared by all warps
            constexpr int TC = 64 / RT;
            const int r = tid / TC;
            const float* arow = As + r * lda;
            for (int j = tid % TC; j < N - nb; j += TC) {
                const float* b = Bt + nb + j;
                float acc = 0.f;
#pragma unroll 8
                for (int k = 0; k < K; ++k) acc = fmaf(arow[k], __ldg(b + (size_t)k * ldb), acc);
                epi(r, nb + j, acc);
            }
            break;
        }
        const int c0 = nb + cg, c1 = c0 + 64;
        const bool v0 = c0 < N, v1 = c1 < N;
        float acc0[RT], acc1[RT];
#pragma unroll
        for (int r = 0; r < RT; ++r) acc0[r] = acc1[r] = 0.f;
        // KU k-steps of 4 per batch (template argument; 4 in the node backward): all 8 KU loads of Bt (L2 latency) are in flight before the first FMA needs one
        for (int k = 0; k < K; k += 4 * KU) {
            float b0[KU][4], b1[KU][4];
#pragma unroll
            for (int u = 0; u < KU; ++u)
#pragma unroll
                for (int kk = 0; kk < 4; ++kk) {
                    const bool kv = k + 4 * u + kk < K;
                    b0[u][kk] = (v0 && kv) ? __ldg(Bt + (size_t)(k + 4 * u + kk) * ldb + c0) : 0.f;
                    b1[u][kk] = (v1 && kv) ? __ldg(Bt + (size_t)(k + 4 * u + kk) * ldb + c1) : 0.f;
                }
#pragma unroll
            for (int u = 0; u < KU; ++u) {
                if (KU > 1 && k + 4 * u >= K) break;  // (the A columns beyond round4(K) are not defined)
#pragma unroll
                for (int r = 0; r < RT; ++r) {
                    const float4 a = *reinterpret_cast<const float4*>(As + (r0 + r) * lda + k + 4 * u);
                    acc0[r] = fmaf(a.x, b0[u][0], acc0[r]);
                    acc0[r] = fmaf(a.y, b0[u][1], acc0[r]);
                    acc0[r] = fmaf(a.z, b0[u][2], acc0[r]);
                    acc0[r] = fmaf(a.w, b0[u][3], acc0[r]);
                    acc1[r] = fmaf(a.x, b1[u][0], acc1[r]);
                    acc1[r] = fmaf(a.y, b1[u][1], acc1[r]);
                    acc1[r] = fmaf(a.z, b1[u][2], acc1[r]);
                    acc1[r] = fmaf(a.w, b1[u][3], acc1[r]);
                }
            }
        }
#pragma unroll
        for (int r = 0; r < RT; ++r) {
            if (v0) epi(r0 + r, c0, acc0[r]);
            if (v1) epi(r0 + r, c1, acc1[r]);
        }
    }
}

// zero the alignment padding columns [N, round4(N)) of a [rows][ld] shared buffer
__device__ __forceinline__ void zero_pad_cols(float* buf, int ld, int rows, int N) {
    const int pad = round4(N) - N;
    if (pad == 0) return;
    for (int idx = threadIdx.x; idx < rows * pad; idx += blockDim.x) buf[(idx / pad) * ld + N + idx % pad] = 0.f;
}

// out[r] = bias + sum_c w[c] * buf[r][c], one warp per row, fixed lane order
template <class Store>
__device__ __forceinline__ void row_dots(const float* buf, int ld, int rows, int N, const float* __restrict__ w,
                                         float bias, Store store) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = blockDim.x >> 5;
    for (int r = warp; r < rows; r += nw) {
        float s = 0.f;
        for (int c = lane; c < N; c += 32) s = fmaf(__ldg(w + c), buf[r * ld + c], s);
        s = warp_sum(s);
        if (lane == 0) store(r, s + bias);
    }
}
