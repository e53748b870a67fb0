// Shared device helpers of the tcgen05 kernels (edge forward / backward, node backward): activation epilogues, bf16 hi/lo
// tile stores, operand descriptors for the K-major and MN-major views of a SWIZZLE_NONE tile, MMA issue loops.
#pragma once
#include "tc_common.cuh"

namespace {

// SiLU / sigmoid of the tensor-core kernels.  FAST: ONE MUFU op per element through tanh.approx.f32
//   z * sigmoid(z) = 0.5 z (1 + tanh(z / 2))
// instead of ex2 + rcp.  The epilogues between dependent MMAs are MUFU-bound (two activations per element and branch),
// so this is worth 14 % of the forward edge kernel.  Measured on B200: per-layer outputs and all gradients of the
// bf16x3 mode are unchanged at the 1e-6 level (2.2e-5 / 3.7e-5 per layer, <= 8e-5 on gradients, same as with ex2 + rcp),
// i.e. the instruction is far more accurate than its documented 2^-11 bound over the range the activations visit.
template <bool FAST>
__device__ __forceinline__ float tc_silu(float z) {
    if (FAST) {
        float t;
        const float h = 0.5f * z;
        asm("tanh.approx.f32 %0, %1;" : "=f"(t) : "f"(h));
        return fmaf(h, t, h);
    }
    return silu_f(z);
}
// SiLU(z) and SiLU'(z) = s + SiLU(z) (1 - s),  s = sigmoid(z) = 0.5 + 0.5 tanh(z / 2): one MUFU, five FMA-pipe ops
__device__ __forceinline__ void tc_silu_and_grad(float z, float& act, float& dact) {
    float t;
    const float h = 0.5f * z;
    asm("tanh.approx.f32 %0, %1;" : "=f"(t) : "f"(h));
    act = fmaf(h, t, h);
    dact = fmaf(act, fmaf(-0.5f, t, 0.5f), fmaf(0.5f, t, 0.5f));
}

// The edge kernels keep their weight images halved, so the accumulators ARE h = z / 2:
__device__ __forceinline__ float tc_silu_h(float h) {
    float t;
    asm("tanh.approx.f32 %0, %1;" : "=f"(t) : "f"(h));
    return fmaf(h, t, h);
}
// SiLU(z) and 2 SiLU'(z) = (1 + t) + SiLU(z) (1 - t) from h = z / 2
__device__ __forceinline__ void tc_silu_and_grad2_h(float h, float& act, float& dact2) {
    float t;
    asm("tanh.approx.f32 %0, %1;" : "=f"(t) : "f"(h));
    act = fmaf(h, t, h);
    dact2 = fmaf(act, 1.f - t, 1.f + t);
}

// this thread's row, columns [c0, c0+32) of a TMEM block holding z / 2 (bias included): SiLU packed as bf16 hi/lo pairs
__device__ __forceinline__ void tc_epi32(uint32_t taddr, uint32_t (&hi)[16], uint32_t (&lo)[16]) {
    float v[32];
    tc::tmem_ld32(taddr, v);
#pragma unroll
    for (int i = 0; i < 16; ++i) tc::split2(tc_silu_h(v[2 * i]), tc_silu_h(v[2 * i + 1]), hi[i], lo[i]);
}
// this thread's row, columns [c0, c0+32) of a TMEM block (z / 2, bias included) -> SiLU -> bf16 hi/lo chunks of a 128-row tile.
// Walks the 32 columns as four 8-column chunks in a REAL loop: each chunk is a burst of MUFU work (two per element)
// followed by ALU work (split, pack, store); keeping the chunks apart stops the scheduler from hoisting all 64 MUFU
// ops to the front, so that across the warps of an SM sub-partition the XU and ALU pipes run concurrently.
// The TMEM load of chunk j + 1 is issued before chunk j is processed (two register sets, explicit wait): with four warps per
// scheduler the TMEM latency of a load-wait-compute chain is otherwise exposed once per chunk.
__device__ __forceinline__ void tc_epi_store_chunked(uint32_t taddr, uint8_t* t_hi, uint8_t* t_lo, int r, int c0, bool with_lo) {
    uint32_t cur[8], nxt[8];
    tc::tmem_ld8_nowait(taddr, cur);
    tc::tmem_wait_ld8(cur);
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        if (j < 3) tc::tmem_ld8_nowait(taddr + 8 * (j + 1), nxt);
        uint32_t hi[4], lo[4];
#pragma unroll
        for (int i = 0; i < 4; ++i)
            tc::split2(tc_silu_h(__uint_as_float(cur[2 * i])), tc_silu_h(__uint_as_float(cur[2 * i + 1])), hi[i], lo[i]);
        const uint32_t off = tc::chunk_off(128, r, (c0 >> 3) + j);
        *reinterpret_cast<uint4*>(t_hi + off) = make_uint4(hi[0], hi[1], hi[2], hi[3]);
        if (with_lo) *reinterpret_cast<uint4*>(t_lo + off) = make_uint4(lo[0], lo[1], lo[2], lo[3]);
        if (j < 3) {
            tc::tmem_wait_ld8(nxt);
#pragma unroll
            for (int i = 0; i < 8; ++i) cur[i] = nxt[i];
        }
    }
}

// backward layer-1 epilogue of columns [c0, c0+32): a1 = SiLU(z1) -> bf16 hi/lo chunks of a 128-row tile, 2 SiLU'(z1)
// back into the same TMEM columns; chunked like tc_epi_store_chunked (8 columns per trip of a REAL loop: light on
// registers, MUFU and ALU work of different warps interleave)
__device__ __forceinline__ void tc_epi_bwd_chunked(uint32_t taddr, uint8_t* t_hi, uint8_t* t_lo, int r, int c0, bool with_lo) {
#pragma unroll 1
    for (int j = 0; j < 4; ++j) {
        float v[8], act[8];
        tc::tmem_ld8(taddr + 8 * j, v);
#pragma unroll
        for (int i = 0; i < 8; ++i) tc_silu_and_grad2_h(v[i], act[i], v[i]);
        uint32_t hi[4], lo[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) tc::split2(act[2 * i], act[2 * i + 1], hi[i], lo[i]);
        const uint32_t off = tc::chunk_off(128, r, (c0 >> 3) + j);
        *reinterpret_cast<uint4*>(t_hi + off) = make_uint4(hi[0], hi[1], hi[2], hi[3]);
        if (with_lo) *reinterpret_cast<uint4*>(t_lo + off) = make_uint4(lo[0], lo[1], lo[2], lo[3]);
        tc::tmem_st8(taddr + 8 * j, v);
    }
    tc::tmem_wait_st();
}
// gz1 = ga1 * (2 SiLU'(z1)) of columns [c0, c0+32) -> bf16 hi/lo chunks, chunked the same way
__device__ __forceinline__ void tc_gz1_chunked(uint32_t t_ga, uint32_t t_d1, uint8_t* t_hi, uint8_t* t_lo, int r, int c0, bool with_lo) {
#pragma unroll 1
    for (int j = 0; j < 4; ++j) {
        float ga[8], d1[8];
        tc::tmem_ld8(t_ga + 8 * j, ga);
        tc::tmem_ld8(t_d1 + 8 * j, d1);
        uint32_t hi[4], lo[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) tc::split2(ga[2 * i] * d1[2 * i], ga[2 * i + 1] * d1[2 * i + 1], hi[i], lo[i]);
        const uint32_t off = tc::chunk_off(128, r, (c0 >> 3) + j);
        *reinterpret_cast<uint4*>(t_hi + off) = make_uint4(hi[0], hi[1], hi[2], hi[3]);
        if (with_lo) *reinterpret_cast<uint4*>(t_lo + off) = make_uint4(lo[0], lo[1], lo[2], lo[3]);
    }
}

__device__ __forceinline__ void tc_store32(uint8_t* a_hi, uint8_t* a_lo, int r, int c0, const uint32_t (&hi)[16],
                                           const uint32_t (&lo)[16], bool with_lo) {
#pragma unroll
    for (int q = 0; q < 4; ++q) {
        const uint32_t off = tc::chunk_off(128, r, (c0 >> 3) + q);
        *reinterpret_cast<uint4*>(a_hi + off) = make_uint4(hi[4 * q], hi[4 * q + 1], hi[4 * q + 2], hi[4 * q + 3]);
        if (with_lo) *reinterpret_cast<uint4*>(a_lo + off) = make_uint4(lo[4 * q], lo[4 * q + 1], lo[4 * q + 2], lo[4 * q + 3]);
    }
}

struct TcOperand {
    uint32_t hi, lo;        // shared-memory addresses of the hi / lo tiles
    uint32_t chunk_stride;  // bytes between consecutive 16-byte chunk columns (rows * 16)
    bool mn;                // viewed MN-major (K index = tile row)
};
__device__ __forceinline__ uint64_t tc_op_desc(const TcOperand& o, bool use_lo, int s) {
    const uint32_t base = use_lo ? o.lo : o.hi;
    return o.mn ? tc::make_desc(base + s * 256, 128, o.chunk_stride)
                : tc::make_desc(base + s * 2 * o.chunk_stride, o.chunk_stride, 128);
}
// one thread: D[128 x N] (+)= A . B^T over K, NPASS split products, into TMEM column dcol.
// The descriptor of k-step s is the descriptor of step 0 plus s * (step bytes >> 4) in the 14-bit address field (shared
// memory is < 256 KB, so the field never carries): two integer adds per MMA instead of a rebuilt descriptor.
// a_has_lo = false: the A tile was stored as bf16 hi only (its lo.hi product is skipped)
template <int NPASS>
__device__ __forceinline__ void tc_issue_op(uint32_t tmem, int dcol, const TcOperand& A, const TcOperand& B, int K, int N,
                                            bool acc, bool a_has_lo = true) {
    const uint32_t idesc = tc::make_idesc_bf16(128, N) | (A.mn ? 1u << 15 : 0u) | (B.mn ? 1u << 16 : 0u);
    const uint64_t astep = (A.mn ? 256u : 2u * A.chunk_stride) >> 4, bstep = (B.mn ? 256u : 2u * B.chunk_stride) >> 4;
#pragma unroll
    for (int p = 0; p < NPASS; ++p) {
        if (p == 1 && !a_has_lo) continue;
        uint64_t ad = tc_op_desc(A, p == 1, 0), bd = tc_op_desc(B, p == 2, 0);
#pragma unroll 8
        for (int s = 0; s < K / 16; ++s) {
            tc::mma_bf16(tmem + dcol, ad, bd, idesc, acc);
            ad += astep;
            bd += bstep;
            acc = true;
        }
    }
}

// warp reduce-scatter of 32 per-thread values over the 32 lanes (fixed order): v[0] of lane L ends as the sum over the
// warp's 32 rows of column L
__device__ __forceinline__ void warp_colsum32(float (&v)[32], int lane) {
#pragma unroll
    for (int st = 0; st < 5; ++st) {
        const int off = 16 >> st, live = 16 >> st;
        const bool hiL = lane & off;
#pragma unroll
        for (int k = 0; k < 16; ++k) {
            if (k < live) {
                const float send = hiL ? v[k] : v[k + live], keep = hiL ? v[k + live] : v[k];
                v[k] = keep + __shfl_xor_sync(0xffffffffu, send, off);
            }
        }
    }
}

// split + store 32 consecutive columns [c0, c0+32) of row r into a 128-row tile
__device__ __forceinline__ void tc_split_store32(uint8_t* t_hi, uint8_t* t_lo, int r, int c0, const float (&v)[32], bool with_lo) {
    uint32_t hi[16], lo[16];
#pragma unroll
    for (int i = 0; i < 16; ++i) tc::split2(v[2 * i], v[2 * i + 1], hi[i], lo[i]);
    tc_store32(t_hi, t_lo, r, c0, hi, lo, with_lo);
}

template <bool FAST>
__device__ __forceinline__ float tc_sigmoid(float z) {
    if (FAST) {
        float t;
        asm("tanh.approx.f32 %0, %1;" : "=f"(t) : "f"(0.5f * z));
        return fmaf(0.5f, t, 0.5f);
    }
    return sigmoid_f(z);
}

}  // namespace
