// tcgen05 (5th-generation tensor core) path of the edge MLPs: operands staged in shared memory as bf16 hi/lo splits,
// fp32 accumulators in TMEM, SiLU epilogues out of TMEM.  This file also holds the one-tile GEMM self-test that pins
// the shared-memory descriptor conventions on real hardware (tests/test_gpu_tc.py).
#include <cstdlib>

#include "tc_tile.cuh"

namespace {

// D[128 x N] = A[128 x K] . B[N x K]^T   (one CTA, 128 threads; K % 16 == 0, N % 16 == 0, N <= 256)
//   npass 1: bf16(A) bf16(B);  npass 3: hi.hi + lo.hi + hi.lo  (fp32-class accuracy)
//   flags bit0: exchange the LBO / SBO fields of K-major descriptors (convention probe)
//         bit1: A is staged MN-major (tile rows = k, 16-byte chunks = 8 consecutive m), bit2: same for B
//         bit3: exchange LBO / SBO of the MN-major descriptors
__global__ void __launch_bounds__(128) tc_selftest_kernel(const float* __restrict__ A, const float* __restrict__ B,
                                                          float* __restrict__ D, int K, int N, int npass, int flags,
                                                          int* __restrict__ status) {
    extern __shared__ __align__(1024) uint8_t smem[];
    __shared__ uint64_t bar;
    __shared__ uint32_t tmem_slot;
    const int tid = threadIdx.x, warp = tid >> 5;
    const bool swapk = flags & 1, amn = flags & 2, bmn = flags & 4, swapmn = flags & 8;
    uint8_t* a_hi = smem;
    uint8_t* a_lo = a_hi + 128 * K * 2;
    uint8_t* b_hi = a_lo + 128 * K * 2;
    uint8_t* b_lo = b_hi + N * K * 2;
    // element (row, col) of an [R x C] row-major matrix; K-major: tile rows = row, chunks over col;
    // MN-major: tile rows = col (the K index), chunks over row
    auto stage = [&](const float* M, int R, int C, bool mn, uint8_t* hi_t, uint8_t* lo_t) {
        const int rows = mn ? C : R, chunks = (mn ? R : C) / 8;
        for (int idx = tid; idx < rows * chunks; idx += 128) {
            const int r = idx % rows, j = idx / rows;
            uint32_t hi[4], lo[4];
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                const float v0 = mn ? M[(j * 8 + 2 * q) * C + r] : M[r * C + j * 8 + 2 * q];
                const float v1 = mn ? M[(j * 8 + 2 * q + 1) * C + r] : M[r * C + j * 8 + 2 * q + 1];
                tc::split2(v0, v1, hi[q], lo[q]);
            }
            *reinterpret_cast<uint4*>(hi_t + tc::chunk_off(rows, r, j)) = make_uint4(hi[0], hi[1], hi[2], hi[3]);
            *reinterpret_cast<uint4*>(lo_t + tc::chunk_off(rows, r, j)) = make_uint4(lo[0], lo[1], lo[2], lo[3]);
        }
    };
    stage(A, 128, K, amn, a_hi, a_lo);
    stage(B, N, K, bmn, b_hi, b_lo);
    if (warp == 0) tc::tmem_alloc(&tmem_slot, 256);
    if (tid == 0) {
        tc::mbar_init(&bar, 1);
        tc::fence_mbar_init();
    }
    tc::fence_async_smem();
    tc::fence_before_sync();
    __syncthreads();
    tc::fence_after_sync();
    const uint32_t tmem = tmem_slot;
    if (tid == 0) {
        const uint32_t idesc = tc::make_idesc_bf16(128, N) | (amn ? 1u << 15 : 0u) | (bmn ? 1u << 16 : 0u);
        // K-major: chunk stride (LBO) = rows*16, 8-row group stride (SBO) = 128, k-step = 2 chunks
        // MN-major: tile rows = K index; stride between 8-wide M/N blocks = K*16, between 8-deep K blocks = 128
        auto desc = [&](const uint8_t* t, int rows, bool mn, int s) {
            if (!mn) {
                const uint32_t lbo = rows * 16, sbo = 128, addr = tc::smem_u32(t) + s * 2 * lbo;
                return swapk ? tc::make_desc(addr, sbo, lbo) : tc::make_desc(addr, lbo, sbo);
            }
            const uint32_t mnstride = K * 16, kstride = 128, addr = tc::smem_u32(t) + s * 2 * kstride;
            return swapmn ? tc::make_desc(addr, kstride, mnstride) : tc::make_desc(addr, mnstride, kstride);
        };
        bool acc = false;
        for (int p = 0; p < npass; ++p) {
            const uint8_t* ta = p == 1 ? a_lo : a_hi;
            const uint8_t* tb = p == 2 ? b_lo : b_hi;
            for (int s = 0; s < K / 16; ++s) {
                tc::mma_bf16(tmem, desc(ta, 128, amn, s), desc(tb, N, bmn, s), idesc, acc);
                acc = true;
            }
        }
        tc::mma_commit(&bar);
    }
    const bool ok = tc::mbar_wait(&bar, 0);
    tc::fence_after_sync();
    if (!ok && tid == 0) *status = 1;
    if (ok) {
        for (int c0 = 0; c0 < N; c0 += 32) {
            float v[32];
            tc::tmem_ld32(tmem + ((uint32_t)(warp * 32) << 16) + c0, v);
#pragma unroll
            for (int i = 0; i < 32; ++i)
                if (c0 + i < N) D[tid * N + c0 + i] = v[i];
        }
    }
    tc::fence_before_sync();
    __syncthreads();
    if (warp == 0) tc::tmem_dealloc(tmem, 256);
}

}  // namespace

extern "C" int egnn_tc_selftest(const float* A, const float* B, float* D, int32_t K, int32_t N, int32_t npass,
                                int32_t flags, int32_t* status_dev, void* stream) {
    EGNN_REQUIRE(A && B && D && status_dev, "egnn_tc_selftest: NULL pointer");
    EGNN_REQUIRE(K >= 16 && K % 16 == 0 && K <= 128 && N >= 16 && N % 16 == 0 && N <= 256 && (npass == 1 || npass == 3),
                 "egnn_tc_selftest: unsupported shape K=%d N=%d npass=%d", K, N, npass);
    const size_t smem = 200 * 1024;  // fixed and ample: the swapped-descriptor probe must stay inside the allocation
    EGNN_SMEM_ATTR(tc_selftest_kernel, 200 * 1024);
    tc_selftest_kernel<<<1, 128, smem, (cudaStream_t)stream>>>(A, B, D, K, N, npass, flags, status_dev);
    EGNN_CHECK_LAUNCH("tc_selftest_kernel");
    return EGNN_OK;
}

// =====================================================================================================================
// Fused EGCL edge kernel on tcgen05 (forward).
//
// Per 128-edge tile (one TMEM lane per edge):   u = [h_i, h_j, |r|^2]  (K padded to 16)
//   GEMM1  [128 x 16] . W1^T   -> TMEM cols [0, 2d)      W1 = [Wm1; Wx1]  (both branches in one N = 2d MMA)
//   GEMM2x [128 x d]  . Wx2^T  -> TMEM cols [3d, 4d)     A = SiLU(GEMM1[:, d:2d] + bx1)  (bf16 hi/lo in smem)
//   GEMM2m [128 x d]  . Wm2^T  -> TMEM cols [2d, 3d)     A = SiLU(GEMM1[:, 0:d] + bm1)
//   phi = wx3 . SiLU(GEMM2x + bx2) + bx3 ;  m = SiLU(GEMM2m + bm2) ;  g = sigmoid(wa . m + ba)
// followed by the same fixed-order segmented sum as the fp32 path, over d/2-edge segments.
// NPASS = 3: every product is hi.hi + lo.hi + hi.lo of bf16 splits (fp32-class accuracy, "bf16x3");  NPASS = 1: plain bf16.
// =====================================================================================================================
namespace {

template <int D>
struct TcLayout {
    // weight image (identical in global memory and at the start of shared memory)
    static constexpr int W1_BYTES = 2 * D * 32;   // [2D rows x 16] bf16
    static constexpr int W2_BYTES = D * D * 2;    // [D rows x D] bf16
    static constexpr int w1_hi = 0, w1_lo = W1_BYTES;
    static constexpr int w2m_hi = 2 * W1_BYTES, w2m_lo = w2m_hi + W2_BYTES;
    static constexpr int w2x_hi = w2m_lo + W2_BYTES, w2x_lo = w2x_hi + W2_BYTES;
    // All four matrices and the four bias vectors are stored HALVED (exact in bf16 / fp32): the MMAs then deliver z / 2,
    // which is what the one-MUFU SiLU (tc_silu_h) consumes -- one multiply per element less in every epilogue.  The
    // layer-1 biases sit in K slot 15 of W1 (rows of U carry a constant 1 there), so GEMM1 adds them for free.
    static constexpr int vec = w2x_lo + W2_BYTES;  // fp32: bm1/2 [D] | bx1/2 [D] | bm2/2 [D] | bx2/2 [D] | wa [D] | wx3 [D] | ba, bx3, 0, 0
    static constexpr int BLOB = vec + (6 * D + 4) * 4;
    static constexpr int blob_pad = (BLOB + 15) / 16 * 16;
    // per-tile operands
    static constexpr int u_hi = blob_pad, u_lo = u_hi + 128 * 32;
    static constexpr int a1_hi = u_lo + 128 * 32, a1_lo = a1_hi + 128 * D * 2;
    static constexpr int stage = a1_hi;  // fp32 [128][D] staging of g*m reuses the A1 tiles (needs D*4 <= 2 * D*2: equal)
    static constexpr int small = a1_lo + 128 * D * 2;
    // small: phiPart [D/32][128] | gatePart [D/32][128] | sDst int[128] | sMask u32[4] | bars | slot
    static constexpr int phiPart = small, gatePart = phiPart + (D / 32) * 512, sDst = gatePart + (D / 32) * 512,
                         sMask = sDst + 512, bars = sMask + 16, slot = bars + 64;
    static constexpr int TOTAL = slot + 16;
    static_assert(TOTAL <= 227 * 1024, "tc_edge_fwd shared memory");
};

__global__ void __launch_bounds__(256) tc_pack_kernel(egnn_net_t net, const float* __restrict__ params,
                                                      float* __restrict__ packed) {
    const LayerOff lo = layer_off(net);
    const PackOff po = pack_off(net);
    const int D = net.Mh, K0 = 2 * net.H + 1;
    const float* P = params + (size_t)blockIdx.y * lo.total;
    uint8_t* blob = reinterpret_cast<uint8_t*>(packed + (size_t)blockIdx.y * po.total + po.tc);
    const int W1B = 2 * D * 32, W2B = D * D * 2;
    const int stride = gridDim.x * blockDim.x, t0 = blockIdx.x * blockDim.x + threadIdx.x;
    // W1 = [Wm1; Wx1], K padded to 16: chunks (n, j), j in {0,1}
    for (int idx = t0; idx < 2 * D * 2; idx += stride) {
        const int n = idx >> 1, j = idx & 1;
        const float* row = n < D ? P + lo.wm1 + n * K0 : P + lo.wx1 + (n - D) * K0;
        const float bias = n < D ? P[lo.bm1 + n] : P[lo.bx1 + n - D];
        uint32_t hi[4], lw[4];
        for (int q = 0; q < 4; ++q) {
            const int k = j * 8 + 2 * q;  // K slot 15 (always >= K0) carries the bias: u[15] is the constant 1
            tc::split2(k < K0 ? 0.5f * row[k] : 0.f, k + 1 < K0 ? 0.5f * row[k + 1] : (k + 1 == 15 ? 0.5f * bias : 0.f), hi[q],
                       lw[q]);
        }
        *reinterpret_cast<uint4*>(blob + tc::chunk_off(2 * D, n, j)) = make_uint4(hi[0], hi[1], hi[2], hi[3]);
        *reinterpret_cast<uint4*>(blob + W1B + tc::chunk_off(2 * D, n, j)) = make_uint4(lw[0], lw[1], lw[2], lw[3]);
    }
    for (int which = 0; which < 2; ++which) {
        const float* W = P + (which == 0 ? lo.wm2 : lo.wx2);
        uint8_t* dst = blob + 2 * W1B + which * 2 * W2B;
        for (int idx = t0; idx < D * (D / 8); idx += stride) {
            const int n = idx / (D / 8), j = idx % (D / 8);
            uint32_t hi[4], lw[4];
            for (int q = 0; q < 4; ++q)
                tc::split2(0.5f * W[n * D + j * 8 + 2 * q], 0.5f * W[n * D + j * 8 + 2 * q + 1], hi[q], lw[q]);
            *reinterpret_cast<uint4*>(dst + tc::chunk_off(D, n, j)) = make_uint4(hi[0], hi[1], hi[2], hi[3]);
            *reinterpret_cast<uint4*>(dst + W2B + tc::chunk_off(D, n, j)) = make_uint4(lw[0], lw[1], lw[2], lw[3]);
        }
    }
    if (tc_w2t_bytes(D) > 0) {  // d = 256: W2^T / 2 as K-major [rows k][K = n] images, per branch hi | lo
        for (int which = 0; which < 2; ++which) {
            const float* W = P + (which == 0 ? lo.wm2 : lo.wx2);
            uint8_t* dst = blob + tc_w2t_offset(D) + which * 2 * W2B;
            for (int idx = t0; idx < D * (D / 8); idx += stride) {
                const int k = idx % D, jn = idx / D;  // consecutive threads read consecutive k of one W2 row
                uint32_t hi[4], lw[4];
                for (int q = 0; q < 4; ++q)
                    tc::split2(0.5f * W[(size_t)(jn * 8 + 2 * q) * D + k], 0.5f * W[(size_t)(jn * 8 + 2 * q + 1) * D + k], hi[q], lw[q]);
                *reinterpret_cast<uint4*>(dst + tc::chunk_off(D, k, jn)) = make_uint4(hi[0], hi[1], hi[2], hi[3]);
                *reinterpret_cast<uint4*>(dst + W2B + tc::chunk_off(D, k, jn)) = make_uint4(lw[0], lw[1], lw[2], lw[3]);
            }
        }
    }
    float* vec = reinterpret_cast<float*>(blob + 2 * W1B + 4 * W2B);
    for (int c = t0; c < D; c += stride) {
        vec[c] = 0.5f * P[lo.bm1 + c];  // (layer-1 biases travel in K slot 15 of W1; kept here for reference only)
        vec[D + c] = 0.5f * P[lo.bx1 + c];
        vec[2 * D + c] = 0.5f * P[lo.bm2 + c];
        vec[3 * D + c] = 0.5f * P[lo.bx2 + c];
        vec[4 * D + c] = P[lo.wa + c];
        vec[5 * D + c] = P[lo.wx3 + c];
    }
    if (tc_node_supported(net)) {  // Wh1 [Hh x (H+Ms)] -> K-major [D x (D+16)] tile, zero-padded columns
        uint8_t* nb = blob + (tc_blob_bytes(D) + 15) / 16 * 16;
        const int KA = net.H + net.Ms, NCH = (D + 16) / 8, WB = D * (D + 16) * 2;
        const float* Wh = P + lo.wh1;
        for (int idx = t0; idx < D * NCH; idx += stride) {
            const int n = idx / NCH, j = idx % NCH;
            uint32_t hi[4], lw[4];
            for (int q = 0; q < 4; ++q) {
                const int k = j * 8 + 2 * q;
                tc::split2(k < KA ? Wh[n * KA + k] : 0.f, k + 1 < KA ? Wh[n * KA + k + 1] : 0.f, hi[q], lw[q]);
            }
            *reinterpret_cast<uint4*>(nb + tc::chunk_off(D, n, j)) = make_uint4(hi[0], hi[1], hi[2], hi[3]);
            *reinterpret_cast<uint4*>(nb + WB + tc::chunk_off(D, n, j)) = make_uint4(lw[0], lw[1], lw[2], lw[3]);
        }
    }
    if (t0 == 0) {
        vec[6 * D + 0] = P[lo.ba];
        vec[6 * D + 1] = P[lo.bx3];
        vec[6 * D + 2] = 0.f;
        vec[6 * D + 3] = 0.f;
    }
}

}  // namespace

namespace {

struct TcFwdArgs {
    egnn_net_t net;
    int N, E, tiles;
    const int32_t *recv, *send, *row_ptr;
    const float *h, *x;
    const uint8_t* blob;  // this layer's weight image
    float *agg, *part;
    float* phi_edge;  // [E] phi_e of this layer (dxu / D are summed per receiver by the node kernel)
};

// 4*D threads: warp w owns TMEM lane quadrant w % 4 (edge rows 32(w%4)..+31) and the 32-column group w / 4 of every
// D-wide block.  Per 128-edge tile:
//   GEMM1 (both branches, issued one tile ahead) -> SiLU of the coordinate branch -> tile A1 -> GEMM2x
//   SiLU of the message branch in registers while GEMM2x runs -> tile A1 -> GEMM2m (+ next tile's GEMM1)
//   phi partial dots while GEMM2m runs -> m, gate -> g*m staged in fp32 (aliases A1) -> fixed-order segmented sums
// The next tile's gather loads are in flight from the top of the tile; its U tile is stored as soon as GEMM1 of the
// current tile has completed.
template <int D, int NPASS>
__global__ void __launch_bounds__(4 * D, 1) tc_edge_fwd(TcFwdArgs a) {
    using L = TcLayout<D>;
    constexpr bool FAST = true;  // one-MUFU SiLU / sigmoid (tanh.approx) in both tensor-core modes, see tc_tile.cuh
    constexpr bool LO = NPASS == 3;
    constexpr int NT = 4 * D, NG = D / 32;
    constexpr int TS = 32;  // edges per segment == partial-buffer granularity (part_tile_for)
    constexpr uint32_t CS = 128 * 16;
    extern __shared__ __align__(1024) uint8_t smem[];
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int q = warp & 3, cg = warp >> 2;
    const int r = q * 32 + lane;  // this thread's edge row == TMEM lane
    const int c0 = cg * 32;
    const int H = a.net.H, AGGW = D + 4;
    float* vec = reinterpret_cast<float*>(smem + L::vec);
    float* phiPart = reinterpret_cast<float*>(smem + L::phiPart);
    float* gatePart = reinterpret_cast<float*>(smem + L::gatePart);
    int* sDst = reinterpret_cast<int*>(smem + L::sDst);
    uint32_t* sMask = reinterpret_cast<uint32_t*>(smem + L::sMask);
    float* stage = reinterpret_cast<float*>(smem + L::stage);
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + L::bars);  // [0] GEMM1, [1] GEMM2x, [2] GEMM2m
    uint32_t* slot = reinterpret_cast<uint32_t*>(smem + L::slot);

    pdl_launch_dependents();  // the node kernel behind this one may start its own prologue now (common.cuh: PDL)
    for (int o = tid * 16; o < L::blob_pad; o += NT * 16)
        *reinterpret_cast<uint4*>(smem + o) = __ldg(reinterpret_cast<const uint4*>(a.blob + o));
    if (warp == 0) tc::tmem_alloc(slot, 4 * D);
    if (tid == 0) {
        for (int b = 0; b < 3; ++b) tc::mbar_init(&bars[b], 1);
        tc::fence_mbar_init();
    }
    const uint32_t sb = tc::smem_u32(smem);
    const TcOperand opU{sb + L::u_hi, sb + L::u_lo, CS, false};
    const TcOperand opA1{sb + L::a1_hi, sb + L::a1_lo, CS, false};
    const TcOperand opW1{sb + L::w1_hi, sb + L::w1_lo, 2 * D * 16, false};
    const TcOperand opW2m{sb + L::w2m_hi, sb + L::w2m_lo, D * 16, false}, opW2x{sb + L::w2x_hi, sb + L::w2x_lo, D * 16, false};

    // The gather of one edge row is split in three stages so that its two dependent global-load latencies hide under
    // other work: (1) edge ids, (2) node data of both ends, (3) u -> bf16 tile, and where the row's finished
    // segmented sum goes (>= 0: agg[node]; -1: head partial; -2: tail partial; -3: the run continues in the next row)
    struct Row {
        int i, inext;             // stage 1 (the sender id is only needed for the addresses of stage 2)
        int j;
        float xi[3], xj[3];       // stage 2
        float u[14];              // h_i | h_j already in their K slots
        int rp0, rp1;
    };
    auto gather_ids = [&](int tile, Row& g) {
        const int e = min(tile * 128 + tid, a.E - 1);  // rows past the end repeat the last edge; they are never flushed
        g.i = a.recv[e];
        g.j = a.send[e];
        g.inext = a.recv[min(e + 1, a.E - 1)];
    };
    auto gather_nodes = [&](Row& g) {
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            g.xi[k] = a.x[3 * g.i + k];
            g.xj[k] = a.x[3 * g.j + k];
        }
        const float* hi_ = a.h + (size_t)g.i * H;
        const float* hjm = a.h + (size_t)g.j * H - H;  // slot k in [H, 2H) reads h_j[k - H]
#pragma unroll
        for (int k = 0; k < 14; ++k) {
            const float* p = k < H ? hi_ : hjm;
            g.u[k] = k < 2 * H ? p[k] : 0.f;
        }
        g.rp0 = a.row_ptr[g.i];
        g.rp1 = a.row_ptr[g.i + 1];
    };
    // returns the flush destination of the row
    auto store_u = [&](int tile, const Row& g) -> int {
        const int e0 = tile * 128, nrows = min(128, a.E - e0);
        const bool valid = tid < nrows;
        const float rx = g.xi[0] - g.xj[0], ry = g.xi[1] - g.xj[1], rz = g.xi[2] - g.xj[2];
        const float nrm = sqrtf(rx * rx + ry * ry + rz * rz);
        const float d2 = nrm * nrm;
        float u[16];
#pragma unroll
        for (int k = 0; k < 14; ++k) u[k] = k == 2 * H ? d2 : g.u[k];
        u[14] = 2 * H == 14 ? d2 : 0.f;
        u[15] = 1.f;  // bias slot
#pragma unroll
        for (int j = 0; j < 2; ++j) {
            uint32_t hi[4], lo[4];
#pragma unroll
            for (int p = 0; p < 4; ++p) tc::split2(u[8 * j + 2 * p], u[8 * j + 2 * p + 1], hi[p], lo[p]);
            *reinterpret_cast<uint4*>(smem + L::u_hi + tc::chunk_off(128, tid, j)) = make_uint4(hi[0], hi[1], hi[2], hi[3]);
            *reinterpret_cast<uint4*>(smem + L::u_lo + tc::chunk_off(128, tid, j)) = make_uint4(lo[0], lo[1], lo[2], lo[3]);
        }
        int dst = -3;
        if (valid) {
            const bool last = tid == nrows - 1 || (tid + 1) % TS == 0 || g.inext != g.i;
            if (last) {
                const int es = e0 + (tid / TS) * TS;
                dst = g.rp0 < es ? -1 : (g.rp1 > es + TS ? -2 : g.i);
            }
        }
        return dst;
    };
    auto store_dst = [&](int dst) {  // tid < 128: one mask word per 32-row segment marks the rows that flush
        sDst[tid] = dst;
        const uint32_t m = __ballot_sync(0xffffffffu, dst != -3);
        if (lane == 0) sMask[warp] = m;
    };

    const int nit = blockIdx.x < a.tiles ? (a.tiles - blockIdx.x + gridDim.x - 1) / gridDim.x : 0;
    pdl_wait();  // everything above read only the packed weights; h, x (and agg / part, written below) belong to the chain
    if (tid < 128 && nit > 0) {
        Row g;
        gather_ids(blockIdx.x, g);
        gather_nodes(g);
        store_dst(store_u(blockIdx.x, g));
    }
    tc::fence_async_smem();
    tc::fence_before_sync();
    __syncthreads();
    tc::fence_after_sync();
    const uint32_t tmem = *slot;
    const uint32_t trow = tmem + ((uint32_t)(q * 32) << 16);
    if (warp == NT / 32 - 1 && nit > 0 && tc::elect_one()) {
        tc_issue_op<NPASS>(tmem, 0, opU, opW1, 16, 2 * D, false);
        tc::mma_commit(&bars[0]);
    }
    bool ok = true;

    for (int it = 0; it < nit; ++it) {
        const int tile = blockIdx.x + it * gridDim.x;
        const int e0 = tile * 128;
        const int nrows = min(128, a.E - e0);
        const uint32_t phase = it & 1;
        const bool has_next = it + 1 < nit;
        Row gnext;
        int dst_next = -3;
        if (tid < 128 && has_next) gather_ids(tile + gridDim.x, gnext);   // stage 1: lands during the GEMM1 wait
        ok = tc::mbar_wait(&bars[0], phase) && ok;
        tc::fence_after_sync();
        if (tid < 128 && has_next) gather_nodes(gnext);                     // stage 2: lands during the epilogue below
        // ---- layer-1 epilogue of the coordinate branch -> A1; the U tile is free again: store the next tile's rows
        tc_epi_store_chunked(trow + D + c0, smem + L::a1_hi, smem + L::a1_lo, r, c0, LO);
        if (tid < 128 && has_next) dst_next = store_u(tile + gridDim.x, gnext);  // stage 3
        tc::fence_async_smem();
        tc::fence_before_sync();
        __syncthreads();
        if (warp == NT / 32 - 1 && tc::elect_one()) {
            tc::fence_after_sync();
            tc_issue_op<NPASS>(tmem, 3 * D, opA1, opW2x, D, D, false);
            tc::mma_commit(&bars[1]);
        }
        // ---- layer-1 epilogue of the message branch: in registers while GEMM2x runs
        uint32_t mhi[16], mlo[16];
        tc_epi32(trow + c0, mhi, mlo);
        ok = tc::mbar_wait(&bars[1], phase) && ok;  // GEMM2x has consumed A1
        tc::fence_after_sync();
        tc_store32(smem + L::a1_hi, smem + L::a1_lo, r, c0, mhi, mlo, LO);
        tc::fence_async_smem();
        tc::fence_before_sync();
        __syncthreads();
        if (warp == NT / 32 - 1 && tc::elect_one()) {
            tc::fence_after_sync();
            tc_issue_op<NPASS>(tmem, 2 * D, opA1, opW2m, D, D, false);
            tc::mma_commit(&bars[2]);
            if (has_next) {  // z1 of both branches has been consumed: next tile's GEMM1 queues behind GEMM2m
                tc_issue_op<NPASS>(tmem, 0, opU, opW1, 16, 2 * D, false);
                tc::mma_commit(&bars[0]);
            }
        }
        // ---- layer-2 epilogue of the coordinate branch while GEMM2m runs: phi partial over this thread's columns
        {
            float s = 0.f;
            float v[32];
            tc::tmem_ld32(trow + 3 * D + c0, v);
#pragma unroll
            for (int i = 0; i < 32; ++i) s = fmaf(vec[5 * D + c0 + i], tc_silu_h(v[i] + vec[3 * D + c0 + i]), s);
            phiPart[cg * 128 + r] = s;
        }
        ok = tc::mbar_wait(&bars[2], phase) && ok;
        tc::fence_after_sync();
        // ---- layer-2 epilogue of the message branch: m in registers, gate partial
        float m[32];
        {
            float s = 0.f;
            float v[32];
            tc::tmem_ld32(trow + 2 * D + c0, v);
#pragma unroll
            for (int i = 0; i < 32; ++i) {
                m[i] = tc_silu_h(v[i] + vec[2 * D + c0 + i]);
                s = fmaf(vec[4 * D + c0 + i], m[i], s);
            }
            gatePart[cg * 128 + r] = s;
        }
        tc::fence_before_sync();
        __syncthreads();  // all column groups of phi / gate are in smem; GEMM2m is done with A1 -> staging may overwrite it
        {
            float gs = vec[6 * D];
#pragma unroll
            for (int k = 0; k < NG; ++k) gs += gatePart[k * 128 + r];
            const float g = sigmoid_f(gs);
            if (cg == 0 && r < nrows) {
                float ph = vec[6 * D + 1];
#pragma unroll
                for (int k = 0; k < NG; ++k) ph += phiPart[k * 128 + r];
                a.phi_edge[e0 + r] = ph;
            }
#pragma unroll
            for (int c = 0; c < 32; ++c) stage[r * D + ((c0 + c) ^ lane)] = g * m[c];
        }
        __syncthreads();
        // ---- fixed-order segmented sums over receivers: thread (segment, column) walks its 32 rows in edge order and
        //      flushes at the rows marked in the segment's mask word
        {
            const int seg = tid / D, c = tid % D;
            const int r0 = seg * TS;
            float* part = a.part + (size_t)(tile * 4 + seg) * 2 * AGGW;
            uint32_t mask = sMask[seg];
            float sum = 0.f;
            int rr = r0;
            while (mask) {
                const int stop = r0 + __ffs(mask) - 1;
                mask &= mask - 1;
#pragma unroll 4
                for (; rr <= stop; ++rr) sum += stage[rr * D + (c ^ (rr & 31))];
                const int dst = sDst[stop];
                if (dst >= 0) a.agg[(size_t)dst * AGGW + c] = sum;
                else part[(dst == -1 ? 0 : AGGW) + c] = sum;
                sum = 0.f;
            }
        }
        __syncthreads();  // staging (aliases A1) and the flush table are free
        if (tid < 128 && has_next) store_dst(dst_next);
    }
    if (!ok) asm volatile("trap;");  // an mbarrier wait timed out: fail loudly instead of returning garbage
    tc::fence_before_sync();
    __syncthreads();
    if (warp == 0) tc::tmem_dealloc(tmem, 4 * D);
}

}  // namespace

// ---- host side -------------------------------------------------------------------------------------------------
int egnn_tc_pack(const egnn_net_t& net, const float* params, float* packed, cudaStream_t st) {
    if (!tc_fwd_supported(net)) return EGNN_OK;
    tc_pack_kernel<<<dim3(16, net.L), 256, 0, st>>>(net, params, packed);
    EGNN_CHECK_LAUNCH("tc_pack_kernel");
    return EGNN_OK;
}

template <int D, int NPASS>
static int launch_tc_fwd(const TcFwdArgs& a, cudaStream_t st, bool pdl) {
    EGNN_SMEM_ATTR((tc_edge_fwd<D, NPASS>), TcLayout<D>::TOTAL);
    const int grid = a.tiles < 148 ? a.tiles : 148;
    egnn_prof_begin(EGNN_K_TC_EDGE_FWD, st);
    EGNN_CUDA(egnn_launch(tc_edge_fwd<D, NPASS>, dim3(grid), dim3(4 * D), TcLayout<D>::TOTAL, st, pdl, a));
    egnn_prof_end(st);
    EGNN_CHECK_LAUNCH("tc_edge_fwd");
    return EGNN_OK;
}

int egnn_tc_edge_forward256(const egnn_net_t& net, const egnn_graph_t& g, const float* packed_layer, const float* h,
                            const float* x, float* agg, float* part, float* phi_edge, int precision, cudaStream_t st);

// edge phase of one layer on the tensor cores; fills agg / part exactly like egcl_edge_fwd (segments of d/2 edges)
int egnn_tc_edge_forward(const egnn_net_t& net, const egnn_graph_t& g, const float* packed_layer, const float* h,
                         const float* x, float* agg, float* part, float* phi_edge, int precision, cudaStream_t st, bool pdl) {
    EGNN_REQUIRE(tc_fwd_supported(net), "tcgen05 path needs Mh = Ms = Xh in {64,128,256} and 2H+1 <= 16 (got Mh=%d Ms=%d Xh=%d H=%d); "
                 "use precision fp32", net.Mh, net.Ms, net.Xh, net.H);
    if (net.Mh == 256) return egnn_tc_edge_forward256(net, g, packed_layer, h, x, agg, part, phi_edge, precision, st);
    const PackOff po = pack_off(net);
    TcFwdArgs a{net, g.N, g.E, (g.E + 127) / 128, g.recv, g.send, g.row_ptr, h, x,
                reinterpret_cast<const uint8_t*>(packed_layer + po.tc), agg, part, phi_edge};
    const bool x3 = precision == EGNN_PREC_BF16X3;
    if (net.Mh == 128) return x3 ? launch_tc_fwd<128, 3>(a, st, pdl) : launch_tc_fwd<128, 1>(a, st, pdl);
    return x3 ? launch_tc_fwd<64, 3>(a, st, pdl) : launch_tc_fwd<64, 1>(a, st, pdl);
}

// =====================================================================================================================
// Fused EGCL edge kernel on tcgen05 (backward).  Each CTA owns ONE branch (blockIdx.x < nM: 0 = mlp_m + gate,
// 1 = mlp_x) and walks its 128-edge tiles with 4*D threads (thread = edge row x 32-column group); per tile
//   recompute   z1 = U.W1^T (block A),  a1 = SiLU(z1+b1) -> tile A1,  z2 = a1.W2^T (block B)
//   upstream    gz2 = dL/dz2 -> tile G            (gate / phi algebra of SURVEY.md Appendix A on the CUDA cores)
//   dgrad       ga1 = gz2 . W2   (block B)        B operand = the SAME W2 image viewed MN-major
//   wgrad       [dW2 | db2] += gz2^T . [a1 | u]   (persistent TMEM accumulators; slot 15 of u is the constant 1)
//   gz1 = ga1 * SiLU'(z1+b1) -> tile G;   gu = gz1 . W1 (16 columns),  [dW1 | db1] += gz1^T . u
// The weight image is stored halved (TcLayout): the MMAs give z / 2, the epilogues produce 2 SiLU' from the same tanh, so
// the gradient tiles hold 2 gz -- dgrad through the halved weights is then exact (2 gz . W / 2) and the persistent
// weight-gradient accumulators are halved once when they are written out.
// Shared-memory tiles are written once by the thread that owns the edge row (16-byte chunks) and serve both as
// K-major operands (row = edge) and as MN-major operands (K = edge) -- same bytes, two descriptor views.
// The two wgrad MMAs run behind the CUDA-core work of the next phase (own mbarriers); the U tile is double-buffered and
// the next tile's gather loads are in flight while GEMM2 runs.  Weight gradients stay in TMEM over all tiles of the
// CTA and are written once to the CTA's private partial row (summed in CTA order by reduce_partials_kernel).
// =====================================================================================================================
namespace {

template <int D>
struct TcBwdLayout {
    static constexpr int NT = 4 * D, NG = D / 32;
    static constexpr int TILE = 128 * D * 2, UT = 128 * 32;  // bytes of one hi (or lo) activation tile / U tile
    static constexpr int g_hi = 0, g_lo = TILE;
    // hi group u0 | a1 | u1 and the same for lo: the wgrad B operand [u0 | a1] (even tiles) or [a1 | u1] (odd tiles) is
    // one MN-major tile of D/8 + 2 chunk columns
    static constexpr int u0_hi = 2 * TILE, a1_hi = u0_hi + UT, u1_hi = a1_hi + TILE;
    static constexpr int u0_lo = u1_hi + UT, a1_lo = u0_lo + UT, u1_lo = a1_lo + TILE;
    static constexpr int w1_hi = u1_lo + UT, w1_lo = w1_hi + D * 32;  // this branch's [D x 16] image
    static constexpr int w2_hi = w1_lo + D * 32, w2_lo = w2_hi + D * D * 2;
    static constexpr int vec = w2_lo + D * D * 2;  // b1 | b2 | wv (D each) | bv, 0, 0, 0
    static constexpr int VEC_BYTES = (3 * D + 4) * 4;
    static constexpr int sI = vec + VEC_BYTES;           // int   [2][128]
    static constexpr int sR = sI + 1024;                 // float4 [2][128]: r_e, valid
    static constexpr int rowA = sR + 4096;               // float [NG][128]   (reused as colred [NT/32][32] at the end)
    static constexpr int rowB = rowA + NG * 512;         // float [NG][128]
    static constexpr int bars = rowB + NG * 512, slot = bars + 64;
    static constexpr int TOTAL = slot + 16;
    static_assert(TOTAL <= 227 * 1024, "tc_edge_bwd shared memory");
};

struct TcBwdArgs {
    egnn_net_t net;
    int N, E, tiles;
    const int32_t *recv, *send, *node_graph;
    const float *h, *x;
    const uint8_t* blob;
    const float *gM, *gdx, *coefS;
    float *Gr, *Gs;  // [2][E][H+3]: branch-major; the gather kernel adds both branches
    float* part;     // [gridDim.x][layer param count]
    int nM;          // CTAs [0, nM) take the message branch, [nM, gridDim.x) the coordinate branch
    int g_lo;        // bit 0 / bit 1: the gradient tile gz2 / gz1 is stored as bf16 hi + lo (else hi only: 2 products per MMA)
    int wg_lo;       // weight-gradient MMAs issue all three split products (small edge counts, see egnn_tc_edge_backward)
};

struct TcEdgeRow {  // one gathered edge held in registers between its global loads and its shared-memory stores
    int i;
    float rx, ry, rz, valid;
    float u[16];
};

template <int D, int NPASS>
__global__ void __launch_bounds__(4 * D, 1) tc_edge_bwd(TcBwdArgs a) {
    using L = TcBwdLayout<D>;
    constexpr bool FAST = true;  // one-MUFU SiLU / sigmoid (tanh.approx) in both tensor-core modes, see tc_tile.cuh
    constexpr bool LO = NPASS == 3;
    const bool GLO2 = LO && (a.g_lo & 1), GLO1 = LO && (a.g_lo & 2);  // lo halves of the gz2 / gz1 tiles
    const bool WGLO2 = GLO2 && a.wg_lo, WGLO1 = GLO1 && a.wg_lo;      // third product (gz_lo . x_hi) of the wgrad MMAs
    constexpr int NT = L::NT, NG = L::NG;
    constexpr int OUTW = NG >= 3 ? 2 : 1;  // warp groups that write the per-edge outputs: cg 1 (Gr) and cg 2 (Gs) at d = 128
    // TMEM columns: z1 -> SiLU'(z1),  z2 | ga1,  db2 (even tiles),  dW2,  db2 (odd tiles),  dW1 + db1,  gu
    constexpr int cA = 0, cB = D, cBias0 = 2 * D, cW2 = 2 * D + 16, cBias1 = 3 * D + 16, cP1 = 3 * D + 32,
                  cGU = 3 * D + 48;
    constexpr uint32_t TS = 128 * 16;  // chunk stride of a 128-row tile
    extern __shared__ __align__(1024) uint8_t smem[];
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int q = warp & 3, cg = warp >> 2;
    const int r = q * 32 + lane;  // edge row == TMEM lane
    const int c0 = cg * 32;       // this thread's 32 columns
    const int branch = (int)blockIdx.x < a.nM ? 0 : 1;
    const int H = a.net.H, K0 = 2 * H + 1, W = H + 3;
    float* vec = reinterpret_cast<float*>(smem + L::vec);
    int* sI = reinterpret_cast<int*>(smem + L::sI);
    float4* sR = reinterpret_cast<float4*>(smem + L::sR);
    float* rowA = reinterpret_cast<float*>(smem + L::rowA);
    float* rowB = reinterpret_cast<float*>(smem + L::rowB);
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + L::bars);  // 0 GEMM1, 1 GEMM2, 2 dgrad, 3 wgrad2, 4 gu, 5 wgrad1
    uint32_t* slot = reinterpret_cast<uint32_t*>(smem + L::slot);

    // ---- one-time setup: this branch's weight images and vectors -> smem; TMEM; barriers
    {
        constexpr int W1B = 2 * D * 32, W2B = D * D * 2;  // sizes inside the packed image (TcLayout)
        for (int o = tid * 16; o < 4 * D * 16; o += NT * 16) {  // [hi|lo][chunk j][D*16 bytes of this branch's rows]
            const int half = o / (2 * D * 16), j = (o / (D * 16)) & 1, off = o % (D * 16);
            *reinterpret_cast<uint4*>(smem + (half ? L::w1_lo : L::w1_hi) + j * D * 16 + off) =
                __ldg(reinterpret_cast<const uint4*>(a.blob + half * W1B + j * 2 * D * 16 + branch * D * 16 + off));
        }
        const uint8_t* w2src = a.blob + 2 * W1B + (size_t)branch * 2 * W2B;
        for (int o = tid * 16; o < 2 * W2B; o += NT * 16)
            *reinterpret_cast<uint4*>(smem + L::w2_hi + o) = __ldg(reinterpret_cast<const uint4*>(w2src + o));
        const float* vsrc = reinterpret_cast<const float*>(a.blob + 2 * W1B + 4 * W2B);
        for (int c = tid; c < D; c += NT) {
            vec[c] = vsrc[branch * D + c];
            vec[D + c] = vsrc[2 * D + branch * D + c];
            vec[2 * D + c] = vsrc[4 * D + branch * D + c];
        }
        if (tid == 0) vec[3 * D] = vsrc[6 * D + branch];
    }
    if (warp == 0) tc::tmem_alloc(slot, 4 * D);
    if (tid == 0) {
        for (int b = 0; b < 6; ++b) tc::mbar_init(&bars[b], 1);
        tc::fence_mbar_init();
    }
    tc::fence_before_sync();
    __syncthreads();
    tc::fence_after_sync();
    const uint32_t tmem = *slot;
    const uint32_t trow = tmem + ((uint32_t)(q * 32) << 16);
    for (int cb = cg; cb < 2 * D / 32; cb += NG) tc::tmem_zero32(trow + 2 * D + 32 * cb);  // persistent accumulators = 0

    const uint32_t sb = tc::smem_u32(smem);
    const TcOperand opA1{sb + L::a1_hi, sb + L::a1_lo, TS, false};
    const TcOperand opG{sb + L::g_hi, sb + L::g_lo, TS, false}, opGmn{sb + L::g_hi, sb + L::g_lo, TS, true};
    const TcOperand opW1{sb + L::w1_hi, sb + L::w1_lo, D * 16, false}, opW1mn{sb + L::w1_hi, sb + L::w1_lo, D * 16, true};
    const TcOperand opW2{sb + L::w2_hi, sb + L::w2_lo, D * 16, false}, opW2mn{sb + L::w2_hi, sb + L::w2_lo, D * 16, true};

    const float* b2 = vec + D;      // b2 / 2 (the image is stored halved, see TcLayout)
    const float* wv = vec + 2 * D;  // wa (gate) or wx3 (phi head)
    float* Grb = a.Gr + (size_t)branch * a.E * W;
    float* Gsb = a.Gs + (size_t)branch * a.E * W;

    auto gather_load = [&](int tile, TcEdgeRow& g) {
        const int e = tile * 128 + tid;
#pragma unroll
        for (int k = 0; k < 16; ++k) g.u[k] = 0.f;
        g.i = 0;
        g.rx = g.ry = g.rz = g.valid = 0.f;
        if (e < a.E) {
            const int i = a.recv[e], j = a.send[e];
            g.i = i;
            g.rx = a.x[3 * i + 0] - a.x[3 * j + 0];
            g.ry = a.x[3 * i + 1] - a.x[3 * j + 1];
            g.rz = a.x[3 * i + 2] - a.x[3 * j + 2];
            g.valid = 1.f;
            const float nrm = sqrtf(g.rx * g.rx + g.ry * g.ry + g.rz * g.rz);
#pragma unroll
            for (int k = 0; k < 15; ++k) {
                if (k < H) g.u[k] = a.h[(size_t)i * H + k];
                else if (k < 2 * H) g.u[k] = a.h[(size_t)j * H + (k - H)];
                else if (k == 2 * H) g.u[k] = nrm * nrm;
            }
            g.u[15] = 1.f;  // turns the wgrad MMAs into bias gradients as well
        }
    };
    auto gather_store = [&](int buf, const TcEdgeRow& g) {
        sI[buf * 128 + tid] = g.i;
        sR[buf * 128 + tid] = make_float4(g.rx, g.ry, g.rz, g.valid);
        uint8_t* uh = smem + (buf ? L::u1_hi : L::u0_hi);
        uint8_t* ul = smem + (buf ? L::u1_lo : L::u0_lo);
#pragma unroll
        for (int j = 0; j < 2; ++j) {
            uint32_t hi[4], lo[4];
#pragma unroll
            for (int p = 0; p < 4; ++p) tc::split2(g.u[8 * j + 2 * p], g.u[8 * j + 2 * p + 1], hi[p], lo[p]);
            *reinterpret_cast<uint4*>(uh + tc::chunk_off(128, tid, j)) = make_uint4(hi[0], hi[1], hi[2], hi[3]);
            *reinterpret_cast<uint4*>(ul + tc::chunk_off(128, tid, j)) = make_uint4(lo[0], lo[1], lo[2], lo[3]);
        }
    };

    const int tstride = branch ? (int)gridDim.x - a.nM : a.nM;
    const int tile0 = branch ? (int)blockIdx.x - a.nM : (int)blockIdx.x;
    const int nit = tile0 < a.tiles ? (a.tiles - tile0 + tstride - 1) / tstride : 0;
    // this branch's per-edge outputs of a finished tile.  Run by the cg == 1 warps (same TMEM lanes as cg == 0), so that they
    // proceed NEXT TO the cg == 0 warps' gather of the following tile instead of behind it; the row's record (rr, i) is read
    // from shared memory by the caller before the gather store may overwrite that buffer
    // part: 0 = both slabs, 1 = the receiver slab Gr only, 2 = the sender slab Gs only (d = 128: cg == 1 / cg == 2 warps)
    auto edge_outputs = [&](int e0p, int nrowsp, const float4 rr, const int i, float phi, int part) {
        float gu[32];
        tc::tmem_ld32(trow + cGU, gu);
        const bool doR = part != 2, doS = part != 1;
        if (r < nrowsp) {
            const size_t e = (size_t)(e0p + r);
            float gd = 0.f;
#pragma unroll
            for (int k = 0; k < 15; ++k) {
                if (k < H) { if (doR) Grb[e * W + k] = gu[k]; }
                else if (k < 2 * H) { if (doS) Gsb[e * W + (k - H)] = gu[k]; }
                else if (k == 2 * H) gd = gu[k];
            }
            float cr = 2.f * gd;
            float gx0 = 0.f, gx1 = 0.f, gx2 = 0.f;
            if (branch == 1) {
                cr += a.coefS[a.node_graph[i]];
                gx0 = a.gdx[(size_t)i * 3 + 0];
                gx1 = a.gdx[(size_t)i * 3 + 1];
                gx2 = a.gdx[(size_t)i * 3 + 2];
            }
            const float g0 = fmaf(gx0, phi, cr * rr.x), g1 = fmaf(gx1, phi, cr * rr.y), g2 = fmaf(gx2, phi, cr * rr.z);
            if (doR) {
                Grb[e * W + H + 0] = g0;
                Grb[e * W + H + 1] = g1;
                Grb[e * W + H + 2] = g2;
            }
            if (doS) {
                Gsb[e * W + H + 0] = -g0;
                Gsb[e * W + H + 1] = -g1;
                Gsb[e * W + H + 2] = -g2;
            }
        }
    };

    if (tid < 128 && nit > 0) {
        TcEdgeRow g;
        gather_load(tile0, g);
        gather_store(0, g);
    }
    tc::fence_async_smem();
    tc::fence_before_sync();
    __syncthreads();
    if (warp == NT / 32 - 1 && nit > 0 && tc::elect_one()) {
        tc::fence_after_sync();
        const TcOperand opU0{sb + L::u0_hi, sb + L::u0_lo, TS, false};
        tc_issue_op<NPASS>(tmem, cA, opU0, opW1, 16, D, false);  // z1 of the first tile
        tc::mma_commit(&bars[0]);
    }

    float accv = 0.f;  // running column sum of coef_e * out_e (grad of wa / wx3): this warp's rows, column c0 + lane
    float accs = 0.f;  // running sum of coef_e of this thread's row (grad of ba / bx3)
    float phi_prev = 0.f;  // phi_e of the previous tile (coordinate branch), consumed by its deferred edge_outputs
    bool ok = true;

    for (int it = 0; it < nit; ++it) {
        const int tile = tile0 + it * tstride;
        const int buf = it & 1;
        const uint32_t phase = it & 1;
        const bool has_next = it + 1 < nit;
        const TcOperand opUmn{sb + (buf ? L::u1_hi : L::u0_hi), sb + (buf ? L::u1_lo : L::u0_lo), TS, true};
        ok = tc::mbar_wait(&bars[0], phase) && ok;  // z1 = U . W1^T of this tile (issued one phase ahead)
        tc::fence_after_sync();
        // ---- a1 = SiLU(z1 + b1) -> tile A1;  SiLU'(z1 + b1) replaces z1 in TMEM (read back by the gz1 phase)
        tc_epi_bwd_chunked(trow + cA + c0, smem + L::a1_hi, smem + L::a1_lo, r, c0, LO);  // TMEM holds z1 / 2 incl. bias
        tc::fence_async_smem();
        tc::fence_before_sync();
        __syncthreads();
        if (warp == NT / 32 - 1 && tc::elect_one()) {
            tc::fence_after_sync();
            tc_issue_op<NPASS>(tmem, cB, opA1, opW2, D, D, false);  // z2 = a1 . W2^T
            tc::mma_commit(&bars[1]);
        }
        // ---- in the shadow of GEMM2: next tile's gather loads, previous tile's per-edge outputs
        TcEdgeRow gnext;
        if (tid < 128 && has_next) gather_load(tile + tstride, gnext);
        if (it > 0 && cg >= 1 && cg <= OUTW) {
            const float4 rrp = sR[(buf ^ 1) * 128 + r];
            const int ip = sI[(buf ^ 1) * 128 + r];
            // records read: the cg == 0 warp of this row quadrant may now overwrite the buffer (named barrier 5 + q)
            if (has_next) asm volatile("bar.arrive %0, %1;" ::"r"(5 + q), "r"(32 * (1 + OUTW)) : "memory");
            ok = tc::mbar_wait(&bars[4], phase ^ 1) && ok;
            tc::fence_after_sync();
            edge_outputs((tile - tstride) * 128, min(128, a.E - (tile - tstride) * 128), rrp, ip, phi_prev, OUTW == 1 ? 0 : cg);
        }
        if (tid < 128 && has_next) {  // still in the shadow: the next tile's U rows (the previous dW1 MMA has left U[buf^1])
            if (it > 0) {
                ok = tc::mbar_wait(&bars[5], phase ^ 1) && ok;
                asm volatile("bar.sync %0, %1;" ::"r"(5 + q), "r"(32 * (1 + OUTW)) : "memory");
            }
            tc::fence_after_sync();
            gather_store(buf ^ 1, gnext);
        }
        ok = tc::mbar_wait(&bars[1], phase) && ok;
        if (it > 0) ok = tc::mbar_wait(&bars[5], phase ^ 1) && ok;  // previous tile's dW1 MMA is done with tile G as well
        tc::fence_after_sync();
        // ---- upstream gradient of z2: out = SiLU(z2 + b2) and SiLU' of this thread's 32 columns
        float cs[32];  // coef_e * out_e[c]: column sums taken after the MMAs below have been issued
        {
            float m[32], d2[32];
            {
                float v[32];
                tc::tmem_ld32(trow + cB + c0, v);
#pragma unroll
                for (int i = 0; i < 32; ++i) tc_silu_and_grad2_h(v[i] + b2[c0 + i], m[i], d2[i]);
            }
            const float4 rr = sR[buf * 128 + r];
            const int i = sI[buf * 128 + r];
            const float* wvh = wv + c0;
            float c_row, c_g = 0.f;  // gz2[c] = (c_g * gM[i][c] + c_row * wv[c]) * SiLU'(z[c]);  d wv[c] += c_row * out[c]
            const float4* gMi = reinterpret_cast<const float4*>(a.gM + (size_t)i * D + c0);
            if (branch == 0) {
                // gate g = sigmoid(wa.m + ba);  t = gM[i].m;  c_row = gs = t g (1-g);  c_g = g
                float sdot = 0.f, tdot = 0.f;
#pragma unroll
                for (int c4 = 0; c4 < 8; ++c4) {
                    const float4 gm = __ldg(gMi + c4);
                    sdot = fmaf(wvh[4 * c4], m[4 * c4], sdot);
                    sdot = fmaf(wvh[4 * c4 + 1], m[4 * c4 + 1], sdot);
                    sdot = fmaf(wvh[4 * c4 + 2], m[4 * c4 + 2], sdot);
                    sdot = fmaf(wvh[4 * c4 + 3], m[4 * c4 + 3], sdot);
                    tdot = fmaf(gm.x, m[4 * c4], tdot);
                    tdot = fmaf(gm.y, m[4 * c4 + 1], tdot);
                    tdot = fmaf(gm.z, m[4 * c4 + 2], tdot);
                    tdot = fmaf(gm.w, m[4 * c4 + 3], tdot);
                }
                rowA[cg * 128 + r] = sdot;
                rowB[cg * 128 + r] = tdot;
                // only the NG warps that share this row quadrant exchange partial dots: named barrier 1 + q
                asm volatile("bar.sync %0, %1;" ::"r"(1 + q), "r"(NG * 32) : "memory");
                float ssum = vec[3 * D], tsum = 0.f;
#pragma unroll
                for (int k = 0; k < NG; ++k) {
                    ssum += rowA[k * 128 + r];
                    tsum += rowB[k * 128 + r];
                }
                const float g = sigmoid_f(ssum);
                c_row = rr.w * tsum * g * (1.f - g);
                c_g = rr.w * g;
            } else {
                // phi = wx3.q + bx3 (needed for the coordinate output);  c_row = gphi = gdx[i].r
                float pdot = 0.f;
#pragma unroll
                for (int c = 0; c < 32; ++c) pdot = fmaf(wvh[c], m[c], pdot);
                rowA[cg * 128 + r] = pdot;
                const float* gd = a.gdx + (size_t)i * 3;
                c_row = rr.w * (gd[0] * rr.x + gd[1] * rr.y + gd[2] * rr.z);
            }
            float gz[32];
#pragma unroll
            for (int c4 = 0; c4 < 8; ++c4) {
                float4 gm = make_float4(0.f, 0.f, 0.f, 0.f);
                if (branch == 0) gm = __ldg(gMi + c4);
                const float gmv[4] = {gm.x, gm.y, gm.z, gm.w};
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    const int c = 4 * c4 + k;
                    gz[c] = fmaf(c_g, gmv[k], c_row * wvh[c]) * d2[c];
                    cs[c] = c_row * m[c];
                }
            }
            tc_split_store32(smem + L::g_hi, smem + L::g_lo, r, c0, gz, GLO2);
            accs += c_row;
        }
        tc::fence_async_smem();
        tc::fence_before_sync();
        __syncthreads();
        if (warp == NT / 32 - 1 && tc::elect_one()) {
            tc::fence_after_sync();
            tc_issue_op<NPASS>(tmem, cB, opG, opW2mn, D, D, false, GLO2);  // ga1 = gz2 . W2
            tc::mma_commit(&bars[2]);
            // [db2 | dW2] += gz2^T . [u | a1] (even tiles) or [dW2 | db2] += gz2^T . [a1 | u] (odd tiles)
            const TcOperand opB{sb + (buf ? L::a1_hi : L::u0_hi), sb + (buf ? L::a1_lo : L::u0_lo), TS, true};
            // weight gradients sum over every edge of the layer: the gz_lo . a1_hi product only carries the bf16 rounding
            // of gz (2^-9, unbiased), which averages out over ~1e5 terms -> two products instead of three; on small
            // edge lists it does not (measured 1.0e-3 on a bias gradient over 834 edges): there all three are issued
            tc_issue_op<NPASS>(tmem, buf ? cW2 : cBias0, opGmn, opB, 128, D + 16, true, WGLO2);
            tc::mma_commit(&bars[3]);
        }
        // ---- in the shadow of dgrad: column sums for d wa / d wx3; phi_e of this tile for its deferred outputs
        warp_colsum32(cs, lane);
        accv += cs[0];
        if (branch == 1 && cg >= 1 && cg <= OUTW) {
            float phi = vec[3 * D];
#pragma unroll
            for (int k = 0; k < NG; ++k) phi += rowA[k * 128 + r];
            phi_prev = phi;
        }
        ok = tc::mbar_wait(&bars[2], phase) && ok;
        tc::fence_after_sync();
        // ---- gz1 = ga1 * SiLU'(z1 + b1), computed while the wgrad MMA still reads tiles G and A1
        {
            float ga[32], d1[32];
            tc::tmem_ld32(trow + cB + c0, ga);
            tc::tmem_ld32(trow + cA + c0, d1);
#pragma unroll
            for (int i = 0; i < 32; ++i) ga[i] *= d1[i];
            ok = tc::mbar_wait(&bars[3], phase) && ok;
            tc::fence_after_sync();
            tc_split_store32(smem + L::g_hi, smem + L::g_lo, r, c0, ga, GLO1);
        }
        tc::fence_async_smem();
        tc::fence_before_sync();
        __syncthreads();
        if (warp == NT / 32 - 1 && tc::elect_one()) {
            tc::fence_after_sync();
            if (has_next) {  // next tile's z1 first: its epilogue overlaps the two narrow MMAs below
                const TcOperand opUn{sb + (buf ? L::u0_hi : L::u1_hi), sb + (buf ? L::u0_lo : L::u1_lo), TS, false};
                tc_issue_op<NPASS>(tmem, cA, opUn, opW1, 16, D, false);
                tc::mma_commit(&bars[0]);
            }
            tc_issue_op<NPASS>(tmem, cGU, opG, opW1mn, D, 16, false, GLO1);  // gu = gz1 . W1   (16 columns)
            tc::mma_commit(&bars[4]);
        } else if (warp == NT / 32 - 2 && tc::elect_one()) {  // independent of the two above: issued by another warp
            tc::fence_after_sync();
            tc_issue_op<NPASS>(tmem, cP1, opGmn, opUmn, 128, 16, true, WGLO1);  // [dW1 | db1] += gz1^T . u
            tc::mma_commit(&bars[5]);
        }
    }
    if (nit > 0) {
        const int lt = tile0 + (nit - 1) * tstride;
        const uint32_t lp = (uint32_t)((nit - 1) & 1);
        if (cg >= 1 && cg <= OUTW) {
            ok = tc::mbar_wait(&bars[4], lp) && ok;
            tc::fence_after_sync();
            edge_outputs(lt * 128, min(128, a.E - lt * 128), sR[((nit - 1) & 1) * 128 + r], sI[((nit - 1) & 1) * 128 + r], phi_prev,
                         OUTW == 1 ? 0 : cg);
        }
        ok = tc::mbar_wait(&bars[5], lp) && ok;
    }
    if (!ok) asm volatile("trap;");
    tc::fence_after_sync();
    // ---- weight gradients of this CTA's tiles: TMEM accumulators / register column sums -> private partial row
    if (nit > 0) {
        const LayerOff lo = layer_off(a.net);
        float* part = a.part + (size_t)blockIdx.x * lo.total;
        const int o_w1 = branch ? lo.wx1 : lo.wm1, o_b1 = branch ? lo.bx1 : lo.bm1;
        const int o_w2 = branch ? lo.wx2 : lo.wm2, o_b2 = branch ? lo.bx2 : lo.bm2;
        const int o_wv = branch ? lo.wx3 : lo.wa, o_bv = branch ? lo.bx3 : lo.ba;
        {
            float v[32];
            tc::tmem_ld32(trow + cW2 + c0, v);
            if (r < D) {
                float* dst = part + o_w2 + (size_t)r * D + c0;  // partial rows are not 16-byte aligned in general
#pragma unroll
                for (int k = 0; k < 32; ++k) dst[k] = 0.5f * v[k];  // the gz tiles carry 2 * dL/dz (see header)
            }
        }
        if (cg == 0) {
            float v0[32], v1[32];
            tc::tmem_ld32(trow + cBias0, v0);  // column 15: db2 of the even tiles
            tc::tmem_ld32(trow + cBias1, v1);  // column 15: db2 of the odd tiles; [16,32): dW1, column 31: db1
            if (r < D) {
                part[o_b2 + r] = 0.5f * (v0[15] + v1[15]);
                part[o_b1 + r] = 0.5f * v1[31];
#pragma unroll
                for (int k = 0; k < 15; ++k)
                    if (k < K0) part[o_w1 + r * K0 + k] = 0.5f * v1[16 + k];
            }
        }
        float* colred = rowA;  // [NT/32][32]
        colred[warp * 32 + lane] = accv;
        if (cg == 0) rowB[r] = accs;
        __syncthreads();
        if (tid < D) {
            const int gc = tid >> 5, cc = tid & 31;
            float s = 0.f;
#pragma unroll
            for (int qq = 0; qq < 4; ++qq) s += colred[(gc * 4 + qq) * 32 + cc];
            part[o_wv + tid] = s;
        }
        if (tid == 0) {
            float s = 0.f;
            for (int rr = 0; rr < 128; ++rr) s += rowB[rr];
            part[o_bv] = s;
        }
    }
    tc::fence_before_sync();
    __syncthreads();
    if (warp == 0) tc::tmem_dealloc(tmem, 4 * D);
}

template <int D, int NPASS>
int launch_tc_bwd(const TcBwdArgs& a, int grid, cudaStream_t st) {
    constexpr int SMEM = TcBwdLayout<D>::TOTAL;
    EGNN_SMEM_ATTR((tc_edge_bwd<D, NPASS>), SMEM);
    egnn_prof_begin(EGNN_K_TC_EDGE_BWD, st);
    tc_edge_bwd<D, NPASS><<<grid, 4 * D, SMEM, st>>>(a);
    egnn_prof_end(st);
    EGNN_CHECK_LAUNCH("tc_edge_bwd");
    return EGNN_OK;
}

}  // namespace

// CTAs (= private partial rows) of the tensor-core edge backward for E edges: nM message-branch CTAs followed by nX
// coordinate-branch CTAs, nM + nX <= 148, each <= tiles.  The message branch does more CUDA-core work per tile (gate,
// two row dots, the gM gather) but the coordinate branch writes more per edge; measured on B200 the even split is best
// (EGNN_TC_BWD_MSHARE percent, default 50; 46 and 54 are 1-8 % slower at cfg2).
void egnn_tc_bwd_split(int64_t E, int* nM, int* nX) {
    const int64_t tiles = (E + 127) / 128;
    if (2 * tiles <= 148) {
        *nM = *nX = (int)tiles;
        return;
    }
    static int share = -1;
    if (share < 0) {
#ifdef EGNN_DBG  // tuning knob of debug builds only: a stray environment variable must not change a release library
        const char* env = getenv("EGNN_TC_BWD_MSHARE");
        share = env ? atoi(env) : 50;
#else
        share = 50;
#endif
        if (share < 10 || share > 90) share = 50;
    }
    *nM = (148 * share + 50) / 100;
    *nX = 148 - *nM;
}
int egnn_tc_bwd_grid(int64_t E) {
    int nM, nX;
    egnn_tc_bwd_split(E, &nM, &nX);
    return nM + nX;
}

// edge phase of one layer's backward on the tensor cores.  Gr / Gs hold [2][E][H+3] (one slab per branch).
int egnn_tc_edge_backward(const egnn_net_t& net, const egnn_graph_t& g, const float* packed_layer, const float* h,
                          const float* x, const float* gM, const float* gdx, const float* coefS, float* Gr, float* Gs,
                          float* part, int precision, cudaStream_t st) {
    EGNN_REQUIRE(tc_supported(net), "tcgen05 path needs Mh = Ms = Xh in {64,128} and 2H+1 <= 16 (got Mh=%d Ms=%d Xh=%d H=%d); "
                 "use precision fp32", net.Mh, net.Ms, net.Xh, net.H);
    const PackOff po = pack_off(net);
    TcBwdArgs a{net, g.N, g.E, (g.E + 127) / 128, g.recv, g.send, g.node_graph, h, x,
                reinterpret_cast<const uint8_t*>(packed_layer + po.tc), gM, gdx, coefS, Gr, Gs, part, 0, 3, 0};
    // Dropping the gz_lo . x_hi product of the weight-gradient MMAs leaves a relative error of 2^-9 / sqrt(E) times the
    // cancellation of the sum: negligible at 1e5 edges, at the 1e-3 bar below ~1e3.  Under 32768 edges (< 2 tiles per
    // CTA: the extra MMA pass is free) all three products are issued.
    a.wg_lo = g.E < 32768 ? 1 : 0;
    {
        static int glo = -1;
        if (glo < 0) {
#ifdef EGNN_DBG  // accuracy knob (which lo-split products the backward issues): debug builds only
            const char* env = getenv("EGNN_TC_BWD_GLO");
            glo = env ? atoi(env) & 3 : 3;
#else
            glo = 3;
#endif
        }
        a.g_lo = glo;
    }
    int nX;
    egnn_tc_bwd_split(g.E, &a.nM, &nX);
    const int grid = a.nM + nX;
    const bool x3 = precision == EGNN_PREC_BF16X3;
    if (net.Mh == 128) return x3 ? launch_tc_bwd<128, 3>(a, grid, st) : launch_tc_bwd<128, 1>(a, grid, st);
    return x3 ? launch_tc_bwd<64, 3>(a, grid, st) : launch_tc_bwd<64, 1>(a, grid, st);
}
