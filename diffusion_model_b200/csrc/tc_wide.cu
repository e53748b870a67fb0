// Generic ("wide") tcgen05 FORWARD path of the edge MLPs, for the network shapes the fused kernels of tc_edge.cu /
// tc_edge256.cu do not cover: the reference's own parameters.yaml network (K0 = 73, m / x hidden 1024, m_size 256, conditional:
// main.py:102-121), ragged widths, any 2H+1 <= 128.  EquivariantGraphNeuralNetwork.py:55-65 as a chain per layer
//     gather            u = [h_i, h_j, |r|^2]                                   -> U image        [E x K0p]
//     GEMM + SiLU       a1m = SiLU(u Wm1^T + bm1)                               -> A1 image       [E x Mhp]
//     GEMM + gate       m = SiLU(a1m Wm2^T + bm2), g = sigmoid(wa.m + ba)       -> g*m fp32       [E x 256]
//     segment sum       M_i = sum_{e in row i} g*m   (edge order, no atomics)   -> agg rows
//     GEMM + SiLU       a1x = SiLU(u Wx1^T + bx1)                               -> A1 image (reused)
//     GEMM + row dot    phi_e = wx3 . SiLU(a1x Wx2^T + bx2) + bx3               -> phi parts per N tile -> phi [E]
// One GEMM kernel (tc_wide_gemm): a CTA computes [128 edges x 256 outputs] with K streamed in 64-wide slices: both operands
// arrive as ready-made bf16 hi + lo tile images by ONE cp.async.bulk each (32 KB + 64 KB per slice) into a 2-stage ring with
// full (complete_tx) / empty (tcgen05.commit) mbarriers; a producer thread and an MMA-issuer thread run the ring while the
// other warps wait for the accumulator; fp32 accumulation in 256 TMEM columns; bf16x3 = hi.hi + lo.hi + hi.lo.  The epilogue
// adds the bias, applies SiLU and either writes the NEXT GEMM's A operand as a tile image (16-byte chunk stores, 512 contiguous
// bytes per warp), or the gated message rows, or the per-row dot with wx3.  Activations between GEMMs go through L2 / HBM:
// for wide layers that traffic (4 K bytes per edge and GEMM) is small against 2 K N FLOPs per edge.
// The node phase runs the FFMA node kernel on plain per-receiver sums; the backward of such a network runs the FFMA kernels.
#include "tc_bulk.cuh"
#include "tc_tile.cuh"

namespace {

// ---- weights -> tile images ---------------------------------------------------------------------------------------------------
// W [N x K] (nn.Linear: out x in) -> for N tile nt and K slice ks the block (nt * KS + ks): hi chunks (n, j) at j * 4096 + n * 16,
// lo at 32768 + the same; rows >= N and columns >= K are zero
__device__ void wide_pack_matrix(const float* __restrict__ W, int N, int K, int NT, int KS, uint8_t* __restrict__ dst, int t0, int stride) {
    const int per_block = WIDE_TN * (WIDE_TK / 8);
    for (int idx = t0; idx < NT * KS * per_block; idx += stride) {
        const int blk = idx / per_block, rem = idx - blk * per_block;
        const int j = rem / WIDE_TN, nl = rem - j * WIDE_TN;
        const int nt = blk / KS, ks = blk - nt * KS;
        const int n = nt * WIDE_TN + nl, k0 = ks * WIDE_TK + j * 8;
        uint32_t hi[4], lo[4];
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            const int k = k0 + 2 * q;
            const float v0 = (n < N && k < K) ? W[(size_t)n * K + k] : 0.f;
            const float v1 = (n < N && k + 1 < K) ? W[(size_t)n * K + k + 1] : 0.f;
            tc::split2(v0, v1, hi[q], lo[q]);
        }
        uint8_t* b = dst + (size_t)blk * WIDE_B_SLICE + j * (WIDE_TN * 16) + nl * 16;
        *reinterpret_cast<uint4*>(b) = make_uint4(hi[0], hi[1], hi[2], hi[3]);
        *reinterpret_cast<uint4*>(b + WIDE_B_SLICE / 2) = make_uint4(lo[0], lo[1], lo[2], lo[3]);
    }
}

__global__ void __launch_bounds__(256) wide_pack_kernel(egnn_net_t net, const float* __restrict__ params, float* __restrict__ packed) {
    const LayerOff lo = layer_off(net);
    const PackOff po = pack_off(net);
    const WideOff wo = wide_off(net);
    const float* P = params + (size_t)blockIdx.y * lo.total;
    uint8_t* img = reinterpret_cast<uint8_t*>(packed + (size_t)blockIdx.y * po.total + po.tc);
    const int K0 = 2 * net.H + 1, t0 = blockIdx.x * blockDim.x + threadIdx.x, stride = gridDim.x * blockDim.x;
    wide_pack_matrix(P + lo.wm1, net.Mh, K0, wo.ntm1, wo.ks0, img + (size_t)wo.m1 * WIDE_B_SLICE, t0, stride);
    wide_pack_matrix(P + lo.wm2, net.Ms, net.Mh, wo.ntm2, wo.ksm, img + (size_t)wo.m2 * WIDE_B_SLICE, t0, stride);
    wide_pack_matrix(P + lo.wx1, net.Xh, K0, wo.ntx1, wo.ks0, img + (size_t)wo.x1 * WIDE_B_SLICE, t0, stride);
    wide_pack_matrix(P + lo.wx2, net.Xh, net.Xh, wo.ntx2, wo.ksx, img + (size_t)wo.x2 * WIDE_B_SLICE, t0, stride);
}

// ---- gather: one thread per (edge row, 8-column chunk) of the U image; rows >= E are zero -----------------------------------------
__global__ void __launch_bounds__(256) wide_gather_kernel(int E, int mtiles, int H, int KS0, const int32_t* __restrict__ recv,
                                                          const int32_t* __restrict__ send, const float* __restrict__ h,
                                                          const float* __restrict__ x, uint8_t* __restrict__ Uimg) {
    const int nchunk = KS0 * (WIDE_TK / 8);
    const long long total = (long long)mtiles * WIDE_TM * nchunk;
    for (long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += (long long)gridDim.x * blockDim.x) {
        // consecutive threads = consecutive rows of one chunk column: 16-byte stores of a warp are contiguous
        const int row_in_tile = (int)(idx % WIDE_TM);
        const long long rest = idx / WIDE_TM;
        const int j = (int)(rest % nchunk);
        const int t = (int)(rest / nchunk);
        const int e = t * WIDE_TM + row_in_tile;
        float v[8];
#pragma unroll
        for (int q = 0; q < 8; ++q) v[q] = 0.f;
        if (e < E) {
            const int i = recv[e], s = send[e];
            const int k0 = j * 8;
#pragma unroll
            for (int q = 0; q < 8; ++q) {
                const int k = k0 + q;
                if (k < H) v[q] = h[(size_t)i * H + k];
                else if (k < 2 * H) v[q] = h[(size_t)s * H + (k - H)];
                else if (k == 2 * H) {
                    const float rx = x[3 * i] - x[3 * s], ry = x[3 * i + 1] - x[3 * s + 1], rz = x[3 * i + 2] - x[3 * s + 2];
                    const float nrm = sqrtf(rx * rx + ry * ry + rz * rz);  // sqrt then square, as :56
                    v[q] = nrm * nrm;
                }
            }
        }
        uint32_t hi[4], lo[4];
#pragma unroll
        for (int q = 0; q < 4; ++q) tc::split2(v[2 * q], v[2 * q + 1], hi[q], lo[q]);
        uint8_t* b = Uimg + ((size_t)t * KS0 + j / 8) * WIDE_A_SLICE + (j % 8) * (WIDE_TM * 16) + row_in_tile * 16;
        *reinterpret_cast<uint4*>(b) = make_uint4(hi[0], hi[1], hi[2], hi[3]);
        *reinterpret_cast<uint4*>(b + WIDE_A_SLICE / 2) = make_uint4(lo[0], lo[1], lo[2], lo[3]);
    }
}

// ---- M_i = sum over the receiver's edges of g*m, in edge order (fixed order, no atomics): one CTA per receiver ----------------------
__global__ void __launch_bounds__(WIDE_TN) wide_segsum_kernel(int Ms, const int32_t* __restrict__ row_ptr, const float* __restrict__ gm,
                                                              float* __restrict__ agg, int AGGW) {
    const int i = blockIdx.x, c = threadIdx.x;
    if (c >= Ms) return;
    float s = 0.f;
    for (int e = row_ptr[i]; e < row_ptr[i + 1]; ++e) s += gm[(size_t)e * WIDE_TN + c];
    agg[(size_t)i * AGGW + c] = s;
}

// ---- phi_e = bx3 + sum of the per-N-tile parts (tile order) ---------------------------------------------------------------------
__global__ void __launch_bounds__(256) wide_phi_kernel(int E, int NT, const float* __restrict__ part, const float* __restrict__ bx3,
                                                       float* __restrict__ phi) {
    const int e = blockIdx.x * 256 + threadIdx.x;
    if (e >= E) return;
    float s = bx3[0];
    for (int t = 0; t < NT; ++t) s += part[(size_t)t * E + e];
    phi[e] = s;
}

// ---- the tile GEMM ------------------------------------------------------------------------------------------------------------
enum { WIDE_IMG = 0, WIDE_GATE = 1, WIDE_ROWDOT = 2 };
struct WideGemmArgs {
    const uint8_t* A;  // activation image: slice (mt, ks) at (mt * a_ks + ks) * WIDE_A_SLICE
    const uint8_t* B;  // weight image of this matrix: block (nt, ks) at (nt * KS + ks) * WIDE_B_SLICE
    int a_ks, KS, NT, mtiles, E, N;
    const float* bias;  // [N]
    uint8_t* out_img;   // WIDE_IMG: the next GEMM's activation image, out_ks slices per tile
    int out_ks;
    const float *vec, *scalar;  // WIDE_GATE: wa [N], ba;  WIDE_ROWDOT: wx3 [N]
    float* out;                 // WIDE_GATE: g*m [E][WIDE_TN];  WIDE_ROWDOT: parts [NT][E]
};
constexpr int WIDE_THREADS = 512, WIDE_STAGES = 2, WIDE_STAGE_BYTES = WIDE_A_SLICE + WIDE_B_SLICE;
constexpr int WIDE_SMEM = WIDE_STAGES * WIDE_STAGE_BYTES + 1024 /*bias, vec*/ * 2 + 4 * WIDE_TM * 4 /*row partials*/ + 64;

template <int MODE, int NPASS>
__global__ void __launch_bounds__(WIDE_THREADS, 1) tc_wide_gemm(WideGemmArgs a) {
    extern __shared__ __align__(1024) uint8_t smem[];
    float* sBias = reinterpret_cast<float*>(smem + WIDE_STAGES * WIDE_STAGE_BYTES);
    float* sVec = sBias + WIDE_TN;
    float* sDot = sVec + WIDE_TN;  // [4][128]
    uint64_t* bars = reinterpret_cast<uint64_t*>(sDot + 4 * WIDE_TM);  // full[2], empty[2], acc
    uint32_t* slot = reinterpret_cast<uint32_t*>(bars + 5);
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, q = warp & 3, cg = warp >> 2;
    if (warp == 0) tc::tmem_alloc(slot, 256);
    if (tid == 0) {
        for (int b = 0; b < 5; ++b) tc::mbar_init(&bars[b], 1);
        tc::fence_mbar_init();
    }
    tc::fence_before_sync();
    __syncthreads();
    tc::fence_after_sync();
    const uint32_t tmem = *slot;
    const uint32_t sb = tc::smem_u32(smem);
    const int tiles = a.mtiles * a.NT;
    const int nit = (int)blockIdx.x < tiles ? (tiles - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x : 0;
    const int total = nit * a.KS;
    bool ok = true;
    int gl = 0;  // producer: slices requested so far (global over this CTA's tiles)
    pdl_launch_dependents();
    pdl_wait();  // the A image comes from the kernel in front

    for (int it = 0; it < nit; ++it) {
        const int tile = blockIdx.x + it * gridDim.x;
        const int nt = tile % a.NT, mt = tile / a.NT;  // the NT tiles of one edge tile run side by side: its A slices are shared in L2
        if (tid == 0) {
            // this tile's slices plus the first two of the next tile (they land while the epilogue runs)
            int target = (it + 1) * a.KS + WIDE_STAGES;
            if (target > total) target = total;
            for (; gl < target; ++gl) {
                const int s = gl % WIDE_STAGES, u = gl / WIDE_STAGES;
                if (u > 0) ok = tc::mbar_wait(&bars[2 + s], (u - 1) & 1) && ok;
                const int lit = gl / a.KS, ks = gl - lit * a.KS;
                const int lt = blockIdx.x + lit * gridDim.x;
                const int lnt = lt % a.NT, lmt = lt / a.NT;
                const uint32_t full = tc::smem_u32(&bars[s]), dst = sb + s * WIDE_STAGE_BYTES;
                tc::mbar_expect_tx(full, WIDE_STAGE_BYTES);
                tc::bulk_g2s(dst, a.A + ((size_t)lmt * a.a_ks + ks) * WIDE_A_SLICE, WIDE_A_SLICE, full);
                tc::bulk_g2s(dst + WIDE_A_SLICE, a.B + ((size_t)lnt * a.KS + ks) * WIDE_B_SLICE, WIDE_B_SLICE, full);
            }
        } else if (tid == 32) {
            for (int ks = 0; ks < a.KS; ++ks) {
                const int g = it * a.KS + ks, s = g % WIDE_STAGES;
                ok = tc::mbar_wait(&bars[s], (g / WIDE_STAGES) & 1) && ok;
                tc::fence_after_sync();
                const uint32_t base = sb + s * WIDE_STAGE_BYTES;
                const TcOperand opA{base, base + WIDE_A_SLICE / 2, WIDE_TM * 16, false};
                const TcOperand opB{base + WIDE_A_SLICE, base + WIDE_A_SLICE + WIDE_B_SLICE / 2, WIDE_TN * 16, false};
                tc_issue_op<NPASS>(tmem, 0, opA, opB, WIDE_TK, WIDE_TN, ks > 0);
                tc::mma_commit(&bars[2 + s]);
            }
            tc::mma_commit(&bars[4]);
        }
        // bias / vector of this N tile (columns past N are zero: their accumulators are zero too, SiLU(0) = 0)
        if (tid >= 64 && tid < 64 + WIDE_TN) {
            const int c = tid - 64, n = nt * WIDE_TN + c;
            sBias[c] = n < a.N ? a.bias[n] : 0.f;
            if (MODE != WIDE_IMG) sVec[c] = n < a.N ? a.vec[n] : 0.f;
        }
        __syncthreads();
        ok = tc::mbar_wait(&bars[4], it & 1) && ok;
        tc::fence_after_sync();
        __syncwarp();

        const int r = q * 32 + lane, e = mt * WIDE_TM + r;
        const uint32_t taddr = tmem + ((uint32_t)(q * 32) << 16) + cg * 64;
        if (MODE == WIDE_IMG) {
            uint8_t* o = a.out_img + ((size_t)mt * a.out_ks + nt * (WIDE_TN / WIDE_TK) + cg) * WIDE_A_SLICE + r * 16;
#pragma unroll 1
            for (int j = 0; j < 8; ++j) {
                float v[8];
                tc::tmem_ld8(taddr + 8 * j, v);
                uint32_t hi[4], lo[4];
#pragma unroll
                for (int i = 0; i < 4; ++i)
                    tc::split2(tc_silu<true>(v[2 * i] + sBias[cg * 64 + 8 * j + 2 * i]),
                               tc_silu<true>(v[2 * i + 1] + sBias[cg * 64 + 8 * j + 2 * i + 1]), hi[i], lo[i]);
                *reinterpret_cast<uint4*>(o + j * (WIDE_TM * 16)) = make_uint4(hi[0], hi[1], hi[2], hi[3]);
                *reinterpret_cast<uint4*>(o + j * (WIDE_TM * 16) + WIDE_A_SLICE / 2) = make_uint4(lo[0], lo[1], lo[2], lo[3]);
            }
        } else {
            // row dot of the activated outputs with vec: 64 columns per thread, then the four column groups in fixed order
            float dot = 0.f;
#pragma unroll 1
            for (int j = 0; j < 8; ++j) {
                float v[8];
                tc::tmem_ld8(taddr + 8 * j, v);
#pragma unroll
                for (int i = 0; i < 8; ++i) dot = fmaf(tc_silu<true>(v[i] + sBias[cg * 64 + 8 * j + i]), sVec[cg * 64 + 8 * j + i], dot);
            }
            sDot[cg * WIDE_TM + r] = dot;
            __syncthreads();
            const float full = ((sDot[r] + sDot[WIDE_TM + r]) + sDot[2 * WIDE_TM + r]) + sDot[3 * WIDE_TM + r];
            if (MODE == WIDE_ROWDOT) {
                if (cg == 0 && e < a.E) a.out[(size_t)nt * a.E + e] = full;
            } else {
                // second pass over the accumulator (tcgen05.ld is warp-collective: every lane runs it, rows past E only skip the store)
                const float gate = tc_sigmoid<false>(full + a.scalar[0]);
                float* o = a.out + (size_t)e * WIDE_TN + cg * 64;
#pragma unroll 1
                for (int j = 0; j < 8; ++j) {
                    float v[8];
                    tc::tmem_ld8(taddr + 8 * j, v);
#pragma unroll
                    for (int i = 0; i < 8; ++i) v[i] = gate * tc_silu<true>(v[i] + sBias[cg * 64 + 8 * j + i]);
                    if (e < a.E) {
                        *reinterpret_cast<float4*>(o + 8 * j) = make_float4(v[0], v[1], v[2], v[3]);
                        *reinterpret_cast<float4*>(o + 8 * j + 4) = make_float4(v[4], v[5], v[6], v[7]);
                    }
                }
            }
        }
        tc::fence_before_sync();
        __syncthreads();  // the accumulator, sBias / sVec / sDot are free for the next tile
        tc::fence_after_sync();
    }
    if (!ok) asm volatile("trap;");  // an mbarrier wait timed out: fail loudly instead of returning garbage
    if (warp == 0) tc::tmem_dealloc(tmem, 256);
}

template <int MODE>
int launch_wide(const WideGemmArgs& a, int precision, cudaStream_t st, const char* what) {
    const int tiles = a.mtiles * a.NT;
    // persistent: one CTA per SM at most; the tile count is rounded so that every CTA runs the same number of tiles when it can
    int grid = tiles < kB200SMs ? tiles : kB200SMs;
    if (tiles > kB200SMs) {
        const int per = wide_cdiv(tiles, kB200SMs);
        grid = wide_cdiv(tiles, per);
    }
    egnn_prof_begin(EGNN_K_TC_EDGE_FWD, st);
    if (precision == EGNN_PREC_BF16X3) {
        EGNN_SMEM_ATTR((tc_wide_gemm<MODE, 3>), WIDE_SMEM);
        tc_wide_gemm<MODE, 3><<<grid, WIDE_THREADS, WIDE_SMEM, st>>>(a);
    } else {
        EGNN_SMEM_ATTR((tc_wide_gemm<MODE, 1>), WIDE_SMEM);
        tc_wide_gemm<MODE, 1><<<grid, WIDE_THREADS, WIDE_SMEM, st>>>(a);
    }
    egnn_prof_end(st);
    EGNN_CHECK_LAUNCH(what);
    return EGNN_OK;
}

}  // namespace

int egnn_wide_pack(const egnn_net_t& net, const float* params, float* packed, cudaStream_t st) {
    if (!wide_fwd_supported(net)) return EGNN_OK;
    wide_pack_kernel<<<dim3(148, net.L), 256, 0, st>>>(net, params, packed);
    EGNN_CHECK_LAUNCH("wide_pack_kernel");
    return EGNN_OK;
}

// edge phase of one layer: leaves per-receiver message sums in agg[i * (Ms + 4) + c] and phi_e in phi_edge[e]
int egnn_wide_edge_forward(const egnn_net_t& net, const egnn_graph_t& g, const float* params_layer, const float* packed_layer,
                           const float* h, const float* x, float* agg, float* phi_edge, float* scratch, int precision,
                           cudaStream_t st) {
    EGNN_REQUIRE(wide_fwd_supported(net), "egnn_wide_edge_forward: network outside the wide path's limits");
    const LayerOff lo = layer_off(net);
    const PackOff po = pack_off(net);
    const WideOff wo = wide_off(net);
    const WideWs ww = wide_ws(net, g.E);
    uint8_t* base = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(scratch) + 1023) / 1024 * 1024);
    uint8_t* U = base + ww.u;
    uint8_t* A1 = base + ww.a1;
    float* gm = reinterpret_cast<float*>(base + ww.gm);
    float* phip = reinterpret_cast<float*>(base + ww.phip);
    const uint8_t* img = reinterpret_cast<const uint8_t*>(packed_layer + po.tc);
    const float* P = params_layer;
    {
        const long long total = (long long)ww.mtiles * WIDE_TM * wo.ks0 * (WIDE_TK / 8);
        const int blocks = (int)((total + 255) / 256 < 148 * 16 ? (total + 255) / 256 : 148 * 16);
        egnn_prof_begin(EGNN_K_TC_EDGE_FWD, st);
        wide_gather_kernel<<<blocks, 256, 0, st>>>(g.E, ww.mtiles, net.H, wo.ks0, g.recv, g.send, h, x, U);
        egnn_prof_end(st);
        EGNN_CHECK_LAUNCH("wide_gather_kernel");
    }
    WideGemmArgs a{};
    a.mtiles = ww.mtiles;
    a.E = g.E;
    // message branch
    a.A = U; a.a_ks = wo.ks0; a.KS = wo.ks0; a.B = img + (size_t)wo.m1 * WIDE_B_SLICE; a.NT = wo.ntm1; a.N = net.Mh;
    a.bias = P + lo.bm1; a.out_img = A1; a.out_ks = ww.a1_ks;
    if (int rc = launch_wide<WIDE_IMG>(a, precision, st, "tc_wide_gemm m1")) return rc;
    a.A = A1; a.a_ks = ww.a1_ks; a.KS = wo.ksm; a.B = img + (size_t)wo.m2 * WIDE_B_SLICE; a.NT = wo.ntm2; a.N = net.Ms;
    a.bias = P + lo.bm2; a.vec = P + lo.wa; a.scalar = P + lo.ba; a.out = gm;
    if (int rc = launch_wide<WIDE_GATE>(a, precision, st, "tc_wide_gemm m2")) return rc;
    egnn_prof_begin(EGNN_K_TC_EDGE_FWD, st);
    wide_segsum_kernel<<<g.N, WIDE_TN, 0, st>>>(net.Ms, g.row_ptr, gm, agg, net.Ms + 4);
    egnn_prof_end(st);
    EGNN_CHECK_LAUNCH("wide_segsum_kernel");
    // coordinate branch
    a.A = U; a.a_ks = wo.ks0; a.KS = wo.ks0; a.B = img + (size_t)wo.x1 * WIDE_B_SLICE; a.NT = wo.ntx1; a.N = net.Xh;
    a.bias = P + lo.bx1; a.out_img = A1; a.out_ks = ww.a1_ks;
    if (int rc = launch_wide<WIDE_IMG>(a, precision, st, "tc_wide_gemm x1")) return rc;
    a.A = A1; a.a_ks = ww.a1_ks; a.KS = wo.ksx; a.B = img + (size_t)wo.x2 * WIDE_B_SLICE; a.NT = wo.ntx2; a.N = net.Xh;
    a.bias = P + lo.bx2; a.vec = P + lo.wx3; a.scalar = nullptr; a.out = phip;
    if (int rc = launch_wide<WIDE_ROWDOT>(a, precision, st, "tc_wide_gemm x2")) return rc;
    egnn_prof_begin(EGNN_K_TC_EDGE_FWD, st);
    wide_phi_kernel<<<(g.E + 255) / 256, 256, 0, st>>>(g.E, wo.ntx2, phip, P + lo.bx3, phi_edge);
    egnn_prof_end(st);
    EGNN_CHECK_LAUNCH("wide_phi_kernel");
    return EGNN_OK;
}
