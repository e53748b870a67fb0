// tcgen05 path of the node backward (mlp_h): one CTA per 128-node tile, 4*D threads (thread = node row x 32 columns).
//   z1 = [h, M, 0.., 1] . Wh1^T                      GEMM1, K = D + 16 (the constant-1 column makes db1 an MMA column)
//   s1 = SiLU(z1 + b1) -> tile S;   dWh2^T = s1^T . gh'          (tile S viewed MN-major, K = 128 nodes)
//   gz1 = SiLU'(z1 + b1) * (gh' . Wh2) -> tile S (reused);  g[h, M] = gz1 . Wh1   (the Wh1 image viewed MN-major)
//   [dWh1 | db1] = gz1^T . [h, M, 0.., 1]
// Every CTA writes its weight-gradient block to its own partial row (no accumulation, no zero-fill needed).
// Reference: EGCL.forward's node update, EquivariantGraphNeuralNetwork.py:26-30, 69 (adjoint of it).
#include "tc_tile.cuh"

namespace {

template <int D>
struct TcNodeLayout {
    static constexpr int NT = 4 * D, NG = D / 32, KP = D + 16, NCH = KP / 8;
    static constexpr int s_hi = 0, s_lo = 128 * D * 2;                 // S, later gz1
    static constexpr int a_hi = 2 * 128 * D * 2, a_lo = a_hi + NCH * 2048;  // [128 x KP]
    static constexpr int w_hi = a_lo + NCH * 2048, w_lo = w_hi + NCH * D * 16;  // Wh1 image [D x KP]
    static constexpr int gh_hi = w_lo + NCH * D * 16, gh_lo = gh_hi + 4096;     // gh' [128 x 16]
    static constexpr int vec = gh_lo + 4096;          // b1 [D] | Wh2 [8][D]
    static constexpr int sGH = vec + 9 * D * 4;       // float [128][8]
    static constexpr int red = sGH + 4096;            // float [NT/32][8]
    static constexpr int bars = red + (NT / 32) * 32, slot = bars + 32;
    static constexpr int TOTAL = slot + 16;
    static_assert(TOTAL <= 227 * 1024, "tc_node_bwd shared memory");
};

struct TcNodeBwdArgs {
    egnn_net_t net;
    int N;
    const float *h, *msum, *gh_out;
    const float* P;        // this layer's fp32 parameters
    const uint8_t* nblob;  // Wh1 image (hi | lo)
    float *gh_node, *gM;
    float* part;           // [gridDim.x][layer param count]
};

template <int D, int NPASS>
__global__ void __launch_bounds__(4 * D, 1) tc_node_bwd(TcNodeBwdArgs a) {
    using L = TcNodeLayout<D>;
    constexpr bool FAST = true;  // one-MUFU SiLU / sigmoid (tanh.approx) in both tensor-core modes, see tc_tile.cuh
    constexpr bool LO = NPASS == 3;
    constexpr int NT = L::NT, NG = L::NG, KP = L::KP;
    constexpr int cZ = 0, cG = D, cW1 = 2 * D + 32, cW2T = 3 * D + 48;  // TMEM columns (512 allocated)
    constexpr uint32_t TS = 128 * 16;
    extern __shared__ __align__(1024) uint8_t smem[];
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int q = warp & 3, cg = warp >> 2;
    const int r = q * 32 + lane, c0 = cg * 32;
    const int H = a.net.H, KA = H + D;
    const LayerOff lo = layer_off(a.net);
    float* vec = reinterpret_cast<float*>(smem + L::vec);
    float* sGH = reinterpret_cast<float*>(smem + L::sGH);
    float* red = reinterpret_cast<float*>(smem + L::red);
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + L::bars);
    uint32_t* slot = reinterpret_cast<uint32_t*>(smem + L::slot);
    const int n = blockIdx.x * 128 + r;
    const bool valid = n < a.N;

    for (int o = tid * 16; o < 2 * L::NCH * D * 16; o += NT * 16)
        *reinterpret_cast<uint4*>(smem + L::w_hi + o) = __ldg(reinterpret_cast<const uint4*>(a.nblob + o));
    for (int c = tid; c < D; c += NT) vec[c] = a.P[lo.bh1 + c];
    for (int c = tid; c < 8 * D; c += NT) vec[D + c] = c < H * D ? a.P[lo.wh2 + c] : 0.f;
    if (warp == 0) tc::tmem_alloc(slot, 512);
    if (tid == 0) {
        for (int b = 0; b < 3; ++b) tc::mbar_init(&bars[b], 1);
        tc::fence_mbar_init();
    }
    // ---- the tile's message sums are one contiguous [rows x D] block: coalesced copy into shared memory (the S tile
    //      region, free until GEMM1 is done), column index XOR-swizzled so that the per-row reads below are conflict-free
    float* stageF = reinterpret_cast<float*>(smem + L::s_hi);
    {
        const int n0 = blockIdx.x * 128, rows = min(128, a.N - n0);
        const float* src = a.msum + (size_t)n0 * D;
        for (int idx = tid; idx < rows * D; idx += NT) {
            const int rr = idx / D, m = idx % D;
            stageF[rr * D + (m ^ (rr & 31))] = __ldg(src + idx);
        }
    }
    __syncthreads();
    // ---- A tile = [h | M | 0 .. | 1] of this thread's row, 32 columns (+ the 16 trailing columns by column group 0)
    {
        float v[32];
#pragma unroll
        for (int k = 0; k < 32; ++k) {
            const int c = c0 + k;
            v[k] = !valid ? 0.f : (c < H ? a.h[(size_t)n * H + c] : stageF[r * D + ((c - H) ^ (r & 31))]);
        }
        tc_split_store32(smem + L::a_hi, smem + L::a_lo, r, c0, v, true);
        if (cg == 0) {
            float t[16];
#pragma unroll
            for (int k = 0; k < 16; ++k) {
                const int c = D + k;
                t[k] = !valid ? 0.f : (c < KA ? stageF[r * D + ((c - H) ^ (r & 31))] : (c == KP - 1 ? 1.f : 0.f));
            }
#pragma unroll
            for (int j = 0; j < 2; ++j) {
                uint32_t hi[4], lw[4];
#pragma unroll
                for (int p = 0; p < 4; ++p) tc::split2(t[8 * j + 2 * p], t[8 * j + 2 * p + 1], hi[p], lw[p]);
                *reinterpret_cast<uint4*>(smem + L::a_hi + tc::chunk_off(128, r, D / 8 + j)) = make_uint4(hi[0], hi[1], hi[2], hi[3]);
                *reinterpret_cast<uint4*>(smem + L::a_lo + tc::chunk_off(128, r, D / 8 + j)) = make_uint4(lw[0], lw[1], lw[2], lw[3]);
            }
        }
        if (cg == NG - 1) {  // gh' row: fp32 copy for the CUDA cores, bf16 tile for the wgrad MMA
            float t[16];
#pragma unroll
            for (int k = 0; k < 16; ++k) t[k] = (valid && k < H) ? a.gh_out[(size_t)n * H + k] : 0.f;
#pragma unroll
            for (int k = 0; k < 8; ++k) sGH[r * 8 + k] = t[k];
#pragma unroll
            for (int j = 0; j < 2; ++j) {
                uint32_t hi[4], lw[4];
#pragma unroll
                for (int p = 0; p < 4; ++p) tc::split2(t[8 * j + 2 * p], t[8 * j + 2 * p + 1], hi[p], lw[p]);
                *reinterpret_cast<uint4*>(smem + L::gh_hi + tc::chunk_off(128, r, j)) = make_uint4(hi[0], hi[1], hi[2], hi[3]);
                *reinterpret_cast<uint4*>(smem + L::gh_lo + tc::chunk_off(128, r, j)) = make_uint4(lw[0], lw[1], lw[2], lw[3]);
            }
            // db2[o] = sum over nodes of gh'[n][o]: warp sums now, cross-warp sum at the end
#pragma unroll
            for (int k = 0; k < 8; ++k) {
                const float s = warp_sum(t[k]);
                if (lane == 0) red[q * 8 + k] = s;
            }
        }
    }
    tc::fence_async_smem();
    tc::fence_before_sync();
    __syncthreads();
    tc::fence_after_sync();
    const uint32_t tmem = *slot;
    const uint32_t trow = tmem + ((uint32_t)(q * 32) << 16);
    const uint32_t sb = tc::smem_u32(smem);
    const TcOperand opA{sb + L::a_hi, sb + L::a_lo, TS, false}, opAmn{sb + L::a_hi, sb + L::a_lo, TS, true};
    const TcOperand opW{sb + L::w_hi, sb + L::w_lo, D * 16, false}, opWmn{sb + L::w_hi, sb + L::w_lo, D * 16, true};
    const TcOperand opS{sb + L::s_hi, sb + L::s_lo, TS, false}, opSmn{sb + L::s_hi, sb + L::s_lo, TS, true};
    const TcOperand opGHmn{sb + L::gh_hi, sb + L::gh_lo, TS, true};
    bool ok = true;
    if (warp == NT / 32 - 1 && tc::elect_one()) {
        tc_issue_op<NPASS>(tmem, cZ, opA, opW, KP, D, false);  // z1
        tc::mma_commit(&bars[0]);
    }
    ok = tc::mbar_wait(&bars[0], 0) && ok;
    tc::fence_after_sync();
    float gz[32];
    {
        float v[32], act[32];
        tc::tmem_ld32(trow + cZ + c0, v);
        float gh[8];
#pragma unroll
        for (int k = 0; k < 8; ++k) gh[k] = sGH[r * 8 + k];
#pragma unroll
        for (int i = 0; i < 32; ++i) {
            float dact;
            tc_silu_and_grad(v[i] + vec[c0 + i], act[i], dact);
            float t = 0.f;
#pragma unroll
            for (int o = 0; o < 8; ++o)
                if (o < H) t = fmaf(gh[o], vec[D + o * D + c0 + i], t);
            gz[i] = t * dact;
        }
        tc_split_store32(smem + L::s_hi, smem + L::s_lo, r, c0, act, LO);
    }
    tc::fence_async_smem();
    tc::fence_before_sync();
    __syncthreads();
    if (warp == NT / 32 - 1 && tc::elect_one()) {
        tc::fence_after_sync();
        tc_issue_op<NPASS>(tmem, cW2T, opSmn, opGHmn, 128, 16, false);  // dWh2^T [k][o] = s1^T . gh'
        tc::mma_commit(&bars[1]);
    }
    ok = tc::mbar_wait(&bars[1], 0) && ok;
    tc::fence_after_sync();
    tc_split_store32(smem + L::s_hi, smem + L::s_lo, r, c0, gz, LO);  // tile S now holds gz1
    tc::fence_async_smem();
    tc::fence_before_sync();
    __syncthreads();
    if (warp == NT / 32 - 1 && tc::elect_one()) {
        tc::fence_after_sync();
        tc_issue_op<NPASS>(tmem, cG, opS, opWmn, D, KP, false);      // g[h, M] = gz1 . Wh1
        tc_issue_op<NPASS>(tmem, cW1, opSmn, opAmn, 128, KP, false);  // [dWh1 | db1] = gz1^T . [h, M, 0.., 1]
        tc::mma_commit(&bars[2]);
    }
    ok = tc::mbar_wait(&bars[2], 0) && ok;
    if (!ok) asm volatile("trap;");
    tc::fence_after_sync();
    float* part = a.part + (size_t)blockIdx.x * lo.total;
    // ---- outputs: every thread owns one row, so direct global stores would hit 32 different lines per instruction.
    //      Stage through shared memory (all MMAs are done: the A tile and the weight image are dead) with odd row
    //      strides (conflict-free for row-per-lane writes), then write the contiguous global blocks coalesced.
    float* stageG = reinterpret_cast<float*>(smem + L::a_hi);  // [128][D+1]   g[M] of the tile's nodes
    float* stageP = reinterpret_cast<float*>(smem + L::w_hi);  // [D][KA]      dWh1 == its global layout
    {
        float v[32];
        tc::tmem_ld32(trow + cG + c0, v);
#pragma unroll
        for (int k = 0; k < 32; ++k) {
            const int c = c0 + k;
            if (c < H) {
                if (valid) a.gh_node[(size_t)n * H + c] = v[k];
            } else {
                stageG[r * (D + 1) + (c - H)] = v[k];
            }
        }
        tc::tmem_ld32(trow + cW1 + c0, v);
        if (r < D) {
#pragma unroll
            for (int k = 0; k < 32; ++k) stageP[r * KA + c0 + k] = v[k];
        }
    }
    if (cg == 0) {
        float v[32];
        tc::tmem_ld32(trow + cG + D, v);
#pragma unroll
        for (int k = 0; k < 16; ++k)
            if (D + k < KA) stageG[r * (D + 1) + (D + k - H)] = v[k];
        tc::tmem_ld32(trow + cW1 + D, v);
        if (r < D) {
#pragma unroll
            for (int k = 0; k < 16; ++k)
                if (D + k < KA) stageP[r * KA + D + k] = v[k];
            part[lo.bh1 + r] = v[15];
        }
    }
    __syncthreads();
    {
        const int n0 = blockIdx.x * 128, rows = min(128, a.N - n0);
        float* dst = a.gM + (size_t)n0 * D;
        for (int idx = tid; idx < rows * D; idx += NT) dst[idx] = stageG[(idx / D) * (D + 1) + idx % D];
        for (int idx = tid; idx < D * KA; idx += NT) part[lo.wh1 + idx] = stageP[idx];
    }
    if (cg == NG - 1) {
        float v[32];
        tc::tmem_ld32(trow + cW2T, v);
        if (r < D) {
#pragma unroll
            for (int o = 0; o < 8; ++o)
                if (o < H) part[lo.wh2 + o * D + r] = v[o];
        }
    }
    if (tid < H) part[lo.bh2 + tid] = red[tid] + red[8 + tid] + red[16 + tid] + red[24 + tid];
    tc::fence_before_sync();
    __syncthreads();
    if (warp == 0) tc::tmem_dealloc(tmem, 512);
}

// ---------------------------------------------------------------------------------------------------------------------
// tcgen05 path of the node FORWARD: one CTA per 128-node tile, 4*D threads.
//   M_i    = per-receiver message sum: the edge kernel's agg row or, for a node cut by 32-edge segment boundaries (the
//            usual case), tail partial of its first segment + head partials of the following ones, in segment order
//   dxu_i  = sum_e r_e phi_e,  D_i = sum_e |r_e|^2      (over the receiver's edges, fixed order)
//   h'     = Wh2 . SiLU(Wh1 [h, M] + b1) + b2           GEMM1 on the tensor cores (K = D + 16), the H-wide second
//                                                       layer as per-thread partial dots over 32 columns
// Reference: EGCL.forward's node update, EquivariantGraphNeuralNetwork.py:26-30, 66-69 (no residual).
// All global reads of the tile's sums are coalesced (a warp reads 32 consecutive columns of one row); the loads of
// several rows are issued together so that the dependent chains (row_ptr -> partial rows, send -> x) overlap.
// ---------------------------------------------------------------------------------------------------------------------
struct TcNodeFwdArgs {
    egnn_net_t net;
    int N, TE;
    const int32_t *row_ptr, *send;
    const float *h, *x;
    const float* P;        // this layer's fp32 parameters
    const uint8_t* nblob;  // Wh1 image (hi | lo)
    const float *agg, *part, *phi_edge;
    float *h_out, *msum, *dxu, *Dn;
};

template <int D, int NPASS>
__global__ void __launch_bounds__(4 * D, 1) tc_node_fwd(TcNodeFwdArgs a) {
    using L = TcNodeLayout<D>;
    constexpr bool LO = NPASS == 3;
    constexpr int NT = L::NT, NG = L::NG, KP = L::KP;
    constexpr int LPN = NT / 128;  // lanes per node of the dxu / D sums
    constexpr uint32_t TS = 128 * 16;
    extern __shared__ __align__(1024) uint8_t smem[];
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int q = warp & 3, cg = warp >> 2;
    const int r = q * 32 + lane, c0 = cg * 32;
    constexpr int TE = 32;  // partial-segment length of the tcgen05 edge kernel (part_tile_for); a shift, not a division
    const int H = a.net.H, KA = H + D, AGGW = D + 4;
    const LayerOff lo = layer_off(a.net);
    float* vec = reinterpret_cast<float*>(smem + L::vec);    // b1 [D] | Wh2 [8][D]
    float* sPart = reinterpret_cast<float*>(smem + L::s_hi);  // float [NG][128][8] partial dots (after the A tile is built)
    int* sRow = reinterpret_cast<int*>(smem + L::sGH);        // int [129] row_ptr of the tile's nodes
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + L::bars);
    uint32_t* slot = reinterpret_cast<uint32_t*>(smem + L::slot);
    const int n0 = blockIdx.x * 128, rows = min(128, a.N - n0);
    const int n = n0 + r;
    const bool valid = r < rows;

    pdl_launch_dependents();  // common.cuh: PDL -- the x-update kernel behind this one may be scheduled now
    // weight image: asynchronous copy, waited for just before the MMA
    {
        const uint32_t dst = tc::smem_u32(smem + L::w_hi);
        for (int o = tid * 16; o < 2 * L::NCH * D * 16; o += NT * 16)
            asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst + o), "l"(a.nblob + o) : "memory");
        asm volatile("cp.async.commit_group;" ::: "memory");
    }
    for (int c = tid; c < D; c += NT) vec[c] = a.P[lo.bh1 + c];
    for (int c = tid; c < 8 * D; c += NT) vec[D + c] = c < H * D ? a.P[lo.wh2 + c] : 0.f;
    if (tid <= 128) sRow[tid] = a.row_ptr[min(n0 + tid, a.N)];
    if (warp == 0) tc::tmem_alloc(slot, D);
    if (tid == 0) {
        tc::mbar_init(&bars[0], 1);
        tc::fence_mbar_init();
    }
    __syncthreads();
    pdl_wait();  // the prologue read weights, fp32 parameters and row_ptr only; agg / part / phi come from the edge kernel

    // ---- message sums of the tile: D/4 lanes own one row (a float4 each: rows of agg / part / msum are 16-byte
    //      aligned).  Per row up to three partial rows are read under predicates (agg row or tail of the first segment,
    //      heads of the next two) so that the loads of UNR rows are in flight together; longer runs take the loop.
    float* stageF = reinterpret_cast<float*>(smem + L::s_hi);
    {
        constexpr int UNR = 4;
        constexpr int LPR = D / 4, RPW = 32 / LPR;           // lanes per row, rows per warp and pass
        constexpr int RPP = (NT / 32) * RPW;                  // rows per CTA pass
        const int m4 = (lane % LPR) * 4;
        const int rbase = warp * RPW + lane / LPR;
        const float4 zero4 = make_float4(0.f, 0.f, 0.f, 0.f);
        for (int rb0 = 0; rb0 < rows; rb0 += UNR * RPP) {
            float4 p0[UNR], p1[UNR], p2[UNR];
#pragma unroll
            for (int u = 0; u < UNR; ++u) {
                const int rr = rb0 + u * RPP + rbase;
                p0[u] = p1[u] = p2[u] = zero4;
                if (rr < rows) {
                    const int ea = sRow[rr], eb = sRow[rr + 1];
                    if (ea < eb) {
                        const int t0 = ea / TE, t1 = (eb - 1) / TE;
                        const float* src = t0 == t1 ? a.agg + (size_t)(n0 + rr) * AGGW : a.part + ((size_t)t0 * 2 + 1) * AGGW;
                        p0[u] = *reinterpret_cast<const float4*>(src + m4);
                        if (t0 + 1 <= t1) p1[u] = *reinterpret_cast<const float4*>(a.part + ((size_t)(t0 + 1) * 2) * AGGW + m4);
                        if (t0 + 2 <= t1) p2[u] = *reinterpret_cast<const float4*>(a.part + ((size_t)(t0 + 2) * 2) * AGGW + m4);
                    }
                }
            }
#pragma unroll
            for (int u = 0; u < UNR; ++u) {
                const int rr = rb0 + u * RPP + rbase;
                if (rr < rows) {
                    float4 sum = make_float4((p0[u].x + p1[u].x) + p2[u].x, (p0[u].y + p1[u].y) + p2[u].y,
                                             (p0[u].z + p1[u].z) + p2[u].z, (p0[u].w + p1[u].w) + p2[u].w);
                    const int ea = sRow[rr], eb = sRow[rr + 1];
                    if (ea < eb) {
                        const int t1 = (eb - 1) / TE;
                        for (int t = ea / TE + 3; t <= t1; ++t) {
                            const float4 ph = *reinterpret_cast<const float4*>(a.part + ((size_t)t * 2) * AGGW + m4);
                            sum.x += ph.x;
                            sum.y += ph.y;
                            sum.z += ph.z;
                            sum.w += ph.w;
                        }
                    }
                    *reinterpret_cast<float4*>(a.msum + (size_t)(n0 + rr) * D + m4) = sum;
                    const int sw = rr & 31;
                    stageF[rr * D + ((m4 + 0) ^ sw)] = sum.x;
                    stageF[rr * D + ((m4 + 1) ^ sw)] = sum.y;
                    stageF[rr * D + ((m4 + 2) ^ sw)] = sum.z;
                    stageF[rr * D + ((m4 + 3) ^ sw)] = sum.w;
                }
            }
        }
    }
    // ---- coordinate sums: LPN lanes per node walk the receiver's edges with stride LPN, four edges in flight per lane
    {
        const int rr = tid / LPN, sub = tid % LPN;
        float sx = 0.f, sy = 0.f, sz = 0.f, sd = 0.f;
        if (rr < rows) {
            const int i = n0 + rr;
            const float xi = a.x[3 * i], yi = a.x[3 * i + 1], zi = a.x[3 * i + 2];
            const int eb = sRow[rr + 1];
            for (int e = sRow[rr] + sub; e < eb; e += 4 * LPN) {
                int j[4];
                float ph[4], xs[4][3];
#pragma unroll
                for (int u = 0; u < 4; ++u) {
                    const int ee = e + u * LPN;
                    j[u] = ee < eb ? a.send[ee] : i;
                    ph[u] = ee < eb ? a.phi_edge[ee] : 0.f;
                }
#pragma unroll
                for (int u = 0; u < 4; ++u) {
                    xs[u][0] = a.x[3 * j[u]];
                    xs[u][1] = a.x[3 * j[u] + 1];
                    xs[u][2] = a.x[3 * j[u] + 2];
                }
#pragma unroll
                for (int u = 0; u < 4; ++u) {  // padding lanes read x_i itself: r = 0 and phi = 0 add exact zeros
                    const float rx = xi - xs[u][0], ry = yi - xs[u][1], rz = zi - xs[u][2];
                    sx = fmaf(rx, ph[u], sx);
                    sy = fmaf(ry, ph[u], sy);
                    sz = fmaf(rz, ph[u], sz);
                    sd += rx * rx + ry * ry + rz * rz;
                }
            }
        }
#pragma unroll
        for (int o = LPN / 2; o > 0; o >>= 1) {
            sx += __shfl_xor_sync(0xffffffffu, sx, o);
            sy += __shfl_xor_sync(0xffffffffu, sy, o);
            sz += __shfl_xor_sync(0xffffffffu, sz, o);
            sd += __shfl_xor_sync(0xffffffffu, sd, o);
        }
        if (rr < rows && sub == 0) {
            const int i = n0 + rr;
            a.dxu[(size_t)i * 3 + 0] = sx;
            a.dxu[(size_t)i * 3 + 1] = sy;
            a.dxu[(size_t)i * 3 + 2] = sz;
            a.Dn[i] = sd;
        }
    }
    __syncthreads();
    // ---- A tile = [h | M | 0 .. | 1] of this thread's row, 32 columns (+ the 16 trailing columns by column group 0)
    {
        float v[32];
#pragma unroll
        for (int k = 0; k < 32; ++k) {
            const int c = c0 + k;
            v[k] = !valid ? 0.f : (c < H ? a.h[(size_t)n * H + c] : stageF[r * D + ((c - H) ^ (r & 31))]);
        }
        tc_split_store32(smem + L::a_hi, smem + L::a_lo, r, c0, v, true);
        if (cg == 0) {
            float t[16];
#pragma unroll
            for (int k = 0; k < 16; ++k) {
                const int c = D + k;
                t[k] = !valid ? 0.f : (c < KA ? stageF[r * D + ((c - H) ^ (r & 31))] : 0.f);
            }
#pragma unroll
            for (int j = 0; j < 2; ++j) {
                uint32_t hi[4], lw[4];
#pragma unroll
                for (int p = 0; p < 4; ++p) tc::split2(t[8 * j + 2 * p], t[8 * j + 2 * p + 1], hi[p], lw[p]);
                *reinterpret_cast<uint4*>(smem + L::a_hi + tc::chunk_off(128, r, D / 8 + j)) = make_uint4(hi[0], hi[1], hi[2], hi[3]);
                *reinterpret_cast<uint4*>(smem + L::a_lo + tc::chunk_off(128, r, D / 8 + j)) = make_uint4(lw[0], lw[1], lw[2], lw[3]);
            }
        }
    }
    asm volatile("cp.async.wait_group 0;" ::: "memory");
    tc::fence_async_smem();
    tc::fence_before_sync();
    __syncthreads();
    tc::fence_after_sync();
    const uint32_t tmem = *slot;
    const uint32_t trow = tmem + ((uint32_t)(q * 32) << 16);
    const uint32_t sb = tc::smem_u32(smem);
    const TcOperand opA{sb + L::a_hi, sb + L::a_lo, TS, false};
    const TcOperand opW{sb + L::w_hi, sb + L::w_lo, D * 16, false};
    if (warp == NT / 32 - 1 && tc::elect_one()) {
        tc_issue_op<NPASS>(tmem, 0, opA, opW, KP, D, false);  // z1
        tc::mma_commit(&bars[0]);
    }
    if (!tc::mbar_wait(&bars[0], 0)) asm volatile("trap;");
    tc::fence_after_sync();
    // ---- s1 = SiLU(z1 + b1); partial dots with the H rows of Wh2 over this thread's 32 columns
    {
        float v[32], pd[8];
        tc::tmem_ld32(trow + c0, v);
#pragma unroll
        for (int o = 0; o < 8; ++o) pd[o] = 0.f;
#pragma unroll
        for (int i = 0; i < 32; ++i) {
            const float s1 = tc_silu<true>(v[i] + vec[c0 + i]);
#pragma unroll
            for (int o = 0; o < 8; ++o)
                if (o < H) pd[o] = fmaf(vec[D + o * D + c0 + i], s1, pd[o]);
        }
#pragma unroll
        for (int o = 0; o < 8; ++o) sPart[(cg * 128 + r) * 8 + o] = pd[o];
    }
    tc::fence_before_sync();
    __syncthreads();
    for (int idx = tid; idx < rows * H; idx += NT) {  // h_out block of the tile is contiguous: coalesced
        const int rr = idx / H, o = idx - rr * H;
        float s = 0.f;
#pragma unroll
        for (int k = 0; k < NG; ++k) s += sPart[(k * 128 + rr) * 8 + o];
        a.h_out[(size_t)n0 * H + idx] = s + a.P[lo.bh2 + o];
    }
    if (warp == 0) tc::tmem_dealloc(tmem, D);
}

template <int D, int NPASS>
int launch_tc_node_fwd(const TcNodeFwdArgs& a, int grid, cudaStream_t st) {
    constexpr int SMEM = TcNodeLayout<D>::TOTAL;
    EGNN_SMEM_ATTR((tc_node_fwd<D, NPASS>), SMEM);
    egnn_prof_begin(EGNN_K_NODE_FWD, st);
    EGNN_CUDA(egnn_launch(tc_node_fwd<D, NPASS>, dim3(grid), dim3(4 * D), SMEM, st, true, a));
    egnn_prof_end(st);
    EGNN_CHECK_LAUNCH("tc_node_fwd");
    return EGNN_OK;
}

template <int D, int NPASS>
int launch_tc_node_bwd(const TcNodeBwdArgs& a, int grid, cudaStream_t st) {
    constexpr int SMEM = TcNodeLayout<D>::TOTAL;
    EGNN_SMEM_ATTR((tc_node_bwd<D, NPASS>), SMEM);
    egnn_prof_begin(EGNN_K_NODE_BWD, st);
    tc_node_bwd<D, NPASS><<<grid, 4 * D, SMEM, st>>>(a);
    egnn_prof_end(st);
    EGNN_CHECK_LAUNCH("tc_node_bwd");
    return EGNN_OK;
}

}  // namespace

int egnn_tc_node_bwd_grid(int64_t N) { return (int)((N + 127) / 128); }

// node phase of one layer's backward on the tensor cores; partial rows [grid][layer param count], mlp_h block fully written
int egnn_tc_node_backward(const egnn_net_t& net, int N, const float* params_layer, const float* packed_layer, const float* h,
                          const float* msum, const float* gh_out, float* gh_node, float* gM, float* part, int precision,
                          cudaStream_t st) {
    EGNN_REQUIRE(tc_node_supported(net), "tcgen05 node path needs Hh = Ms = d in {64,128}");
    const PackOff po = pack_off(net);
    const uint8_t* blob = reinterpret_cast<const uint8_t*>(packed_layer + po.tc);
    TcNodeBwdArgs a{net, N, h, msum, gh_out, params_layer, blob + (tc_blob_bytes(net.Mh) + 15) / 16 * 16, gh_node, gM, part};
    const int grid = egnn_tc_node_bwd_grid(N);
    const bool x3 = precision == EGNN_PREC_BF16X3;
    if (net.Mh == 128) return x3 ? launch_tc_node_bwd<128, 3>(a, grid, st) : launch_tc_node_bwd<128, 1>(a, grid, st);
    return x3 ? launch_tc_node_bwd<64, 3>(a, grid, st) : launch_tc_node_bwd<64, 1>(a, grid, st);
}

// node phase of one layer's forward on the tensor cores (the edge phase must have run on the tcgen05 path: 32-edge
// partial segments, phi per edge)
int egnn_tc_node_forward(const egnn_net_t& net, const egnn_graph_t& g, const float* params_layer, const float* packed_layer,
                         const float* h, const float* x, const float* agg, const float* part, const float* phi_edge,
                         float* h_out, float* msum, float* dxu, float* Dn, int precision, cudaStream_t st) {
    EGNN_REQUIRE(tc_node_supported(net), "tcgen05 node path needs Hh = Ms = d in {64,128}");
    EGNN_REQUIRE(part_tile_for(net) == 32, "tcgen05 node forward expects 32-edge partial segments");
    EGNN_REQUIRE(((reinterpret_cast<uintptr_t>(agg) | reinterpret_cast<uintptr_t>(part) | reinterpret_cast<uintptr_t>(msum)) & 15) == 0,
                 "workspace must be 16-byte aligned");
    (void)precision;  // K = d + 16 only: always three split-bf16 products, in the bf16 mode as well (h feeds every later layer)
    const PackOff po = pack_off(net);
    const uint8_t* blob = reinterpret_cast<const uint8_t*>(packed_layer + po.tc);
    TcNodeFwdArgs a{net, g.N, part_tile_for(net), g.row_ptr, g.send, h, x, params_layer,
                    blob + (tc_blob_bytes(net.Mh) + 15) / 16 * 16, agg, part, phi_edge, h_out, msum, dxu, Dn};
    const int grid = (g.N + 127) / 128;
    if (net.Mh == 128) return launch_tc_node_fwd<128, 3>(a, grid, st);
    return launch_tc_node_fwd<64, 3>(a, grid, st);
}
