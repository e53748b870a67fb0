// Tile-GEMM ("wide") tcgen05 path for the network shapes the fused kernels of tc_edge.cu / tc_edge256.cu do not cover: the
// reference's own parameters.yaml network (K0 = 73, m / x / h hidden 1024, m_size 256, conditional: main.py:102-121), ragged
// widths, any 2H+1 <= 128 -- every Linear of an EGCL layer (EquivariantGraphNeuralNetwork.py:13-34, 55-71), forward AND
// backward, and the node MLP of the d = 256 networks of the fused edge kernels.  Forward of the edge phase per layer:
//     gather            u = [h_i, h_j, |r|^2]                                   -> U image        [E x K0p]
//     GEMM + SiLU       a1m = SiLU(u Wm1^T + bm1)                               -> A1 image       [E x Mhp]
//     GEMM + gate       m = SiLU(a1m Wm2^T + bm2), g = sigmoid(wa.m + ba)       -> g*m fp32       [E x 256]
//     segment sum       M_i = sum_{e in row i} g*m   (edge order, no atomics)   -> agg rows
//     GEMM + SiLU       a1x = SiLU(u Wx1^T + bx1)                               -> A1 image (reused)
//     GEMM + row dot    phi_e = wx3 . SiLU(a1x Wx2^T + bx2) + bx3               -> phi parts per N tile -> phi [E]
// node phase: [h | M] image -> GEMM + SiLU -> GEMM + bias -> h'; coordinate sums from phi_e (wide_node_geom_kernel).
// ONE GEMM kernel (tc_wide_gemm): a CTA computes [128 rows x 256 outputs] with K streamed in 64-wide slices: both operands
// arrive as ready-made bf16 hi + lo tile images by ONE cp.async.bulk each (32 KB + 64 KB per slice) into a 2-stage ring with
// full (complete_tx) / empty (tcgen05.commit) mbarriers, run by a producer warp and an MMA-issuer warp; fp32 accumulation in
// TMEM, double-buffered (2 x 256 columns) so that the 16 epilogue warps drain tile t while tile t + 1 is being multiplied;
// bf16x3 = hi.hi + lo.hi + hi.lo.  The epilogue adds the bias, applies SiLU and writes the NEXT GEMM's A operand as a tile
// image (16-byte chunk stores, 512 contiguous bytes per warp), or the gated message rows, or the per-row dot with wx3.
// Activations between GEMMs go through L2 / HBM: 4 bytes per element and GEMM against 2 K FLOPs per element.
// Backward (egnn_wide_edge_backward, egnn_wide_node_backward): both Linears are recomputed by the same GEMMs with epilogues
// that also leave SiLU' rows and TRANSPOSED images (rows = features, K = edges); gz2 comes out of the gate / phi epilogues;
// dgrad = the same GEMM on transposed weight images; weight gradients = split-K GEMMs over the edges whose partial tiles are
// stored in the parameter layout of the partial buffer reduce_partials_kernel sums; bias gradients = column sums of the
// transposed images.  Every reduction has a fixed order.  DESIGN.md 4.2 has the measurements and what bounds each GEMM.
#include "tc_bulk.cuh"
#include "tc_tile.cuh"

namespace {

// ---- weights -> tile images ---------------------------------------------------------------------------------------------------
// Image of a matrix B [N rows x K] for N tile nt and K slice ks: block (nt * KSstride + ks_off + ks): hi chunks (n, j) at
// j * 4096 + n * 16, lo at 32768 + the same; rows >= N and columns >= K are zero.  B(n, k) = W[n * ld + k] (forward: nn.Linear
// stores out x in) or, transposed = true, W[k * ld + n] (dgrad images: rows = input features)
__device__ void wide_pack_matrix(const float* __restrict__ W, int N, int K, int ld, bool transposed, int KSstride, int ks_off,
                                 uint8_t* __restrict__ dst, int t0, int stride) {
    const int per_block = WIDE_TN * (WIDE_TK / 8);
    const int NT = wide_cdiv(N, WIDE_TN), KS = wide_cdiv(K, WIDE_TK);
    for (int idx = t0; idx < NT * KS * per_block; idx += stride) {
        const int blk = idx / per_block, rem = idx - blk * per_block;
        const int j = rem / WIDE_TN, nl = rem - j * WIDE_TN;
        const int nt = blk / KS, ks = blk - nt * KS;
        const int n = nt * WIDE_TN + nl, k0 = ks * WIDE_TK + j * 8;
        uint32_t hi[4], lo[4];
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            const int k = k0 + 2 * q;
            const float v0 = (n < N && k < K) ? (transposed ? W[(size_t)k * ld + n] : W[(size_t)n * ld + k]) : 0.f;
            const float v1 = (n < N && k + 1 < K) ? (transposed ? W[(size_t)(k + 1) * ld + n] : W[(size_t)n * ld + k + 1]) : 0.f;
            tc::split2(v0, v1, hi[q], lo[q]);
        }
        uint8_t* b = dst + (size_t)(nt * KSstride + ks_off + ks) * WIDE_B_SLICE + j * (WIDE_TN * 16) + nl * 16;
        *reinterpret_cast<uint4*>(b) = make_uint4(hi[0], hi[1], hi[2], hi[3]);
        *reinterpret_cast<uint4*>(b + WIDE_B_SLICE / 2) = make_uint4(lo[0], lo[1], lo[2], lo[3]);
    }
}

__global__ void __launch_bounds__(256) wide_pack_kernel(egnn_net_t net, const float* __restrict__ params, float* __restrict__ packed) {
    const LayerOff lo = layer_off(net);
    const PackOff po = pack_off(net);
    const WideOff wo = wide_off(net);
    const float* P = params + (size_t)blockIdx.y * lo.total;
    uint8_t* img = reinterpret_cast<uint8_t*>(packed + (size_t)blockIdx.y * po.total + po.wide);
    const int K0 = 2 * net.H + 1, t0 = blockIdx.x * blockDim.x + threadIdx.x, stride = gridDim.x * blockDim.x;
    auto at = [&](int first) { return img + (size_t)first * WIDE_B_SLICE; };
    if (wide_fwd_supported(net)) {
        wide_pack_matrix(P + lo.wm1, net.Mh, K0, K0, false, wo.ks0, 0, at(wo.m1), t0, stride);
        wide_pack_matrix(P + lo.wm2, net.Ms, net.Mh, net.Mh, false, wo.ksm, 0, at(wo.m2), t0, stride);
        wide_pack_matrix(P + lo.wx1, net.Xh, K0, K0, false, wo.ks0, 0, at(wo.x1), t0, stride);
        wide_pack_matrix(P + lo.wx2, net.Xh, net.Xh, net.Xh, false, wo.ksx, 0, at(wo.x2), t0, stride);
        // dgrad images: ga1 = gz2 . W2 -> rows = columns of W2;  gu = [gz1m | gz1x] . [Wm1 ; Wx1] -> rows = the K0 input columns
        wide_pack_matrix(P + lo.wm2, net.Mh, net.Ms, net.Mh, true, wo.kss, 0, at(wo.m2t), t0, stride);
        wide_pack_matrix(P + lo.wx2, net.Xh, net.Xh, net.Xh, true, wo.ksx, 0, at(wo.x2t), t0, stride);
        wide_pack_matrix(P + lo.wm1, K0, net.Mh, K0, true, wo.ksm + wo.ksx, 0, at(wo.w1t), t0, stride);
        wide_pack_matrix(P + lo.wx1, K0, net.Xh, K0, true, wo.ksm + wo.ksx, wo.ksm, at(wo.w1t), t0, stride);
    }
    if (wide_node_supported(net)) {  // mlp_h: forward images, then the dgrad images (rows = input features)
        const int Kn = net.H + net.Ms;
        wide_pack_matrix(P + lo.wh1, net.Hh, Kn, Kn, false, wo.ksn, 0, at(wo.h1), t0, stride);
        wide_pack_matrix(P + lo.wh2, net.H, net.Hh, net.Hh, false, wo.ksh, 0, at(wo.h2), t0, stride);
        wide_pack_matrix(P + lo.wh2, net.Hh, net.H, net.Hh, true, wo.ksH, 0, at(wo.h2t), t0, stride);
        wide_pack_matrix(P + lo.wh1, Kn, net.Hh, Kn, true, wo.ksh, 0, at(wo.h1t), t0, stride);
    }
}

// ---- gather: one thread per (edge row, 8-column chunk) of the U image; rows >= E are zero -----------------------------------------
// UT != nullptr (backward): also the transposed image (rows = the K0 input columns, K = edges), the B operand of dW1 = gz1^T u
__global__ void __launch_bounds__(256) wide_gather_kernel(int E, int mtiles, int H, int KS0, const int32_t* __restrict__ recv,
                                                          const int32_t* __restrict__ send, const float* __restrict__ h,
                                                          const float* __restrict__ x, uint8_t* __restrict__ Uimg,
                                                          uint8_t* __restrict__ UT) {
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
        if (UT) {
            uint8_t* tb = UT + (size_t)(e / WIDE_TK) * WIDE_B_SLICE + ((e % WIDE_TK) / 8) * (WIDE_TN * 16) + (e % 8) * 2;
#pragma unroll
            for (int q = 0; q < 8; ++q) {
                const uint32_t hw = hi[q >> 1], lw = lo[q >> 1];
                uint8_t* d = tb + (j * 8 + q) * 16;
                *reinterpret_cast<uint16_t*>(d) = (uint16_t)(q & 1 ? hw >> 16 : hw & 0xffffu);
                *reinterpret_cast<uint16_t*>(d + WIDE_B_SLICE / 2) = (uint16_t)(q & 1 ? lw >> 16 : lw & 0xffffu);
            }
        }
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
__global__ void __launch_bounds__(256) wide_phi_kernel(int E, int NT, int64_t part_stride, const float* __restrict__ part,
                                                       const float* __restrict__ bx3, float* __restrict__ phi) {
    const int e = blockIdx.x * 256 + threadIdx.x;
    if (e >= E) return;
    float s = bx3[0];
    for (int t = 0; t < NT; ++t) s += part[(size_t)t * part_stride + e];
    phi[e] = s;
}

// ---- backward helpers ---------------------------------------------------------------------------------------------------------
// gphi_e = gdx[i] . r_e  (egcl_bwd.cu: the coordinate branch's row coefficient); rows >= E are zero
__global__ void __launch_bounds__(256) wide_gphi_kernel(int E, int Ep, const int32_t* __restrict__ recv, const int32_t* __restrict__ send,
                                                        const float* __restrict__ x, const float* __restrict__ gdx,
                                                        float* __restrict__ gphi) {
    const int e = blockIdx.x * 256 + threadIdx.x;
    if (e >= Ep) return;
    float v = 0.f;
    if (e < E) {
        const int i = recv[e], j = send[e];
        v = gdx[3 * i] * (x[3 * i] - x[3 * j]) + gdx[3 * i + 1] * (x[3 * i + 1] - x[3 * j + 1]) + gdx[3 * i + 2] * (x[3 * i + 2] - x[3 * j + 2]);
    }
    gphi[e] = v;
}

// column sums of a transposed (rows = features) hi | lo image over the 64-edge slices of split blockIdx.y: bias gradients.
// 1024 threads = the 128 feature rows of tile blockIdx.x x the 8 chunk columns of a slice; fixed order inside a split, the
// splits are summed by reduce_partials_kernel
__global__ void __launch_bounds__(1024) wide_colsum_T_kernel(const uint8_t* __restrict__ T, int kst, int per, int N, float* __restrict__ part,
                                                             int64_t part_stride) {
    __shared__ float red[8][128];
    const int nl = threadIdx.x & 127, j = threadIdx.x >> 7, n = blockIdx.x * 128 + nl, s = blockIdx.y;
    const int k0 = s * per, k1 = min(k0 + per, kst);
    float sum = 0.f;
#pragma unroll 4
    for (int ks = k0; ks < k1; ++ks) {
        const uint8_t* b = T + ((size_t)blockIdx.x * kst + ks) * WIDE_A_SLICE + j * (WIDE_TM * 16) + nl * 16;
        const uint4 h = *reinterpret_cast<const uint4*>(b);
        const uint4 l = *reinterpret_cast<const uint4*>(b + WIDE_A_SLICE / 2);
        const uint32_t hw[4] = {h.x, h.y, h.z, h.w}, lw[4] = {l.x, l.y, l.z, l.w};
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            sum += __uint_as_float(hw[q] << 16) + __uint_as_float(lw[q] << 16);
            sum += __uint_as_float(hw[q] & 0xffff0000u) + __uint_as_float(lw[q] & 0xffff0000u);
        }
    }
    red[j][nl] = sum;
    __syncthreads();
    if (j == 0 && n < N) {
        float t = red[0][nl];
#pragma unroll
        for (int k = 1; k < 8; ++k) t += red[k][nl];
        part[(size_t)s * part_stride + n] = t;
    }
}

// column sums of fp32 rows [E x ld] over the edges of split blockIdx.y (32 columns x 8 row lanes per CTA, fixed order)
__global__ void __launch_bounds__(256) wide_colsum_rows_kernel(const float* __restrict__ rows, int ld, int ncols, int E, int per_edges,
                                                               float* __restrict__ part, int64_t part_stride) {
    __shared__ float red[8][32];
    const int cl = threadIdx.x & 31, rl = threadIdx.x >> 5, c = blockIdx.x * 32 + cl, s = blockIdx.y;
    const int e0 = s * per_edges, e1 = min(e0 + per_edges, E);
    float sum = 0.f;
    if (c < ncols) {
#pragma unroll 4
        for (int e = e0 + rl; e < e1; e += 8) sum += rows[(size_t)e * ld + c];
    }
    red[rl][cl] = sum;
    __syncthreads();
    if (rl == 0 && c < ncols) {
        float t = red[0][cl];
#pragma unroll
        for (int k = 1; k < 8; ++k) t += red[k][cl];
        part[(size_t)s * part_stride + c] = t;
    }
}
// sum of a per-edge vector over the edges of split blockIdx.x (d ba, d bx3): 1024 strided partial sums, fixed-order tree
__global__ void __launch_bounds__(1024) wide_sum_vec_kernel(const float* __restrict__ v, int E, int per_edges, float* __restrict__ part,
                                                            int64_t part_stride) {
    __shared__ float red[1024];
    const int s = blockIdx.x, e0 = s * per_edges, e1 = min(e0 + per_edges, E);
    float sum = 0.f;
#pragma unroll 4
    for (int e = e0 + threadIdx.x; e < e1; e += 1024) sum += v[e];
    red[threadIdx.x] = sum;
    __syncthreads();
    for (int w = 512; w > 0; w >>= 1) {
        if (threadIdx.x < w) red[threadIdx.x] += red[threadIdx.x + w];
        __syncthreads();
    }
    if (threadIdx.x == 0) part[(size_t)s * part_stride] = red[0];
}
// the same for wide row matrices (ld a multiple of 4, 16-byte aligned rows): 128 columns x 32 row lanes per CTA, float4 loads
__global__ void __launch_bounds__(1024) wide_colsum_rows4_kernel(const float* __restrict__ rows, int ld, int ncols, int E, int per_edges,
                                                                 float* __restrict__ part, int64_t part_stride) {
    __shared__ float4 red[32][32];
    const int cl = threadIdx.x & 31, rl = threadIdx.x >> 5, c = blockIdx.x * 128 + 4 * cl, s = blockIdx.y;
    const int e0 = s * per_edges, e1 = min(e0 + per_edges, E);
    float4 sum = make_float4(0.f, 0.f, 0.f, 0.f);
    if (c < ncols) {
#pragma unroll 4
        for (int e = e0 + rl; e < e1; e += 32) {
            const float4 v = *reinterpret_cast<const float4*>(rows + (size_t)e * ld + c);
            sum.x += v.x, sum.y += v.y, sum.z += v.z, sum.w += v.w;
        }
    }
    red[rl][cl] = sum;
    __syncthreads();
    if (rl == 0 && c < ncols) {
        float4 t = red[0][cl];
#pragma unroll 1
        for (int k = 1; k < 32; ++k) t.x += red[k][cl].x, t.y += red[k][cl].y, t.z += red[k][cl].z, t.w += red[k][cl].w;
        float* o = part + (size_t)s * part_stride + c;
        o[0] = t.x;
        if (c + 1 < ncols) o[1] = t.y;
        if (c + 2 < ncols) o[2] = t.z;
        if (c + 3 < ncols) o[3] = t.w;
    }
}

// per-edge outputs of the layer backward (egcl_bwd.cu: egcl_edge_bwd, "per-edge outputs"): receiver / sender contributions
__global__ void __launch_bounds__(256) wide_edge_out_kernel(int E, int H, const int32_t* __restrict__ recv, const int32_t* __restrict__ send,
                                                            const int32_t* __restrict__ node_graph, const float* __restrict__ x,
                                                            const float* __restrict__ gu, const float* __restrict__ gdx,
                                                            const float* __restrict__ coefS, const float* __restrict__ phi,
                                                            float* __restrict__ Gr, float* __restrict__ Gs) {
    const int W = H + 3;
    const long long idx = (long long)blockIdx.x * 256 + threadIdx.x;
    if (idx >= (long long)E * W) return;
    const int e = (int)(idx / W), c = (int)(idx - (long long)e * W);
    if (c < H) {
        Gr[idx] = gu[(size_t)e * 128 + c];
        Gs[idx] = gu[(size_t)e * 128 + H + c];
    } else {
        const int k = c - H, i = recv[e], j = send[e];
        const float gd = gu[(size_t)e * 128 + 2 * H];  // grad of d_e = |r_e|^2
        const float rk = x[3 * i + k] - x[3 * j + k];
        const float gr = gdx[(size_t)i * 3 + k] * phi[e] + (coefS[node_graph[i]] + 2.f * gd) * rk;
        Gr[idx] = gr;
        Gs[idx] = -gr;
    }
}

// per-receiver sum of column c from the direct rows and the head / tail partials of the edge tiles that cut the receiver's
// edge range (egcl_fwd.cu: combine_agg, same order)
__device__ __forceinline__ float wide_combine(const float* __restrict__ agg, const float* __restrict__ part, const int32_t* __restrict__ row_ptr,
                                              int i, int c, int AGGW, int TE) {
    const int ra = row_ptr[i], rb = row_ptr[i + 1];
    if (ra == rb) return 0.f;
    const int t0 = ra / TE, t1 = (rb - 1) / TE;
    if (t0 == t1) return agg[(size_t)i * AGGW + c];
    float s = part[((size_t)t0 * 2 + 1) * AGGW + c];
    for (int t = t0 + 1; t <= t1; ++t) s += part[((size_t)t * 2) * AGGW + c];
    return s;
}

// ---- node phase (mlp_h on the tile GEMMs) ---------------------------------------------------------------------------------------
// fp32 rows -> operand image (+ transposed image): value(i, k) = k < W1 ? src1[i * W1 + k] : src2[i * ld2 + (k - W1)] for k < W1 + W2.
// The node MLP's input [h | M] (W1 = H, W2 = Ms; forward: src2 = the per-receiver sums, copied to copy2 = msum), and with
// W2 = 0 any row matrix (gh_out).  t_rows = 256: transposed image in the B layout (weight-gradient B operand), 128: A layout
// part != nullptr: src2 / part are the per-receiver sums of a fused edge kernel, cut at TE-edge segment boundaries (wide_combine)
__global__ void __launch_bounds__(256) wide_rows_to_img_kernel(int N, int mtiles, int W1, int W2, int KS, const float* __restrict__ src1,
                                                               const float* __restrict__ src2, int ld2, const float* __restrict__ part,
                                                               const int32_t* __restrict__ row_ptr, int TE, float* __restrict__ copy2,
                                                               uint8_t* __restrict__ img, uint8_t* __restrict__ T, int t_rows, int kst) {
    const int nchunk = KS * (WIDE_TK / 8);
    const long long total = (long long)mtiles * WIDE_TM * nchunk;
    for (long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += (long long)gridDim.x * blockDim.x) {
        const int row_in_tile = (int)(idx % WIDE_TM);
        const long long rest = idx / WIDE_TM;
        const int j = (int)(rest % nchunk);
        const int t = (int)(rest / nchunk);
        const int i = t * WIDE_TM + row_in_tile;
        float v[8];
#pragma unroll
        for (int q = 0; q < 8; ++q) {
            const int k = j * 8 + q;
            float val = 0.f;
            if (i < N) {
                if (k < W1) val = src1[(size_t)i * W1 + k];
                else if (k < W1 + W2) {
                    val = part ? wide_combine(src2, part, row_ptr, i, k - W1, ld2, TE) : src2[(size_t)i * ld2 + (k - W1)];
                    if (copy2) copy2[(size_t)i * W2 + (k - W1)] = val;
                }
            }
            v[q] = val;
        }
        uint32_t hi[4], lo[4];
#pragma unroll
        for (int q = 0; q < 4; ++q) tc::split2(v[2 * q], v[2 * q + 1], hi[q], lo[q]);
        uint8_t* b = img + ((size_t)t * KS + j / 8) * WIDE_A_SLICE + (j % 8) * (WIDE_TM * 16) + row_in_tile * 16;
        *reinterpret_cast<uint4*>(b) = make_uint4(hi[0], hi[1], hi[2], hi[3]);
        *reinterpret_cast<uint4*>(b + WIDE_A_SLICE / 2) = make_uint4(lo[0], lo[1], lo[2], lo[3]);
        if (T) {
            const int slice_bytes = t_rows * WIDE_TK * 4;  // hi | lo
            const int k0 = j * 8;
            uint8_t* tb = T + ((size_t)(k0 / t_rows) * kst + i / WIDE_TK) * slice_bytes + ((i % WIDE_TK) / 8) * (t_rows * 16) + (i % 8) * 2;
#pragma unroll
            for (int q = 0; q < 8; ++q) {
                const uint32_t hw = hi[q >> 1], lw = lo[q >> 1];
                uint8_t* d = tb + (k0 % t_rows + q) * 16;
                *reinterpret_cast<uint16_t*>(d) = (uint16_t)(q & 1 ? hw >> 16 : hw & 0xffffu);
                *reinterpret_cast<uint16_t*>(d + slice_bytes / 2) = (uint16_t)(q & 1 ? lw >> 16 : lw & 0xffffu);
            }
        }
    }
}

// un-normalised coordinate sums and sum of |r_e|^2 per receiver from phi_e (egcl_fwd.cu: egcl_node_fwd, the phi_edge branch):
// 8 lanes per node over its edges, fixed order
__global__ void __launch_bounds__(256) wide_node_geom_kernel(int N, const int32_t* __restrict__ row_ptr, const int32_t* __restrict__ send,
                                                             const float* __restrict__ x, const float* __restrict__ phi_edge,
                                                             float* __restrict__ dxu, float* __restrict__ D) {
    const int i = blockIdx.x * 32 + (threadIdx.x >> 3), sub = threadIdx.x & 7;
    float sx = 0.f, sy = 0.f, sz = 0.f, sd = 0.f;
    if (i < N) {
        const float xi = x[3 * i], yi = x[3 * i + 1], zi = x[3 * i + 2];
        for (int e = row_ptr[i] + sub; e < row_ptr[i + 1]; e += 8) {
            const int j = send[e];
            const float rx = xi - x[3 * j], ry = yi - x[3 * j + 1], rz = zi - x[3 * j + 2];
            const float ph = phi_edge[e];
            sx = fmaf(rx, ph, sx);
            sy = fmaf(ry, ph, sy);
            sz = fmaf(rz, ph, sz);
            sd += rx * rx + ry * ry + rz * rz;
        }
    }
#pragma unroll
    for (int o = 4; o > 0; o >>= 1) {
        sx += __shfl_xor_sync(0xffffffffu, sx, o);
        sy += __shfl_xor_sync(0xffffffffu, sy, o);
        sz += __shfl_xor_sync(0xffffffffu, sz, o);
        sd += __shfl_xor_sync(0xffffffffu, sd, o);
    }
    if (i < N && sub == 0) {
        dxu[(size_t)i * 3 + 0] = sx;
        dxu[(size_t)i * 3 + 1] = sy;
        dxu[(size_t)i * 3 + 2] = sz;
        D[i] = sd;
    }
}

// g[h | M] rows -> gh_node [N x H], gM [N x Ms]
__global__ void __launch_bounds__(256) wide_split_gin_kernel(int N, int H, int Ms, const float* __restrict__ gin, int ld,
                                                             float* __restrict__ gh_node, float* __restrict__ gM) {
    const long long idx = (long long)blockIdx.x * 256 + threadIdx.x;
    const int W = H + Ms;
    if (idx >= (long long)N * W) return;
    const int i = (int)(idx / W), c = (int)(idx - (long long)i * W);
    const float v = gin[(size_t)i * ld + c];
    if (c < H) gh_node[(size_t)i * H + c] = v;
    else gM[(size_t)i * Ms + (c - H)] = v;
}

// ---- the tile GEMM ------------------------------------------------------------------------------------------------------------
// forward epilogues:  IMG (SiLU -> next operand image), GATE (gated messages), ROWDOT (phi parts)
// backward epilogues: IMG_BWD (IMG + transposed image + SiLU' rows), GATE_BWD / ROWDOT_BWD (recompute the second Linear and
//                     turn it into gz2: operand image + transposed image + the rows whose column sums are d wa / d wx3),
//                     DGRAD (gz1 = acc * SiLU'(z1): image + transposed image), F32 (plain rows: gu), WGRAD (split-K partial
//                     tile of a weight gradient, stored in the parameter layout of one row of the partial buffer)
enum { WIDE_IMG = 0, WIDE_GATE = 1, WIDE_ROWDOT = 2, WIDE_IMG_BWD = 3, WIDE_GATE_BWD = 4, WIDE_ROWDOT_BWD = 5, WIDE_DGRAD = 6,
       WIDE_F32 = 7, WIDE_WGRAD = 8 };
struct WideGemmArgs {
    const uint8_t *A, *A2;  // A operand images (rows x K slices): slice (mt, ks) at (mt * a_ks + ks) * WIDE_A_SLICE; K slices >= ks_split
    int a_ks, a2_ks, ks_split;  // come from A2 (slice ks - ks_split, row stride a2_ks)
    const uint8_t* B;       // B operand image: block (nt, ks) at (nt * b_ks + ks) * WIDE_B_SLICE
    int b_ks;
    int KS, KS_total;       // K slices per tile, and in all: split s of a WGRAD tile covers [s * KS, min((s + 1) * KS, KS_total))
    int NT, mtiles, splits;
    int E, N;               // valid rows, valid output columns (WGRAD: valid feature rows n, valid input columns k)
    const float* bias;      // [N]
    uint8_t* out_img;       // operand image written by IMG / IMG_BWD / GATE_BWD / ROWDOT_BWD / DGRAD, out_ks slices per row tile
    int out_ks;
    uint8_t* out_T;         // transposed image (kst = 64-edge slices): IMG_BWD rows = 256-row B tiles, the others 128-row A tiles
    int kst;
    const float *vec, *scalar;  // GATE*: wa [N], ba;  ROWDOT*: wx3 [N]
    float* out;                 // GATE: g*m [E][WIDE_TN]; ROWDOT*: phi parts [NT][out_ld]; F32: rows [E][out_ld]; WGRAD: partial rows
    int64_t out_ld;             // WGRAD: row stride of the weight matrix; its split stride is out_split
    int64_t out_split;
    float* rows;                // IMG_BWD: SiLU' rows;  GATE_BWD: gs * m;  ROWDOT_BWD: gphi * a2;  DGRAD (in): SiLU' rows
    int rows_ld;
    float* q;                   // GATE_BWD: gs per edge
    const float* gM;            // GATE_BWD: [N nodes][Ms] gradient of the message sums
    const int32_t* recv;
    const float* gphi;          // ROWDOT_BWD: per edge
};
constexpr int WIDE_EPI_THREADS = 512, WIDE_THREADS = WIDE_EPI_THREADS + 64, WIDE_STAGES = 2, WIDE_STAGE_BYTES = WIDE_A_SLICE + WIDE_B_SLICE;
constexpr int WIDE_SMEM = WIDE_STAGES * WIDE_STAGE_BYTES + 1024 /*bias, vec*/ * 2 + 2 * 4 * WIDE_TM * 4 /*row partials*/ + 128;

// 8 consecutive columns of row r -> chunk j of a [128-row] operand slice (hi | lo)
__device__ __forceinline__ void wide_store_chunk(uint8_t* slice_row, int j, const float (&v)[8]) {
    uint32_t hi[4], lo[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) tc::split2(v[2 * i], v[2 * i + 1], hi[i], lo[i]);
    *reinterpret_cast<uint4*>(slice_row + j * (WIDE_TM * 16)) = make_uint4(hi[0], hi[1], hi[2], hi[3]);
    *reinterpret_cast<uint4*>(slice_row + j * (WIDE_TM * 16) + WIDE_A_SLICE / 2) = make_uint4(lo[0], lo[1], lo[2], lo[3]);
}
// the same 8 values into the transposed image (rows = features, 16-byte chunks of 8 consecutive edges).  The 8 lanes of a
// group hold 8 consecutive edges: an 8 x 8 transpose through the shuffle network (3 butterfly steps) leaves lane s of the
// group with feature row0 + s of all 8 edges, i.e. one whole chunk -- a 16-byte store per lane, 128 contiguous bytes per
// group, instead of sixteen 2-byte stores.  tchunk = the slice's chunk column of this edge group; v is consumed.
__device__ __forceinline__ void wide_store_T(uint8_t* tchunk, int row0, int lo_off, float (&v)[8], int lane) {
#pragma unroll
    for (int d = 4; d > 0; d >>= 1) {
        const bool up = lane & d;
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            if (i & d) continue;
            const float send = up ? v[i] : v[i + d];
            const float recv = __shfl_xor_sync(0xffffffffu, send, d);
            if (up) v[i] = recv;
            else v[i + d] = recv;
        }
    }
    uint32_t hi[4], lo[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) tc::split2(v[2 * i], v[2 * i + 1], hi[i], lo[i]);
    uint8_t* d = tchunk + (row0 + (lane & 7)) * 16;
    *reinterpret_cast<uint4*>(d) = make_uint4(hi[0], hi[1], hi[2], hi[3]);
    *reinterpret_cast<uint4*>(d + lo_off) = make_uint4(lo[0], lo[1], lo[2], lo[3]);
}

// this thread's accumulator columns [c0 + 8 j, +8) of the tile; the MMAs only write the first ncols columns (a multiple of 16
// covering the valid outputs of this N tile): the rest of the TMEM buffer holds an older tile and reads as zero
__device__ __forceinline__ void wide_ld8(uint32_t taddr, int j, int c0, int ncols, float (&v)[8]) {
    tc::tmem_ld8(taddr + 8 * j, v);
    if (c0 + 8 * j >= ncols) {
#pragma unroll
        for (int i = 0; i < 8; ++i) v[i] = 0.f;
    }
}

// 8 consecutive floats to global memory with the widest store the address allows
__device__ __forceinline__ void wide_store8(float* o, const float (&v)[8]) {
    const uintptr_t a = reinterpret_cast<uintptr_t>(o);
    if ((a & 15) == 0) {
        *reinterpret_cast<float4*>(o) = make_float4(v[0], v[1], v[2], v[3]);
        *reinterpret_cast<float4*>(o + 4) = make_float4(v[4], v[5], v[6], v[7]);
    } else if ((a & 7) == 0) {
#pragma unroll
        for (int i = 0; i < 4; ++i) *reinterpret_cast<float2*>(o + 2 * i) = make_float2(v[2 * i], v[2 * i + 1]);
    } else {
#pragma unroll
        for (int i = 0; i < 8; ++i) o[i] = v[i];
    }
}

// 18 warps: 16 epilogue warps (warp w reads TMEM lanes 32 (w & 3) .., columns 64 (w >> 2) ..), one producer warp (bulk copies)
// and one MMA-issuer warp.  The accumulator is double-buffered in TMEM (2 x 256 columns): while the epilogue warps drain
// tile t from buffer t & 1, the issuer already runs tile t + 1 into the other buffer -- the CUDA-core epilogue (SiLU, bf16
// splits, image stores) and the tensor pipe overlap.  Barriers: full / empty per ring stage (complete_tx / tcgen05.commit),
// acc_full per buffer (tcgen05.commit), acc_empty per buffer (one arrival per epilogue warp); the epilogue warps synchronise
// among themselves on named barrier 1.
__device__ __forceinline__ void wide_epi_sync() { asm volatile("bar.sync 1, %0;" ::"r"(WIDE_EPI_THREADS) : "memory"); }

template <int MODE, int NPASS>
__global__ void __launch_bounds__(WIDE_THREADS, 1) tc_wide_gemm(WideGemmArgs a) {
    extern __shared__ __align__(1024) uint8_t smem[];
    float* sBias = reinterpret_cast<float*>(smem + WIDE_STAGES * WIDE_STAGE_BYTES);
    float* sVec = sBias + WIDE_TN;
    float* sDot = sVec + WIDE_TN;      // [4][128]
    float* sDot2 = sDot + 4 * WIDE_TM;  // [4][128]
    uint64_t* bars = reinterpret_cast<uint64_t*>(sDot2 + 4 * WIDE_TM);  // full[2], empty[2], acc_full[2], acc_empty[2]
    uint32_t* slot = reinterpret_cast<uint32_t*>(bars + 8);
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, q = warp & 3, cg = (warp >> 2) & 3;
    if (warp == 0) tc::tmem_alloc(slot, 512);
    if (tid == 0) {
        for (int b = 0; b < 6; ++b) tc::mbar_init(&bars[b], 1);
        for (int b = 6; b < 8; ++b) tc::mbar_init(&bars[b], WIDE_EPI_THREADS / 32);
        tc::fence_mbar_init();
    }
    tc::fence_before_sync();
    __syncthreads();
    tc::fence_after_sync();
    const uint32_t tmem = *slot;
    const uint32_t sb = tc::smem_u32(smem);
    const int tiles = a.mtiles * a.NT * a.splits;
    const int nit = (int)blockIdx.x < tiles ? (tiles - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x : 0;
    bool ok = true;
    // tile number -> (N tile, row tile, K range).  The NT tiles of one row tile run side by side: its A slices are shared in L2
    auto decode = [&](int it, int& nt, int& mt, int& kb, int& kc) {
        const int tile = blockIdx.x + it * gridDim.x;
        nt = tile % a.NT;
        const int rest = tile / a.NT;
        mt = rest % a.mtiles;
        const int s = rest / a.mtiles;
        kb = s * a.KS;
        kc = min(a.KS, a.KS_total - kb);
    };
    pdl_launch_dependents();
    pdl_wait();  // the A image comes from the kernel in front

    if (warp == WIDE_EPI_THREADS / 32) {  // ---- producer: one thread runs the ring as far ahead as the empty barriers allow
        if (lane == 0) {
            int g = 0;
            for (int it = 0; it < nit; ++it) {
                int nt, mt, kb, kc;
                decode(it, nt, mt, kb, kc);
                for (int k = 0; k < kc; ++k, ++g) {
                    const int s = g % WIDE_STAGES, u = g / WIDE_STAGES;
                    if (u > 0) ok = tc::mbar_wait(&bars[2 + s], (u - 1) & 1) && ok;
                    const int ks = kb + k;
                    const uint8_t* srcA = ks < a.ks_split ? a.A + ((size_t)mt * a.a_ks + ks) * WIDE_A_SLICE
                                                          : a.A2 + ((size_t)mt * a.a2_ks + (ks - a.ks_split)) * WIDE_A_SLICE;
                    const uint32_t full = tc::smem_u32(&bars[s]), dst = sb + s * WIDE_STAGE_BYTES;
                    tc::mbar_expect_tx(full, WIDE_STAGE_BYTES);
                    tc::bulk_g2s(dst, srcA, WIDE_A_SLICE, full);
                    tc::bulk_g2s(dst + WIDE_A_SLICE, a.B + ((size_t)nt * a.b_ks + ks) * WIDE_B_SLICE, WIDE_B_SLICE, full);
                }
            }
        }
    } else if (warp == WIDE_EPI_THREADS / 32 + 1) {  // ---- MMA issuer
        if (lane == 0) {
            int g = 0;
            for (int it = 0; it < nit; ++it) {
                int nt, mt, kb, kc;
                decode(it, nt, mt, kb, kc);
                const int b = it & 1;
                const int ncols = min(WIDE_TN, (a.N - nt * WIDE_TN + 15) & ~15);  // MMA N: the valid outputs of this tile, in 16s
                if (it >= 2) ok = tc::mbar_wait(&bars[6 + b], ((it >> 1) - 1) & 1) && ok;  // the epilogue has drained this buffer
                tc::fence_after_sync();
                for (int k = 0; k < kc; ++k, ++g) {
                    const int s = g % WIDE_STAGES;
                    ok = tc::mbar_wait(&bars[s], (g / WIDE_STAGES) & 1) && ok;
                    tc::fence_after_sync();
                    const uint32_t base = sb + s * WIDE_STAGE_BYTES;
                    const TcOperand opA{base, base + WIDE_A_SLICE / 2, WIDE_TM * 16, false};
                    const TcOperand opB{base + WIDE_A_SLICE, base + WIDE_A_SLICE + WIDE_B_SLICE / 2, WIDE_TN * 16, false};
                    tc_issue_op<NPASS>(tmem, b * WIDE_TN, opA, opB, WIDE_TK, ncols, k > 0);
                    tc::mma_commit(&bars[2 + s]);
                }
                tc::mma_commit(&bars[4 + b]);
            }
        }
    } else
    for (int it = 0; it < nit; ++it) {  // ---- epilogue warps
        int nt, mt, kb, kc;
        decode(it, nt, mt, kb, kc);
        const int b = it & 1;
        // bias / vector of this N tile (columns past N are zero: their accumulators are zero too, SiLU(0) = 0)
        if (MODE != WIDE_DGRAD && MODE != WIDE_F32 && MODE != WIDE_WGRAD && tid < WIDE_TN) {
            const int c = tid, n = nt * WIDE_TN + c;
            sBias[c] = n < a.N ? a.bias[n] : 0.f;
            if (MODE == WIDE_GATE || MODE == WIDE_ROWDOT || MODE == WIDE_GATE_BWD || MODE == WIDE_ROWDOT_BWD)
                sVec[c] = n < a.N ? a.vec[n] : 0.f;
        }
        wide_epi_sync();
        ok = tc::mbar_wait(&bars[4 + b], (it >> 1) & 1) && ok;
        tc::fence_after_sync();
        __syncwarp();

        const int r = q * 32 + lane, e = mt * WIDE_TM + r;
        const int ncols = min(WIDE_TN, (a.N - nt * WIDE_TN + 15) & ~15);
        const bool live = e < a.E;
        const int c0 = cg * 64;                // this thread's 64 columns of the tile
        const int n0 = nt * WIDE_TN + c0;      // ... of the matrix
        const uint32_t taddr = tmem + ((uint32_t)(q * 32) << 16) + b * WIDE_TN + c0;
        if (MODE == WIDE_IMG || MODE == WIDE_IMG_BWD) {
            uint8_t* o = a.out_img + ((size_t)mt * a.out_ks + n0 / WIDE_TK) * WIDE_A_SLICE + r * 16;
            uint8_t* t = nullptr;
            float* d1 = nullptr;
            if (MODE == WIDE_IMG_BWD) {
                t = a.out_T + ((size_t)nt * a.kst + e / WIDE_TK) * WIDE_B_SLICE + ((e % WIDE_TK) / 8) * (WIDE_TN * 16);
                d1 = a.rows + (size_t)e * a.rows_ld + n0;
            }
#pragma unroll 1
            for (int j = 0; j < 8; ++j) {
                float v[8];
                wide_ld8(taddr, j, c0, ncols, v);
                if (MODE == WIDE_IMG) {
#pragma unroll
                    for (int i = 0; i < 8; ++i) v[i] = tc_silu<true>(v[i] + sBias[c0 + 8 * j + i]);
                } else {
                    float d[8];
#pragma unroll
                    for (int i = 0; i < 8; ++i) {
                        tc_silu_and_grad(v[i] + sBias[c0 + 8 * j + i], v[i], d[i]);
                        if (!live) v[i] = 0.f, d[i] = 0.f;
                    }
                    *reinterpret_cast<float4*>(d1 + 8 * j) = make_float4(d[0], d[1], d[2], d[3]);
                    *reinterpret_cast<float4*>(d1 + 8 * j + 4) = make_float4(d[4], d[5], d[6], d[7]);
                }
                wide_store_chunk(o, j, v);
                if (MODE == WIDE_IMG_BWD) wide_store_T(t, c0 + 8 * j, WIDE_B_SLICE / 2, v, lane);
            }
        } else if (MODE == WIDE_GATE || MODE == WIDE_ROWDOT || MODE == WIDE_GATE_BWD || MODE == WIDE_ROWDOT_BWD) {
            // row dot of the activated outputs with vec: 64 columns per thread, then the four column groups in fixed order
            const float* gMi = nullptr;
            float gph = 0.f;
            uint8_t *o = nullptr, *t = nullptr;
            if (MODE == WIDE_GATE_BWD) gMi = a.gM + (size_t)(live ? a.recv[e] : 0) * a.N;
            if (MODE == WIDE_ROWDOT_BWD) gph = live ? a.gphi[e] : 0.f;
            if (MODE == WIDE_GATE_BWD || MODE == WIDE_ROWDOT_BWD) {
                o = a.out_img + ((size_t)mt * a.out_ks + n0 / WIDE_TK) * WIDE_A_SLICE + r * 16;
                t = a.out_T + ((size_t)(n0 / WIDE_TM) * a.kst + e / WIDE_TK) * WIDE_A_SLICE + ((e % WIDE_TK) / 8) * (WIDE_TM * 16);
            }
            float dot = 0.f, dot2 = 0.f;
#pragma unroll 1
            for (int j = 0; j < 8; ++j) {
                float v[8];
                wide_ld8(taddr, j, c0, ncols, v);
                if (MODE == WIDE_ROWDOT_BWD) {
                    // a2 = SiLU(z2), phi part += wx3 . a2;  gz2 = gphi wx3 SiLU'(z2);  rows = gphi a2 (column sums = d wx3)
                    float g2[8], ga[8];
#pragma unroll
                    for (int i = 0; i < 8; ++i) {
                        float act, dact;
                        tc_silu_and_grad(v[i] + sBias[c0 + 8 * j + i], act, dact);
                        const float w = sVec[c0 + 8 * j + i];
                        dot = fmaf(act, w, dot);
                        g2[i] = gph * w * dact;
                        ga[i] = gph * act;
                    }
                    float* ro = a.rows + (size_t)e * a.rows_ld + n0 + 8 * j;
                    *reinterpret_cast<float4*>(ro) = make_float4(ga[0], ga[1], ga[2], ga[3]);
                    *reinterpret_cast<float4*>(ro + 4) = make_float4(ga[4], ga[5], ga[6], ga[7]);
                    wide_store_chunk(o, j, g2);
                    wide_store_T(t, (c0 % WIDE_TM) + 8 * j, WIDE_A_SLICE / 2, g2, lane);
                } else {
#pragma unroll
                    for (int i = 0; i < 8; ++i) {
                        const float m = tc_silu<true>(v[i] + sBias[c0 + 8 * j + i]);
                        dot = fmaf(m, sVec[c0 + 8 * j + i], dot);
                        if (MODE == WIDE_GATE_BWD) dot2 = fmaf(m, c0 + 8 * j + i < a.N ? gMi[c0 + 8 * j + i] : 0.f, dot2);
                    }
                }
            }
            sDot[cg * WIDE_TM + r] = dot;
            if (MODE == WIDE_GATE_BWD) sDot2[cg * WIDE_TM + r] = dot2;
            wide_epi_sync();
            const float full = ((sDot[r] + sDot[WIDE_TM + r]) + sDot[2 * WIDE_TM + r]) + sDot[3 * WIDE_TM + r];
            if (MODE == WIDE_ROWDOT || MODE == WIDE_ROWDOT_BWD) {
                if (cg == 0 && live) a.out[(size_t)nt * a.out_ld + e] = full;
            } else if (MODE == WIDE_GATE) {
                // second pass over the accumulator (tcgen05.ld is warp-collective: every lane runs it, rows past E only skip the store)
                const float gate = tc_sigmoid<false>(full + a.scalar[0]);
                float* o2 = a.out + (size_t)e * WIDE_TN + c0;
#pragma unroll 1
                for (int j = 0; j < 8; ++j) {
                    float v[8];
                    wide_ld8(taddr, j, c0, ncols, v);
#pragma unroll
                    for (int i = 0; i < 8; ++i) v[i] = gate * tc_silu<true>(v[i] + sBias[c0 + 8 * j + i]);
                    if (live) {
                        *reinterpret_cast<float4*>(o2 + 8 * j) = make_float4(v[0], v[1], v[2], v[3]);
                        *reinterpret_cast<float4*>(o2 + 8 * j + 4) = make_float4(v[4], v[5], v[6], v[7]);
                    }
                }
            } else {
                // GATE_BWD: g = sigmoid(wa.m + ba); t = gM[i].m; gs = t g (1-g); gz2 = (gM[i] g + gs wa) SiLU'(z2); rows = gs m
                const float tt = ((sDot2[r] + sDot2[WIDE_TM + r]) + sDot2[2 * WIDE_TM + r]) + sDot2[3 * WIDE_TM + r];
                const float gate = tc_sigmoid<false>(full + a.scalar[0]);
                const float gs = live ? tt * gate * (1.f - gate) : 0.f;
                if (cg == 0) a.q[e] = gs;
                float* ro = a.rows + (size_t)e * a.rows_ld + c0;
#pragma unroll 1
                for (int j = 0; j < 8; ++j) {
                    float v[8], g2[8];
                    wide_ld8(taddr, j, c0, ncols, v);
#pragma unroll
                    for (int i = 0; i < 8; ++i) {
                        float m, dm;
                        tc_silu_and_grad(v[i] + sBias[c0 + 8 * j + i], m, dm);
                        const int c = c0 + 8 * j + i;
                        const float gm = c < a.N && live ? gMi[c] : 0.f;
                        g2[i] = fmaf(gm, gate, gs * sVec[c]) * dm;
                        v[i] = gs * m;
                    }
                    *reinterpret_cast<float4*>(ro + 8 * j) = make_float4(v[0], v[1], v[2], v[3]);
                    *reinterpret_cast<float4*>(ro + 8 * j + 4) = make_float4(v[4], v[5], v[6], v[7]);
                    wide_store_chunk(o, j, g2);
                    wide_store_T(t, (c0 % WIDE_TM) + 8 * j, WIDE_A_SLICE / 2, g2, lane);
                }
            }
        } else if (MODE == WIDE_DGRAD) {
            uint8_t* o = a.out_img + ((size_t)mt * a.out_ks + n0 / WIDE_TK) * WIDE_A_SLICE + r * 16;
            uint8_t* t = a.out_T + ((size_t)(n0 / WIDE_TM) * a.kst + e / WIDE_TK) * WIDE_A_SLICE + ((e % WIDE_TK) / 8) * (WIDE_TM * 16);
            const float* d1 = a.rows + (size_t)e * a.rows_ld + n0;
#pragma unroll 1
            for (int j = 0; j < 8; ++j) {
                float v[8];
                wide_ld8(taddr, j, c0, ncols, v);
                const float4 da = *reinterpret_cast<const float4*>(d1 + 8 * j), db = *reinterpret_cast<const float4*>(d1 + 8 * j + 4);
                const float d[8] = {da.x, da.y, da.z, da.w, db.x, db.y, db.z, db.w};
#pragma unroll
                for (int i = 0; i < 8; ++i) v[i] = live ? v[i] * d[i] : 0.f;
                wide_store_chunk(o, j, v);
                wide_store_T(t, (c0 % WIDE_TM) + 8 * j, WIDE_A_SLICE / 2, v, lane);
            }
        } else if (MODE == WIDE_F32) {
            float* o = a.out + (size_t)e * a.out_ld + n0;
#pragma unroll 1
            for (int j = 0; j < 8; ++j) {
                float v[8];
                wide_ld8(taddr, j, c0, ncols, v);
                if (live) {
                    if (a.bias) {
#pragma unroll
                        for (int i = 0; i < 8; ++i) v[i] += n0 + 8 * j + i < a.N ? a.bias[n0 + 8 * j + i] : 0.f;
                    }
                    if (n0 + 8 * j + 8 <= a.N) {
                        wide_store8(o + 8 * j, v);
                    } else {
#pragma unroll
                        for (int i = 0; i < 8; ++i)
                            if (n0 + 8 * j + i < a.N) o[8 * j + i] = v[i];
                    }
                }
            }
        } else {  // WIDE_WGRAD: row e = feature n of the weight matrix, columns = its input features
            const int s = kb / a.KS;
            float* o = a.out + (size_t)s * a.out_split + (size_t)e * a.out_ld + n0;
#pragma unroll 1
            for (int j = 0; j < 8; ++j) {
                float v[8];
                wide_ld8(taddr, j, c0, ncols, v);
                if (live) {
                    if (n0 + 8 * j + 8 <= a.N) {
                        wide_store8(o + 8 * j, v);
                    } else {
#pragma unroll
                        for (int i = 0; i < 8; ++i)
                            if (n0 + 8 * j + i < a.N) o[8 * j + i] = v[i];
                    }
                }
            }
        }
        tc::fence_before_sync();
        __syncwarp();
        if (lane == 0) tc::mbar_arrive(&bars[6 + b]);  // this warp has read its part of the accumulator
        wide_epi_sync();                                // sBias / sVec / sDot are free for the next tile
    }
    if (!ok) asm volatile("trap;");  // an mbarrier wait timed out: fail loudly instead of returning garbage
    tc::fence_before_sync();
    __syncthreads();
    if (warp == 0) tc::tmem_dealloc(tmem, 512);
}

template <int MODE>
int launch_wide(const WideGemmArgs& a, int precision, cudaStream_t st, const char* what, int prof_kind) {
    const int tiles = a.mtiles * a.NT * a.splits;
    // persistent: one CTA per SM at most; the grid is rounded so that every CTA runs the same number of tiles when it can
    int grid = tiles < kB200SMs ? tiles : kB200SMs;
    if (tiles > kB200SMs) grid = wide_cdiv(tiles, wide_cdiv(tiles, kB200SMs));
    egnn_prof_begin(prof_kind, st);
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

WideGemmArgs wide_args(int mtiles, int E) {
    WideGemmArgs a{};
    a.mtiles = mtiles;
    a.E = E;
    a.splits = 1;
    a.ks_split = 1 << 30;
    return a;
}
// A image with KS slices per row tile, all of them used; B = KS slices per N tile
void wide_operands(WideGemmArgs& a, const uint8_t* A, int a_ks, const uint8_t* B, int KS, int NT, int N) {
    a.A = A;
    a.a_ks = a_ks;
    a.B = B;
    a.b_ks = KS;
    a.KS = a.KS_total = KS;
    a.NT = NT;
    a.N = N;
}

int wide_gather(const egnn_net_t& net, const egnn_graph_t& g, int mtiles, int ks0, const float* h, const float* x, uint8_t* U,
                uint8_t* UT, cudaStream_t st, int prof_kind) {
    const long long total = (long long)mtiles * WIDE_TM * ks0 * (WIDE_TK / 8);
    const int blocks = (int)((total + 255) / 256 < 148 * 16 ? (total + 255) / 256 : 148 * 16);
    egnn_prof_begin(prof_kind, st);
    wide_gather_kernel<<<blocks, 256, 0, st>>>(g.E, mtiles, net.H, ks0, g.recv, g.send, h, x, U, UT);
    egnn_prof_end(st);
    EGNN_CHECK_LAUNCH("wide_gather_kernel");
    return EGNN_OK;
}

uint8_t* wide_align(float* scratch) { return reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(scratch) + 1023) / 1024 * 1024); }

// split-K weight gradients / bias column sums of one phase (edges or nodes): rows [0, S) of `part`, parameter layout
struct WideGrad {
    int kst, per, S;   // 64-row slices, slices per split, splits
    float* part;
    int64_t pstride;
    int precision;
    cudaStream_t st;
    int prof;
    void colsum_T(const uint8_t* T, int N, int off) const {
        egnn_prof_begin(prof, st);
        wide_colsum_T_kernel<<<dim3(wide_cdiv(N, WIDE_TM), S), 1024, 0, st>>>(T, kst, per, N, part + off, pstride);
        egnn_prof_end(st);
    }
    void colsum_rows(const float* r, int ld, int ncols, int nrows, int off) const {
        egnn_prof_begin(prof, st);
        if (ncols == 1 && ld == 1)
            wide_sum_vec_kernel<<<S, 1024, 0, st>>>(r, nrows, per * WIDE_TK, part + off, pstride);
        else if (ld % 4 == 0 && ncols >= 64)
            wide_colsum_rows4_kernel<<<dim3(wide_cdiv(ncols, 128), S), 1024, 0, st>>>(r, ld, ncols, nrows, per * WIDE_TK, part + off, pstride);
        else
            wide_colsum_rows_kernel<<<dim3(wide_cdiv(ncols, 32), S), 256, 0, st>>>(r, ld, ncols, nrows, per * WIDE_TK, part + off, pstride);
        egnn_prof_end(st);
    }
    // dW [Nout x Kin] = G^T X: A = transposed G image (128-row feature tiles), B = transposed X image (256-row feature tiles)
    int wgrad(const uint8_t* GT, int Nout, const uint8_t* XT, int Kin, int off, const char* what) const {
        WideGemmArgs a = wide_args(wide_cdiv(Nout, WIDE_TM), Nout);
        a.A = GT; a.a_ks = kst; a.B = XT; a.b_ks = kst;
        a.KS = per; a.KS_total = kst; a.splits = S;
        a.NT = wide_cdiv(Kin, WIDE_TN); a.N = Kin;
        a.out = part + off; a.out_ld = Kin; a.out_split = pstride;
        return launch_wide<WIDE_WGRAD>(a, precision, st, what, EGNN_K_TC_WGRAD);
    }
};

int wide_rows_to_img(int N, int mtiles, int W1, int W2, int KS, const float* src1, const float* src2, int ld2, float* copy2, uint8_t* img,
                     uint8_t* T, int t_rows, int kst, cudaStream_t st, int prof, const float* part = nullptr,
                     const int32_t* row_ptr = nullptr, int TE = 1) {
    const long long total = (long long)mtiles * WIDE_TM * KS * (WIDE_TK / 8);
    const int blocks = (int)((total + 255) / 256 < 148 * 16 ? (total + 255) / 256 : 148 * 16);
    egnn_prof_begin(prof, st);
    wide_rows_to_img_kernel<<<blocks, 256, 0, st>>>(N, mtiles, W1, W2, KS, src1, src2, ld2, part, row_ptr, TE, copy2, img, T, t_rows, kst);
    egnn_prof_end(st);
    EGNN_CHECK_LAUNCH("wide_rows_to_img_kernel");
    return EGNN_OK;
}

}  // namespace

int egnn_wide_pack(const egnn_net_t& net, const float* params, float* packed, cudaStream_t st) {
    if (!wide_any(net)) return EGNN_OK;
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
    const WideWs ww = wide_ws(net, g.E, g.N);
    uint8_t* base = wide_align(scratch);
    uint8_t* U = base + ww.u;
    uint8_t* A1 = base + ww.a1;
    float* gm = reinterpret_cast<float*>(base + ww.gm);
    float* phip = reinterpret_cast<float*>(base + ww.phip);
    const uint8_t* img = reinterpret_cast<const uint8_t*>(packed_layer + po.wide);
    auto blk = [&](int first) { return img + (size_t)first * WIDE_B_SLICE; };
    const float* P = params_layer;
    const int K = EGNN_K_TC_EDGE_FWD;
    if (int rc = wide_gather(net, g, ww.mtiles, wo.ks0, h, x, U, nullptr, st, K)) return rc;
    WideGemmArgs a = wide_args(ww.mtiles, g.E);
    // message branch
    wide_operands(a, U, wo.ks0, blk(wo.m1), wo.ks0, wo.ntm1, net.Mh);
    a.bias = P + lo.bm1; a.out_img = A1; a.out_ks = ww.a1_ks;
    if (int rc = launch_wide<WIDE_IMG>(a, precision, st, "tc_wide_gemm m1", K)) return rc;
    wide_operands(a, A1, ww.a1_ks, blk(wo.m2), wo.ksm, wo.ntm2, net.Ms);
    a.bias = P + lo.bm2; a.vec = P + lo.wa; a.scalar = P + lo.ba; a.out = gm;
    if (int rc = launch_wide<WIDE_GATE>(a, precision, st, "tc_wide_gemm m2", K)) return rc;
    egnn_prof_begin(K, st);
    wide_segsum_kernel<<<g.N, WIDE_TN, 0, st>>>(net.Ms, g.row_ptr, gm, agg, net.Ms + 4);
    egnn_prof_end(st);
    EGNN_CHECK_LAUNCH("wide_segsum_kernel");
    // coordinate branch
    wide_operands(a, U, wo.ks0, blk(wo.x1), wo.ks0, wo.ntx1, net.Xh);
    a.bias = P + lo.bx1; a.out_img = A1; a.out_ks = ww.a1_ks;
    if (int rc = launch_wide<WIDE_IMG>(a, precision, st, "tc_wide_gemm x1", K)) return rc;
    wide_operands(a, A1, ww.a1_ks, blk(wo.x2), wo.ksx, wo.ntx2, net.Xh);
    a.bias = P + lo.bx2; a.vec = P + lo.wx3; a.scalar = nullptr; a.out = phip; a.out_ld = g.E;
    if (int rc = launch_wide<WIDE_ROWDOT>(a, precision, st, "tc_wide_gemm x2", K)) return rc;
    egnn_prof_begin(K, st);
    wide_phi_kernel<<<(g.E + 255) / 256, 256, 0, st>>>(g.E, wo.ntx2, g.E, phip, P + lo.bx3, phi_edge);
    egnn_prof_end(st);
    EGNN_CHECK_LAUNCH("wide_phi_kernel");
    return EGNN_OK;
}

int egnn_wide_bwd_rows(const egnn_net_t& net, int64_t E) { return wide_bwd_ws(net, E).splits; }
int64_t egnn_wide_bwd_bytes(const egnn_net_t& net, int64_t E) { return wide_bwd_ws(net, E).total + 1024; }

// Edge phase of one layer's backward on the tile GEMMs: recompute both branches (keeping SiLU' and the transposed images),
// gz2 from the gate / phi gradients, dgrad and wgrad GEMMs, bias gradients as column sums.  Weight-gradient partials go to
// rows [0, egnn_wide_bwd_rows) of `part` in the parameter layout (summed by reduce_partials_kernel); per-edge receiver /
// sender contributions to Gr / Gs as the FFMA kernel leaves them (egcl_gather_bwd sums them per node).
int egnn_wide_edge_backward(const egnn_net_t& net, const egnn_graph_t& g, const float* params_layer, const float* packed_layer,
                            const float* h, const float* x, const float* gM, const float* gdx, const float* coefS, float* Gr,
                            float* Gs, float* part, float* scratch, int precision, cudaStream_t st) {
    EGNN_REQUIRE(wide_fwd_supported(net), "egnn_wide_edge_backward: network outside the wide path's limits");
    const LayerOff lo = layer_off(net);
    const PackOff po = pack_off(net);
    const WideOff wo = wide_off(net);
    const WideBwdWs w = wide_bwd_ws(net, g.E);
    uint8_t* base = wide_align(scratch);
    auto f32 = [&](int64_t off) { return reinterpret_cast<float*>(base + off); };
    uint8_t *U = base + w.u, *UT = base + w.ut, *A1 = base + w.a1, *A1T = base + w.a1t, *G2 = base + w.g2, *G2T = base + w.g2t;
    uint8_t *G1m = base + w.g1m, *G1x = base + w.g1x, *G1T = base + w.g1t;
    float *D1 = f32(w.d1), *rows = f32(w.rows), *q = f32(w.q), *gphi = f32(w.gphi), *phip = f32(w.phip), *phi = f32(w.phi), *gu = f32(w.gu);
    const uint8_t* img = reinterpret_cast<const uint8_t*>(packed_layer + po.wide);
    auto blk = [&](int first) { return img + (size_t)first * WIDE_B_SLICE; };
    const float* P = params_layer;
    const int K = EGNN_K_TC_EDGE_BWD, K0 = 2 * net.H + 1;
    const int Ep = w.mtiles * WIDE_TM, S = w.splits;
    const int64_t pstride = lo.total;

    const WideGrad G{w.kst, w.per, S, part, pstride, precision, st, K};

    if (int rc = wide_gather(net, g, w.mtiles, wo.ks0, h, x, U, UT, st, K)) return rc;
    egnn_prof_begin(K, st);
    wide_gphi_kernel<<<wide_cdiv(Ep, 256), 256, 0, st>>>(g.E, Ep, g.recv, g.send, x, gdx, gphi);
    egnn_prof_end(st);
    EGNN_CHECK_LAUNCH("wide_gphi_kernel");

    for (int branch = 0; branch < 2; ++branch) {  // 0: message, 1: coordinate
        const bool M = branch == 0;
        const int D1w = M ? net.Mh : net.Xh, D2w = M ? net.Ms : net.Xh;
        const int ks1 = M ? wo.ksm : wo.ksx, ks2 = M ? wo.kss : wo.ksx;
        uint8_t* G1 = M ? G1m : G1x;
        // recompute the first Linear: a1 image, its transpose, SiLU'(z1) rows
        WideGemmArgs a = wide_args(w.mtiles, g.E);
        wide_operands(a, U, wo.ks0, blk(M ? wo.m1 : wo.x1), wo.ks0, wide_cdiv(D1w, WIDE_TN), D1w);
        a.bias = P + (M ? lo.bm1 : lo.bx1); a.out_img = A1; a.out_ks = w.w_ks; a.out_T = A1T; a.kst = w.kst; a.rows = D1; a.rows_ld = w.np;
        if (int rc = launch_wide<WIDE_IMG_BWD>(a, precision, st, "tc_wide_gemm bwd 1", K)) return rc;
        // recompute the second Linear and turn it into gz2 (image + transpose); rows = the row-dot vector's gradient terms
        wide_operands(a, A1, w.w_ks, blk(M ? wo.m2 : wo.x2), ks1, wide_cdiv(D2w, WIDE_TN), D2w);
        a.bias = P + (M ? lo.bm2 : lo.bx2); a.vec = P + (M ? lo.wa : lo.wx3); a.scalar = M ? P + lo.ba : nullptr;
        a.out_img = G2; a.out_ks = w.w_ks; a.out_T = G2T; a.rows = rows; a.rows_ld = w.np;
        a.q = q; a.gM = gM; a.recv = g.recv; a.gphi = gphi; a.out = phip; a.out_ld = Ep;
        if (M) {
            if (int rc = launch_wide<WIDE_GATE_BWD>(a, precision, st, "tc_wide_gemm bwd gate", K)) return rc;
        } else {
            if (int rc = launch_wide<WIDE_ROWDOT_BWD>(a, precision, st, "tc_wide_gemm bwd phi", K)) return rc;
            egnn_prof_begin(K, st);
            wide_phi_kernel<<<wide_cdiv(g.E, 256), 256, 0, st>>>(g.E, wo.ntx2, Ep, phip, P + lo.bx3, phi);
            egnn_prof_end(st);
            EGNN_CHECK_LAUNCH("wide_phi_kernel");
        }
        // d vec = column sums of the rows; d (its bias) = sum of the row coefficients
        G.colsum_rows(rows, w.np, D2w, g.E, M ? lo.wa : lo.wx3);
        G.colsum_rows(M ? q : gphi, 1, 1, g.E, M ? lo.ba : lo.bx3);
        // second Linear: dW2 = gz2^T a1, db2 = column sums of gz2
        if (int rc = G.wgrad(G2T, D2w, A1T, D1w, M ? lo.wm2 : lo.wx2, "tc_wide_gemm wgrad 2")) return rc;
        G.colsum_T(G2T, D2w, M ? lo.bm2 : lo.bx2);
        // ga1 = gz2 . W2;  gz1 = ga1 * SiLU'(z1): image (for gu) + transpose (for dW1)
        WideGemmArgs d = wide_args(w.mtiles, g.E);
        wide_operands(d, G2, w.w_ks, blk(M ? wo.m2t : wo.x2t), ks2, wide_cdiv(D1w, WIDE_TN), D1w);
        d.out_img = G1; d.out_ks = w.w_ks; d.out_T = G1T; d.kst = w.kst; d.rows = D1; d.rows_ld = w.np;
        if (int rc = launch_wide<WIDE_DGRAD>(d, precision, st, "tc_wide_gemm dgrad 2", K)) return rc;
        // first Linear: dW1 = gz1^T u, db1 = column sums of gz1
        if (int rc = G.wgrad(G1T, D1w, UT, K0, M ? lo.wm1 : lo.wx1, "tc_wide_gemm wgrad 1")) return rc;
        G.colsum_T(G1T, D1w, M ? lo.bm1 : lo.bx1);
    }
    EGNN_CHECK_LAUNCH("wide colsum kernels");
    {   // gu = gz1m . Wm1 + gz1x . Wx1: one GEMM over the concatenated K
        WideGemmArgs a = wide_args(w.mtiles, g.E);
        a.A = G1m; a.a_ks = w.w_ks; a.A2 = G1x; a.a2_ks = w.w_ks; a.ks_split = wo.ksm;
        a.B = blk(wo.w1t); a.b_ks = wo.ksm + wo.ksx; a.KS = a.KS_total = wo.ksm + wo.ksx; a.NT = 1; a.N = K0;
        a.out = gu; a.out_ld = 128;
        if (int rc = launch_wide<WIDE_F32>(a, precision, st, "tc_wide_gemm dgrad 1", K)) return rc;
    }
    const long long tot = (long long)g.E * (net.H + 3);
    egnn_prof_begin(K, st);
    wide_edge_out_kernel<<<(unsigned)((tot + 255) / 256), 256, 0, st>>>(g.E, net.H, g.recv, g.send, g.node_graph, x, gu, gdx, coefS, phi, Gr, Gs);
    egnn_prof_end(st);
    EGNN_CHECK_LAUNCH("wide_edge_out_kernel");
    return EGNN_OK;
}

// ---- node phase ---------------------------------------------------------------------------------------------------------------
// h' = mlp_h([h, M]) (EquivariantGraphNeuralNetwork.py:69, no residual) on the tile GEMMs, plus the coordinate sums from phi_e.
// agg (row stride Ms + 4) holds complete per-receiver sums (egnn_wide_edge_forward, TE = 2^30) or, after a fused edge kernel,
// the sums of receivers inside one TE-edge segment with head / tail partials in part; the combined sums are saved to msum.
int egnn_wide_node_forward(const egnn_net_t& net, const egnn_graph_t& g, const float* params_layer, const float* packed_layer,
                           const float* h, const float* x, const float* agg, const float* part, int TE, const float* phi_edge,
                           float* h_out, float* msum, float* dxu, float* Dn, float* scratch, int precision, cudaStream_t st) {
    EGNN_REQUIRE(wide_node_supported(net), "egnn_wide_node_forward: network outside the wide path's limits");
    const LayerOff lo = layer_off(net);
    const PackOff po = pack_off(net);
    const WideOff wo = wide_off(net);
    const WideWs ww = wide_ws(net, g.E, g.N);
    uint8_t* base = wide_align(scratch);
    uint8_t *NI = base + ww.ni, *A1 = base + ww.a1n;
    const uint8_t* img = reinterpret_cast<const uint8_t*>(packed_layer + po.wide);
    auto blk = [&](int first) { return img + (size_t)first * WIDE_B_SLICE; };
    const int K = EGNN_K_NODE_FWD, a1_ks = wo.nth * (WIDE_TN / WIDE_TK);
    egnn_prof_begin(K, st);
    wide_node_geom_kernel<<<wide_cdiv(g.N, 32), 256, 0, st>>>(g.N, g.row_ptr, g.send, x, phi_edge, dxu, Dn);
    egnn_prof_end(st);
    EGNN_CHECK_LAUNCH("wide_node_geom_kernel");
    if (int rc = wide_rows_to_img(g.N, ww.mtiles_n, net.H, net.Ms, wo.ksn, h, agg, net.Ms + 4, msum, NI, nullptr, 0, 0, st, K,
                                  TE < (1 << 30) ? part : nullptr, g.row_ptr, TE))
        return rc;
    WideGemmArgs a = wide_args(ww.mtiles_n, g.N);
    wide_operands(a, NI, wo.ksn, blk(wo.h1), wo.ksn, wo.nth, net.Hh);
    a.bias = params_layer + lo.bh1; a.out_img = A1; a.out_ks = a1_ks;
    if (int rc = launch_wide<WIDE_IMG>(a, precision, st, "tc_wide_gemm h1", K)) return rc;
    wide_operands(a, A1, a1_ks, blk(wo.h2), wo.ksh, 1, net.H);
    a.bias = params_layer + lo.bh2; a.out = h_out; a.out_ld = net.H;
    return launch_wide<WIDE_F32>(a, precision, st, "tc_wide_gemm h2", K);
}

int egnn_wide_node_bwd_rows(const egnn_net_t& net, int64_t N) { return wide_node_bwd_ws(net, N).splits; }
int64_t egnn_wide_node_bwd_bytes(const egnn_net_t& net, int64_t N) { return wide_node_bwd_ws(net, N).total + 1024; }

// backward of mlp_h (egcl_bwd.cu: egcl_node_bwd): gh_node = d/dh, gM = d/dM, weight-gradient partials in rows [0, rows) of part
int egnn_wide_node_backward(const egnn_net_t& net, int N, const float* params_layer, const float* packed_layer, const float* h,
                            const float* msum, const float* gh_out, float* gh_node, float* gM, float* part, float* scratch,
                            int precision, cudaStream_t st) {
    EGNN_REQUIRE(wide_node_supported(net), "egnn_wide_node_backward: network outside the wide path's limits");
    const LayerOff lo = layer_off(net);
    const PackOff po = pack_off(net);
    const WideOff wo = wide_off(net);
    const WideNodeBwdWs w = wide_node_bwd_ws(net, N);
    uint8_t* base = wide_align(scratch);
    uint8_t *NI = base + w.ni, *NIT = base + w.nit, *A1 = base + w.a1, *A1T = base + w.a1t, *GO = base + w.go, *GOT = base + w.got;
    uint8_t *G1 = base + w.g1, *G1T = base + w.g1t;
    float *D1 = reinterpret_cast<float*>(base + w.d1), *gin = reinterpret_cast<float*>(base + w.gin);
    const uint8_t* img = reinterpret_cast<const uint8_t*>(packed_layer + po.wide);
    auto blk = [&](int first) { return img + (size_t)first * WIDE_B_SLICE; };
    const int K = EGNN_K_NODE_BWD, Kn = net.H + net.Ms;
    const WideGrad G{w.kst, w.per, w.splits, part, (int64_t)lo.total, precision, st, K};
    // [h | M] image + transpose; gh_out image + transpose (gz2 = gh_out: the second Linear has no activation)
    if (int rc = wide_rows_to_img(N, w.mtiles, net.H, net.Ms, wo.ksn, h, msum, net.Ms, nullptr, NI, NIT, WIDE_TN, w.kst, st, K)) return rc;
    if (int rc = wide_rows_to_img(N, w.mtiles, net.H, 0, wo.ksH, gh_out, nullptr, 0, nullptr, GO, GOT, WIDE_TM, w.kst, st, K)) return rc;
    // recompute the first Linear: a1 image + transpose + SiLU' rows
    WideGemmArgs a = wide_args(w.mtiles, N);
    wide_operands(a, NI, wo.ksn, blk(wo.h1), wo.ksn, wo.nth, net.Hh);
    a.bias = params_layer + lo.bh1; a.out_img = A1; a.out_ks = w.h_ks; a.out_T = A1T; a.kst = w.kst; a.rows = D1; a.rows_ld = w.hp;
    if (int rc = launch_wide<WIDE_IMG_BWD>(a, precision, st, "tc_wide_gemm node bwd 1", K)) return rc;
    // dWh2 = gh_out^T a1, dbh2 = column sums of gh_out
    if (int rc = G.wgrad(GOT, net.H, A1T, net.Hh, lo.wh2, "tc_wide_gemm node wgrad 2")) return rc;
    G.colsum_T(GOT, net.H, lo.bh2);
    // ga1 = gh_out . Wh2;  gz1 = ga1 * SiLU'(z1)
    WideGemmArgs d = wide_args(w.mtiles, N);
    wide_operands(d, GO, wo.ksH, blk(wo.h2t), wo.ksH, wo.nth, net.Hh);
    d.out_img = G1; d.out_ks = w.h_ks; d.out_T = G1T; d.kst = w.kst; d.rows = D1; d.rows_ld = w.hp;
    if (int rc = launch_wide<WIDE_DGRAD>(d, precision, st, "tc_wide_gemm node dgrad 2", K)) return rc;
    // dWh1 = gz1^T [h | M], dbh1 = column sums of gz1
    if (int rc = G.wgrad(G1T, net.Hh, NIT, Kn, lo.wh1, "tc_wide_gemm node wgrad 1")) return rc;
    G.colsum_T(G1T, net.Hh, lo.bh1);
    // g[h | M] = gz1 . Wh1
    WideGemmArgs f = wide_args(w.mtiles, N);
    wide_operands(f, G1, w.h_ks, blk(wo.h1t), wo.ksh, wo.ntn, Kn);
    f.out = gin; f.out_ld = w.gin_ld;
    if (int rc = launch_wide<WIDE_F32>(f, precision, st, "tc_wide_gemm node dgrad 1", K)) return rc;
    const long long tot = (long long)N * Kn;
    egnn_prof_begin(K, st);
    wide_split_gin_kernel<<<(unsigned)((tot + 255) / 256), 256, 0, st>>>(N, net.H, net.Ms, gin, w.gin_ld, gh_node, gM);
    egnn_prof_end(st);
    EGNN_CHECK_LAUNCH("wide_split_gin_kernel");
    return EGNN_OK;
}
