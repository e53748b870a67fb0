// EGCL edge BACKWARD on tcgen05 for hidden width d = 256 (BASELINE configs[3]).  At this width TMEM (512 columns) is
// filled by z1 | z2 of ONE 128-edge tile, so the design of tc_edge_bwd (weight gradients persistent in TMEM over the
// CTA's life) does not fit.  The adjoint is split in two kernels per layer instead:
//
//   tc_edge_bwd256 (one branch per CTA, 128-edge tiles, 512 threads)
//     recompute   z1 = U.W1^T -> a1 = SiLU -> tile R;  z2 = a1.W2^T with W2 streamed in 16 K-slices (cp.async.bulk ring)
//     upstream    gz2 = dL/dz2 (gate / phi algebra of SURVEY.md Appendix A on the CUDA cores) -> tile R
//     dgrad       ga1 = gz2.W2 with the TRANSPOSED image W2T streamed the same way (over z2 in TMEM)
//     gz1 = ga1 * SiLU'(z1) -> tile R;  gu = gz1.W1 (16 columns) -> per-edge outputs Gr / Gs
//     [dW1 | db1] of the tile = gz1^T.u (two M = 128 halves, 16 columns each, next to gu) is read back and summed over the
//     CTA's tiles in 8 registers per thread -- the narrow product needs no image;
//     the operand tiles of the wide product -- a1 (hi + lo), gz2 (hi + lo), u (hi + lo) -- leave shared memory as ready-made
//     tcgen05 tile images through bulk async stores (no transposition, no second layout): 270 KB per tile and branch, which
//     makes this kernel HBM-WRITE bound for a third of its time (the stores run at the SM's share of the HBM bandwidth);
//     d wa / d wx3 / d ba / d bx3 are register column sums written to the CTA's private partial row.
//   tc_wgrad256 (split-K GEMM over the edge tiles; CTA = branch x half of the output rows x K split)
//     [dW2 | db2] += gz2^T.[a1 | u]   (tile images viewed MN-major: K = the 128 edges of a tile; slot 15 of u is the
//     constant 1, so the bias gradient is an MMA column); accumulators persistent in TMEM over the CTA's tiles, written
//     once to its partial row.  This kernel is HBM-bound by construction (it streams the images).
// Conventions as in tc_edge.cu: weight images are stored halved, gradient tiles carry 2 dL/dz, results are halved once.
#include <cstdlib>

#include "tc_bulk.cuh"
#include <algorithm>
#include <cmath>
#include <cstdlib>
#include "tc_tile.cuh"

#ifdef EGNN_DBG_TIMING  // phase time stamps of tile #3 of the first (mlp_m) and last (mlp_x) CTA, scripts/dbg_phases256.py
__device__ long long g_dbg_stamp_bwd[2][16];
#define DBG_STAMP_T(k) do { if (it == 3 && (blockIdx.x == 0 || blockIdx.x == gridDim.x - 1)) g_dbg_stamp_bwd[blockIdx.x ? 1 : 0][k] = clock64(); } while (0)
#define DBG_STAMP(k) do { if (it == 3 && tid == 0 && (blockIdx.x == 0 || blockIdx.x == gridDim.x - 1)) g_dbg_stamp_bwd[blockIdx.x ? 1 : 0][k] = clock64(); } while (0)
extern "C" int egnn_dbg_read_stamps_bwd(long long* host16x2) {
    return cudaMemcpyFromSymbol(host16x2, g_dbg_stamp_bwd, sizeof(g_dbg_stamp_bwd)) == cudaSuccess ? 0 : -2;
}
#else
#define DBG_STAMP(k) do { } while (0)
#define DBG_STAMP_T(k) do { } while (0)
#endif

namespace {

constexpr int D = 256, NT = 512;
constexpr int NSLICE = 16, NSLOT = 4, SLICE_B = D * 16 * 2;
constexpr int TILE_B = 128 * D * 2;  // one bf16 [128 x 256] tile (hi or lo)
constexpr int U_B = 128 * 32;

// bytes of tile images per 128-edge tile and branch:  a1 hi | a1 lo | gz2 hi | gz2 lo | u hi | u lo
constexpr int IMG_A1 = 0, IMG_G2 = 2 * TILE_B, IMG_U = 4 * TILE_B, IMG_TILE = 4 * TILE_B + 2 * U_B;

struct LB {
    static constexpr int w1_hi = 0, w1_lo = D * 32;
    static constexpr int u_hi = 2 * D * 32, u_lo = u_hi + U_B;
    static constexpr int R_hi = u_lo + U_B, R_lo = R_hi + TILE_B;  // a1 -> gz2 -> gz1
    static constexpr int ring = R_lo + TILE_B;
    static constexpr int vec = ring + NSLOT * 2 * SLICE_B;  // b2/2 [256] | wv [256] | bv, 0, 0, 0
    static constexpr int rowA = vec + (2 * D + 4) * 4;      // float [4][128]
    static constexpr int rowB = rowA + 2048;                // float [4][128]
    static constexpr int sI = rowB + 2048;                  // int [128]
    static constexpr int sR = sI + 512;                     // float4 [128]
    static constexpr int bars = sR + 2048;                  // 0 GEMM1, 1 GEMM2, 2 dgrad, 3 gu, 4..7 full, 8..11 empty
    static constexpr int slot = bars + 96;
    static constexpr int TOTAL = slot + 16;
    static_assert(TOTAL <= 227 * 1024, "tc_edge_bwd256 shared memory");
};

struct Bwd256Args {
    egnn_net_t net;
    int N, E, tiles;
    const int32_t *recv, *send, *node_graph;
    const float *h, *x;
    const uint8_t* blob;  // this layer's weight image; w2t = its transposed-W2 part
    const uint8_t* w2t;   // [branch][hi | lo] K-major images of W2^T / 2
    const float *gM, *gdx, *coefS;
    float *Gr, *Gs;       // [2][E][H+3]
    float* part;          // [rows][layer param count]
    uint8_t* img;         // [branch][tiles][IMG_TILE]
    int nM;
};

struct EdgeRow {
    int i;
    float rx, ry, rz, valid;
    float u[16];
};

template <int NPASS>
__global__ void __launch_bounds__(NT, 1) tc_edge_bwd256(Bwd256Args a) {
    using L = LB;
    constexpr bool LO = NPASS == 3;
    constexpr int cA = 0, cB = D;
    constexpr uint32_t CS = 128 * 16;
    extern __shared__ __align__(1024) uint8_t smem[];
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int q = warp & 3, cg = warp >> 2;
    const int r = q * 32 + lane;
    const int c0 = cg * 64;
    const int branch = (int)blockIdx.x < a.nM ? 0 : 1;
    const int H = a.net.H, W = H + 3;
    float* vec = reinterpret_cast<float*>(smem + L::vec);
    float* rowA = reinterpret_cast<float*>(smem + L::rowA);
    float* rowB = reinterpret_cast<float*>(smem + L::rowB);
    int* sI = reinterpret_cast<int*>(smem + L::sI);
    float4* sR = reinterpret_cast<float4*>(smem + L::sR);
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + L::bars);
    uint32_t* slot = reinterpret_cast<uint32_t*>(smem + L::slot);

    constexpr int W1B = 2 * D * 32, W2B = D * D * 2;
    {
        for (int o = tid * 16; o < 4 * D * 16; o += NT * 16) {
            const int half = o / (2 * D * 16), j = (o / (D * 16)) & 1, off = o % (D * 16);
            *reinterpret_cast<uint4*>(smem + (half ? L::w1_lo : L::w1_hi) + j * D * 16 + off) =
                __ldg(reinterpret_cast<const uint4*>(a.blob + half * W1B + j * 2 * D * 16 + branch * D * 16 + off));
        }
        const float* vsrc = reinterpret_cast<const float*>(a.blob + 2 * W1B + 4 * W2B);
        for (int c = tid; c < D; c += NT) {
            vec[c] = vsrc[2 * D + branch * D + c];
            vec[D + c] = vsrc[4 * D + branch * D + c];
        }
        if (tid == 0) vec[2 * D] = vsrc[6 * D + branch];
    }
    const uint8_t* w2_g = a.blob + 2 * W1B + (size_t)branch * 2 * W2B;   // hi | lo, K-major over k
    const uint8_t* w2t_g = a.w2t + (size_t)branch * 2 * W2B;             // hi | lo, K-major over n
    if (warp == 0) tc::tmem_alloc(slot, 512);
    if (tid == 0) {
        for (int b = 0; b < 4 + 2 * NSLOT; ++b) tc::mbar_init(&bars[b], 1);
        tc::fence_mbar_init();
    }
    const uint32_t sb = tc::smem_u32(smem);
    const TcOperand opU{sb + L::u_hi, sb + L::u_lo, CS, false};
    const TcOperand opR{sb + L::R_hi, sb + L::R_lo, CS, false};
    const TcOperand opW1{sb + L::w1_hi, sb + L::w1_lo, D * 16, false}, opW1mn{sb + L::w1_hi, sb + L::w1_lo, D * 16, true};
    const float* b2 = vec;
    const float* wv = vec + D;
    float* Grb = a.Gr + (size_t)branch * a.E * W;
    float* Gsb = a.Gs + (size_t)branch * a.E * W;

    const int tstride = branch ? (int)gridDim.x - a.nM : a.nM;
    const int tile0 = branch ? (int)blockIdx.x - a.nM : (int)blockIdx.x;
    const int nit = tile0 < a.tiles ? (a.tiles - tile0 + tstride - 1) / tstride : 0;
    const int total_slices = nit * 2 * NSLICE;  // per tile: 16 slices of W2, then 16 slices of W2T

    auto gather_load = [&](int tile, EdgeRow& g) {
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
            g.u[15] = 1.f;
        }
    };
    auto store_u = [&](const EdgeRow& g) {
#pragma unroll
        for (int j = 0; j < 2; ++j) {
            uint32_t hi[4], lo[4];
#pragma unroll
            for (int p = 0; p < 4; ++p) tc::split2(g.u[8 * j + 2 * p], g.u[8 * j + 2 * p + 1], hi[p], lo[p]);
            *reinterpret_cast<uint4*>(smem + L::u_hi + tc::chunk_off(128, tid, j)) = make_uint4(hi[0], hi[1], hi[2], hi[3]);
            *reinterpret_cast<uint4*>(smem + L::u_lo + tc::chunk_off(128, tid, j)) = make_uint4(lo[0], lo[1], lo[2], lo[3]);
        }
    };
    auto store_row = [&](const EdgeRow& g) {
        sI[tid] = g.i;
        sR[tid] = make_float4(g.rx, g.ry, g.rz, g.valid);
    };
    auto load_slice = [&](int g) {
#ifdef EGNN_DBG_NOSTREAM
        return;
#endif
        const int sl = g % (2 * NSLICE), st = g % NSLOT;
        const uint8_t* src = sl < NSLICE ? w2_g + (size_t)sl * SLICE_B : w2t_g + (size_t)(sl - NSLICE) * SLICE_B;
        const uint32_t full = tc::smem_u32(&bars[4 + st]);
        const uint32_t dst = sb + L::ring + st * 2 * SLICE_B;
        tc::mbar_expect_tx(full, LO ? 2 * SLICE_B : SLICE_B);
        tc::bulk_g2s(dst, src, SLICE_B, full);
        if (LO) tc::bulk_g2s(dst + SLICE_B, src + W2B, SLICE_B, full);
    };

    if (tid < 128 && nit > 0) {
        EdgeRow g;
        gather_load(tile0, g);
        store_u(g);
        store_row(g);
    }
    tc::fence_async_smem();
    tc::fence_before_sync();
    __syncthreads();
    tc::fence_after_sync();
    const uint32_t tmem = *slot;
    const uint32_t trow = tmem + ((uint32_t)(q * 32) << 16);
    const bool issuer_warp = warp == NT / 32 - 1;
    // ring producer = lane 0 of the warp before the issuer's (see tc_edge256.cu): the MMA issuer never waits for a free slot
    const bool producer = warp == NT / 32 - 2 && lane == 0;
    int gload = 0;
    bool ok = true;
    auto produce = [&](int limit) {
        for (; gload < limit; ++gload) {
            // slots are released in pairs (one tcgen05.commit per two slices, see tc_edge256.cu)
            if (gload >= NSLOT && (gload & 1) == 0)
                ok = tc::mbar_wait(&bars[4 + NSLOT + (gload / 2) % 2], ((gload - NSLOT) / NSLOT) & 1) && ok;
            load_slice(gload);
        }
    };
    if (producer && nit > 0) produce(min(total_slices, NSLOT));
    if (issuer_warp && nit > 0 && tc::elect_one()) {
        tc_issue_op<NPASS>(tmem, cA, opU, opW1, 16, D, false);
        tc::mma_commit(&bars[0]);
    }
    // one elected thread: stream 16 slices into D[128 x 256] at dcol, A = tile R (k-step s = slice s)
    auto stream_gemm = [&](int gbase, int dcol, bool a_lo, bool& ok) {
        const uint32_t idesc = tc::make_idesc_bf16(128, D);
        for (int s = 0; s < NSLICE; ++s) {
            const int g = gbase + s, st = g % NSLOT;
#ifndef EGNN_DBG_NOSTREAM  // timing experiment only (wrong results): MMAs without the weight stream
            ok = tc::mbar_wait(&bars[4 + st], (g / NSLOT) & 1) && ok;
#endif
            const uint32_t wb = sb + L::ring + st * 2 * SLICE_B;
            const uint64_t ah = tc::make_desc(sb + L::R_hi + s * 2 * CS, CS, 128), al = tc::make_desc(sb + L::R_lo + s * 2 * CS, CS, 128);
            const uint64_t bh = tc::make_desc(wb, D * 16, 128), bl = tc::make_desc(wb + SLICE_B, D * 16, 128);
            tc::mma_bf16(tmem + dcol, ah, bh, idesc, s > 0);
            if (LO) {
                if (a_lo) tc::mma_bf16(tmem + dcol, al, bh, idesc, true);
                tc::mma_bf16(tmem + dcol, ah, bl, idesc, true);
            }
            if (g & 1) tc::mma_commit(&bars[4 + NSLOT + (g / 2) % 2]);
        }
    };

    float accv0 = 0.f, accv1 = 0.f;  // running column sums of coef_e * out_e (d wa / d wx3): columns c0 + lane, c0 + 32 + lane
    float accs = 0.f;                // running sum of coef_e of this thread's row (d ba / d bx3)
    float accW1[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};  // running [dW1 | db1] entries of this thread (see the tile end)

    for (int it = 0; it < nit; ++it) {
        const int tile = tile0 + it * tstride;
        const int e0 = tile * 128, nrows = min(128, a.E - e0);
        const uint32_t phase = it & 1;
        const bool has_next = it + 1 < nit;
        uint8_t* img = a.img + ((size_t)branch * a.tiles + tile) * IMG_TILE;
        DBG_STAMP(0);
        ok = tc::mbar_wait(&bars[0], phase) && ok;
        tc::fence_after_sync();
        DBG_STAMP(1);
        // ---- a1 = SiLU(z1) -> tile R;  2 SiLU'(z1) replaces z1 in TMEM
        tc_epi_bwd_chunked(trow + cA + c0, smem + L::R_hi, smem + L::R_lo, r, c0, LO);
        tc_epi_bwd_chunked(trow + cA + c0 + 32, smem + L::R_hi, smem + L::R_lo, r, c0 + 32, LO);
        tc::fence_async_smem();
        tc::fence_before_sync();
        __syncthreads();
        DBG_STAMP(10);  // a1 barrier passed
        if (issuer_warp && tc::elect_one()) {
            tc::fence_after_sync();
            tc::bulk_s2g(img + IMG_A1, sb + L::R_hi, LO ? 2 * TILE_B : TILE_B);
            tc::bulk_s2g(img + IMG_U, sb + L::u_hi, 2 * U_B);
            tc::bulk_commit();
            // (draining the store before the MMAs start was measured: the store then takes 5.8 k cycles on its own -- 136 KB at
            //  the SM's share of the HBM write bandwidth -- and the sum stays the same; the overlap is kept)
            DBG_STAMP_T(11);  // image stores issued
            stream_gemm(it * 2 * NSLICE, cB, true, ok);  // z2 = a1 . W2^T
            tc::mma_commit(&bars[1]);
            DBG_STAMP_T(12);  // GEMM2 issued
        } else if (producer) {
            produce(min(total_slices, it * 2 * NSLICE + NSLICE + NSLOT));
        }
        __syncwarp();
        // ---- in the shadow of GEMM2: next tile's gather loads
        EdgeRow gnext;
        if (tid < 128 && has_next) gather_load(tile + tstride, gnext);
        DBG_STAMP(2);  // a1 epilogue + barrier + gather loads done
        ok = tc::mbar_wait(&bars[1], phase) && ok;
        tc::fence_after_sync();
        DBG_STAMP(3);  // GEMM2 observed
        // ---- upstream gradient of z2.  message branch: two passes over z2 (the gate needs the whole row first)
        const float4 rr = sR[r];
        const int i = sI[r];
        float c_row = 0.f, c_g = 0.f;
        if (branch == 0) {
            float sdot = 0.f, tdot = 0.f;
#pragma unroll 1
            for (int hh = 0; hh < 2; ++hh) {
                float v[32];
                tc::tmem_ld32(trow + cB + c0 + 32 * hh, v);
                const float4* gMi = reinterpret_cast<const float4*>(a.gM + (size_t)i * D + c0 + 32 * hh);
                const float* bb = b2 + c0 + 32 * hh;
                const float* ww = wv + c0 + 32 * hh;
#pragma unroll
                for (int c4 = 0; c4 < 8; ++c4) {
                    const float4 gm = __ldg(gMi + c4);
                    const float gmv[4] = {gm.x, gm.y, gm.z, gm.w};
#pragma unroll
                    for (int k = 0; k < 4; ++k) {
                        const float m = tc_silu_h(v[4 * c4 + k] + bb[4 * c4 + k]);
                        sdot = fmaf(ww[4 * c4 + k], m, sdot);
                        tdot = fmaf(gmv[k], m, tdot);
                    }
                }
            }
            rowA[cg * 128 + r] = sdot;
            rowB[cg * 128 + r] = tdot;
        } else {
            const float* gd = a.gdx + (size_t)i * 3;
            c_row = rr.w * (gd[0] * rr.x + gd[1] * rr.y + gd[2] * rr.z);
        }
        if (issuer_warp && tc::elect_one()) tc::bulk_wait_read();  // the a1 / u images have left shared memory
        __syncwarp();
        __syncthreads();
        DBG_STAMP(4);  // gate pass + barrier
        if (branch == 0) {
            const float ssum = vec[2 * D] + rowA[r] + rowA[128 + r] + rowA[256 + r] + rowA[384 + r];
            const float tsum = rowB[r] + rowB[128 + r] + rowB[256 + r] + rowB[384 + r];
            const float g = sigmoid_f(ssum);
            c_row = rr.w * tsum * g * (1.f - g);
            c_g = rr.w * g;
        }
        float pdot = 0.f;
#pragma unroll 1
        for (int hh = 0; hh < 2; ++hh) {
            float v[32], cs[32];
            tc::tmem_ld32(trow + cB + c0 + 32 * hh, v);
            const float4* gMi = reinterpret_cast<const float4*>(a.gM + (size_t)i * D + c0 + 32 * hh);
            const float* bb = b2 + c0 + 32 * hh;
            const float* ww = wv + c0 + 32 * hh;
#pragma unroll
            for (int c4 = 0; c4 < 8; ++c4) {
                float4 gm = make_float4(0.f, 0.f, 0.f, 0.f);
                if (branch == 0) gm = __ldg(gMi + c4);
                const float gmv[4] = {gm.x, gm.y, gm.z, gm.w};
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    const int c = 4 * c4 + k;
                    float m, d2;
                    tc_silu_and_grad2_h(v[c] + bb[c], m, d2);
                    pdot = fmaf(ww[c], m, pdot);
                    v[c] = fmaf(c_g, gmv[k], c_row * ww[c]) * d2;  // 2 gz2
                    cs[c] = c_row * m;
                }
            }
            tc_split_store32(smem + L::R_hi, smem + L::R_lo, r, c0 + 32 * hh, v, LO);
            warp_colsum32(cs, lane);
            if (hh == 0) accv0 += cs[0];
            else accv1 += cs[0];
        }
        accs += c_row;
        if (branch == 1) rowA[cg * 128 + r] = pdot;
        tc::fence_async_smem();
        tc::fence_before_sync();
        __syncthreads();
        if (issuer_warp && tc::elect_one()) {
            tc::fence_after_sync();
            tc::bulk_s2g(img + IMG_G2, sb + L::R_hi, LO ? 2 * TILE_B : TILE_B);
            tc::bulk_commit();
            stream_gemm(it * 2 * NSLICE + NSLICE, cB, LO, ok);  // ga1 = gz2 . W2  (over z2)
            tc::mma_commit(&bars[2]);
        } else if (producer) {
            produce(min(total_slices, (it + 1) * 2 * NSLICE + NSLOT));
        }
        __syncwarp();
        float phi = 0.f;
        if (branch == 1 && cg == 0) phi = vec[2 * D] + rowA[r] + rowA[128 + r] + rowA[256 + r] + rowA[384 + r];
        DBG_STAMP(5);  // gz2 pass + barrier
        ok = tc::mbar_wait(&bars[2], phase) && ok;
        tc::fence_after_sync();
        DBG_STAMP(6);  // dgrad observed
        if (issuer_warp && tc::elect_one()) tc::bulk_wait_read();  // the gz2 image has left tile R
        __syncwarp();
        __syncthreads();
        // ---- gz1 = ga1 * SiLU'(z1) -> tile R
        tc_gz1_chunked(trow + cB + c0, trow + cA + c0, smem + L::R_hi, smem + L::R_lo, r, c0, LO);
        tc_gz1_chunked(trow + cB + c0 + 32, trow + cA + c0 + 32, smem + L::R_hi, smem + L::R_lo, r, c0 + 32, LO);
        tc::fence_async_smem();
        tc::fence_before_sync();
        __syncthreads();
        if (issuer_warp && tc::elect_one()) {
            tc::fence_after_sync();
            tc_issue_op<NPASS>(tmem, cB, opR, opW1mn, D, 16, false, LO);  // gu = gz1 . W1  (16 columns, over ga1)
            // [dW1 | db1] of this tile = gz1^T . u: tile R viewed MN-major (K = the 128 edges), one MMA group per half of n
            const TcOperand opUmn{sb + L::u_hi, sb + L::u_lo, CS, true};
#pragma unroll
            for (int hf = 0; hf < 2; ++hf) {
                const TcOperand opRh{sb + L::R_hi + hf * 16 * CS, sb + L::R_lo + hf * 16 * CS, CS, true};
                tc_issue_op<NPASS>(tmem, cB + 32 + 16 * hf, opRh, opUmn, 128, 16, false, LO);
            }
            tc::mma_commit(&bars[3]);
        }
        __syncwarp();
        DBG_STAMP(7);  // gz1 pass + barrier
        ok = tc::mbar_wait(&bars[3], phase) && ok;
        tc::fence_after_sync();
        DBG_STAMP(8);  // gu observed
        if (cg == 0) {  // this branch's per-edge outputs
            float gu[32];
            tc::tmem_ld32(trow + cB, gu);
            if (r < nrows) {
                const size_t e = (size_t)(e0 + r);
                float gd = 0.f;
#pragma unroll
                for (int k = 0; k < 15; ++k) {
                    if (k < H) Grb[e * W + k] = gu[k];
                    else if (k < 2 * H) Gsb[e * W + (k - H)] = gu[k];
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
                Grb[e * W + H + 0] = g0;
                Grb[e * W + H + 1] = g1;
                Grb[e * W + H + 2] = g2;
                Gsb[e * W + H + 0] = -g0;
                Gsb[e * W + H + 1] = -g1;
                Gsb[e * W + H + 2] = -g2;
            }
        }
        {   // thread (row r, column group cg): rows n = 128 (cg / 2) + r of [dW1 | db1], columns 8 (cg % 2) .. + 7
            float w[8];
            tc::tmem_ld8(trow + cB + 32 + 16 * (cg >> 1) + 8 * (cg & 1), w);
#pragma unroll
            for (int k = 0; k < 8; ++k) accW1[k] += w[k];
        }
        tc::fence_before_sync();
        __syncthreads();  // sI / sR / tile R / the U tile / the gu and dW1 columns are free
        DBG_STAMP(9);  // tile end
        if (tid < 128 && has_next) {
            store_u(gnext);
            store_row(gnext);
        }
        tc::fence_async_smem();
        __syncthreads();
        if (has_next && issuer_warp && tc::elect_one()) {
            tc::fence_after_sync();
            tc_issue_op<NPASS>(tmem, cA, opU, opW1, 16, D, false);  // z1 of the next tile
            tc::mma_commit(&bars[0]);
        }
    }
    if (issuer_warp && tc::elect_one()) tc::bulk_wait_all();
    if (!ok) asm volatile("trap;");
    if (nit > 0) {
        const LayerOff lo = layer_off(a.net);
        float* part = a.part + (size_t)blockIdx.x * lo.total;
        const int o_wv = branch ? lo.wx3 : lo.wa, o_bv = branch ? lo.bx3 : lo.ba;
        {
            const int K0 = 2 * H + 1, o_w1 = branch ? lo.wx1 : lo.wm1, o_b1 = branch ? lo.bx1 : lo.bm1;
            const int n = 128 * (cg >> 1) + r;
#pragma unroll
            for (int k = 0; k < 8; ++k) {
                const int col = 8 * (cg & 1) + k;
                if (col < K0) part[o_w1 + n * K0 + col] = 0.5f * accW1[k];  // the gz tiles carry 2 dL/dz
                else if (col == 15) part[o_b1 + n] = 0.5f * accW1[k];
            }
        }
        float* colred = rowA;  // [16 warps][2][32] = 1024 floats: rowA and rowB are contiguous
        colred[(warp * 2 + 0) * 32 + lane] = accv0;
        colred[(warp * 2 + 1) * 32 + lane] = accv1;
        float* rowsum = reinterpret_cast<float*>(smem + L::sR);
        __syncthreads();
        if (cg == 0) rowsum[r] = accs;
        __syncthreads();
        if (tid < D) {  // column c = 64 cgc + 32 hh + cc: sum over the four row quadrants of column group cgc
            const int cgc = tid >> 6, hh = (tid >> 5) & 1, cc = tid & 31;
            float s = 0.f;
#pragma unroll
            for (int qq = 0; qq < 4; ++qq) s += colred[((cgc * 4 + qq) * 2 + hh) * 32 + cc];
            part[o_wv + tid] = s;
        }
        if (tid == 0) {
            float s = 0.f;
            for (int k = 0; k < 128; ++k) s += rowsum[k];
            part[o_bv] = s;
        }
    }
    tc::fence_before_sync();
    __syncthreads();
    if (warp == 0) tc::tmem_dealloc(tmem, 512);
}

// ---------------------------------------------------------------------------------------------------------------------
// Weight-gradient GEMM over the tile images (split-K over the edge tiles, accumulators persistent in TMEM).
//   CTA = (branch, half mh of the output rows n, K split):   cols [0,256)   += gz2[:, n-half]^T . a1
//                                                            cols [256,272) += gz2[:, n-half]^T . u    (col 15: db2)
// All three split products (hi.hi + lo.hi + hi.lo): with only the hi part of gz the error of a weight gradient whose
// per-edge terms have random signs is ~2^-9 of the LARGEST entry (measured 3e-3 on the cfg4 fixture), above the 1e-3 bar.
// 128 threads; thread t owns TMEM lane t = output row n = 128 mh + t in the epilogue.
// ---------------------------------------------------------------------------------------------------------------------
struct Wg256Args {
    egnn_net_t net;
    int tiles, ns0;  // K splits
    const uint8_t* img;
    float* part;  // rows [row0, row0 + grid)
    int row0;
};

// Software pipeline: the three split products of a tile are issued as separate groups in the order
//   P_A = gz_hi . a1_lo,   P_B = gz_hi . a1_hi,   P_C = gz_lo . a1_hi      (each with its u-column product)
// so that every operand buffer is free as early as possible: a producer thread refills a1_lo (+ u) as soon as P_A of the
// current tile has completed, gz_hi after P_B, a1_hi and gz_lo after P_C, while the issuer thread only waits for "landed".
// Loads of the next tile therefore overlap the MMAs of this one without a second set of buffers (208 KB in use).
struct WgSmem {
    static constexpr int HALF_B = TILE_B / 2;  // 16 chunk columns x 128 rows x 16 B: one half (128 n) of a gz tile
    static constexpr int gh = 0, gl = HALF_B, ah = 2 * HALF_B, al = ah + TILE_B, u0 = al + TILE_B, u1 = u0 + 2 * U_B;
    static constexpr int bars = u1 + 2 * U_B;  // F0, F1, DA, DB, DC
    static constexpr int slot = bars + 64;
    static constexpr int TOTAL = slot + 16;
    static_assert(TOTAL <= 227 * 1024, "tc_wgrad256 shared memory");
};

template <int NPASS>
__global__ void __launch_bounds__(128, 1) tc_wgrad256(Wg256Args a) {
    using S = WgSmem;
    constexpr bool LO = NPASS == 3;
    constexpr int HALF_B = S::HALF_B;
    constexpr uint32_t CS = 128 * 16;
    extern __shared__ __align__(1024) uint8_t smem[];
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + S::bars);
    uint32_t* slot = reinterpret_cast<uint32_t*>(smem + S::slot);
    const int tid = threadIdx.x, warp = tid >> 5;
    constexpr int role = 0;  // (a second role, gz1^T . u from a gz1 image, was folded into tc_edge_bwd256)
    const int id = (int)blockIdx.x, ns = a.ns0;
    const int sp = id % ns, mh = (id / ns) & 1, branch = id / (2 * ns);
    if (warp == 0) tc::tmem_alloc(slot, 512);
    if (tid == 0) {
        for (int b = 0; b < 5; ++b) tc::mbar_init(&bars[b], 1);
        tc::fence_mbar_init();
    }
    tc::fence_before_sync();
    __syncthreads();
    tc::fence_after_sync();
    const uint32_t tmem = *slot;
    const uint32_t sb = tc::smem_u32(smem);
    const int nit = sp < a.tiles ? (a.tiles - sp + ns - 1) / ns : 0;
    const bool has_f1 = role == 0 || LO;  // a1_hi and / or gz_lo exist
    const int g_off = IMG_G2;
    const uint32_t F0 = tc::smem_u32(&bars[0]), F1 = tc::smem_u32(&bars[1]);
    auto img_of = [&](int it) { return a.img + ((size_t)branch * a.tiles + sp + it * ns) * IMG_TILE; };
    const uint32_t f0_bytes = HALF_B + 2 * U_B + (role == 0 && LO ? TILE_B : 0);
    const uint32_t f1_bytes = (role == 0 ? TILE_B : 0) + (LO ? HALF_B : 0);
    // producer-side pieces (one thread)
    auto load_al_u = [&](int it) {  // first load of a tile's F0 group: arms the barrier for the whole group
        const uint8_t* img = img_of(it);
        tc::mbar_expect_tx(F0, f0_bytes);
        if (role == 0 && LO) tc::bulk_g2s(sb + S::al, img + IMG_A1 + TILE_B, TILE_B, F0);
        tc::bulk_g2s(sb + ((it & 1) ? S::u1 : S::u0), img + IMG_U, 2 * U_B, F0);
    };
    auto load_gh = [&](int it) { tc::bulk_g2s(sb + S::gh, img_of(it) + g_off + mh * HALF_B, HALF_B, F0); };
    auto load_ah_gl = [&](int it) {
        const uint8_t* img = img_of(it);
        tc::mbar_expect_tx(F1, f1_bytes);
        if (role == 0) tc::bulk_g2s(sb + S::ah, img + IMG_A1, TILE_B, F1);
        if (LO) tc::bulk_g2s(sb + S::gl, img + g_off + TILE_B + mh * HALF_B, HALF_B, F1);
    };
    bool ok = true;
    if (tid == 32 && nit > 0) {  // producer
        load_al_u(0);
        load_gh(0);
        if (has_f1) load_ah_gl(0);
        for (int it = 0; it + 1 < nit; ++it) {
            const uint32_t ph = it & 1;
            // F0 may only be re-armed once the issuer has consumed its current phase, i.e. after a commit that follows
            // the issuer's F0 wait: P_A's (bf16x3) or P_B's (bf16)
            ok = tc::mbar_wait(&bars[LO ? 2 : 3], ph) && ok;  // P_A(it) done: a1_lo (and the other u buffer) are free
            load_al_u(it + 1);
            if (LO) ok = tc::mbar_wait(&bars[3], ph) && ok;   // P_B(it) done: gz_hi is free
            load_gh(it + 1);
            if (has_f1) {
                ok = tc::mbar_wait(&bars[4], ph) && ok;      // P_C(it) done: a1_hi and gz_lo are free
                load_ah_gl(it + 1);
            }
        }
        if (!ok) asm volatile("trap;");
    }
    if (tid == 0 && nit > 0) {  // MMA issuer
        bool accW = false, accU = false;
        auto product = [&](uint32_t g_addr, uint32_t a_addr, uint32_t u_addr) {
            const TcOperand G{sb + g_addr, sb + g_addr, CS, true}, U{sb + u_addr, sb + u_addr, CS, true};
            if (role == 0) {
                const TcOperand A1{sb + a_addr, sb + a_addr, CS, true};
                tc_issue_op<1>(tmem, 0, G, A1, 128, D, accW);
                accW = true;
                tc_issue_op<1>(tmem, D, G, U, 128, 16, accU);
            } else {
                tc_issue_op<1>(tmem, 0, G, U, 128, 16, accU);
            }
            accU = true;
        };
        for (int it = 0; it < nit; ++it) {
            const uint32_t ph = it & 1;
            const uint32_t ub = (it & 1) ? S::u1 : S::u0;
            ok = tc::mbar_wait(&bars[0], ph) && ok;  // gz_hi, a1_lo, u
            if (LO) {
                product(S::gh, S::al, ub + U_B);      // P_A: gz_hi . [a1_lo | u_lo]
                tc::mma_commit(&bars[2]);
            }
            if (has_f1) ok = tc::mbar_wait(&bars[1], ph) && ok;  // a1_hi, gz_lo
            product(S::gh, S::ah, ub);                // P_B: gz_hi . [a1_hi | u_hi]
            tc::mma_commit(&bars[3]);
            if (LO) product(S::gl, S::ah, ub);        // P_C: gz_lo . [a1_hi | u_hi]
            tc::mma_commit(&bars[4]);
        }
        ok = tc::mbar_wait(&bars[4], (nit - 1) & 1) && ok;
        if (!ok) asm volatile("trap;");
    }
    __syncthreads();
    tc::fence_after_sync();
    if (nit > 0) {
        const LayerOff lo = layer_off(a.net);
        const int K0 = 2 * a.net.H + 1;
        float* part = a.part + (size_t)(a.row0 + blockIdx.x) * lo.total;
        const int o_w1 = branch ? lo.wx1 : lo.wm1, o_b1 = branch ? lo.bx1 : lo.bm1;
        const int o_w2 = branch ? lo.wx2 : lo.wm2, o_b2 = branch ? lo.bx2 : lo.bm2;
        const int n = 128 * mh + tid;
        const uint32_t trow = tmem + ((uint32_t)(warp * 32) << 16);
        if (role == 0) {
#pragma unroll 1
            for (int cb = 0; cb < D; cb += 32) {
                float v[32];
                tc::tmem_ld32(trow + cb, v);
                float* dst = part + o_w2 + (size_t)n * D + cb;
#pragma unroll
                for (int k = 0; k < 32; ++k) dst[k] = 0.5f * v[k];
            }
            float v[32];
            tc::tmem_ld32(trow + D, v);
            part[o_b2 + n] = 0.5f * v[15];
        } else {
            float v[32];
            tc::tmem_ld32(trow, v);
            part[o_b1 + n] = 0.5f * v[15];
#pragma unroll
            for (int k = 0; k < 15; ++k)
                if (k < K0) part[o_w1 + n * K0 + k] = 0.5f * v[k];
        }
    }
    tc::fence_before_sync();
    __syncthreads();
    if (warp == 0) tc::tmem_dealloc(tmem, 512);
}

}  // namespace

// rows of the partial-gradient buffer the two kernels write (reduce_partials_kernel reads exactly these, nothing is
// zero-filled):  tc_edge_bwd256 CTA b -> row b, [W1 | b1 | row-dot vector and its bias] of its branch: rows [0, nM) the
// message branch, [nM, nM + nX) the coordinate branch;  tc_wgrad256 CTA id -> row 148 + id, the 128 rows `mh` of
// [W2 | b2] of its branch: rows 148 + (2 branch + mh) ns + [0, ns)
int egnn_tc_bwd256_rows(int64_t E) { return E > 0 ? 296 : 0; }
void egnn_tc_bwd256_split(int64_t E, int* nM, int* nX, int* ns) {
    static int fixed = -1;  // CTAs of the message branch (it makes two passes over z2 for the gate); 0 = choose per call
    if (fixed < 0) {
#ifdef EGNN_DBG  // tuning knob of debug builds only
        const char* env = getenv("EGNN_TC256_BWD_MCTAS");
        fixed = env ? atoi(env) : 0;
#else
        fixed = 0;
#endif
        if (fixed < 16 || fixed > 132) fixed = 0;
    }
    const int64_t tiles = (E + 127) / 128;
    int share = fixed;
    if (!share) {
        // a message-branch tile costs ~4 % more than a coordinate-branch tile (measured: 74/74 and 76/72 CTAs bracket the
        // optimum at cfg4); pick the split with the smallest max over branches of (tiles per CTA) x (tile cost)
        double best = 1e30;
        for (int m = 60; m <= 88; ++m) {
            const double c = std::max(1.04 * (double)((tiles + m - 1) / m), (double)((tiles + 147 - m) / (148 - m)));
            if (c < best - 1e-9 || (c < best + 1e-9 && std::abs(m - 74) < std::abs(share - 74))) best = c, share = m;
        }
    }
    *nM = (int)(tiles < share ? tiles : share);
    *nX = (int)(tiles < 148 - share ? tiles : 148 - share);
    *ns = (int)(tiles < 37 ? tiles : 37);  // 2 branches x 2 halves x 37 K splits = 148 CTAs
}
int64_t egnn_tc_bwd256_img_bytes(int64_t E) { return 2 * ((E + 127) / 128) * (int64_t)IMG_TILE; }

int egnn_tc_edge_backward256(const egnn_net_t& net, const egnn_graph_t& g, const float* packed_layer, const float* h,
                             const float* x, const float* gM, const float* gdx, const float* coefS, float* Gr, float* Gs,
                             float* part, uint8_t* img, int precision, cudaStream_t st) {
    EGNN_REQUIRE(net.Mh == 256 && tc_fwd_supported(net), "egnn_tc_edge_backward256: needs Mh = Ms = Xh = 256 and 2H+1 <= 16");
    EGNN_REQUIRE((reinterpret_cast<uintptr_t>(img) & 15) == 0, "egnn_tc_edge_backward256: image workspace must be 16-byte aligned");
    const PackOff po = pack_off(net);
    const uint8_t* blob = reinterpret_cast<const uint8_t*>(packed_layer + po.tc);
    const int tiles = (g.E + 127) / 128;
    int nM, nX, ns0;
    egnn_tc_bwd256_split(g.E, &nM, &nX, &ns0);
    Bwd256Args a{net, g.N, g.E, tiles, g.recv, g.send, g.node_graph, h, x, blob, blob + tc_w2t_offset(net.Mh), gM, gdx, coefS,
                 Gr, Gs, part, img, nM};
    const bool x3 = precision == EGNN_PREC_BF16X3;
    constexpr int WG_SMEM = WgSmem::TOTAL;
    EGNN_SMEM_ATTR(tc_edge_bwd256<3>, LB::TOTAL);
    EGNN_SMEM_ATTR(tc_edge_bwd256<1>, LB::TOTAL);
    EGNN_SMEM_ATTR(tc_wgrad256<3>, WG_SMEM);
    EGNN_SMEM_ATTR(tc_wgrad256<1>, WG_SMEM);
    egnn_prof_begin(EGNN_K_TC_EDGE_BWD, st);
    if (x3) tc_edge_bwd256<3><<<nM + nX, NT, LB::TOTAL, st>>>(a);
    else tc_edge_bwd256<1><<<nM + nX, NT, LB::TOTAL, st>>>(a);
    egnn_prof_end(st);
    EGNN_CHECK_LAUNCH("tc_edge_bwd256");
    Wg256Args w{net, tiles, ns0, img, part, 148};
    egnn_prof_begin(EGNN_K_TC_WGRAD, st);
    if (x3) tc_wgrad256<3><<<4 * ns0, 128, WG_SMEM, st>>>(w);
    else tc_wgrad256<1><<<4 * ns0, 128, WG_SMEM, st>>>(w);
    egnn_prof_end(st);
    EGNN_CHECK_LAUNCH("tc_wgrad256");
    return EGNN_OK;
}
