/*
 * egnn_b200.h -- C ABI of libegnn_b200.so: the B200 (sm_100a) implementation of the EGNN-diffusion hot path of
 * Ren-Okubo/diffusion_model.  Citations are file:line in the reference checkout.
 *
 * Conventions
 *   - every pointer is a DEVICE pointer owned by the caller unless the name ends in _host; nothing is allocated
 *     inside the library; all work is enqueued on the cudaStream_t passed as `stream` (void* here so that the
 *     header is plain C); calls are re-entrant per stream;
 *   - every entry point returns 0 on success or a negative egnn_status; egnn_b200_last_error() returns a
 *     thread-local description of the last failure; no exception ever crosses this boundary;
 *   - all floating point tensors are fp32, row-major, contiguous; all index tensors are int32 except the
 *     reference-facing edge_index (int64 [2,E], the layout PyG / the reference scripts use).
 *
 * Parameter buffer ("flat params"): one fp32 array holding, for l = 0..L-1 and in the order of the reference
 * module's named_parameters() (EquivariantGraphNeuralNetwork.py:13-34):
 *     mlp_m.0.weight [Mh,K0]  mlp_m.0.bias [Mh]  mlp_m.2.weight [Ms,Mh]  mlp_m.2.bias [Ms]
 *     mlp_x.0.weight [Xh,K0]  mlp_x.0.bias [Xh]  mlp_x.2.weight [Xh,Xh]  mlp_x.2.bias [Xh]
 *     mlp_x.4.weight [1,Xh]   mlp_x.4.bias [1]
 *     mlp_h.0.weight [Hh,H+Ms] mlp_h.0.bias [Hh] mlp_h.2.weight [H,Hh]   mlp_h.2.bias [H]
 *     attention.0.weight [1,Ms] attention.0.bias [1]
 * with K0 = 2H+1.  Gradients use the same layout.
 */
#ifndef EGNN_B200_H
#define EGNN_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define EGNN_B200_ABI_VERSION 1

typedef enum {
    EGNN_OK = 0,
    EGNN_ERR_ARG = -1,       /* bad argument / unsupported shape */
    EGNN_ERR_CUDA = -2,      /* a CUDA runtime call or launch failed */
    EGNN_ERR_WORKSPACE = -3, /* workspace too small */
    EGNN_ERR_UNSUPPORTED = -4
} egnn_status;

/* arithmetic of the MLP GEMMs.  The tcgen05 modes cover two families of network shapes and refuse anything else
 * (EGNN_ERR_ARG with a message, never a silent fallback):
 *   fused edge kernels   Mh = Ms = Xh = d in {64, 128, 256} and 2H+1 <= 16 (BASELINE.json cfg1-4);
 *   tile-GEMM path       Mh, Xh, Hh <= 1024, Ms <= 256, 2H+1 <= 128, ragged widths allowed: the reference's own
 *                        parameters.yaml network (main.py:102-121: K0 = 73, hidden 1024, m_size 256). */
typedef enum {
    EGNN_PREC_FP32 = 0,   /* CUDA-core FFMA, true fp32 like the reference */
    EGNN_PREC_BF16X3 = 1, /* tcgen05 kind::f16, operands split hi+lo bf16, 3 products, fp32 accumulate (~2^-16) */
    EGNN_PREC_BF16 = 2    /* tcgen05 kind::f16, single bf16 product, fp32 accumulate */
} egnn_precision;

/* scope of the coordinate-message normaliser S = ||x_i - x_j||_F (EquivariantGraphNeuralNetwork.py:64) */
typedef enum {
    EGNN_NORM_CALL = 0, /* one scalar over every edge of the call: what a batched reference call computes */
    EGNN_NORM_GRAPH = 1 /* one scalar per graph: what the reference computes sampling one structure per call */
} egnn_norm_scope;

/* network sizes: EquivariantGNN(L, 2H+1, Mh, Ms, 2H+1, Xh, 1, H+Ms, Hh, H)  (EquivariantGraphNeuralNetwork.py:74-83) */
typedef struct {
    int32_t L, H, Mh, Ms, Xh, Hh;
} egnn_net_t;

/* a receiver-sorted directed edge list plus the derived index structures (all built by egnn_graph_prepare) */
typedef struct {
    int32_t N, E, G;
    const int32_t* recv;       /* [E]   edge_index[0]  (aggregation target, flow='target_to_source', :10-11) */
    const int32_t* send;       /* [E]   edge_index[1] */
    const int32_t* row_ptr;    /* [N+1] edges of receiver i are [row_ptr[i], row_ptr[i+1]) */
    const int32_t* col_ptr;    /* [N+1] sender CSR: edges whose sender is i ... */
    const int32_t* col_edge;   /* [E]   ... listed by ascending edge id (needed by backward only) */
    const int32_t* node_ptr;   /* [G+1] nodes of graph g are [node_ptr[g], node_ptr[g+1]) */
    const int32_t* node_graph; /* [N]   graph id per node (PyG `batch`) */
} egnn_graph_t;

const char* egnn_b200_last_error(void);
int egnn_b200_version(void);

/* ---- instrumentation (bench.py): number of kernels the library has launched in this process; optional CUDA-event
 *      brackets around one kernel family (kind: 0 edge fwd, 1 node fwd, 2 x-update, 3 edge bwd, 4 node bwd,
 *      5 gather bwd, 7 tcgen05 edge fwd, 8 tcgen05 edge bwd), summed by egnn_profile_read into HOST variables ---- */
int64_t egnn_launch_count(void);
int egnn_profile_enable(int32_t on);
int egnn_profile_read(int32_t kind, double* total_ms_host, int64_t* launches_host);

/* one-tile tcgen05 GEMM used by the tests to pin the shared-memory descriptor conventions on hardware:
 * D[128,N] = A[128,K] . B[N,K]^T, fp32 in/out, bf16 operands (npass 1) or hi/lo split with 3 products (npass 3);
 * flags: 1 swap LBO/SBO (K-major), 2 stage A MN-major, 4 stage B MN-major, 8 swap LBO/SBO of MN-major descriptors */
int egnn_tc_selftest(const float* A, const float* B, float* D, int32_t K, int32_t N, int32_t npass,
                     int32_t flags, int32_t* status_dev, void* stream);

/* ---- sizes ----------------------------------------------------------------------------------------------- */
int64_t egnn_param_count(const egnn_net_t* net);          /* floats in the flat parameter buffer */
int64_t egnn_packed_count(const egnn_net_t* net);         /* floats in the packed (transposed) weight buffer */
int64_t egnn_forward_workspace_floats(const egnn_net_t* net, int32_t N, int32_t E, int32_t G);
int64_t egnn_backward_workspace_floats(const egnn_net_t* net, int32_t N, int32_t E, int32_t G);
int32_t egnn_edge_tile(const egnn_net_t* net);            /* edges per forward tile (partials granularity) */

/* ---- graph construction: replaces the python double loops
 *      parts/train_per_iretation.py:308-313, make_dataset.py:131-136 (fully connected, row-major)
 *      something.ipynb:6340-6365 (strict '<' cutoff on fp32 torch.norm, lexicographic order) ------------------ */
/* fully connected: edge_index_out int64 [2,E], E = sum n_g (n_g - 1); node_ptr int32 [G+1] */
int egnn_build_edges_fc(const int32_t* node_ptr, int32_t G, int32_t N, int64_t E, int64_t* edge_index_out,
                        void* stream);
/* cutoff graph, two passes.  box: fp32 [G,3] orthorhombic box lengths for periodic minimum image, or NULL (open).
 * pass 1: deg_out int32 [N] = number of neighbours of every node.  The caller scans deg into row_ptr [N+1]
 * (egnn_exclusive_scan_i32) and reads E = row_ptr[N].  pass 2 fills edge_index_out int64 [2,E]. */
int egnn_cutoff_count(const float* pos, const int32_t* node_ptr, const int32_t* node_graph, const float* box,
                      float r_c, int32_t N, int32_t* deg_out, void* stream);
int egnn_cutoff_fill(const float* pos, const int32_t* node_ptr, const int32_t* node_graph, const float* box,
                     float r_c, int32_t N, const int32_t* row_ptr, int64_t E, int64_t* edge_index_out, void* stream);
int egnn_exclusive_scan_i32(const int32_t* in, int32_t n, int32_t* out /* [n+1] */, void* stream);
/* fallback of the same loop (something.ipynb:6346-6358): a graph with no pair inside the cutoff gets the single closest
 * ordered pair with r_c < dist < 100 (first strict minimum of the row-major scan).  Call egnn_cutoff_fallback between
 * the count pass and the scan (it sets deg[i*] = 1 and records fb_pairs int32 [G][2], -1 when unused), and
 * egnn_cutoff_fallback_fill after egnn_cutoff_fill. */
int egnn_cutoff_fallback(const float* pos, const int32_t* node_ptr, const float* box, float r_c, int32_t G,
                         int32_t* deg, int32_t* fb_pairs, void* stream);
int egnn_cutoff_fallback_fill(const int32_t* fb_pairs, const int32_t* row_ptr, int32_t G, int64_t E,
                              int64_t* edge_index_out, void* stream);

/* derived structures from a reference-layout edge_index (int64 [2,E]).  Requires edge_index[0] non-decreasing
 * (every builder of the reference produces that); flags_out[0] is set to 1 on the device if it is not, flags_out[1]
 * if an index is out of range.  Writes recv/send [E], row_ptr [N+1], col_ptr [N+1], col_edge [E];
 * scratch: int32 [N+1+E]. */
int egnn_graph_prepare(const int64_t* edge_index, int32_t N, int64_t E, int32_t* recv, int32_t* send,
                       int32_t* row_ptr, int32_t* col_ptr, int32_t* col_edge, int32_t* scratch, int32_t* flags_out,
                       void* stream);
/* Allocation-free variant of egnn_graph_prepare.  It has NO counterpart in the reference: it was added
 * for callers that cannot call cudaMalloc (code under CUDA-graph capture, tight serving loops, harnesses
 * that forbid it).  Same inputs, outputs and flags as egnn_graph_prepare, but the int32 [N+1+E] scratch lives in a static
 * __device__ array inside the library (EGNN_GRAPH_PREPARE_NOALLOC_MAX_INTS int32 = 4 MiB, held by every GPU that loads
 * the library until exit) and the call performs no cudaMalloc / cudaFree / allocating library routine.
 *   - N + 1 + E > EGNN_GRAPH_PREPARE_NOALLOC_MAX_INTS is rejected with EGNN_ERR_ARG and a message naming the limit;
 *   - every call on a device shares the ONE array, which keeps the previous call's contents; the counters that must
 *     start at zero are cleared on `stream` inside the call.  Calls that overlap on the GPU (different streams, or two
 *     host threads on one device) RACE on it: serialise them, or use egnn_graph_prepare with private scratch buffers. */
#define EGNN_GRAPH_PREPARE_NOALLOC_MAX_INTS (1 << 20)
int egnn_graph_prepare_noalloc(const int64_t* edge_index, int32_t N, int64_t E, int32_t* recv, int32_t* send,
                               int32_t* row_ptr, int32_t* col_ptr, int32_t* col_edge, int32_t* flags_out, void* stream);
/* node_graph [N] (non-decreasing) -> node_ptr [G+1] */
int egnn_node_ptr_from_batch(const int64_t* batch, int32_t N, int32_t G, int32_t* node_ptr, int32_t* node_graph,
                             void* stream);

/* ---- EGNN denoiser: replaces EquivariantGNN.forward / EGCL.forward / EGCL.message
 *      (EquivariantGraphNeuralNetwork.py:55-71, 85-88) and their autograd backward ----------------------------- */
int egnn_pack_weights(const egnn_net_t* net, const float* params, float* packed, void* stream);
/* h_in [N,H], x_in [N,3] -> h_out [N,H], x_out [N,3].  `ws` (egnn_forward_workspace_floats) keeps every layer's
 * input and per-node aggregates, which egnn_net_backward reads. */
int egnn_net_forward(const egnn_net_t* net, const egnn_graph_t* g, const float* params, const float* packed,
                     const float* h_in, const float* x_in, float* h_out, float* x_out, float* ws, int64_t ws_floats,
                     int32_t norm_scope, int32_t precision, void* stream);
/* gh_out [N,H], gx_out [N,3] (grad of the outputs) -> gh_in, gx_in (grad of the inputs) and grad_params (same
 * layout as params, overwritten).  fwd_ws is the workspace the forward call filled. */
int egnn_net_backward(const egnn_net_t* net, const egnn_graph_t* g, const float* params, const float* packed,
                      const float* h_in, const float* x_in, const float* fwd_ws, const float* gh_out,
                      const float* gx_out, float* gh_in, float* gx_in, float* grad_params, float* ws,
                      int64_t ws_floats, int32_t norm_scope, int32_t precision, void* stream);

/* egnn_net_backward that also records one CUDA event per layer: layer_events_host is a HOST array of L cudaEvent_t
 * (entries may be NULL); event l is recorded on `stream` right after layer l's slice of grad_params
 * ([l * P_layer, (l+1) * P_layer), P_layer = egnn_param_count / L) is final.  Backward walks l = L-1 .. 0, so a
 * communication stream that waits on event l can all-reduce that slice over NCCL while the layers below are still in
 * backward (data-parallel training, SURVEY.md 8e: the gradient all-reduce is the path's only collective). */
int egnn_net_backward_events(const egnn_net_t* net, const egnn_graph_t* g, const float* params, const float* packed,
                             const float* h_in, const float* x_in, const float* fwd_ws, const float* gh_out,
                             const float* gx_out, float* gh_in, float* gx_in, float* grad_params, float* ws,
                             int64_t ws_floats, int32_t norm_scope, int32_t precision, void* const* layer_events_host,
                             void* stream);

/* ---- diffusion process: replaces remove_mean (diffusion_x_h.py:5-14), diffuse_zero_to_t (:51-59) inside
 *      diffuse_as_batch (parts/train_per_iretation.py:36-92), the loss tail (:161-169),
 *      calculate_mu + reverse_diffuse_one_step (diffusion_x_h.py:61-90) inside generate (:335-428) ------------- */
/* x [N,C] -= per-graph column mean, in place (node_ptr [G+1]) */
int egnn_remove_mean(float* x, int32_t C, const int32_t* node_ptr, int32_t G, void* stream);
/* per graph g with time t[g]:  eps_x = noise_x - mean_g(noise_x);  z_x = a_t*pos + s_t*eps_x;
 *                              z_h = a_t*x0 + s_t*noise_h;  h_in = [z_h | cond | t/T]        (H = A + C + 1)
 * alpha/sigma: the [T+1] schedule tables.  cond may be NULL (C = 0). */
int egnn_diffuse_batch(const float* pos, const float* x0, const float* noise_x, const float* noise_h,
                       const float* cond, const int32_t* t, const float* alpha, const float* sigma, int32_t T,
                       const int32_t* node_ptr, int32_t G, int32_t A, int32_t C, float* z_x, float* eps_x,
                       float* h_in, void* stream);
/* loss = sum_g loss_g, loss_g = (||P(x_out - z_x) - eps_x||^2 + ||h_out[:, :A] - eps_h||^2) / G;
 * writes loss_g [G] and the gradients gx_out [N,3], gh_out [N,H] of that loss. */
int egnn_loss_fwd_bwd(const float* x_out, const float* h_out, const float* z_x, const float* eps_x,
                      const float* eps_h, const int32_t* node_ptr, int32_t G, int32_t H, int32_t A, float* loss_g,
                      float* gx_out, float* gh_out, void* stream);
/* one ancestral step for every graph of the batch, in place on z_x [N,3] and h_in [N,H] (= [s*z_h | cond | t/T]):
 *   eps_x = P(x_out - z_x);  z_x <- z_x/a_ts - s2_ts*eps_x/a_ts/sig_t + std*P(noise_x)
 *   z_h   <- h_in[:, :A]/a_ts - s2_ts*h_out[:, :A]/a_ts/sig_t + std*noise_h;   h_in <- [s*z_h | cond | (t-1)/T]
 * coef: DEVICE table [T+1][8], row t = {a_ts, s2_ts, sig_t, std, time column of t-1, alpha_0, sigma_0, 0}, filled on
 * the host by the reference's own fp32 formulae (diffusion_x_h.py:61-84).  The current t is read from *t_dev so that
 * one captured CUDA graph of a step can be replayed for every t (egnn_counter_add decrements it); the noise of step t
 * is read at noise + ((t_base - t) mod noise_period) * noise_stride (noise_period 0: no wrap; the caller refills the
 * ring between replays).
 * final_step != 0 selects the t = 0 form (parts/train_per_iretation.py:416-426): z <- z/a0 - s0*eps/a0 + s0*noise/a0,
 * positions in place, atom-type logits to z_h_out [N,A]; t_dev is not read.
 * finite_flag [G] is cleared for a graph when anything written for it is not finite (:376-389). */
int egnn_sampler_step(float* z_x, float* h_in, const float* x_out, const float* h_out, const float* noise_x,
                      const float* noise_h, const float* coef, const int32_t* t_dev, int32_t t_base,
                      int64_t noise_stride_x, int64_t noise_stride_h, int32_t noise_period, float onehot_scale,
                      const int32_t* node_ptr, int32_t G, int32_t H, int32_t A, int32_t final_step, float* z_h_out, int32_t* finite_flag,
                      void* stream);
int egnn_counter_add(int32_t* counter, int32_t delta, void* stream);

/* fused Adam on the flat buffers (torch.optim.Adam semantics, parts/def_for_main.py:133-134) */
int egnn_adam_step(float* params, const float* grads, float* exp_avg, float* exp_avg_sq, int64_t n, float lr,
                   float beta1, float beta2, float eps, float weight_decay, int32_t step, float grad_scale,
                   void* stream);

/* egnn_adam_step with the 1-based step number read from DEVICE memory (*step_dev): the form a captured CUDA graph of a
 * training step replays (the host-side counter of egnn_adam_step would freeze the bias corrections at capture time). */
int egnn_adam_step_dev(float* params, const float* grads, float* exp_avg, float* exp_avg_sq, int64_t n, float lr,
                       float beta1, float beta2, float eps, float weight_decay, const int32_t* step_dev, float grad_scale,
                       void* stream);

/* schedulefree.RAdamScheduleFree (parts/def_for_main.py:137-138; `optimizer: RAdamScheduleFree` in parameters.yaml) on flat
 * buffers.  The package is not part of the reference checkout (un-vendored, unpinned dependency): the update is restated
 * from its published algorithm (facebookresearch/schedule_free, radam_schedulefree.py); oracle/radam_sf_ref.py is the CPU
 * restatement the tests compare against.  One launch does, per element and with host-computed scalars of step t,
 *   v <- beta2 v + (1-beta2) g^2;   g^ = adam_phase ? g / (sqrt(v / bias_correction2) + eps) : g;   g^ += weight_decay y
 *   y <- lerp(y, z, ckp1) + adaptive_y_lr g^;      z <- z - lr g^
 * (y: the parameters in train mode, z: the base sequence, lr: the rectified learning rate of the step). */
int egnn_radam_sf_step(float* y, const float* grads, float* z, float* exp_avg_sq, int64_t n, float lr, float beta2,
                       float eps, float weight_decay, float ckp1, float adaptive_y_lr, float bias_correction2,
                       int32_t adam_phase, float grad_scale, void* stream);
/* p <- lerp(p, end, weight) (torch.lerp): the train() / eval() switch of the schedule-free optimizer (y <-> x) */
int egnn_lerp(float* p, const float* end, int64_t n, float weight, void* stream);

/* ---- structure analysis ("sample -> evaluate"): one CTA per structure; structures are the node
 *      ranges node_ptr[s] .. node_ptr[s+1] of pos [N,3]; atom 0 of a structure is its first node ------------------- */
/* Local RDF around atom 0 (replaces evaluate_RDF.py:39-60, RDF + length_from_exO).
 *   counts [S,nbins] int32: #{i >= 1 : edge_lo[k] < |p_i - p_0| < edge_hi[k]}, compared in fp32 like the reference
 *                           (edge_lo/hi = fp32 images of r_k and r_k + dR, r = np.arange(dR, R + dR, dR));
 *   rdf    [S,nbins] f64  : gaussian_filter1d(counts / (4 pi rho r_k^2 dR)), rho = n_s / vol, correlation weights
 *                           `weights` [2*radius+1] (host-computed Gaussian, boundary mode 'reflect');
 *   dists  [N] f32 or NULL: |p_i - p_0| of every atom (0 for atom 0).
 * r_sq [nbins] f64 = r_k^2.  nbins <= 4096. */
int egnn_rdf_local(const float* pos, const int32_t* node_ptr, int32_t S, const float* edge_lo, const float* edge_hi,
                   const double* r_sq, int32_t nbins, double dR, double vol, const double* weights, int32_t radius,
                   int32_t* counts, double* rdf, float* dists, void* stream);
/* Two-coordinated site of atom 0 (replaces the selection loop of evaluate_Si-O-Si.py:28-39 and
 * CN2_evaluate.py:12-21 calculate_angle_for_CN2 / calculate_bond_length_for_CN2).
 *   radius > 0 : count [S] = atoms i >= 1 with |p_i - p_0| < radius; idx [S,2] = the first two of them in index order;
 *   radius <= 0: idx = the two atoms nearest to atom 0 (index order), count = min(2, n - 1);
 *   angle_deg [S] f64 = degrees(acos(cos)) at atom 0, bonds [S,2] f32 = the two bond lengths (NaN when fewer than two
 *   atoms were selected; -1 in idx). */
int egnn_cn2_sites(const float* pos, const int32_t* node_ptr, int32_t S, float radius, int32_t* count, int32_t* idx,
                   double* angle_deg, float* bonds, void* stream);
/* Kabsch alignment of P onto Q per structure (replaces evaluate_rmsd_for_pos_generate.py:11-51 kabsch_torch):
 * R [S,3,3], t [S,3] = centroid(Q) - centroid(P), rmsd [S].  Keeps the reference's handedness quirk (:42-45): for a
 * left-handed V U^T it returns diag(1,1,-1) V U^T. */
int egnn_kabsch(const float* P, const float* Q, const int32_t* node_ptr, int32_t S, float* R, float* t, float* rmsd,
                void* stream);

#ifdef __cplusplus
}
#endif
#endif /* EGNN_B200_H */
