/* fedml_b200 native library — C ABI.
 *
 * One shared object (fedml_b200/native/libfedml_b200.so, sm_100a only) holds every hand-written
 * kernel of the framework.  All entry points take raw device pointers and a CUDA stream, return 0
 * on success or a negative FB_ERR_* code (text via fb_last_error()).  Kernels never allocate device
 * memory: scratch buffers are supplied by the caller, sized by the *_scratch_bytes helpers (the
 * only allocating calls are the explicit fb_dev_malloc / IPC helpers of the host section).
 * Python binds these with ctypes (fedml_b200/ops/native.py); C/C++ callers include this header.
 *
 * The reference (Cossack9989/FedML, python/fedml) has no native code at all; each
 * declaration cites the Python call site whose work the kernel takes over.
 */
#ifndef FEDML_B200_H
#define FEDML_B200_H
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct CUstream_st* fb_stream_t;

const char* fb_last_error(void);
int fb_version(void);
int fb_device_sm_count(void);

/* ---------------------------------------------------------------------------------------------
 * Host runtime: shareable device allocations (CUDA IPC), peer access, pinned staging memory.
 * These back fedml_b200.parallel.symmetric when torch's symmetric-memory rendezvous is unavailable;
 * the reference's counterpart is ProcessGroupManager (cross_silo/hierarchical/process_group_manager.py:30-34).
 * ------------------------------------------------------------------------------------------- */
int fb_dev_malloc(void** out, int64_t bytes);      /* zero-filled cudaMalloc */
int fb_dev_free(void* p);
int fb_ipc_handle_bytes(void);
int fb_ipc_get_handle(void* dev_ptr, void* handle_out);
int fb_ipc_open_handle(const void* handle, void** out);
int fb_ipc_close_handle(void* p);
int fb_enable_peer_access(int peer_device);
int fb_host_alloc_pinned(void** out, int64_t bytes);
int fb_host_free_pinned(void* p);

/* ---------------------------------------------------------------------------------------------
 * K1+K2+K3+K4  Fused FL round kernel: per-client weighting, optional DP clip + Gaussian noise,
 * local accumulation, cross-GPU reduction over peer (NVLink/NVSwitch) memory, server-optimizer
 * step and redistribution of the new global model — one cooperative launch per rank, no NCCL.
 * Replaces: FedAvgAPI._aggregate (fedavg/fedavg_api.py:206-221), FedAVGAggregator.aggregate
 * (mpi_p2p_mp/fedavg/FedAVGAggregator.py:68-97), the MPI model send/recv (communication/mpi/*.py),
 * RobustAggregator.norm_diff_clipping/add_noise (core/robustness/robust_aggregation.py:47-63),
 * FedOpt server step (fedopt/fedopt_api.py:134-186), FedNova aggregate (fednova_trainer.py:136-164).
 * ------------------------------------------------------------------------------------------- */
enum { FB_OPT_NONE = 0, FB_OPT_SGD = 1, FB_OPT_ADAM = 2, FB_OPT_ADAGRAD = 3, FB_OPT_YOGI = 4 };

typedef struct {
  /* this rank's clients */
  const float* clients;        /* [C][ld] fp32 stacked client models (NULL iff C == 0) */
  int64_t ld;                  /* row stride in floats, multiple of 4 */
  int32_t C;                   /* clients resident on this rank */
  int32_t client_base;         /* global index of this rank's first client (keys the DP noise stream) */
  const float* weights;        /* [C] aggregation weight of each client for [0, P_train) (e.g. n_i) */
  const float* weights_buf;    /* [C] weights for [P_train, P) (BN buffers); NULL = same as weights */
  int64_t P;                   /* model length in floats, multiple of 4 */
  int64_t P_train;             /* leading trainable-parameter region (clip / noise / optimizer), multiple of 4 */
  float inv_total;             /* scale applied to the summed weights for [0,P_train) (1 / sum n_i over ALL ranks) */
  float inv_total_buf;         /* same for [P_train, P) */
  float self_coef;             /* avg += self_coef * w_global on [0,P_train) (FedNova: 1 - sum coef_i; else 0) */
  /* global model and cross-rank plumbing */
  float* global_model;         /* [P] in: w(t), out: w(t+1); peers write into it in two-shot mode */
  float* partial;              /* [P] this rank's partial sum, readable by peers */
  float* const* peer_partial;  /* device array [G] of every rank's `partial` (NULL iff G == 1) */
  float* const* peer_global;   /* device array [G] of every rank's `global_model` */
  uint32_t* const* peer_flags; /* device array [G] of every rank's flag buffer: uint32[2][G][FB_ROUND_MAX_BLOCKS] */
  int32_t rank, G;
  uint32_t epoch;              /* strictly increasing across calls, same on all ranks, starts at 1 */
  int32_t two_shot;            /* 0: every rank reduces the whole vector; 1: rank r reduces shard r and pushes it */
  /* DP */
  float clip_bound;            /* <= 0 disables clipping */
  float noise_std;             /* <= 0 disables noise */
  uint64_t seed;
  uint32_t round;
  /* server optimizer on g = (w - avg) * grad_scale */
  int32_t opt;                 /* FB_OPT_* */
  int32_t nesterov;
  float lr, momentum, beta1, beta2, eps, bias_corr1, bias_corr2, grad_scale;
  float* m;                    /* [P_train] first moment / momentum buffer (NULL if unused) */
  float* v;                    /* [P_train] second moment (NULL if unused) */
  /* scratch */
  float* sqnorm_part;          /* [nblocks][C] per-block squared-norm partials (only when clipping) */
  float* clip_scale_out;       /* [C] optional: the clip factor applied to every client (NULL to skip) */
  uint32_t* grid_bar;          /* [1] zero-initialised counter for the intra-rank barrier */
  uint32_t grid_bar_target;    /* nblocks * (number of clipped calls so far + 1) */
  /* NVLS (optional, two-shot only): multicast addresses of `partial` and `global_model` bound over all G ranks.
   * When both are set the shard reduce is one multimem.ld_reduce (in-switch sum) and the redistribution one
   * multimem.st, instead of G peer loads and G peer stores. */
  float* mc_partial;
  float* mc_global;
  float agg_noise_std;         /* > 0: add N(0, agg_noise_std^2) once to the aggregate of [0,P_train) (same distribution
                                  as per-client noise when set to sigma * sqrt(sum_i (w_i*inv_total)^2)) */
} FbRoundParams;

#define FB_ROUND_MAX_BLOCKS 1024

/* Grid used for a model of P floats on G ranks (identical on every rank by construction). */
int fb_round_grid(int64_t P, int32_t G, int32_t emulated_ranks);
/* Bytes of the per-rank flag buffer (zero it once at set-up). */
int64_t fb_round_flag_bytes(int32_t G);
/* One rank of the round.  Every rank must launch with the same P, G, epoch, two_shot. */
int fb_round_launch(const FbRoundParams* p, fb_stream_t stream);
/* Test/emulation entry: run all G ranks as ONE cooperative launch on one GPU (blockIdx.y = rank).
 * `params_dev` is a device array of G FbRoundParams whose peer_* arrays point at local buffers. */
int fb_round_launch_emulated(const FbRoundParams* params_dev, int32_t G, int64_t P, fb_stream_t stream);

/* Allocation-free single-GPU round.  No counterpart in the reference (which has no native code and
 * aggregates on the CPU); it is fb_round_launch for one rank with the clipping scratch built in.
 *
 * Who needs it: hosts that may not allocate device memory when a round is aggregated.  Examples are an
 * embedding application that handed all free HBM to its own pool at start-up (a later cudaMalloc fails or
 * fragments that pool), a loop that cannot accept the device-wide stall cudaMalloc/cudaFree imply, and
 * plug-in code whose host forbids allocator calls after initialisation.  Such callers cannot provide the
 * `sqnorm_part` / `grid_bar` scratch that fb_round_launch asks for, and this entry point does not ask:
 * the scratch is a fixed 256.25 KiB array inside the library image, present on every GPU that loads the
 * library until the process exits.  Inputs and outputs still belong to the caller.  Nothing here calls
 * cudaMalloc, cudaFree or any allocating library routine.
 *
 *   global_model[j] <- sum_c weights[c] * inv_total * (g + s_c * (clients[c][j] - g))   for j <  P_train
 *   global_model[j] <- sum_c weights[c] * inv_total * clients[c][j]                     for j >= P_train
 *   with g = global_model[j] on entry and s_c = 1 / max(1, ||clients[c][:P_train] - g||_2 / clip_bound)
 *   (clip_bound <= 0: s_c = 1).  clip_scale_out (optional, [C]) receives s_c.
 *
 * Limits and rules:
 *  - at most fb_round_single_noalloc_max_clients() (= 64) clients per call; more is refused with
 *    FB_ERR_LIMIT and a message naming fb_round_launch as the way out.  P, P_train, ld as for fb_round_launch.
 *  - The first call on each device looks up the device address of the built-in array (with that device
 *    current) and remembers it; make that first call outside any stream capture.
 *  - All calls on a device share the one array, and it keeps what the previous call left: the entry point
 *    clears the barrier counter itself on `stream` before the launch.  Two calls whose kernels can overlap
 *    on the GPU (different streams, no dependency between them) race on it — serialise them.
 */
int fb_round_single_noalloc_max_clients(void);
int fb_round_single_noalloc(const float* clients, int64_t ld, int32_t C, const float* weights, float inv_total,
                            float clip_bound, float* global_model, int64_t P, int64_t P_train,
                            float* clip_scale_out, fb_stream_t stream);

/* K4 stand-alone: out[j] = beta*out[j] + scale * sum_c w[c] * X[c][j]  (local streaming accumulate). */
int fb_weighted_accumulate(const float* clients, int64_t ld, int32_t C, const float* weights, float scale,
                           float beta, float* out, int64_t P, fb_stream_t stream);
/* bf16 client rows, fp32 accumulate. */
int fb_weighted_accumulate_bf16(const void* clients_bf16, int64_t ld, int32_t C, const float* weights, float scale,
                                float beta, float* out, int64_t P, fb_stream_t stream);
/* rows[c][j] = src[j] for all c: broadcast the global model into the stacked client arena. */
int fb_broadcast_rows(const float* src, float* rows, int64_t ld, int32_t C, int64_t P, fb_stream_t stream);

/* K3 stand-alone server optimizer on one rank: w <- opt(w, avg). */
int fb_server_opt(float* w, const float* avg, float* m, float* v, int64_t n, int32_t opt, int32_t nesterov,
                  float lr, float momentum, float beta1, float beta2, float eps, float bias_corr1, float bias_corr2,
                  float grad_scale, fb_stream_t stream);

/* K2 stand-alone pieces (also used by tests as oracles for the fused path). */
int fb_sqnorm_diff(const float* clients, int64_t ld, int32_t C, const float* ref, int64_t n, float* out_sq,
                   fb_stream_t stream);                                   /* out_sq[c] = ||X[c][:n] - ref||^2 */
int fb_philox_normal(float* out, int64_t n, uint64_t seed, uint32_t stream_id, uint32_t substream,
                     fb_stream_t stream);                                 /* n multiple of 4 */

/* Cross-rank barrier over peer flag buffers (1 block).  `slot` in [0, 8). */
int fb_peer_barrier(uint32_t* const* peer_flags, int32_t rank, int32_t G, uint32_t epoch, int32_t slot,
                    fb_stream_t stream);
int64_t fb_peer_barrier_flag_bytes(int32_t G);

/* ---------------------------------------------------------------------------------------------
 * K5/K6  Coordinate-wise median / trimmed mean over K client rows that may live on different GPUs.
 * Replaces RobustAggregator.coordinate_median_agg (robust_aggregation.py:65-99); trimmed mean is
 * absent from the reference (Yin et al. 2018).  rows[k] may be peer-mapped pointers, so the
 * "transpose" that brings all K values of a coordinate together happens inside the kernel loads.
 * mode 0: lower median (torch.median tie rule); mode 1: mean of the values with rank in [trim, K-trim).
 * Coordinates [j0, j1) are processed and written to every pointer in outs[0..n_out).
 * ------------------------------------------------------------------------------------------- */
int fb_robust_select(const float* const* rows_dev, int32_t K, int64_t j0, int64_t j1, int32_t mode, int32_t trim,
                     float* const* outs_dev, int32_t n_out, fb_stream_t stream);
/* Many clients (K > 128: BASELINE config "512 clients per round"): sorted runs + merge-select.
 * fb_robust_sort_runs sorts, for every coordinate in [j0, j1), each run of `run_len` (<= 128) consecutive rows of the
 * table IN PLACE (register sorting network; the rows must be writable from this GPU — a rank sorts the rows it trained).
 * fb_robust_select_runs then finds the lower median (mode 0) or the trimmed mean (mode 1) with one binary search per run
 * and bit instead of a pass over all K keys.  K rows are read; only the K_valid smallest values per coordinate count
 * (padding rows must hold +inf).  Semantics as fb_robust_select (robust_aggregation.py:65-99, lower median). */
int fb_robust_sort_runs(float* const* rows_dev, int32_t K, int32_t run_len, int64_t j0, int64_t j1, fb_stream_t stream);
int fb_robust_select_runs(const float* const* rows_dev, int32_t K, int32_t K_valid, int32_t run_len, int64_t j0, int64_t j1,
                          int32_t mode, int32_t trim, float* const* outs_dev, int32_t n_out, fb_stream_t stream);

/* ---------------------------------------------------------------------------------------------
 * K7  Krum: pack (centre + bf16 hi/lo split) and Gram matrix on tcgen05 tensor cores, then scores.
 * Absent from the reference (Blanchard et al. 2017).
 * ------------------------------------------------------------------------------------------- */
/* dst_hi/lo[k][j - j0] = split_bf16(rows[k][j] - ref[j]) for j in [j0, j1); dst row stride ld_dst (elements). */
int fb_krum_pack(const float* const* rows_dev, int32_t K, const float* ref, int64_t j0, int64_t j1, void* dst_hi,
                 void* dst_lo, int64_t ld_dst, fb_stream_t stream);
/* gram[K][K] (fp32) += A A^T over the packed columns; split-K partials reduced inside via atomics-free
 * workspace of fb_gram_workspace_bytes.  Uses hi*hi + hi*lo + lo*hi when lo != NULL. */
int64_t fb_gram_workspace_bytes(int32_t K, int64_t n_cols);
int fb_gram_bf16(const void* a_hi, const void* a_lo, int32_t K, int64_t n_cols, int64_t ld, float* gram,
                 void* workspace, fb_stream_t stream);
/* scores[i] = sum of the (K - f - 2) smallest d2(i, j), j != i, with d2 from the Gram matrix. */
int fb_krum_scores(const float* gram, int32_t K, int32_t f, float* scores, fb_stream_t stream);

/* ---------------------------------------------------------------------------------------------
 * K8  Secure-aggregation arithmetic in GF(p), p < 2^31 (reference: turboaggregate/mpc_function.py:4-281).
 * ------------------------------------------------------------------------------------------- */
/* q[i] = round(x[i] * scale) mod p  (negative values wrap to p - |.|);  inverse for dequantize. */
int fb_ff_quantize(const float* x, int64_t n, float scale, uint32_t p, uint32_t* q, fb_stream_t stream);
int fb_ff_dequantize(const uint32_t* q, int64_t n, float inv_scale, uint32_t p, float* x, fb_stream_t stream);
/* out[i][d] = sum_j U[i][j] * X[j][d] mod p  — LCC/BGW encode and decode (mpc_function.py:62-75,111-136,201-217). */
int fb_ff_matmul(const uint32_t* U, int32_t n_out, int32_t n_in, const uint32_t* X, int64_t d, int64_t ldx,
                 uint32_t p, uint32_t* out, int64_t ldo, fb_stream_t stream);
/* y = x + sum_{s} sign[s] * PRG(seed[s], nonce) mod p — additive masks expanded in registers; `nonce` (the round index)
 * enters the Philox counter so no mask is ever used twice. */
int fb_ff_mask(const uint32_t* x, int64_t n, const uint64_t* seeds_dev, const int32_t* signs_dev, int32_t n_seeds,
               uint32_t p, uint32_t nonce, uint32_t* y, fb_stream_t stream);
/* acc = (acc + x) mod p */
int fb_ff_add(uint32_t* acc, const uint32_t* x, int64_t n, uint32_t p, fb_stream_t stream);

/* ---------------------------------------------------------------------------------------------
 * K11  Fused multi-client optimizer step and losses (reference: one torch.optim kernel per tensor per
 * step, my_model_trainer_classification.py:26-37; CrossEntropyLoss, BCELoss, KLDivLoss fedgkt/utils.py:78-124).
 * ------------------------------------------------------------------------------------------- */
/* For every client c and j < n:  d = g + wd*w (+ momentum) + mu*(w - anchor[j]);  w -= lr[c]*d.
 * Optional FedNova accumulators: cum[c][j] += lr[c]*d.  lr[c] == 0 freezes a client. Rows share stride ld. */
int fb_sgd_step_stacked(float* w, const float* g, int64_t ld, int32_t C, int64_t n, const float* lr_dev,
                        float weight_decay, float momentum, float dampening, int32_t nesterov, float* mom_buf,
                        int32_t first_step, float mu, const float* anchor, float* cum_grad, fb_stream_t stream);
/* Same for bf16 gradients (fp32 master weights). */
int fb_sgd_step_stacked_bf16g(float* w, const void* g_bf16, int64_t ld, int64_t ld_g, int32_t C, int64_t n,
                              const float* lr_dev, float weight_decay, float mu, const float* anchor,
                              fb_stream_t stream);
/* Softmax cross-entropy over rows: logits [R][ncls] fp32, labels int64 [R] (ignore_index skips a row).
 * loss_rows[R] (optional), dlogits[R][ncls] = (softmax - onehot) * row_scale[r / rows_per_group]
 * (row_scale NULL -> 1/R). pre_sigmoid != 0 applies sigmoid to the logits first (the reference's LR model,
 * model/linear/lr.py:10) and back-propagates through it. */
int fb_softmax_xent(const float* logits, const int64_t* labels, int64_t R, int32_t ncls, int64_t ignore_index,
                    int32_t pre_sigmoid, const float* row_scale, int64_t rows_per_group, float* loss_rows,
                    float* dlogits, fb_stream_t stream);
/* One-launch staging of every conv / linear weight of every client for a training step: fp32 master rows of
 * the stacked arena -> bf16 GEMM operands in (kh,kw,cin) order (+ transposed copies for dgrad).  table_dev: int64
 * [n_layers][8] = {src_off, cout, cin, cin_pad, k, dst_off, dstT_off or -1, cout*k*k*cin_pad}.  Replaces ~3 small
 * cast/permute kernels per layer per step. */
int fb_stage_conv_weights(const float* arena, int64_t ld, int32_t C, void* stage_bf16, int64_t stage_ld,
                          const int64_t* table_dev, int32_t n_layers, int32_t max_n, fb_stream_t stream);
/* Plain-SGD optimizer step fused with the re-staging of the next step's GEMM operands: for every listed segment of every
 * client  w -= lr[c] * g, then the updated weights are written back AND cast / re-laid-out into the bf16 staging stack in
 * the same pass; k x k weight gradients are read in the GEMM layout from the scratch stack of the weight-gradient
 * kernels (no un-staging pass).  table_dev: int64 [n_layers][10] = {w_off, cout, cin, cin_pad, k, dst | -1, dstT | -1,
 * n, g_off, g_kind (0 arena / 1 scratch)}.  Replaces torch.optim.SGD.step() of every client
 * (my_model_trainer_classification.py:26-37) plus this library's own staging and un-staging launches. */
int fb_sgd_restage(float* arena, const float* grads, int64_t ld, int32_t C, const float* scratch, int64_t scratch_ld,
                   void* stage, int64_t stage_ld, const int64_t* table_dev, int32_t n_layers, int32_t total_elems,
                   const float* lr_dev, fb_stream_t stream);
/* The inverse direction for k x k weight gradients: GEMM-layout fp32 [cout][(kh,kw,cin_pad)] rows of a scratch stack
 * -> parameter layout [cout][cin][kh][kw] inside the gradient arena, all listed layers of all clients in one launch.
 * table_dev: int64 [n_layers][8] = {src_off, cout, cin, cin_pad, k, dst_off, 0, 0}; max_tiles = largest
 * ceil(cout/32)*ceil(cin/32) of a layer. */
int fb_unstage_conv_grads(const float* scratch, int64_t scratch_ld, float* grads, int64_t ld, int32_t C,
                          const int64_t* table_dev, int32_t n_layers, int32_t max_tiles, fb_stream_t stream);
/* K15 fused evaluation: counters[0]=correct, [1]=loss_sum, [2]=total, then tp[ncls], fp[ncls], fn[ncls]
 * (all float64-free: uint32 counts in `counts`, float loss in `loss_sum`). Accumulates into the outputs. */
int fb_eval_counts(const float* logits, const int64_t* labels, int64_t R, int32_t ncls, int32_t pre_sigmoid,
                   uint32_t* counts /* [2 + 3*ncls]: correct,total,tp..,fp..,fn.. */, float* loss_sum,
                   fb_stream_t stream);

/* ---------------------------------------------------------------------------------------------
 * K9  Grouped GEMM on tcgen05 tensor cores with TMEM accumulators and TMA-fed operands.
 * For every group g:  D_g[M][N] = A_g[M][K] * B_g[N][K]^T  (bf16 in, fp32 accumulate, bf16 or fp32 out).
 * Replaces per-client cuBLAS nn.Linear / 1x1-conv calls.
 * ------------------------------------------------------------------------------------------- */
typedef struct {
  const void* a;      /* bf16 [G][M][K], K contiguous; group stride a_gs elements */
  const void* b;      /* bf16 [G][N][K], K contiguous; group stride b_gs elements */
  void* d;            /* [G][M][N], N contiguous; group stride d_gs elements */
  const float* bias;  /* optional fp32 [G][N] (group stride bias_gs) */
  int64_t a_gs, b_gs, d_gs, bias_gs;
  int32_t G, M, N, K; /* K multiple of 8 (TMA 16-byte rows); any M, N */
  int32_t lda, ldb, ldd; /* row strides in elements (multiples of 8 for bf16, 4 for fp32 out) */
  int32_t d_fp32;     /* 1: fp32 output, 0: bf16 output */
  int32_t relu;       /* fused ReLU epilogue */
  int32_t accumulate; /* D += result (fp32 output only; TMA reduce-store, so concurrent groups may target one output) */
  int32_t out_group_div; /* > 1: groups g and g' with g / div == g' / div write the same output block (split-K slices) */
  int32_t group_inner;   /* > 0: two-level groups, g = outer * group_inner + inner; operand base = outer * x_gs2 +
                          * inner * x_gs.  Lets attention heads be addressed in place inside [batch][seq][heads*dh]
                          * buffers (4-D tensor maps) without split/merge-heads copies.  0: one level (x_gs only). */
  int64_t a_gs2, b_gs2, d_gs2;
  /* optional per-column statistics of the output, accumulated (atomic adds; zero it first) by the epilogue:
   * stats[g][0][n] += sum_m D_g[m][n], stats[g][1][n] += sum_m D_g[m][n]^2 over the values as stored; group stride stats_gs
   * floats (>= 2*N).  These are the BatchNorm batch statistics of a convolution's output, so the normalisation
   * that follows needs no statistics pass of its own (reference op: nn.BatchNorm2d after every conv,
   * model/cv/resnet.py:116-136).  Needs a 16-byte aligned output, one-level groups, no split-K.  NULL: off. */
  float* stats;
  int64_t stats_gs;
} FbGroupedGemm;
int fb_grouped_gemm_bf16(const FbGroupedGemm* p, fb_stream_t stream);
/* Weight-gradient form: D_g[M][N] = A_g^T * B_g with A_g stored [K][M] and B_g stored [K][N] (M resp. N contiguous,
 * row strides lda/ldb); the operands are fed to the tensor cores MN-major, so activations need no transpose. */
int fb_grouped_gemm_tn_bf16(const FbGroupedGemm* p, fb_stream_t stream);
/* Implicit 3x3 convolution (pad 1) on the same tensor-core pipeline: no im2col / col2im buffers.  For every group g
 * (client) and image b < B:  y[g][b][ho][wo][:] = sum_{kh,kw} x[g][b][ho*s+kh-1][wo*s+kw-1][:] * W_g[kh][kw].
 * The producer warp gathers each tap's shifted input window with one 4-D TMA box (padding = out-of-bounds zero fill,
 * stride = element strides of the tensor map); nine taps x channel blocks accumulate into one TMEM tile.
 *   forward:        w = bf16 [G][cout][9*cin]  ((kh,kw,cin) column order, as staged for the explicit path)
 *   data gradient:  call with x = dy [.., cout_fwd], cin = cout_fwd, cout = cin_fwd, dgrad = 1 and
 *                   w = bf16 [G][9*cin_fwd][cout_fwd] (the transposed staging copy); taps are flipped (stride 1 only).
 * cin in {16, 32, 64k}; cout % 8 == 0; output plane sides powers of two, width <= 128.  Replaces
 * nn.Conv2d(3x3) forward / input-gradient calls of the reference's ResNets (model/cv/resnet.py:48-51). */
typedef struct {
  const void* x;     /* bf16 [G*B][H][W][cin] */
  const void* w;     /* see above; group stride w_gs elements */
  void* y;           /* bf16 [G*B][Ho][Wo][cout] */
  int64_t w_gs;
  int32_t G, B, H, W, cin, cout, stride, dgrad;
  float* stats;      /* optional [G][2][cout] output statistics, as in FbGroupedGemm; NULL: off */
  int64_t stats_gs;
} FbConv3x3;
int fb_conv3x3_bf16(const FbConv3x3* p, fb_stream_t stream);
/* Weight gradient of the same convolution, also without an im2col buffer:
 *   dw[g][co][(kh,kw,ci)] (+)= sum_{b,ho,wo} dy[g][b][ho][wo][co] * x[g][b][ho*s+kh-1][wo*s+kw-1][ci]
 * dy is fed MN-major as in fb_grouped_gemm_tn_bf16, the shifted input windows are gathered MN-major by TMA; a tile is one
 * (tap, channel chunk) of one client.  split > 1 cuts the pixel reduction into slices that reduce-add (TMA) into dw,
 * which must then be zero on entry; accumulate != 0 adds to dw as well.  Output plane sides powers of two, width <= 64;
 * B*Ho*Wo % (64*split) == 0. */
typedef struct {
  const void* x;     /* bf16 [G*B][H][W][cin] */
  const void* dy;    /* bf16 [G*B][Ho][Wo][cout] */
  float* dw;         /* fp32 [G][cout][9*cin], group stride dw_gs elements */
  int64_t dw_gs;
  int32_t G, B, H, W, cin, cout, stride, split, accumulate;
} FbConv3x3Wgrad;
int fb_conv3x3_wgrad_bf16(const FbConv3x3Wgrad* p, fb_stream_t stream);
/* Row-major x row-major form: D_g[M][N] = A_g[M][K] * B_g[K][N] (A K-major, B fed MN-major, row stride ldb).
 * Data gradients of nn.Linear (dX = dY * W with W stored [out][in]) and attention's P*V / dS*K products use
 * it, so neither weights nor activations are ever transposed in memory. */
int fb_grouped_gemm_nn_bf16(const FbGroupedGemm* p, fb_stream_t stream);

/* ---------------------------------------------------------------------------------------------
 * K12  Client-stacked transformer-encoder kernels (DistilBERT: reference model/nlp via FedNLP,
 * python/fedml/model/nlp + app/fednlp).  Activations are bf16 [C clients][R = B*S rows][D]; parameters are
 * fp32 rows of the stacked arena (row stride ld elements).  The projections and the attention products run
 * on the grouped GEMMs above (two-level (batch, head) groups address Q/K/V heads in place); these kernels
 * are the single-pass glue between them.  Replaces the per-client nn.Embedding / nn.LayerNorm /
 * softmax / GELU / dropout ATen launches of `MyModelTrainer.train` (my_model_trainer_classification.py:21-70
 * driving a transformers DistilBertForSequenceClassification).
 * ------------------------------------------------------------------------------------------- */
/* out[c][r][:] = word[c][ids[c][r]][:] + pos[c][r % S][:]   (ids int64 [C][R], tables fp32 in the arena) */
int fb_embed_gather(const int64_t* ids, const float* word, const float* pos, int64_t ld, void* out_bf16, int32_t C,
                    int64_t R, int32_t S, int32_t D, int32_t vocab, fb_stream_t stream);
/* gword[c][ids[c][r]][:] += word_scale[c] * dz[c][r][:] (fp32 atomics; word_scale NULL = 1);  gpos[c][s][:] =
 * sum_b dz.  With gword = the (pre-zeroed) gradient rows this is the embedding backward; with gword = the
 * word-embedding PARAMETER rows and word_scale[c] = -lr[c] it is the fused sparse SGD update of plain SGD: only the
 * rows of tokens that occur are touched and the dense vocab x D gradient is never materialised. */
int fb_embed_bwd(const int64_t* ids, const void* dz_bf16, float* gword, float* gpos, int64_t ld, int32_t C, int64_t R,
                 int32_t S, int32_t D, int32_t vocab, const float* word_scale, fb_stream_t stream);
/* z = a + b (b, z optional); y = LayerNorm(z) * gamma + beta; saves per-row mean / rstd.  D % 8 == 0, D <= 1024. */
int fb_add_ln_fwd(const void* a, const void* b, void* z, void* y, float* mean, float* rstd, const float* gamma,
                  const float* beta, int64_t ld_p, int32_t C, int64_t R, int32_t D, float eps, fb_stream_t stream);
int64_t fb_ln_bwd_scratch_floats(int32_t C, int64_t R, int32_t D);
/* dz from dy (+ dy2, an optional second gradient stream summed on load) and the saved z / mean / rstd;
 * dgamma / dbeta fp32 [C][D] with row stride ld_g (e.g. rows of the gradient arena); deterministic. */
int fb_ln_bwd(const void* dy, const void* dy2, const void* z, const float* mean, const float* rstd, const float* gamma,
              int64_t ld_p, void* dz, float* dgamma, float* dbeta, int64_t ld_g, float* scratch, int32_t C, int64_t R,
              int32_t D, fb_stream_t stream);
/* P = softmax(scale * scores + key mask) over rows of fp32 scores [groups][S][S]; group = (c*B + b) * H + h;
 * key_mask fp32 [C*B][S] (1 = attend) or NULL.  S % 8 == 0, S <= 512.  With p > 0 the attention-dropout copy
 * Pdrop = dropout(P) is written as well (mask keyed (seed, *step_dev, site), see fb_dropout_bf16). */
int fb_attn_softmax_fwd(const float* scores, void* P_bf16, void* Pdrop_bf16, const float* key_mask, int64_t groups,
                        int32_t S, int32_t H, float scale, float p, uint64_t seed, uint32_t site,
                        const uint32_t* step_dev, fb_stream_t stream);
/* dS = scale * P * (dP' - rowsum(dP' * P)), dP' = dropout-backward(dP) under the same key */
int fb_attn_softmax_bwd(const float* dP, const void* P_bf16, void* dS_bf16, int64_t groups, int32_t S, float scale,
                        float p, uint64_t seed, uint32_t site, const uint32_t* step_dev, fb_stream_t stream);
/* elementwise activation on bf16 (mode 0 ReLU, 1 exact GELU) and its backward from the pre-activation x */
int fb_act_fwd(const void* x, void* y, int64_t n, int32_t mode, fb_stream_t stream);
int fb_act_bwd(const void* dy, const void* x, void* dx, int64_t n, int32_t mode, fb_stream_t stream);
/* y = dropout(x (+ x2)): inverted dropout, Philox mask keyed (seed, *step_dev, site) and indexed by element, so
 * calling it on the gradient with the same key is the backward pass.  step_dev (device uint32, may be NULL) lets
 * a captured CUDA graph draw fresh masks on every replay. */
int fb_dropout_bf16(const void* x, const void* x2, void* y, int64_t n, float p, uint64_t seed, uint32_t site,
                    const uint32_t* step_dev, fb_stream_t stream);
/* bias gradients: out[c][n] = sum_r x[c][r][n], x bf16 rows of stride ldx, clients client_stride apart */
int fb_colsum_bf16(const void* x, int32_t C, int64_t R, int32_t N, int64_t ldx, int64_t client_stride, float* out,
                   int64_t ld_o, fb_stream_t stream);
/* fp32 forms of the four element-wise kernels above (tf32 / fp32 activation pipelines, e.g. the FEMNIST CNN). */
int fb_act_fwd_f32(const void* x, void* y, int64_t n, int32_t mode, fb_stream_t stream);
int fb_act_bwd_f32(const void* dy, const void* x, void* dx, int64_t n, int32_t mode, fb_stream_t stream);
int fb_dropout_f32(const void* x, const void* x2, void* y, int64_t n, float p, uint64_t seed, uint32_t site,
                   const uint32_t* step_dev, fb_stream_t stream);
int fb_colsum_f32(const void* x, int32_t C, int64_t R, int32_t N, int64_t ldx, int64_t client_stride, float* out,
                  int64_t ld_o, fb_stream_t stream);

/* ---------------------------------------------------------------------------------------------
 * K10 helpers: NHWC im2col / col2im for 3x3 convolutions batched over clients, fused per-client
 * BatchNorm (+ReLU, +residual) forward/backward on [C][M][ch] activations.
 * ------------------------------------------------------------------------------------------- */
int fb_im2col3x3_nhwc_bf16(const void* x, void* cols, int32_t n_img, int32_t H, int32_t W, int32_t ch,
                           int32_t stride, fb_stream_t stream);
int fb_col2im3x3_nhwc_bf16(const void* cols, void* dx, int32_t n_img, int32_t H, int32_t W, int32_t ch,
                           int32_t stride, fb_stream_t stream);
/* y = relu?(gamma*(x-mean)*rstd + beta (+res)); statistics per (client, channel) over M rows.
 * x,y,res: bf16 [C][M][ch] (ch a power of two in [8, 2048]); gamma/beta: fp32 rows with stride ld_p;
 * running_mean/var: fp32 rows with stride ld_r, updated when training (scaled by active[c], NULL = all active);
 * save_mean/save_rstd: fp32 [C][ch]; scratch: fb_bn_scratch_floats(C, M, ch) floats (training only).
 * Replaces nn.BatchNorm2d + ReLU + residual add of every reference block (model/cv/resnet.py:116-136). */
int64_t fb_bn_scratch_floats(int32_t C, int64_t M, int32_t ch);
int fb_bn_fwd_stacked(const void* x, const void* res, void* y, int32_t C, int64_t M, int32_t ch, const float* gamma,
                      const float* beta, int64_t ld_p, float* running_mean, float* running_var, int64_t ld_r,
                      const float* active, float momentum, float eps, int32_t relu, int32_t training,
                      float* save_mean, float* save_rstd, float* scratch, float* num_batches_tracked /* fp32 [C] rows, stride ld_n; may be NULL */,
                      int64_t ld_n, fb_stream_t stream);
/* dx (and dres = ReLU-masked dy, optional) from dy; dgamma/dbeta fp32 rows with stride ld_g.  The ReLU mask comes
 * from y (forward output; required when the forward had a residual input) or, when y is NULL, is recomputed from
 * x, gamma and beta so that the output tensor is not streamed again. */
int fb_bn_bwd_stacked(const void* dy, const void* dy2 /* optional second gradient summed on load */, const void* x,
                      const void* y, void* dx, void* dres, int32_t C, int64_t M,
                      int32_t ch, const float* gamma, const float* beta, int64_t ld_p, const float* save_mean,
                      const float* save_rstd,
                      int32_t relu, float* dgamma, float* dbeta, int64_t ld_g, float* scratch, fb_stream_t stream);

/* ---------------------------------------------------------------------------------------------
 * K13  Cheetah data-parallel gradient all-reduce fused with the SGD(+momentum) step
 * (reference: DDP bucketed NCCL all-reduce + torch.optim, cross_silo/hierarchical/trainer_dist_adapter.py:66).
 * Rank r reduces shard r of the gradient bucket from all peers, updates the fp32 master shard and
 * pushes the new parameters to every rank — one launch, no NCCL.
 * ------------------------------------------------------------------------------------------- */
typedef struct {
  float* const* peer_grad;    /* [G] each rank's gradient bucket (fp32 [n]) */
  float* const* peer_param;   /* [G] each rank's parameter bucket (fp32 [n]) */
  uint32_t* const* peer_flags;/* [G] flag buffers (fb_round_flag_bytes) */
  float* mom;                 /* [n] momentum (only shard of this rank is touched), may be NULL */
  int64_t n;                  /* multiple of 4 */
  int32_t rank, G;
  uint32_t epoch;
  float lr, momentum, weight_decay, inv_world;
  int32_t nesterov;
  int32_t max_blocks;         /* > 0: cap on the grid (a bucket reduced while the backward pass still runs) */
  int64_t offset;             /* elements: the launch covers [offset, offset + n) of the buckets, momentum included */
} FbDpStepParams;
int fb_dp_allreduce_step(const FbDpStepParams* p, fb_stream_t stream);
int fb_dp_allreduce_step_emulated(const FbDpStepParams* params_dev, int32_t G, int64_t n, fb_stream_t stream);

/* ---------------------------------------------------------------------------------------------
 * K14  HS-FedAvg amplitude normalisation (reference: single_process/hs_fedavg/hs_fft.py:8-84):
 * per image/channel 2-D DFT, swap amplitude with the running mean amplitude, inverse DFT.
 * H, W powers of two up to 512. amp_run [ch][H][W] updated with `momentum` when update_running != 0.
 * ------------------------------------------------------------------------------------------- */
int fb_amp_normalize(const float* x, float* y, int32_t n_img, int32_t ch, int32_t H, int32_t W, float* amp_run,
                     float* amp_batch_mean, float momentum, int32_t update_running, int32_t have_running,
                     void* workspace, fb_stream_t stream);
int64_t fb_amp_normalize_workspace_bytes(int32_t n_img, int32_t ch, int32_t H, int32_t W);

/* ---------------------------------------------------------------------------------------------
 * fp32-activation / TF32 tensor-core variants (the precision the reference trains at: stock nn.Conv2d under
 * cuDNN's TF32 default, fp32 activations; simulation/single_process/fedavg/my_model_trainer_classification.py:18-91,
 * model/cv/resnet.py:116-136).  Same structs and conventions as the bf16 entry points; operands and outputs are fp32,
 * leading dimensions / group strides are in fp32 elements (multiples of 4), products run on tcgen05 kind::tf32.
 *   - grouped GEMMs: k-blocks of 32 elements; MN-major operands use the 128-byte-swizzle / 32-byte-atom layout, the
 *     only shared-memory layout tcgen05 accepts for 32-bit MN-major operands;
 *   - fb_conv3x3_tf32: cin in {8, 16} or a multiple of 32; fb_conv3x3_wgrad_tf32: cin a multiple of 32
 *     (<= 256, or a multiple of 256), B*Ho*Wo % (32*split) == 0;
 *   - BatchNorm / im2col / col2im / staging kernels: the same kernels instantiated for fp32 activations; the staged
 *     GEMM operands of fb_stage_conv_weights_f32 / fb_sgd_restage_f32 are fp32 (1x1-convolution and fc weights need
 *     no staging in this mode: the [cout][cin] rows of the arena are the GEMM operand).
 * ------------------------------------------------------------------------------------------- */
int fb_grouped_gemm_tf32(const FbGroupedGemm* p, fb_stream_t stream);
int fb_grouped_gemm_tn_tf32(const FbGroupedGemm* p, fb_stream_t stream);
int fb_grouped_gemm_nn_tf32(const FbGroupedGemm* p, fb_stream_t stream);
int fb_conv3x3_tf32(const FbConv3x3* p, fb_stream_t stream);
int fb_conv3x3_wgrad_tf32(const FbConv3x3Wgrad* p, fb_stream_t stream);
int fb_bn_fwd_stacked_f32(const void* x, const void* res, void* y, int32_t C, int64_t M, int32_t ch, const float* gamma,
                          const float* beta, int64_t ld_p, float* running_mean, float* running_var, int64_t ld_r,
                          const float* active, float momentum, float eps, int32_t relu, int32_t training, float* save_mean,
                          float* save_rstd, float* scratch, float* num_batches_tracked, int64_t ld_n, fb_stream_t stream);
int fb_bn_bwd_stacked_f32(const void* dy, const void* dy2, const void* x, const void* y, void* dx, void* dres, int32_t C,
                          int64_t M, int32_t ch, const float* gamma, const float* beta, int64_t ld_p, const float* save_mean,
                          const float* save_rstd, int32_t relu, float* dgamma, float* dbeta, int64_t ld_g, float* scratch,
                          fb_stream_t stream);
int fb_im2col3x3_nhwc_f32(const void* x, void* cols, int32_t n_img, int32_t H, int32_t W, int32_t ch, int32_t stride,
                          fb_stream_t stream);
int fb_col2im3x3_nhwc_f32(const void* cols, void* dx, int32_t n_img, int32_t H, int32_t W, int32_t ch, int32_t stride,
                          fb_stream_t stream);
int fb_stage_conv_weights_f32(const float* arena, int64_t ld, int32_t C, void* stage_f32, int64_t stage_ld,
                              const int64_t* table_dev, int32_t n_layers, int32_t max_n, fb_stream_t stream);
int fb_sgd_restage_f32(float* arena, const float* grads, int64_t ld, int32_t C, const float* scratch, int64_t scratch_ld,
                       void* stage_f32, int64_t stage_ld, const int64_t* table_dev, int32_t n_layers, int32_t total_elems,
                       const float* lr_dev, fb_stream_t stream);

/* Race-detection aid for the round kernel's flag protocol: every (rank, block, handshake) stalls a pseudo-random time below
 * max_ns (Philox of seed) before it signals; 0 turns it off.  Results must not depend on it. */
int fb_round_set_skew(uint32_t seed, uint32_t max_ns);

/* Debugging aid: when `buf_u64` (device memory, 8 x uint64 per CTA of the next resident BatchNorm forward launch) is set,
 * thread 0 of every CTA stores %globaltimer at [start, loaded, statistics, barrier, coefficients, applied, stored] and its
 * SM id; NULL turns it off.  tools/trace_bn.py prints the phase table. */
int fb_bn_set_trace(void* buf_u64);

/* 2x2 / stride-2 max pooling on NHWC bf16 or fp32 (act_fp32) maps [n_img][H][W][ch], ch a multiple of 8.  The pooled
 * windows start at (off_h, off_w): y [n_img][Ho][Wo][ch] = max of x[off_h + 2i + {0,1}][off_w + 2j + {0,1}].  Backward
 * writes the whole dx [n_img][H][W][ch] (zero outside the windows), routes dy to the first maximal element of a window and,
 * with relu_mask, drops it where that maximum is <= 0.  Reference: nn.MaxPool2d(2, 2) in model/cv/cnn.py:74-142. */
int fb_maxpool2x2_fwd(const void* x, void* y, int32_t act_fp32, int32_t n_img, int32_t H, int32_t W, int32_t ch,
                      int32_t off_h, int32_t off_w, int32_t Ho, int32_t Wo, fb_stream_t stream);
int fb_maxpool2x2_bwd(const void* dy, const void* x, void* dx, int32_t act_fp32, int32_t n_img, int32_t H, int32_t W,
                      int32_t ch, int32_t off_h, int32_t off_w, int32_t Ho, int32_t Wo, int32_t relu_mask, fb_stream_t stream);

/* GroupNorm (+residual, +ReLU) forward / backward for client-stacked conv nets, on the BatchNorm kernels: a row is one
 * (client, sample), rows = C*B, M = H*W pixels, statistics pool `cpg` adjacent channels of a row (reference
 * model/cv/group_normalization.py:7-93 and resnet_gn.py:108-196, two channels per group for ResNet-18-GN); gamma / beta
 * belong to client row / rows_per_client.  x, res, y, dy, dx, dres: bf16 (act_fp32 = 0) or fp32 [rows][M][ch];
 * save_mean / save_rstd fp32 [rows][ch]; scratch as fb_bn_scratch_floats(rows, M, ch); backward writes per-row
 * dgamma_rows / dbeta_rows [rows][ch] which the caller sums over the samples of a client. */
int fb_gn_fwd_stacked(const void* x, const void* res, void* y, int32_t act_fp32, int32_t rows, int64_t M, int32_t ch,
                      const float* gamma, const float* beta, int64_t ld_p, int32_t rows_per_client, int32_t cpg, float eps,
                      int32_t relu, float* save_mean, float* save_rstd, float* scratch, fb_stream_t stream);
int fb_gn_bwd_stacked(const void* dy, const void* dy2, const void* x, const void* y, void* dx, void* dres, int32_t act_fp32,
                      int32_t rows, int64_t M, int32_t ch, const float* gamma, const float* beta, int64_t ld_p,
                      int32_t rows_per_client, int32_t cpg, const float* save_mean, const float* save_rstd, int32_t relu,
                      float* dgamma_rows, float* dbeta_rows, float* scratch, fb_stream_t stream);

/* Train-time augmentation of uint8 NHWC image batches: RandomCrop(H, padding = pad, black border) + horizontal flip +
 * Cutout(cutout x cutout, filled with fill_rgba = the per-channel dataset mean, byte c in bits [8c, 8c+8)).  Every
 * random choice of sample i comes from Philox4x32-10(counter = (first_index + i, tag, nonce), key = seed); when
 * nonce_dev is not NULL the nonce is read from that device int64 (a captured CUDA graph then augments differently at
 * every replay).  Out of place.  Replaces the torchvision transform chain the reference runs per sample on the host
 * (data/cifar10/data_loader.py:58-103). */
int fb_augment_cifar(const uint8_t* x, uint8_t* y, int64_t n_img, int32_t H, int32_t W, int32_t C, uint64_t seed, uint32_t nonce,
                     const int64_t* nonce_dev, int32_t pad, int32_t cutout, uint32_t fill_rgba, int64_t first_index,
                     fb_stream_t stream);

/* Fused classifier head of a client-stacked conv net, forward AND backward in one launch (one CTA per client):
 * global average pool -> fc (+bias) -> softmax cross-entropy -> d(logits) -> fc weight / bias gradients -> d(features)
 * broadcast back over the H x W positions.  h / dh: [C][B][HW][ch] in bf16 (act_fp32 = 0) or fp32 (1); W, bias, gW, gb:
 * fp32 rows of the parameter / gradient arenas (client strides ld_*); labels int64 [C][B], rows with a label outside
 * [0, ncls) are padding; row_scale [C] (NULL: 1/B); loss [C] = row_scale * sum of the rows' cross-entropy.
 * Replaces avgpool + fc + F.cross_entropy + their autograd nodes (model/cv/resnet.py:226-232,
 * my_model_trainer_classification.py:44-52) and keeps the 10-column fc products off the library GEMM. */
int fb_head_fwd_bwd(const void* h, int32_t act_fp32, int32_t C, int32_t B, int32_t HW, int32_t ch, const float* W, int64_t ld_w,
                    const float* bias, int64_t ld_b, int32_t ncls, const int64_t* labels, const float* row_scale, float* loss,
                    void* dh, float* gW, int64_t ld_gw, float* gb, int64_t ld_gb, fb_stream_t stream);

/* ---------------------------------------------------------------------------------------------
 * Bounded cross-rank waits (failure detection; the reference has none on this path: a dead MPI worker hangs
 * FedAVGAggregator.check_whether_all_receive forever, simulation/mpi_p2p_mp/fedavg/FedAVGAggregator.py:58-66).
 * Every wait on a flag another rank or block writes (fb_round_launch, fb_peer_barrier, fb_dp_allreduce_step and
 * their emulated forms) gives up after the time-out (default 20 s), records who it was waiting for and lets the
 * kernel retire.  *_spin_status synchronises with the device, copies {wait kind, rank waited for, expected flag
 * value, block} to out4 (may be NULL), clears the record when `clear` != 0 and returns FB_ERR_TIMEOUT if a wait
 * gave up since the last clear — the data produced by that launch must then be discarded.
 * ------------------------------------------------------------------------------------------- */
#define FB_ERR_TIMEOUT (-4)
int fb_round_set_spin_timeout_ms(int64_t ms);
int fb_round_spin_status(uint32_t* out4, int32_t clear);
int fb_dp_set_spin_timeout_ms(int64_t ms);
int fb_dp_spin_status(uint32_t* out4, int32_t clear);

#ifdef __cplusplus
}
#endif
#endif /* FEDML_B200_H */
