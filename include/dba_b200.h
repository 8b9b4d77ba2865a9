/* C ABI of libdba_b200.so - the hand-written sm_100a kernels of the dba_mod_b200 framework.
 *
 * Conventions: every function returns 0 on success, non-zero on failure (dba_last_error() has the text).  All device
 * pointers are plain CUDA device addresses; `stream` is a cudaStream_t passed as void*.  Nothing here allocates device
 * memory: callers own every buffer (sizes are stated per function).  "Reference" = ehsan886/DBA_mod, file:line.
 */
#ifndef DBA_B200_H
#define DBA_B200_H

#ifdef __cplusplus
extern "C" {
#endif

const char* dba_last_error(void);
unsigned long long dba_launch_count(void);       /* kernels launched by this library since the last reset */
void dba_reset_launch_count(void);
void dba_add_launch_count(unsigned long long n); /* kernels launched through a replayed CUDA graph */

/* ---- flat-slab kernels (csrc/flat_ops.cu) ------------------------------------------------------------------- */

/* SGD with momentum + weight decay for K client rows at once, optional FoolsGold raw-gradient accumulation.
 * Reference: torch.optim.SGD use in image_train.py:33-35,102 and the grad-sum loop image_train.py:94-100. */
int dba_sgd_step(float* W, long long ldw, float* V, const float* Gr, float* S, long long ldv, const float* lr,
                 const unsigned char* active, int K, long long Dp, float mu, float wd, void* stream);

/* Gradient of the distance-regularised poison loss alpha*CE + (1-alpha)*||w - w_anchor|| on top of the CE gradient.
 * Reference: image_train.py:85-90, helper.py:111-123. */
int dba_dist_grad(float* Gr, long long ldv, const float* W, long long ldw, const float* A, long long lda,
                  const double* dist, const long long* poison_n, const unsigned char* active, const long long* slot,
                  float* mf_dist, int ld_m, float alpha, int K, long long Dp, void* stream);

/* Poison scaling W[k] = A[k] + (W[k]-A[k])*s[k].  Reference: image_train.py:166-171, loan_train.py:154-160. */
int dba_scale_rows(float* W, long long ldw, const float* A, long long lda, const float* s, int K, long long D,
                   void* stream);

/* out[k] = ||X[k] - Y[k] - Z|| (or its square), fp64, deterministic two-stage reduction through `scratch`
 * (dba_rows_sqdist_scratch_elems(K) doubles).  Reference: helper.py:60-71 (norms), 376-381 (l2dist). */
long long dba_rows_sqdist_scratch_elems(int K);
int dba_rows_sqdist(const float* X, long long ldx, const float* Y, long long ldy, const float* Z, int K, long long n,
                    double* out, int take_sqrt, double* scratch, long long scratch_elems, void* stream);

/* dba_rows_sqdist without the scratch argument, for callers that cannot allocate device memory (all of it is in
 * their own pool, or cudaMalloc's device-wide synchronisation is not acceptable where they run).  The partial sums
 * go to a 64 KiB array built into the library, so: at most dba_rows_sqdist_noalloc_max_rows() (64) rows per call,
 * larger calls are refused; the first call on a device resolves the array's address and should happen before any
 * stream capture; calls that may overlap on the GPU race on the shared array; the device owning the buffers / stream
 * must be the calling thread's current device; callable from several host threads (address cache idempotent, launch
 * counter atomic, error text per thread).  No counterpart in the reference
 * (it has no native code and no scratch buffers). */
int dba_rows_sqdist_noalloc_max_rows(void);
int dba_rows_sqdist_noalloc(const float* X, long long ldx, const float* Y, long long ldy, const float* Z, int K,
                            long long n, double* out, int take_sqrt, void* stream);

/* FedAvg: G += coef * sum_k (W[k]-G) (+ sigma*N(0,1)), optionally writes the new G back into every row of W.
 * Reference: helper.py:193-257 (accumulate_weight + average_shrink_models), dp_noise helper.py:187-191. */
int dba_fedavg_update(float* G, float* W, long long ldw, int K, long long D, float coef, float sigma,
                      unsigned long long seed, int bcast, void* stream);

/* RFA / Weiszfeld loop run entirely on the device.  Reference: helper.py:295-418. */
int dba_weiszfeld(const float* W, long long ldw, const float* G, int n, long long D, const float* alphas, int maxiter,
                  float eps, double ftol, float* median, double* dist, float* wts, float* wv, double* obj, int* state,
                  double* scratch, long long scratch_elems, void* stream);

/* FoolsGold: history update + fp64 Gram + cosine / pardoning / logit weights; then the weighted gradient reduce fused
 * with the server SGD step.  Reference: helper.py:527-607 and 259-293. */
int dba_foolsgold_weights(double* history, long long ldh, const int* ids, const float* feat, long long ldf, int n,
                          long long L, int use_memory, double* work_rows, double* gram, double* cs, double* wv,
                          double* alpha, void* stream);
/* Variant whose feature rows are given by (possibly peer-mapped) pointers: the gather of the clients' features from the
 * other GPUs happens inside the kernel's loads. */
int dba_foolsgold_weights_ptrs(double* history, long long ldh, const int* ids, const float* const* feat_rows, int n,
                               long long L, int use_memory, double* work_rows, double* gram, double* cs, double* wv,
                               double* alpha, void* stream);
int dba_foolsgold_apply(float* G, const float* S, long long lds, const double* wv, int K, long long Dp,
                        float eta_over_n, float lr, float decay, void* stream);

/* ---- tensor-core convolution (csrc/conv_tc.cu, csrc/wgrad_tc.cu) ----------------------------------------------- */

/* Rows of one planar-8 / shared-halo activation plane for B images of H x W. */
long long dba_conv_rows(int B, int H, int W);

/* Batched-client implicit-GEMM convolution (3x3 pad 1 or 1x1; stride via sub = 2) on tcgen05 / TMEM with fused
 * scale-shift (folded BatchNorm) + residual + ReLU epilogue, or raw output + BatchNorm statistics.
 * Reference: nn.Conv2d / BatchNorm2d / relu / residual in models/resnet_cifar.py:20-35,89. */
int dba_conv_tc(const void* X, const void* Wt, void* Y, const void* Res, const float* scale, const float* shift,
                float* stats, int K, int B, int H, int W, int Cin, int Cout, int ksize, int sub, int relu,
                int x_shared, int reserved, void* stream);
int dba_conv_tc_strided(const void* X, const void* Wt, long long w_client_stride, void* Y, const void* Res,
                        const float* scale, const float* shift, long long ss_stride, float* stats,
                        long long stats_stride, int K, int B, int H, int W, int Cin, int Cout, int ksize, int sub,
                        int relu, int x_shared, void* stream);
/* Same, with space-to-depth operands for stride-2 convolutions (models/resnet_cifar.py:20,27: stride=2 conv1 and
 * shortcut): in_s2d - X is [K][4*Cin/8][R'][8], parity plane p = 2*(h&1)+(w&1) of the real input on the (H, W) OUTPUT
 * grid, pass sub = 1; out_s2d - Y is written in that form ((H/2, W/2) grid, 4*Cout channels) for the next stage's
 * stride-2 consumers while Res stays plain.  Only computed positions are real outputs: no 4x full-grid waste. */
int dba_conv_tc_ex(const void* X, const void* Wt, long long w_client_stride, void* Y, const void* Res,
                   const float* scale, const float* shift, long long ss_stride, float* stats,
                   long long stats_stride, int K, int B, int H, int W, int Cin, int Cout, int ksize, int sub,
                   int relu, int x_shared, int in_s2d, int out_s2d, void* stream);

/* 'Valid' (unpadded) convolution as the centre crop of the padded one (MnistNet's 5x5 convolutions,
 * models/MnistNet.py:11-12): outputs are stored only for vlo <= h, w < vhi, at their position on the input grid.
 * shift without scale = bias epilogue.  ksize 5 needs (W+2) zero rows of slack before / after X. */
int dba_conv_tc_valid(const void* X, const void* Wt, long long w_client_stride, void* Y, const void* Res,
                      const float* scale, const float* shift, long long ss_stride, float* stats,
                      long long stats_stride, int K, int B, int H, int W, int Cin, int Cout, int ksize, int relu,
                      int x_shared, int vlo, int vhi, void* stream);

/* Convolution weight gradient (autograd of nn.Conv2d; image_train.py:91,209) added into the fp32 gradient slab. */
int dba_wgrad_tc(const void* dY, const void* X, float* G, long long ldg, int w_off, int K, int B, int H, int W,
                 int Cin, int Cout, int cin_real, int cout_real, int ksize, void* stream);

int dba_wgrad_tc_ex(const void* dY, const void* X, long long x_client_stride, float* G, long long ldg, int w_off, int K,
                    int B, int H, int W, int Cin, int Cout, int cin_real, int cout_real, int ksize, void* stream);
int dba_wgrad_tc_s2d(const void* dY, const void* X, float* G, long long ldg, int w_off, int K, int B, int H, int W,
                     int Cin, int Cout, int cin_real, int cout_real, void* stream);
/* Weight gradient of a linear layer over a flattened [c_real][npos] map kept as channels n = pos*c_pad + c of a 1x1
 * grid (MnistNet fc1 after x.view(-1, 4*4*50), models/MnistNet.py:22-23). */
int dba_wgrad_linear_chw(const void* dY, const void* X, float* G, long long ldg, int w_off, int K, int B, int Cout,
                         int cout_real, int c_pad, int c_real, int npos, void* stream);

/* ---- BatchNorm-free MnistNet path (csrc/mnist_ops.cu; models/MnistNet.py:17-30) ------------------------------- */
int dba_maxpool2_fwd(const void* x, void* y, int K, int C, int B, int Hi, int Wi, int off, int Ho, int Wo, int flatten,
                     long long Ri, long long Ro, void* stream);
int dba_maxpool2_bwd_relu_bias(const void* g, const void* x, void* dz, float* G, long long ldg, int b_off, int c_real,
                               int K, int C, int B, int Hi, int Wi, int off, int Ho, int Wo, int flatten, long long Ri,
                               long long Ro, void* stream);
int dba_relu_bwd_bias(const void* g, const void* act, void* dz, float* G, long long ldg, int b_off, int c_real, int K,
                      int C, int B, int H, int W, long long R, void* stream);
/* The same three with an optional bf16 hi/lo split copy of the output (tf32 mode: the weight-gradient operand form). */
int dba_maxpool2_fwd_split(const void* x, void* y, void* y_split, int K, int C, int B, int Hi, int Wi, int off, int Ho,
                           int Wo, int flatten, long long Ri, long long Ro, void* stream);
int dba_maxpool2_bwd_relu_bias_split(const void* g, const void* x, void* dz, void* dz_split, float* G, long long ldg,
                                     int b_off, int c_real, int K, int C, int B, int Hi, int Wi, int off, int Ho, int Wo,
                                     int flatten, long long Ri, long long Ro, void* stream);
int dba_relu_bwd_bias_split(const void* g, const void* act, void* dz, void* dz_split, float* G, long long ldg, int b_off,
                            int c_real, int K, int C, int B, int H, int W, long long R, void* stream);

/* ---- network plumbing (csrc/net_ops.cu, csrc/bn_ops.cu) -------------------------------------------------------- */

/* fp32 master weights ([Cout][taps][Cin]) -> packed bf16 operand layouts (forward + transposed). */
int dba_pack_weights(const float* W, long long ldw, const int* table, int n_layers, int K, void* dst,
                     long long dst_client_stride, void* stream);
/* Evaluation-mode BatchNorm folded to per-channel scale / shift. */
int dba_bn_fold(const float* W, long long ldw, const int* table, int n_layers, int K, float* scale, float* shift,
                long long ld_out, float eps, void* stream);
/* uint8 gather + /255 + trigger stamping + flip + stem im2col.  Reference: image_helper.py:289-350. */
int dba_input_stage_image(const void* data, const long long* labels, const long long* idx, long long ld_idx,
                          const void* trig, const long long* trig_row, const long long* poison_n, const void* flip,
                          void* out, int* labels_out, int K, int B, int Hi, int Wi, int C, int ks, int stride, int pad,
                          int Ho, int Wo, int KP, int label_swap, long long R, void* stream);
/* avg-pool + linear + cross-entropy + accuracy counters (+ backward).  Reference: resnet_cifar.py:94-96,
 * image_train.py:85,104-105, test.py:20-24. */
int dba_head(const void* act, const float* Wflat, long long ldw, int w_off, int b_off, float* Gflat, long long ldg,
             void* dact, const int* labels, int shared_labels, const int* nvalid, const long long* slot,
             float* metrics_f, int* metrics_i, int ld_m, unsigned char* nan_flag, int K, int B, int C, int H, int W,
             int ncls, long long R, void* stream);
/* c_pad: channel count of the planar act / dact buffers when the layer is padded (MnistNet fc1: 500 -> 512). */
int dba_head_ex(const void* act, const float* Wflat, long long ldw, int w_off, int b_off, float* Gflat, long long ldg,
             void* dact, const int* labels, int shared_labels, const int* nvalid, const long long* slot,
             float* metrics_f, int* metrics_i, int ld_m, unsigned char* nan_flag, int K, int B, int C, int c_pad, int H, int W,
             int ncls, long long R, void* stream);
/* Training-mode BatchNorm forward (apply) and backward.  Reference: nn.BatchNorm2d in resnet_cifar.py:20-35. */
int dba_bn_apply(const void* y, const void* res, void* out, float* stats, long long ld_stats, int stats_off,
                 float* saved, long long ld_saved, int saved_off, float* Wflat, long long ldw, int g_off, int b_off,
                 int rm_off, int rv_off, const int* nvalid, const unsigned char* active, int K, int C, int B, int H,
                 int W, long long R, int relu, float momentum, float eps, void* stream);
/* Same; out_s2d (nullable) receives a second copy of the activation in space-to-depth form for the next stage's
 * stride-2 convolutions (see dba_conv_tc_ex). */
int dba_bn_apply_ex(const void* y, const void* res, void* out, void* out_s2d, long long R_s2d, float* stats,
                    long long ld_stats, int stats_off, float* saved, long long ld_saved, int saved_off, float* Wflat,
                    long long ldw, int g_off, int b_off, int rm_off, int rv_off, const int* nvalid,
                    const unsigned char* active, int K, int C, int B, int H, int W, long long R, int relu,
                    float momentum, float eps, void* stream);
int dba_bn_backward(const void* g, const void* y, const void* act, void* dy, void* gmask, float* sums,
                    long long ld_sums, int sums_off, float* saved, long long ld_saved, int saved_off, float* Wflat,
                    long long ldw, float* Gflat, long long ldg, int g_off, int b_off, const int* nvalid, int K, int C,
                    int B, int H, int W, long long R, int up, long long Rup, void* stream);

/* tf32 mode only: the same three kernels also storing bf16 hi/lo "split" copies ([K][2][C/8][R][8]: hi = bf16(x),
 * lo = bf16(x - hi)) of what they write - the operand form of the weight-gradient kernel, which then issues
 * hi*hi + hi*lo + lo*hi (16 mantissa bits; the tensor core takes MN-major 32-bit operands only in a swizzled layout
 * that planar-4 rows cannot be loaded into).  Null split pointers = no copy. */
int dba_bn_apply_split(const void* y, const void* res, void* out, void* out_s2d, long long R_s2d, void* out_split,
                       void* out_s2d_split, float* stats, long long ld_stats, int stats_off, float* saved,
                       long long ld_saved, int saved_off, float* Wflat, long long ldw, int g_off, int b_off, int rm_off,
                       int rv_off, const int* nvalid, const unsigned char* active, int K, int C, int B, int H, int W,
                       long long R, int relu, float momentum, float eps, void* stream);
int dba_bn_backward_split(const void* g, const void* y, const void* act, void* dy, void* dy_split, void* gmask,
                          float* sums, long long ld_sums, int sums_off, float* saved, long long ld_saved, int saved_off,
                          float* Wflat, long long ldw, float* Gflat, long long ldg, int g_off, int b_off,
                          const int* nvalid, int K, int C, int B, int H, int W, long long R, int up, long long Rup,
                          void* stream);
/* Same with sums_ready: non-zero = the reduction sums were accumulated by the convolution that produced g
 * (dba_conv_tc_dgrad_bn), only the apply kernel runs. */
int dba_bn_backward_ex(const void* g, const void* y, const void* act, void* dy, void* dy_split, void* gmask,
                       float* sums, long long ld_sums, int sums_off, float* saved, long long ld_saved, int saved_off,
                       float* Wflat, long long ldw, float* Gflat, long long ldg, int g_off, int b_off,
                       const int* nvalid, int K, int C, int B, int H, int W, long long R, int up, long long Rup,
                       int sums_ready, void* stream);
/* Input-gradient convolution G = conv(dY, Wt) (+ Res) whose epilogue also accumulates the BatchNorm-backward sums of
 * the layer consuming G (sum g', sum g' * xhat with g' = bn_act > 0 ? g : 0) into sums[k][2c], sums[k][2c + 1]; saved =
 * that layer's (mean, rstd) pairs.  Pointers sums / saved are already offset to the layer.  Reference counterpart:
 * cuDNN's BatchNorm backward inside autograd (image_train.py:91,209). */
int dba_conv_tc_dgrad_bn(const void* dY, const void* Wt, long long w_client_stride, void* G, const void* Res,
                         const void* bn_y, const void* bn_act, const float* saved, long long ld_saved, float* sums,
                         long long ld_sums, int K, int B, int H, int W, int Cin, int Cout, int ksize, void* stream);
int dba_maxpool_fwd_split(const void* x, void* y, void* y_split, int K, int C, int B, int H, int W, long long Rin,
                          long long Rout, void* stream);
int dba_input_stage_image_split(const void* data, const long long* labels, const long long* idx, long long ld_idx,
                                const void* trig, const long long* trig_row, const long long* poison_n, const void* flip,
                                void* out, void* out_split, int* labels_out, int K, int B, int Hi, int Wi, int C, int ks,
                                int stride, int pad, int Ho, int Wo, int KP, int label_swap, long long R, void* stream);

/* Activation element mode of all network kernels above (library-wide): 0 = bf16 activations, 8 channels per 16-byte
 * chunk, tcgen05 kind::f16; 1 = fp32 activations, 4 channels per chunk, kind::tf32 - the reference's precision class
 * (fp32 tensors with TF32 convolutions, cuDNN's default).  No counterpart in the reference (it has no native code). */
int dba_set_act_dtype(int mode);
/* Programmatic dependent launch between consecutive network kernels (default on): a kernel's blocks are scheduled, and
 * run their prologue, while its predecessor drains.  No counterpart in the reference. */
/* Lock-step staging: one launch copies the schedule entries of `step` (sample indices [K][T][B], flip flags, valid
 * count, active flag, log slot, learning rate [K][T]) of clients rows[0..n) (null = 0..n-1) into the step slab's
 * st_* rows.  Reference counterpart: the per-batch DataLoader / scheduler bookkeeping of image_train.py:191-230. */
int dba_stage_step(const long long* idx, const unsigned char* flip, const int* nvalid, const unsigned char* active,
                   const long long* slot, const float* lr, const long long* rows, int n, int T, int B, int step,
                   long long* st_idx, unsigned char* st_flip, int* st_nvalid, unsigned char* st_active,
                   long long* st_slot, float* st_lr, void* stream);
void dba_set_pdl(int on);
int dba_get_pdl(void);
/* BatchNorm backward as one thread-block-cluster launch per layer (1; partial sums exchanged through distributed
 * shared memory, deterministic, `sums` scratch unused) or as reduce + apply kernels (0, default - measured faster;
 * `sums` must be zeroed by the caller). */
void dba_set_wgrad_vstack(int on);   /* 3x3 weight gradients with Cout <= 32: three vertical taps stacked along M (default on) */
void dba_set_bn_pair(int on);        /* tf32 mode: BatchNorm apply handles 8-channel planes per thread (default on) */
void dba_set_bn_fused(int on);
int dba_get_bn_fused(void);
int dba_get_act_dtype(void);

/* MaxPool2d(3, 2, 1) forward / backward on planar-8 activations (Tiny-ImageNet stem, resnet_tinyimagenet.py:147). */
int dba_maxpool_fwd(const void* x, void* y, int K, int C, int B, int H, int W, long long Rin, long long Rout,
                    void* stream);
int dba_maxpool_bwd(const void* x, const void* gy, void* gx, int K, int C, int B, int H, int W, long long Rin,
                    long long Rout, void* stream);

/* ---- LOAN MLP (csrc/loan_mlp.cu) ------------------------------------------------------------------------------- */

/* Whole local training of K LOAN clients, one persistent CTA per client (gather + feature trigger + forward with
 * fused dropout + CE + backward + SGD for all T steps).  Reference: models/loan_model.py:10-27, loan_train.py:11-261. */
int dba_loan_train(float* W, long long ldw, float* gsum, long long ldg, const float* X, const long long* Y,
                   const long long* idx, const float* lr, const unsigned char* active, const long long* epoch_of,
                   const long long* poison_n, const long long* trig_row, const unsigned char* trig_mask,
                   const float* trig_val, float* loss_sum, int* correct, int* count, int* pcount,
                   unsigned char* nan_flag, int K, int T, int B, int E, int label_swap, float mu, float wd,
                   unsigned int seed, int dropout, void* stream);
/* M LOAN models on the rows idx[0..n) (optionally with one feature trigger applied).  Reference: test.py:7-239. */
int dba_loan_eval(const float* W, long long ldw, int M, const float* X, const long long* Y, const long long* idx,
                  long long n, const unsigned char* trig_mask, const float* trig_val, int label_swap, double* loss,
                  long long* correct, void* stream);

/* ---- peer-memory fused collectives (csrc/p2p_collectives.cu) ---------------------------------------------------- */

/* One communicator context per process (one process per GPU).  Registers the symmetric buffers (host arrays of `world`
 * peer-mapped device pointers: stage / replica / aux [D] floats, signal 64 uint32, mail 128 doubles, table 2*table_cap
 * floats; *_mc = multicast addresses or null; counters 8 uint32, err 1 int, table_dev `world` pointers: local). */
int dba_peer_init(int rank, int world, const void* const* stage, const void* const* replica, const void* const* aux,
                  const void* const* signal, const void* const* mail, const void* const* table, int table_cap,
                  void* stage_mc, void* replica_mc, void* aux_mc, unsigned int* counters, int* err, void* table_dev);
/* 0 = fine, 1 = a cross-GPU barrier timed out on this GPU (a peer died or never launched the matching kernel). */
int dba_peer_error(void);
/* F1, FedAvg over all ranks in ONE cooperative launch per rank: local K-way reduce of (W[k] - G) -> barrier -> owner
 * reduces its slice over NVLink peer loads (or multimem.ld_reduce when nvls != 0) -> update (+DP noise) -> peer /
 * multicast stores into every replica -> barrier.  Reference: helper.py:193-257 + the copy image_train.py:26-32. */
int dba_peer_fedavg(const float* W, long long ldw, int K, long long D, float coef, float sigma,
                    unsigned long long seed, int nvls, void* stream);
/* target (0 = replica, 1 = aux) = beta*target + alpha * sum over all ranks of (K > 0 ? sum_k wv[k]*S[k] : staging
 * vector).  F3b: FoolsGold's weighted gradient reduce + server step (helper.py:561-570,280-290) with the local
 * weighted reduce as phase 1 of the launch; K = 0: plain fused all-reduce of the staging vector. */
int dba_peer_axpby(const float* S, long long lds, int K, const double* wv, long long D, float alpha, float beta,
                   int target, int nvls, void* stream);
/* F2, RFA / Weiszfeld over the deltas (W[k] - G) of all ranks (helper.py:295-418): one launch per iteration per rank
 * (local weights + weighted partial sum, scalars in the ticket mailbox, normalised median into every rank's aux
 * buffer, local distances, device-side convergence test), maxiter + 2 launches, no host synchronisation, no NCCL. */
int dba_peer_weiszfeld(const float* W, long long ldw, int K, long long D, const float* G, const float* alphas,
                       int maxiter, float eps, double ftol, int slot0, int n_total, double* dist2, int kmax, float* wts,
                       float* wv, double* obj, int* state, void* stream);
/* Stream-ordered barrier across all ranks. */
int dba_peer_barrier(void* stream);
/* out[j][:] = *rows[j] (rows: device array of local or peer-mapped pointers): NVLink gather into local memory. */
int dba_gather_rows_ptrs(const float* const* rows, int n, long long D, float* out, long long ldo, void* stream);
/* Gram matrix over feature rows given by (possibly peer-mapped) pointers: FoolsGold's gather fused into the cosine
 * GEMM (helper.py:544-555,574-577). */
int dba_gram_ptrs(const double* const* rows, int n, long long L, double* gram, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* DBA_B200_H */
