/* b200ddp.h — C ABI of libb200ddp.so: the B200 (sm_100a) kernels behind the DDP training step of
 * annalena-k/tutorial-torch-distributed-data-parallel.
 *
 * The reference has no FFI of its own: its hot path is Python calling torch / torchvision / NCCL
 * (SURVEY.md 8(b)). Each entry point below therefore cites the reference call site it serves and the
 * upstream (torch 2.11.0, torchvision 0.26.0, NCCL 2.28.9) routine it replaces. INTEGRATION.md shows
 * the ctypes binding the Python layer uses.
 *
 * Conventions
 *   - plain C types only; every pointer is a BORROWED device pointer unless stated otherwise;
 *   - `device` is the CUDA ordinal the buffers live on, `stream` a cudaStream_t passed as void*;
 *   - all work is stream ordered and asynchronous; nothing synchronises the device;
 *   - return value 0 = success; otherwise b200ddp_last_error() (thread local) describes the failure;
 *     1 = invalid argument, 2 = not initialised, 1000+e = cudaError_t e, 2000+r = ncclResult_t r,
 *     3000+r = CUresult r;
 *   - the library never allocates or frees device memory: scratch / workspace buffers are provided
 *     by the caller, their sizes come from the *_bytes() queries;
 *   - entry points may be called from any host thread (autograd runs hooks on worker threads).
 */
#ifndef B200DDP_H_
#define B200DDP_H_

#ifdef __cplusplus
extern "C" {
#endif

/* ---- library ---------------------------------------------------------------------------------- */
const char* b200ddp_last_error(void);
int b200ddp_version(void);
/* Number of CUDA kernels this library has launched in this process (evidence for bench `gpu_launches`). */
unsigned long long b200ddp_launch_count(void);
int b200ddp_device_info(int device, int* sm_count, int* cc_major, int* cc_minor);
/* Number of SMs the grid planners (split counts of the wgrad and GEMM kernels) may assume are free; 0 or > 148 = all.
 * Returns the previous value. The DDP wrapper lowers it by the number of NCCL channels: an all-reduce that is
 * resident on 8 SMs turns a 148-CTA one-wave grid into two waves (measured: 2-GPU step 4.2 ms instead of 2.9 ms). */
int b200ddp_set_sm_budget(int sms);

/* ---- O1: optimizer ---------------------------------------------------------------------------- */
/* Replaces torch.optim.Adam._multi_tensor_adam (torch/optim/adam.py:554-800), reached from
 * reference multi-GPU-training-torch.py:126 `optimizer.step()` and :249 `optim.Adam(..., lr=0.001)`.
 * One pass over flat arenas: p,g,m,v of n floats (16-byte aligned). `step` is the 1-based step count
 * AFTER increment. grad_scale_dev (nullable): device scalar multiplied into g first. */
int b200ddp_adam_step_flat(float* p, const float* g, float* m, float* v, long long n, double lr, double beta1,
                           double beta2, double eps, double weight_decay, long long step,
                           const float* grad_scale_dev, int device, void* stream);

/* ---- R1: bucket plan (host only) ---------------------------------------------------------------- */
/* Replaces c10d::compute_bucket_assignment_by_size (torch reducer.hpp:592-598; called from
 * torch/nn/parallel/distributed.py:1237-1244 when reference :245 constructs DDP). nbytes[i] = size of
 * tensor i in ready order, keys[i] (nullable) = dtype/device key, caps = byte caps (e.g. {1 MiB, 25 MiB}).
 * Writes bucket_of[i] and *nbuckets. */
int b200ddp_plan_buckets(const long long* nbytes, const int* keys, int n, const long long* caps, int ncaps,
                         int* bucket_of, int* nbuckets);

/* ---- R2/R3: pack, scale, clip before aggregation ------------------------------------------------ */
/* dst = clamp(src, +-clip_value) * scale (clip_value <= 0: no clamp); dst may alias src.
 * Replaces Reducer::mark_variable_ready_dense's grad*1/ws copy (reducer.hpp:276) and
 * torch.nn.utils.clip_grad_value_ (clip_grad.py:291-299) placed before the all-reduce (README.md:92-95). */
int b200ddp_pack_scale_clip(float* dst, const float* src, long long n, float scale, float clip_value,
                            int nan_to_zero, int device, void* stream);
/* g *= *coef_dev; no memory traffic when *coef_dev == 1; exact_zero: *coef_dev == 0 stores zeros. */
int b200ddp_scale_by_device_scalar(float* g, long long n, const float* coef_dev, int exact_zero, int device,
                                   void* stream);
/* out_sq[0] (+)= sum(g^2). Local part of clip_grad_norm_ (clip_grad.py:50-118). scratch: 256-byte
 * aligned, >= b200ddp_sqnorm_scratch_bytes() bytes, its first 4 bytes zero before the first call. */
long long b200ddp_sqnorm_scratch_bytes(void);
int b200ddp_sqnorm(const float* g, long long n, float* out_sq, int accumulate, void* scratch,
                   long long scratch_bytes, int device, void* stream);
/* coef = min(1, max_norm / (sqrt(*sq_dev) + 1e-6)) (clip_grad.py:165-169); nonfinite_policy 0 = stock
 * (NaN propagates), 1 = coef 0 (this rank contributes nothing). norm_out_dev nullable. */
int b200ddp_clip_coef(const float* sq_dev, float max_norm, int nonfinite_policy, float* coef_out_dev,
                      float* norm_out_dev, int device, void* stream);

/* ---- R4: collectives (NCCL over NVLink 5 / NVSwitch) ------------------------------------------- */
/* Replaces ProcessGroupNCCL::allreduce / allgather / broadcast as used by c10d::Reducer
 * (reducer.hpp:276-286), SyncBatchNorm (_functions.py:74-76,165) and DDP's parameter broadcast
 * (distributed.py:883-890). The comm handle owns an ncclComm_t, a high-priority communication stream
 * and the events that order it against the caller's compute stream. */
int b200ddp_nccl_load(const char* libnccl_path_or_null);
int b200ddp_nccl_version(int* version);
int b200ddp_comm_unique_id(void* out, int nbytes /* >= 128 */);
int b200ddp_comm_create(void** handle_out, const void* unique_id, int rank, int nranks, int device);
int b200ddp_comm_destroy(void* handle);
int b200ddp_comm_rank(void* handle, int* rank, int* nranks);
/* In place sum over ranks of (premul * buf) on the caller's stream. */
int b200ddp_comm_allreduce(void* handle, float* buf, long long n, float premul, void* stream);
int b200ddp_comm_allgather(void* handle, const float* send, float* recv, long long n_per_rank, void* stream);
int b200ddp_comm_broadcast(void* handle, void* buf, long long nbytes, int root, void* stream);
/* Bucket all-reduce overlapped with backward: valid-at = current tail of compute_stream; runs on the
 * handle's communication stream; join makes compute_stream wait for every launched bucket. */
int b200ddp_bucket_reduce_launch(void* handle, float* buf, long long n, float premul, void* compute_stream);
int b200ddp_bucket_reduce_join(void* handle, void* compute_stream);
/* `stream` waits for the bucket collective launched last (a per-bucket optimizer launch follows its bucket). */
int b200ddp_bucket_wait_last(void* handle, void* stream);
/* arena_register: the gradient arena as NCCL-registered, peer-mapped memory (replaces the unregistered bucket tensors
 * of c10d::Reducer, reducer.hpp:330-334). COLLECTIVE: every rank calls it with the same size at the same point.
 * The buffer comes from ncclMemAlloc (owned by the communicator, freed by comm_destroy, zero-filled) and is registered
 * as an NCCL symmetric window: *mode_out = 2 peer + NVSwitch multicast mappings, 1 peer mappings only, 0 plain
 * ncclCommRegister (NCCL collectives only). Set-up call: synchronises the device. */
int b200ddp_comm_arena_alloc(void* handle, long long nbytes, void** ptr_out, int* mode_out);
/* Which all-reduce bucket_reduce_launch runs: 0 NCCL ring with PreMulSum(1/world) (any buffer), 1 this library's own
 * kernel over the arena's peer mappings (multimem.ld_reduce / multimem.st through the NVSwitch, or peer loads / stores;
 * `ctas` CTAs of 256 threads, <= 256; buffers outside an arena fall back to 0), 2 NCCL ncclSum on pre-scaled data.
 * Returns 1 if the peer path is available, 0 if not, < 0 on a bad argument. Same value on every rank. */
int b200ddp_comm_set_algo(void* handle, int algo, int ctas);
/* CTA shape of that kernel: 0 = 256 threads x 32 registers (fits beside a running convolution CTA), 1 = 1024 threads
 * (owns its SM; leave `ctas` SMs out of the compute grids with b200ddp_set_sm_budget). */
int b200ddp_comm_set_cta_shape(void* handle, int big);
/* B1: the two tiny collectives of SyncBatchNorm (torch _functions.py:50-75 all_gather of [mean, invstd, count], :160-170
 * all_reduce of [sum_dy, sum_dy_xmu]) as ONE single-CTA kernel over the communicator's peer / multicast mappings: every
 * rank deposits its n <= 2048 floats in every rank's scratch, one flag round, then sum_mode 0: out [nranks][n] (all-gather),
 * 1: out [n] = scale * sum in rank order (bit-identical everywhere). Caller's stream. Returns 2 (not initialised) when the
 * peer path is unavailable: use comm_allgather / comm_allreduce then. COLLECTIVE on first use (set-up). */
int b200ddp_comm_small_collective(void* handle, const float* send, float* out, int n, int sum_mode, float scale, void* stream);
/* Device time of each bucket collective of the last joined backward pass (for the NVLink roofline: busBW =
 * 4 n * 2 (ws-1)/ws / t). set_timing(1) first; synchronise before reading. Returns the number of entries. */
int b200ddp_comm_set_timing(void* handle, int on);
int b200ddp_comm_bucket_times(void* handle, float* ms_out, long long* n_out, int max_entries);

/* ---- F1/F2/F3: Linear layers --------------------------------------------------------------------- */
/* Replace nn.Linear forward/backward (cuBLAS sgemm) + nn.ReLU + nn.Dropout of torchvision
 * alexnet.py:38-43 and the head data_and_toy_model.py:44, reached from reference :121 / :125.
 * Layouts: x [B][K], w [N][K] (torch Linear weight), y / dz [B][N], row-major fp32.
 * mode 0 = strict fp32 SIMT, mode 1 = TF32 tcgen05 tensor cores. N <= 16 uses the head kernels.
 * Epilogue of fwd: y = dropout(relu(x w^T + bias)); gate[b][n] (nullable) = 0 off, 1 on, 2 on*1/(1-p).
 * ws: workspace of b200ddp_linear_workspace_bytes(B,N,K) bytes for the split-K partials. */
long long b200ddp_linear_workspace_bytes(int B, int N, int K);
int b200ddp_linear_fwd(const float* x, const float* w, const float* bias, float* y, unsigned char* gate, int B, int N,
                       int K, int relu, float dropout_p, unsigned long long seed, int mode, void* ws,
                       long long ws_bytes, int device, void* stream);
int b200ddp_linear_bwd_dx(const float* dz, const float* w, float* dx, int B, int N, int K, int mode, void* ws,
                          long long ws_bytes, int device, void* stream);
/* dw [N][K] = dz^T x (written, not accumulated); db [N] (nullable) = column sums of dz. */
int b200ddp_linear_bwd_dw(const float* dz, const float* x, float* dw, float* db, int B, int N, int K, int mode,
                          int device, void* stream);
/* dz = dy * gate (0 / 1 / 1/(1-p)); db[n] (nullable) = sum_b dz[b][n]; dz nullable. */
int b200ddp_gate_bias_bwd(const float* dy, const unsigned char* gate, float dropout_p, float* dz, float* db, int B,
                          int N, int device, void* stream);

/* ---- F3: loss ---------------------------------------------------------------------------------------- */
/* Replaces nn.CrossEntropyLoss() (reference :248,:122): loss[0] = mean CE over the rows whose label is not
 * ignore_index (-100, torch's default), dlogits (nullable) = (softmax - onehot)/n_valid, zero rows for ignored
 * labels. labels int64. A label outside [0, C) other than -100 sets err_flag_dev (nullable; may be mapped host
 * memory) to 1 and is treated as ignored: its gradient row is zero, never uninitialised. */
int b200ddp_softmax_ce(const float* logits, const long long* labels, float* loss, float* dlogits, int B, int C,
                       int* err_flag_dev, int device, void* stream);

/* ---- T1: native step driver ------------------------------------------------------------------------------- */
/* Replaces the layer-by-layer Python / autograd walk of `outputs = model(inputs)` and `loss.backward()` in reference
 * train() (multi-GPU-training-torch.py:121,125) over torchvision AlexNet (alexnet.py:17-52, head of
 * data_and_toy_model.py:44) by ONE call each, which enqueue the same kernels as the per-layer entry points above, in
 * the same order with the same arguments (bit-identical results), out of one step-persistent workspace.
 * Standard geometry only: x [B][3][224][224]; TF32 kernels. w / b: conv1..conv5, fc1, fc2, head (torch layouts).
 * conv1_ktab: the k_order-0 offset table of b200ddp_conv2d_fwd for conv1 (Cin 3, 224x224, 11x11). */
typedef struct b200ddp_alexnet_desc {
  int B, hidden, num_classes;
  float dropout_p;
  const float* w[8];
  const float* b[8];
  const int* conv1_ktab;
  int conv1_order;             /* k_order of the first layer's forward: 0 (offset table) or 2 (slab kernel, conv1.cu) */
  void* workspace;             /* b200ddp_alexnet_workspace_bytes(B, hidden, num_classes) bytes, 256-byte aligned */
  long long workspace_bytes;
  int device;
} b200ddp_alexnet_desc;
typedef void (*b200ddp_ready_fn)(void* user, int group);
long long b200ddp_alexnet_workspace_bytes(int B, int hidden, int num_classes);
int b200ddp_alexnet_plan_create(void** plan_out, const b200ddp_alexnet_desc* desc);
int b200ddp_alexnet_plan_destroy(void* plan);
/* train != 0 keeps gates / argmax codes for backward and applies dropout (seed0: classifier.0, seed1: classifier.1);
 * out = logits [B][classes], or the head's input [B][hidden] when stop_at_hidden != 0 (evaluate()). */
int b200ddp_alexnet_forward(void* plan, const float* x, float* out, int train, unsigned long long seed0,
                            unsigned long long seed1, int stop_at_hidden, void* stream);
/* Backward of the last training forward (x = its input). dw[i] / db[i] are WRITTEN. ready(user, g) is called on the
 * host once group g's gradients are enqueued on `stream` (0 head+fc2, 1 fc1, 2 conv5..conv2, 3 conv1): the DDP
 * reducer launches its bucket all-reduce from there. use_side_stream: weight gradients of conv5..conv2 on the plan's
 * own stream (joined before group 2 is reported). */
int b200ddp_alexnet_backward(void* plan, const float* x, const float* dlogits, float* const* dw, float* const* db,
                             float dropout_p_used, b200ddp_ready_fn ready, void* user, int use_side_stream, void* stream);

/* ---- f1: evaluate() ------------------------------------------------------------------------------------- */
/* Replaces, per test batch of reference evaluate() (multi-GPU-training-torch.py:136-153, -accelerate.py:60-75), the
 * head Linear (classifier.6) + criterion + loss.item() + torch.max(outputs, 1) + (predicted == labels).sum().item():
 * metrics[0] += sum_i CE_i, metrics[1] += #(argmax_i == y_i), metrics[2] += B — on the device, no host sync.
 * h [B][K] (input of the head), w [N][K], bias [N] or null, N <= 16, labels int64; logits_out [B][N] nullable.
 * scratch: b200ddp_eval_scratch_bytes(B) bytes, 256-byte aligned, first 4 bytes zero before the first call
 * (the kernel resets them). Deterministic (row results summed in row order by the last CTA). */
long long b200ddp_eval_scratch_bytes(int B);
int b200ddp_head_eval(const float* h, const float* w, const float* bias, const long long* labels, float* logits_out, int B,
                      int N, int K, float* metrics, void* scratch, long long scratch_bytes, int device, void* stream);
/* Same three sums from logits [B][C] that already exist (any model). */
int b200ddp_logits_eval(const float* logits, const long long* labels, int B, int C, float* metrics, void* scratch,
                        long long scratch_bytes, int device, void* stream);

/* ---- f2: input transform --------------------------------------------------------------------------------- */
/* Replaces transforms.Resize -> RandomHorizontalFlip -> ToTensor -> Normalize of reference data_and_toy_model.py:13-36
 * (PIL on DataLoader workers) and the 77 MB blocking H2D of multi-GPU-training-torch.py:110: the batch arrives as the
 * dataset stores it, img uint8 [B][H][W][3], and out fp32 [B][3][OH][OW] is bit-identical to the torchvision result
 * (Pillow's two-pass 8-bit bilinear resample, then x/255, then (x - mean)/std). flip [B] bytes (nullable): 1 = mirror
 * the resized image. tab_h / tab_v: DEVICE tables from b200ddp_resize_coeffs for (W -> OW) and (H -> OH);
 * mean3 / std3: HOST arrays of 3 floats. Enlargements only (H <= OH <= 512, H, W <= 64). */
int b200ddp_resize_coeffs(int in_size, int out_size, int* table_host /* 5 * out_size ints: first index | taps | 3 coefs */);
int b200ddp_image_transform(const unsigned char* img, const unsigned char* flip, float* out, int B, int H, int W, int OH,
                            int OW, const int* tab_h, const int* tab_v, const float* mean3, const float* std3, int device,
                            void* stream);

/* ---- C1-C5: convolutions ------------------------------------------------------------------------------ */
/* Replace nn.Conv2d forward / backward (cuDNN TF32 implicit GEMM + NCHW<->NHWC conversions) of torchvision
 * alexnet.py:22,25,28,30,32, reached from reference :121 / :125. x [Nb][Cin][H][W], y [Nb][Cout][OH][OW], fp32 NCHW.
 * wmat: weight matrix [Cout][ldw] (torch's [Cout][Cin][KH][KW] flattened; ldw % 4 == 0, use weight_prep mode 0 to
 * pad). ktab: device table of int32 pairs, one per k = (ci,kh,kw), padded to a multiple of 384 pairs:
 * {ci*H*W + kh*W + kw, 1 << kh | 1 << (16 + kw)} and {0, 1 << 31} beyond K (k_order 0; KH, KW <= 15). k_order 1 = reduction index
 * ordered channel-block-major (ci/32, kh, kw, ci%32) — needs Cin % 32 == 0, no table, weights permuted by weight_prep
 * mode 2 / 3; dw is always
 * returned in torch's [Cout][Cin][KH][KW] order. TF32 tensor cores (gathered operands rounded), fp32 accumulate.
 * dgrad of a stride-1 convolution = conv2d_fwd on dy with weight_prep mode 1 weights and pad' = KH-1-pad. */
int b200ddp_conv2d_fwd(const float* x, const float* wmat, long long ldw, const float* bias, float* y, const int* ktab,
                       int k_order, int Nb, int Cin, int H, int W, int Cout, int KH, int KW, int stride, int pad, int relu,
                       int device, void* stream);
/* dw [Cout][Cin*KH*KW] = sum over pixels of dy x im2col(x) (written, not accumulated). k_order 2 = the first layer's own
 * kernel (conv1.cu, same geometry limits as the forward); the workspace query covers it. */
long long b200ddp_conv2d_wgrad_workspace_bytes(int Nb, int Cin, int H, int W, int Cout, int KH, int KW, int stride,
                                               int pad);
int b200ddp_conv2d_wgrad(const float* x, const float* dy, float* dw, const int* ktab, int k_order, int Nb, int Cin, int H,
                         int W, int Cout, int KH, int KW, int stride, int pad, void* ws, long long ws_bytes, int device,
                         void* stream);
/* Strict mode of the same three operations ("rtol 1e-5 strict" against torch with
 * torch.backends.cudnn.allow_tf32 = False): direct fp32 SIMT convolutions, no TF32 rounding, fp64 across partial rows,
 * deterministic. w / dw in torch's [Cout][Cin][KH][KW] layout, any stride / padding. Exactness path, not a fast one. */
int b200ddp_conv2d_fwd_fp32(const float* x, const float* w, const float* bias, float* y, int Nb, int Cin, int H, int W,
                            int Cout, int KH, int KW, int stride, int pad, int relu, int device, void* stream);
int b200ddp_conv2d_dgrad_fp32(const float* dy, const float* w, float* dx, int Nb, int Cin, int H, int W, int Cout, int KH,
                              int KW, int stride, int pad, int device, void* stream);
int b200ddp_conv2d_wgrad_fp32(const float* x, const float* dy, float* dw, int Nb, int Cin, int H, int W, int Cout, int KH,
                              int KW, int stride, int pad, int device, void* stream);
/* k_order 2 of b200ddp_conv2d_fwd = the first layer's own kernel (conv1.cu; Cin 3, 11x11, stride 4, Cout 64, W + 2 pad <=
 * 256): wmat from weight_prep mode 5, out[co][ci*136 + kh*12 + kw] = w[co][ci][kh][kw], zero elsewhere, ld_out >= 416. */
/* mode 0: out[co][0..ld_out) = w[co][:] zero padded; mode 1: flipped / transposed weights for dgrad,
 * out[ci][co*KH*KW + (KH-1-kh)*KW + (KW-1-kw)] = w[co][ci][kh][kw], row stride ld_out; modes 2 / 3: the same two
 * matrices with the reduction index in channel-block-major order (k_order 1). */
int b200ddp_conv2d_weight_prep(const float* w, float* out, int Cout, int Cin, int KH, int KW, long long ld_out, int mode,
                               int device, void* stream);
/* mode 4 | S << 8: forward weights for the space-to-depth form of a stride-S convolution (Cin*S*S <= 64): the
 * equivalent stride-1 kernel with ceil(KH/S) x ceil(KW/S) taps over 64 channels c' = (ci*S + a)*S + b,
 * w'[co][c'][q][r] = w[co][ci][S*q + a][S*r + b], channel-block-major; ld_out >= 64*ceil(KH/S)*ceil(KW/S). */
/* out[n][(c*S + a)*S + b][i][j] = x[n][c][S*i + a - pad][S*j + b - pad] (0 elsewhere; out: [Nb][Cp][Hs][Ws], Cp >=
 * C*S*S, zeroed once by the caller - positions no input maps to are never written). Input transform of the first
 * layer: the stride-4 11x11 convolution over 3 channels (alexnet.py:22) becomes a stride-1 3x3 over 48(64). */
int b200ddp_space_to_depth(const float* x, float* out, int Nb, int C, int H, int W, int S, int pad, int Cp, int Hs, int Ws,
                           int device, void* stream);
/* Several weight_prep jobs in one launch (at most 16): what a training step needs for all conv layers at once. */
typedef struct b200ddp_weight_prep_job {
  const float* w;      /* [Cout][Cin][KH][KW] */
  float* out;
  int Cout, Cin, KH, KW;
  long long ld_out;
  int mode;            /* as b200ddp_conv2d_weight_prep */
} b200ddp_weight_prep_job;
int b200ddp_conv2d_weight_prep_multi(const b200ddp_weight_prep_job* jobs, int njobs, int device, void* stream);
/* out[c] = sum over (n, pix) dy[n][c][pix] — the conv bias gradient. scratch: 256-byte aligned, zeroed once. */
long long b200ddp_channel_sum_scratch_bytes(int Nb, int C);
int b200ddp_channel_sum(const float* dy, float* out, int Nb, int C, int HW, void* scratch, long long scratch_bytes,
                        int device, void* stream);

/* Same result as b200ddp_channel_sum, for callers that cannot hand in scratch memory (no allocator within reach,
 * or cudaMalloc forbidden on their path). No counterpart in the reference: this exists only because of the
 * library's own "caller provides scratch" rule. The partial sums live in a 258 KB area built into the library
 * (held by every GPU that loads it, until exit). Limits: C <= 512 channels, otherwise an error is returned; all
 * calls on a device share that area, so calls that overlap on the GPU race — keep them on one stream. The first
 * call on each device also resolves the area's device address; make that call before any stream capture begins.
 * Allocates nothing (no cudaMalloc / cudaFree, no allocating library call). */
int b200ddp_channel_sum_static(const float* dy, float* out, int Nb, int C, int HW, int device, void* stream);

/* ---- C6: activations / pooling / dropout ------------------------------------------------------------ */
/* Replace nn.ReLU + nn.MaxPool2d(3,2) (alexnet.py:23-24,26-27,33-34), nn.ReLU (:29,:31) and
 * nn.Dropout (:38,:41). NCHW fp32. idx: one byte per output, window slot 0..8, 255 = ReLU closed. */
int b200ddp_relu_maxpool_fwd(const float* x, float* y, unsigned char* idx, int B, int C, int H, int W, int relu,
                             int device, void* stream);
int b200ddp_relu_maxpool_bwd(const float* dy, const unsigned char* idx, float* dx, int B, int C, int H, int W,
                             int device, void* stream);
/* The same two backward passes, additionally producing the bias gradient of the convolution whose output gradient
 * they compute (torch: aten::convolution_backward's grad_bias = dY.sum((0,2,3)), reached from reference :125):
 * db[c] = sum over batch and plane of dx. Per-plane sums in fp64 scratch, folded over the batch in a fixed order by
 * a second small kernel (deterministic). scratch: b200ddp_bias_grad_scratch_bytes(B, C) bytes, 8-byte aligned, need
 * not be initialised. relu_bwd_bias: dx = dy * (y > 0) on [B][C][HW]. */
long long b200ddp_bias_grad_scratch_bytes(int B, int C);
int b200ddp_relu_maxpool_bwd_bias(const float* dy, const unsigned char* idx, float* dx, float* db, int B, int C, int H,
                                  int W, void* scratch, long long scratch_bytes, int device, void* stream);
int b200ddp_relu_bwd_bias(const float* dy, const float* y, float* dx, float* db, int B, int C, int HW, void* scratch,
                          long long scratch_bytes, int device, void* stream);
int b200ddp_relu_fwd(float* x, long long n, int device, void* stream);
int b200ddp_relu_bwd(const float* dy, const float* y, float* dx, long long n, int device, void* stream);
int b200ddp_dropout_fwd(const float* x, float* y, unsigned char* gate, long long n, float p, unsigned long long seed,
                        int device, void* stream);

/* ---- B1: SyncBatchNorm --------------------------------------------------------------------------------- */
/* Replace aten::batch_norm_stats / _gather_stats_with_counts / _elemt / _backward_reduce /
 * _backward_elemt as called by torch/nn/modules/_functions.py:7-209 (README.md:79-81 of the reference).
 * x,y,dy,dx: [B][C][HW] fp32. scratch: 256-byte aligned, b200ddp_bn_scratch_bytes(B,C) bytes, zeroed
 * once before first use (the kernels reset their tickets). y_relu (nullable): post-ReLU output, masks dy. */
long long b200ddp_bn_scratch_bytes(int B, int C);
int b200ddp_bn_stats(const float* x, int B, int C, int HW, float eps, float* mean, float* invstd, void* scratch,
                     long long scratch_bytes, int device, void* stream);
int b200ddp_bn_gather_stats(const float* all_stats /* [world][2C+1] */, int world, int C, float eps, float momentum,
                            float* running_mean, float* running_var, float* mean, float* invstd, int device,
                            void* stream);
int b200ddp_bn_elemt(const float* x, const float* mean, const float* invstd, const float* weight, const float* bias,
                     float* y, int B, int C, int HW, int relu, int device, void* stream);
int b200ddp_bn_bwd_reduce(const float* dy, const float* x, const float* y_relu, const float* mean, const float* invstd,
                          int B, int C, int HW, float* sum_dy, float* sum_dy_xmu, float* grad_weight, float* grad_bias,
                          void* scratch, long long scratch_bytes, int device, void* stream);
int b200ddp_bn_bwd_elemt(const float* dy, const float* x, const float* y_relu, const float* mean, const float* invstd,
                         const float* weight, const float* sum_dy, const float* sum_dy_xmu, float inv_count, float* dx,
                         int B, int C, int HW, int device, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* B200DDP_H_ */
