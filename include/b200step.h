/*
 * b200step.h — C ABI of libb200step.so: the sm_100a kernels behind the data-parallel training
 * step of the reference's task.py (hot loop task.py:303-312).
 *
 * The reference has no FFI of its own (it is one Python script that calls torch); each entry
 * point below names the reference call site whose work it replaces.  Conventions:
 *   - every function returns 0 on success, non-zero on error; b200_last_error() gives the text;
 *   - all pointers are DEVICE pointers unless the comment says "host";
 *   - the library never allocates the caller's data: the host (PyTorch) owns every buffer and
 *     passes tensor.data_ptr(); work is enqueued on the caller's stream and does not synchronise;
 *   - kernels are launched with programmatic dependent launch (a kernel's prologue may overlap the tail of the
 *     previous kernel on the stream; it touches global memory only after that kernel has completed);
 *     B200STEP_PDL=0 in the environment turns this off;
 *   - matrices are row-major; `ld*` are row strides in ELEMENTS;
 *   - dtype codes: B200_F32 (fp32 storage, TF32 tensor-core math), B200_BF16 (bf16 storage, fp32 accumulate).
 */
#ifndef B200STEP_H_
#define B200STEP_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef void* b200_stream_t; /* a cudaStream_t */

#define B200_F32 0
#define B200_BF16 1

#define B200_FLAG_RELU 1      /* fwd: y = max(y,0);  dgrad: dx *= (act > 0) */
#define B200_FLAG_ACCUMULATE 2 /* wgrad: dw += ... instead of dw = ... */
#define B200_FLAG_REVERSE_M 4  /* fwd / dgrad: process the row tiles last-to-first.  Same result; a consumer that starts
                                  with the rows its producer wrote LAST finds them still in the 126 MB L2 */
#define B200_FLAG_DETERMINISTIC 8 /* wgrad: do not split the batch dimension across CTAs: every output tile is summed
                                  by one CTA pair in a fixed order, so the result is bit-reproducible run to run (the
                                  default splits the batch for occupancy and combines the slices with fp32 atomics) */

/* ---- errors / build info ------------------------------------------------------------------ */
const char* b200_last_error(void);           /* host string, thread-local */
int b200_abi_version(void);
int b200_device_sm_count(int* out_host);     /* SM count of the current device */
/* Size the persistent grids of every later launch (GEMMs, K-pre, K-loss, K-opt) for `sms` SMs instead of the whole
 * device; 0 restores the default.  Host-side state read at launch (or graph-capture) time, one caller thread.
 * No counterpart in the reference: DDP overlaps its all-reduce with cuBLAS/cuDNN kernels that are not persistent
 * (task.py:189).  Here a collective on another stream needs free SMs, otherwise its CTAs wait for a grid boundary
 * and then push part of the next persistent grid into a second wave.  Results do not depend on the budget. */
int b200_set_sm_budget(int sms);
/* Where later launches report input that torch would have rejected with an exception (the kernels cannot raise):
 * a zero-initialised uint32 in device memory onto which they OR the bits below; NULL (the default) turns reporting
 * off.  The host reads the word whenever it synchronises anyway.  Offending values are made harmless first:
 *   B200_STATUS_BAD_CATEGORICAL_ID  b200_preprocess met an id outside [0, table_rows) (nn.Embedding raises
 *                                   IndexError): that id is replaced by row 0;
 *   B200_STATUS_BAD_LABEL           b200_ce_loss_grad / b200_linear_ce met a label outside [0, num_classes) that is
 *                                   not ignore_index (nn.CrossEntropyLoss raises): that row is dropped like an
 *                                   ignored one. */
#define B200_STATUS_BAD_CATEGORICAL_ID 1u
#define B200_STATUS_BAD_LABEL 2u
/*   B200_STATUS_BAD_ROW_INDEX       b200_image_augment met a sampler index outside [0, n_rows) (Dataset.__getitem__
 *                                   raises IndexError): row 0 is read instead. */
#define B200_STATUS_BAD_ROW_INDEX 4u
int b200_set_status_word(uint32_t* device_word);

/* ---- K-pre: fused batch load + standardise + categorical gather ---------------------------
 * Replaces the reference's host-side input pipeline for one batch:
 *   DistributedSampler/DataLoader batching (task.py:260-263: batch b = rows row_idx[cursor*batch + b]),
 *   transforms.Normalize (task.py:250-251 -> tensor.sub_(mean).div_(std), IEEE fp32 subtract then divide),
 *   the two H2D copies (task.py:305,307), and the model's embedding lookup + concat.
 * Column order of x_out (embeddings first, so that every gathered float4 lands on an aligned chunk of the row;
 * the host keeps the first Linear's weight columns in the same order, b200step/model.py):
 * x_out[b, t*dim : (t+1)*dim]                          = tables[t][cat[r,t]]      (bit-exact copy / RNE to bf16)
 * x_out[b, num_cat*dim : num_cat*dim + num_dense]      = (dense[r] - mean) / std  (bit-exact fp32, RNE to bf16 if x_dtype=BF16)
 * x_out[b, num_cat*dim + num_dense : ldx]              = 0                        (K padding for the GEMMs)
 * y_out[b] = label[r];  idx_out[b,t] = (int32) cat[r,t]  (kept for the embedding backward)
 * ds_cat holds int64 ids (cat_id_bytes = 8, torch's index dtype) or int32 ids (cat_id_bytes = 4, the narrowed form
 * the host-fed path sends over PCIe).
 * cursor may be NULL (treated as 0); it is read on the device so a captured graph can be replayed.  The batch is
 * rows row_idx[*cursor * cursor_stride + b], b < batch; cursor_stride <= 0 means `batch` (a step may be processed
 * in several row chunks: each call then gets row_idx advanced by its chunk offset and cursor_stride = full batch).
 * label_stats (nullable): 4 zero-initialised int32 of device memory; after the kernel, label_stats[0] = number of
 * rows of this batch whose label != ignore_index (what nn.CrossEntropyLoss's mean divides by, loss.py:1379-1387);
 * words 1..3 are the kernel's own accumulator / ticket and return to zero, so a captured graph can be replayed.
 * max_ctas > 0 caps the grid (256-thread CTAs, no shared memory): a launch that is meant to run UNDER another
 * kernel (the next batch's preprocessing under this step's wgrad GEMMs) takes one or two CTA slots per SM instead of
 * filling the device, so the persistent GEMM grid launched after it still finds room on every SM.
 */
int b200_preprocess(const float* ds_dense, const void* ds_cat, int cat_id_bytes, const int64_t* ds_label,
                    const int64_t* row_idx, const int32_t* cursor,
                    const float* mean, const float* std, const float* tables,
                    void* x_out, int64_t* y_out, int32_t* idx_out,
                    int64_t batch, int num_dense, int num_cat, int64_t table_rows, int embed_dim,
                    int64_t ldx, int x_dtype, int64_t cursor_stride, int32_t* label_stats, int64_t ignore_index,
                    int max_ctas, b200_stream_t stream);

/* ---- K-gemm: tcgen05/TMEM linear layers ----------------------------------------------------
 * M = batch rows, N = out_features, K = in_features for all three.
 * fwd   (task.py:309, nn.Linear inside model(inputs)):  y[M,N] = act(x[M,K] · w[N,K]^T + bias[N])
 *        with optional inverted dropout after the activation: keep iff philox(seed,*step,layer,elem) >= p, where
 *        elem = (dropout_row0 + m) * N + n (dropout_row0 = index of x's first row inside the whole batch when a
 *        batch is processed in row chunks).
 * dgrad (task.py:311 loss.backward, addmm backward wrt input):
 *        dx[M,K] = (dy[M,N] · w[N,K]) * (act[M,K] > 0 if RELU) * dscale     (act = the stored layer input)
 *        relu_mask (nullable, either side): 1 bit per element, word [row, col/32], row stride ldmask words; the
 *        forward writes (y > 0) and the backward of the NEXT layer reads it instead of re-reading `act`.
 *        dbias (nullable) [K] = column sums of the stored dx = bias gradient of the layer that produced `act`
 *        (addmm backward wrt bias).  With `scratch` (b200_dgrad_scratch_floats): deterministic two-stage reduction,
 *        dbias is overwritten.  With scratch == NULL: every CTA adds its column sums onto dbias with float
 *        atomics — dbias must hold zeros (or a value to accumulate onto) and the sum order is not fixed.
 * wgrad (task.py:311, addmm backward wrt weight): dw[N,K] (+)= dy[M,N]^T · x[M,K], fp32 result.
 * K and ld* must make every row 16-byte aligned (K%4==0 for F32, K%8==0 for BF16; same for N where N is a row length).
 */
int b200_linear_fwd(const void* x, int64_t ldx, const void* w, int64_t ldw, const float* bias,
                    void* y, int64_t ldy, int64_t M, int64_t N, int64_t K, int dtype, int flags,
                    float dropout_p, uint64_t seed, const int32_t* step, int layer_id,
                    int64_t dropout_row0, uint32_t* relu_mask, int64_t ldmask, b200_stream_t stream);
int b200_linear_dgrad(const void* dy, int64_t lddy, const void* w, int64_t ldw,
                      const void* act, int64_t ldact, void* dx, int64_t lddx,
                      int64_t M, int64_t N, int64_t K, int dtype, int flags, float dscale,
                      const uint32_t* relu_mask, int64_t ldmask,
                      float* dbias, float* scratch, int64_t scratch_floats, b200_stream_t stream);
/* dgrad of the FIRST Linear fused with the embedding backward (task.py:311; addmm backward wrt input followed by
 * embedding_dense_backward): the input row of that layer starts with the concatenated embedding rows
 * [num_tables x embed_dim], so
 *     dtables[t][idx[m,t]][:] += (dy[M,N] . w[N, 0 : num_tables*embed_dim])[m, t*embed_dim : (t+1)*embed_dim]
 * straight from the fp32 accumulator (red.add; dtables must hold zeros or the value to add onto).  The input
 * gradient itself is never stored — it has no other consumer — which removes its HBM round trip and the separate
 * b200_emb_bwd launch.  idx: the int32 ids b200_preprocess wrote ([M, ld_idx]); embed_dim: a power of two >= 4. */
int b200_linear_dgrad_emb(const void* dy, int64_t lddy, const void* w, int64_t ldw, const int32_t* idx,
                          int64_t ld_idx, float* dtables, int64_t M, int64_t N, int num_tables,
                          int64_t table_rows, int embed_dim, int dtype, b200_stream_t stream);
/* wgrad of `count` (<= 8) layers that share the batch dimension M, in ONE launch: dw[j] += dy[j]^T . x[j].  All
 * arrays are HOST arrays of length count.  Results are accumulated with red.add: every dw[j] must hold zeros (or the
 * value to add onto).  One launch instead of `count`, and the epilogue of one work unit overlaps the MMAs of the next. */
int b200_linear_wgrad_grouped(int count, const void* const* dy, const int64_t* lddy, const void* const* x,
                              const int64_t* ldx, float* const* dw, const int64_t* lddw, int64_t M,
                              const int64_t* N, const int64_t* K, int dtype, b200_stream_t stream);
/* scratch floats b200_linear_dgrad needs when dbias != NULL */
int64_t b200_dgrad_scratch_floats(int64_t M, int64_t K);
int b200_linear_wgrad(const void* dy, int64_t lddy, const void* x, int64_t ldx,
                      float* dw, int64_t lddw, int64_t M, int64_t N, int64_t K, int dtype, int flags,
                      b200_stream_t stream);

/* ---- K-loss: fused cross-entropy loss + gradient (+ last-layer bias gradient) --------------
 * Replaces nn.CrossEntropyLoss()(outputs, labels) (task.py:211,310; mean reduction) and its backward
 * (task.py:311): loss_out[0] = mean_b(logsumexp(z_b) - z_b[y_b]);
 * dlogits[b,c] = (softmax(z_b)[c] - [c==y_b]) * grad_scale   (grad_scale = 1/(batch*world): DDP's mean, task.py:189)
 * dbias (nullable) [C] = sum_b dlogits[b,:] as stored (after rounding to dtype), fp32.
 * scratch: b200_ce_scratch_floats(C) floats of device memory, contents need not be initialised.
 * loss_scale <= 0 means 1/batch; accumulate != 0 adds loss and dbias onto the existing values (a batch processed in
 * row chunks: grad_scale = 1/(full batch * world), loss_scale = 1/full batch, accumulate on every chunk but the first).
 * ignore_index (torch's default: -100): rows with that label add neither loss nor gradient (their dlogits row is 0).
 * n_valid (nullable): device int32 = number of rows of the batch whose label != ignore_index (label_stats[0] of
 * b200_preprocess); when given, loss and gradient are divided by it instead of by `batch` (torch's mean over the
 * non-ignored targets; 0 valid rows -> loss NaN, gradient 0, like torch).  With NULL the divisor stays `batch`.
 * A label outside [0, num_classes) that is not ignore_index sets B200_STATUS_BAD_LABEL and is treated as ignored.
 */
int64_t b200_ce_scratch_floats(int num_classes);
int b200_ce_loss_grad(const void* logits, int64_t ldl, const int64_t* labels,
                      void* dlogits, int64_t lddl, float* loss_out, float* dbias,
                      float* scratch, int64_t scratch_floats,
                      int64_t batch, int num_classes, float grad_scale, float loss_scale, int accumulate,
                      int64_t ignore_index, const int32_t* n_valid, int dtype, b200_stream_t stream);

/* ---- K-gemm-ce: the output layer and the loss as ONE kernel ------------------------------------
 * Replaces b200_linear_fwd (last nn.Linear, task.py:309) followed by b200_ce_loss_grad (task.py:310-311) without
 * ever storing the logits: z = x[M,K]·w[N,K]^T + bias kept in fp32 (never rounded to `dtype`, because never stored),
 * loss_out[0] = mean_b(logsumexp(z_b) - z_b[y_b]), dlogits[M,N] = (softmax(z) - onehot(y)) * grad_scale,
 * dbias[N] (nullable) = column sums of dlogits as stored.  Same semantics for grad_scale / loss_scale / accumulate /
 * ignore_index / n_valid / out-of-range labels as b200_ce_loss_grad.  N <= 1024 classes (a cluster of up to four
 * CTA pairs holds one softmax row in tensor memory); scratch: b200_linear_ce_scratch_floats(N) floats, contents
 * need not be initialised.  Loss and dbias are summed in a fixed order (run-to-run deterministic). */
int64_t b200_linear_ce_scratch_floats(int64_t N);
int b200_linear_ce(const void* x, int64_t ldx, const void* w, int64_t ldw, const float* bias,
                   const int64_t* labels, void* dlogits, int64_t lddl, float* loss_out, float* dbias,
                   float* scratch, int64_t scratch_floats, int64_t M, int64_t N, int64_t K, int dtype,
                   float grad_scale, float loss_scale, int accumulate, int64_t ignore_index,
                   const int32_t* n_valid, b200_stream_t stream);

/* ---- column sum: bias gradients of the hidden layers (task.py:311, addmm backward wrt bias) -
 * out[N] = sum_m dy[m, :]  (fp32, deterministic two-stage reduction through scratch) */
int64_t b200_colsum_scratch_floats(int64_t N);
int b200_colsum(const void* dy, int64_t lddy, float* out, float* scratch, int64_t scratch_floats,
                int64_t M, int64_t N, int dtype, b200_stream_t stream);

/* Column sum for callers that can neither pass scratch nor tolerate an allocation inside the call (a bare C
 * binding with no allocator of its own; a process whose framework memory pool already owns the card, where a raw
 * cudaMalloc can fail or fragment the pool; a step that overlaps communication, which the device-wide
 * synchronisation of cudaMalloc/cudaFree would stall).  No counterpart in the reference: it exists only for
 * those callers.  Same result as b200_colsum (the partials may be grouped differently, so sums agree to fp32
 * rounding, not bit for bit).  The partial sums live in a 1 MiB device array that is part of the library and
 * stays resident on every GPU that loads it; the call allocates nothing.
 *   - limit: N <= 4096 columns, otherwise the call fails with an explanatory error;
 *   - the first call on each device looks up the array's device address; make that first call outside any
 *     stream capture;
 *   - all calls share the one array: two calls that can overlap on the GPU (different streams, no ordering)
 *     race on it — order them, or use b200_colsum with separate scratch buffers. */
int b200_colsum_static(const void* dy, int64_t lddy, float* out, int64_t M, int64_t N, int dtype,
                       b200_stream_t stream);

/* ---- K-embbwd: embedding gradient scatter-add (task.py:311; embedding_dense_backward) -------
 * dtables[t][idx[b,t]][:] += dx[b, emb_col0 + t*dim : +dim]   (fp32 atomics; dtables must be zeroed by the caller;
 * emb_col0 % 4 == 0 — it is 0 in the layout b200_preprocess writes) */
int b200_emb_bwd(const void* dx, int64_t lddx, const int32_t* idx, float* dtables,
                 int64_t batch, int emb_col0, int num_cat, int64_t table_rows, int embed_dim,
                 int dtype, b200_stream_t stream);

/* ---- K-opt: fused optimizer over the flat parameter buffer ----------------------------------
 * SGD (torch.optim.SGD, task.py:212-214 + .step() task.py:312; zero_grad task.py:308 is implicit —
 *      gradients are overwritten by the next backward):
 *      g' = g + wd*p;  buf = momentum*buf + g'  (buf starts at 0 == torch's first-step clone);  p -= lr*buf
 * Adam (additive --optimizer adam; torch.optim.Adam semantics, coupled L2): *step is the 0-based count of
 *      completed steps, the kernel uses t = *step + 1 for the bias corrections.
 * Hyper-parameters are doubles (python floats) and are rounded to fp32 where torch rounds them.
 * zero_grad != 0: the consumed gradient is overwritten with zeros in the same pass (the reference's
 *      optimizer.zero_grad(), task.py:308, moved behind the update) so the next backward's split-K / scatter
 *      accumulations need no separate memset.
 * adv_ticket (nullable): a zero-initialised device counter; when given, the last block of this launch to finish
 *      does what b200_advance_step does (cursor = (cursor+1) mod steps_per_epoch, step += 1; either pointer may be
 *      NULL) — pass it to the LAST optimizer launch of a step to save the separate bookkeeping kernel.
 * shadow_bf16 (nullable): bf16 copy of p refreshed for the first n_shadow elements (GEMM operand copy).
 */
int b200_sgd_step(float* p, float* g, float* momentum_buf, void* shadow_bf16,
                  int64_t n, int64_t n_shadow, double lr, double momentum, double weight_decay,
                  int zero_grad, int32_t* adv_cursor, int32_t* adv_step, int32_t steps_per_epoch,
                  uint32_t* adv_ticket, b200_stream_t stream);
int b200_adam_step(float* p, float* g, float* exp_avg, float* exp_avg_sq, void* shadow_bf16,
                   int64_t n, int64_t n_shadow, double lr, double beta1, double beta2, double eps,
                   double weight_decay, const int32_t* step, int zero_grad, int32_t* adv_cursor,
                   int32_t* adv_step, int32_t steps_per_epoch, uint32_t* adv_ticket, b200_stream_t stream);
/* Operands for fp32-accurate linear layers on TF32 tensor cores ("3xTF32"; task.py:309 computes in IEEE fp32):
 * every value is split into hi = tf32(x) and lo = x - hi, and three copies are laid out along the contraction
 * dimension so that ONE TF32 GEMM over the 3x longer contraction computes a_hi.b_hi + a_hi.b_lo + a_lo.b_hi
 * (relative error ~2^-22 instead of 2^-11).  src[rows, cols] (row stride ld) ->
 *   kcat   [rows, 3*cols]  column block j = lo if bit j of kmask else hi     (contraction along the row)
 *   rstack [3*rows, cols]  row block j    = lo if bit j of rmask else hi     (contraction over rows)
 * either output may be NULL.  Pairings: left operand mask 0b100 (hi,hi,lo), right operand mask 0b010 (hi,lo,hi). */
int b200_split_tf32(const float* src, int64_t ld, int64_t rows, int64_t cols, float* kcat, int kmask,
                    float* rstack, int rmask, b200_stream_t stream);
/* fp32 -> bf16 RNE copy (initial shadow weights) */
int b200_cast_bf16(const float* src, void* dst_bf16, int64_t n, b200_stream_t stream);
/* zero `bytes` bytes of device memory on the stream (gradient regions that are accumulated with atomics) */
int b200_zero(void* ptr, int64_t bytes, b200_stream_t stream);
/* advance the device-side batch cursor (wraps at steps_per_epoch) and the global step counter */
int b200_advance_step(int32_t* cursor, int32_t* step, int32_t steps_per_epoch, b200_stream_t stream);

/* ---- K-comm: gradient all-reduce (DistributedDataParallel, task.py:189/194) -----------------
 * NCCL over NVLink; SUM in place on a bucket of the flat fp32 gradient buffer (the 1/world factor is folded
 * into grad_scale of b200_ce_loss_grad).  unique id: 128 host bytes made on rank 0 and shared by the host
 * (torch.distributed store) — the counterpart of DDP's process-group bootstrap at task.py:148. */
int b200_comm_unique_id(void* out_host_128);
int b200_comm_init(const void* unique_id_host_128, int rank, int world);
int b200_comm_allreduce_sum(float* buf, int64_t count, b200_stream_t stream);
int b200_comm_broadcast(float* buf, int64_t count, int root, b200_stream_t stream); /* DDP ctor param sync */
int b200_comm_destroy(void);
/* Optional: gradient memory that NCCL can reduce in place.  b200_comm_mem_alloc returns device memory from
 * ncclMemAlloc (the one allocation this library makes for the caller: NVSwitch multicast / in-switch reduction only
 * maps memory NCCL allocated itself); b200_comm_register registers a buffer with the communicator (after
 * b200_comm_init; deregistered by b200_comm_destroy).  The flat gradient buffer is the intended tenant:
 * DistributedDataParallel's buckets (task.py:189) are copies, here the all-reduce runs on the gradients themselves. */
int b200_comm_mem_alloc(int64_t bytes, void** out_host);
int b200_comm_mem_free(void* ptr);
int b200_comm_register(void* buf, int64_t bytes);

/* ---- streams / events / graphs (host plumbing for the step schedule) ------------------------ */
int b200_stream_create(b200_stream_t* out_host, int high_priority);
int b200_stream_destroy(b200_stream_t s);
int b200_stream_sync(b200_stream_t s);
int b200_event_create(void** out_host, int timing);
int b200_event_destroy(void* ev);
int b200_event_record(void* ev, b200_stream_t s);
int b200_stream_wait_event(b200_stream_t s, void* ev);
int b200_event_elapsed_ms(void* start, void* stop, float* out_host);
int b200_graph_begin(b200_stream_t s);
int b200_graph_end(b200_stream_t s, void** out_exec_host);
int b200_graph_launch(void* exec, b200_stream_t s);
int b200_graph_destroy(void* exec);

/* ---- convolutional path: ResNet-18 on 3x32x32 images (task.py:171 `models.__dict__[args.arch]()`, task.py:247-257) ----
 * Activations are NHWC = row-major [B*H*W, C] matrices, so every convolution is an im2col gather followed by the
 * tcgen05 GEMMs above: conv fwd = b200_im2col + b200_linear_fwd (bias NULL); conv backward wrt input =
 * b200_linear_dgrad + b200_col2im; wrt filter = b200_linear_wgrad(dy, col).  The filter is stored as
 * Wg[Cout][ky][kx][c] with row stride ldcol (torch's OIHW permuted to O,H,W,I; columns past kh*kw*C are zero).
 * All tensors: 16-byte aligned device pointers; `dtype` as above (storage type of activations). */

/* K-aug: the reference's input pipeline for one batch (task.py:247-263): DistributedSampler order (row_idx, cursor as
 * in b200_preprocess: sample b of the step is row_idx[cursor[0]*cursor_stride + b]; both nullable), then torchvision's
 * RandomCrop(W, padding=pad) + RandomHorizontalFlip + ToTensor + Normalize(mean3, std3) (transforms.py:653,720):
 * draws[p*4 + {0,1,2}] = {crop top, crop left, flip} for sample position p, drawn on the host (NULL: centre crop, no
 * flip = the evaluation transform); images uint8 [n_rows, H, W, 3]; x_out [batch, H, W, 3] in `dtype`, bit-identical
 * to torchvision in fp32; y_out[b] = labels[row] (nullable).  mean3 / std3 are HOST pointers, read during the call.
 * A row index outside [0, n_rows) sets B200_STATUS_BAD_ROW_INDEX. */
int b200_image_augment(const uint8_t* images, const int64_t* labels, const int64_t* row_idx, const int32_t* cursor,
                       int64_t cursor_stride, const int8_t* draws, int64_t n_rows, const float* mean3,
                       const float* std3, int pad, void* x_out, int64_t* y_out, int64_t batch, int H, int W,
                       int dtype, b200_stream_t stream);
/* col[m, (ky*kw+kx)*C + c] = x[b, ho*stride-pad+ky, wo*stride-pad+kx, c] (0 outside), m = (b*Ho+ho)*Wo+wo,
 * Ho = (H+2*pad-kh)/stride+1; columns [kh*kw*C, ldcol) are zero-filled.  (F.conv2d's unfold, torch/nn/modules/conv.py) */
int b200_im2col(const void* x, void* col, int64_t ldcol, int64_t B, int H, int W, int C, int kh, int kw, int stride,
                int pad, int dtype, b200_stream_t stream);
/* Convolution forward as an IMPLICIT GEMM (no col matrix): y[B*Ho*Wo, Cout] = conv2d(x NHWC, w) with w stored as
 * Wg[Cout][ky][kx][c] (row stride ldw).  The tcgen05 GEMM's A tiles are fetched by im2col-mode TMA straight from x
 * (padding halo zero-filled by the hardware).  C must be a multiple of 64 (bf16) / 32 (fp32); flags: B200_FLAG_RELU,
 * B200_FLAG_DETERMINISTIC.  ws (nullable, >= B*Ho*Wo*Cout floats): fp32 workspace that lets a convolution with few
 * output tiles but a long contraction (a ResNet's last stages) split the contraction over the idle SMs; the partial
 * tiles are combined with fp32 atomics (not with B200_FLAG_DETERMINISTIC) and cast to `dtype` by a second launch. */
int b200_conv_fwd(const void* x, const void* w, int64_t ldw, void* y, int64_t ldy, int64_t B, int H, int W, int C,
                  int kh, int kw, int stride, int pad, int64_t Cout, int dtype, int flags, float* ws,
                  int64_t ws_floats, b200_stream_t stream);
/* Input gradient of a STRIDE-1 convolution without the dcol matrix and the col2im pass: dx[B*H*W, C] = conv2d(dy,
 * flipped filter, padding k-1-pad); dy is [B, Ho, Wo, Cout] NHWC, w the same Wg[Cout][ky][kx][c] the forward uses (read
 * as an MN-major operand at the flipped tap).  Cout a multiple of 64 (bf16) / 32 (fp32), square filters.  flags / ws
 * (>= B*H*W*C floats) as for b200_conv_fwd. */
int b200_conv_dgrad(const void* dy, const void* w, int64_t ldw, void* dx, int64_t lddx, int64_t B, int H, int W, int C,
                    int kh, int kw, int pad, int64_t Cout, int dtype, int flags, float* ws, int64_t ws_floats,
                    b200_stream_t stream);
/* Filter gradient of the same convolution, also without a col matrix: dw[Cout, kh*kw*C] (+)= dy^T . unfold(x), fp32
 * (row stride lddw); flags: B200_FLAG_ACCUMULATE, B200_FLAG_DETERMINISTIC as for b200_linear_wgrad. */
int b200_conv_wgrad(const void* dy, int64_t lddy, const void* x, float* dw, int64_t lddw, int64_t B, int H, int W,
                    int C, int kh, int kw, int stride, int pad, int64_t Cout, int dtype, int flags,
                    b200_stream_t stream);
/* The adjoint, as a gather: dx[b,h,w,c] = sum of the dcol entries that im2col filled from x[b,h,w,c], plus
 * add[b,h,w,c] when add != NULL (the gradient arriving over a residual connection).  C % 4 == 0. */
int b200_col2im(const void* dcol, int64_t ldcol, const void* add, void* dx, int64_t B, int H, int W, int C, int kh,
                int kw, int stride, int pad, int dtype, b200_stream_t stream);

/* nn.BatchNorm2d in training mode (torch/nn/modules/batchnorm.py; F.batch_norm with training=True) over y[M, C],
 * M = B*H*W.  scratch: b200_bn_scratch_floats(C) floats that must be ZERO before the first use (each call leaves
 * them ready for the next).  C = 4 * (a divisor of 256): 64, 128, 256, 512 ...
 * _stats: mean[c], invstd[c] = 1/sqrt(biased var + eps); when running_mean != NULL also
 *         running = (1-momentum)*running + momentum*{mean, unbiased var} and num_batches[0] += 1 (nullable).
 *         Sums are folded in a fixed order (run-to-run deterministic).
 * _apply: out = act((y - mean)*invstd*gamma + beta (+ res)), act = ReLU when relu != 0.  With the running statistics
 *         passed as mean / invstd this is the eval-mode forward. */
int64_t b200_bn_scratch_floats(int64_t C);
int b200_bn_fwd_stats(const void* y, int64_t M, int64_t C, float eps, float momentum, float* mean, float* invstd,
                      float* running_mean, float* running_var, int64_t* num_batches, float* scratch,
                      int64_t scratch_floats, int dtype, b200_stream_t stream);
int b200_bn_fwd_apply(const void* y, const void* res, void* out, int64_t M, int64_t C, const float* mean,
                      const float* invstd, const float* gamma, const float* beta, int relu, int dtype,
                      b200_stream_t stream);
/* Backward of the same (task.py:311).  The incoming gradient is dout (+ dout2 when dout2 != NULL: the two branches that
 * meet at a residual connection, summed on the fly instead of by a separate pass).
 * dz = dout * (out > 0) when out != NULL (the ReLU that followed), else dout.
 * _reduce: dbeta[c] = sum_m dz, dgamma[c] = sum_m dz * xhat (fixed order);
 * _apply:  dy = gamma*invstd*(dz - dbeta/M - xhat*dgamma/M); dz_out (nullable) receives dz for the residual branch. */
int b200_bn_bwd_reduce(const void* dout, const void* dout2, const void* out, const void* y, int64_t M, int64_t C, const float* mean,
                       const float* invstd, float* dgamma, float* dbeta, float* scratch, int64_t scratch_floats,
                       int dtype, b200_stream_t stream);
int b200_bn_bwd_apply(const void* dout, const void* dout2, const void* out, const void* y, void* dy, void* dz_out, int64_t M, int64_t C,
                      const float* mean, const float* invstd, const float* gamma, const float* dgamma,
                      const float* dbeta, int dtype, b200_stream_t stream);

/* The same two pairs as ONE launch each: statistics + apply (b200_bn_fwd) and reduce + apply (b200_bn_bwd).  Every
 * block waits inside the kernel for the last block's per-channel result and then treats its own rows, which are still in
 * cache: one launch and one DRAM pass instead of two.  Same arguments, results and scratch rule as the pairs above. */
int b200_bn_fwd(const void* y, const void* res, void* out, int64_t M, int64_t C, float eps, float momentum, float* mean,
                float* invstd, const float* gamma, const float* beta, int relu, float* running_mean,
                float* running_var, int64_t* num_batches, float* scratch, int64_t scratch_floats, int dtype,
                b200_stream_t stream);
int b200_bn_bwd(const void* dout, const void* dout2, const void* out, const void* y, void* dy, void* dz_out, int64_t M,
                int64_t C, const float* mean, const float* invstd, const float* gamma, float* dgamma, float* dbeta,
                float* scratch, int64_t scratch_floats, int dtype, b200_stream_t stream);

/* nn.MaxPool2d(k, stride, pad) on NHWC; taps[m, c] (one byte per output element) records the winning window position
 * (first maximum in row-major window order, NaN wins — aten's rule) for the backward gather. */
int b200_maxpool_fwd(const void* x, void* out, uint8_t* taps, int64_t B, int H, int W, int C, int k, int stride,
                     int pad, int dtype, b200_stream_t stream);
int b200_maxpool_bwd(const void* dout, const void* dout2, const uint8_t* taps, void* dx, int64_t B, int H, int W, int C, int k,
                     int stride, int pad, int dtype, b200_stream_t stream);
/* nn.AdaptiveAvgPool2d((1,1)) on NHWC: out[b, c] = mean over the HW pixels; bwd: dx[b, p, c] = dout[b, c] / HW */
int b200_avgpool_fwd(const void* x, void* out, int64_t B, int HW, int C, int dtype, b200_stream_t stream);
int b200_avgpool_bwd(const void* dout, void* dx, int64_t B, int HW, int C, int dtype, b200_stream_t stream);

#ifdef __cplusplus
}
#endif
#endif /* B200STEP_H_ */
