/*
 * cvt_b200.h — C-ABI of libcvt_b200.so, the B200 (sm_100a) implementation of the CvT
 * forward/backward hot path of POCHUN-CHEN/Transformer-STM.
 *
 * The reference has no FFI of its own: its boundary for this path is the Keras-2 Python surface its
 * scripts call.  Each entry point below cites the reference call it stands in for (paths relative
 * to the reference repo).  The Python mirror of that surface (transformer-stm_b200/keras_api.py)
 * binds these symbols with ctypes; INTEGRATION.md shows the binding a reference maintainer would add.
 *
 * Conventions
 *   - every function returns 0 on success or a negative cvt_status; cvt_last_error() gives the message
 *     of the calling thread's last failure.  Nothing throws, nothing falls back to the CPU.
 *   - one cvt_model per GPU; a model is not thread-safe; all device work runs on the model's stream.
 *   - `*_dev` pointers are device pointers, every other pointer is host memory owned by the caller.
 *   - tensors are fp32, NHWC / [B, T, D] row-major exactly as in the reference.
 */
#ifndef CVT_B200_H
#define CVT_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(__GNUC__)
#define CVT_API __attribute__((visibility("default")))
#else
#define CVT_API
#endif

#define CVT_MAX_STAGES 4
#define CVT_DATASET_SLOTS 4
#define CVT_MAX_PEERS 8          /* ranks of one NVSwitch box whose workspaces a model can map (cvt_comm_peer_open) */
#define CVT_ABI_VERSION 3

typedef enum {
  CVT_OK = 0,
  CVT_ERR_INVALID = -1,      /* bad argument / unsupported configuration */
  CVT_ERR_CUDA = -2,         /* a CUDA runtime call or kernel launch failed */
  CVT_ERR_NO_DEVICE = -3,    /* no CUDA device / wrong architecture: there is no CPU fallback */
  CVT_ERR_STATE = -4,        /* call order (e.g. backward without a training forward) */
  CVT_ERR_NCCL = -5
} cvt_status;

/* qkv_method of a stage — Projection.__init__, models/CvT(Par).py:92-100; LINEAR is the identity there (self.proj = None, :101),
 * LINEAR_DENSE is the Dense(dim) of the Grad-CAM script's Projection (tools/grad_cam_CvT.py:103-104,117) */
enum { CVT_PROJ_DW_BN = 0, CVT_PROJ_AVG = 1, CVT_PROJ_LINEAR = 2, CVT_PROJ_LINEAR_DENSE = 3 };
/* head of create_cvt_model.  AUTO = models/CvT(Par).py:326-337 (LayerNorm of the cls token when the last stage has one, else
 * LayerNorm + mean over the tokens); CLS / MEAN = the un-normalised heads of tools/model_plot.py:289-292; MEAN is also the
 * GlobalAveragePooling2D head of tools/grad_cam_CvT.py:336 (the cls token, if any, is dropped first) */
enum { CVT_HEAD_AUTO = 0, CVT_HEAD_CLS_LN = 1, CVT_HEAD_LN_MEAN = 2, CVT_HEAD_CLS = 3, CVT_HEAD_MEAN = 4 };
/* arithmetic of the dense contractions */
enum { CVT_PREC_FP32 = 0,   /* fp32 FFMA everywhere (exact-parity mode) */
       CVT_PREC_TF32 = 1,   /* tcgen05 kind::tf32, fp32 accumulate in TMEM (what TF does on sm_80+) */
       CVT_PREC_BF16 = 2,   /* tcgen05 kind::f16 (bf16 operands), fp32 accumulate: NOT implemented, cvt_create refuses it */
       CVT_PREC_TF32X3 = 3 }; /* error-compensated tcgen05: every dense product as hi.hi + lo.hi + hi.lo (operand tiles split in shared memory),
                                 activations kept unrounded; attention stays single-pass TF32.  Between TF32 speed and fp32 accuracy. */
/* dropout source */
enum { CVT_DROPOUT_OFF = 0, CVT_DROPOUT_PHILOX = 1, CVT_DROPOUT_MASKS = 2 };

/* One entry of the reference's `spec['stages']` list, models/CvT(Par).py:66-72 */
typedef struct {
  int32_t embed_dim;      /* multiple of 64 (key_dim = 64 in every reference spec) */
  int32_t patch_size;     /* ConvEmbed kernel (7 / 3 / 3) */
  int32_t stride;         /* ConvEmbed stride (4 / 2 / 2) */
  int32_t num_heads;      /* embed_dim / 64 */
  int32_t qkv_method;     /* CVT_PROJ_* */
  int32_t with_cls_token; /* 0/1 per stage: models/CvT(Par).py sets it on the last stage only, tools/grad_cam_CvT.py:324-332 on every
                             stage (the token is concatenated at block entry and split off again at block exit, :264-268,283-286) */
} cvt_stage_config;

/* Mirrors the module-level constants + `spec` that create_cvt_model reads, models/CvT(Par).py:25-72,292 */
typedef struct {
  int32_t image_h, image_w, in_chans;     /* 128,128,1  (models/CvT(Par).py:40-42) */
  int32_t n_stages;
  cvt_stage_config stage[CVT_MAX_STAGES];
  int32_t proc_dim;                       /* 5; 0 = CvT(Img) variant (models/CvT(Img).py:291) */
  int32_t proc_hidden;                    /* 256 (Proc_Dense_1/2, models/CvT(Par).py:343-344) */
  int32_t num_classes;                    /* 1 (models/CvT(Par).py:44) */
  int32_t mlp_ratio;                      /* 4 (Mlp_dim_factor, :232) */
  float   drop_rate;                      /* 0.1 (:116,232) */
  float   ln_eps;                         /* 1e-6 (:248,328) */
  float   bn_eps;                         /* 1e-3 (Keras BatchNormalization default) */
  float   bn_momentum;                    /* 0.99 */
  int32_t precision;                      /* CVT_PREC_* */
  int32_t max_batch;                      /* largest per-call batch; bigger host batches are chunked */
  int32_t device;                         /* CUDA ordinal */
  /* variants of the tools/ scripts; all-zero = models/CvT(Par).py */
  int32_t head_kind;                      /* CVT_HEAD_* */
  int32_t skip_proj_qkv;                  /* 1: proj_q/k/v are not applied and own no weights (tools/grad_cam_CvT.py:146-149 vs :168-201) */
  float   attn_drop_rate;                 /* > 0: Dropout on the MultiHeadAttention output before self.proj (tools/grad_cam_CvT.py:200) */
} cvt_config;

typedef struct cvt_model cvt_model;

/* ---- lifetime ------------------------------------------------------------------------------- */
CVT_API int  cvt_abi_version(void);
/* fills `cfg` with the reference configuration (spec at models/CvT(Par).py:66-72, dw_bn, cls on stage 3) */
CVT_API void cvt_default_config(cvt_config* cfg);
/* create_cvt_model(image_height, image_width, num_channels, proc_dim, num_classes) — models/CvT(Par).py:292 */
CVT_API int  cvt_create(const cvt_config* cfg, cvt_model** out);
CVT_API void cvt_destroy(cvt_model* m);
CVT_API const char* cvt_last_error(void);

/* ---- weights: model.get_weights / set_weights / save_weights / load_weights (models/CvT(Par).py:489,
 *      models/CvT_test(Par).py:513).  Canonical names in DESIGN.md; shapes are Keras' own. ------------ */
CVT_API int  cvt_param_count(const cvt_model* m);
CVT_API int  cvt_param_info(const cvt_model* m, int index, const char** name, int32_t* rank, int64_t shape[4],
                    int32_t* trainable, int64_t* numel);
CVT_API int  cvt_param_index(const cvt_model* m, const char* name);          /* <0 if unknown */
CVT_API int  cvt_set_param(cvt_model* m, int index, const float* host_src);
CVT_API int  cvt_get_param(cvt_model* m, int index, float* host_dst);
CVT_API int  cvt_get_grad(cvt_model* m, int index, float* host_dst);          /* after cvt_loss_backward; trainable only */
/* glorot_uniform / zeros / ones exactly as the Keras layers of the reference initialise */
CVT_API int  cvt_init_params(cvt_model* m, uint64_t seed);

/* ---- model.predict / model(x, training=...) — models/CvT_test(Par).py:522, Keras train_step forward -- */
/* img_dev [B,H,W,C], proc_dev [B,proc_dim] or NULL (Img variant), out_dev [B,num_classes]; B <= max_batch */
CVT_API int  cvt_forward(cvt_model* m, const float* img_dev, const float* proc_dev, int batch, int training, float* out_dev);
/* host-buffer predict: H2D, forward(training=0) in chunks of max_batch, D2H.  n may exceed max_batch. */
CVT_API int  cvt_predict_host(cvt_model* m, const float* img, const float* proc, int64_t n, float* out);

/* ---- model.compile(Adam(1e-3), 'mean_squared_error', ['mae']) + train_step — models/CvT(Par).py:464-476 */
/* MSE / MAE of the last training forward against y_dev [B,nc]; fills every gradient (no optimizer step). */
CVT_API int  cvt_loss_backward(cvt_model* m, const float* y_dev, float* loss_host, float* mae_host);
/* Keras Adam step on all trainable weights with the gradients currently held (eps 1e-7, bias-corrected). */
CVT_API int  cvt_adam_step(cvt_model* m, float lr);
/* one full train step on device buffers: forward(training) + loss + backward + [all-reduce] + Adam +
 * BN moving-stat update.  loss/mae are written to device scalars readable with cvt_read_metrics. */
CVT_API int  cvt_train_step(cvt_model* m, const float* img_dev, const float* proc_dev, const float* y_dev, int batch, float lr);
/* same from host buffers (pinned staging inside): the call `model.fit` makes per batch */
CVT_API int  cvt_train_step_host(cvt_model* m, const float* img, const float* proc, const float* y, int batch, float lr,
                         float* loss, float* mae);
/* the batch loop of one fit epoch (models/CvT(Par).py:475-476) from host arrays in the caller's (shuffled) order: batches of
 * `batch` samples, last partial batch kept; loss_out / mae_out [ceil(n / batch)].  Host->device copies of batch i+1 overlap the
 * step on batch i (pinned host memory makes them asynchronous). */
CVT_API int  cvt_train_epoch_host(cvt_model* m, const float* img, const float* proc, const float* y, int64_t n, int batch, float lr,
                          float* loss_out, float* mae_out);
/* evaluation (validation pass of fit): forward(training=0) + MSE/MAE sums over n host samples */
CVT_API int  cvt_evaluate_host(cvt_model* m, const float* img, const float* proc, const float* y, int64_t n,
                       float* loss, float* mae);
CVT_API int  cvt_read_metrics(cvt_model* m, float* loss, float* mae);         /* synchronises the model stream */
CVT_API int  cvt_optimizer_reset(cvt_model* m);                                /* zero m, v, t (a fresh compile) */
CVT_API int64_t cvt_optimizer_step_count(const cvt_model* m);

/* ---- input pipeline and resident datasets ------------------------------------------------------------------
 * preprocess_data (models/CvT(Par).py:363-428) decodes 200 JPEGs per part on the host, resizes them with cv2 to 128x128, converts to
 * grey, divides by 255 and hands model.fit float64 arrays [N,H,W], [N,5], [N]; Keras then shuffles and copies one batch per step.
 * Here the decoded 8-bit images can be resized / converted on the device (bit-exact with cv2's 8-bit fixed-point path), the whole set
 * stays in HBM (40 000 x 128 x 128 fp32 = 2.6 GB) and an epoch is a permutation: batches are gathered by index on the device. */
typedef struct {
  const void* img;   /* [n, H, W, C]  float32, or float64 when img_f64  (the reference passes float64 [N,H,W])   */
  const void* proc;  /* [n, proc_dim] float32 / float64; NULL for the Img variant                                 */
  const void* y;     /* [n, num_classes] float32 / float64; NULL where no targets are needed                      */
  int32_t img_f64, proc_f64, y_f64;
  int64_t n;
} cvt_host_arrays;
CVT_API int  cvt_dataset_alloc(cvt_model* m, int slot, int64_t n);          /* slot in [0, CVT_DATASET_SLOTS) */
CVT_API int  cvt_dataset_free(cvt_model* m, int slot);
CVT_API int64_t cvt_dataset_count(const cvt_model* m, int slot);
/* rows [offset, offset + a->n) <- host arrays (pageable or pinned; any member may be NULL = left as is) */
CVT_API int  cvt_dataset_write(cvt_model* m, int slot, int64_t offset, const cvt_host_arrays* a);
/* images [offset, offset + count) <- decoded 8-bit pictures [count, src_h, src_w, src_c] (cv2.imread layout, src_c = 3 BGR or 1):
 * cv2.resize(INTER_LINEAR) to (image_w, image_h), cv2.cvtColor(BGR2GRAY), / 255 — models/CvT(Par).py:420-423.  in_chans must be 1. */
CVT_API int  cvt_dataset_write_u8(cvt_model* m, int slot, int64_t offset, int64_t count, const uint8_t* pixels, int src_h, int src_w, int src_c);
CVT_API int  cvt_dataset_read(cvt_model* m, int slot, int64_t offset, int64_t count, float* img, float* proc, float* y);   /* any may be NULL */
/* one epoch of model.fit (models/CvT(Par).py:475) over host arrays, batch i = rows order[i*batch .. ) (order == NULL: 0..n-1).  The rows of a
 * batch are gathered into a pinned ring (shuffle and pageable->pinned staging are one pass over 4 MB, not a copy of the data set), copied
 * to the device while the previous step runs, and - when fill_slot >= 0 - also stored at their own index of that resident dataset, so
 * that the first epoch of a fit uploads the set as a side effect and later epochs call cvt_train_epoch_resident. */
CVT_API int  cvt_train_epoch_stream(cvt_model* m, const cvt_host_arrays* a, const int32_t* order, int64_t n_order, int batch, float lr,
                                    int fill_slot, float* loss_out, float* mae_out);
CVT_API int  cvt_train_epoch_resident(cvt_model* m, int slot, const int32_t* order, int64_t n_order, int batch, float lr,
                                      float* loss_out, float* mae_out);
/* model.predict / model.evaluate on the caller's arrays as they are (float32 or float64, pageable or pinned), chunks of max_batch */
CVT_API int  cvt_predict_arrays(cvt_model* m, const cvt_host_arrays* a, float* out);
CVT_API int  cvt_evaluate_arrays(cvt_model* m, const cvt_host_arrays* a, float* loss, float* mae);
/* validation pass / predict over a resident dataset: forward(training=0) in chunks of max_batch */
CVT_API int  cvt_evaluate_resident(cvt_model* m, int slot, float* loss, float* mae);
CVT_API int  cvt_predict_resident(cvt_model* m, int slot, float* out);

/* ---- dropout (layers.Dropout at models/CvT(Par).py:189,255,257) ------------------------------- */
/* PHILOX: keep-masks come from Philox4x32-10(seed, step, site) — reproducible by the oracle.
 * MASKS : the caller supplies uint8 keep-masks per site with cvt_set_dropout_mask (parity tests). */
CVT_API int  cvt_set_dropout(cvt_model* m, int mode, uint64_t seed);
/* site = stage*4 + {0: attention proj, 1: mlp hidden, 2: mlp out}; count = B*Tc*width */
CVT_API int  cvt_set_dropout_mask(cvt_model* m, int site, const uint8_t* host_mask, int64_t count);

/* ---- Grad-CAM taps: Model(inputs, [get_layer(name).output, output]) + tape.gradient
 *      (tools/grad_cam_CvT.py:430-457) and the per-layer parity tests --------------------------------- */
/* copies a named activation of the last forward to host; returns its element count (<0 on error).
 * names: "s{i}.embed|x0|xn|conv_qkv|qkv0|qkv1|qkv2|attn_o|a1|x1|y|hpre|hact|x2|out", "head.feat", "pred" */
CVT_API int64_t cvt_get_tap(cvt_model* m, const char* name, float* host_dst, int64_t capacity);
/* d(sum_b dout[b,:]·pred[b,:]) / d tap for the block output "s{i}.x2"; forward must have run.  */
CVT_API int  cvt_backward_to_tap(cvt_model* m, const float* dout_host, const char* name, float* host_dst, int64_t capacity);

/* ---- data-parallel training: one process per GPU, gradients summed over NCCL/NVLink ------- */
/* unique_id: 128 bytes from cvt_comm_unique_id on rank 0, broadcast by the host (torch.distributed).  */
CVT_API int  cvt_comm_unique_id(uint8_t id_out[128]);
CVT_API int  cvt_comm_init(cvt_model* m, const uint8_t unique_id[128], int rank, int world);
CVT_API int  cvt_comm_destroy(cvt_model* m);
/* Peer-memory data parallelism inside one NVSwitch box (after cvt_comm_init, world <= CVT_MAX_PEERS): every rank exports a 64-byte
 * cudaIpc handle of its workspace, the host gathers the `world` handles (torch.distributed all_gather) and hands all of them to every
 * rank.  From then on cvt_train_step ends in ONE kernel that reduce-scatters the gradients over NVLink, applies Adam to the rank's slice
 * and all-gathers the new weights straight into the peers' arenas (csrc/dp_peer.cu) instead of ncclAllReduce + a full Adam pass.
 * sync_grads != 0 also leaves the reduced gradient in every rank's arena (cvt_get_grad in tests). */
CVT_API int  cvt_comm_peer_export(cvt_model* m, uint8_t handle_out[64]);
CVT_API int  cvt_comm_peer_open(cvt_model* m, const uint8_t* handles /* [world][64] */, int sync_grads);
/* back to the NCCL path (used when some rank of the job could not map its peers: all ranks must take the same path) */
CVT_API int  cvt_comm_peer_close(cvt_model* m);
/* diagnostics: how many dense contractions issued by the calling thread went to {the TMA-fed tcgen05 kernel, the older tcgen05 kernel, the
 * fp32 FFMA kernel} since the last reset.  In TF32 mode the last one should stay 0 for the shapes of the reference specs. */
CVT_API int  cvt_gemm_dispatch_counts(int64_t out[3], int reset);
/* flat trainable arenas so the host may run its own collective instead (torch.distributed plumbing) */
CVT_API int  cvt_grad_arena(cvt_model* m, float** grads_dev, int64_t* count);
CVT_API int  cvt_param_arena(cvt_model* m, float** params_dev, int64_t* count);

/* ---- streams / graphs / bookkeeping ----------------------------------------------------------------------- */
CVT_API int  cvt_synchronize(cvt_model* m);
CVT_API void* cvt_stream(cvt_model* m);                     /* cudaStream_t the model launches on */
CVT_API int  cvt_set_graphs(cvt_model* m, int enable);      /* capture train/predict steps into CUDA graphs */
CVT_API int64_t cvt_kernel_launches(const cvt_model* m);    /* kernels this model has launched so far (own kernels only) */
CVT_API int64_t cvt_workspace_bytes(const cvt_model* m);
/* name of the kernel group timed last + its device time; used by bench.py's roofline leg */
CVT_API int  cvt_profile_step(cvt_model* m, const float* img_dev, const float* proc_dev, const float* y_dev, int batch,
                      int training, int max_entries, const char** names, float* ms, int* n_entries);

/* (group name, kernels launched) of the last cvt_profile_step in issue order: lets a profiler's launch list be cut into the groups */
CVT_API int  cvt_profile_group_launches(cvt_model* m, int max_entries, const char** names, int* launches, int* n_entries);

/* diagnostics: C[M,N] = A x B through the fp32 FFMA kernel (use_tc = 0) or the tcgen05 kernel (use_tc = 1) on device 0.
 * a_mode 0: A is [M,K] row-major, 1: A is [K,M] (read transposed);  b_mode 0: B is [K,N], 1: B is [N,K]. */
CVT_API int  cvt_debug_gemm(int a_mode, int b_mode, int M, int N, int K, int splitk, int use_tc, const float* A, const float* B, float* C);

/* diagnostics: the tcgen05 flash-attention kernels on host arrays.  Q (already scaled by 1/sqrt(64)), K, V, dO, O, dQ, dK, dV are
 * [B, Tc, heads*64]; LSE is [B*heads, Tc].  dO == NULL runs the forward only.  dQ is w.r.t. the un-scaled query (x 0.125). */
CVT_API int  cvt_debug_attention(int B, int Tc, int heads, const float* Q, const float* K, const float* V, const float* dO,
                         float* O, float* LSE, float* dQ, float* dK, float* dV);

/* Allocation-free entry point.  NO COUNTERPART IN THE REFERENCE: added for callers that cannot
 * call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * The tcgen05 flash-attention backward (the kernels cvt_debug_attention / the model's backward pass run) on caller-owned DEVICE buffers:
 * Q (already scaled by 1/sqrt(64)), K, V, O, dO, dQ, dK, dV are [B, Tc, heads*64]; LSE is [B*heads, Tc] (O and LSE as the forward left
 * them); dQ is w.r.t. the un-scaled query (x 0.125).  Operands are expected rounded to TF32 (the tensor core truncates otherwise).
 * The kernels' scratch, Dd[B*heads, Tc] = rowsum(dO * O), lives in a static __device__ array of CVT_STATIC_ATTN_MAX_ROWS floats (4 MiB,
 * held by every GPU that loads the library until exit): B * heads * Tc beyond that is refused with CVT_ERR_INVALID.  The call performs
 * no cudaMalloc / cudaFree, launches on `stream` (a cudaStream_t; NULL = the default stream) and does not synchronise.
 * Every call on a device shares the one array: calls that overlap on the GPU race - serialise them on one stream.  Nothing in the array
 * needs clearing between calls (every element read is rewritten first). */
#define CVT_STATIC_ATTN_MAX_ROWS (1 << 20)
CVT_API int  cvt_attention_backward_static(int B, int Tc, int heads, const float* Q_dev, const float* K_dev, const float* V_dev,
                                           const float* O_dev, const float* LSE_dev, const float* dO_dev, float* dQ_dev, float* dK_dev,
                                           float* dV_dev, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* CVT_B200_H */
