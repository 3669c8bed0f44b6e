/*
 * ctr_b200.h -- C-ABI of libctr_b200.so: the B200 (sm_100a) train/predict hot path of the
 * RodJohn/deep_learning CTR zoo (DeepFM / Wide&Deep / DNN).
 *
 * The reference has NO native/FFI interface: every op on its hot path is a TensorFlow-1.x
 * Python op call inside models/<alg>.py.  Each entry point below therefore cites the reference
 * GRAPH LINES it replaces (paths relative to the reference root).  A maintainer binds this
 * library from the model modules with ctypes.
 *
 * Conventions
 *   - plain C, extern "C"; pointers are DEVICE pointers unless named host_*; sizes are explicit.
 *   - the caller owns every buffer, workspaces included; the library owns only the handle.
 *   - every call takes a cudaStream_t (passed as void*), enqueues and returns: no hidden
 *     synchronisation, no allocation, CUDA-graph capturable.
 *   - return 0 on success; negative = CTR_E_*; ctr_last_error(h) gives the message.  Never aborts.
 *   - one handle per device; a handle is not thread-safe.
 *   - there is NO CPU fallback anywhere.
 *
 * Layouts (floats unless noted)
 *   table row   : [ v_0 .. v_{k-1} | w | pad ]  `table_stride` floats, 32-byte multiple.  `w` (the
 *                 FM first-order weight of the same id, reference tables feats / fm_first_order_emb)
 *                 is present iff has_first.  Adam slots m, v use the same layout.
 *   X0 [B,x0_ld]: deep-tower input, canonical order [ single emb Fc*k | pooled emb Fm*k | cont C |
 *                 vector V | 0-pad ].  (The reference order is [cont|vector|emb]; the host permutes
 *                 the rows of deep_0 once at injection time.)
 *   fm  [B,fm_ld]: [ first single Fc | first multi Fm | first cont C | second k ].
 *   grad row    : [ g_v0 .. g_v{k-1} | g_w | pad ]  `grad_stride` floats.
 */
#ifndef CTR_B200_H
#define CTR_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define CTR_ABI_VERSION 4

#define CTR_OK              0
#define CTR_E_INVALID      -1   /* bad argument / unsupported shape */
#define CTR_E_CUDA         -2   /* CUDA runtime or launch error */
#define CTR_E_WORKSPACE    -3   /* workspace too small */
#define CTR_E_UNSUPPORTED  -4   /* feature not built (e.g. wrong arch) */

#define CTR_MAX_LAYERS 8
typedef struct ctr_handle ctr_handle;
typedef void* ctr_stream;        /* cudaStream_t */

/* shape/layout of the embedding side of one model; shared by forward and backward.
 * Mirrors the placeholders of DeepModel.__init__graph (models/deepfm.py:39-54,
 * models/deepfm_multi.py:46-79) plus the layouts above. */
typedef struct ctr_embed_cfg {
    int32_t B;             /* samples in this batch (input_data_size) */
    int32_t id_bytes;      /* 4 (models/deepfm.py:44 int32) or 8 (utils/data_loader.py:13 int64) */
    int32_t ids_ld;        /* columns of cate_feats = Fc + M */
    int32_t Fc;            /* single-valued categorical fields (cate_field_size) */
    int32_t Fm;            /* multi-valued fields (multi_field_size) */
    int32_t M;             /* total padded multi slots (multi_feats_size) */
    int32_t C;             /* continuous fields (cont_field_size) */
    int32_t V;             /* vector_feats width */
    int32_t k;             /* embedding_size: 4, 8, 16 or 32 */
    int32_t has_first;     /* rows carry the FM first-order weight at [k] */
    int32_t use_fm;        /* compute FM first/second order (DeepFM) */
    int32_t fm_cont;       /* continuous fields own table rows in the FM part (models/deepfm.py:63-75) */
    int32_t zero_row0;     /* id 0 reads as zeros and gets no gradient (models/deepfm_multi.py:130-133) */
    int32_t count_rule;    /* multi-valued mean: 0 = count rows whose sum != 0 (reference,
                              models/deepfm_multi.py:83-88), 1 = count ids != 0 */
    int32_t combiner;      /* 0 mean, 1 sum */
    int32_t table_stride;  /* floats per table row */
    int32_t x0_ld;         /* floats per X0 row (multiple of 4) */
    int32_t fm_ld;         /* floats per fm row */
    int32_t grad_stride;   /* floats per gradient row (multiple of 4, >= k + has_first) */
    int32_t round_tf32;    /* TF32 tower: write X0 rounded-to-nearest tf32 (tensor cores truncate fp32 operands) */
    int32_t reserved0;
    int64_t table_rows;    /* rows a fed id (+ id_offset) may address: [0, table_rows).  tf.nn.embedding_lookup raises
                              InvalidArgumentError for anything else (models/deepfm.py:78,85,98); here such a lookup reads
                              no memory, yields NaN (so logit, loss and score of the sample are NaN) and sets the status
                              word.  0 = unchecked. */
    uint32_t* status;      /* device, CTR_STATUS_WORDS words (see below) or NULL */
} ctr_embed_cfg;

/* Device status words shared by the kernels of one engine (all zero = healthy).  Kernels only ever set them; the host
 * clears them.  The optimizer entry points take the same array as ctr_adam_hp.guard and leave every weight untouched
 * while any word is non-zero: a batch with a bad id, a step whose peer barrier timed out or an exchange that would
 * overflow a buffer must not move the model. */
#define CTR_STATUS_WORDS       4
#define CTR_STATUS_STEP_BARRIER 0   /* 1 + rank that did not arrive (ctr_peer_barrier on the step's stream) */
#define CTR_STATUS_PLAN_BARRIER 1   /* same, id-plan barriers */
#define CTR_STATUS_BAD_ID       2   /* 1 + field (column of cate_feats / wide_feats) of an id outside its table */

/* ---- handle ------------------------------------------------------------------------- */
int  ctr_abi_version(void);
/* hex SHA-256 of the sources (csrc, this header, compiler flags) the library was built from, "unknown" for a build that
 * did not go through deep_learning_b200/build.py; the Python loader refuses a library whose hash differs from the sources
 * beside it (no counterpart in the reference: a build-hygiene hook) */
const char* ctr_source_hash(void);
int  ctr_sizeof_embed_cfg(void);
int  ctr_create(int device, ctr_handle** out);
int  ctr_destroy(ctr_handle* h);
const char* ctr_last_error(ctr_handle* h);
int  ctr_sm_count(ctr_handle* h);
/* kernels launched through this handle since ctr_create (bench.py's gpu_launches) */
int64_t ctr_launch_count(ctr_handle* h);
/* 1 if the tcgen05/TMA TF32 GEMM path can run on this device (sm_100) */
int  ctr_has_tcgen05(ctr_handle* h);

/* ---- table packing (host-side parameter injection / read-back) ------------------------
 * reference variables feats_emb [R,k] + feats [R,1] (models/deepfm.py:57-60) <-> packed rows */
int ctr_table_pack(ctr_handle* h, ctr_stream s, const float* emb /*[R,k]*/, const float* first /*[R] or NULL*/,
                   int64_t R, int k, int table_stride, float* table);
int ctr_table_unpack(ctr_handle* h, ctr_stream s, const float* table, int64_t R, int k, int table_stride,
                     float* emb /*[R,k]*/, float* first /*[R] or NULL*/);

/* ---- K2: multi-valued pooled gather ----------------------------------------------------
 * replaces multi_cate_emb()/nonzero_reduce_mean (models/deepfm_multi.py:81-121,
 * models/dnn_multi.py:83-104).  For every (sample, multi field j) pools the rows of slots
 * [multi_start[j], multi_start[j+1]) of the multi block of `ids`; writes the pooled vector into
 * X0[b, (Fc+j)*k ..], the pooled first-order weight into pooled_first[b,j] and the two divisors
 * (1/count, 0 when count==0) into inv_cnt_emb / inv_cnt_first [B,Fm].  `slot_weight` ([B,M] or
 * NULL) is the weighted form of models/dnn_multi_textline.py:94-103. */
int ctr_pooled_gather_fwd(ctr_handle* h, ctr_stream s, const ctr_embed_cfg* cfg,
                          const float* table, const void* ids, const int32_t* multi_start /*[Fm+1]*/,
                          const float* slot_weight, float* x0, float* pooled_first,
                          float* inv_cnt_emb, float* inv_cnt_first);

/* ---- K1: fused field-wise gather + FM first/second order + deep-input tile --------------
 * replaces models/deepfm.py:63-100 (index/value assembly, both embedding_lookups on feats_emb,
 * the lookup on feats, the FM reductions and the dense_vector concat) and the same lines of
 * models/deepfm_multi.py:137-190, models/wdl.py:132-179, models/dnn.py:54-57.
 * Each table row is read from HBM once.  `cont_rows` points at the C table rows of the
 * continuous fields (table + cont_base*stride on one GPU).  Outputs: X0 (single-emb block + cont +
 * vector columns; pooled block must already be there when Fm>0), fm [B,fm_ld] and
 * sum_e [B,k] (= sum over FM fields of e; kept for backward) when use_fm. */
int ctr_fm_gather_fwd(ctr_handle* h, ctr_stream s, const ctr_embed_cfg* cfg,
                      const float* table, const void* ids, int64_t id_offset,
                      const float* cont, const float* vec, const float* cont_rows,
                      const float* pooled_first,
                      float* x0, float* fm, float* sum_e);

/* scalar-table gather-sum: out[b] = sum_j table[ids[b,j]]   (wide part, models/wdl.py:240-246).  ids outside
 * [0, rows) contribute NaN and set status word CTR_STATUS_BAD_ID (rows = 0: unchecked; status may be NULL). */
int ctr_scalar_gather_sum(ctr_handle* h, ctr_stream s, const float* table, const void* ids, int id_bytes,
                          int B, int W, int64_t rows, uint32_t* status, float* out);

/* ---- K3: MLP tower ---------------------------------------------------------------------
 * replaces the matmul/add/relu/dropout loop (models/deepfm.py:121-126) and what
 * AdamOptimizer.minimize differentiates through it (models/deepfm.py:161).
 * precision: 0 = FP32 SIMT (tight 1e-5 checks), 1 = TF32 tcgen05/TMA (2e-3; outputs are written rounded to
 * tf32 and the caller passes tf32-rounded weights, so no operand is truncated by the tensor core).
 * W is the reference layout [K, N] row-major (ldw = N). */
#define CTR_PREC_FP32 0
#define CTR_PREC_TF32 1
/* Y[B,N] = dropout(relu(X[B,K] @ W[K,N] + bias[N])); relu/bias optional; keep==1 -> no dropout */
int ctr_mlp_fwd(ctr_handle* h, ctr_stream s, int precision, int B, int K, int N,
                const float* X, int64_t ldx, const float* W, const float* Wt /*[N, roundup(K,4)] or NULL*/,
                const float* bias, int relu, float keep, uint64_t seed, float* Y, int64_t ldy);
/* dX[B,K] = (dZ[B,N] @ W^T) * (mask[B,K] > 0 ? mask_scale : 0)   (mask = the layer's input
 * activation, NULL for the tower input) */
int ctr_mlp_bwd_data(ctr_handle* h, ctr_stream s, int precision, int B, int K, int N,
                     const float* dZ, int64_t lddz, const float* W,
                     const float* mask, int64_t ldmask, float mask_scale, float* dX, int64_t lddx);
/* (same reference lines as ctr_mlp_fwd / ctr_mlp_bwd_data: models/deepfm.py:121-126 and what minimize() differentiates
 * through them.)  TF32 path only: the forward pass of a hidden layer can leave ONE BIT per output element (1 = the stored activation
 * is > 0) next to Y, and the data gradient of the NEXT layer can take those bits as its ReLU mask instead of re-reading
 * the fp32 activation (105 MB per layer at config 4, and the latency of those loads was the data-gradient GEMM's
 * epilogue).  Results are identical to the mask form.  The bit layout is private to the kernels (it follows the
 * epilogue's thread <-> element mapping) and depends on (B, N) only, so a buffer written for a [B, N] activation is
 * read back by the data gradient whose output is that [B, N] matrix; ctr_mlp_relu_bits_bytes(B, N) is its size.
 * ctr_mlp_fwd_bits = ctr_mlp_fwd(precision TF32, relu = 1, bias required) + relu_bits;
 * ctr_mlp_bwd_data_bits = ctr_mlp_bwd_data(precision TF32) with relu_bits (of the [B, K] activation) for the mask. */
int64_t ctr_mlp_relu_bits_bytes(ctr_handle* h, int B, int N);
int ctr_mlp_fwd_bits(ctr_handle* h, ctr_stream s, int B, int K, int N, const float* X, int64_t ldx,
                     const float* Wt, const float* bias, float keep, uint64_t seed, float* Y, int64_t ldy,
                     uint32_t* relu_bits);
int ctr_mlp_bwd_data_bits(ctr_handle* h, ctr_stream s, int B, int K, int N, const float* dZ, int64_t lddz,
                          const float* W, const uint32_t* relu_bits, float mask_scale, float* dX, int64_t lddx);
/* dW[K,N] = X^T @ dZ ; dbias[N] = colsum(dZ).  ws: ctr_mlp_bwd_weight_ws_bytes().
 * x_ones_col != 0: the caller keeps X[:, K] == 1.0 (ldx > K); the TF32 path then gets dbias as row K of the same
 * GEMM instead of a separate pass over dZ (ignored by the FP32 path). */
int64_t ctr_mlp_bwd_weight_ws_bytes(ctr_handle* h, int precision, int B, int K, int N);
int ctr_mlp_bwd_weight(ctr_handle* h, ctr_stream s, int precision, int B, int K, int N,
                       const float* X, int64_t ldx, const float* dZ, int64_t lddz,
                       float* dW, float* dbias, void* ws, int64_t ws_bytes, int x_ones_col);
/* out[N, ld_out] = in[K,N]^T, ld_out >= K: the K-major copy the tcgen05 path reads (TMA needs 16-byte
 * row pitches, so ld_out = K rounded up to 4; pad columns are left untouched) */
int ctr_transpose(ctr_handle* h, ctr_stream s, const float* in, int K, int N, float* out, int64_t ld_out,
                  int round_tf32);
/* out[i] = rna_tf32(in[i]): the rounded weight copy the TF32 data-gradient GEMM reads */
int ctr_round_tf32_copy(ctr_handle* h, ctr_stream s, const float* in, float* out, int64_t n);

/* ---- cross network of DCN -------------------------------------------------------------
 * replaces the loop of test_model/DCN.py:84-86:  xl = tensordot(matmul(x0, xl, transpose_b=True), w_l, axes=1) + b_l + xl,
 * i.e.  x_{l+1} = x0 * (x_l . w_l) + b_l + x_l  (the [B, D, D] outer product is never formed).  One warp per sample;
 * D <= 128, n_layers <= 4 (the reference uses 2).
 *   ctr_cross_fwd   out[b, :D] = x_L
 *   ctr_cross_bwd   gw[l][:] = d loss / d w_l, gb[l][:] = d loss / d b_l for d loss / d x_L [b, :] = dlogit[b] * w_proj[:]
 *                   (written, not accumulated; deterministic two-pass sum over the batch).  ws: ctr_cross_bwd_ws_bytes(). */
typedef struct ctr_cross_cfg {
    int32_t B, D, n_layers, reserved;
    int64_t ldx;                                /* floats per row of x0 */
    const float* w[CTR_MAX_LAYERS];             /* cross_layer_l [D] */
    const float* b[CTR_MAX_LAYERS];             /* cross_bias_l  [D] */
    float* gw[CTR_MAX_LAYERS];                  /* gradients (ctr_cross_bwd only) */
    float* gb[CTR_MAX_LAYERS];
} ctr_cross_cfg;
int ctr_cross_fwd(ctr_handle* h, ctr_stream s, const ctr_cross_cfg* cfg, const float* x0, float* out, int64_t ldo);
int64_t ctr_cross_bwd_ws_bytes(ctr_handle* h, const ctr_cross_cfg* cfg);
int ctr_cross_bwd(ctr_handle* h, ctr_stream s, const ctr_cross_cfg* cfg, const float* x0, const float* dlogit,
                  const float* w_proj, void* ws, int64_t ws_bytes);

/* ---- K5: head + sigmoid log-loss --------------------------------------------------------
 * replaces concat + matmul + sigmoid (models/deepfm.py:128-145; models/dnn.py:82-84; the deep half
 * of models/wdl.py:240-247) and tf.losses.log_loss (models/deepfm.py:151-152, eps = 1e-7).
 *   logit[b] = fm[b,:n_fm].w_fm + deep[b,:H].w_deep + (extra ? extra[b] : 0) + bias[0]
 *   score    = sigmoid(logit)
 *   loss     = mean_b( -y log(score+eps) - (1-y) log(1-score+eps) )        (label != NULL)
 *   dlogit[b]= dloss/dlogit[b]                                              (label != NULL)
 *   dz_last[b,j] = dlogit[b] * w_deep[j] * (deep[b,j] > 0)   (optional: the gradient through the last ReLU,
 *                                                            written while the row is cache-hot)
 *   g_fm[j] = sum_b dlogit[b]*fm[b,j], g_deep[j] = sum_b dlogit[b]*deep[b,j], g_bias = sum_b dlogit[b]   (optional, with
 *             label: the head-weight gradients, accumulated while the rows are in registers; H <= 512, n_fm <= 128)
 * loss_out: one float, written by a second deterministic reduction pass. ws: ctr_head_ws_bytes() */
#define CTR_HEAD_SCE_ON_SCORE 1
#define CTR_HEAD_NO_RELU_GATE 2
int64_t ctr_head_ws_bytes(ctr_handle* h, int B, int n_fm, int H);
int ctr_head_fwd(ctr_handle* h, ctr_stream s, int B,
                 int64_t B_global /* samples the loss is a mean over: B on one GPU, the sum of all ranks' batches when
                 the batch is sharded (loss and dlogit are then this rank's SHARE, so a sum-allreduce gives the global
                 mean and its gradient); <= 0 means B */,
                 const float* fm, int64_t fm_ld, int n_fm, const float* w_fm,
                 const float* deep, int64_t deep_ld, int H, const float* w_deep,
                 const float* extra, const float* bias,
                 const float* label, float* logit, float* score, float* dlogit, float* dz_last, int64_t lddz,
                 int round_tf32 /* round dz_last to tf32 */, float* loss_out,
                 float* g_fm /*[n_fm]*/, float* g_deep /*[H]*/, float* g_bias /*[1]*/, void* ws, int64_t ws_bytes,
                 int flags /* 0: every models/*.py.  CTR_HEAD_SCE_ON_SCORE: the loss of test_model/wdl.py:119-121,
                 tf.nn.sigmoid_cross_entropy_with_logits applied to the ALREADY-sigmoided score.  CTR_HEAD_NO_RELU_GATE:
                 dz_last = dlogit * w_deep without the (deep > 0) gate -- the tower of test_model/DCN.py:66-70 has no
                 activation */);
/* dZ_last[b,j] = dlogit[b] * w_deep[j] * (deep[b,j] > 0)        (through the last ReLU) */
int ctr_head_bwd_deep(ctr_handle* h, ctr_stream s, int B, int H, const float* dlogit, const float* w_deep,
                      const float* deep, int64_t deep_ld, float* dZ, int64_t lddz);
/* out[j] = sum_b (w ? w[b] : 1) * X[b,j]  for j < ncols; deterministic two-pass column reduction
 * (head-weight and bias gradients).  ws: ctr_colsum_ws_bytes() */
int64_t ctr_colsum_ws_bytes(ctr_handle* h, int B, int ncols);
int ctr_colsum(ctr_handle* h, ctr_stream s, const float* X, int64_t ldx, int B, int ncols,
               const float* w, float* out, void* ws, int64_t ws_bytes);
/* out[0] = sum_i f(x[i]);  mode 0: x, 1: x^2, 2: |x|   (regulariser terms, models/deepfm.py:155-156,
 * models/dnn.py:90).  ws >= 4*1024 B */
int ctr_reduce(ctr_handle* h, ctr_stream s, const float* x, int64_t n, int mode, float* out,
               void* ws, int64_t ws_bytes);
/* ctr_reduce without a workspace argument and WITHOUT ANY ALLOCATION.  No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving
 * loops, harnesses that forbid it).  The block partials live in a file-scope `static __device__` array of
 * CTR_REDUCE_STATIC_MAX_PARTIALS floats (4 KB; every GPU that loads the library holds it until the process exits), whose
 * device address is taken once per handle at ctr_create (cudaGetSymbolAddress with that device current).  Same kernels
 * and block slices as ctr_reduce: bit-identical results.  Limits: n <= CTR_REDUCE_STATIC_MAX_N, larger n is rejected
 * with CTR_E_INVALID and a message.  Every call on a device shares the array (it keeps the last call's partials, which
 * are never read: each call overwrites the partials it sums), so calls that OVERLAP on the GPU race -- issue them on one
 * stream or order them with events. */
#define CTR_REDUCE_STATIC_MAX_PARTIALS 1024
#define CTR_REDUCE_STATIC_MAX_N ((int64_t) 1 << 28)   /* 268 M floats (1 GiB): at most 1024 addends per thread, as ctr_reduce's callers have */
int ctr_reduce_static(ctr_handle* h, ctr_stream s, const float* x, int64_t n, int mode, float* out);

/* ---- K4: sparse backward = sort / segment-reduce / Adam ----------------------------------
 * replaces the IndexedSlices gradient of tf.nn.embedding_lookup, TF's unique +
 * unsorted_segment_sum de-duplication and AdamOptimizer._apply_sparse/_apply_dense
 * (all built by `.minimize`, models/deepfm.py:161).
 *
 * ctr_sort_ids: key[i] = shard_key(ids[i] + id_offset) sorted ascending (stable LSD radix sort),
 *   perm[i] = flat position (b*ids_ld + slot) of the i-th smallest.  With n_shards>1 the key is
 *   (id % n_shards) << local_bits | (id / n_shards): owner-major, so one sort also buckets by
 *   owner for the all-to-all.  ids == 0 become key 0xFFFFFFFF when drop_zero (padding: no grad).
 *   ids are read as a [rows, ld] matrix restricted to columns [col0, col0+ncols). */
int64_t ctr_sort_ws_bytes(ctr_handle* h, int64_t n);
int ctr_sort_ids(ctr_handle* h, ctr_stream s, const void* ids, int id_bytes, int64_t rows, int64_t ld,
                 int col0, int ncols, int64_t id_offset, int drop_zero,
                 int64_t extra_base, int n_extra /* appends ids extra_base+c with perm = rows*ld + c: the
                 continuous-field rows, so their gradient rides the same segment-reduce */,
                 int n_shards, int local_bits,
                 int key_bits, int64_t id_limit /* ids (+ id_offset) outside [0, id_limit) become padding keys (no
                 gradient) and set status word CTR_STATUS_BAD_ID; 0 = unchecked */, uint32_t* status /* or NULL */,
                 uint32_t* keys_out, uint32_t* perm_out, void* ws, int64_t ws_bytes);
/* generic stable pair sort (also used by ctr_auc) */
int ctr_sort_pairs(ctr_handle* h, ctr_stream s, uint32_t* keys, uint32_t* vals, int64_t n, int key_bits,
                   void* ws, int64_t ws_bytes);
/* run-length structure of sorted keys: uniq[u], seg_start[u] (seg_start[U] = n), uidx[i] = index of
 * the run containing i, *n_unique = U (device).  Bit-exact against numpy.unique. ws >= 4*(n/2048+2) B.
 * n_dev (device word or NULL): the element count when only the device knows it; n is then the capacity. */
int ctr_unique_segments(ctr_handle* h, ctr_stream s, const uint32_t* sorted_keys, int64_t n,
                        uint32_t* uniq, uint32_t* seg_start, uint32_t* uidx, uint32_t* n_unique,
                        void* ws, int64_t ws_bytes, const uint32_t* n_dev);
/* counts[o] = number of unique keys owned by shard o (keys >> local_bits == o), o < n_shards;
 * keys equal to 0xFFFFFFFF (dropped) are not counted. */
int ctr_shard_counts(ctr_handle* h, ctr_stream s, const uint32_t* uniq, const uint32_t* n_unique,
                     int n_shards, int local_bits, uint32_t* counts);

/* Per-run partial gradients.  The sorted lookup list is cut into fixed chunks of CTR_BWD_CHUNK
 * entries; one sub-warp walks a chunk in order and writes the sum of each run it meets to
 * part[(uidx + chunk) * grad_stride ..] (injective, a row's runs are consecutive).  A gradient row is
 *     [0,k) embedding coordinates | [k] first-order weight | [k+1] D
 * and the gradient of lookup (b, field) is formed on the fly (models/deepfm.py:78-98 differentiated):
 *   single-valued field:  g_v = dX0[b, field*k + :] + use_fm * dlogit[b] * w_second * sum_e[b]        D += dlogit[b]
 *   multi-valued slot:    g_v = scale_e * ( dX0[..] + use_fm * dlogit[b] * w_second * (sum_e[b] - X0[b, field*k + :]) )
 *                         g_w = scale_w * dlogit[b] * w_first[field]       (scale = the saved 1/count; 1 for single-valued)
 * The remaining FM term of a single-valued row, -w_second * e_row * D, is the same for every lookup of the row and is
 * applied by ctr_adam_rows (fm_w2), which reads e_row anyway: the X0 tile is not re-read for single-valued fields.
 * grad_stride: a multiple of 4, >= k + 2.  part must hold (n + n/CTR_BWD_CHUNK + 2) * grad_stride floats. */
#define CTR_BWD_CHUNK 32
int ctr_embed_bwd_partials(ctr_handle* h, ctr_stream s, const ctr_embed_cfg* cfg,
                           const uint32_t* sorted_keys, const uint32_t* perm, const uint32_t* uidx, int64_t n,
                           const int32_t* field_of_slot /*[M] or NULL*/,
                           const float* x0, const float* dx0, int64_t dx0_ld, const float* sum_e, const float* dlogit,
                           const float* w_fm /*[Fc+Fm+C+k] internal order*/,
                           const float* inv_cnt_emb, const float* inv_cnt_first,
                           const float* cont_grad /*[C, grad_stride] or NULL: gradient of the appended
                           continuous-field entries (perm >= B*ids_ld), from ctr_fm_cont_bwd */,
                           float* part);
/* same with per-slot weights of the multi-valued block (slot_weight [B, M] or NULL): the gradient of slot (b, t) is
 * additionally scaled by slot_weight[b, t] -- the backward of the weighted pooling of
 * models/dnn_multi_textline.py:96-102 (sum_t w[b,t] * E[b,t,:] / count). */
int ctr_embed_bwd_partials_weighted(ctr_handle* h, ctr_stream s, const ctr_embed_cfg* cfg,
                                    const uint32_t* sorted_keys, const uint32_t* perm, const uint32_t* uidx, int64_t n,
                                    const int32_t* field_of_slot, const float* x0, const float* dx0, int64_t dx0_ld,
                                    const float* sum_e, const float* dlogit, const float* w_fm,
                                    const float* inv_cnt_emb, const float* inv_cnt_first, const float* cont_grad,
                                    const float* slot_weight, float* part);
/* grads[u, :] = sum of the partials of run u, in chunk order (deterministic) */
int ctr_embed_bwd_finalize(ctr_handle* h, ctr_stream s, const uint32_t* seg_start, const uint32_t* n_unique,
                           int64_t n_max, int width, int grad_stride, const float* part, float* grads);
/* gradient of the C continuous-field rows (models/deepfm.py:63-93 with value = cont):
 *   g_v[c] = sum_b x * dlogit * w_second * (sum_e - x*theta_c),  g_w[c] = sum_b dlogit * w_first_cont[c] * x
 * ws: ctr_cont_bwd_ws_bytes() */
int64_t ctr_cont_bwd_ws_bytes(ctr_handle* h, const ctr_embed_cfg* cfg);
int ctr_fm_cont_bwd(ctr_handle* h, ctr_stream s, const ctr_embed_cfg* cfg, const float* cont,
                    const float* cont_rows, const float* sum_e, const float* dlogit, const float* w_fm,
                    float* cont_grad /*[C, grad_stride]*/, void* ws, int64_t ws_bytes);

/* scalar-table scatter-add, the gradient of ctr_scalar_gather_sum (wide part, models/wdl.py:240-246 as
 * differentiated by models/wdl.py:282):  dst[uniq[u]] += sum_{i in run u} src[perm[i] / ld]   (perm = flat position
 * b*ld + j of a wide lookup, src = dlogit [B]).  Runs come from ctr_sort_ids + ctr_unique_segments; the sum order
 * is fixed.  Keys 0xFFFFFFFF and keys >= dst_rows are skipped. */
int ctr_scalar_segment_add(ctr_handle* h, ctr_stream s, const uint32_t* uniq, const uint32_t* seg_start,
                           const uint32_t* n_unique, int64_t n_max, const uint32_t* perm, const float* src,
                           int ld, int64_t dst_rows, float* dst, void* ws, int64_t ws_bytes);
/* ws of ctr_scalar_segment_add (may be NULL): with it, runs of more than 2048 lookups (the hottest ids of a Zipf-distributed
 * wide field) are cut into pieces summed by one block each.  The caller zeroes the buffer ONCE; every call leaves its
 * counters at zero again.  One buffer per stream that calls the function. */
int64_t ctr_scalar_segment_add_ws_bytes(ctr_handle* h);

/* l2 regulariser over GATHERED scalar weights (models/wdl_textline.py:264-269 overwrites weights["wdl_weights"] with
 * the [B, W+H] lookup result before :284-286 regularises it):  for run u of the sorted wide ids with c = run length,
 *   grad[uniq[u]] += scale * c * w[uniq[u]],  sq[u] = c * w^2   (sq[u] = 0 for u >= *n_unique);
 * then for the dense_n rows from dense_lo (the deep-projection rows, each looked up dense_count = B times):
 *   grad[r] += scale * dense_count * w[r],  sq[n_max + j] = dense_count * w[r]^2.
 * loss term = scale/2 * sum(sq[0 .. n_max + dense_n))  (ctr_reduce mode 0).  Deterministic: one writer per row. */
int ctr_scalar_gathered_l2(ctr_handle* h, ctr_stream s, const uint32_t* uniq, const uint32_t* seg_start,
                           const uint32_t* n_unique, int64_t n_max, const float* w, float scale, int64_t rows,
                           int64_t dense_lo, int dense_n, float dense_count, float* grad, float* sq);

/* owner-side reduction of row gradients received from several peers (multi-GPU):
 *   dst[u, :width] = sum_{i in run u} src[perm[i], :width]   in run order.  The runs come from a STABLE sort of
 * the received row keys, so contributions are added in fixed peer order: deterministic. */
int ctr_segment_sum_rows(ctr_handle* h, ctr_stream s, const uint32_t* seg_start, const uint32_t* n_unique,
                         int64_t n_max, const uint32_t* perm, const float* src, int src_stride, int width,
                         float* dst, int dst_stride);

/* TF1 Adam on the rows named by row_ids (unique):  g = grads[u];
 *   m <- b1*m + (1-b1) g ; v <- b2*v + (1-b2) g^2 ; theta <- theta - lr_t * m / (sqrt(v) + eps)
 * rows with id 0xFFFFFFFF are skipped.  `touched` ([R] bytes or NULL) is set to 1 for every updated
 * row (used by the exact dense sweep).  row = (key & local_mask) when local_bits > 0.
 * fm_w2 ([emb_k] second-order head weights, or NULL): g[j] -= fm_w2[j] * theta[j] * grads[u, emb_k + 1] for j < emb_k
 * before the update (the e-term ctr_embed_bwd_partials leaves out).  emb_k > 0 with width == emb_k (+1) selects the
 * 128-bit row kernel; emb_k = 0: generic width. */
/* c1 = 1 - beta1 and c2 = 1 - beta2 are passed in, rounded ONCE from the exact values: formed on the device as
 * 1.f - 0.999f they would be 0.00099998713 (1.3e-5 low), which scales v by the same factor and -- because lr_t's own
 * sqrt(1 - beta2^t) is computed exactly on the host -- leaves a systematic 6.5e-6 * lr drift per step (TensorFlow's
 * float32 kernel makes the same rounding in BOTH places, where it cancels).  Entry points reject c1 <= 0 or c2 <= 0. */
typedef struct ctr_adam_hp {
    float lr_t, beta1, beta2, eps, c1, c2;
    const uint32_t* guard;   /* device status words (CTR_STATUS_WORDS) or NULL: the update is skipped while one of the
                                first three is non-zero */
    int32_t kind;            /* CTR_OPT_ADAM (0): tf.train.AdamOptimizer, every models/*.py (models/deepfm.py:161).
                                CTR_OPT_FTRL (1): tf.train.FtrlOptimizer with its defaults (test_model/wdl.py:136); slot m
                                holds `accum` (initialise to 0.1), slot v holds `linear`, lr_t is the decayed learning
                                rate; beta1 / beta2 / eps / c1 / c2 are not read.  The row kernels are the SPARSE apply
                                (looked-up rows only), ctr_adam_dense / ctr_dense_update the dense one. */
    int32_t reserved;
} ctr_adam_hp;
#define CTR_OPT_ADAM 0
#define CTR_OPT_FTRL 1
int64_t ctr_adam_rows_ws_bytes(ctr_handle* h, int64_t n_max);
int ctr_adam_rows(ctr_handle* h, ctr_stream s, const uint32_t* row_ids, const uint32_t* n_rows, int64_t n_max,
                  int local_bits, const float* grads, const uint32_t* seg_start /* non-NULL: `grads` is the
                  partials buffer of ctr_embed_bwd_partials and the run sums are formed inline (fused finalize) */,
                  int grad_stride, int width,
                  float* table, float* m, float* v, int table_stride, int64_t table_rows,
                  uint8_t* touched, void* ws, int64_t ws_bytes /* ctr_adam_rows_ws_bytes(); needed when seg_start != NULL:
                  runs spread over many chunks (hot rows) are queued and reduced by one warp each */,
                  const float* fm_w2, int emb_k, const ctr_adam_hp* hp);
/* Owner side of the sharded step in one launch: run u of the (stable-sorted) request list is the gradient rows
 * src[perm[i]], i in [seg_start[u], seg_start[u+1]) -- one per requesting peer, in fixed peer order; they are summed in
 * that order and TF1 Adam is applied to table row (row_ids[u] & mask).  == ctr_segment_sum_rows + ctr_adam_rows bit for
 * bit, without the intermediate [U, stride] buffer.  Embedding-row shapes only (emb_k = 4/8/16/32, width = emb_k [+1]). */
int ctr_adam_rows_gathered(ctr_handle* h, ctr_stream s, const uint32_t* row_ids, const uint32_t* n_rows, int64_t n_max,
                           int local_bits, const float* src, const uint32_t* perm, const uint32_t* seg_start,
                           int src_stride, int width, float* table, float* m, float* v, int table_stride,
                           int64_t table_rows, uint8_t* touched, const float* fm_w2, int emb_k, const ctr_adam_hp* hp);
/* dense TF1 Adam over n floats: g_eff = (g ? g[i] : 0) + l2*theta + l1*sign(theta).
 * skip ([n/skip_group] bytes or NULL): elements whose group flag is 1 are left alone and the flag is
 * cleared (rows already updated by ctr_adam_rows).  This is the exact `tf_dense` mode: TF1 Adam
 * decays m,v and moves EVERY row each step. */
int ctr_adam_dense(ctr_handle* h, ctr_stream s, float* theta, float* m, float* v, const float* g, int64_t n,
                   float l2, float l1, uint8_t* skip, int skip_group, const ctr_adam_hp* hp);

/* The whole dense side of a step in one launch: TF1 Adam over the flat parameter buffer [0, group_end[2]) with the
 * regulariser gradient chosen by position -- [0, group_end[0]) none, [group_end[0], group_end[1]) + reg_scale * theta (l2),
 * [group_end[1], group_end[2]) + reg_scale * sign(theta) (l1: models/dnn.py:90) -- and, for each tower weight
 * W_l [K,N] at w_off[l], the copies the tcgen05 GEMMs read next step: w_rounded[l][K,N] = rna_tf32(W) and
 * w_kmajor[l][N, ld_kmajor] = rna_tf32(W)^T (either may be NULL).  Same arithmetic as ctr_adam_dense + ctr_transpose +
 * ctr_round_tf32_copy. */
typedef struct ctr_dense_layout {
    int64_t group_end[3];
    float reg_scale;
    int32_t n_layers;
    int64_t w_off[CTR_MAX_LAYERS];
    int32_t K[CTR_MAX_LAYERS], N[CTR_MAX_LAYERS];
    int64_t ld_kmajor[CTR_MAX_LAYERS];
    float* w_kmajor[CTR_MAX_LAYERS];
    float* w_rounded[CTR_MAX_LAYERS];
} ctr_dense_layout;
int ctr_dense_update(ctr_handle* h, ctr_stream s, float* theta, float* m, float* v, const float* g,
                     const ctr_dense_layout* lay, const ctr_adam_hp* hp);

/* ---- AUC -------------------------------------------------------------------------------
 * replaces sklearn.metrics.roc_auc_score(labels, scores) (models/deepfm.py:231,253): exact
 * tie-aware rank statistic in integers; *auc_out is a device double (NaN if one class only). */
int64_t ctr_auc_ws_bytes(ctr_handle* h, int64_t n);
int ctr_auc(ctr_handle* h, ctr_stream s, const float* scores, const float* labels, int64_t n,
            double* auc_out, void* ws, int64_t ws_bytes);

/* ---- small utilities ---------------------------------------------------------------------*/
/* dst[i, :width] = src[idx[i], :width] (row gather, e.g. owner-side rows for the all-to-all) */
int ctr_gather_rows(ctr_handle* h, ctr_stream s, const float* src, int src_stride, const uint32_t* idx,
                    const uint32_t* n_rows, int64_t n_max, int idx_mask_bits, int width, float* dst, int dst_stride);
/* inv[perm[i]] = uidx[i] : lookup position -> index of its row in the unique list */
int ctr_invert_perm(ctr_handle* h, ctr_stream s, const uint32_t* perm, const uint32_t* uidx, int64_t n, uint32_t* inv);

/* ---- peer memory over NVLink 5 / NVSwitch (multi-GPU) ------------------------------------------------
 * The reference has no multi-device code (one tf.Session, one device); these entry points are what a row-sharded
 * deployment of models/deepfm.py:78 (embedding_lookup) / :151 (minimize) needs between the shards.  Each rank
 * allocates ONE slab, exports it (CUDA IPC) and maps the slabs of its peers; kernels then store straight into the
 * peer's slab, so "owner gathers rows -> all-to-all -> requester" is ONE kernel whose stores travel over NVLink while
 * its loads are still in flight, and so is "sum the gradient partials of a row -> all-to-all -> owner".
 * Lifetime: alloc/open/close/free are host-synchronous set-up calls (not for the step loop). */
#define CTR_MAX_PEERS 16
#define CTR_IPC_HANDLE_BYTES 64
int ctr_peer_alloc(ctr_handle* h, int64_t bytes, void** slab, void* ipc_handle /*[CTR_IPC_HANDLE_BYTES] out*/);
int ctr_peer_open(ctr_handle* h, const void* ipc_handle, void** slab);
int ctr_peer_close(ctr_handle* h, void* slab);
int ctr_peer_free(ctr_handle* h, void* slab);

/* where the rows of one source list go: the list is a concatenation of `world` segments, segment p = rows
 * [seg_end[p-1], seg_end[p]) -> rows dst_row[p], dst_row[p]+1, ... of the buffer at byte offset dst_off in peer p's slab */
typedef struct ctr_peer_route {
    int32_t  world, rank;
    void*    slab[CTR_MAX_PEERS];      /* peer p's slab as mapped into THIS process (slab[rank] = the own slab) */
    uint32_t seg_end[CTR_MAX_PEERS];
    uint32_t dst_row[CTR_MAX_PEERS];
    int64_t  dst_off;
    const uint32_t* dyn;               /* NULL, or a DEVICE route block of CTR_ROUTE_WORDS words written by
                                          ctr_peer_plan_routes: [0,16) seg_end, [16,32) dst_row, [32] total rows.  When set,
                                          the kernels take the placement (and the row count) from it and the host never
                                          needs to know the split sizes: the sharded step has no host synchronisation */
} ctr_peer_route;
#define CTR_ROUTE_WORDS 48
/* layout of the plan block ctr_peer_plan_routes writes (uint32 words) */
#define CTR_PLAN_VEC    0     /* route of the vectors: source = rows peers requested of this rank (peer-major) */
#define CTR_PLAN_GRAD   48    /* route of the gradients and of the requested keys: source = this rank's unique list */
#define CTR_PLAN_NREQ   96    /* rows requested of this rank, all peers */
#define CTR_PLAN_U      97    /* this rank's unique rows (padding run excluded) */
#define CTR_PLAN_WORDS  128
#define CTR_STATUS_EXCHANGE 3  /* status word: 1 + owner whose receive / request buffer would overflow */

/* Device-side routing.  mat[p * world + o] = unique rows rank p requests of owner o (every rank holds the same matrix
 * once the per-owner counts have been broadcast: ctr_peer_bcast + barrier).  Writes this rank's plan block:
 *   vectors    seg_end = prefix sums of column `rank`, dst_row[p] = sum_{o < rank} mat[p][o]   (p's unique-list offset)
 *   gradients  seg_end = prefix sums of row `rank`,    dst_row[o] = sum_{p < rank} mat[p][o]   (behind the ranks before)
 * -- the layout an all_to_all_single with these split sizes would produce -- plus n_req and U.  If an owner would
 * receive more than grad_cap rows, or this rank more than req_cap requests, status word CTR_STATUS_EXCHANGE is set
 * (every rank sees the same matrix, so every rank reaches the same verdict) and the optimizer kernels stand down. */
int ctr_peer_plan_routes(ctr_handle* h, ctr_stream s, const uint32_t* mat, int world, int rank, int64_t req_cap,
                         int64_t grad_cap, uint32_t* plan, uint32_t* status);
/* buf[row_index[0], :width] = 0 (row index read on the device: the all-zero staging row of the padding ids sits
 * behind the unique rows, whose count only the device knows) */
int ctr_zero_row_at(ctr_handle* h, ctr_stream s, float* buf, int stride, int width, const uint32_t* row_index);

/* all-ranks barrier on `stream` (one 32-thread CTA): releases everything this rank stored into peer slabs before it,
 * and returns once every peer has done the same.  flags_off: byte offset of CTR_MAX_PEERS uint32 arrival words in
 * every slab (zero-initialised once); epoch must grow by 1 per barrier on every rank.  A peer that does not arrive
 * within timeout_ms sets *status (device uint32, own memory) to 1 + the peer's rank instead of hanging the GPU. */
int ctr_peer_barrier(ctr_handle* h, ctr_stream s, const ctr_peer_route* rt, int64_t flags_off, uint32_t epoch,
                     uint32_t* status, int timeout_ms);

/* The three small exchanges of a sharded step, as stores / loads over peer memory instead of NCCL collectives (whose
 * host-side enqueue cost, ~0.3 ms each through torch.distributed, made the multi-GPU step host-bound):
 *   ctr_peer_bcast       the same n_words 32-bit words into EVERY slab at rt->dst_off (replaces the all-gather of the
 *                        per-owner unique counts, SURVEY 8e collective #1a); seg_end / dst_row are not used
 *   ctr_peer_push_rows   row i of `src` ([n_rows, width] 32-bit words) to its routed row of the buffer at rt->dst_off in the
 *                        owning peer's slab (replaces the all-to-all of the de-duplicated ids, #1b: width 1)
 *   ctr_peer_reduce_sum  out[i] = sum_p slab[p][dst_off + 4 i], p in rank order (replaces the dense-gradient allreduce, #4;
 *                        same order on every rank: replicas stay bitwise identical).  n % 4 == 0.
 * All need a ctr_peer_barrier between the writer's store and the reader's load. */
int ctr_peer_bcast(ctr_handle* h, ctr_stream s, const uint32_t* src, int n_words, const ctr_peer_route* rt);
int ctr_peer_push_rows(ctr_handle* h, ctr_stream s, const uint32_t* src, int64_t n_rows, int width,
                       const ctr_peer_route* rt);
int ctr_peer_reduce_sum(ctr_handle* h, ctr_stream s, const ctr_peer_route* rt, int64_t n, float* out);

/* fused owner gather + vector all-to-all: row i of the request list (keys of the rows peers asked this rank for,
 * grouped by peer) is read from the table and stored into the requesting peer's staging buffer:
 *   peer(i).buf[dst_row + (i - seg_begin), :width] = table[idx[i] & mask, :width], columns [width, dst_stride) = 0 */
int ctr_gather_rows_push(ctr_handle* h, ctr_stream s, const float* table, int table_stride, const uint32_t* idx,
                         int64_t n_rows, int idx_mask_bits, int width, int dst_stride, const ctr_peer_route* rt);

/* fused ctr_embed_bwd_finalize + gradient all-to-all: the summed gradient row of run u (u < rt->seg_end[world-1];
 * later runs = the padding key, dropped) is stored into the owner's receive buffer */
int ctr_embed_bwd_finalize_push(ctr_handle* h, ctr_stream s, const uint32_t* seg_start, const uint32_t* n_unique,
                                int64_t n_max, int width, int grad_stride, const float* part,
                                const ctr_peer_route* rt);

/* owner side of the gradient exchange: `keys` is a concatenation of `world` segments (peer-major, segment p ends at
 * seg_end[p], a HOST array), each ascending and duplicate-free (a slice of the requester's unique list).  Writes their
 * stable merge -- out_keys ascending, equal keys in peer order, out_idx[i] = position of that entry in `keys` -- by
 * ranking every entry with one binary search per segment: one launch instead of a radix sort of the list.
 * seg_end_dev (device, world words) replaces seg_end and n (then only a capacity for the launch) when non-NULL. */
int ctr_merge_sorted_segments(ctr_handle* h, ctr_stream s, const uint32_t* keys, int64_t n, const uint32_t* seg_end,
                              int world, uint32_t* out_keys, uint32_t* out_idx, const uint32_t* seg_end_dev);

/* ---- timing experiments (not part of the product path) -----------------------------------------------------
 * with CTR_GEMM_DBG & 8 in the environment the TF32 forward/dgrad GEMM records %globaltimer at its phase boundaries
 * (8 slots per CTA: start, setup done, last MMA committed, accumulator ready, epilogue done, CTA synced, TMEM freed);
 * this copies the last launch's records to the host. */
int ctr_debug_gemm_trace(unsigned long long* host, int n_ctas);

/* ---- host-side input decoding ---------------------------------------------
 * No handle, no stream, no device work: bytes in host memory -> row-major arrays in caller-owned (pinned) host
 * memory.  The caller then issues one cudaMemcpyAsync per tensor.
 *
 * ctr_parse_lines replaces the per-line Python of utils/data_loader_load.py:94-139 (load_process), :142-184
 * (load_test) and the tf.data graphs of utils/data_loader_load.py:187-243 / utils/data_loader_textline.py:4-36.
 * Line format: `label,c c c,id id id[,third]` -- columns split on ',', values on ONE space (str.split(" ")).
 *   n_cont     floats in column 1 (the reference's cont_field_size: continuous values followed by the vector block)
 *   n_cate     integer ids in column 2
 *   third_kind 0: no column 3;  1: n_third integer ids (wide_feats, :125-127);  2: n_third `id=value` pairs
 *              (get_libsvm_value, :12-21) -> host_third_ids + host_third_vals
 *   id_bytes   4 or 8: element size of host_cate / host_third_ids
 * Floats are converted like Python's float() followed by a cast to float32.  Only complete lines (terminated by a
 * newline) are consumed, at most max_rows of them; blank lines are skipped.  *host_consumed = bytes to drop from the
 * front of the buffer before the next call.  n_threads <= 0: all host threads.
 * Returns CTR_E_INVALID on a malformed line, with *host_bad_line = its index among the non-blank lines. */
int ctr_parse_lines(const char* host_buf, int64_t len, int n_cont, int n_cate, int third_kind, int n_third,
                    int id_bytes, int64_t max_rows, int n_threads, float* host_label, float* host_cont,
                    void* host_cate, void* host_third_ids, float* host_third_vals, int64_t* host_n_rows,
                    int64_t* host_consumed, int64_t* host_bad_line);

/* TFRecord framing + tf.Example decoding for the FixedLenFeature spec of utils/data_loader.py:7-24:
 *   label [1] f32, cont_feats [n_cont] f32 (n_cont = 0: key not in the spec, the *_cate generation), vector_feats
 *   [n_vec] f32, cate_feats [n_cate] i64.  Unknown keys are ignored, a missing key is an error (no defaults in the
 *   reference spec).  verify_crc != 0 checks both masked CRC-32C words of every record.
 * Only complete records are consumed, at most max_rows. */
int ctr_parse_tfrecords(const char* host_buf, int64_t len, int n_cont, int n_vec, int n_cate, int verify_crc,
                        int64_t max_rows, float* host_label, float* host_cont, float* host_vec,
                        int64_t* host_cate, int64_t* host_n_rows, int64_t* host_consumed);

#ifdef __cplusplus
}
#endif
#endif /* CTR_B200_H */
