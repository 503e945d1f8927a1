/*
 * obst_b200.h -- C-ABI of the B200-native OBST mixer training-step hot path.
 *
 * The reference (HomebrewNLP/HomebrewNLP-MTF) has no FFI: its "operator API" is the Python
 * registries LAYER_FUNCTIONS (src/model/frontend.py:58-75) and OPTIMIZERS
 * (src/optimizer/optimizers.py:131-151); every FLOP is a TF/XLA kernel picked by Mesh-TensorFlow.
 * This header is what a binding for that path would bind instead: one entry point per fused device
 * op, each citing the reference lines whose arithmetic it replaces.
 *
 * Conventions (all entry points):
 *   - plain pointers and sizes only; every pointer is a DEVICE pointer unless named host_* (descriptor
 *     structs passed by pointer -- obst_gemm_desc, obst_opt_stage_args -- are read on the host);
 *   - `stream` is a cudaStream_t passed as void*; work is enqueued, never synchronised;
 *   - nothing is allocated or freed (the caller owns every buffer), except a host-side cache of
 *     TMA tensor maps inside obst_gemm (host memory, keyed by operand geometry);
 *   - return 0 on success, a negative OBST_E* code otherwise; obst_last_error() gives the text;
 *   - dtype codes: OBST_F32 = 0, OBST_BF16 = 1.  "AT" below = the activation dtype
 *     (calculation_dtype, src/dataclass.py:253-255).
 */
#ifndef OBST_B200_H
#define OBST_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define OBST_F32 0
#define OBST_BF16 1

#define OBST_OK 0
#define OBST_EINVAL (-1)   /* bad argument / unsupported shape */
#define OBST_ECUDA (-2)    /* CUDA runtime or driver error at launch */
#define OBST_ENOTSUP (-3)  /* requested backend cannot run this problem */

/* ---- library plumbing (no reference counterpart) ------------------------------------------ */
int obst_version(void);
const char* obst_last_error(void);
/* number of kernels this library has launched since load / since the last reset (bench.py's
 * gpu_launches). */
int64_t obst_launch_count(void);
void obst_launch_count_reset(void);
/* How many calls since load were routed to a specialised kernel: lets a test (or a user) confirm that the shape it
 * runs takes the kernel it means to measure, not the generic fallback.  Returns -1 for an unknown id. */
enum {
  OBST_PATH_GEMM_TC = 0,    /* obst_gemm -> tcgen05 / TMEM / TMA kernel            */
  OBST_PATH_GEMM_SIMT,      /* obst_gemm -> fp32-FMA kernel                        */
  OBST_PATH_GN_FWD_RING,    /* obst_gn_fwd -> streaming kernel (gn_fwd_ring.cu)    */
  OBST_PATH_GN_BWD_RING,    /* obst_gn_bwd -> streaming kernel (gn_bwd_ring.cu)    */
  OBST_PATH_OPT_RING,       /* obst_opt_update -> streaming kernel (optimizer_ring.cu) */
  OBST_PATH_COUNT
};
int64_t obst_path_count(int32_t which);

/* ---- host-side checksum --------------------------------------------------------------------------
 * CRC-32C (Castagnoli) of HOST memory, continuing from `crc` (0 to start): the checksum of TFRecord frames
 * (tf.data.TFRecordDataset behind src/inputs.py:231-268) and of TensorBundle checkpoint tensors and table blocks
 * (tf.train.Saver, src/run/run.py:158-175).  SSE4.2 crc32 instruction where the host has it, slicing-by-8 tables
 * otherwise (obst_crc32c_portable: always the tables).  No device work, no stream. */
uint32_t obst_crc32c(const void* host_data, int64_t nbytes, uint32_t crc);
uint32_t obst_crc32c_portable(const void* host_data, int64_t nbytes, uint32_t crc);

/* ---- GEMM family -------------------------------------------------------------------------------
 * One descriptor covers every contraction on the path:
 *   dense  W_in / W_mid (+ReLU)            src/model/basic.py:33-34,47-57; src/model/backend.py:108-110
 *   grouped W_out (per head)               src/model/basic.py:122-126
 *   causal token mixing (lower-triangular) src/model/spatial.py:19-23,73,81; src/utils_mtf.py:411-415
 *   input linear / LM head                 src/model/__init__.py:82,144-145
 *   all dgrad / wgrad einsums              mtf.einsum gradients driven by src/optimizer/__init__.py:151-174
 *                                          and src/model/revnet.py:91-120
 *
 * Semantics, for g in [0, GLO*GHI) with g_lo = g % GLO, g_hi = g / GLO:
 *   acc[m,n] = sum_{ko<KO} sum_{k<KI} A(g,m,ko,k) * B(g,n,ko,k)       (fp32 accumulate)
 *   X(g,r,ko,k) = X.ptr[g_lo*s_glo + g_hi*s_ghi + r*s_mn + ko*s_ko + k*s_k]
 *   tri: 0 none
 *        1 promise: A(g,m,k) == 0 for k > m, the kernel may skip those products (KO must be 1)
 *        2 promise: A(g,m,k) == 0 for k < m, likewise
 *        3 only outputs with n <= m are accumulated (requires accumulate != 0)
 *   epi: 0 D=acc | 1 D=relu(acc) | 2 D=R+acc | 3 D=R-acc | 4 D=acc*[R>0]
 *        5/6 (tcgen05 backend, bf16 out, GHI == 1): a group norm of the A operand folded into this GEMM --
 *            D = R +/- ( rstd[m,g]*acc - rstd[m,g]*mean[m,g]*c[g,n] + d[g,n] )
 *            aux0 = float2 (mean, rstd) per (m, g_lo) [M, GLO]; aux1 = c, aux2 = d: fp32 [GLO, N]
 *            (norm -> linear, src/model/basic.py:55-56,122-126, evaluated as r*(h@(gamma.W)) - r*mu*(gamma^T W)
 *            + beta^T W: B holds gamma.W, c = gamma^T W, d = beta^T W -- see obst_fold_norm_weight)
 *        7 (tcgen05 backend, bf16 out, one group, N % 256 == 0): D = relu(acc), and aux0 (float2
 *            [M, N/128]) receives per row and 128-column slice the (sum, sum of squares) of the stored values:
 *            entry [m, 2*(n/256) + w] covers columns 256*(n/256) + 64*(w + 2i) .. +63, i = 0, 1
 *        8 (tcgen05 backend, accumulate): D += aux0[g_lo, m] * acc, aux0 fp32 [GLO, M].  With aux1 (bf16, D's
 *            layout) and aux2 (fp32 [GLO, M]) set, also aux2[g_lo, m] += sum_n aux1[g_lo, m, n] * acc[m, n]
 *            (atomics): the gamma gradient of a folded group norm, d gamma_j = sum_k W_jk (G_jk - ...), taken
 *            where the weight gradient's accumulator G is (src/model/normalization.py:22-34 differentiated)
 *        9/10 = 5/6, plus two row dot products of the accumulator's tile with a second tile aux3 (bf16, D's
 *            layout: the gradient that will reach this output): aux4 (float2 [M, GLO, 2 * ceil(N/256)]) receives
 *            ( sum_n aux3*c , sum_n aux3*acc ) per row, group, 256-column tile and epilogue warp pair; their
 *            sums are the two reductions of the folded norm's backward (obst_norm_bwd_coeffs)
 *        11 (tcgen05 backend, bf16 out, GHI == 1): the folded norm's backward applied to this GEMM's output:
 *            D = [R > 0] * ( a[m,g] * acc + b[m,g] * R + c[m,g] ) with acc = dy @ (gamma . W)^T = gamma * dn (B is
 *            obst_fold_item.wg), R = the norm's input h (a ReLU output: the mask is its own derivative),
 *            aux0 = float4 (a, b, c, -) [M, GLO] from obst_norm_bwd_coeffs
 *        12 (tcgen05 backend, bf16 out, one group, N == 256 = the whole vocabulary in one tile, M > 128): the LM head
 *            with the softmax cross-entropy in its epilogue (src/model/__init__.py:144-145 -> src/mtf_wrapper.py:64-71,
 *            accuracy src/model/__init__.py:185-187): the logits are never written.  Per row m, on the accumulator
 *            rounded to bf16 (what obst_xent_fwd_bwd would have read back): logZ = log sum_n exp(l_n),
 *            D[m, n] = (softmax_n - [n == t_m]) * scalar0 + 2 * scalar1 * scalar0 * logZ * softmax_n  (the gradient of
 *            scalar0 * sum_m (logZ_m - l_{m,t_m} + scalar1 * logZ_m^2), as obst_xent_fwd_bwd with inv_n = scalar0,
 *            z_loss = scalar1), aux0 = int32 targets [M] (clamped to the vocabulary),
 *            aux4 = double[3] that receives += (sum nll, sum logZ^2, number of rows whose first arg-max is t_m)
 *   accumulate!=0: D (fp32 only) += value with atomics (split-K allowed; tri==3 masks the store)
 *   D/R address: ptr[g_lo*d_glo + g_hi*d_ghi + m*d_ld + n]; R may alias D.
 * backend: 0 auto (tcgen05 when dtype_in is bf16 and the geometry is TMA-legal, else SIMT),
 *          1 SIMT fp32-FMA kernel (any strides, fp32 or bf16 inputs), 2 tcgen05/TMEM/TMA kernel.
 */
typedef struct obst_operand {
  const void* ptr;
  int64_t s_mn;  /* element stride along M (operand A) or N (operand B) */
  int64_t s_k;   /* element stride along the inner reduction index k    */
  int64_t s_ko;  /* element stride along the outer reduction index ko   */
  int64_t s_glo; /* element stride along g_lo (0 = broadcast)           */
  int64_t s_ghi; /* element stride along g_hi (0 = broadcast)           */
} obst_operand;

typedef struct obst_gemm_desc {
  int32_t dtype_in;  /* OBST_F32 | OBST_BF16 (A and B)          */
  int32_t dtype_out; /* OBST_F32 | OBST_BF16 (D and R)          */
  int32_t M, N, KI, KO;
  int32_t GLO, GHI;
  obst_operand A, B;
  void* D;
  const void* R;
  int64_t d_ld, d_glo, d_ghi;
  int32_t tri;
  int32_t epi;
  int32_t accumulate;
  int32_t splits; /* split-K factor for accumulate mode; 0 = choose */
  const void* aux0; /* epilogues 5..12, else NULL */
  const void* aux1;
  const void* aux2;
  const void* aux3;
  void* aux4;
  float scalar0; /* epilogue 12 */
  float scalar1;
} obst_gemm_desc;

int obst_gemm(const obst_gemm_desc* desc, int backend, void* stream);

/* ---- group norm (per token, per head), src/model/normalization.py:22-34 ------------------------
 * x, y: [rows, width] contiguous AT, rows = tokens * heads (row r belongs to head r % heads);
 * gamma/beta: [heads, width] AT (may be NULL = no
 * scale / no shift).  y = (x-mean)*rsqrt(mean((x-mean)^2)+1e-5)*gamma + beta.
 * act: 0 none | 1 OBST gelu applied to y (src/model/activation.py:158-161; fuses the
 * "norm -> activation-gelu" layer pair of block_config[1], configs/32big_mixer.json:13-19).
 * stats: NULL, or fp32 [rows, 2] receiving (mean, rstd) of every normalised row for the
 * backward pass (8 bytes per row of `width` elements). */
int obst_gn_fwd(int dtype, const void* x, const void* gamma, const void* beta, void* y,
                float* stats, int64_t rows, int32_t heads, int32_t width, int32_t act, void* stream);

/* Backward of the above (full LayerNorm gradient).  dy: grad of the output
 * (after the activation when act=1); dx written (dx_accumulate=0) or added to (1);
 * dgamma/dbeta: fp32 [heads,width], accumulated with atomics.
 * relu_mask: 1 = x is the output of a ReLU (W_mid path, src/model/basic.py:50,55-56): the
 * gradient is additionally multiplied by [x > 0] so the caller gets d(pre-ReLU).
 * stats: what obst_gn_fwd wrote for the same x, or NULL (the statistics are then recomputed). */
int obst_gn_bwd(int dtype, const void* dy, const void* x, const void* gamma, const void* beta,
                void* dx, float* dgamma, float* dbeta, const float* stats, int64_t rows,
                int32_t heads, int32_t width, int32_t act, int32_t relu_mask, int32_t dx_accumulate,
                void* stream);

/* ---- group norm folded into the neighbouring GEMMs (bf16 activations) -------------------------------
 * bottleneck_group_linear's mid/out pair (src/model/basic.py:47-57,122-126): out = norm(h) @ W per head with
 * h = relu(h1 @ W_mid).  With n = (h - mu) r gamma + beta:
 *   n @ W = r (h @ (gamma.W)) - r mu (gamma^T W) + beta^T W,
 *   dW[j,k] = gamma_j ( sum_t h[t,j] r_t dy[t,k] - sum_t r_t mu_t dy[t,k] ) + beta_j sum_t dy[t,k],
 * so the normalised tensor is never written: obst_gemm epilogue 7 emits the statistics with h, epilogues 5/6
 * apply the norm to W_out's accumulator, epilogue 8 scales the weight gradient's rows, and:
 *   obst_norm_stats_finalize  partials float2 [rows, heads, per_head] (epilogue 7's layout with rows of
 *                             heads * per_head slices of 128 columns; width = 128 * per_head) -> stats float2
 *                             (mean, rstd) [rows, heads], the array obst_gn_bwd accepts
 *   obst_fold_norm_weight     for each item of a DEVICE table (all weights of one shape in one launch, once per
 *                             weight update): w AT [heads, kin, kout], gamma / beta AT [heads, kin] (NULL = 1 / 0)
 *                             -> wg_t AT [heads, kout, kin] = (gamma.W)^T, cvec = gamma^T W and dvec = beta^T W
 *                             fp32 [heads, kout], gamma_f fp32 [heads, kin], wg AT [heads, kin, kout] = gamma.W
 *   obst_rowscale_colsum      dyr = rstd . dy (AT [tokens, heads, width]), v1[h,:] += sum_t rstd mean dy,
 *                             v2[h,:] += sum_t dy (fp32 [heads, width], atomics; width 256 or 512)
 *   obst_norm_wgrad_rank1     dw[h,j,k] += -gamma[h,j] v1[h,k] + beta[h,j] v2[h,k]   (dw fp32 [heads, kin, kout])
 * AT is bfloat16 here. */
int obst_norm_stats_finalize(const float* partials, float* stats, int64_t rows, int32_t heads, int32_t per_head,
                             int32_t width, void* stream);
typedef struct obst_fold_item {
  const void* w;
  const void* gamma;
  const void* beta;
  void* wg_t;
  float* cvec;
  float* dvec;
  float* gamma_f;
  void* wg; /* AT [heads, kin, kout] = gamma . W in W's own layout (the dgrad's B operand for epilogue 11), or NULL */
} obst_fold_item;
int obst_fold_norm_weight(const obst_fold_item* items, int32_t nitems, int32_t heads, int32_t kin, int32_t kout,
                          void* stream);
int obst_rowscale_colsum(const void* dy, const float* stats, void* dyr, float* v1, float* v2, int64_t tokens,
                         int32_t heads, int32_t width, void* stream);
int obst_norm_wgrad_rank1(float* dw, const void* gamma, const void* beta, const float* v1, const float* v2,
                          int32_t heads, int32_t kin, int32_t kout, void* stream);
/* The norm's own backward folded into the GEMMs as well (src/model/normalization.py:22-34 differentiated; used by
 * the reversible backward, where the gradient dy of the block's output is known while W_out is recomputed):
 *   obst_norm_bwd_coeffs    partials float2 [rows, heads, nparts] from obst_gemm epilogues 9 / 10 + stats float2
 *                           (mean, rstd) [rows, heads] -> coeffs float4 (a, b, c, 0) [rows, heads] for epilogue 11:
 *                           dh = [h > 0] (a gamma dn + b h + c), gamma dn = dy @ (gamma . W)^T the dgrad accumulator
 *   obst_norm_affine_grad   dgamma[h,j] -= sum_k W[h,j,k] v1[h,k]; dbeta[h,j] += sum_k W[h,j,k] v2[h,k]
 *                           (w AT [heads, kin, kout]; v1, v2 as obst_rowscale_colsum leaves them; dgamma / dbeta fp32
 *                           [heads, kin], either may be NULL) -- together with epilogue 8's row dots (aux1 = W,
 *                           aux2 = dgamma) this is sum_t dn xhat and sum_t dn without dn ever being stored */
int obst_norm_bwd_coeffs(const float* partials, const float* stats, float* coeffs, int64_t rows, int32_t heads,
                         int32_t nparts, int32_t width, void* stream);
int obst_norm_affine_grad(const void* w, const float* v1, const float* v2, float* dgamma, float* dbeta, int32_t heads,
                          int32_t kin, int32_t kout, void* stream);

/* ---- activations, src/model/activation.py:158-169,201-211 ---------------------------------------
 * kind: 0 relu | 1 OBST gelu 0.5x(1+tanh(0.044715x^3+sqrt(2/pi)x)) | 2.. the other entries of ACTIVATIONS
 * (activation.py:13-155,168-200).  silu and mish use the reference's own backward operations, which are not
 * the analytic derivatives: silu' := (x-1)*sigmoid(x)+1, mish' := t + 1 - t^2*x*sigmoid(x), t = tanh(softplus x);
 * mtf_mish is differentiated analytically (MTF autodiff). */
enum {
  OBST_ACT_RELU = 0, OBST_ACT_GELU, OBST_ACT_SIGMOID, OBST_ACT_TANH, OBST_ACT_LECUN_TANH, OBST_ACT_SILU,
  OBST_ACT_MISH, OBST_ACT_MTF_MISH, OBST_ACT_SOFTSIGN, OBST_ACT_EXP
};
int obst_act_fwd(int dtype, int32_t kind, const void* x, void* y, int64_t n, void* stream);
int obst_act_bwd(int dtype, int32_t kind, const void* x, const void* dy, void* dx, int64_t n,
                 void* stream);
/* GLU gate of activated_linear (src/model/basic.py:51-54): out = a * sigmoid(b);
 * backward: da = dy * sigmoid(b), db = dy * a * sigmoid(b) * (1 - sigmoid(b)). */
int obst_gate_fwd(int dtype, const void* a, const void* b, void* out, int64_t n, void* stream);
int obst_gate_bwd(int dtype, const void* a, const void* b, const void* dy, void* da, void* db,
                  int64_t n, void* stream);

/* ---- running sums along the sequence axis, src/model/spatial.py:26-39 ------------------------------
 * x, out: AT [outer, seq, inner], inner contiguous (a multiple of 8 bf16 / 4 fp32), x != out.
 *   reverse == 0: out[o,s,:] = sum_{t<=s} x[o,t,:]   (cumsum; tf.cumsum)
 *   reverse != 0: out[o,s,:] = sum_{t>=s} x[o,t,:]   (its gradient, _cumsum_grad: tf.cumsum(reverse=True))
 *   scale 1: the sum is divided by (s + 1) on the way out   (cummean = cumsum / (1 + range), spatial.py:37-39)
 *   scale 2: x[o,t,:] is divided by (t + 1) on the way in    (cummean's gradient, with reverse != 0)
 * The running sum is kept in fp32. */
int obst_cumsum(int dtype, const void* x, void* out, int64_t outer, int32_t seq, int64_t inner,
                int32_t reverse, int32_t scale, void* stream);

/* ---- registered layers off the mixer hot path ---------------------------------
 * sum_heads (src/model/basic.py:77-78): out[r,w] = sum_h x[r,h,w]; its gradient broadcasts back. */
int obst_sum_heads(int dtype, const void* x, void* out, int64_t rows, int32_t heads, int32_t width, void* stream);
int obst_broadcast_heads(int dtype, const void* dy, void* dx, int64_t rows, int32_t heads, int32_t width,
                         void* stream);
/* transpose_sequence_features (src/model/basic.py:81-86): x [outer, a, mid, b] -> out [outer, b, mid, a] with
 * out[o,j,m,i] = x[o,i,m,j] (x != out).  The gradient is the same call with a and b exchanged. */
int obst_swap_axes(int dtype, const void* x, void* out, int64_t outer, int32_t a, int32_t mid, int32_t b,
                   void* stream);
/* element-wise product (split_path's `multiply`, src/model/frontend.py:39-55); out may alias a or b. */
int obst_mul(int dtype, const void* a, const void* b, void* out, int64_t n, void* stream);
/* rezero (src/model/basic.py:21-22): out = x * scalar[0] (scalar: AT, one element in device memory); the scalar's
 * gradient is obst_dot_accumulate(x, dy): out[0] += sum_i a[i] * b[i] in fp32. */
int obst_scale(int dtype, const void* x, const void* scalar, void* out, int64_t n, void* stream);
int obst_dot_accumulate(int dtype, const void* a, const void* b, float* out, int64_t n, void* stream);
/* dropout (src/model/basic.py:25-30 -> mtf.dropout): out = x * [u_i < keep] / keep with u_i a counter-based
 * uniform of (seed, i); calling it on dy with the same seed applies the same mask (the backward). */
int obst_dropout(int dtype, const void* x, void* out, int64_t n, float keep, uint64_t seed, void* stream);

/* ---- dot-product attention's soft-max stage, src/model/spatial.py:42-81 ------------------------------
 * logits, soft, probs: AT [batch, seq, heads, keys] (keys contiguous; the q.k^T, p.v contractions and their
 * gradients are obst_gemm calls).  Per (batch, position s, head) row over key positions t:
 *   z_t = scale * logit_t + bias[h,s,t];  z_t = -inf for t > s (the reference always adds the causal -2e38, :71);
 *   soft = softmax(z);  p_t = (soft_t + add_map[h,s,t] * m) * (mul_map[h,s,t] * m),  m = [s >= t] if masked else 1
 * bias / add_map / mul_map: AT [heads, seq, keys] or NULL ('biased_softmax', 'biased_attention_map',
 * 'scale_attention_map': the learned maps of _masked_map, :19-23).  soft is always written (the backward needs
 * it); probs only when a map follows the soft-max (otherwise p == soft; pass NULL).
 * Backward: dprobs -> dlogits (= scale * dz; may alias dprobs), map gradients (fp32 [heads, seq, keys], atomics,
 * NULL to skip). */
int obst_attn_softmax_fwd(int dtype, const void* logits, const void* bias, const void* add_map, const void* mul_map,
                          void* soft, void* probs, int64_t batch, int32_t seq, int32_t heads, int32_t keys, float scale,
                          int32_t masked, void* stream);
int obst_attn_softmax_bwd(int dtype, const void* dprobs, const void* soft, const void* add_map, const void* mul_map,
                          void* dlogits, float* dbias, float* dadd, float* dmul, int64_t batch, int32_t seq,
                          int32_t heads, int32_t keys, float scale, int32_t masked, void* stream);
/* a position embedding added to every batch row (key += embed(...), :57-58): out[o,i] = x[o,i] + bias[i]; its
 * gradient out[i] += sum_o x[o,i] (fp32). */
int obst_bias_add(int dtype, const void* x, const void* bias, void* out, int64_t outer, int64_t n, void* stream);
int obst_colsum_accumulate(int dtype, const void* x, float* out, int64_t outer, int64_t n, void* stream);

/* ---- residual stream arithmetic, src/model/revnet.py:23,89,120; src/model/__init__.py:129 ------
 * out = a + sign*b (sign = +1 or -1), all AT, n elements; out may alias a or b. */
int obst_axpby(int dtype, const void* a, const void* b, float sign, void* out, int64_t n,
               void* stream);

/* ---- embedding gather / scatter-add, src/model/embedding.py:39-125 ------------------------------
 * gather: out[t,:] = table[idx[t],:]  (table AT [vocab,width]);
 * scatter: dtable[idx[t],:] += dout[t,:]  (dtable fp32 [vocab,width], atomics). */
int obst_embed_gather(int dtype, const int32_t* idx, const void* table, void* out, int64_t tokens,
                      int32_t vocab, int32_t width, void* stream);
int obst_embed_scatter_add(int dtype, const int32_t* idx, const void* dout, float* dtable,
                           int64_t tokens, int32_t vocab, int32_t width, void* stream);

/* ---- softmax cross-entropy + z-loss + accuracy, forward and backward in one pass ---------------
 * src/mtf_wrapper.py:64-71, src/model/__init__.py:185-187.
 * logits AT [tokens,vocab]; targets int32 [tokens]; inv_n = 1/(global target count).
 * dlogits (AT, may alias logits) = (softmax-onehot)*inv_n + 2*z_loss*inv_n*logZ*softmax.
 * sums (double[3], accumulated with atomics): {sum(logZ-logit[tgt]), sum(logZ^2), #argmax==tgt}. */
int obst_xent_fwd_bwd(int dtype, const void* logits, const int32_t* targets, void* dlogits,
                      double* sums, int64_t tokens, int32_t vocab, float inv_n, float z_loss,
                      void* stream);

/* ---- parameter shadow copies --------------------------------------------------------------------
 * The graph reads every variable through a cast to the activation dtype
 * (mtf.Variable.value).  dst(AT) = cast(src fp32); tril_s > 0 additionally
 * zeroes entries with col > row of each trailing [tril_s, tril_s] matrix (the causal mask of
 * src/model/spatial.py:19-23 folded into the weight the kernels see). */
int obst_cast_param(int dtype_dst, const float* src, void* dst, int64_t n, int32_t tril_s,
                    void* stream);

/* dst[b, c, r] = src[b, r, c] for `batch` row-major [rows, cols] matrices of dtype AT (src != dst).
 * The reference's einsums read one weight in whatever orientation the contraction needs
 * (src/model/backend.py:108-110); the tcgen05 GEMM prefers its B operand K-major, so the host keeps a
 * transposed shadow of each weight that is a forward B operand and refreshes it after the update. */
int obst_transpose_param(int dtype, const void* src, void* dst, int64_t batch, int32_t rows,
                         int32_t cols, void* stream);

/* ---- fused optimizer update ---------------------------------------------------------------------
 * adaptive_clip:<clip> - sm3 - momentum:<beta>:1:<nesterov> - learning_rate, plus the
 * weight-decay rule and assign_sub:  src/optimizer/__init__.py:31-66,
 * src/optimizer/optimizers.py:13-14,60-84,114-128, src/optimizer/backend.py:23-25.
 * The arithmetic is the reference's op sequence evaluated op by op in fp32 (IEEE sqrt / reciprocal, no
 * fused multiply-add), the two norms in float64: for identical gradients the slots and the master
 * weights are bit-identical to such an evaluation (oracle/obst_oracle.py: optimizer_step).
 *
 * One obst_opt_tensor per variable (table lives in DEVICE memory, built by the caller once):
 *   p      fp32 master slice (updated in place)
 *   shadow AT copy refreshed with the new value (may be NULL); tril_s as in obst_cast_param
 *   g      the gradient handed to the optimizer (post-allreduce): fp32, or -- g_bf16 != 0 -- bfloat16,
 *          the dtype the reference hands over when calculation_dtype is bfloat16
 *          (src/optimizer/__init__.py:180).  With an fp32 g, grad_bf16 != 0 rounds it to bf16 on read.
 *   g_acc  fp32 buffer the backward pass accumulates into when g is bf16 (obst_grad_finalize rounds it
 *          into g), else NULL
 *   m      momentum state, state dtype (optimizer_slice_dtype)
 *   acc[i] SM3 accumulator of dim i (state dtype), length dims[i]; ndim in 1..4
 *   acc_new[i] fp32 staging for the new accumulator (scratch, length dims[i])
 * Launches: [grad_finalize ->] norms -> update (+ new-accumulator max) -> finalize accumulators.
 */
typedef struct obst_opt_tensor {
  float* p;
  void* shadow;
  void* g;
  void* m;
  void* acc[4];
  float* acc_new[4];
  float* g_acc;
  int64_t numel;
  int32_t dims[4];
  int32_t ndim;
  int32_t decay;   /* 1 = "large tensor" per src/optimizer/__init__.py:49-61 */
  int32_t tril_s;
  int32_t g_bf16;  /* 1 = g points to bfloat16 */
} obst_opt_tensor;

/* Work list (DEVICE arrays, built by the caller once): item i covers rows
 * [item_row0[i], item_row0[i]+item_rows[i]) of tensor item_tensor[i] viewed as [numel/cols, cols]
 * with cols = dims[ndim-1]; an item covers at most 2048 rows.  A call may pass any sub-range of the
 * list (e.g. the items of one depth).
 * norms: double[2*ntensors] scratch: {sum g^2, sum cast_AT(p)^2} per tensor.  Entries
 * [first_tensor, first_tensor+tensor_count) are cleared first unless flags&2.
 * flags&1 ("carry"): sum cast_AT(p)^2 is NOT recomputed (4 B/param less traffic) but moved from
 * wnext[t], where the previous obst_opt_update accumulated it while writing p; wnext[t] is reset.
 * flags&4: sum g^2 is already in norms (obst_grad_finalize put it there): only the weight half is
 * produced (with flags&1 no gradient or weight is read at all).
 * flags&8: the weight half is taken from the shadows (macro-batching chain, see obst_opt_update's preview). */
int obst_opt_norms(const obst_opt_tensor* table, int32_t ntensors, const int32_t* item_tensor,
                   const int32_t* item_row0, const int32_t* item_rows, int32_t nitems,
                   int32_t dtype_act, int32_t grad_bf16, double* norms, int32_t flags, double* wnext,
                   int32_t first_tensor, int32_t tensor_count, void* stream);
/* The hand-over of src/optimizer/__init__.py:180 for tensors with a bf16 g: g = bf16(g_acc) for every
 * item passed (tensors whose g_acc is NULL are skipped).  flags&1: norms[2t] += sum g^2 of the rounded
 * values (the caller clears norms at the start of a step; skip this when the gradient still has to be
 * all-reduced, and let obst_opt_norms take the sum afterwards). */
int obst_grad_finalize(const obst_opt_tensor* table, int32_t ntensors, const int32_t* item_tensor,
                       const int32_t* item_row0, const int32_t* item_rows, int32_t nitems,
                       double* norms, int32_t flags, void* stream);
/* reads norms, old accumulators and m; writes p, shadow, m and max-reduces the new SM3
 * accumulators into acc_new (which must be zero on entry; obst_opt_finalize leaves it zero).
 * max_cols = largest last dim over the items passed; vec4 != 0 promises that every last dim is a
 * multiple of 4 (the streaming kernel; otherwise a slower generic kernel runs).  clip <= 0 skips
 * adaptive_clip (norms may then be NULL).
 * wnext (double[ntensors], may be NULL): += sum cast_AT(p_new)^2 per tensor, for the next step's carry.
 * preview: the reference's macro-batching chain (src/run/train.py:35-74).  bit 0: a NON-FINAL macro-batch -- the
 * optimizer state (m, accumulators) is updated, the master p is not, and `shadow` receives what the next
 * macro-batch's graph reads: subtract(var.value, cast(update)) evaluated in the activation dtype
 * (src/optimizer/__init__.py:65-66, src/optimizer/optimizers.py:25-29).  bit 1: the weight the graph sees (weight
 * decay here, the clip norm via obst_opt_norms flags&8) is the shadow's current value instead of cast_AT(p) --
 * macro-batches after the first read the chained weights.  0 = the ordinary committed update. */
int obst_opt_update(const obst_opt_tensor* table, int32_t ntensors, const int32_t* item_tensor,
                    const int32_t* item_row0, const int32_t* item_rows, int32_t nitems,
                    int32_t max_cols, int32_t vec4, int32_t dtype_act, int32_t dtype_state,
                    int32_t grad_bf16, const double* norms, double* wnext, float clip, float beta,
                    int32_t nesterov, float lr, float weight_decay, int32_t preview, void* stream);
/* acc[i] = cast_state(acc_new[i]); acc_new[i] = 0. */
int obst_opt_finalize(const obst_opt_tensor* table, int32_t ntensors, int32_t dtype_state,
                      void* stream);

/* Batched form of obst_transpose_param: every transposed weight shadow in one launch.  items: DEVICE
 * array; item i is `batch` row-major [rows, cols] matrices at src, written as [cols, rows] at dst;
 * tile0 = number of 64x64 tiles of all earlier items (exclusive prefix sum of
 * batch * ceil(rows/64) * ceil(cols/64)); total_tiles = that sum over all items. */
typedef struct obst_transpose_item {
  const void* src;
  void* dst;
  int32_t batch, rows, cols, tile0;
} obst_transpose_item;
int obst_transpose_batched(int dtype, const obst_transpose_item* items, int32_t nitems, int32_t total_tiles,
                           void* stream);

/* ---- autoregressive sampling, one step -------------------------------------------------------------
 * src/run/inference.py:76-95 (language branch of autoregressive_model's loop body), after the forward pass:
 * tokens[b, position] = argmax_v(logits[b, position-1, v] + log(-log(U)) * (-temperature)),
 * U ~ Uniform(1e-9, 1) -- the one entry of shift(argmax(...), +1, wrap=False) the loop keeps; position 0
 * receives the shift's zero.  logits AT [batch, seq, vocab]; tokens int32 [batch, seq], updated in place.
 * uniform: NULL (counter-based generator keyed by seed / position / row / entry) or fp32 [batch, vocab]
 * samples to use instead (tests).  temperature 0 = greedy. */
int obst_sample_step(int dtype, const void* logits, int32_t* tokens, const float* uniform, int32_t batch,
                     int32_t seq, int32_t vocab, int32_t position, float temperature, uint64_t seed,
                     void* stream);

/* ---- generic optimizer chains ---------------------------------------------------------------------
 * Every stage the reference registers in OPTIMIZERS (src/optimizer/optimizers.py:131-151), one launch per
 * stage over the same tensor table and work list, for optimizer strings other than the hot path's
 * (which obst_opt_update runs fused).  A stage rewrites the working gradient `work` in place -- as the
 * reference's stage functions return the new ctx.grad -- and keeps its slots in flat state buffers laid
 * out like the flat gradient buffer (element i of tensor t lives at (table[t].g - grad_base) + i).
 *   REDUCE             sums[t] += {sum g^2, sum g, sum cast_AT(p)^2, sum cast_AT(p)} (flags&2: clear first)
 *   ROUND_BF16         g <- bf16(g): the gradient dtype the graph hands over (src/optimizer/__init__.py:180)
 *   COPY               work2 <- work
 *   ADAPTIVE_CLIP      f[0]=clip                                   optimizers.py:79-84    (reads sums)
 *   L2NORM_CLIP        f[0]=clip                                   optimizers.py:87-90    (reads sums)
 *   GLOBAL_L2NORM_CLIP f[0]=clip; sums = REDUCE of the RAW grads   optimizers.py:93-98
 *   VALUE_CLIP         f[0]=clip                                   optimizers.py:101-103
 *   GRAD_CENTRALISE / WEIGHT_CENTRALISE                            optimizers.py:106-111  (reads sums)
 *   LEARNING_RATE      f[0]=lr                                     optimizers.py:114-115
 *   MOMENTUM           f[0]=momentum mult, f[1]=grad mult, flags&1 nesterov; state_a   optimizers.py:118-128
 *   ADAM               f={beta1, beta2, 1/(1-beta1^t), 1/(1-beta2^t)}; state_a=exp_avg_p1, state_b=exp_avg_p2
 *                                                                  optimizers.py:32-43
 *   NOVOGRAD           f={beta1, beta2, -, 1/(1-beta2^t)}; state_a=exp_avg_p1; scalars_in/out = exp_avg_p2
 *                      (one state-dtype scalar per tensor, in != out)    optimizers.py:46-57 (reads sums)
 *   GRAFT_COMBINE      g <- g * rsqrt(sum g^2) * sqrt(sum inner^2); sums of g, sums2 of the inner step
 *                                                                  optimizers.py:147-151
 *   APPLY              p <- p - (g + [decay] cast_AT(p) * f[0]); shadow refreshed;  f[0] = lr * weight_decay
 *                                                                  src/optimizer/__init__.py:49-66
 * SM3 as a stage of such a chain is obst_opt_sm3_stage (+ obst_opt_finalize). */
enum {
  OBST_STAGE_REDUCE = 0, OBST_STAGE_ROUND_BF16, OBST_STAGE_COPY, OBST_STAGE_ADAPTIVE_CLIP,
  OBST_STAGE_L2NORM_CLIP, OBST_STAGE_GLOBAL_L2NORM_CLIP, OBST_STAGE_VALUE_CLIP,
  OBST_STAGE_GRAD_CENTRALISE, OBST_STAGE_WEIGHT_CENTRALISE, OBST_STAGE_LEARNING_RATE,
  OBST_STAGE_MOMENTUM, OBST_STAGE_ADAM, OBST_STAGE_NOVOGRAD, OBST_STAGE_GRAFT_COMBINE, OBST_STAGE_APPLY
};
typedef struct obst_opt_stage_args {
  int32_t stage, dtype_act, dtype_state, flags;
  float f[4];
  const float* grad_base; /* start of the flat gradient buffer the table's g pointers point into */
  float* work;            /* buffer the stage rewrites (normally == grad_base) */
  float* work2;           /* COPY destination */
  void* state_a;
  void* state_b;
  const void* scalars_in;
  void* scalars_out;
  double* sums;           /* [ntensors][4] */
  const double* sums2;    /* [ntensors][4] */
} obst_opt_stage_args;
int obst_opt_stage(const obst_opt_tensor* table, int32_t ntensors, const int32_t* item_tensor,
                   const int32_t* item_row0, const int32_t* item_rows, int32_t nitems,
                   const obst_opt_stage_args* args, void* stream);
/* g <- g * opt_rsqrt(nu) for every tensor of the work list (src/optimizer/optimizers.py:60-76); the new
 * accumulators are max-reduced into acc_new, obst_opt_finalize commits them.  Reads table[t].g in place. */
int obst_opt_sm3_stage(const obst_opt_tensor* table, int32_t ntensors, const int32_t* item_tensor,
                       const int32_t* item_row0, const int32_t* item_rows, int32_t nitems,
                       int32_t max_cols, int32_t vec4, int32_t dtype_state, void* stream);

/* ---- the same update for ONE tensor, without any device allocation -----------------------------------
 * No counterpart in the reference (its update always runs inside the Mesh-TensorFlow graph).
 * obst_opt_norms/update/finalize need a tensor table, a work list, norm sums and SM3 staging in device
 * memory that the caller has to allocate.  A host that cannot allocate -- extension code inside a process
 * whose framework already owns all of the GPU's memory, or a runtime that bars driver allocation calls --
 * uses this entry point instead: the table, work list, norm sums and staging live in a fixed ~114 KB block
 * inside the library image (a file-scope __device__ array; every GPU that loads the library holds it until
 * the process exits).  The caller passes only the tensor itself:
 *   p, g fp32 [numel], shadow (activation dtype, may be NULL), m (state dtype) -- DEVICE pointers, all
 *   16-byte aligned when the last dim is a multiple of 4;
 *   host_acc: HOST array of ndim DEVICE pointers, host_acc[i] = SM3 accumulator of dim i (state dtype,
 *   host_dims[i] entries); host_dims: HOST array of the ndim dims; ndim 1..4;
 *   remaining arguments as obst_opt_update.
 * Limits (rejected with OBST_EINVAL and a message): the dims, each rounded up to 64, may add up to at most
 * 16640; the tensor may need at most 4096 work items of 32768/min(last_dim, 512) rows (134 M elements).
 * The first call on each device looks up the block's device address; make that call before capturing a
 * stream.  Every call on a device shares the block (it is cleared on entry where the update needs zeros),
 * so two calls that overlap on the GPU -- different streams, or graph branches -- race: serialise them.
 * Calls no cudaMalloc / cudaFree or other allocating routine; enqueues 3-4 kernels on `stream`. */
int obst_opt_step_noalloc(float* p, void* shadow, float* g, void* m, void* const* host_acc,
                          const int32_t* host_dims, int32_t ndim, int32_t decay, int32_t tril_s,
                          int32_t dtype_act, int32_t dtype_state, int32_t grad_bf16, float clip,
                          float beta, int32_t nesterov, float lr, float weight_decay, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* OBST_B200_H */
