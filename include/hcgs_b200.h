/*
 * hcgs_b200.h -- C ABI of the B200-native HCGS acoustic-model hot path.
 *
 * The reference (hellboywyh/pytorch-kaldi-CGS) is 100 % Python: it has no FFI of its own.
 * Every entry point below therefore cites the reference *Python* call site it replaces
 * (file:line relative to the reference root).  The reference-side binding is a ctypes stub
 * (see INTEGRATION.md); `pytorch-kaldi-cgs_b200/b200/native.py` is that stub.
 *
 * Conventions
 *   - plain C types only: device pointers, sizes, a cudaStream_t passed as void*.
 *   - every function returns 0 on success, non-zero on failure; hcgs_last_error() gives the text.
 *     No C++ exception crosses the ABI.
 *   - no ownership transfer: the caller (PyTorch) owns every buffer; the library owns only the
 *     opaque hcgs_plan (device copies of the block index) until hcgs_plan_destroy().
 *   - all matrices are row-major fp32 unless stated; "ld" = leading dimension in elements.
 *   - gate order everywhere: 0=f (forget), 1=i (input), 2=o (output), 3=c (candidate).
 *   - functions are asynchronous on `stream`; a plan is thread-compatible (one user at a time).
 */
#ifndef HCGS_B200_H
#define HCGS_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define HCGS_ABI_VERSION 1
#define HCGS_MAX_GROUPS 4 /* gates of one LSTM layer share a mask and are processed together */

typedef struct hcgs_plan hcgs_plan;
typedef void* hcgs_stream; /* cudaStream_t */

/* activation codes = reference act_fun() names (neural_networks.py:54-78) */
enum {
    HCGS_ACT_LINEAR = 0,
    HCGS_ACT_RELU = 1,
    HCGS_ACT_TANH = 2,
    HCGS_ACT_HTANH = 3,
    HCGS_ACT_SIGMOID = 4,
    HCGS_ACT_LEAKY_RELU = 5, /* slope 0.2 */
    HCGS_ACT_ELU = 6
};

/* ---- library ---------------------------------------------------------------------------- */
int hcgs_abi_version(void);
const char* hcgs_last_error(void);
/* info[0]=SM count, [1]=cc major, [2]=cc minor, [3]=max opt-in smem/block, [4]=cluster launch ok */
int hcgs_device_info(int device, int* info);

/* ---- plan: HCGS/pattern mask -> compressed block index ------------------------------------
 * Replaces nothing 1:1 -- the reference multiplies dense weights by the dense mask on every
 * forward (neural_networks.py:858-861, 980-983, 258).  The plan is built ONCE per mask from the
 * host copy of the 0/1 mask (row-major [n_out, n_in]); a non-zero entry marks a kept weight.
 * It holds (a) the 32x32-tile CSR/CSC index used by the block-sparse GEMMs and (b), when
 * recurrent != 0 (requires n_out == n_in), the per-row ELL index used by the persistent
 * recurrent kernels.  cluster_size: 0 = choose automatically, else 1/2/4/8/16.
 */
int hcgs_plan_create(const float* mask_host, int n_out, int n_in, int recurrent, int cluster_size,
                     hcgs_plan** plan);
int hcgs_plan_destroy(hcgs_plan* plan);
/* info[0]=n_out [1]=n_in [2]=non-empty 32x32 tiles [3]=total tiles [4]=nnz(mask)
 * [5]=cluster size [6]=rows per CTA [7]=block-ELL width L [8]=fwd pack floats [9]=bwd pack floats
 * [10]=fwd smem bytes [11]=bwd smem bytes [12]=threads per CTA [13]=4x4 blocks [14]=weights in smem */
int hcgs_plan_info(const hcgs_plan* plan, int64_t* info);

/* ---- elementwise weight preparation -------------------------------------------------------
 * hcgs_mask_apply: W *= mask in place for n matrices sharing one mask
 *   (neural_networks.py:858-861, 980-983 LSTM; :258 MLP).                                    */
int hcgs_mask_apply(float* const* w, int n, const float* mask, int64_t numel, hcgs_stream stream);
/* hcgs_quantize_w: quantized_modules.Quantize(balanced=False) (quantized_modules.py:77-97):
 *   master is clamped to [-1,1] IN PLACE; q = sign(w)*ceil(|w|*2^(bits-1))/2^(bits-1) is written
 *   to `q` (q may alias master for the if_forward=True in-place variant).                     */
int hcgs_quantize_w(float* master, float* q, int64_t numel, int bits, hcgs_stream stream);
/* hcgs_quantize_act: quantized_modules.Quantize_inp (quantized_modules.py:99-119) applied
 *   `repeat` times in a row (the four gate linears each re-quantize the same tensor).
 *   x is [rows, cols] with leading dimension ldx; scratch = 2 floats of device memory.       */
int hcgs_quantize_act(const float* x, int ldx, float* y, int ldy, int64_t rows, int cols, int bits,
                      int repeat, float* scratch, hcgs_stream stream);

/* ---- block-sparse GEMM family (time-batched projections and their gradients) --------------
 * n_w weight matrices W_g [n_out, n_in] (ld = ldw) share the plan's mask.
 * hcgs_bsgemm_nt:  Y[m, g*n_out + n] = sum_k X[m,k] * W_g[n,k] (+ bias_g[n])
 *   replaces the four F.linear calls neural_networks.py:951-954 and MLP wx(x) :306-317.
 * hcgs_bsgemm_nn:  dX[m,k] = sum_g sum_n dY[m, g*n_out + n] * W_g[n,k]      (autograd of the above)
 * hcgs_bs_sddmm:   dW_g[n,k] = mask[n,k] * sum_m dY[m, g*n_out + n] * X[m,k]  only inside non-empty
 *   tiles (mask-preserving weight gradient; entries outside non-empty tiles are left untouched,
 *   so the caller passes a zero-initialised buffer).  mask may be NULL (no re-masking).
 * bias / dW are host arrays of n_w device pointers (bias may be NULL).                        */
int hcgs_bsgemm_nt(const hcgs_plan* plan, int n_w, const float* const* w, int ldw, const float* x,
                   int ldx, int64_t m, const float* const* bias, float* y, int ldy,
                   hcgs_stream stream);
int hcgs_bsgemm_nn(const hcgs_plan* plan, int n_w, const float* const* w, int ldw, const float* dy,
                   int lddy, int64_t m, float* dx, int lddx, hcgs_stream stream);
int hcgs_bs_sddmm(const hcgs_plan* plan, int n_w, const float* dy, int lddy, const float* x, int ldx,
                  int64_t m, float* const* dw, int lddw, const float* mask, hcgs_stream stream);
/* column sums: out_g[n] = sum_m dY[m, g*n_out+n]  (bias gradient) */
int hcgs_colsum(const float* dy, int lddy, int64_t m, int n_groups, int group_cols, float* const* out,
                hcgs_stream stream);

/* ---- batch norm over the time-batched rows (nn.BatchNorm1d, momentum 0.05) ------------------
 * C = n_groups*group_cols columns; group g uses gamma[g]/beta[g]/running_*[g] (each [group_cols]).
 * Replaces neural_networks.py:957-976 (LSTM, 4 groups) and :309-314 (MLP, 1 group).
 * training!=0: batch statistics (biased var for the normalisation, unbiased for running_var),
 * running stats updated in place, save_mean/save_invstd [C] written for the backward.
 * training==0: running statistics.  y may alias x.  workspace: >= 2*C floats of device scratch
 * (more lets the statistics use more CTAs: up to 2*C*256 floats are used).                    */
int hcgs_bn_fwd(const float* x, int ldx, float* y, int ldy, int64_t m, int n_groups, int group_cols,
                const float* const* gamma, const float* const* beta, float* const* running_mean,
                float* const* running_var, float eps, float momentum, int training,
                float* save_mean, float* save_invstd, float* workspace, int64_t workspace_floats,
                hcgs_stream stream);
/* hcgs_bn_fwd_static: hcgs_bn_fwd WITHOUT a caller workspace and without any allocation.  It has NO counterpart in
 * the reference: it was added for callers that cannot call cudaMalloc (code
 * under CUDA graph capture, tight serving loops, harnesses that forbid it).  The statistics scratch is a file-scope
 * device array of the library (2 * HCGS_BN_STATIC_MAX_COLS * 64 floats = 2 MiB per GPU that loads the library, held
 * until exit).  Limit: n_groups * group_cols <= HCGS_BN_STATIC_MAX_COLS, anything wider is rejected with an error
 * (use hcgs_bn_fwd).  Calls no cudaMalloc / cudaFree / allocating library routine; the first call on a device
 * resolves the array's address (cudaGetSymbolAddress) -- make that call outside a capture.  Every call shares the
 * array (nothing in it has to be cleared: each call writes what it reads), so calls that OVERLAP on the GPU race;
 * serialise them on one stream or with events.  Same arithmetic and results as hcgs_bn_fwd with a workspace of
 * 2 * C * 64 floats.                                                                                           */
#define HCGS_BN_STATIC_MAX_COLS 4096
int hcgs_bn_fwd_static(const float* x, int ldx, float* y, int ldy, int64_t m, int n_groups, int group_cols,
                       const float* const* gamma, const float* const* beta, float* const* running_mean,
                       float* const* running_var, float eps, float momentum, int training,
                       float* save_mean, float* save_invstd, hcgs_stream stream);
int hcgs_bn_bwd(const float* dy, int lddy, const float* x, int ldx, int64_t m, int n_groups,
                int group_cols, const float* const* gamma, const float* save_mean,
                const float* save_invstd, float* dx, int lddx, float* const* dgamma,
                float* const* dbeta, float* workspace, int64_t workspace_floats, hcgs_stream stream);

/* gate-interleaved variants for the tensor-core build: x / y / dy / dx are [m, n_groups*ld_units] with
 * column u*n_groups+g (u < units real, pad units up to ld_units stay zero); hcgs_bn_bwd_il reads dy as
 * fp32 (dy) or bf16 (dy_bf16, the copy hcgs_lstm_tc_bwd writes for the GEMMs) - exactly one of the two
 * is non-NULL - and writes dx as fp32 and/or bf16 (either pointer may be NULL).                          */
int hcgs_bn_fwd_il(const float* x, float* y, int64_t m, int n_groups, int units, int ld_units,
                   const float* const* gamma, const float* const* beta, float* const* running_mean,
                   float* const* running_var, float eps, float momentum, int training, float* save_mean,
                   float* save_invstd, float* workspace, int64_t workspace_floats, hcgs_stream stream);
int hcgs_bn_bwd_il(const float* dy, const void* dy_bf16, const float* x, int64_t m, int n_groups, int units, int ld_units,
                   const float* const* gamma, const float* save_mean, const float* save_invstd, float* dx,
                   void* dx_bf16, float* const* dgamma, float* const* dbeta, float* workspace,
                   int64_t workspace_floats, hcgs_stream stream);

/* statistics only (tensor-core build; the recurrent kernel applies them): training != 0 merges `splits`
 * per-tile {mean, M2} partials ([splits][n_groups*ld_units][2], rows_per_split rows each, m rows in all)
 * with Chan's formula in a fixed order, writes save_mean / save_invstd ([n_groups*ld_units]) and updates the
 * running statistics like nn.BatchNorm1d; training == 0 writes the running statistics into save_*.   */
int hcgs_bn_stats_il(const float* partial, int64_t splits, int64_t rows_per_split, int64_t m, int n_groups,
                     int units, int ld_units, float* const* running_mean, float* const* running_var, float eps,
                     float momentum, int training, float* save_mean, float* save_invstd, hcgs_stream stream);

/* ---- the reference's LayerNorm (neural_networks.py:40-51; NOT nn.LayerNorm) ---------------------
 * y = gamma * (x - mean) / (std + eps) + beta over the last dimension, std unbiased (n-1).
 * Used by *_use_laynorm_inp (:825-826, :248-249) and dnn_use_laynorm (:306-314).  save_mean / save_std
 * ([m], optional, both or none) feed hcgs_layernorm_bwd, which writes dx (optional) and ADDS the column sums
 * to dgamma / dbeta (caller zero-initialises; either may be NULL).                                       */
int hcgs_layernorm_fwd(const float* x, int64_t ldx, float* y, int64_t ldy, int64_t m, int n,
                       const float* gamma, const float* beta, float eps, float* save_mean, float* save_std,
                       hcgs_stream stream);
int hcgs_layernorm_bwd(const float* dy, int64_t lddy, const float* x, int64_t ldx, int64_t m, int n,
                       const float* gamma, float eps, const float* save_mean, const float* save_std, float* dx,
                       int64_t lddx, float* dgamma, float* dbeta, hcgs_stream stream);

/* ---- persistent recurrent LSTM kernels ------------------------------------------------------
 * hcgs_lstm_pack: gather the four masked U_g [H,H] (ld = ldu) into the plan's compressed
 *   resident layouts (fwd_pack: info[8] floats, bwd_pack: info[9] floats; either may be NULL).
 * hcgs_lstm_fwd: the whole time loop of one layer (neural_networks.py:1063-1100):
 *     f,i,o = sigmoid(pre_{f,i,o}[t] + U_{f,i,o} h) ; g = act(pre_c[t] + U_c h)
 *     c = i*g*drop + f*c ; h = o*act(c)
 *   pre [T,B,4H] (column g*H+u), drop [B,H] or NULL (then drop_scalar is used, eval mode),
 *   h_out [T,B,H]; c_out [T,B,H] and gates_out [T,B,4H] (post-activation f,i,o,g) are saved for
 *   the backward when non-NULL.  act_bits>0 enables the per-step Quantize_inp of h
 *   (quantized_modules.py:212/219 through QuantizeLinear on h_t, 4 applications per step).  Its scale is the
 *   max |h_t| over the WHOLE batch (quantized_modules.py:104-108): with more than 8 sequences the clusters exchange
 *   their maxima once per time step through a static device scratch, so every cluster must be resident at the same
 *   time -- the call fails with the limit in hcgs_last_error() when batch exceeds 8 x the co-resident clusters -- and
 *   only one such launch may run on a device at a time.
 * hcgs_lstm_bwd: BPTT of the same loop; dh_out [T,B,H] is the gradient arriving from above,
 *   dpre [T,B,4H] the gradient w.r.t. the gate pre-activations.                  */
int hcgs_lstm_pack(const hcgs_plan* plan, const float* const* u, int ldu, float* fwd_pack,
                   float* bwd_pack, hcgs_stream stream);
int hcgs_lstm_fwd(const hcgs_plan* plan, const float* fwd_pack, const float* pre, const float* drop,
                  float drop_scalar, int act, int t_len, int batch, float* h_out, float* c_out,
                  float* gates_out, int act_bits, hcgs_stream stream);
int hcgs_lstm_bwd(const hcgs_plan* plan, const float* bwd_pack, const float* dh_out,
                  const float* gates, const float* c, const float* drop, float drop_scalar, int act,
                  int t_len, int batch, float* dpre, hcgs_stream stream);
/* LayerNorm of h_t INSIDE the time loop (lstm_use_laynorm: neural_networks.py:1092-1095 with the reference's own class at
 * :40-51 -- gamma * (h - mean) / (std + eps) + beta over the H units of a sequence, UNBIASED std).  The normalised h_t is
 * what the next step's U h product, Quantize_inp (act_bits) and the layer output see.  Each cluster normalises its 8
 * sequences on the gathered h_t; the backward takes dh through the LayerNorm backward at every step (two cluster-wide sums
 * per sequence) before the gate derivatives.
 * hcgs_lstm_fwd_ln: hcgs_lstm_fwd + ln_gamma / ln_beta [H], ln_eps; ln_stats [T,B,2] = {mean, std + eps} (NULL when no
 *   backward follows).
 * hcgs_lstm_bwd_ln: hcgs_lstm_bwd + the same gamma / stats / eps; ln_dgamma / ln_dbeta: [ceil(B/8)][H] partial parameter
 *   gradients, one row per cluster (deterministic; the caller sums the rows).                                        */
int hcgs_lstm_fwd_ln(const hcgs_plan* plan, const float* fwd_pack, const float* pre, const float* drop,
                     float drop_scalar, int act, int t_len, int batch, float* h_out, float* c_out, float* gates_out,
                     int act_bits, const float* ln_gamma, const float* ln_beta, float ln_eps, float* ln_stats,
                     hcgs_stream stream);
int hcgs_lstm_bwd_ln(const hcgs_plan* plan, const float* bwd_pack, const float* dh_out, const float* gates,
                     const float* c, const float* drop, float drop_scalar, int act, int t_len, int batch, float* dpre,
                     const float* ln_gamma, const float* ln_stats, float ln_eps, float* ln_dgamma, float* ln_dbeta,
                     hcgs_stream stream);

/* ---- persistent recurrent kernels of the GRU family (fp32 build) ----------------------------
 * One kernel pair for the four gated / plain cells of the toolkit.  The reference has them dense only
 * (neural_networks.py GRU :1240-1426, liGRU :1429-1599, minimalGRU :1602-1777, RNN :1780-1931); here the
 * recurrent matrices take an HCGS mask like the LSTM's, and the dense reference is the all-ones mask.
 *   GRU        z = s(wz + Uz h)  r = s(wr + Ur h)  a = wh + Uh (r*h)      (:1395-1408)
 *   minimalGRU z = s(wz + Uz h)                    a = wh + Uh (z*h)      (:1752-1763)
 *   liGRU      z = s(wz + Uz h)                    a = wh + Uh h          (:1574-1585)
 *   RNN                                            a = wh + Uh h          (:1907-1916)
 *   hcand = act(a) * drop ; h = z*h + (1-z)*hcand (RNN: h = hcand); drop [B,H] or NULL (then drop_scalar, eval).
 * Time-batched buffers hold NG = 3 / 2 / 2 / 1 groups of H columns, gate-major, in the order z, r, candidate
 * (absent gates left out): pre / gates_out / dpre are [T, B, NG*H].
 * Weights: hcgs_lstm_pack with u = {Uz, Ur, Uh, any} -- slot 0 = update, 1 = reset, 2 = candidate; a cell without
 *   a gate passes any valid matrix in that slot (never read).
 * hcgs_gru_fwd: h_out [T,B,H]; saved for the backward when non-NULL: gates_out (post-activation z, r, act(a)) and,
 *   for GRU / minimalGRU, m_out [T,B,H] = the gated copy r*h_{t-1} (z*h_{t-1}) that Uh multiplied (operand of dUh).
 * hcgs_gru_bwd: BPTT; dh_out [T,B,H] gradient from above, h = the forward's h_out; dpre [T,B,NG*H].          */
#define HCGS_CELL_GRU 0
#define HCGS_CELL_MINGRU 1
#define HCGS_CELL_LIGRU 2
#define HCGS_CELL_RNN 3
int hcgs_gru_fwd(const hcgs_plan* plan, int cell, const float* fwd_pack, const float* pre, const float* drop,
                 float drop_scalar, int act, int t_len, int batch, float* h_out, float* m_out, float* gates_out,
                 hcgs_stream stream);
int hcgs_gru_bwd(const hcgs_plan* plan, int cell, const float* bwd_pack, const float* dh_out, const float* gates,
                 const float* h, const float* drop, float drop_scalar, int act, int t_len, int batch, float* dpre,
                 hcgs_stream stream);

/* ---- tensor-core recurrent kernels (bf16 operands, fp32 accumulate and state) ---------------
 * Same loop as hcgs_lstm_fwd / hcgs_lstm_bwd (neural_networks.py:1063-1100 and its autograd), for
 * the "bf16 train" configurations: U and h_{t-1} are rounded to bf16 for the product, gates, cell
 * state and all saved tensors stay fp32.  A cluster of `cluster_size` CTAs (4, 8 or 16; 0 = the
 * smallest that fits) runs `nb` = 16 sequences (0 = 16; no other value is built) through all T steps.
 * Layout of this build's tensors (ld_h = H rounded up to a multiple of 8, pad columns zero):
 *   pre / gates / dpre : [T*B, 4*ld_h] fp32, gate-INTERLEAVED: (row, unit u, gate g) at row*4*ld_h + u*4 + g
 *   c                  : [T*B, ld_h] fp32          hb (bf16 copy of h for the GEMMs): [T*B, ld_h]
 *   h_out              : [T*B, ld_h] fp32 (pad columns zero; the reference's [T,B,H] tensor is its [:, :H] view)
 *   dh_out             : [T*B, ld_h] fp32 (pad columns are ignored), drop: [B,H]
 * hcgs_lstm_tc_info: info[0]=cluster size [1]=nb [2]=fwd pack uint32s [3]=bwd pack uint32s
 *   [4]=fwd smem bytes [5]=bwd smem bytes [6]=fwd warps [7]=strips/CTA [8]=k-steps [9]=4x4 blocks
 *   [10]=strips/warp [11]=padded hidden size of the cluster.
 * In hcgs_lstm_tc_fwd h_out, hb_out, c_out, gates_out are each optional (at least one of h_out /
 * hb_out); hcgs_lstm_tc_bwd optionally writes a bf16 copy of dpre (same layout) for the GEMMs.
 * Fused batch norm (neural_networks.py:957-976): with bn_mean / bn_invstd ([4*ld_h], gate-interleaved, from
 * hcgs_bn_stats_il) the kernel uses (pre - mean) * invstd * gamma_g + beta_g in place of pre, so the
 * normalised projections never touch HBM; bn_gamma / bn_beta are four per-gate [H] pointers (NULL = 1 / 0).
 * All four NULL = pre is taken as is.                                                             */
int hcgs_lstm_tc_info(const hcgs_plan* plan, int cluster_size, int nb, int64_t* info);
int hcgs_lstm_tc_pack(const hcgs_plan* plan, int cluster_size, const float* const* u, int ldu,
                      uint32_t* fwd_pack, uint32_t* bwd_pack, hcgs_stream stream);
int hcgs_lstm_tc_fwd(const hcgs_plan* plan, int cluster_size, int nb, const uint32_t* fwd_pack,
                     const float* pre, int ld_h, const float* drop, float drop_scalar, int act,
                     int t_len, int batch, float* h_out, void* hb_out, float* c_out, float* gates_out,
                     const float* bn_mean, const float* bn_invstd, const float* const* bn_gamma,
                     const float* const* bn_beta, hcgs_stream stream);
/* hcgs_lstm_tc_fwd_split: the same launch also writing hl_out = bf16(h - float(bf16(h))) next to hb_out = bf16(h)
 * ([t_len*batch, ld_h] each): the two-term operand of the NEXT layer's projection (hcgs_tc_gemm_nt_split) straight out of
 * the recurrence, instead of a separate hcgs_split_bf16 pass over the fp32 h.                                         */
int hcgs_lstm_tc_fwd_split(const hcgs_plan* plan, int cluster_size, int nb, const uint32_t* fwd_pack,
                           const float* pre, int ld_h, const float* drop, float drop_scalar, int act,
                           int t_len, int batch, float* h_out, void* hb_out, void* hl_out, float* c_out,
                           float* gates_out, const float* bn_mean, const float* bn_invstd,
                           const float* const* bn_gamma, const float* const* bn_beta, hcgs_stream stream);
/* tuning aid: with HCGS_TC_DEBUG=64 in the environment the recurrent kernels accumulate clock64 deltas per
 * phase (thread 0 of CTA 0); out16 receives the 16 counters of the last traced launch.              */
int hcgs_tc_trace_read(int64_t* out16);
/* tuning aid: with HCGS_TC_DEBUG=128 hcgs_lstm_tc_bwd records clock64 stamps of the warps of CTA 0 at the middle
 * time step (20 warps x 16 stamps; no counterpart in the reference).                                   */
int hcgs_tc_timeline_read(int64_t* out320);
int hcgs_lstm_tc_bwd(const hcgs_plan* plan, int cluster_size, int nb, const uint32_t* bwd_pack,
                     const float* dh_out, const float* gates, const float* c, int ld_h, const float* drop,
                     float drop_scalar, int act, int t_len, int batch, float* dpre, void* dpre_bf16,
                     hcgs_stream stream);

/* ---- tcgen05 block-sparse bf16 GEMM family (tensor-core build) ------------------------------
 * Same products as hcgs_bsgemm_nt / _nn / hcgs_bs_sddmm with bf16 operands and fp32 accumulation
 * in TMEM.  Activations are bf16 row-major [m, ld] (ld*2 a multiple of 16 bytes, pad columns
 * zero), loaded by TMA; weights are re-packed per step into pre-swizzled bf16 tiles.
 * With n_w = 4 (LSTM gates) the output / dY columns are gate-INTERLEAVED: (unit u, gate g) at
 * column u*4+g; with n_w = 1 (MLP) they are the plain unit index.
 * hcgs_tc_pack_nt: wpack needs nz_tiles * 32*n_w * 32 bf16 (plan info[2] tiles).
 * hcgs_tc_gemm_nt stats (optional, n_w = 4 only): [ceil(m/128)][ldy][2] floats; the epilogue writes, per
 *   128-row tile and output column, {mean, M2} of that tile's rows -- the batch-norm statistics of the
 *   projections at no extra pass over them (merged by hcgs_bn_stats_il).
 * hcgs_cast_bf16 : fp32 [rows, cols] -> bf16 [rows, ldo], pad columns zeroed.
 * Two-term ("split") forward products -- the projections W{f,i,o,c}x.x (neural_networks.py:951-954) and the
 *   hidden MLP layers (:306) of the tensor-core build: each fp32 operand is carried as hi + lo with
 *   hi = bf16(v), lo = bf16(v - hi) (16 mantissa bits) and the product is x_hi w_hi + x_lo w_hi + x_hi w_lo
 *   (three tcgen05 products per kept k tile, fp32 accumulation; the lo*lo term, 2^-16 relative, is dropped).
 *   This is what keeps the log-posteriors within 1e-4 of the reference's fp32 path (tools/bf16_attribution.py:
 *   the single-term projections alone cost 1.8e-4).
 *   hcgs_tc_pack_nt_split: wpack2 needs 2 * nz_tiles * 32*n_w * 32 bf16 ([hi tile | lo tile] per kept tile).
 *   hcgs_tc_gemm_nt_split: x_hi / x_lo bf16 [m, ldx_hi] / [m, ldx_lo]; otherwise as hcgs_tc_gemm_nt.
 *   hcgs_split_bf16: fp32 [rows, cols] -> hi (may be NULL) and lo, bf16 [rows, ldo], pad columns zeroed. */
int hcgs_tc_pack_nt(const hcgs_plan* plan, int n_w, const float* const* w, int ldw, void* wpack,
                    hcgs_stream stream);
int hcgs_tc_gemm_nt(const hcgs_plan* plan, int n_w, const void* wpack, const void* x_bf16, int64_t ldx,
                    int64_t m, const float* const* bias, float* y, int64_t ldy, float* stats,
                    hcgs_stream stream);
int hcgs_cast_bf16(const float* x, int64_t ldx, int64_t rows, int cols, void* out, int64_t ldo,
                   hcgs_stream stream);
int hcgs_tc_pack_nt_split(const hcgs_plan* plan, int n_w, const float* const* w, int ldw, void* wpack2,
                          hcgs_stream stream);
int hcgs_tc_gemm_nt_split(const hcgs_plan* plan, int n_w, const void* wpack2, const void* x_hi, int64_t ldx_hi,
                          const void* x_lo, int64_t ldx_lo, int64_t m, const float* const* bias, float* y,
                          int64_t ldy, float* stats, hcgs_stream stream);
int hcgs_split_bf16(const float* x, int64_t ldx, int64_t rows, int cols, void* hi, void* lo, int64_t ldo,
                    hcgs_stream stream);
/* hcgs_tc_pack_nn / hcgs_tc_gemm_nn: dX[m,k] = sum_{u,g} dY[m,(u,g)] W_g[u,k]; wpack same size as for nt;
 *   dY bf16 [m, lddy] (lddy = padded n_out * n_w, pad columns zero); dX fp32 [m, lddx].
 * hcgs_tc_sddmm: dW_g[u,k] += mask[u,k] * sum_m dY[m,(u,g)] X[m,k] inside kept tiles only (fp32 atomics:
 *   the caller zero-initialises dW; entries outside kept tiles stay untouched).                  */
int hcgs_tc_pack_nn(const hcgs_plan* plan, int n_w, const float* const* w, int ldw, void* wpack,
                    hcgs_stream stream);
int hcgs_tc_gemm_nn(const hcgs_plan* plan, int n_w, const void* wpack, const void* dy_bf16, int64_t lddy,
                    int64_t m, float* dx, int64_t lddx, hcgs_stream stream);
int hcgs_tc_sddmm(const hcgs_plan* plan, int n_w, const void* dy_bf16, int64_t lddy, const void* x_bf16,
                  int64_t ldx, int64_t m, float* const* dw, int64_t lddw, const float* mask,
                  hcgs_stream stream);

/* ---- MLP pointwise pieces ------------------------------------------------------------------
 * y = drop_mask * act(x)  (neural_networks.py:306-317; drop_mask already holds 1/(1-p) scaling,
 * NULL = no dropout).  Backward: dx = dy * drop_mask * act'(.) using the saved output y.       */
int hcgs_act_fwd(const float* x, float* y, const float* drop_mask, int64_t numel, int act,
                 hcgs_stream stream);
int hcgs_act_bwd(const float* dy, const float* y, const float* drop_mask, float* dx, int64_t numel,
                 int act, hcgs_stream stream);

/* ---- output layer epilogue: log-softmax (+ prior normalisation, + NLL/err) ------------------
 * hcgs_logsoftmax_fwd: out[m,:] = z[m,:] - logsumexp(z[m,:]) - (log_prior ? log_prior[:] : 0)
 *   = nn.LogSoftmax(dim=1) (neural_networks.py:73-74) fused with core.py:241-245
 *   (`out - np.log(counts/np.sum(counts))`).  out may alias z.
 * hcgs_logsoftmax_bwd: dz = dlogp - exp(logp) * rowsum(dlogp)       (autograd of LogSoftmax)
 * hcgs_nll_fwd_bwd: fused cost_nll + cost_err (utils.py:1935-1952, 1993-2011) on log-probs:
 *   loss_sum += -logp[m,label[m]], err_sum += (argmax != label); when dz != NULL also writes
 *   dz[m,:] = (exp(logp[m,:]) - onehot) * scale  (gradient of mean NLL w.r.t. the logits).
 *   labels are fp32 (the reference stores them as the last float column, utils.py:1940-1944),
 *   with leading dimension ldlab.  stats = 2 floats {loss_sum, err_sum}, accumulated atomically. */
int hcgs_logsoftmax_fwd(const float* z, int ldz, float* out, int ldo, int64_t m, int p,
                        const float* log_prior, hcgs_stream stream);
int hcgs_logsoftmax_bwd(const float* dlogp, const float* logp, float* dz, int64_t m, int p,
                        hcgs_stream stream);
/* tensor-core build: the same backward fused with what follows it in the output layer -- the bias gradient
 * db[p] += sum_m dz[m,p] (fp32, caller zero-initialises; NULL = skip) and the bf16 copy of dz that the dW / dX
 * products read ([m, ldo], pad columns zeroed).  The fp32 dz is never written.  p % 4 == 0, p <= 4096.   */
int hcgs_logsoftmax_bwd_fused(const float* dlogp, const float* logp, int64_t m, int p, void* dz_bf16,
                              int64_t ldo, float* db, hcgs_stream stream);
/* the same with the incoming gradient in its sparse form: ONE entry per row, entry i = (rows[i] == i, cols[i], vals[i])
 * -- what the backward of nn.NLLLoss is (utils.forward_model cost_nll, utils.py:1931-1940: -1/m at the label column)
 * before it is scattered into a dense [m, p] matrix.  dz[i, j] = vals[i] * ((j == cols[i]) - exp(logp[i, j])); a column
 * outside [0, p) gives a zero row (ignore_index).  rows / cols: DEVICE int64 (the COO indices), vals: DEVICE fp32.
 * The kernel traps when rows[i] != i.                                                                                 */
int hcgs_logsoftmax_bwd_fused_onehot(const int64_t* rows, const int64_t* cols, const float* vals, const float* logp,
                                     int64_t m, int p, void* dz_bf16, int64_t ldo, float* db, hcgs_stream stream);
/* hcgs_nll_err_rows: cost_nll and cost_err of one output in one pass (utils.py:1931-1952): loss_rows[i] = -logp[i, label_i]
 * (0 for a label outside [0, p)), err_rows[i] = (first arg-max of row i != label_i); the caller takes the two means.  */
int hcgs_nll_err_rows(const float* logp, int64_t ldl, int64_t m, int p, const int64_t* labels, float* loss_rows,
                      float* err_rows, hcgs_stream stream);
int hcgs_nll_fwd_bwd(const float* logp, int64_t m, int p, const float* labels, int ldlab, float scale,
                     float* dz, float* stats, hcgs_stream stream);

/* ---- fused output layer (tensor-core build) -------------------------------------------------
 * out[m, :] = log_softmax(x[m, :] W^T + bias) - (log_prior ? log_prior : 0): the MLP output layer
 * (neural_networks.py:306-317 with act = LogSoftmax(dim=1), :73-74) and the prior normalisation of the
 * forward pass (core.py:241-245) as ONE kernel.  The logits stay in TMEM: a cluster of CTAs owns a 128-row
 * tile, each CTA accumulates a slice of the output columns over the kept 32-wide k tiles (tcgen05, bf16
 * operands, fp32 accumulate), the row max / sum-of-exponentials partials are exchanged through
 * distributed shared memory and the log-posteriors are written once.  HBM traffic per row: n_in*2 + n_out*4 B.
 * hcgs_out_info: info[0]=cluster size (0: n_out or n_in > 4096, use hcgs_tc_gemm_nt + hcgs_logsoftmax_fwd)
 *   [1]=strips of 64 columns per CTA [2]=packed tiles [3]=TMEM columns [4]=smem bytes [5]=k items
 *   [6]=bytes of the packed weights [7]=ring stages.
 * hcgs_out_pack: masked dense fp32 W [n_out, n_in] -> pre-swizzled bf16 tiles (info[6] bytes).
 * x_bf16: [m, ldx] bf16 (pad columns zero, ldx*2 a multiple of 16); out: [m, ldo] fp32.            */
int hcgs_out_info(const hcgs_plan* plan, int64_t* info);
int hcgs_out_pack(const hcgs_plan* plan, const float* w, int ldw, void* wpack, hcgs_stream stream);
int hcgs_out_logsoftmax_prior_fwd(const hcgs_plan* plan, const void* wpack, const void* x_bf16, int64_t ldx,
                                  int64_t m, const float* bias, const float* log_prior, float* out,
                                  int64_t ldo, hcgs_stream stream);

/* ---- optimizers on flat (compressed or dense) buffers -------------------------------------
 * torch.optim semantics as configured by utils.optimizer_init (utils.py:1833-1881).
 * kind: 0 = SGD(momentum, weight_decay, dampening, nesterov), 1 = RMSprop(alpha, eps,
 * weight_decay, momentum, centered=False), 2 = Adam(beta1, beta2, eps, weight_decay, step).
 * hp = 8 floats: {lr, a, b, eps, weight_decay, c, d, step}; see optimizer.cu.                */
int hcgs_opt_step(int kind, float* param, const float* grad, float* state1, float* state2,
                  int64_t numel, const float* hp, float grad_scale, hcgs_stream stream);
/* hcgs_opt_step_multi: the same update for a whole parameter group in one launch per 48 tensors (torch.optim walks
 * the group tensor by tensor, utils.py:1833-1881 hands it every parameter of a net): HOST arrays of n_tensors device
 * pointers / element counts; state1 / state2 may be NULL when the optimizer kind does not use them.  hp: the same eight
 * values in DOUBLE precision (1 - beta, the bias corrections and lr / bias_correction1 are formed in double and rounded
 * once, as torch.optim does with its Python floats).                                                               */
int hcgs_opt_step_multi(int kind, int n_tensors, float* const* params, const float* const* grads,
                        float* const* state1, float* const* state2, const int64_t* numels, const double* hp,
                        float grad_scale, hcgs_stream stream);

/* ---- minibatch assembly from a chunk resident in HBM ---------------------------------------
 * Replaces the per-minibatch host loop of core.py:183-200 (torch.zeros(max_len, batch, dim) + one slice copy per
 * sentence behind `random.randint(0, max_len - snt_len)` leading zero rows, then `.cuda()`): the chunk's
 * [chunk_rows, dim] fp32 matrix (features and label columns, as read_lab_fea packs them) is uploaded once, the host
 * draws the offsets in the reference's order and this entry point writes the whole [max_len, batch, dim] tensor,
 * padding included, in one launch:
 *   out[t][k][:] = chunk[beg[k] + t - left[k]][:]  for left[k] <= t < left[k] + len[k], else 0.
 * beg / len / left: DEVICE arrays of `batch` entries (first chunk row, length and leading zero rows of sentence k).
 * Rows that would fall outside the chunk are written as zeros (the host wrapper checks the tables before upload).  */
int hcgs_assemble_batch(const float* chunk, int64_t chunk_rows, int dim, const int64_t* beg, const int32_t* len,
                        const int32_t* left, int batch, int max_len, float* out, hcgs_stream stream);

/* hcgs_unpad_rows: the opposite movement after a batched forward (b200/export.py): rows t * batch + k, t < len[k], of
 * src [t_max * batch, ld_src] go to rows row_off[k] + t of dst [sum(len), cols] -- every utterance one contiguous
 * [len, cols] matrix, i.e. what core.py:238-249 hands to data_io.write_mat one utterance at a time; padding rows are
 * not copied.  row_off (int64) / len (int32): DEVICE arrays of `batch` entries.                                      */
int hcgs_unpad_rows(const float* src, int64_t ld_src, int cols, const int64_t* row_off, const int32_t* len, int batch,
                    int t_max, float* dst, hcgs_stream stream);

/* ---- guided HCGS masks on the device (guided_hcgs.py:9-76, guided_choices.py:4-30) -------------------------------------
 * hcgs_window_abs_mean: means[i] = mean |w| over the window {r0, r1, c0, c1} = windows[4i..4i+3] (half-open), accumulated
 *   in fp32 in row-major order and divided by the window area -- the arithmetic of the CPU AvgPool2d call the reference
 *   makes once per block-row strip on a host copy of the weights; the block ranking (np.argsort of a few hundred means)
 *   stays on the host, the weights never leave the device.  windows: DEVICE int32 [n_windows][4]; w: [rows, ldw] fp32.
 * hcgs_fill_rects: out[r0:r1, c0:c1] = value for every rectangle (the kept blocks of the last level -> the 0/1 mask).   */
int hcgs_window_abs_mean(const float* w, int64_t ldw, const int32_t* windows, int n_windows, float* means,
                         hcgs_stream stream);
int hcgs_fill_rects(float* out, int64_t ld, const int32_t* rects, int n_rects, float value, hcgs_stream stream);

#ifdef __cplusplus
}
#endif
#endif /* HCGS_B200_H */
