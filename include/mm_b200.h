/*
 * mm_b200.h — C-ABI of the B200-native modality-embedding hot path.
 *
 * One shared library (multimeditron_b200/csrc/libmm_b200.so) exports every symbol
 * declared here.  All pointers are raw DEVICE pointers owned by the caller, all
 * sizes are element counts unless named *_bytes, `stream` is a cudaStream_t passed
 * as void*.  No entry point synchronises, allocates, or keeps global mutable state
 * (the only process state is write-once: per-device cudaFuncSetAttribute flags and the
 * SM count, both behind atomics, and two environment switches read once at first use);
 * every call is re-entrant across host threads, streams and devices.  Scratch memory is
 * always caller-supplied, so two streams never share a buffer unless the caller hands
 * them the same one.  (One documented exception to all three: mm_attach_positions_static,
 * for callers that cannot allocate, keeps its scratch in a fixed array inside the library.)
 * Return value: MM_OK or an
 * mm_status error (host-detectable argument errors).  Data-dependent errors (an
 * index out of range, attachment-token count != number of modality rows) are
 * reported through the caller-supplied device `status` word (see mm_status_bits) so
 * that the path never forces a host sync; the host wrapper checks it lazily and
 * raises the same Python exception class the reference raises.
 *
 * Each entry point cites the reference interface (relative to /root/reference, or
 * HF: = transformers 5.5.0) it replaces.
 */
#ifndef MM_B200_H_
#define MM_B200_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef enum mm_status {
  MM_OK = 0,
  MM_ERR_BAD_ARG = 1,        /* shape/alignment/dtype combination not supported            */
  MM_ERR_CUDA = 2,           /* a CUDA runtime/driver call failed (mm_last_cuda_error())   */
  MM_ERR_UNSUPPORTED_ARCH = 3/* device is not sm_100                                       */
} mm_status;

/* bits OR-ed into the device status word by kernels */
enum mm_status_bits {
  MM_STATUS_INDEX_OOB = 1,      /* batch_idx/token_range outside [ -size, size )  -> IndexError   */
  MM_STATUS_COUNT_MISMATCH = 2  /* #attachment tokens != n_rows                  -> RuntimeError */
};

typedef enum mm_dtype { MM_BF16 = 0, MM_F32 = 1 } mm_dtype;

/* epilogue activation of mm_gemm_bias_act */
typedef enum mm_act {
  MM_ACT_NONE = 0,
  MM_ACT_QUICK_GELU = 1,  /* x*sigmoid(1.702x)   HF:activations.py:117-123 (CLIP MLP)          */
  MM_ACT_GELU_ERF = 2,    /* exact GELU          src/multimeditron/model/projectors/mlp.py:35,37 */
  MM_ACT_GELU_TANH = 3,   /* gelu_pytorch_tanh   HF:models/siglip (hidden_act)                 */
  MM_ACT_RELU = 4,        /* max(x, 0) applied LAST, i.e. AFTER the optional residual add: relu(bf16(acc+bias) + res) —
                             torchvision Bottleneck (`out += identity; out = relu(out)`), the MoE gate's ResNet-50   */
  /* per-call tuning flag OR-ed into `act` of mm_gemm_bias_act / mm_gemm_ln_bias_act: run the 1-CTA kernel
   * (128 x 256 tiles) instead of the default CTA-pair kernel (tcgen05 cta_group::2, 256 x 256 tiles, M > 128).
   * Both kernels accumulate every output in the same k order: results are bit-identical (tested).            */
  MM_GEMM_FLAG_1CTA = 0x100
} mm_act;

/* Library / build info. */
const char* mm_version(void);
int mm_last_cuda_error(void);       /* last cudaError_t seen by this thread (0 = none)         */
int mm_device_sm_count(int device); /* multiprocessor count, used to size grids                */

/* ------------------------------------------------------------------------------------------
 * Splice (bit-exact integer/byte work; HBM-bound)
 * ------------------------------------------------------------------------------------------ */

/*
 * mm_attach_positions — device prefix-sum of (input_ids == attachment_id).
 * Replaces the collator-side index build  src/multimeditron/model/data_loader.py:124-141 +
 * src/multimeditron/model/prompt_tokenizers.py:255-286 : the k-th attachment token in
 * row-major order of input_ids[B_seq,S] receives row k of the flattened projected stack.
 *   dest_rows[k]   = flat position (b*S+s) of the k-th attachment token     (int32, capacity n_rows_cap)
 *   src_of_pos[r]  = k if position r holds the k-th attachment token else -1 (int32, B_seq*S entries; may be NULL)
 *   count[0]       = total number of attachment tokens                        (int32)
 *   status        |= MM_STATUS_COUNT_MISMATCH when count != expected_rows (expected_rows >= 0)
 * workspace: at least mm_attach_positions_workspace_bytes(n_ids) bytes.
 */
int64_t mm_attach_positions_workspace_bytes(int64_t n_ids);
int mm_attach_positions(const int64_t* input_ids, int64_t n_ids, int64_t attachment_id,
                        int32_t* dest_rows, int64_t n_rows_cap, int32_t* src_of_pos,
                        int32_t* count, int64_t expected_rows, int32_t* status,
                        void* workspace, void* stream);

/*
 * mm_attach_positions_static — mm_attach_positions for callers that cannot allocate device memory.
 * No counterpart in the reference: it exists only because of how this library may be hosted.
 *
 * Who needs it: a host process that handed all of HBM to its own pool at start-up and may not call
 * cudaMalloc afterwards (inference servers commonly work this way), a plugin or foreign-language
 * binding that receives device pointers for the inputs and outputs but has no allocator of its own,
 * or any caller that wants this step to have a fixed, known memory footprint.  Such a caller cannot
 * produce the `workspace` argument of mm_attach_positions, so this variant drops it.
 *
 * Same inputs, outputs, status bits and results as mm_attach_positions (the same three kernels run).  The
 * per-block counters live in a fixed array inside the library image (MM_ATTACH_STATIC_MAX_IDS / 2048
 * int32 = 32 KB, 256-byte aligned) which every GPU that loads the library keeps until the process
 * exits.  The call itself allocates and frees nothing — no cudaMalloc, no cudaFree, no library
 * routine that allocates.
 *
 *   limit     n_ids <= MM_ATTACH_STATIC_MAX_IDS, otherwise MM_ERR_BAD_ARG and nothing is launched.
 *             Longer id vectors need mm_attach_positions with a caller-owned workspace.
 *   set-up    the first call on each device (the device that is current, which must be the one
 *             `stream` belongs to) looks up the array's device address with cudaGetSymbolAddress
 *             and remembers it.  Make that first call outside any stream capture.
 *   sharing   every call on a device uses the same array and the array keeps whatever the last
 *             call left in it.  Nothing needs to start at zero (each counter is written before it is
 *             read), so back-to-back calls on one stream are fine.  Two calls that can run at the
 *             same time on one GPU — different streams, or different host threads on different
 *             streams — race on the array and return garbage: order them yourself (one stream, or
 *             events), or give the concurrent ones mm_attach_positions with their own workspace.
 */
#define MM_ATTACH_STATIC_MAX_IDS 16777216 /* 2^24 ids, e.g. 2048 sequences of 8192 tokens */
int mm_attach_positions_static(const int64_t* input_ids, int64_t n_ids, int64_t attachment_id,
                               int32_t* dest_rows, int64_t n_rows_cap, int32_t* src_of_pos,
                               int32_t* count, int64_t expected_rows, int32_t* status, void* stream);

/*
 * mm_splice_scatter — embedded[dest_rows[k], :] = src[k, :]  (with dtype cast).
 * Replaces  src/multimeditron/model/model.py:441-444  (ATen index_put_, non-accumulate).
 * dest is [n_dest_rows, hidden] of dst_dtype, src is [n_rows, hidden] of src_dtype.
 * Rows with dest_rows[k] < 0 or >= n_dest_rows set MM_STATUS_INDEX_OOB and are skipped.
 */
int mm_splice_scatter(void* dest, int64_t n_dest_rows, int64_t hidden, int dst_dtype,
                      const void* src, int64_t n_rows, int src_dtype,
                      const int32_t* dest_rows, int32_t* status, void* stream);

/*
 * mm_splice_scatter_idx — same statement, taking the collator's own int64 index vectors
 * (processed_multimodal_inputs['batch_idx'/'token_range'], model.py:441-442) and doing the
 * b*S+t arithmetic (with Python negative-index wrap) on device.
 */
int mm_splice_scatter_idx(void* dest, int64_t batch, int64_t seq, int64_t hidden, int dst_dtype,
                          const void* src, int64_t n_rows, int src_dtype,
                          const int64_t* batch_idx, const int64_t* token_range,
                          int32_t* status, void* stream);

/*
 * mm_embed_splice_fused — one pass that produces inputs_embeds from ids:
 *   out[r,:] = src_of_pos[r] >= 0 ? src[src_of_pos[r], :] : table[input_ids[r], :]
 * Replaces  model.py:433 (nn.Embedding lookup) + model.py:441-444 (splice) in one kernel.
 */
int mm_embed_splice_fused(void* out, int64_t n_pos, int64_t hidden, int dst_dtype,
                          const void* table, int64_t vocab,
                          const int64_t* input_ids, const int32_t* src_of_pos,
                          const void* src, int64_t n_rows, int src_dtype,
                          int32_t* status, void* stream);

/* mm_embed_gather — out[r,:] = table[input_ids[r],:]   (model.py:433, ATen embedding). */
int mm_embed_gather(void* out, int64_t n_pos, int64_t hidden, int dtype, const void* table,
                    int64_t vocab, const int64_t* input_ids, int32_t* status, void* stream);

/* mm_embed_gather_skip — mm_embed_gather that leaves rows with input_ids[r] == skip_id untouched: the caller's
 * splice (mm_splice_scatter or the projector GEMM's scatter epilogue) writes every attachment position, so the
 * attachment token's own embedding row (model.py:433) is never observable and need not be copied. */
int mm_embed_gather_skip(void* out, int64_t n_pos, int64_t hidden, int dtype, const void* table,
                         int64_t vocab, const int64_t* input_ids, int64_t skip_id, int32_t* status,
                         void* stream);

/* mm_gather_rows — out[k,:] = src[rows[k],:]  (k < n_rows; rows int32, bounds-checked against n_src_rows ->
 * MM_STATUS_INDEX_OOB).  The backward of the splice statement model.py:441-444: the gradient of the projected
 * tokens is the gradient of inputs_embeds at their positions. */
int mm_gather_rows(void* out, int64_t n_rows, int64_t hidden, int dst_dtype, const void* src,
                   int64_t n_src_rows, int src_dtype, const int32_t* rows, int32_t* status, void* stream);

/* ------------------------------------------------------------------------------------------
 * ViT / projector building blocks (bf16 storage, fp32 accumulate)
 * ------------------------------------------------------------------------------------------ */

/*
 * mm_layernorm — y[m,:] = (x[m,:]-mean)*rstd*gamma+beta, fp32 statistics, bf16 in/out.
 * Replaces nn.LayerNorm call sites HF:models/clip/modeling_clip.py:371,380,677 and
 * HF:models/siglip/modeling_siglip.py:346,355,617.
 * Optional row remap: when rows_in_group>0, input row m is read from
 *   (m / rows_out_group) * rows_in_group + in_row_offset + (m % rows_out_group)
 * (used to drop the CLS row: image_modality.py:133).
 */
int mm_layernorm(const void* x, void* y, const void* gamma, const void* beta, int64_t M,
                 int64_t D, float eps, int64_t rows_in_group, int64_t rows_out_group,
                 int64_t in_row_offset, void* stream);

/*
 * mm_gemm_bias_act — C = act(A · Wᵀ + bias) [+ residual], tcgen05/TMEM GEMM fed by TMA.
 * Replaces nn.Linear (+activation, + residual add) call sites:
 *   HF:models/clip/modeling_clip.py:310-312 (q,k,v as one fused N=3D GEMM), :334 (+:377 residual),
 *   :347-351 (fc1+QuickGELU, fc2 +:382 residual); projectors/mlp.py:33-39.
 *   A [M,K] bf16 row-major (lda elements), W [N,K] bf16 row-major (nn.Linear.weight layout, ldw),
 *   bias [N] bf16 or NULL, C [*,N] bf16 (ldc).
 *   residual (optional, bf16, ldr): row used = (res_mod>0 ? m % res_mod : m) + res_row_off.
 *     The sum is formed as bf16(bf16(acc+bias) + residual), the reference's rounding points.
 *   out_row_map (optional, int32[M]): C row written for GEMM row m; <0 or >= n_out_rows = row not stored.
 *     (drop-CLS image_modality.py:133; splice fusion model.py:441-444)
 *   Requirements: K % 8 == 0, N % 8 == 0, 16-byte aligned pointers/leading dimensions.
 */
int mm_gemm_bias_act(const void* A, int64_t lda, const void* W, int64_t ldw, const void* bias,
                     void* C, int64_t ldc, int64_t M, int64_t N, int64_t K, int act,
                     const void* residual, int64_t ldr, int64_t res_mod, int64_t res_row_off,
                     const int32_t* out_row_map, int64_t n_out_rows, void* stream);

/*
 * mm_gemm_ln_bias_act — the same GEMM with LayerNorm folded in on either side, so that the encoder layer's two
 * LayerNorm passes (HF:models/clip/modeling_clip.py:370,379; siglip :341,350) cost no HBM traffic of their own:
 *     LN(x)·Wᵀ + b  =  rstd · (x·W''ᵀ) + b',    W'' = gamma ⊙ W with every row re-centred to sum to zero
 *                                               (so the −mean·Σ_k W'' term vanishes inside the MMA),
 *                                               b'_n = b_n + Σ_k beta_k·W[n,k]
 *   producer side  stats_out != NULL (act must be NONE): after bias/residual, every stored row also emits, per 128
 *                  output columns, (Σ v, Σ v²) of the bf16 values it stored: stats_out [rows of C][16] float2 — one
 *                  128-byte line per row, N <= 2048 (mm_gemm_row_stats_slots(N) = 16, or -1 when N is too wide).
 *                  The buffer must be ZERO-initialised once; slots beyond ceil(N/128) are never written.
 *   consumer side  ln_stats != NULL: A is the RAW residual stream x, W is W'' and `bias` is b' (both packed once at
 *                  load time, bf16; W'' is rounded with error feedback along k so that its bf16 row sums stay ~0),
 *                  the epilogue rebuilds rstd of row m from ln_stats [M][16] float2 (no residual in this mode; summed in a
 *                  fixed order: bit-reproducible) and forms bf16(rstd·acc + b').  Rounding then follows
 *                  mm_gemm_bias_act.
 * Numerics: the reference rounds LN(x) to bf16 before the GEMM; here gamma ⊙ W is rounded to bf16 instead and
 * LN itself stays in fp32 — same error scale (2^-9 relative per operand), covered by the stated bf16 tolerance.
 */
int64_t mm_gemm_row_stats_slots(int64_t N);

/*
 * mm_fold_layernorm — load-time weight preparation for the consumer side of mm_gemm_ln_bias_act (no reference
 * counterpart: the reference keeps nn.LayerNorm as its own kernel, HF:models/clip/modeling_clip.py:371,380).
 *   W_out[n,k] = bf16 of gamma[k]*W[n,k] - mean_k(gamma*W[n,:]), rounded with error feedback along k (the bf16 row
 *                sums stay within half an ulp of zero, so the LayerNorm mean term cancels inside the MMA)
 *   bias_out[n] = bf16(bias[n] + sum_k beta[k]*W[n,k])      (bias may be NULL)
 * W [N, K] bf16 row-major (ldw elements per row), gamma/beta [K] bf16.  One thread per row; K <= 12000.
 */
int mm_fold_layernorm(const void* W, int64_t ldw, const void* bias, const void* gamma, const void* beta,
                      void* W_out, int64_t ldo, void* bias_out, int64_t N, int64_t K, void* stream);

int mm_gemm_ln_bias_act(const void* A, int64_t lda, const void* W, int64_t ldw, const void* bias,
                        void* C, int64_t ldc, int64_t M, int64_t N, int64_t K, int act,
                        const void* residual, int64_t ldr, const int32_t* out_row_map,
                        int64_t n_out_rows, void* stats_out, const void* ln_stats, float ln_eps,
                        void* stream);
/*
 * mm_gemm_grouped — `groups` GEMMs of IDENTICAL shape in one persistent launch (the MoE modalities' experts:
 * image_modality_moe.py:160-162 runs every expert tower on every image, one after the other).  Everything is stacked:
 *   A [groups*M, K], C / residual [groups*M, N], stats [groups*M, slots] along the rows; W [groups*N, K] and bias
 *   [groups*N] along the rows of W.  Row tile r multiplies the weight block of group r / (M / tile_rows); M must be a
 *   multiple of the row tile (256 for the CTA-pair kernel, 128 with MM_GEMM_FLAG_1CTA), so no tile straddles two groups.
 * One launch walks groups*tiles tiles round-robin over the CTA pairs: 5 experts x 150 tiles fill 10.1 waves instead of
 * 5 x 2.03.  Options as mm_gemm_ln_bias_act (act, residual, out_row_map over the stacked rows, stats_out, ln_stats).
 */
int mm_gemm_grouped(const void* A, int64_t lda, const void* W, int64_t ldw, const void* bias, void* C, int64_t ldc,
                    int64_t M, int64_t N, int64_t K, int64_t groups, int act, const void* residual, int64_t ldr,
                    const int32_t* out_row_map, int64_t n_out_rows, void* stats_out, const void* ln_stats,
                    float ln_eps, void* stream);

/*
 * mm_patch_embed — im2col-free implicit GEMM of the patch convolution
 *   (Conv2d k=s=patch; HF:models/clip/modeling_clip.py:148-154,208-210, siglip :124-130,163-164)
 * fused with flatten/transpose, +position embedding and the CLS row (modeling_clip.py:210-218).
 *   pixels [B,3,img,img] fp32 or bf16 (cast to bf16 like :208-209),
 *   weight [D, ldw] bf16 = conv weight flattened to [D, 3*patch*patch] and zero-padded (once, at load
 *   time) to ldw >= round_up(3*patch*patch, 64) so that rows are 16-byte aligned for TMA,
 *   conv_bias [D] bf16 or NULL, pos_emb [T,D] bf16, cls [D] bf16 or NULL (NULL => no CLS row, T=P),
 *   out [B,T,D] bf16: out[b,0]=cls+pos[0]; out[b,cls_rows+p]=conv(patch p)+bias+pos[cls_rows+p].
 */
int mm_patch_embed(const void* pixels, int pixel_dtype, int64_t B, int64_t img, int64_t patch,
                   const void* weight, int64_t ldw, const void* conv_bias, const void* pos_emb,
                   const void* cls, void* out, int64_t D, void* stream);

/*
 * mm_patch_embed_ws — the same operator in two passes (the faster one; same inputs, outputs and rounding points):
 *   1. patch rows: pixels -> bf16 [B*P, Kpad] (k = (c, ky, kx), zero padded), each pixel read once from HBM;
 *   2. the TMA-fed CTA-pair GEMM (mm_gemm_bias_act's kernel) with the position-embedding / CLS-offset epilogue.
 * The in-kernel gather of mm_patch_embed re-reads every patch once per 256 output columns through 4-byte loads and
 * runs 8x slower at cfg2 shapes; it is kept for callers that cannot provide scratch memory.
 *   workspace: device scratch of at least mm_patch_embed_workspace_bytes(B, img, patch) bytes, 16-byte aligned
 *   (contents undefined afterwards).  Results are bit-identical to mm_patch_embed.
 */
int64_t mm_patch_embed_workspace_bytes(int64_t B, int64_t img, int64_t patch);
int mm_patch_embed_ws(const void* pixels, int pixel_dtype, int64_t B, int64_t img, int64_t patch,
                      const void* weight, int64_t ldw, const void* conv_bias, const void* pos_emb,
                      const void* cls, void* out, int64_t D, void* workspace, int64_t workspace_bytes,
                      void* stream);

/*
 * mm_attention_fwd — non-causal softmax(Q·Kᵀ·scale)·V for every (image, head), no mask,
 * tcgen05 QKᵀ / PV with TMEM accumulators and an online softmax in registers.
 * Replaces HF:models/clip/modeling_clip.py:318-331 and HF:models/siglip/modeling_siglip.py:285-300
 * (SDPA, HF:integrations/sdpa_attention.py:40).
 *   head_dim 64 (CLIP) or 72 (SigLIP so400m); T <= 1024.
 *   hs = mm_attention_head_stride(head_dim) is the column slot of one head in the qkv buffer: 64 -> 64,
 *   72 -> 80 (the 8 pad columns of every slot must be zeros: the fused QKV weight/bias is packed once, at
 *   load time, with zero rows there, so the QKV GEMM itself writes them).
 *   qkv [B*T, 3*heads*hs] bf16, columns [q | k | v], head h occupies columns h*hs.. of each third;
 *   out [B*T, heads*head_dim] bf16 (compact: pad columns are not stored).
 */
int64_t mm_attention_head_stride(int64_t head_dim); /* -1 when head_dim is not supported */
int mm_attention_fwd(const void* qkv, void* out, int64_t B, int64_t T, int64_t heads,
                     int64_t head_dim, float scale, void* stream);

/*
 * mm_weighted_expert_sum — out[b,p,:] = sum_e w[b, perm[e]] * x[e,b,p,:]
 * (MoE 'weighted_average' fusion: image_modality_moe.py:171-177, image_modality_moe_pep.py:212-216
 *  where perm is the identity).  x [E,B*P,D] bf16, w [B,E] fp32, out [B*P,D] bf16.
 */
int mm_weighted_expert_sum(const void* x, const float* w, const int32_t* perm, void* out,
                           int64_t E, int64_t B, int64_t P, int64_t D, void* stream);

/*
 * Row staging and pooling for the MoE gate's ResNet-50 (src/multimeditron/model/modalities/moe/gating.py:60-65,85 —
 * torchvision resnet50).  Every convolution runs as mm_gemm_bias_act on NHWC activations
 * ([B*H*W, C] bf16), BatchNorm folded into weight/bias at load time, MM_ACT_RELU in the epilogue.
 *   mm_conv_rows_nhwc : x [B,H,W,C] bf16 -> out [B*Ho*Wo, Kpad] bf16, k = (ky*kw + kx)*C + c, zeros outside the image
 *                       and in the K padding; Ho = (H + 2*pad - kh)/stride + 1 (3x3 and strided 1x1 convolutions)
 *   mm_conv_rows_stem : x [B,3,H,W] fp32|bf16 (NCHW pixel_values) -> out [B*Ho*Wo, Kpad], k = (ky*kw + kx)*3 + c
 *   mm_maxpool_nhwc   : 3x3, stride 2, pad 1 (-inf padding), bf16 [B,H,W,C] -> [B,Ho,Wo,C]
 *   mm_avgpool_nhwc   : mean over the HW pixels, fp32 accumulation: [B,HW,C] -> [B,C] bf16
 */
int mm_conv_rows_nhwc(const void* x, void* out, int64_t B, int64_t H, int64_t W, int64_t C, int64_t kh, int64_t kw,
                      int64_t stride, int64_t pad, int64_t Kpad, void* stream);
int mm_conv_rows_stem(const void* x, int x_dtype, void* out, int64_t B, int64_t H, int64_t W, int64_t kh, int64_t kw,
                      int64_t stride, int64_t pad, int64_t Kpad, void* stream);
int mm_maxpool_nhwc(const void* x, void* out, int64_t B, int64_t H, int64_t W, int64_t C, void* stream);
int mm_avgpool_nhwc(const void* x, void* out, int64_t B, int64_t HW, int64_t C, void* stream);

/*
 * mm_cross_attention — the MoE 'cross_attn' fusion's attention: generalist queries over the
 * concatenated, gate-weighted specialist tokens.  Replaces src/multimeditron/model/attention.py:48-101
 * (CrossAttention.forward between its q/k/v projections and its output projection; eval mode, no mask — the call sites
 * image_modality_moe.py:205 and image_modality_moe_pep.py:245 pass none).
 *   q [B, Nq, C] bf16; kv [B, Nkv, 2C] bf16 with columns [K | V] (one fused k_proj|v_proj GEMM); out [B, Nq, C] bf16;
 *   head_dim = C / heads must be a multiple of 32 and <= 512; scale = head_dim**-0.5 (attention.py:33).
 */
int mm_cross_attention(const void* q, const void* kv, void* out, int64_t B, int64_t Nq, int64_t Nkv,
                       int64_t heads, int64_t C, float scale, void* stream);

/*
 * mm_position_ids — the collator's position ids on the device (bit-exact integer work).
 * Replaces src/multimeditron/model/data_loader.py:155-188:
 *   use_2d == 0:  out [batch, seq] int64 = cumsum(attention_mask) - 1, 0 where the mask is 0            (:156-157)
 *   use_2d != 0:  out [batch, seq, 2] int64: both components start as the 1-D ids; every image (a run of
 *                 tokens_per_image = grid*grid attachment tokens, found from input_ids) gets (row, col) grid ids
 *                 offset by the id at its first token, and everything behind it — padding included — is shifted by
 *                 (largest id of the image's last token) - (1-D id of its last token, or 0 when the image opens the
 *                 row: the reference's `token_range[0] > 0` clause, :178)                                 (:159-186)
 *   attention_mask: int64 (HF) or uint8/bool (mask_is_u8).  At most 256 images per row; more, or an image block
 *   that runs past the end of the row, sets MM_STATUS_COUNT_MISMATCH in `status` (may be NULL).
 */
int mm_position_ids(const int64_t* input_ids, const void* attention_mask, int mask_is_u8, int64_t batch,
                    int64_t seq, int64_t attachment_id, int64_t tokens_per_image, int64_t grid, int use_2d,
                    int64_t* out, int32_t* status, void* stream);

/* ------------------------------------------------------------------------------------------
 * Image pre-processing on the device — bit-exact byte/integer work, HBM-bound
 * ------------------------------------------------------------------------------------------ */

/*
 * mm_preprocess_images — raw uint8 HWC images (ragged batch) -> the pixel tensor the image embedder consumes.
 * Replaces the CPU work of `ImageProcessor.process` (src/multimeditron/model/modalities/image_modality.py:96):
 * transformers 5.5.0 CLIPImageProcessor / SiglipImageProcessor on the torchvision backend
 * (HF:image_processing_backends.py:372-420): uint8 antialiased bicubic resize (ATen's separable int16-weight
 * kernel: horizontal pass, uint8 intermediate, vertical pass), center crop, (x - mean*255) / (std*255) in fp32;
 * with out_dtype MM_BF16 the fp32 value is rounded RNE exactly as HF:models/clip/modeling_clip.py:208-209 casts it.
 *   src    device bytes, 4-byte aligned; image i is [H,W,3] uint8 at src + desc[i][0]
 *   desc   device int64 [B][8] = {src_off, H, W, out_h, out_w, crop_top, crop_left, tmp_off}: the image is resized
 *          to (out_h,out_w) and the S x S window at (crop_top,crop_left) is produced (out_h,out_w >= S);
 *          tmp_off = byte offset of this image's [H,S,3] intermediate inside `tmp`
 *   kmax   >= 2*ceil(2*max(1, in/out)) + 1 over every image and axis (taps per output pixel)
 *   max_h, max_w   largest H / W in the batch (grid and shared-memory sizing)
 *   ws     mm_preprocess_workspace_bytes(B,S,kmax) device bytes; tmp  sum_i H_i*S*3 device bytes
 *   mean255, std255   HOST float[3]: image_mean*(1/rescale_factor), image_std*(1/rescale_factor) in fp32
 *   out    [B,3,S,S] fp32 (MM_F32, the reference's `pixel_values`) or bf16 (MM_BF16)
 */
int64_t mm_preprocess_workspace_bytes(int64_t B, int64_t S, int64_t kmax);
int mm_preprocess_images(const void* src, const int64_t* desc, int64_t B, int64_t S, int64_t kmax,
                         int64_t max_h, int64_t max_w, void* ws, void* tmp, const float* mean255,
                         const float* std255, void* out, int out_dtype, void* stream);

/* ------------------------------------------------------------------------------------------
 * Projector training step (reference: freeze_for_alignment, model.py:310-326,
 * MLPProjector projectors/mlp.py:33-39 under torch.autograd).  The backward GEMMs are mm_gemm_bias_act calls on
 * transposed operands:  dX = dY·W  ->  A = dY [M,N], "W" = Wᵀ [K,N];   dW = dYᵀ·X -> A = dYᵀ [N,M], "W" = Xᵀ [K,M].
 * ------------------------------------------------------------------------------------------ */

/* a = GELU_erf(z), element-wise bf16 (n % 8 == 0); bit-identical to the fused MM_ACT_GELU_ERF epilogue. */
int mm_gelu_fwd(const void* z, void* a, int64_t n, void* stream);

/* bytes of fp32 scratch mm_gelu_bwd_colsum needs for an [M, N] operand */
int64_t mm_colsum_workspace_bytes(int64_t M, int64_t N);

/* dz = da * GELU_erf'(z) (bf16 [M,N], fp32 math, one rounding) and bias_grad[n] = sum_m dz[m,n] (fp32 [N],
 * deterministic two-stage sum over the bf16 values as stored).  z == dz == NULL: column sum of `da` only. */
int mm_gelu_bwd_colsum(const void* da, const void* z, void* dz, float* bias_grad, void* workspace, int64_t M,
                       int64_t N, void* stream);

/* out [C, R] = in [R, C]ᵀ, bf16, R % 8 == C % 8 == 0. */
int mm_transpose(const void* in, void* out, int64_t R, int64_t C, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* MM_B200_H_ */
