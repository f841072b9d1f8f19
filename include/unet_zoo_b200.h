/*
 * unet_zoo_b200 — C ABI of the B200-native (sm_100a) kernel library behind the UNet hot path of
 * irfanfadhullah/unet_zoo.
 *
 * The reference has NO FFI / plugin / operator-registration layer of its own: every arithmetic op is a stock
 * torch.nn module call.  Each entry point below therefore cites the torch.nn call site in the reference that it
 * replaces (paths relative to the reference repo root).
 *
 * Conventions
 *  - Plain pointers and sizes only; all pointers are DEVICE pointers unless stated otherwise.
 *  - Activations are NHWC bf16 ("pixels x channels"); `ld` is the pixel stride in elements, so a channel slice of a
 *    wider buffer is addressed as (ptr + c_off, ld = C_total).  Channel counts and ld must be multiples of 8.
 *  - A virtual channel concat (torch.cat(..., 1)) is a list of segments (parallel arrays seg_ptr/seg_c/seg_ld).
 *  - Parameters and parameter gradients stay in the reference's fp32 layouts (Conv2d OIHW, ConvTranspose2d IOHW,
 *    BatchNorm [C]); the tensor-core kernels consume bf16 copies made by uz_pack_weight.
 *  - The caller owns every buffer including workspaces (uz_*_workspace_bytes); the library allocates nothing and
 *    keeps no mutable global state.  Calls are stream-ordered on `stream` (a cudaStream_t) with no implicit syncs,
 *    and may be issued from any host thread that has the right device current.
 *  - Return value: 0 on success, negative on error; uz_last_error() describes the last failure on this thread.
 *  - There is no CPU path: every function fails with UZ_ERR_CUDA if no sm_100 device is current.
 */
#ifndef UNET_ZOO_B200_H_
#define UNET_ZOO_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define UZ_OK 0
#define UZ_ERR_INVALID (-1)
#define UZ_ERR_UNSUPPORTED (-2)
#define UZ_ERR_CUDA (-3)
#define UZ_ERR_WORKSPACE (-4)

const char* uz_last_error(void);
int uz_version(void);
/* Number of SMs of the current device (grid sizing of the persistent kernels); <0 on error. */
int uz_device_sm_count(void);

/* ------------------------------------------------------------------------------------------------------------
 * Weight packing: fp32 parameter -> bf16 operand
 *   dst[a*dst_pitch + t*B + b] = bf16(src[a*sa + b*sb + t*st] * (row_scale ? row_scale[a] : 1))   (a<A, t<T, b<B)
 *   Conv2d fwd  (OIHW [Co,Ci,kh,kw])      : A=Co T=kh*kw B=Ci  sa=Ci*T sb=T  st=1
 *   Conv2d dgrad                           : A=Ci T=kh*kw B=Co  sa=T    sb=Ci*T st=1
 *   ConvTranspose2d fwd (IOHW [Ci,Co,2,2]) : A=Co T=4     B=Ci  sa=4    sb=Co*4 st=1
 *   ConvTranspose2d dgrad                  : A=Ci T=4     B=Co  sa=Co*4 sb=4   st=1
 *   stem (im2col'd 3x3, K = 9*Ci padded)   : A=Co T=9     B=Ci  sa=Ci*9 sb=9   st=1  dst_pitch=kp (pad zeroed by caller)
 * dst_pitch >= T*B (elements).  row_scale folds an inference-mode BatchNorm scale into the output-channel rows.
 * Replaces: the implicit weight reads of nn.Conv2d / nn.ConvTranspose2d (common_layers.py:28,31,104). */
int uz_pack_weight(const float* src, void* dst_bf16, int A, int T, int B, long long sa, long long sb, long long st,
                   long long dst_pitch, const float* row_scale, void* stream);
/* All weights of a model in one launch: jobs (DEVICE memory) = njobs x {src, dst, A, T, B, sa, sb, st, dst_pitch, first}
 * as 64-bit integers - the arguments of uz_pack_weight (no row scale) plus `first`, the number of (a, b) pairs of all
 * earlier jobs; total_pairs = sum of A * B.  A training step otherwise spends one ~5 us launch per convolution and layout. */
int uz_pack_weights_batched(const long long* jobs_device, int njobs, long long total_pairs, void* stream);

/* ------------------------------------------------------------------------------------------------------------
 * Input staging.  The reference feeds NCHW fp32 images (training_loop.py:109, unet.py:31).
 *  uz_stem_im2col : out[n,y,x, t*cin+ci] = x[n,ci,y+r-1,x+s-1] (t=r*3+s, zero outside / for columns >= 9*cin), NHWC
 *                   bf16 with kp in {16,32,64} channels; the low-channel stem conv then runs as a 1x1 GEMM over it
 *                   (common_layers.py:28 with in_channels=3; resunet.py:26,32).
 *  uz_nchw_to_nhwc: layout + dtype change with zero channel padding up to a multiple of 8 (in_channels >= 8). */
int uz_stem_im2col(const float* x, int N, int cin, int H, int W, void* out, int kp, void* stream);
int uz_nchw_to_nhwc(const float* x, int N, int C, int H, int W, void* out, int out_ld, void* stream);

/* ------------------------------------------------------------------------------------------------------------
 * BatchNorm2d, training mode (common_layers.py:29,32; attention_unet.py:13,20,27; nested_unet.py:8,10).
 *  partial statistics: fp32 [rows][2][C] (sum, sum of squares) - produced by uz_conv_igemm(UZ_FLAG_STATS) with
 *  rows = uz_conv_stats_rows(), or by uz_bn_stats for an arbitrary tensor with rows = uz_reduce_rows().
 *  uz_bn_finalize: bn[4][C] = {mean, invstd, scale = gamma*invstd, shift = beta - mean*scale}; running_mean /
 *    running_var updated exactly like nn.BatchNorm2d (EMA with `momentum`, unbiased variance), may be NULL.
 *  uz_bn_apply   : y = act(z*scale + shift [+ residual]); `pooled` (optional) receives MaxPool2d(2,2)(y)
 *    (common_layers.py:90,94); y may be NULL when only the pooled tensor is wanted; scale==shift==NULL -> identity. */
int uz_reduce_rows(void);
int uz_bn_stats(const void* x, int ld, long long npix, int C, float* partial, void* stream);
int uz_bn_finalize(const float* partial, int rows, int C, long long count, const float* gamma, const float* beta,
                   float eps, float momentum, float* running_mean, float* running_var, float* bn,
                   long long* num_batches_tracked /* int64 device scalar, += 1 (nn.BatchNorm2d's counter); may be NULL */,
                   void* stream);
int uz_bn_apply(const void* z, int z_ld, int N, int H, int W, int C, const float* scale, const float* shift, int relu,
                const void* residual, int res_ld, void* y, int y_ld, void* pooled, int pooled_ld, void* stream);

/* Backward of y = [ReLU](BN(z)) fused with the MaxPool2d(2,2) gradient routing:
 *   g  = (dy + route(dpool)) * [relu ? (y > 0) : 1]       route(): first maximum of each 2x2 window (row-major scan)
 *   dz = gamma*invstd * (g - mean(g) - x^ * mean(g*x^)),  dgamma = sum g*x^,  dbeta = sum g,  dbias = sum dz
 * dy and dpool are each optional (not both NULL).  dgamma/dbeta/dbias are fp32 [C] (NULL to skip); `accumulate`
 * bit 0 adds into them, bit 1 makes dz accumulate (dz += ..., for tensors with several consumers; not with dbias).  Replaces autograd through BatchNorm2d/ReLU/MaxPool2d for the call sites above. */
size_t uz_bn_bwd_workspace_bytes(int C);
int uz_bn_bwd(const void* z, int z_ld, const void* dy, int dy_ld, const void* dpool, int dpool_ld, int N, int H, int W,
              int C, const float* bn, int relu, void* dz, int dz_ld, float* dgamma, float* dbeta, float* dbias,
              int accumulate, void* workspace, size_t workspace_bytes, void* stream);

/* out[c] (+)= sum over pixels x[p, c]  (bias gradient of ConvTranspose2d / plain Conv2d). */
size_t uz_colsum_workspace_bytes(int C);
int uz_colsum(const void* x, int ld, long long npix, int C, float* out, int accumulate, void* workspace,
              size_t workspace_bytes, void* stream);
/* uz_colsum_static: uz_colsum WITHOUT a workspace argument, for callers that cannot allocate one (CUDA graph capture,
 * serving loops, harnesses that forbid cudaMalloc).  No counterpart in the reference.
 * The per-block partial sums live in a file-scope device array inside the library (2.5 MiB per device that
 * loads the library, held until exit); the call performs no cudaMalloc / cudaFree / allocating library call.
 * Limit: C <= uz_colsum_static_max_channels() (1024); larger C returns UZ_ERR_WORKSPACE - use uz_colsum.
 * All calls on one device share the array: calls that OVERLAP on the GPU (different streams, no ordering) race;
 * calls ordered on one stream, or by events, are safe.  Nothing in the array needs to start at zero. */
int uz_colsum_static_max_channels(void);
int uz_colsum_static(const void* x, int ld, long long npix, int C, float* out, int accumulate, void* stream);

/* ------------------------------------------------------------------------------------------------------------
 * 1x1 classifier head: logits[n,k,h,w] (fp32, NCHW) = b[k] + sum_c w[k,c] * y[n,h,w,c]   (no activation)
 * Replaces OutConv (common_layers.py:125-128), attention_unet.py:70, nested_unet.py:61-69, resunet.py:49-52.
 * Backward: dy[p,c] = sum_k dlogits[p,k] w[k,c] (bf16), dw[k,c] = sum_p dlogits[p,k] y[p,c], db[k] = sum_p dlogits. */
int uz_head_fwd(const void* y, int ld, int N, int H, int W, int C, int K, const float* w, const float* b,
                float* logits, void* stream);
size_t uz_head_bwd_workspace_bytes(int C);
int uz_head_bwd(const void* y, int ld, const float* dlogits, int N, int H, int W, int C, int K, const float* w,
                void* dy, int dy_ld, float* dw, float* db, int accumulate, void* workspace, size_t workspace_bytes,
                void* stream);

/* ------------------------------------------------------------------------------------------------------------
 * Last DoubleConv half fused with the classifier head (training): the full-resolution activation y = ReLU(BN(z))
 * between DoubleConv (common_layers.py:28-33) and OutConv (common_layers.py:125-128; attention_unet.py:70,
 * nested_unet.py:69) is never written to memory.
 *  uz_bn_head_fwd: logits[n,k,h,w] = b[k] + sum_c w[k,c] * act(scale[c]*z[n,h,w,c] + shift[c])   (fp32 NCHW)
 *  uz_bn_head_bwd: given dlogits (fp32 NCHW) rebuilds y and dy[p,c] = sum_k dlogits[p,k] w[k,c] on the fly and runs the
 *                  BatchNorm backward of uz_bn_bwd on it: dz (bf16, may alias z), dgamma, dbeta, dbias (conv bias in
 *                  front of the BN), and the head's dw[k,c] = sum_p dlogits[p,k] y[p,c], db[k] = sum_p dlogits[p,k].
 *                  bn = [4][C] {mean, invstd, scale, shift} of uz_bn_finalize; accumulate bit0: += into the parameter
 *                  gradients.  C a power of two in [8, 256], K in [1, 4]. */
size_t uz_bn_head_workspace_bytes(int C, int K);
int uz_bn_head_fwd(const void* z, int z_ld, int N, int H, int W, int C, const float* scale, const float* shift,
                   int relu, int K, const float* w, const float* b, float* logits, void* stream);
int uz_bn_head_bwd(const void* z, int z_ld, int N, int H, int W, int C, const float* bn, int relu, int K,
                   const float* w, const float* dlogits, void* dz, int dz_ld, float* dgamma, float* dbeta,
                   float* dbias, float* dw, float* db, int accumulate, void* workspace, size_t workspace_bytes,
                   void* stream);

/* ------------------------------------------------------------------------------------------------------------
 * Additive attention gate (AttentionBlock, attention_unet.py:6-40): the two 1x1 convolutions W_g, W_x run through
 * uz_conv_igemm(UZ_CONV1X1, UZ_FLAG_STATS) -> zg, zx (bf16 [npix, F], F = f_int <= 256) and uz_bn_finalize -> bn_g,
 * bn_x ([4][F] = mean, invstd, scale, shift).  Then
 *  uz_gate_psi_fwd: psi_raw[p] = b_psi + sum_c w_psi[c] * ReLU(BN_g(zg)[p,c] + BN_x(zx)[p,c])   (fp32 [npix])
 *                   partial[uz_reduce_rows()][2][1] = sum / sum of squares of psi_raw -> uz_bn_finalize(C = 1) -> psi_bn[4]
 *  uz_gate_apply  : out[p, c] = x[p, c] * sigmoid(psi_raw[p] * psi_bn.scale + psi_bn.shift)       (attention_unet.py:38-40)
 *  uz_gate_bwd    : from dout = dL/dout:  dx (+)= a * dout;  through sigmoid, BN_psi (dgamma_psi, dbeta_psi), w_psi
 *                   (dw_psi [F], db_psi [1]) and the ReLU into dr (bf16 [npix, F]) - the common dL/d(g1) = dL/d(x1),
 *                   to be fed to uz_bn_bwd(zg, dy = dr, relu = 0) and uz_bn_bwd(zx, dy = dr, relu = 0).
 * Inference: pass running-statistics scale/shift in bn_g, bn_x, psi_bn. */
int uz_gate_psi_fwd(const void* zg, int zg_ld, const void* zx, int zx_ld, long long npix, int F, const float* bn_g,
                    const float* bn_x, const float* w_psi, const float* b_psi, float* psi_raw, float* partial,
                    void* stream);
int uz_gate_apply(const void* x, int x_ld, const float* psi_raw, const float* psi_bn, long long npix, int C, void* out,
                  int out_ld, void* stream);
size_t uz_gate_bwd_workspace_bytes(long long npix, int F);
int uz_gate_bwd(const void* x, int x_ld, const void* dout, int dout_ld, const float* psi_raw, const float* psi_bn,
                long long npix, int C, void* dx, int dx_ld, int accumulate_dx, const void* zg, int zg_ld,
                const void* zx, int zx_ld, int F, const float* bn_g, const float* bn_x, const float* w_psi, void* dr,
                int dr_ld, float* dw_psi, float* db_psi, float* dgamma_psi, float* dbeta_psi, void* workspace,
                size_t workspace_bytes, void* stream);

/* ------------------------------------------------------------------------------------------------------------
 * nn.Upsample(scale_factor=2, mode='bilinear', align_corners=True) (nested_unet.py:33,81-93).
 *  fwd: x [N,H,W,C] -> out [N,2H,2W,C] written at (out, out_ld) - typically a channel slice of a concat buffer.
 *  bwd: dx [N,H,W,C] (+)= gather of dout [N,2H,2W,C]; H, W are the LOW-resolution sizes in both calls. */
int uz_bilinear2x_fwd(const void* x, int x_ld, int N, int H, int W, int C, void* out, int out_ld, void* stream);
/* Attention bottlenecks of the neighbouring models (transatt_unet.py:26-107, unet_transformer.py:115-229).
 * uz_bmm_tn: C[n][m][j] = sum_k A[n][m][k] * B[n][j][k], bf16 operands (both K-contiguous: A rows lda, B rows K), fp32
 *   accumulation, bf16 output rows ldc - replaces torch.bmm / torch.matmul.  Image strides in elements.  Shapes with
 *   M, K, J multiples of 8 run on the tcgen05 1x1 kernel with a per-image B operand; others on a plain kernel.
 * uz_softmax_rows / _bwd: y = softmax(scale * x) over the last axis (F.softmax(dim=-1), the temperature division folded in
 *   as `scale`) and dx = scale * y * (dy - sum(dy * y)).
 * uz_transpose_bf16: dst[n][c][r] = src[n][r][c] (the TN product wants K-contiguous operands). */
int uz_bmm_tn(const void* A, long long lda, long long a_img, const void* B, long long b_img, void* C, long long ldc,
              long long c_img, int N, int M, int K, int J, void* stream);
int uz_softmax_rows(const void* x, long long x_ld, void* y, long long y_ld, long long rows, int cols, float scale,
                    void* stream);
int uz_softmax_rows_bwd(const void* y, long long y_ld, const void* dy, long long dy_ld, void* dx, long long dx_ld,
                        long long rows, int cols, float scale, void* stream);
/* softmax over the ROW index of every column of N contiguous [rows][cols] bf16 matrices: nn.Softmax(dim=1) on the
 * [b, queries, keys] scores of unet_transformer.py:122,130,216 (normalised over the queries), and its backward. */
int uz_softmax_cols(const void* x, void* y, int N, int rows, int cols, float scale, void* stream);
int uz_softmax_cols_bwd(const void* y, const void* dy, void* dx, int N, int rows, int cols, float scale, void* stream);
int uz_transpose_bf16(const void* src, long long src_ld, long long src_img, void* dst, long long dst_ld, long long dst_img,
                      int N, int R, int C, void* stream);

/* Bandwidth operators around unet_transformer's attention blocks (unet_zoo/models/unet_transformer.py), NHWC bf16:
 * uz_addpe_fwd / _bwd: y = x + pe with pe a fixed fp32 [H][W][C] table (PositionalEncodingPermute2D, :68-113; built once
 *   on the host side), optionally followed by MaxPool2d(2) of the sum without storing it (Sconv_process[0], :153).
 *   Backward: dx (+)= dy + the pooled gradient routed to the first maximum of bf16(x + pe) per 2x2 window.
 * uz_adaptive_avgpool_fwd / _bwd: F.adaptive_avg_pool2d (:201-203), ATen windows [floor(o*I/O), ceil((o+1)*I/O)).
 * uz_bilinear_resize_fwd / _bwd: F.interpolate(size=(OH, OW), mode="bilinear", align_corners=...) and its exact adjoint in
 *   gather form (no atomics): align_corners = 1 for unet_transformer.py:222, 0 for u2net.py:19-23 (_upsample_like). */
int uz_addpe_fwd(const void* x, int x_ld, const float* pe, int N, int H, int W, int C, void* y, int y_ld, void* pooled,
                 int pooled_ld, void* stream);
int uz_addpe_bwd(const void* x, int x_ld, const float* pe, int N, int H, int W, int C, const void* dy, int dy_ld,
                 const void* dpooled, int dpooled_ld, void* dx, int dx_ld, int accumulate, void* stream);
int uz_adaptive_avgpool_fwd(const void* x, int x_ld, int N, int H, int W, int C, void* out, int out_ld, int OH, int OW,
                            void* stream);
int uz_adaptive_avgpool_bwd(const void* dout, int dout_ld, int N, int H, int W, int C, int OH, int OW, void* dx, int dx_ld,
                            int accumulate, void* stream);
int uz_bilinear_resize_fwd(const void* x, int x_ld, int N, int H, int W, int C, void* out, int out_ld, int OH, int OW,
                           int align_corners, void* stream);
size_t uz_bilinear_resize_bwd_workspace_bytes(int N, int H, int W, int C, int OH, int OW); /* 0 below 4x magnification */
int uz_bilinear_resize_bwd(const void* dout, int dout_ld, int N, int H, int W, int C, int OH, int OW, void* dx, int dx_ld,
                           int accumulate, int align_corners, void* workspace, size_t workspace_bytes, void* stream);

/* Residual add of U^2-Net's RSU blocks, `return hx1d + hxin` (u2net.py:78,127,171,210,237), and - with relu = 1 - the
 * `x = x + temp; x = F.relu(x)` of MultiResUnet's Multiresblock / Respath (multiresunet.py:82-83,125-126,132-133), NHWC bf16:
 * uz_residual_add_fwd: y = [relu](a + b); `pooled` (optional) receives MaxPool2d(2, stride 2) of y (u2net.py:222-232).
 * uz_residual_add_bwd: g = dy + the pooled gradient routed to the first maximum of the stored y per 2x2 window, masked
 *   by y > 0 when relu; da (+)= g and db (+)= g (acc_a / acc_b select accumulation; either incoming gradient may be
 *   NULL, not both; y is read when dpooled is given or relu is set). */
int uz_residual_add_fwd(const void* a, int a_ld, const void* b, int b_ld, int N, int H, int W, int C, void* y, int y_ld,
                        void* pooled, int pooled_ld, int relu, void* stream);
int uz_residual_add_bwd(const void* y, int y_ld, int N, int H, int W, int C, const void* dy, int dy_ld, const void* dpooled,
                        int dpooled_ld, void* da, int da_ld, int acc_a, void* db, int db_ld, int acc_b, int relu,
                        void* stream);

/* Two train-mode BatchNorms, add and ReLU in one pass from the two RAW (pre-BatchNorm) tensors, NHWC bf16:
 *   y = relu( actA(a * scaleA + shiftA) + (b * scaleB + shiftB) ),  actA = ReLU when relu_a.
 * Replaces `x = self.batch_norm1(cat); x = x + temp; x = F.relu(x)` of Multiresblock (multiresunet.py:79-83; a = the
 * concatenation, b = the 1x1 shortcut convolution's output) and `x = F.relu(conv3x3_bn_relu(x) + conv1x1_bn(x))` of
 * Respath (multiresunet.py:122-126,130-133; relu_a = 1); neither normalised branch is stored.
 * bnA / bnB: [4][C] {mean, invstd, scale, shift} of uz_bn_finalize.  y_stats (optional): [uz_reduce_rows()][2][C] partial
 * sums / sums of squares of the stored y, i.e. what uz_bn_stats(y) would produce - for the BatchNorm that follows.
 * uz_dual_bn_add_relu_bwd: g = dy * [y > 0] (y recomputed from a and b), gA = g * [actA'], gB = g; both BatchNorm
 *   backwards from one reduce pass and one apply pass: dA, dB (bf16; may alias a / b; accA / accB: += into the
 *   destination), dbiasA / dbiasB (optional, overwritten): gradient of a convolution bias in front of the BatchNorm.
 *   bn_outer / coef_outer (optional, together): y is followed by out = BatchNorm_o(y) (multiresunet.py:84,127,134); then
 *   `dy` is d(out) and the BatchNorm_o backward dy' = scale_o * d(out) + coef_outer[0] + coef_outer[1] * y is applied on
 *   the fly to the recomputed y (coef_outer from uz_bn_bwd_coef) - d(y) is never stored. */
size_t uz_dual_bn_bwd_workspace_bytes(int C);
int uz_dual_bn_add_relu_fwd(const void* a, int a_ld, const float* bnA, int relu_a, const void* b, int b_ld,
                            const float* bnB, int N, int H, int W, int C, void* y, int y_ld, float* y_stats, void* stream);
int uz_dual_bn_add_relu_bwd(const void* a, int a_ld, const float* bnA, int relu_a, const void* b, int b_ld,
                            const float* bnB, int N, int H, int W, int C, const void* dy, int dy_ld,
                            const float* bn_outer, const float* coef_outer, void* dA, int dA_ld, int accA, float* dbiasA,
                            void* dB, int dB_ld, int accB, float* dbiasB, void* workspace, size_t workspace_bytes,
                            void* stream);
/* First half of uz_bn_bwd (no pooling): the reduce pass and the finalize, leaving coef[2][C] = {A, B} with
 * dz = scale * g + A + B * z for a consumer that applies it on the fly (uz_dual_bn_add_relu_bwd's bn_outer / coef_outer);
 * dgamma / dbeta / dbias as in uz_bn_bwd (optional).  Workspace: uz_bn_bwd_workspace_bytes(C). */
int uz_bn_bwd_coef(const void* z, int z_ld, const void* dy, int dy_ld, int N, int H, int W, int C, const float* bn, int relu,
                   float* coef, float* dgamma, float* dbeta, float* dbias, int accumulate, void* workspace,
                   size_t workspace_bytes, void* stream);

/* Channels [0, K), K <= 8, of an NHWC bf16 map as fp32 NCHW logits [N][K][H][W]: the output of MultiResUnet's conv_final
 * (Conv2d 1x1 -> BatchNorm2d(affine=False), no activation; multiresunet.py:200-201,238), which runs as a conv-BN pair
 * with the classes zero-padded to 8 channels.  The way back (fp32 NCHW gradient -> NHWC bf16) is uz_nchw_to_nhwc. */
int uz_nhwc_to_nchw_f32(const void* x, int ld, int N, int H, int W, int K, float* out, void* stream);

/* Output stage of U2NET / U2NETP (u2net.py:273-295).  `fuse`: NHWC bf16, side map i (already at full resolution) in
 * channels [8 i, 8 i + K), i < nsides <= 8, K <= 4 classes; w = outconv.weight [K][nsides * K], b = outconv.bias [K].
 * uz_side_head_fwd: logits fp32 NCHW [N][(1 + nsides) K][H][W]: rows [0, K) = outconv(cat(sides)) ('main'), rows
 *   [K (1 + i), K (2 + i)) = side map i ('side<i+1>').
 * uz_side_head_bwd: dfuse slot i = dlogits[side i] + outconv^T dlogits[main] (channels K..7 of a slot: 0),
 *   dw [K][nsides * K], db [K] (overwritten). */
size_t uz_side_head_bwd_workspace_bytes(int nsides, int K);
int uz_side_head_fwd(const void* fuse, int ld, int N, int H, int W, int nsides, int K, const float* w, const float* b,
                     float* logits, void* stream);
int uz_side_head_bwd(const void* fuse, int ld, const float* dlogits, int N, int H, int W, int nsides, int K, const float* w,
                     void* dfuse, int dfuse_ld, float* dw, float* db, void* workspace, size_t workspace_bytes, void* stream);

/* F.pad with zeros on the bottom / right, or its backward (crop), between two NHWC pixel grids: replaces the
 * `F.pad(x1, [dx // 2, dx - dx // 2, dy // 2, dy - dy // 2])` of common_layers.py:110-113 / :173-178 (dx, dy in {0, 1} for
 * the odd input sizes the reference accepts).  dst[n,y,x,:] = (y < Hs && x < Ws) ? src[n,y,x,:] : 0 over (Hd, Wd);
 * `row_bytes` bytes per pixel are moved (multiple of 16), pixel strides in bytes - bf16 and fp32 tensors alike. */
int uz_pad_copy(const void* src, long long src_pixel_bytes, int Hs, int Ws, void* dst, long long dst_pixel_bytes, int Hd,
                int Wd, int N, int row_bytes, void* stream);

/* nn.Upsample(scale_factor=2) (nearest, common_layers.py:70) materialised: out [N,2H,2W,C] = x[n, y>>1, x>>1, :].
 * Used only for the weight gradient of UpConvBlock's convolution (3x3 wgrad kernel on the upsampled operand). */
int uz_nearest2x_fwd(const void* x, int x_ld, int N, int H, int W, int C, void* out, int out_ld, void* stream);
int uz_bilinear2x_bwd(const void* dout, int dout_ld, int N, int H, int W, int C, void* dx, int dx_ld, int accumulate,
                      void* stream);

/* ------------------------------------------------------------------------------------------------------------
 * Implicit-GEMM convolutions on tcgen05 (forward-type).  `mode` selects the gather / scatter pattern. */
#define UZ_CONV3X3 0        /* Conv2d k3 s1 p1        common_layers.py:28,31,47,52 ; nested_unet.py:7,9          */
#define UZ_CONV3X3_DGRAD 1  /* its data gradient (weights packed with the dgrad strides)                         */
#define UZ_CONV1X1 2        /* Conv2d k1 s1           attention_unet.py:10,17,24                                  */
#define UZ_CONV3X3_S2 3     /* Conv2d k3 s2 p1        common_layers.py:188 (ResidualConv, stride 2)               */
#define UZ_CONV3X3_S2_DGRAD 4
#define UZ_CONV1X1_S2 5     /* Conv2d k1 s2           common_layers.py:194 (ResidualConv skip)                    */
#define UZ_CONV1X1_S2_DGRAD 6
#define UZ_CONVT2X2 7       /* ConvTranspose2d k2 s2  common_layers.py:104,204                                    */
#define UZ_CONVT2X2_DGRAD 8
#define UZ_UP2X_CONV3X3 9   /* nn.Upsample(scale 2, nearest) + Conv2d k3 p1   common_layers.py:70-73              */
#define UZ_UP2X_CONV3X3_DGRAD 10 /* its data gradient w.r.t. the LOW-res input (weights packed with the dgrad strides) */

/* Dilated 3x3 convolution, Conv2d(k3, padding = dilation = d) of REBNCONV (u2net.py:10): OR into UZ_CONV3X3 /
 * UZ_CONV3X3_DGRAD (uz_conv_igemm) and UZ_CONV3X3 (uz_conv_wgrad, uz_conv_wgrad_workspace_bytes).  d = 1 is the plain mode. */
#define UZ_DILATION(d) ((d) << 8)

#define UZ_FLAG_RELU 1
#define UZ_FLAG_ACCUMULATE 2 /* out += result (used when a tensor has several consumers in backward) */
#define UZ_FLAG_STATS 4      /* also emit per-channel sum / sum-of-squares partials for BatchNorm */
/* Inference form of ResidualConv (common_layers.py:182-199): with UZ_CONV3X3 and >= 2 segments, the LAST segment is the
 * operand of the block's 1x1 skip convolution - only the centre tap of its weight columns is read; _S2: that segment is a
 * 2H x 2W tensor sampled at its even pixels (stride-2 skip, common_layers.py:194).  out = conv3x3(first segments) +
 * conv1x1(last segment) + bias in one accumulator.  Halo-box kernel only (output >= 16 x 8 pixels). */
#define UZ_FLAG_RES1X1 8
#define UZ_FLAG_RES1X1_S2 16

/* out[n, oy, ox, 0:cout] (bf16, pixel stride out_ld) = conv(concat(segments), w_packed) [+ bias] [ReLU].
 * N,H,W: batch and spatial size of the INPUT activation(s).  Output spatial size follows from `mode`.
 * stats: when UZ_FLAG_STATS, fp32 [uz_conv_stats_rows()][2][cout] partial sums (zeroed by this call). */
int uz_conv_igemm(int mode, int N, int H, int W, int nseg, const void* const* seg_ptr, const int* seg_c,
                  const int* seg_ld, const void* w_packed, int cout, const float* bias, void* out, int out_ld,
                  int out_h, int out_w, float* stats, int flags, void* stream);
int uz_conv_stats_rows(void);

/* Nearest-x2 upsample + conv3x3 (UpConvBlock, common_layers.py:69-73) runs in PHASE form on images of at least 16 x 8
 * low-resolution pixels with cout a multiple of 64: each of the four output phases is a 2x2 convolution of the
 * low-resolution input with pre-summed weights (16 tap-pixels per input pixel instead of 36; the 4x upsampled tensor
 * never exists).  uz_up2x_phase_form tells which weight packing UZ_UP2X_CONV3X3 / _DGRAD expect for a layer:
 * 1 -> uz_pack_up2x ([cout][16][cin], transpose = 0; data gradient [cin][16][cout], transpose = 1),
 * 0 -> uz_pack_weight with nine taps (first-generation kernel, small images). */
int uz_up2x_phase_form(int H_low, int W_low, int cout);
int uz_pack_up2x(const float* w_oihw, void* dst_bf16, int cout, int cin, int transpose, void* stream);

/* ------------------------------------------------------------------------------------------------------------
 * Weight gradients on tcgen05 (reduction over pixels).  dY / X are NHWC bf16.
 *   UZ_CONV3X3 / UZ_CONV1X1 / *_S2 / UZ_UP2X_CONV3X3 : dW[co, c_off + ci, r, s] (OIHW fp32, full Cin = cin_total)
 *   UZ_CONVT2X2                                      : dW[ci, co, dy, dx]      (IOHW fp32)
 * N,H,W: size of the conv INPUT x.  `accumulate` != 0 adds into dw instead of overwriting.
 * Replaces: autograd's conv weight-gradient for the same call sites as uz_conv_igemm. */
size_t uz_conv_wgrad_workspace_bytes(int mode, int N, int H, int W, int cin, int cout);
int uz_conv_wgrad(int mode, int N, int H, int W, const void* x, int cin, int x_ld, const void* dy, int cout,
                  int dy_ld, float* dw, int cin_total, int c_off, int accumulate, void* workspace,
                  size_t workspace_bytes, void* stream);

/* ------------------------------------------------------------------------------------------------------------
 * fp32 parity mode (forward only).  BASELINE.json: "fp32 mode within 1e-4 relative on logits"; configs[0] = unet,
 * batch 1, 256x256 fp32 forward (README.md:83-90).  Same operators and reference call sites as above, but on NHWC
 * fp32 activations (`ld` in floats, any channel count) with CUDA-core FFMA arithmetic - the bf16 tensor-core
 * kernels cannot meet 1e-4.  BatchNorm partial statistics have the uz_bn_stats row layout (rows = uz_reduce_rows(),
 * fp64 accumulation inside) and go through uz_bn_finalize unchanged.
 *  uz_f32_pack_weight: dst[t][ci][co] (fp32, [T][cin_total][coutp], zero padded) = src[co*s_co + ci*s_ci + t]
 *                      * (row_scale ? row_scale[co] : 1).  Conv2d OIHW: s_co=Ci*T, s_ci=T; ConvTranspose2d IOHW
 *                      [Ci,Co,2,2]: s_co=4, s_ci=Co*4 (T=4 output phases).  coutp = cout rounded up to 4.
 *  uz_f32_conv       : modes UZ_CONV3X3, UZ_CONV1X1, UZ_CONV3X3_S2, UZ_CONV1X1_S2, UZ_UP2X_CONV3X3, UZ_CONVT2X2 over a
 *                      virtual concat of up to UZ_F32_MAX_SEGS segments; flags UZ_FLAG_RELU / UZ_FLAG_ACCUMULATE;
 *                      output size follows from `mode` (N,H,W describe the INPUT).
 *  uz_f32_affine_act : y = act(z*scale + shift + residual) (+ MaxPool2d(2,2) second output; y may be NULL then);
 *                      scale == shift == NULL -> identity.
 *  uz_f32_bilinear2x : nn.Upsample(scale_factor=2, mode='bilinear', align_corners=True) (nested_unet.py:33).
 *  uz_f32_gate_apply : out = x * sigmoid(psi_raw*bn[2] + bn[3])   (attention_unet.py:38-40).
 *  uz_f32_nchw_to_nhwc / uz_f32_nhwc_to_nchw: layout changes at the model boundary (input image, multi-class logits). */
#define UZ_F32_MAX_SEGS 4
int uz_f32_pack_weight(const float* src, float* dst, int T, int cin, int cin_total, int cout, int coutp,
                       long long s_co, long long s_ci, const float* row_scale, void* stream);
int uz_f32_conv(int mode, int N, int H, int W, int nseg, const void* const* seg_ptr, const int* seg_c,
                const int* seg_ld, const float* w_packed, int cout, int coutp, const float* bias, float* out,
                int out_ld, int flags, void* stream);
int uz_f32_nchw_to_nhwc(const float* x, int N, int C, int H, int W, float* out, int out_ld, void* stream);
int uz_f32_nhwc_to_nchw(const float* in, int in_ld, int N, int C, int H, int W, float* out, void* stream);
int uz_f32_bn_stats(const float* x, int ld, long long npix, int C, float* partial, void* stream);
int uz_f32_affine_act(const float* z, int z_ld, int N, int H, int W, int C, const float* scale, const float* shift,
                      int relu, const float* residual, int res_ld, float* y, int y_ld, float* pooled, int pooled_ld,
                      void* stream);
int uz_f32_bilinear2x(const float* x, int x_ld, int N, int H, int W, int C, float* out, int out_ld, void* stream);
int uz_f32_gate_apply(const float* x, int x_ld, const float* psi_raw, const float* bn_psi, long long npix, int C,
                      float* out, int out_ld, void* stream);

/* ------------------------------------------------------------------------------------------------------------
 * Tail of the training step (training_loop.py:115-121): loss + metric + clip + optimizer as device-side passes with
 * no host synchronisation.  Workspace for both calls: uz_tail_workspace_bytes().
 *  uz_bce_dice  : result[0] = nn.BCEWithLogitsLoss()(logits, target) (mean; scripts/train.py:135),
 *                 result[1] = dice_coefficient(logits, target) (metrics.py:7-24: sigmoid -> > 0.5 ->
 *                 (2I + eps) / (U + eps), 1.0 when U == 0), result[2..4] = intersection, prediction sum, target sum;
 *                 dlogits (optional) = (sigmoid(logits) - target) * grad_scale, i.e. the loss gradient for
 *                 grad_scale = 1/n.  One pass over logits and target (fp32, n elements, 16-byte aligned).
 *  uz_adamw_flat: torch.nn.utils.clip_grad_norm_(params, max_norm) + torch.optim.AdamW.step() over FLAT fp32 buffers
 *                 (n elements each, 16-byte aligned; padding elements must be zero in params and grads).
 *                 hyper[6] (device) = {lr, beta1, beta2, eps, weight_decay, max_norm (<= 0: no clipping)};
 *                 state[5] (device) = {step, total_norm, clip_coef, bias_correction1, sqrt(bias_correction2)} -
 *                 state[0] must persist between calls (zero before the first); grads are scaled by clip_coef in
 *                 place like clip_grad_norm_ does. */
size_t uz_tail_workspace_bytes(void);
int uz_bce_dice(const float* logits, const float* target, long long n, float grad_scale, float dice_eps,
                float* dlogits, float* result, void* workspace, size_t workspace_bytes, void* stream);
int uz_adamw_flat(float* params, float* grads, float* exp_avg, float* exp_avg_sq, long long n, const float* hyper,
                  float* state, void* workspace, size_t workspace_bytes, void* stream);

/* transforms.Resize((H, W)) on the decoded uint8 image / mask (unet_zoo/data/datasets.py:41,47 -> PIL Image.resize,
 * BILINEAR with Pillow's support scaling), bit-exact with Pillow's ImagingResample (8-bit fixed-point, two passes).
 * src [N][Hin][Win][C] (C <= 4) -> dst [N][Hout][Wout][C]; tmp [N][Hin][Wout][C] when both axes change.  The window
 * and coefficient tables come from unet_zoo_b200.data.resample_coeffs (device int32 arrays). */
int uz_resize_bilinear_u8(const uint8_t* src, int N, int Hin, int Win, int C, uint8_t* tmp, uint8_t* dst, int Hout,
                          int Wout, const int* bounds_h, const int* kk_h, int ksize_h, const int* bounds_v,
                          const int* kk_v, int ksize_v, void* stream);

/* ------------------------------------------------------------------------------------------------------------
 * Input pipeline tail on the device (datasets.py:40-44,59; training_loop.py:109-110): the decoded bytes travel
 * (3 + 1 bytes per pixel instead of 12 + 4) and the device normalises.
 *  uz_stem_im2col_u8: img NHWC uint8 [N,H,W,cin] (as PIL decodes it) -> ToTensor (/255) -> Normalize((v-mean)/std)
 *                     -> bf16 -> the 3x3 im2col operand that uz_stem_im2col writes for the normalised fp32 tensor.
 *                     mean / std are HOST arrays of cin floats (cin <= 7).
 *  uz_mask_u8_to_f32: out = (mask/255 > 0.5) ? 1 : 0  (fp32; the mask transform of datasets.py:46-49,59). */
int uz_stem_im2col_u8(const uint8_t* img, int N, int cin, int H, int W, const float* mean_host, const float* std_host,
                      void* out, int kp, void* stream);
int uz_mask_u8_to_f32(const uint8_t* mask, long long n, float* out, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* UNET_ZOO_B200_H_ */
