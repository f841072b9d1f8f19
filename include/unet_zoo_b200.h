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
 * Weight packing: fp32 parameter -> bf16 operand  dst[a][t][b] = src[a*sa + b*sb + t*st]   (a<A, t<T, b<B)
 *   Conv2d fwd  (OIHW [Co,Ci,kh,kw])      : A=Co T=kh*kw B=Ci  sa=Ci*T sb=T  st=1
 *   Conv2d dgrad                           : A=Ci T=kh*kw B=Co  sa=T    sb=Ci*T st=1
 *   ConvTranspose2d fwd (IOHW [Ci,Co,2,2]) : A=Co T=4     B=Ci  sa=4    sb=Co*4 st=1
 *   ConvTranspose2d dgrad                  : A=Ci T=4     B=Co  sa=Co*4 sb=4   st=1
 * Replaces: the implicit weight reads of nn.Conv2d / nn.ConvTranspose2d (common_layers.py:28,31,104). */
int uz_pack_weight(const float* src, void* dst_bf16, int A, int T, int B, long long sa, long long sb, long long st,
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

#define UZ_FLAG_RELU 1
#define UZ_FLAG_ACCUMULATE 2 /* out += result (used when a tensor has several consumers in backward) */
#define UZ_FLAG_STATS 4      /* also emit per-channel sum / sum-of-squares partials for BatchNorm */

/* out[n, oy, ox, 0:cout] (bf16, pixel stride out_ld) = conv(concat(segments), w_packed) [+ bias] [ReLU].
 * N,H,W: batch and spatial size of the INPUT activation(s).  Output spatial size follows from `mode`.
 * stats: when UZ_FLAG_STATS, fp32 [uz_conv_stats_rows()][2][cout] partial sums (zeroed by this call). */
int uz_conv_igemm(int mode, int N, int H, int W, int nseg, const void* const* seg_ptr, const int* seg_c,
                  const int* seg_ld, const void* w_packed, int cout, const float* bias, void* out, int out_ld,
                  int out_h, int out_w, float* stats, int flags, void* stream);
int uz_conv_stats_rows(void);

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

#ifdef __cplusplus
}
#endif
#endif /* UNET_ZOO_B200_H_ */
