/*
 * cxxnet_b200 kernel library - C ABI of libcxxnet_b200.so (sm_100a only).
 *
 * Every entry point takes raw device pointers and a cudaStream_t, launches asynchronously on
 * that stream, allocates nothing (unless its comment says otherwise) and returns 0 on success,
 * a cudaError_t (> 0) or a negative library error.  bf16 tensors are passed as void*.
 * Activations are NHWC; 2-D matrices are row-major [rows][ld] with `cols` valid entries.
 * Citations name the reference code whose BEHAVIOUR the kernel reproduces
 * (paths relative to the cxxnet reference tree).
 */
#ifndef CXXNET_B200_H_
#define CXXNET_B200_H_

#include <cuda_runtime_api.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ---- GEMM flags --------------------------------------------------------------------------- */
#define CXN_GEMM_RELU     1  /* max(x,0) after bias                                  */
#define CXN_GEMM_OUT_BF16 2  /* D is bf16 (else fp32)                                */
#define CXN_GEMM_ATOMIC   4  /* fp32 D += partial with red.global.add (split-K)      */
#define CXN_GEMM_BETA1    8  /* fp32 D = D + result (read-modify-write, no atomics)  */

/* K5/K6 + the GEMMs inside K1-K3: D[z] = A[z] * B[z]^T (+bias) on tcgen05 tensor cores, bf16 in,
 * fp32 accumulate in TMEM, operands by TMA.  a_mn / b_mn = 0: operand stored [MN][K] (K
 * contiguous); 1: stored [K][MN].  Leading dimensions and strides in elements (multiples of 8).
 * Reference: dot() calls in src/layer/fullc_layer-inl.hpp:108-129 and
 * src/layer/convolution_layer-inl.hpp:93-96,135-142 (cuBLAS SGEMM behind mshadow). */
int cxn_gemm_bf16(const void* A, long long lda, long long strideA, int a_mn, const void* B, long long ldb,
                  long long strideB, int b_mn, void* D, long long ldd, long long strideD, const float* bias,
                  long long strideBias, int M, int N, int K, int batch, int splits, int flags, int bn,
                  cudaStream_t stream);
/* experiment knob: skinny K-major-A problems (the fully connected forward / backward-data GEMMs) run as clusters of
 * cl = 2 | 4 column tiles that share the activation tile by TMA multicast (0 = off, -1 = re-read CXN_GEMM_CLUSTER).
 * No reference counterpart. */
int cxn_gemm_cluster(int cl);
/* CTA-pair route of cxn_gemm_bf16 (tcgen05 cta_group::2: two SMs compute one 256 x 256 tile, each staging its own
 * rows of A and half of the B tile): 1 = automatic for problems with at least one tile per pair of SMs (default),
 * 0 = never, 2 = whenever the shape allows, -1 = re-read CXN_GEMM_PAIR.  No reference counterpart. */
int cxn_gemm_pair(int mode);
/* The internally split GEMMs (skinny fully connected shapes) meet in a zeroed fp32 scratch buffer; a device has
 * four independent ones ("lanes").  Selects the lane used by the calling thread's following cxn_gemm_bf16 calls
 * (default 0) and returns the previous one, -1 for an invalid lane.  GEMMs that can run concurrently (different
 * streams, or parallel branches of a captured graph) must use different lanes.  No reference counterpart. */
int cxn_gemm_workspace_lane(int lane);
/* K6 + K15 in one kernel: the weight gradient of a fully connected layer, g (M x N) = A . B^T with both operands
 * stored [K][MN] (dy as [batch][nout], x as [batch][nin]), whose epilogue applies the updater (kind 0 sgd, 1 nag,
 * 2 adam - the arithmetic of cxn_update) to the fp32 master w and the state m (m2 for adam) at [row * ldw + column] and
 * rewrites the bf16 copy at [row * shadow_ld + column] (shadow may be NULL).  The gradient is never stored: 18 bytes
 * per parameter cross HBM instead of 26.  hyper = {lr, wd, momentum, clip, adam_lr_t, beta1, beta2} in device memory
 * (the layer's entry of the per-step segment table), read when the kernel runs.  Returns -7 when the arguments do not
 * fit the fused path (the caller then uses cxn_gemm_bf16 + cxn_update).
 * Reference: src/layer/fullc_layer-inl.hpp:121-123 followed by src/updater/sgd_updater-inl.hpp:73-84. */
int cxn_gemm_wgrad_update(const void* A, long long lda, const void* B, long long ldb, int M, int N, int K, float* w,
                          float* m, float* m2, void* shadow, long long ldw, long long shadow_ld, const float* hyper,
                          int kind, cudaStream_t stream);

/* K1 (and K2 for stride-1 layers): implicit-GEMM convolution forward, bias (+ReLU) fused.
 * Reference: src/layer/convolution_layer-inl.hpp:79-105. */
int cxn_conv_fprop(const void* src, int N, int H, int W, int Cs, int Cg, int G, int gran, const void* w,
                   long long ldw, int Ng, int KH, int KW, int stride, int pad_y, int pad_x, int OH, int OW, void* y,
                   long long ldy, const float* bias, int relu, cudaStream_t stream);
/* K2 for strided layers: backward-data as stride * stride implicit-GEMM problems, one per stride phase - the positions
 * (ih + pad) % s == ph only meet the taps kh = ph (mod s), and over that sub-grid the layer is a stride-1 correlation
 * of dy with the flipped sub-filter, written through a strided view of dx; no im2col buffer, no multiplication by an
 * inserted zero.  `w` is the forward filter [G][ocg][KH][KW][icg] (row pitch ldw), `mask` an optional tensor laid out
 * like dx (results become zero where it is <= 0), `scratch` bf16 room of cxn_conv_dgrad_strided_scratch() elements.
 * -24: a phase would need negative padding (cxn_conv_dgrad_strided_ok answers 0) - use the explicit path.
 * Reference: src/layer/convolution_layer-inl.hpp:140-153. */
/* Limit the grid of the calling thread's following weight-gradient launches to `ctas` CTAs (0 = the whole GPU);
 * returns the previous limit.  For running a weight gradient on one stream beside a backward-data problem that fills
 * only part of the GPU on another (both kernels are persistent, one CTA per SM).  No reference counterpart. */
int cxn_conv_wgrad_grid_cap(int ctas);
/* Number of forward launches of this process that took the FLAT route so far: a stride-1 problem without padding and with
 * at most 128 output columns per group is computed over the INPUT's pixel grid, where a filter tap is a constant row
 * offset and a tile's patch rows are one tiled-mode TMA box; rows outside the output are discarded (CXN_CONV_FLAT /
 * cxn_conv_tune(12, v)).  For tests.  No reference counterpart. */
long long cxn_conv_flat_launches(void);
/* The next cxn_conv_fprop_view / cxn_conv_wgrad_view call of the calling thread takes group g's channels from channel
 * g * stride on instead of g * Cg: groups as overlapping channel windows, each padded up to a multiple of 64 channels
 * that meet structural zeros of the filter (channel-padded groups; one launch for all groups).  stride % 8 == 0.
 * No reference counterpart. */
int cxn_conv_view_group_stride(int stride);
long long cxn_conv_dgrad_strided_scratch(int G, int ocg, int icg, int KH, int KW, int stride);
int cxn_conv_dgrad_strided_ok(int KH, int KW, int stride, int pad_y, int pad_x);
int cxn_conv_dgrad_strided(const void* dy, int N, int OH, int OW, long long lddy, int ocg, int G, const void* w,
                           long long ldw, int icg, int KH, int KW, int stride, int pad_y, int pad_x, int H, int W, void* dx,
                           long long lddx, const void* mask, void* scratch, cudaStream_t stream);
/* K3: implicit-GEMM weight gradient, split-K over pixels, fp32 atomics into dw.
 * Reference: src/layer/convolution_layer-inl.hpp:121-138. */
int cxn_conv_wgrad(const void* src, int N, int H, int W, int Cs, int Cg, int G, int gran, const void* dy,
                   long long lddy, int Ng, int KH, int KW, int stride, int pad_y, int pad_x, int OH, int OW, float* dw,
                   long long lddw, int splits, cudaStream_t stream);
/* filter flip + channel transpose used to run backward-data as a forward convolution. */
/* cxn_conv_wgrad that leaves columns n with (n % mask_period) >= mask_valid untouched (structural zeros of a
 * padded filter layout; both multiples of 4, 0 = no mask).  In every conv entry point `stride` may carry an
 * anisotropic stride as stride_y | (stride_x << 16). */
int cxn_conv_wgrad_masked(const void* src, int N, int H, int W, int Cs, int Cg, int G, int gran, const void* dy,
                          long long lddy, int Ng, int KH, int KW, int stride, int pad_y, int pad_x, int OH, int OW,
                          float* dw, long long lddw, int splits, int mask_period, int mask_valid, cudaStream_t stream);
/* Convolution over a strided view of the input: position (n,h,w) starts at src + n*img_stride + h*row_stride + w*Cs
 * and exposes `cdim` >= G*Cg contiguous elements (positions may overlap in memory).  Only the im2col-TMA route can
 * read such a view (Cg % 64 == 0 ..., else -17).  The first layer uses it: 16 pixels x 4 channels of a pitch-4 image
 * are one 64-channel position every `stride` pixels, which turns an 11x11x3 filter into an 11x1x64 one. */
int cxn_conv_fprop_view(const void* src, int N, int H, int W, int Cs, int Cg, int G, const void* w, long long ldw, int Ng,
                        int KH, int KW, int stride, int pad_y, int pad_x, int OH, int OW, void* y, long long ldy,
                        const float* bias, int relu, long long cdim, long long row_stride, long long img_stride,
                        cudaStream_t stream);
int cxn_conv_wgrad_view(const void* src, int N, int H, int W, int Cs, int Cg, int G, const void* dy, long long lddy, int Ng,
                        int KH, int KW, int stride, int pad_y, int pad_x, int OH, int OW, float* dw, long long lddw,
                        int splits, int mask_period, int mask_valid, long long cdim, long long row_stride,
                        long long img_stride, cudaStream_t stream);
/* cxn_conv_fprop (no bias / activation) whose outputs are zeroed where mask[row][col] <= 0; mask is bf16 laid out
 * like y.  Backward-data of a layer whose input went through a ReLU, in one kernel. */
int cxn_conv_fprop_masked(const void* src, int N, int H, int W, int Cs, int Cg, int G, int gran, const void* w,
                          long long ldw, int Ng, int KH, int KW, int stride, int pad_y, int pad_x, int OH, int OW,
                          void* y, long long ldy, const void* mask, cudaStream_t stream);
int cxn_conv_flip_weights(const void* w, void* wt, int G, int ocg, int icg, int KH, int KW, int ld_in, int ld_out,
                          cudaStream_t stream);
/* The same from a filter whose taps are cpitch >= icg elements apart (channel-padded internal layout; the padding
 * channels are not part of the backward-data problem).  No reference counterpart. */
int cxn_conv_flip_weights_pitch(const void* w, void* wt, int G, int ocg, int icg, int KH, int KW, int ld_in, int ld_out,
                                int cpitch, cudaStream_t st);
/* experiment / tuning knobs of the implicit-GEMM convolution (0, or -1 for switches, = automatic):
 * key 0 row halves per forward tile (1|2), 1 ring stages, 2 unused,
 * 3 unused, 4 forward tile-width class (128|256), 5 timing-experiment bits,
 * 6 im2col-mode TMA for the patch operand (0|1), 7 = 2: forward layers on the im2col-TMA route run as clusters of
 * two CTAs that multicast the filter tile to each other (also CXN_CONV_CLUSTER=2), 8 = 1: narrower column tiles when
 * a forward grid would leave SMs idle (also CXN_CONV_FILL=1), 9 = 1|2: halo-tile kernel for stride-1 problems with at
 * most 64 output columns per group (also CXN_CONV_HALO).  No reference counterpart. */
int cxn_conv_tune(int key, int value);
/* host-only planning query of the halo-tile kernel (key 9): 1 and out[0..7] = {padded width, output rows per half, tiles
 * per image, patch rows, patch stage bytes, patch stages, filter stages, dynamic shared memory}, or 0 if the kernel
 * would not take the problem.  No reference counterpart. */
int cxn_conv_halo_plan(int N, int H, int W, int Cs, int Cg, int G, int Ng, int KH, int KW, int stride, int pad_y, int pad_x,
                       int OH, int OW, int halves, int* out);
/* hardware probe, not on any training path (csrc/kernels/probe_sm100.cu): one tcgen05.mma whose 128B-swizzled
 * operand starts `shift_rows` rows (0..32) into a shared-memory patch; out = device float [128][16];
 * base_offset_mode 0 | 1 = descriptor base-offset field left 0 | set from the start address.
 * cxn_probe_expected(m, n, shift_rows) is the value out[m][n] should hold.  No reference counterpart. */
int cxn_probe_shifted_desc(float* out, int shift_rows, int base_offset_mode, cudaStream_t stream);
float cxn_probe_expected(int m, int n, int shift_rows);
/* per-role cycle counters of the forward implicit-GEMM kernel: buf = device int64 [148][16], NULL = off */
int cxn_conv_set_prof(void* buf);
/* explicit NHWC patch gather / scatter (fallback path).  Reference: unpack_patch2col /
 * pack_col2patch use in src/layer/convolution_layer-inl.hpp:86-92,143-151. */
int cxn_im2col_nhwc(const void* x, void* col, int N, int H, int W, int C, int G, int KH, int KW, int stride, int pad_y,
                    int pad_x, int OH, int OW, int ldk, long long gstride, cudaStream_t stream);
int cxn_col2im_nhwc(const void* col, void* dx, int N, int H, int W, int C, int G, int KH, int KW, int stride,
                    int pad_y, int pad_x, int OH, int OW, int ldk, long long gstride, cudaStream_t stream);

/* ---- layout (API boundary: public layout is NCHW fp32, src/layer/layer.h:33-54) ------------- */
int cxn_nchw_f32_to_nhwc_bf16(const float* src, void* dst, int N, int C, int HW, int ldc, cudaStream_t stream);
/* C <= 4 input image -> NHWC bf16 with pixel pitch 4 and an even row pitch Wp >= W (pad pixels zero) */
int cxn_nchw_f32_to_nhwc4_pitched(const float* src, void* dst, int N, int C, int H, int W, int Wp, cudaStream_t stream);
/* NCHW input with C <= 4 -> the space-to-depth layout [N][ceil(H/4)][ceil(W/4)][64] (channel (dy*4 + dx)*4 + c of
 * block (Y, X) = pixel (4Y + dy, 4X + dx), channel c; zero past the image and for c >= C): a stride-4 first layer reads
 * it as a stride-1 convolution over 64 channels whose patch rows are aligned 128-byte lines.  kind: 0 fp32 source,
 * 1 bf16, 2 8-bit pixels normalised as (pixel - mean[c]) * scale.  No reference counterpart (the reference unpacks
 * patches of the NCHW tensor, src/layer/convolution_layer-inl.hpp:88-92). */
int cxn_nchw_to_s2d4(const void* src, int kind, void* dst, int N, int C, int H, int W, float mean0, float mean1, float mean2,
                     float mean3, float scale, cudaStream_t stream);
/* g[r * ld + c] = 0 where mask[c] == 0: the structural zeros of an internal filter layout collect no gradient. */
int cxn_mask_cols_f32(float* g, long long rows, int cols, long long ld, const float* mask, cudaStream_t stream);
int cxn_nchw_bf16_to_nhwc4_pitched(const void* src, void* dst, int N, int C, int H, int W, int Wp, cudaStream_t stream);
/* the same from 8-bit pixels, with the input normalisation (pixel - mean[c]) * scale applied in fp32 before the bf16
 * rounding (`input_dtype = u8` batches; no reference counterpart) */
int cxn_nchw_u8_to_nhwc4_pitched(const void* src, void* dst, int N, int C, int H, int W, int Wp, float mean0, float mean1,
                                 float mean2, float mean3, float scale, cudaStream_t stream);
int cxn_nhwc_bf16_to_nchw_f32(const void* src, float* dst, int N, int C, int HW, int ldc, cudaStream_t stream);
/* flatten in (c,h,w) order and its inverse.  Reference: src/layer/flatten_layer-inl.hpp:10-39. */
int cxn_bf16_hwc_chw(const void* src, void* dst, int N, int C, int HW, long long src_img, long long dst_img,
                     int ld_hwc, int ld_chw, int to_chw, cudaStream_t stream);
/* kind: 0 bf16->bf16, 1 f32->bf16, 2 bf16->f32, 3 f32->f32.  Concat / split copies (K14). */
int cxn_copy2d(const void* src, long long lds, void* dst, long long ldd, long long rows, int cols, int kind,
               int accumulate, cudaStream_t stream);
int cxn_fill_f32(float* p, long long n, float v, cudaStream_t stream);

/* ---- K7 activations: op 0 relu, 1 sigmoid, 2 tanh, 3 xelu(b).  Backward takes the ACTIVATED
 * value, as the reference keeps it in the input node.  src/layer/op.h:27-73,
 * src/layer/activation_layer-inl.hpp:28-37, src/layer/xelu_layer-inl.hpp:35-44. */
int cxn_act_fwd(const void* x, void* y, long long rows, int cols, long long ld, int op, float b, cudaStream_t stream);
int cxn_act_bwd(const void* y, const void* dy, void* dx, long long rows, int cols, long long ld, int op, float b,
                cudaStream_t stream);
/* K12 dropout: y = x * (U<pkeep)/pkeep with a Philox mask regenerated from (seed, *counter).
 * src/layer/dropout_layer-inl.hpp:44-58. */
int cxn_dropout(const void* x, void* y, long long rows, int cols, long long ld, float pkeep, unsigned long long seed,
                const unsigned long long* counter, cudaStream_t stream);
int cxn_counter_add(unsigned long long* counter, unsigned long long inc, cudaStream_t stream);
/* K4 bias gradient: out[c] (+)= sum_r x[r][c].  src/layer/fullc_layer-inl.hpp:123-125,
 * src/layer/convolution_layer-inl.hpp:117-119. */
int cxn_colsum_bf16(const void* x, long long rows, int cols, long long ld, float* out, int accumulate,
                    cudaStream_t stream);
/* the vectorised kernel behind cxn_colsum_bf16 (cols % 8 == 0, 16-byte aligned rows): out[c] += sum_r x[r][c] */
int cxn_colsum_bf16_vec(const void* x, long long rows, int cols, long long ld, float* out, cudaStream_t stream);

/* ---- K8/K9 pooling: mode 0 max, 1 sum, 2 avg; ceil-style output, clipped windows, avg / k*k,
 * max-unpool to every tie; relu=1 is relu_max_pooling.  src/layer/pooling_layer-inl.hpp:45-106. */
/* ---- bias / prelu / batch_norm (csrc/kernels/misc_layers.cu); matrices are bf16 [rows][ld], channels = columns.
 * reference: src/layer/bias_layer-inl.hpp, prelu_layer-inl.hpp, batch_norm_layer-inl.hpp */
int cxn_bias_add(const void* x, void* y, long long rows, int cols, long long ld, const float* bias, cudaStream_t stream);
/* scratch: 2*C floats; mean / istd: C floats each, kept for backward; xhat may be NULL (evaluation) */
int cxn_batchnorm_fwd(const void* x, void* y, void* xhat, long long rows, int C, long long ldx, long long ldy,
                      long long ldxhat, const float* slope, const float* bias, float eps, float* scratch, float* mean,
                      float* istd, cudaStream_t stream);
/* gslope += sum dy*xhat, gbias += sum dy, dx = slope*istd*(dy - mean(dy) - xhat*mean(dy*xhat)); scratch: 5*C floats */
int cxn_batchnorm_bwd(const void* dy, const void* xhat, void* dx, long long rows, int C, long long lddy, long long ldxhat,
                      long long lddx, const float* slope, const float* istd, float* gslope, float* gbias, float* scratch,
                      int prop_grad, cudaStream_t stream);
/* y = x > 0 ? x : x * clamp(slope[c] * (1 + u*2*noise - noise), 0, 1), u from Philox(seed, *counter) */
int cxn_prelu_fwd(const void* x, void* y, long long rows, int cols, long long ldx, long long ldy, const float* slope,
                  float noise, unsigned long long seed, const unsigned long long* counter, cudaStream_t stream);
int cxn_prelu_bwd(const void* x, const void* dy, void* dx, long long rows, int cols, long long ldx, long long ldg,
                  const float* slope, float* gslope, float noise, unsigned long long seed,
                  const unsigned long long* counter, cudaStream_t stream);
/* insanity (randomised leaky ReLU): y = x > 0 ? x : x / s, s ~ U[lb,ub) per element when `random`, else (lb+ub)/2;
 * backward regenerates s.  Reference: src/layer/insanity_layer-inl.hpp:13-102 */
int cxn_insanity_fwd(const void* x, void* y, long long rows, int cols, long long ldx, long long ldy, float lb, float ub,
                     int random, unsigned long long seed, const unsigned long long* counter, cudaStream_t stream);
int cxn_insanity_bwd(const void* x, const void* dy, void* dx, long long rows, int cols, long long ldx, long long ldg,
                     float lb, float ub, int random, unsigned long long seed, const unsigned long long* counter,
                     cudaStream_t stream);
/* insanity pooling: out[n][y][x][c] = in at (y,x) with probability keep, else at the upper / lower / left / right
 * neighbour ((1-keep)/4 each, clamped); NHWC bf16 with pixel pitch ld.  Reference: insanity_pooling_layer-inl.hpp:63-94 */
int cxn_insanity_displace(const void* in, void* out, int N, int H, int W, int C, long long ld, float keep,
                          unsigned long long seed, const unsigned long long* counter, cudaStream_t stream);
int cxn_pool_fwd(const void* x, void* y, int N, int H, int W, int C, int ky, int kx, int stride, int OH, int OW,
                 int mode, int relu, cudaStream_t stream);
int cxn_pool_bwd(const void* x, const void* y, const void* dy, void* dx, int N, int H, int W, int C, int ky, int kx,
                 int stride, int OH, int OW, int mode, int relu, cudaStream_t stream);
/* K9 + K4: max-pool backward that also accumulates the column sums of dx into bias_grad[C] (fp32; zeroed, or the
 * value to add to) - the bias gradient of the convolution whose ReLU-masked output is the pooled node
 * (src/layer/convolution_layer-inl.hpp:127-129).  -8 = shape outside the fused kernel (use cxn_pool_bwd + a column sum). */
int cxn_pool_bwd_bias(const void* x, const void* y, const void* dy, void* dx, float* bias_grad, int N, int H, int W, int C,
                      int ky, int kx, int stride, int OH, int OW, int mode, int relu, cudaStream_t stream);
/* ---- K10/K11 local response normalisation across channels.  src/layer/lrn_layer-inl.hpp:53-75. */
int cxn_lrn_fwd(const void* x, void* y, long long pixels, int C, int nsize, float alpha, float beta, float knorm,
                cudaStream_t stream);
int cxn_lrn_bwd(const void* x, const void* dy, void* dx, long long pixels, int C, int nsize, float alpha, float beta,
                float knorm, cudaStream_t stream);

/* ---- K13 loss layers: in-place softmax / sigmoid on the fp32 output node and the gradient
 * (p - target) * scale as bf16.  kind 0: class-index label, 1: dense label.
 * src/layer/loss/softmax_layer-inl.hpp:21-32, loss_layer_base-inl.hpp:52-63. */
/* error metric on the device: accum (double) += #rows whose first-maximum score index != (int)label[row*label_ld]
 * (cols == 1: score > 0 is class 1).  Reference: src/utils/metric.h:92-121 (host loop after a D2H copy). */
int cxn_metric_error(const float* p, long long ld, const float* label, long long label_ld, int rows, int cols,
                     double* accum, cudaStream_t stream);
int cxn_softmax_rows(float* x, int rows, int cols, long long ld, cudaStream_t stream);
int cxn_sigmoid_rows(float* x, long long rows, int cols, long long ld, cudaStream_t stream);
int cxn_loss_grad(const float* p, long long ld, const float* label, long long label_ld, void* g, long long ldg,
                  long long rows, int cols, float scale, int kind, cudaStream_t stream);

/* ---- K15 updaters on the flat parameter arena: kind 0 sgd, 1 nag, 2 adam; zeroes the gradient
 * and refreshes the bf16 GEMM shadow in the same pass.  `segs` is an array of 64-byte segment
 * records (cxn_segment_bytes()), `chunk_seg` maps each cxn_update_chunk()-element chunk to its
 * segment.  src/updater/{sgd,nag,adam}_updater-inl.hpp. */
int cxn_segment_bytes(void);
int cxn_update_chunk(void);
int cxn_update(float* w, float* g, float* m, float* m2, void* shadow, const void* segs, const int* chunk_seg,
               long long chunk_begin, long long chunk_end, int kind, cudaStream_t stream);
int cxn_refresh_shadow(const float* w, void* shadow, const void* segs, const int* chunk_seg, long long nchunks,
                       cudaStream_t stream);

/* ---- K17/K19 multi-GPU: fused gradient reduce-scatter + update + weight all-gather over
 * NVLink peer memory, one launch per layer bucket; weight broadcast.  `peers` is a HOST array
 * of every rank's symmetric-block base address as mapped in the calling process.
 * src/updater/async_updater-inl.hpp:94-127, src/nnet/nnet_impl-inl.hpp:70-81. */
int cxn_fused_rs_update_ag(const unsigned long long* peers, int nranks, int rank, long long off_w, long long off_g,
                           long long off_m, long long off_m2, long long off_shadow, long long off_flags,
                           const void* segs, const int* chunk_seg, long long chunk_begin, long long chunk_end, int kind,
                           unsigned seq, int slot, unsigned* counter, int grid, cudaStream_t stream);
/* Same exchange for hosts that must not allocate device memory (no counterpart in the reference).
 * Meant for a process that hands the whole GPU to its own pool at start-up, so that a later
 * cudaMalloc by a library fails, and for latency-critical loops that cannot take the device-wide
 * synchronisation of cudaMalloc / cudaFree.  The per-bucket CTA counter that the entry point above
 * takes from the caller lives here in a 1 KiB array inside the library image (256 slots, resident
 * on every GPU that loads the library until exit): `slot` >= cxn_fused_rs_update_ag_noalloc_max_slots()
 * is refused with an error and nothing is launched.  The slot's word is cleared on `stream` before
 * each launch.  All calls share the array, so two calls for the same slot that overlap on the GPU
 * race.  The first call on each device resolves the array's device address; make that call
 * outside any stream capture.  Calls no cudaMalloc / cudaFree or allocating library routine. */
int cxn_fused_rs_update_ag_noalloc(const unsigned long long* peers, int nranks, int rank, long long off_w,
                                   long long off_g, long long off_m, long long off_m2, long long off_shadow,
                                   long long off_flags, const void* segs, const int* chunk_seg, long long chunk_begin,
                                   long long chunk_end, int kind, unsigned seq, int slot, int grid,
                                   cudaStream_t stream);
int cxn_fused_rs_update_ag_noalloc_max_slots(void);
int cxn_symm_broadcast(const unsigned long long* peers, int nranks, int rank, long long off_flags, long long off,
                       long long bytes, int root, unsigned seq, int slot, unsigned* counter, int grid,
                       cudaStream_t stream);
/* K18: all-gather of one or two local pieces into every rank's block (fullc_gather rows, weight
 * shards), with entry / exit flag barriers.  src/updater/async_updater-inl.hpp:67-93,115-126. */
int cxn_symm_allgather(const unsigned long long* peers, int nranks, int rank, long long off_flags, const void* src0,
                       long long dst_off0, long long slot_bytes0, long long valid_bytes0, const void* src1,
                       long long dst_off1, long long slot_bytes1, long long valid_bytes1, int npieces, unsigned seq,
                       int slot, unsigned* counter, int grid, cudaStream_t stream);
/* optional device word added to the `seq` argument of every exchange kernel at kernel start (so that a
 * captured CUDA graph advances the flag barriers on replay); NULL = off.  No reference counterpart. */
int cxn_xchg_set_epoch_ptr(const void* dev_word);
int cxn_xchg_flag_bytes(void);
/* NVSwitch multicast address of the symmetric block (or NULL): with it the fused exchange reduces the gradients with
 * multimem.ld_reduce (the sum is formed inside the switch) and publishes the weights with one multimem.st per 16 bytes
 * instead of one load / store per rank.  One GPU per process.  No reference counterpart. */
int cxn_xchg_set_multicast(void* mc_base);
int cxn_xchg_max_ranks(void);
int cxn_xchg_max_slots(void);
/* Collectives of the column-sharded (model-parallel) fully connected layers (csrc/kernels/tp_collectives.cu): the
 * reference moves whole weight gradients / weights of these layers through its parameter server
 * (src/updater/async_updater-inl.hpp:94-127); here the weights stay with their owner and activations move.
 *   cxn_tp_push         one 2-D sub-block per destination rank (desc: nranks x {src_off, dst_off, nrows, valid_rows,
 *                       row_bytes} as long long; rows >= valid_rows are written as zeros) from local memory into every
 *                       rank's symmetric block by peer stores, then ONE flag barrier (signal, wait for all).
 *   cxn_tp_reduce_pull  flag barrier ("my partial is complete"), then dst = sum over ranks of the same sub-block of
 *                       every rank's bf16 partial matrix (peer loads, fp32 sum, bf16 result).
 * `peers`: HOST array of block base addresses as mapped here; off_flags: cxn_tp_flag_bytes() zeroed bytes inside every
 * block; counters: cxn_tp_counter_bytes() zeroed LOCAL device bytes (the per-slot sequence numbers live there, so a
 * launch captured in a CUDA graph replays correctly); all offsets / pitches / row_bytes multiples of 16. */
int cxn_tp_set_multicast(void* mc_base);   /* NVSwitch multicast address of the block (or NULL): multimem.st / ld_reduce */
int cxn_tp_flag_bytes(void);
int cxn_tp_counter_bytes(void);
int cxn_tp_max_slots(void);
int cxn_tp_push(const unsigned long long* peers, int nranks, int rank, long long off_flags, int slot, unsigned* counters,
                const void* src, long long src_ld, long long dst_ld, const long long* desc, int grid, cudaStream_t stream);
int cxn_tp_reduce_pull(const unsigned long long* peers, int nranks, int rank, long long off_flags, int slot,
                       unsigned* counters, long long part_off, long long part_ld, long long pull_off, int rows,
                       int row_bytes, void* dst, long long dst_ld, int grid, cudaStream_t stream);
/* CUDA IPC helpers for the symmetric block (these DO allocate / map device memory). */
int cxn_ipc_alloc(unsigned long long bytes, unsigned long long* ptr, unsigned char* handle64);
int cxn_ipc_open(const unsigned char* handle64, unsigned long long* ptr);
int cxn_ipc_close(unsigned long long ptr);
int cxn_ipc_free(unsigned long long ptr);
/* 1 when [p, p + bytes) is plain device memory of the current GPU allocated by this process (not an IPC / peer /
 * VMM mapping of another GPU's memory, not host memory), else 0.  cxn_gemm_bf16 and cxn_conv_fprop* consult it
 * before taking a cluster-multicast route: a multicast tensor map must cover this GPU's memory alone; every
 * other operand runs on the non-cluster kernels.  No reference counterpart. */
int cxn_memory_is_local(const void* p, unsigned long long bytes);
/* Report a range that maps, or multicasts to, another GPU's memory (VMM imports, NVSwitch multicast addresses obtained
 * outside cxn_ipc_open): cxn_memory_is_local answers 0 for everything that overlaps it.  No reference counterpart. */
int cxn_note_peer_range(const void* p, unsigned long long bytes);

/* step trace: stores the GPU's %globaltimer (ns) into buf[slot]; one thread, capturable.  No reference counterpart
 * (the reference has no profiling facility). */
int cxn_timestamp(unsigned long long* buf, int slot, cudaStream_t stream);

/* number of kernel launches issued by this library in this process */
unsigned long long cxn_launch_count(void);
unsigned long long cxn_launch_count_add(unsigned long long n);   /* account for kernels replayed from a CUDA graph */

#ifdef __cplusplus
}
#endif
#endif /* CXXNET_B200_H_ */
