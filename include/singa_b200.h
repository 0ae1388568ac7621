/* singa_b200 native library — C ABI.
 *
 * Every entry point launches sm_100a kernels on the CUDA stream passed as `stream` (a cudaStream_t
 * cast to void*; NULL = legacy default stream) and returns 0 on success or a non-zero code; the text
 * of the last failure on the calling thread is returned by sg_last_error().  Pointers are device
 * pointers unless stated otherwise.  "bf16" buffers hold __nv_bfloat16.  Leading dimensions (ld*) are in
 * elements.  Unless an entry point says otherwise it allocates no memory.
 *
 * Reference citations (path:line under the reference tree) name the code whose behaviour the entry
 * point provides; none of that code is reused.
 */
#ifndef SINGA_B200_H_
#define SINGA_B200_H_

#ifdef __cplusplus
extern "C" {
#endif

const char* sg_last_error(void);
/* Library/ABI version and device capability probe (returns compute capability major*10+minor, or <0). */
int sg_version(void);
int sg_device_cc(int device);

/* ------------------------------------------------------------------------------------------------
 * tcgen05 GEMM  (reference: mshadow dot() -> cblas/cublas sgemm, include/mshadow/tensor_expr_engine-inl.hpp:271-357;
 *                used by InnerProduct src/worker/layer.cc:186-213 and Convolution layer.cc:63-123)
 *
 *   acc[m][n] = sum_k A(m,k) * B(n,k)          bf16 operands, fp32 accumulation in TMEM
 *   A: a_mn_major == 0 -> stored [M][K] (row stride lda), == 1 -> stored [K][M] (row stride lda)
 *   B: b_mn_major == 0 -> stored [N][K] (row stride ldb), == 1 -> stored [K][N] (row stride ldb)
 *   lda/ldb must be multiples of 8 elements and the base pointers 16-byte aligned (TMA).
 *
 * Epilogue, applied in registers after tcgen05.ld, in this order:
 *   v = acc (+ bias[n]); v = act(v); v *= act'(aux[m][n]); v += d32_old (accumulate); store fp32 and/or bf16.
 * ---------------------------------------------------------------------------------------------- */
enum { SG_ACT_NONE = 0, SG_ACT_RELU = 1, SG_ACT_STANH = 2 };

typedef struct SgGemmEpilogue {
  float* d32;            /* optional fp32 output [M][ldd32] */
  long long ldd32;
  void* d16;             /* optional bf16 output [M][ldd16] */
  long long ldd16;
  const void* bias;      /* optional [N], fp32 (or bf16 when bias_bf16) */
  int act;               /* SG_ACT_* applied to acc+bias */
  int grad_act;          /* SG_ACT_*: multiply by the activation derivative evaluated from aux (= activation OUTPUT) */
  const void* aux;       /* [M][ldaux], bf16 if aux_bf16 else fp32 */
  long long ldaux;
  int aux_bf16;
  int accumulate;        /* d32 += result instead of = */
  int perm_c;            /* >0: column j=(kk_idx*perm_c + c) is stored at column c*perm_kk + kk_idx (conv wgrad -> {F, C*k*k}) */
  int perm_kk;
  /* Parameter pull fused into the GEMM (src/utils/param_manager.cc:204-234 WaitUpdate): before the first
   * operand load the TMA producer acquire-spins until every *wait_flags[i] >= *wait_seq — the version counters
   * the parameter-server push kernel bumps after broadcasting fresh weights into this GPU's bf16 mirror.
   * wait_flags is a DEVICE array of nwait flag addresses. */
  const unsigned long long* const* wait_flags;
  int nwait;
  const unsigned long long* wait_seq;
  int bias_bf16;         /* bias points at bf16 values (the parameter server's weight mirror) instead of fp32 */
  int tile_bn;           /* 0 = library's choice, else 64 / 128 / 256 (set by the shape autotuner) */
  int tile_pair;         /* 0 = library's choice, 1 = single-CTA tiles, 2 = CTA pairs */
  int split_k;           /* fp32-output GEMMs without bias/activation only: 0 = library's choice (splits the K loop
                            across CTAs when the output has too few tiles, partial tiles are added atomically into the
                            zeroed output), 1 = never, n > 1 = n slices */
} SgGemmEpilogue;

/* Operand precision.  SG_DTYPE_TF32: A and B hold fp32 values (lda/ldb multiples of 4) that the tensor core
 * reads as tf32 (sign, 8-bit exponent, top 10 mantissa bits) with fp32 accumulation — the parity path for the
 * reference's fp32 sgemm; outputs, bias and aux are fp32.  SG_DTYPE_BF16: bf16 operands (lda/ldb multiples of 8). */
enum { SG_DTYPE_BF16 = 0, SG_DTYPE_TF32 = 1 };
int sg_gemm(int dtype, const void* A, int a_mn_major, long long lda, const void* B, int b_mn_major,
            long long ldb, int M, int N, int K, const SgGemmEpilogue* ep, void* stream);
/* sg_gemm with SG_DTYPE_BF16 */
int sg_gemm_bf16(const void* A, int a_mn_major, long long lda, const void* B, int b_mn_major,
                 long long ldb, int M, int N, int K, const SgGemmEpilogue* ep, void* stream);

/* Test hook: force the N-tile width (64/128/256) of sg_gemm_bf16; 0 restores the heuristic. */
int sg_gemm_set_tile(int bn);
/* Tile shape policy: 0 = one CTA per 128 x BN tile, 1 = CTA pairs (cta_group::2) on 256 x BN tiles,
 * 2 = pairs for large problems only. */
int sg_gemm_set_pair(int mode);

/* ------------------------------------------------------------------------------------------------
 * Element-wise family (reference: every mshadow expression lowers to MapPlanKernel,
 * include/mshadow/cuda/tensor_gpu-inl.cuh:56-92).  is_bf16 selects bf16 vs fp32 element type.
 * ---------------------------------------------------------------------------------------------- */
/* ReLU / scaled tanh forward and backward-from-output (src/worker/layer.cc:556-569,688-701). */
int sg_act_fwd(const void* x, void* y, long long n, int act, int is_bf16, void* stream);
int sg_act_bwd(const void* dy, const void* y, void* dx, long long n, int act, int is_bf16, void* stream);
/* Dropout: y = x * (philox(seed', i) < pkeep) / pkeep.  Backward is the same call on dy (layer.cc:142-160).
 * seed_dev (optional, device memory) is mixed into the seed at run time so that a captured CUDA graph
 * draws a new mask per replay. */
int sg_dropout(const void* x, void* y, long long n, float pkeep, unsigned long long seed,
               const unsigned long long* seed_dev, int is_bf16, void* stream);
/* out[c] (+)= sum_r x[r][c]  — bias gradients (sum_rows / sumall_except_dim, layer.cc:107,206). */
int sg_colsum(const void* x, long long ld, int rows, int cols, float* out, int accumulate, int is_bf16,
              void* stream);
/* fp32 [rows][lds] -> bf16 [rows][ldd] (padding columns zeroed), optional conv-weight column permutation
 * (c,ky,kx) -> (ky,kx,c) given perm_c = C, perm_kk = k*k. */
int sg_cast_pack(const float* src, long long lds, void* dst, long long ldd, int rows, int cols,
                 int perm_c, int perm_kk, void* stream);
/* the same re-layout from a bf16 source (a rank that only holds the parameter server's bf16 weight mirror) */
int sg_repack_bf16(const void* src, long long lds, void* dst, long long ldd, int rows, int cols, int perm_c,
                   int perm_kk, void* stream);
int sg_convert(const void* src, int src_bf16, void* dst, int dst_bf16, long long n, void* stream);
/* [rows][lds] -> [rows][ldd] of one element type, padding columns zeroed (re-striding to a TMA-legal leading
 * dimension). */
int sg_copy2d(const void* src, long long lds, void* dst, long long ldd, int rows, int cols, int is_bf16, void* stream);
int sg_copy2d_bf16(const void* src, long long lds, void* dst, long long ldd, int rows, int cols, void* stream);
/* Inner product with 1..16 outputs (classifier layers) on CUDA cores; operands and layouts as for sg_gemm:
 * x [M][K], w [K][N] (the reference's {vdim, hdim} weight), y [M][N]; every operand fp32 or bf16 by its flag.
 *   fwd  : y = act(x . w + bias)            dgrad: dx = (dy . w^T) * act'(aux)   (aux = output of the layer below)
 *   wgrad: dw (+)= x^T . dy (fp32) and, when db is given, db (+)= column sums of dy */
int sg_ip_small_fwd(const void* x, long long ldx, int x_bf16, const void* w, long long ldw, int w_bf16, const void* bias,
                    int bias_bf16, void* y, long long ldy, int y_bf16, int M, int K, int N, int act, void* stream);
int sg_ip_small_dgrad(const void* dy, long long lddy, int dy_bf16, const void* w, long long ldw, int w_bf16,
                      const void* aux, long long ldaux, int aux_bf16, int grad_act, void* dx, long long lddx, int dx_bf16,
                      int M, int K, int N, void* stream);
/* ws: zero-initialised device buffer of sg_ip_small_wgrad_ws_bytes(M, K) bytes, private to the calls that may run
 * concurrently (partial sums per 32-row slice, added up in slice order by the last block: deterministic). */
long long sg_ip_small_wgrad_ws_bytes(int M, int K);
int sg_ip_small_wgrad(const void* x, long long ldx, int x_bf16, const void* dy, long long lddy, int dy_bf16, float* dw,
                      long long lddw, float* db, int M, int K, int N, int accumulate, void* ws, void* stream);
/* Space-to-depth for strided convolutions over thin (grey / RGB) inputs:
 *   xs[n][Y][X][(dy*s + dx)*C + c] = x[n][Y*s + dy - pad][X*s + dx - pad][c], zero outside the image and for the
 *   padding channels [s*s*C, Cs).  x: [N][H][W][C], xs: [N][Hs][Ws][Cs], both of one element type. */
int sg_space_to_depth(const void* x, void* xs, int N, int H, int W, int C, int Hs, int Ws, int Cs, int s, int pad,
                      int is_bf16, void* stream);
/* Column re-layouts of a [rows][*] matrix through an index list (idx[j] < 0: no source / dropped):
 *   gather : dst[r][j] = src[r][idx[j]] (or 0), fp32 or bf16 source, fp32 destination
 *   scatter: dst[r][idx[j]] (+)= src[r][j]; idx must not repeat a non-negative entry */
int sg_gather_cols(const void* src, int src_bf16, long long lds, float* dst, long long ldd, int rows, int ncols,
                   const int* idx, void* stream);
int sg_scatter_cols(const float* src, long long lds, float* dst, long long ldd, int rows, int ncols, const int* idx,
                    int accumulate, void* stream);
/* Per image [rows][cols] -> [cols][rows] (NCHW <-> NHWC). */
int sg_transpose_planes(const void* src, int src_bf16, void* dst, int dst_bf16, int batch, int rows,
                        int cols, void* stream);
/* out[0] = sum |x|  (Blob::asum_data, src/utils/blob.cc:251-255; the caller divides by count). */
int sg_asum(const void* x, long long n, float* out, int is_bf16, void* stream);
/* Philox fills for Param::Init (src/utils/param.cc:89-127): kind 0 uniform[a,b)*scale, 1 gaussian(a,b)*scale. */
int sg_fill_random(float* dst, long long n, int kind, float a, float b, float scale,
                   unsigned long long seed, void* stream);
/* Parsers (src/worker/layer.cc:382-448, 573-624): uint8 records -> normalised activations. */
int sg_mnist_parse(const void* src_u8, void* dst, long long n, float norm_a, float norm_b, int dst_bf16,
                   void* stream);
int sg_rgb_parse(const void* src_u8, const float* mean, void* dst, int batch, int C, int H, int W,
                 int crop, const int* off_h, const int* off_w, const int* mirror, float scale,
                 int dst_bf16, void* stream);
/* MNIST distortion (the reference's disabled block, layer.cc:406-438): dst[n][y][x] = bilinear(src[n]; A_n (x,y,1) +
 * disp[n][:, y, x]) / norm_a - norm_b with A_n = affine[n][0..5] (2x3) and disp (optional) an [n][2][out][out]
 * displacement field in source pixels. */
int sg_mnist_distort(const void* src, int src_f32, void* dst, int n_img, int in_size, int out_size,
                     const float* affine, const float* disp, float norm_a, float norm_b, int dst_bf16, void* stream);
/* the same with src_f32 != 0: the staged records are fp32 (float-valued records such as mean-subtracted images,
 * which the reference copies unchanged, layer.cc:595-598) */
int sg_mnist_parse_src(const void* src, int src_f32, void* dst, long long n, float norm_a, float norm_b,
                       int dst_bf16, void* stream);
int sg_rgb_parse_src(const void* src, int src_f32, const float* mean, void* dst, int batch, int C, int H, int W,
                     int crop, const int* off_h, const int* off_w, const int* mirror, float scale,
                     int dst_bf16, void* stream);

/* ------------------------------------------------------------------------------------------------
 * Softmax loss forward+backward in one pass (src/worker/layer.cc:718-764).
 *   metric[0] += scale * mean(-log max(p_label, FLT_MIN)),  metric[1] += scale * top-k accuracy
 *   grad = (p - onehot) * scale / rows   (grad, prob and metric are each optional)
 * ---------------------------------------------------------------------------------------------- */
int sg_softmax_loss(const float* logits, long long ldl, const int* labels, int rows, int dim, int topk,
                    float scale, float* prob, long long ldp, void* grad, long long ldg, int grad_bf16,
                    float* metric, void* stream);

/* Pooling (MAX / AVE) and across-channel LRN on channels-last activations
 * (src/worker/layer.cc:476-540 and 331-378; mshadow pool/unpool/chpool). */
int sg_pool_fwd(const void* x, void* y, int N, int H, int W, int C, int OH, int OW, int k, int stride,
                int pad, int is_max, int is_bf16, void* stream);
int sg_pool_bwd(const void* x, const void* y, const void* dy, void* dx, int N, int H, int W, int C,
                int OH, int OW, int k, int stride, int pad, int is_max, int is_bf16, void* stream);
int sg_lrn_fwd(const void* x, void* y, float* norm, long long pixels, int C, int lsize, float alpha,
               float beta, float knorm, int is_bf16, void* stream);
int sg_lrn_bwd(const void* x, const float* norm, const void* dy, void* dx, long long pixels, int C,
               int lsize, float alpha, float beta, int is_bf16, void* stream);

/* Explicit convolution lowering (src/worker/layer.cc:63-123): NHWC patches -> col[pixels][ld] with column
 * (ky*k+kx)*C + c, and the gather-form inverse for the data gradient (optionally fused with the activation gradient
 * of the layer below; aux = that layer's output, NHWC).  Only two cases still use it — everything else runs as
 * implicit GEMM (sg_conv_rows / sg_conv_wgrad): the weight gradient of convolutions whose pixels are narrower than
 * 64 bytes (RGB / grey inputs), and the data gradient of strided convolutions. */
int sg_im2col(const void* x, void* col, long long ld, int N, int H, int W, int C, int OH, int OW,
              int k, int stride, int pad, int is_bf16, void* stream);
int sg_col2im(const void* dcol, long long ld, void* dx, int N, int H, int W, int C, int OH, int OW,
              int k, int stride, int pad, int grad_act, const void* aux, int is_bf16, void* stream);
/* the bf16 forms */
int sg_im2col_nhwc(const void* x, void* col, long long ld, int N, int H, int W, int C, int OH, int OW,
                   int k, int stride, int pad, void* stream);
int sg_col2im_nhwc(const void* dcol, long long ld, void* dx, int N, int H, int W, int C, int OH, int OW,
                   int k, int stride, int pad, int grad_act, const void* aux, void* stream);

/* Implicit-GEMM convolution (src/worker/layer.cc:63-123): the patch matrix is built tile by tile inside the GEMM by
 * im2col-mode TMA loads from the channels-last activation; nothing is materialised in global memory.
 *   source tensor  [N][H][W][C] with pixels ldc elements apart (ldc * element size a multiple of 16 bytes)
 *   row grid       OH x OW positions per image; position (oy, ox) reads source pixel (oy*stride - pad_lo + ky, ...)
 * sg_conv_rows: acc[pixel][n] = sum_{ky,kx,c} src(pixel; ky,kx,c) * filt[n][(ky*k + kx) * Cp + c], followed by the
 *   sg_gemm epilogue on D[pixels][ncols].  filt is packed K-major with Cp = nchunk * chunk channels per tap (zero
 *   padded; rows padded to whole 128-byte k-blocks; ldf elements apart), chunk = chunk_bytes / element size,
 *   chunk_bytes in {16, 32, 64, 128} (sg_conv_chunk_bytes picks one for a channel count).
 *   forward:        src = X,  filt = W as [F][(ky,kx,c)],  pad_lo = pad
 *   data gradient   (stride 1): src = dY, row grid = the input image, pad_lo = k-1-pad,
 *                   filt = [C][(jy,jx,f)] with W's taps flipped
 * sg_conv_wgrad: dw[f][c*k*k + ky*k + kx] (+)= sum_pixel dy[pixel][f] * src(pixel; ky,kx,c)   ({F, C*k*k} layout of
 *   the reference), split over the pixels with fp32 atomics when there are few output tiles. */
typedef struct SgConvGeom {
  int N, H, W, C;        /* source tensor */
  long long ldc;         /* elements between pixels of the source (>= C) */
  int OH, OW;            /* row grid per image */
  int k, stride, pad_lo;
  int chunk_bytes;       /* sg_conv_rows only */
} SgConvGeom;
int sg_conv_chunk_bytes(int channels, int elem_bytes);
int sg_conv_rows(int dtype, const void* src, const SgConvGeom* g, const void* filt, long long ldf, int ncols,
                 const SgGemmEpilogue* ep, void* stream);
int sg_conv_wgrad(int dtype, const void* src, const SgConvGeom* g, const void* dy, long long lddy, int F, float* dw,
                  long long lddw, int accumulate, void* stream);
/* Filter packing for sg_conv_rows from the reference layout w[F][C*k*k] (column (c*k + ky)*k + kx; fp32 or bf16 source,
 * ldw elements per row): fwd[f][(ky*k+kx)*Cp_f + c] = w[f][c][ky][kx] and, when dgrad != NULL,
 * dgrad[c][(jy*k+jx)*Fp_d + f] = w[f][c][k-1-jy][k-1-jx]; padding written as zeros; outputs bf16 or fp32 (out_bf16). */
int sg_conv_pack(const void* w, int w_bf16, long long ldw, int F, int C, int k, void* fwd, long long ldfwd, int cp_f,
                 void* dgrad, long long lddgrad, int fp_d, int out_bf16, void* stream);

/* ------------------------------------------------------------------------------------------------
 * Updaters over the flat parameter arena (src/utils/updater.cc:54-182) and PS exchange rules
 * (src/utils/param.cc:130-284).  Arena offsets are multiples of SG_ARENA_CHUNK elements; chunk_param maps a
 * chunk to its parameter index (or -1 for padding); lr_mult / wd_mult are indexed by parameter.
 * ---------------------------------------------------------------------------------------------- */
#define SG_ARENA_CHUNK 1024
enum { SG_UPD_SGD = 0, SG_UPD_NESTEROV = 1, SG_UPD_ADAGRAD = 2, SG_UPD_RMSPROP = 3, SG_UPD_ADADELTA = 4 };
typedef struct SgUpdater {
  int kind;
  float lr;            /* LR(step), before the per-parameter multiplier */
  float momentum;
  float weight_decay;
  float delta;
  float rho;
  float grad_scale;
  const float* lr_dev; /* optional device scalar overriding `lr` (lets a captured CUDA graph follow the LR schedule) */
} SgUpdater;
int sg_update_arena(float* w, const float* g, float* h, float* u, void* w16, const int* chunk_param,
                    const float* lr_mult, const float* wd_mult, long long first_chunk,
                    long long nchunks, const SgUpdater* up, void* stream);
/* EASGD: d = (w - server) * alpha; server += d (atomic); w -= d.  `server` may be peer memory. */
int sg_easgd_sync(float* w, float* server, long long n, float alpha, void* w16, void* stream);
/* RandomSync on the shared-seed subset {i : philox(seed,i) < ratio}. */
int sg_randomsync(float* w, float* snap, float* server, long long n, float ratio,
                  unsigned long long seed, void* w16, unsigned long long* nselected, void* stream);

/* Sharded forms: ONE launch exchanges the whole local arena (n elements) with every server shard — shard k holds
 * elements [k*per, (k+1)*per) behind server[k], a (peer-mapped) arena base indexed with the same element index.
 * seed_dev (optional device word) is mixed into the RandomSync seed so a captured CUDA graph draws a new subset per
 * replay.  SG_PS_MAX_WORLD bounds the table. */
typedef struct SgShardTable {
  float* server[8];
  int nshards;
  long long per;
} SgShardTable;
int sg_easgd_sync_sharded(float* w, const SgShardTable* table, long long n, float alpha, void* w16, void* stream);
int sg_randomsync_sharded(float* w, float* snap, const SgShardTable* table, long long n, float ratio,
                          unsigned long long seed, const unsigned long long* seed_dev, void* w16,
                          unsigned long long* nselected, void* stream);
/* Server-group neighbour synchronisation (src/server/pm_server.cc:105-127,296-354, src/utils/param_shard.cc:52-58):
 * mine[lo, lo+n) moves to the mean of itself and the same elements of the peers->nshards neighbour groups' copies
 * (peer loads; applied as an atomic add of the difference, so concurrent worker exchanges are kept). */
int sg_group_average(float* mine, const SgShardTable* peers, long long lo, long long n, void* stream);

/* ------------------------------------------------------------------------------------------------
 * Parameter-server data plane over NVSwitch peer memory (gradient push = reduce-scatter + server Updater +
 * weight broadcast in ONE kernel; device-side version flags instead of kSync / WaitUpdate messages:
 * src/utils/param_manager.cc:163-234, src/server/server.cc:18-43, src/server/pm_server.cc:262-294).
 * All pointers in the per-rank arrays are device pointers valid on the calling GPU (local or IPC-mapped peer
 * memory).  Flag value = (*step_dev if given else 0) + value_add.
 * ---------------------------------------------------------------------------------------------- */
#define SG_PS_MAX_WORLD 8
#define SG_PS_MAX_WAIT 32
typedef struct SgPsPush {
  int world;
  const float* grad[SG_PS_MAX_WORLD];               /* every rank's gradient arena base */
  void* w16[SG_PS_MAX_WORLD];                       /* every rank's weight mirror base: bf16, or fp32 when mirror_fp32
                                                       (NULL = that rank keeps no mirror) */
  const unsigned long long* ready[SG_PS_MAX_WORLD]; /* rank r's "gradients of this layer are ready" flag (or NULL) */
  unsigned long long* version[SG_PS_MAX_WORLD];     /* rank r's version slot for (layer, this owner) (or NULL) */
  float* w; float* h; float* u;                     /* local fp32 master / history / update arenas */
  const int* chunk_param; const float* lr_mult; const float* wd_mult;
  long long first_chunk, nchunks;                   /* chunk range updated by this launch (owned by this rank) */
  SgUpdater up;
  const long long* step_dev; long long value_add;   /* explicit flag value (used when self_count == 0) */
  unsigned long long* scratch;                      /* 16 bytes of device memory, zero-initialised once:
                                                       [0] CTA arrival counter, [1] launch sequence number */
  int self_count;                                   /* 1: flag value = scratch[1] + 1 (n-th launch waits for ready >= n,
                                                       publishes version = n); independent of any step scalar */
  int max_ctas;                                     /* CTA budget (0 = 64) so backward keeps the SMs */
  const float* grad_mc;                             /* optional NVLS multicast address of all ranks' gradient arenas:
                                                       the reduction then happens inside the NVSwitch (multimem.ld_reduce) */
  void* w16_mc;                                     /* optional NVLS multicast address of all ranks' bf16 mirrors:
                                                       one multimem.st broadcasts the fresh weights */
  unsigned long long* bump_ready;                   /* optional: this rank's own ready flag; the kernel increments it
                                                       (and *bump_seq, the consumer's wait sequence) before waiting,
                                                       replacing a separate sg_ps_flag_bump launch */
  unsigned long long* bump_seq;
  int mirror_fp32;                                  /* mirrors (w16, w16_mc) hold fp32 values: the tf32 compute path reads
                                                       its GEMM operands in fp32 */
  unsigned long long* trace;                        /* optional (profiling): 8 device words that receive %globaltimer
                                                       stamps of the launch: [0] start, [1] own ready flag published,
                                                       [2] every rank ready, [3] CTA 0 done with its chunks, [4] its stores
                                                       fenced, [5] last CTA arrived, [6] version flags published */
} SgPsPush;
int sg_ps_push_update(const SgPsPush* args, void* stream);
/* Allocation-free variant of sg_ps_push_update — no counterpart in the reference (its parameter server keeps
 * all state in host objects).  For callers that cannot create device memory (a host application whose memory
 * manager owns all of HBM, a callback where cudaMalloc is forbidden, a loop that must never block on the
 * allocator): the kernel's scratch words live in a device array built into the library, SG_PS_NOALLOC_SLOTS
 * slots of 256 bytes (16 KB per GPU for the life of the process).  `slot` selects the words; slots beyond the
 * maximum are rejected.  args->scratch is ignored; args->self_count must be 0 (explicit flag value).  The
 * first call on a device looks up the array's device address — make it before any stream capture.  The
 * scratch is cleared on `stream` before each launch; launches that share a slot and may overlap on the GPU
 * race.  Calls no cudaMalloc/cudaFree or other allocating routine. */
#define SG_PS_NOALLOC_SLOTS 64
int sg_ps_push_update_noalloc(const SgPsPush* args, int slot, void* stream);
/* release-store / acquire-wait on 64-bit flags in (peer) device memory */
int sg_ps_flag_set(unsigned long long* flag, const long long* step_dev, long long value_add, void* stream);
/* flag += 1 (single writer); local_seq (optional, local memory) += 1 as well — the per-layer sequence number
 * that sg_gemm_bf16's fused weight wait reads (SgGemmEpilogue.wait_seq). */
int sg_ps_flag_bump(unsigned long long* flag, unsigned long long* local_seq, void* stream);
/* seq != NULL: wait for flags >= *seq (the sequence number sg_ps_flag_bump advances), else for the explicit value */
int sg_ps_flag_wait(const unsigned long long* const* flags, int n, const long long* step_dev,
                    long long value_add, const unsigned long long* seq, void* stream);
/* CUDA-IPC arenas: these (and only these) entry points allocate / free device memory. */
int sg_ipc_alloc(long long bytes, void** out);
int sg_ipc_free(void* p);
int sg_ipc_get_handle(void* p, void* handle64);
int sg_ipc_open_handle(const void* handle64, void** out);
int sg_ipc_close_handle(void* p);

/* ------------------------------------------------------------------------------------------------
 * Shard record file (host code; include/utils/shard.h:33-142, src/utils/shard.cc:24-206):
 * <folder>/shard.dat = [u64 keylen][key][u64 vallen][val]...; mode 'r' read, 'c' create, 'a' append (recovers
 * from a torn tail, ignores duplicate keys).  sg_shard_read_image_batch decodes singa.Record image records
 * straight into batch buffers (ShardDataLayer, src/worker/layer.cc:646-661).
 * ---------------------------------------------------------------------------------------------- */
void* sg_shard_open(const char* folder, char mode, long long capacity);
void sg_shard_close(void* shard);
int sg_shard_insert(void* shard, const void* key, long long klen, const void* val, long long vlen);
int sg_shard_flush(void* shard);
int sg_shard_next(void* shard);
const void* sg_shard_key(void* shard, long long* len);
const void* sg_shard_val(void* shard, long long* len);
void sg_shard_seek_to_first(void* shard);
long long sg_shard_count(void* shard);
int sg_shard_read_image_batch(void* shard, void* pixels_u8, int* labels, int batch, long long dim);
/* fp32 batch: byte-valued and float-valued records alike, values unchanged.  The uint8 form fails (-4) on a
 * float-valued record instead of narrowing it. */
int sg_shard_read_image_batch_f32(void* shard, float* pixels, int* labels, int batch, long long dim);

/* ------------------------------------------------------------------------------------------------
 * Read-only LMDB environment (host code, no liblmdb): B+tree walk over an mmap of <dir>/data.mdb in key order,
 * Caffe Datum values decoded straight into batch buffers, wrap-around at the end (LMDBDataLayer,
 * src/worker/layer.cc:237-328).  sg_lmdb_next: 1 = record current, 0 = end (next call restarts), -1 = corrupt.
 * ---------------------------------------------------------------------------------------------- */
void* sg_lmdb_open(const char* dir_or_file);
void sg_lmdb_close(void* env);
long long sg_lmdb_count(void* env);
void sg_lmdb_seek_to_first(void* env);
int sg_lmdb_next(void* env);
const void* sg_lmdb_key(void* env, long long* len);
const void* sg_lmdb_val(void* env, long long* len);
int sg_lmdb_first_shape(void* env, int* chw);
int sg_lmdb_read_datum_batch(void* env, void* pixels_u8, int* labels, int batch, long long dim);
/* fp32 batch (Datum.data bytes widened, Datum.float_data unchanged); the uint8 form returns 3 on float_data.
 * sg_lmdb_first_is_float: 1 / 0 for the first record, <0 on error. */
int sg_lmdb_read_datum_batch_f32(void* env, float* pixels, int* labels, int batch, long long dim);
int sg_lmdb_first_is_float(void* env);

#ifdef __cplusplus
}
#endif
#endif /* SINGA_B200_H_ */
