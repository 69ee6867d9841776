/* arbius_b200 native C ABI (libarbius_b200.so, built from csrc/ for sm_100a).
 *
 * Conventions
 *   - every entry point returns 0 on success, non-zero on error; ab_last_error() gives the text.
 *   - `stream` is a cudaStream_t passed as void* (torch.cuda.current_stream().cuda_stream).
 *   - activations are channels-last everywhere: images are [N, H, W, C], token tensors [M, C].
 *   - dtype: 0 = fp16, 1 = bf16 (storage); accumulation and statistics are fp32.
 *   - no entry point allocates device memory unless its comment says so: callers pass workspaces.
 *   - reductions run in a fixed order that does not depend on grid size or GPU count
 *     (the canonical 8-shard rule); there are no atomics on value paths.
 *
 * The reference has no native code: "ref" notes below name the op the external
 * Cog containers would dispatch to (SURVEY §2.9 row ids K1..K17).
 */
#ifndef ARBIUS_B200_H
#define ARBIUS_B200_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

const char* ab_last_error(void);
int ab_version(void);
/* number of kernel launches issued through this library since load (bench.py gpu_launches) */
unsigned long long ab_launch_count(void);
/* programmatic dependent launch between consecutive kernels of this library (default off; AB_PDL=1 enables) */
int ab_pdl_enabled(void);
void ab_set_pdl(int on);

enum { AB_F16 = 0, AB_BF16 = 1 };
enum {
  AB_ACT_NONE = 0, AB_ACT_SILU = 1, AB_ACT_GELU = 2, AB_ACT_QUICK_GELU = 3, AB_ACT_GELU_TANH = 4,
  AB_ACT_SIGMOID = 5, AB_ACT_TANH = 6, AB_ACT_RELU = 7, AB_ACT_HARDSWISH = 8, AB_ACT_HARDSIGMOID = 9,
  AB_ACT_RELU6 = 10
};

/* ---- K2/K3/K1: tcgen05 GEMM / implicit-GEMM conv ------------------------------------------------
 * out[M, Nout] = epilogue( A[M, K] * W[N, K]^T )
 *   linear mode (conv == 0): A = a [M, K1] (and optionally a2 [M, K2] appended along K: virtual concat)
 *   conv mode   (conv == 1): A is the implicit im2col of the NHWC image a [nimg, H, W, K1] (+ a2 with K2
 *                channels, virtual channel concat) for a 3x3 / pad 1 / stride 1 kernel. M = nimg*H*W.
 *                W is [N, 9*(K1+K2)] with K ordered (kh, kw, cin).
 *   epilogue: v = acc + bias[n] + rowbias[m / rows_per_group, n]; v = act(v);
 *             geglu: out[m, j] = v[value j] * gelu(v[gate j]) with value/gate columns interleaved in
 *             blocks of 16 inside W (see ops/gemm.py: pack_geglu); Nout = N/2.
 *             v = v * out_scale + residual[m, n]
 *   splitk > 1: K is cut into `splitk` equal chunks, fp32 partial tiles go to `workspace`
 *               (>= splitk*M*N*4 bytes) and are summed in split order by a second kernel.
 * ref: cuBLAS HGEMM + ATen elementwise / cuDNN implicit GEMM conv.            */
typedef struct {
  const void* a; const void* a2; const void* w; void* out;
  const void* bias; const void* rowbias; const void* residual;
  void* workspace;
  int M, N, K1, K2;
  int conv, nimg, H, W;
  int rows_per_group;       /* rows sharing one rowbias row (H*W for a time-embedding add) */
  int act, geglu;
  float out_scale;          /* 0 is treated as 1 */
  int dtype;
  int splitk;
  int block_n;              /* 0 = auto; else one of 32, 64, 96, 128, 160, 256 */
  int ldo;                  /* output row stride in elements, 0 = Nout */
  int ldr;                  /* residual row stride in elements, 0 = Nout */
  int ldrb;                 /* rowbias row stride in elements, 0 = N */
  int halo;                 /* conv modes: A (and A2) hold `halo` extra rows of H before and after every image
                             * ([nimg, H + 2*halo, W, K], H counts output rows) that are read as data instead of the
                             * zero padding: a row band of an image (patch parallelism) or the local frames of a clip
                             * (frame parallelism) with the neighbours' edge rows copied in from the GPUs next door.
                             * 0 or 1. */
  /* optional cross-GPU wait folded into the kernel (see "cross-GPU flags" below): before its first load the TMA
   * producer thread spins until every rank in sync_wait_mask has published *sync_tick in this GPU's copy of the
   * site — the neighbours' halo rows, pushed straight into A's halo slots, have then arrived. NULL = no wait. */
  const unsigned int* sync_tick; const unsigned int* sync_wait; unsigned int sync_wait_mask;
  unsigned int sync_top_mask;   /* conv + halo: the ranks of sync_wait_mask that fill the TOP halo row (the rest fill the
                                 * bottom one); only tiles that touch a halo row wait, and only for that neighbour */
  /* conv == 1 only: resampling folded into the conv's TMA coordinates (no gathered or upsampled copy of the input).
   *   0  none (3x3, stride 1, pad 1)
   *   1  stride 2: H, W are the OUTPUT size, A is [nimg, Hin, Win, K]; tap (dy, dx) of output (y, x) reads input
   *      (2y + dy - pad_lo, 2x + dx - pad_lo) through a tensor map whose W and H traversal strides are 2.
   *      pad_lo = 1: pad 1 all round (Hin = 2H or 2H - 1); pad_lo = 0: pad right / bottom only (Hin = 2H + 1 ... ).
   *   2  nearest 2x upsample before the conv: H, W are the INPUT size, the output is [nimg, 2H, 2W, N]. A tile holds the
   *      outputs of one parity class (Y = 2y + py, X = 2x + px) of a low-resolution box, so that every tap reads a plain
   *      box of the low-resolution input at offset ((py + dy - 1) >> 1, (px + dx - 1) >> 1).
   * Both give the bytes of the unfused op (same taps, same K order, same split-K). Not combined with halo / a2 / sync. */
  int resample; int pad_lo; int Hin; int Win;
} ab_gemm_args;
int ab_gemm(const ab_gemm_args* args, void* stream);
/* Multi-GPU forms of the same kernel (one NVSwitch box; pointers may be peer-mapped addresses).
 *   all-gather -> GEMM:  nsrc > 1: the rows of A live in nsrc buffers of rows_per_src rows each (args->a is
 *                        ignored, M = nsrc*rows_per_src); tiles are fetched by TMA from whichever GPU owns them.
 *   GEMM -> reduce-scatter: rs_world >= 1: K is cut into 8 canonical shards, this rank (rs_rank) computes
 *                        8/rs_world of them over its local K slice and stores each fp32 partial row into the
 *                        workspace of the rank that owns the row (ws_peers[owner], >= 8*(M/rs_world)*N*4 bytes).
 *                        phase 1 = partials only; (peer barrier); phase 2 = owners sum shards 0..7 in order and
 *                        apply the epilogue to their M/rs_world rows (args->out / residual are that local band).
 *                        phase 0 = both (rs_world == 1). Results are bit-identical for rs_world in {1,2,4,8}.
 * No counterpart in the reference (single GPU, no collectives: SURVEY section 2.8).                         */
typedef struct {
  const void* a_srcs[8]; int nsrc; int rows_per_src;
  void* ws_peers[8]; int rs_world; int rs_rank;
} ab_gemm_dist_args;
int ab_gemm_dist(const ab_gemm_args* args, const ab_gemm_dist_args* dist, int phase, void* stream);
size_t ab_gemm_workspace_bytes(const ab_gemm_args* args);

/* ---- K1 small-channel direct conv (conv_in / conv_out, Cin or Cout < 64) ------------------------ */
/* x [N,H,W,Cin] -> out [N,Ho,Wo,Cout], pad = KH/2, stride 1 or 2. Weight layout by shape class:
 * Cout <= 8 and Cin % 8 == 0: [Cout, KH, KW, Cin]; otherwise (Cin <= 16, Cout % 8 == 0): [KH, KW, Cin, Cout]. */
int ab_conv_direct(const void* x, const void* w, const void* bias, void* out, int n, int h, int wd, int cin,
                   int cout, int ksize, int stride, int act, int dtype, void* stream);

/* ---- K4: GroupNorm (+SiLU) on channels-last tensors --------------------------------------------
 * x [N, HW, C1] (++ x2 [N, HW, C2] virtual channel concat) -> y [N, HW, C1+C2].
 * Statistics: HW is cut into 8 canonical row shards; each shard is reduced by a fixed number of
 * blocks in a fixed order, partials go to `partials` (>= ab_groupnorm_partials_bytes) and are
 * combined in (shard, block) order by every apply block.  ref: ATen group_norm + silu.            */
typedef struct {
  const void* x; const void* x2; void* y;
  const void* gamma; const void* beta;
  void* partials;
  int n, hw, c1, c2, groups;
  float eps;
  int act;
  int dtype;
  /* optional per-(n, channel) scale/shift applied after the affine: y = y*(1+scale)+shift */
  const void* scale; const void* shift;
  int ss_stride;            /* row stride of scale/shift in elements, 0 = C1+C2 */
} ab_groupnorm_args;
int ab_groupnorm(const ab_groupnorm_args* args, void* stream);
size_t ab_groupnorm_partials_bytes(int n, int groups);
/* One sample's rows spread over world = 2, 4 or 8 GPUs (rank r holds canonical shards [r*8/world, (r+1)*8/world)):
 * stats writes this GPU's partial moments into peer-visible memory ([n][8][sub][groups] x 16 B; see
 * ab_groupnorm_partials_bytes for an upper bound), the GPUs meet at a barrier, apply merges all partials in canonical
 * order (part_srcs[r] = rank r's array, r < world) and normalises the local rows. args->hw is the local row count;
 * args->partials the usual local workspace; y_batch_stride (elements, 0 = hw*C) lets y be the interior of a padded
 * buffer. Output bytes equal ab_groupnorm on the undivided tensor. */
int ab_groupnorm_sharded_stats(const ab_groupnorm_args* args, int world, int rank, void* partials_out, void* stream);
int ab_groupnorm_sharded_apply(const ab_groupnorm_args* args, int world, void* const* part_srcs, long long y_batch_stride,
                               void* stream);
/* ---- cross-GPU flags: patch parallelism without barrier launches (csrc/patch.cu) --------------------------------
 * A row-band split of one image over 2/4/8 GPUs needs ~200 producer->consumer hand-offs per UNet step. Instead of a
 * barrier kernel on each side of every one, producers publish a flag behind their data and consumers wait inside the
 * kernel that needs the data. A "site" is 16 u32 at the same heap offset on every GPU (word r: written only by rank r,
 * word 8: local arrival counter); flags carry a "tick" that every GPU advances once per UNet step (ab_sync_tick), so a
 * captured CUDA graph replays correctly. No counterpart in the reference (single GPU). */
int ab_sync_tick(unsigned int* tick, void* stream);
/* publish *tick into word `rank` of each non-NULL site in peer_sites (entry d = rank d's copy of the site) */
int ab_sync_signal(const unsigned int* tick, unsigned int* const* peer_sites, int world, int rank, void* stream);
/* spin until every rank in mask has published *tick in this GPU's copy of the site */
int ab_sync_wait(const unsigned int* tick, const unsigned int* site, unsigned int mask, void* stream);
/* Strided copies whose source or destination may be peer memory, with the hand-off folded in: every CTA first waits
 * (wait_site/wait_mask, optional), the segments are copied (rows x width_bytes each, 16-byte units), and the last CTA
 * to finish publishes *tick to the sites in signal_sites (optional). `local_site` is this GPU's copy of the signalled
 * site (its word 8 counts arrivals). Used for halo rows, for all-gathering the UNet's output bands and for gathering a
 * level that is too small to split. */
typedef struct { void* dst; const void* src; long long dst_pitch, src_pitch, width_bytes; int rows; } ab_copy_seg;
typedef struct {
  ab_copy_seg seg[16]; int nseg;
  const unsigned int* tick;
  const unsigned int* wait_site; unsigned int wait_mask;
  unsigned int* signal_sites[8]; int world, rank;
  unsigned int* local_site;
} ab_patch_copy_args;
int ab_patch_copy(const ab_patch_copy_args* args, void* stream);
/* NVLS broadcast: copy nbytes (16-byte units) from src to the MULTICAST mapping dst_mc of a buffer bound on every GPU
 * of the subset (parallel/nvls.py: cuMulticastCreate / cuMulticastBindMem); the NVSwitch replicates each multimem.st
 * into every GPU's copy, so the producer's links carry each byte once. Stores only — values are copied, never reduced in
 * the switch (its reduction order is the hardware's, and every exchange here feeds the CID). */
int ab_multimem_copy(void* dst_mc, const void* src, long long nbytes, void* stream);
/* GroupNorm of a row band in ONE launch (patch parallelism): the partial moments of this GPU's canonical shards go to
 * every GPU's array as self-validating 16-byte entries {mean, tick, M2, tick} (the flag travels inside the data: no
 * fence, no separate publish, no grid barrier), every GPU polls the entries of all 8 shards, merges them in canonical
 * order and normalises its rows. y may be the interior of a [n, H+2, W, C] halo buffer (y_batch_stride); with
 * halo_rows > 0 the band's first / last halo_rows rows are also stored into the neighbours' halo slots (up_halo /
 * down_halo: address of sample 0's slot on that GPU, batch stride y_batch_stride) and the last CTA publishes the tick
 * to done_sites (the consumers' wait: the next conv's producer warp). args->hw is the LOCAL row count; args->partials
 * is not used. Output bytes equal ab_groupnorm on the undivided tensor. dry != 0: no waits, no signals (module warm-up). */
typedef struct {
  int world, rank;
  void* part_peers[8];            /* every rank's [n][8][sub][groups] x 16 B array (peer-mapped; [rank] = local), zero-filled once */
  const unsigned int* tick;
  long long y_batch_stride;       /* elements; 0 = hw * C */
  void* up_halo; void* down_halo; int halo_rows;
  unsigned int* done_sites[8];    /* who gets the "rows written" tick: neighbours (halo) or everyone (gather); NULL = skip */
  unsigned int* done_local;       /* this GPU's copy of the done site */
  int dry;
} ab_gn_patch_args;
int ab_groupnorm_patch(const ab_groupnorm_args* args, const ab_gn_patch_args* patch, void* stream);

/* Same operation as ab_groupnorm for callers that cannot allocate device memory (a host whose arena
 * already owns all of HBM, a plain C caller with no allocator, a loop that cannot take the device-wide
 * sync of cudaMalloc). args->partials is ignored: the statistics scratch is a ~0.5 MB array compiled
 * into the library, resident on every GPU that loads it until exit. Limits: n <= 4 images and
 * groups <= 32 per call, otherwise an error is returned. The first call on each device resolves the
 * array's device address; make that call outside any stream capture. All calls share the array: the
 * counters are cleared on `stream` before each launch, and calls that overlap on one GPU race, so
 * serialise them. Calls no cudaMalloc/cudaFree. No counterpart in the reference (it has no native code). */
int ab_groupnorm_static(const ab_groupnorm_args* args, void* stream);

/* ---- K11/K14/K16 helpers (csrc/elementwise2.cu), all NHWC, C % 8 == 0 ----------------------------------- */
/* 2x2 average pooling (guided-diffusion "resblock down") */
int ab_avgpool2x(const void* x, void* y, int n, int h, int w, int c, int dtype, void* stream);
/* MoVQ SpatialNorm tail: out = act(x * ymap[up] + bmap[up]); ymap/bmap [N, lh, lw, C] nearest-upsampled to [h, w] */
int ab_spatial_modulate(const void* x, const void* ymap, const void* bmap, void* out, int n, int h, int w, int c,
                        int lh, int lw, int act, int dtype, void* stream);
/* RVM ConvGRU: out = (1 - z) * hprev + z * tanh(cand) */
int ab_gru_update(const void* z, const void* cand, const void* hprev, void* out, long long n, int dtype, void* stream);
/* depthwise KxK conv (K = 3 | 5), stride 1 | 2, dilation, pad = (K/2)*dilation; weights [K*K, C]; fused bias + act */
int ab_dwconv(const void* x, const void* w, const void* bias, void* y, int n, int h, int wd, int c, int ksize,
              int stride, int dilation, int act, int dtype, void* stream);
/* global average pool [N, HW, C] -> [N, C]; per-(n, channel) scaling y = x * s[n, c] (squeeze-excite) */
int ab_global_avgpool(const void* x, void* y, int n, int hw, int c, int dtype, void* stream);
int ab_channel_scale(const void* x, const void* s, void* y, int n, int hw, int c, int dtype, void* stream);

/* ---- K5: LayerNorm over the last dim ------------------------------------------------------------ */
int ab_layernorm(const void* x, const void* gamma, const void* beta, void* y, int rows, int cols, float eps,
                 int dtype, void* stream);

/* row softmax over the last dim (VAE single-head attention with head_dim 512), in place allowed */
int ab_softmax_rows(const void* x, void* y, int rows, int cols, float scale, int dtype, void* stream);
/* out[c, r] = x[r, c] for a [rows, cols] matrix (batched over `batch` matrices) */
int ab_transpose2d(const void* x, void* y, int batch, int rows, int cols, int dtype, void* stream);
/* token + position embedding gather: out[b, l, :] = tok[ids[b,l], :] + pos[l, :] */
int ab_embedding(const long long* ids, const void* tok, const void* pos, void* out, int batch, int len, int dim,
                 int dtype, void* stream);

/* ---- K6/K7: flash attention (tcgen05), q [B, Lq, H, D], k/v [B, Lk, H, D] token-major ------------
 * strides are in elements: element (b, l, h, d) of q lives at q + b*q_bs + l*q_ls + h*D + d.       */
typedef struct {
  const void* q; const void* k; const void* v; void* out;
  int batch, heads, lq, lk, head_dim;
  long long q_bs, q_ls, k_bs, k_ls, v_bs, v_ls, o_bs, o_ls;
  float scale;
  int causal;               /* bit 0: causal mask. bit 1: pick the kernel as for lq > 128 (two query tiles per CTA) even when
                             * this call's lq is not: a rank that holds part of a longer query sequence (frames of a clip)
                             * must run the kernel the whole sequence gets, the two kernels round differently */
  int dtype;
} ab_attn_args;
int ab_flash_attn(const ab_attn_args* args, void* stream);
/* patch-parallel attention: K/V of source s at k_ptrs[s]/v_ptrs[s] ([B, lk_per_src, H, D], strides from args->k_*),
 * global key order = source 0, 1, ...; head_dim <= 64; bit-identical to ab_flash_attn on the concatenation.
 * No counterpart in the reference. */
int ab_flash_attn_multi(const ab_attn_args* args, const void* const* k_ptrs, const void* const* v_ptrs, int nsrc,
                        int lk_per_src, void* stream);
/* Segmented keys: [prefix | pixels]. The prefix (text / image-embedding tokens, replicated on every GPU) fills its
 * own KV tiles; the pixel keys follow in tiles of 128 counted from pixel 0 and live in nsrc buffers of rows_per_src
 * rows each (row bands of a patch-parallel subset; rows_per_src % 8 == 0; a tile that straddles two GPUs is fetched
 * in 8-row boxes). Every buffer is a fused projection [B, rows, heads_total, D]: K of head h is head k_head0 + h, V
 * is v_head0 + h. wait_site/tick (optional): the producer warp waits for source s's flag before its first load from
 * it (s != self_rank). The tile sequence does not depend on nsrc, so the bytes do not either. head_dim <= 64.
 * No counterpart in the reference. */
typedef struct {
  const void* q; void* out;
  int batch, heads, lq, head_dim;
  long long q_bs, q_ls, o_bs, o_ls;
  float scale; int dtype;
  const void* prefix; int prefix_len; long long prefix_bs, prefix_ls; int prefix_heads_total, prefix_k_head0, prefix_v_head0;
  const void* pix[8]; int nsrc, rows_per_src; long long pix_bs, pix_ls; int heads_total, k_head0, v_head0;
  const unsigned int* tick; const unsigned int* wait_site; int self_rank;
} ab_attn_segs_args;
int ab_flash_attn_segs(const ab_attn_segs_args* args, void* stream);
/* Profiling aids (never needed for results). ab_gemm_set_debug: a device buffer of >= 8 * grid_size int64; every CTA of
 * the next tcgen05 GEMM / conv launches writes clock64() offsets {setup, first operands, MMAs issued, accumulator
 * ready, epilogue done, TMEM released, first chunk in registers, first chunk stored} (scripts/gemm_phases.py).
 * ab_attn_set_debug: >= 4096 int64; CTA (0,0,0) of the head-dim <= 64 flash attention kernel writes a timeline of its
 * first 32 KV tiles (scripts/attn_timeline.py). NULL switches either off. Not thread-safe; not for use under capture. */
void ab_gemm_set_debug(long long* buf);
/* Large plain GEMMs (no conv / concat / row bias / split-K, M >= min_rows) run on the two-SM kernel (csrc/gemm_2sm.cu:
 * tcgen05.mma.cta_group::2 on 256-row tiles, M-persistent, double-buffered accumulators). Same bytes as the one-SM kernel.
 * mode 0 / 1 = off / on, -1 = the default (on unless AB_GEMM_2SM=0); min_rows < 0 = the default (12288). */
void ab_gemm_set_2sm(int mode, int min_rows);
void ab_attn_set_debug(long long* buf);
/* ---- K8: temporal self-attention of the video UNets (csrc/temporal_attn.cu) -----------------------------
 * q/k/v/out are channels-last video tensors [B, F_s, HW, heads*D] (token rows of `row_stride` elements, so q, k, v
 * may be column slices of one fused QKV projection). Every pixel attends over its F frames. With nsrc > 1 the
 * frame axis is sharded over nsrc buffers (sequence parallelism; pointers may be peer-mapped): the kernel pulls
 * each source's frames into its tile and scatters output rows back to the buffer that owns the frame.
 * ref: SDPA on a permuted [B*HW, F, C] tensor inside diffusers' TransformerTemporalModel.                    */
typedef struct {
  const void* q[8]; const void* k[8]; const void* v[8]; void* out[8];
  int nsrc;                 /* 0 or 1 = single buffer */
  int batch, heads, frames, hw, head_dim;
  long long row_stride;     /* elements between consecutive tokens of q/k/v */
  long long out_row_stride;
  float scale;              /* 0 = 1/sqrt(head_dim) */
  int dtype;
  int pix_begin, pix_count; /* this launch handles pixels [pix_begin, pix_begin + pix_count) of hw (0, 0 = all): with
                             * frames sharded over GPUs, every GPU takes a pixel band, reads that band's Q/K/V frames
                             * from all sources and stores each frame's output to the GPU that owns the frame */
} ab_temporal_attn_args;
int ab_temporal_attn(const ab_temporal_attn_args* args, void* stream);

/* peer-memory runtime (csrc/peer.cu): cudaMalloc'ed heap + CUDA IPC mapping + device-side barrier */
int ab_heap_alloc(void** ptr, size_t bytes);
int ab_heap_free(void* ptr);
int ab_ipc_get_handle(void* ptr, unsigned char* handle64);
int ab_ipc_open_handle(const unsigned char* handle64, void** ptr);
int ab_ipc_close_handle(void* ptr);
int ab_peer_barrier(void* const* flag_ptrs, int rank, int world, unsigned int epoch, void* stream);
/* same barrier with the epoch kept in device memory (own flag words), so it can be captured in a CUDA graph and replayed */
int ab_peer_barrier_auto(void* const* flag_ptrs, int rank, int world, void* stream);
/* rows x width_bytes strided copy (16-byte units); src or dst may be a peer-mapped address, src == NULL writes zeros */
int ab_peer_copy2d(void* dst, long long dst_pitch, const void* src, long long src_pitch, long long width_bytes,
                   int rows, void* stream);

/* ---- K12/K15/G1/G3/S11: elementwise ------------------------------------------------------------- */
/* counter-based N(0,1): out[i] = normal(philox4x32-10(seed, stream_id, offset + i)); i is the GLOBAL
 * element index so any shard of the tensor draws the same values. out is fp32. */
int ab_philox_normal(float* out, long long n, unsigned long long seed, unsigned long long stream_id,
                     long long offset, void* stream);
/* eps = uncond + g*(cond - uncond); x_out = ca*x + cb*eps + cc*aux + cn*noise (all fp32 math).
 * eps_c/eps_u are model dtype [n]; x, aux, noise, x_out fp32; xm_out is x_out cast to model dtype. */
int ab_cfg_step(const void* eps_c, const void* eps_u, const float* x, const float* aux, const float* noise,
                float* x_out, void* xm_out, float* eps_out, long long n, float g, float ca, float cb, float cc,
                float cn, int dtype, void* stream);
/* scheduler update: dst[i] = sum_k coefs[k]*srcs[k][i] (fp32, left to right, nsrc <= 6); dst_m (optional)
 * receives dst*m_scale in the model dtype (the next step's scaled model input). dst may alias a source. */
int ab_lincomb(float* dst, void* dst_m, float m_scale, const float* const* srcs, const float* coefs, int nsrc,
               long long n, int dtype, void* stream);
/* (x/2+0.5).clamp(0,1)*255 -> uint8, NHWC in, NHWC (HWC) out; round-half-to-even */
int ab_finalize_u8(const void* x, unsigned char* out, long long n, int dtype, void* stream);
/* nearest-neighbour 2x upsample, NHWC */
int ab_upsample_nearest2x(const void* x, void* y, int n, int h, int w, int c, int dtype, void* stream);
/* bilinear resize NHWC (align_corners = 0 or 1) */
int ab_resize_bilinear(const void* x, void* y, int n, int h, int w, int c, int oh, int ow, int align_corners,
                       int dtype, void* stream);
/* im2col gather for strided 3x3 conv: x [N,H,W,C] -> col [N*Ho*Wo, 9*C] (kh,kw,c), pad 1 */
int ab_im2col3x3(const void* x, void* col, int n, int h, int w, int c, int stride, int pad_mode, int dtype,
                 void* stream);
/* y = act(a + b) or a*b etc.  op: 0 add, 1 mul, 2 a + b*s */
int ab_binary(const void* a, const void* b, void* y, long long n, int op, float s, int act, int dtype,
              void* stream);
/* sinusoidal timestep embedding: out[b, dim] fp16/bf16, flip_sin_to_cos & freq_shift as in SD */
int ab_timestep_embedding(const float* t, void* out, int batch, int dim, int flip_sin_to_cos, float freq_shift,
                          float max_period, int dtype, void* stream);
/* RVM: green-screen / alpha / foreground compositing, fgr [N,H,W,3], pha [N,H,W,1] -> u8 [N,H,W,3] */
int ab_matting_composite(const void* fgr, const void* pha, unsigned char* out, long long npix, int mode,
                         int dtype, void* stream);

/* ---------------------------------------------------------------- matting (csrc/matting.cu) ---
 * Small-channel direct convolution on the CUDA cores for the shapes the tensor-core path cannot tile
 * (RVM decoder: 171/107/59/35 input channels, 3-channel image sources). NHWC, up to three virtually
 * concatenated sources, each a (pointer, channels, row stride) triple so channel slices of a wider tensor
 * are sources without a copy. Optional per-source multiplicative gate (ConvGRU r*h), optional input affine
 * on source 0 (ImageNet normalisation; applied to in-bounds pixels only, padding stays zero), optional
 * ConvGRU epilogue out = (1-z)*h + z*act(acc).
 * w: [ks*ks*(c0+c1+c2), cout_pad] with cout_pad = round_up(cout, 16), model dtype. bias: [cout] model dtype. */
typedef struct ab_conv_generic_args {
    const void* src[3];
    const void* gate[3];
    int c[3], ld[3], gate_ld[3];
    int nsrc;
    const void* w;
    const void* bias;
    const float* in_scale;
    const float* in_shift;
    void* out;
    int out_ld;
    const void* gru_z;
    const void* gru_h;
    int gru_z_ld, gru_h_ld;
    int n, h, w_in, ks, stride, dilation, cout, act, dtype;
} ab_conv_generic_args;
int ab_conv_generic(const ab_conv_generic_args* args, void* stream);
/* bilinear resize with explicit source-step (PyTorch align_corners=False with a given scale_factor):
 * src index = (dst + 0.5) * step - 0.5, clamped. in_u8 = 1: x is uint8 and the output is x/255. */
int ab_resize_bilinear_ex(const void* x, int ldx, void* y, int ldy, int n, int h, int w, int c, int oh, int ow,
                          float step_h, float step_w, int in_u8, int dtype, void* stream);
/* 2x2 average pool, ceil_mode, count_include_pad=False, any h/w/c */
int ab_avgpool2x_ceil(const void* x, void* y, int n, int h, int w, int c, int dtype, void* stream);
/* rows x cols strided copy to one or two destinations (dst2 may be NULL) */
int ab_copy_cols(const void* src, int ld_src, void* dst1, int ld1, void* dst2, int ld2, long long rows, int cols,
                 int dtype, void* stream);
/* deep guided filter, low-res half: base_src [n,h,w,3], proj [n,h,w,4] -> stats [n,h,w,8] (cov_xy, var_x),
 * mean_x [n,h,w,4], mean_y [n,h,w,4]; 3x3 box filter with zero padding (divide by 9 everywhere) */
int ab_guided_stats(const void* base_src, const void* proj, void* stats, void* mean_x, void* mean_y, int n, int h,
                    int w, int dtype, void* stream);
/* deep guided filter, full-res half + output stage: src u8 [N,H,W,3]; refine=1: A/mean_x/mean_y [n,h,w,4] are
 * bilinearly upsampled (b = mean_y - A*mean_x at low res), out = A*[src,mean(src)] + b; refine=0: proj [N,H,W,4]
 * at full res. fgr = clamp(out[:3]+src), pha = clamp(out[3]). Writes fgr [N,H,W,3] / pha [N,H,W,1] (model dtype)
 * when non-NULL and the composited u8 frame (mode as ab_matting_composite) when out_u8 is non-NULL. */
int ab_guided_apply(const unsigned char* src, const void* a_or_proj, const void* mean_x, const void* mean_y,
                    void* fgr, void* pha, unsigned char* out_u8, int n, int H, int W, int h, int w, int refine,
                    int mode, int dtype, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* ARBIUS_B200_H */
