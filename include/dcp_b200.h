// Public C ABI of the native libraries (libdcp_host.so: host plumbing, g++; libdcp_cuda.so: sm_100a kernels, nvcc).
// Everything is plain extern "C": Python binds it with ctypes (_native.py, _cuda_sigs.py, ops/cuda.py, ops/tc.py) and
// C/C++ hosts can link it directly.  Device pointers are raw addresses; `stream` arguments are cudaStream_t.
// All int-returning entry points return 0 on success and -1 on failure (message via dcp_last_error()).
// GENERATED from the sources by tools/gen_header.py - the comments are the ones next to each definition.
#pragma once
#include <stddef.h>
#include <stdint.h>

#ifdef __CUDACC__
#include <cuda_runtime.h>
#else
typedef struct CUstream_st* cudaStream_t;
#endif

#ifdef __cplusplus
extern "C" {
#endif

struct DcpDevCtx;                   /* csrc/common.cuh */
typedef struct ConvP { int N, H, W, C, K, R, S, stride, pad, Ho, Wo; } ConvP;

/* ---- csrc/coll/coll.cu ---- */
size_t dcp_coll_oneshot_max_bytes(void* h);
size_t dcp_coll_bounce_max_bytes(void* h);
// algo: 1 one-shot, 2 two-shot P2P, 3 NVLS.  In-place on `buf` (n elements).  If `buf` lies inside the
// symmetric heap the peers read it directly; otherwise it is bounced through the channel's staging area, in as many
// chunks as its size needs.
int dcp_all_reduce(void* h, void* buf, size_t n, int dtype, int op, int algo, int chan, int max_ctas, cudaStream_t stream);
int dcp_broadcast(void* h, void* buf, size_t nbytes, int root, int chan, int max_ctas, cudaStream_t stream);
// out: [W][nbytes] contiguous.  root < 0: all_gather; root >= 0: gather to root (out may be null elsewhere).
// A shard larger than the bounce capacity goes through in chunks (shard p's chunk lands at out + p * nbytes + done).
int dcp_all_gather(void* h, const void* in, void* out, size_t nbytes, int root, int chan, int max_ctas, cudaStream_t stream);
// in: [W][nbytes] contiguous on root.  Chunked like dcp_all_gather when a shard exceeds the bounce capacity.
int dcp_scatter(void* h, const void* in, void* out, size_t nbytes, int root, int chan, int max_ctas, cudaStream_t stream);
// root < 0: reduce_scatter (in: W*n_out elements, out: n_out).  root >= 0: reduce (in: n_out elements, out on root).
// A non-symmetric input larger than the bounce capacity is reduced in chunks of the output range.
int dcp_reduce_scatter(void* h, const void* in, void* out, size_t n_out, int dtype, int op, int root, int chan, int max_ctas, cudaStream_t stream);
int dcp_send(void* h, const void* in, size_t nbytes, int dst, cudaStream_t stream);
int dcp_recv(void* h, void* out, size_t nbytes, int src, cudaStream_t stream);
int dcp_barrier_launch(void* h, int chan, cudaStream_t stream);

/* ---- csrc/coll/fused_opt.cu ---- */
// Elements of the shard rank `rank` owns for a bucket of n elements of `grad_dtype` (state tensors need this many floats).
size_t dcp_fused_shard_elems(size_t n, int world, int grad_dtype);
// grads / params: device pointers INSIDE the symmetric heap (same offsets on every rank); the parameter buffer must
// be padded to a multiple of 8 floats.  s1 / s2: this rank's state shards (dcp_fused_shard_elems floats, or null).
// kind: 0 Adadelta, 1 SGD, 2 Adam.  hp: device hyper-parameter block (see ops/optim.cu; call dcp_optim_advance first).
int dcp_fused_rs_opt_ag(void* h, const void* grads, float* params, float* s1, float* s2, size_t n, const float* hp, int kind, int grad_dtype, int use_nvls, int chan, int max_ctas, cudaStream_t stream);

/* ---- csrc/ddp/reducer.cpp ---- */
// Plan buckets.  `nbytes[i]` / `group_key[i]` describe tensor i in the order given (dtype+device folded
// into group_key by the caller).  `limits` is the per-group sequence of size limits; the last entry
// repeats.  A bucket is closed once its size reaches the current limit.  Buckets are numbered by
// the position of their first tensor.  Returns the number of buckets; out_bucket_of[i] = bucket of i.
int dcp_bucket_assign(const int64_t* nbytes, const int32_t* group_key, int n, const int64_t* limits, int nlimits, int32_t* out_bucket_of);
void* dcp_reducer_create(const int32_t* bucket_of, int n, int nbuckets);
void dcp_reducer_destroy(void* h);
// Start a new backward pass.
void dcp_reducer_reset(void* h);
// Gradient of parameter `idx` has landed.  Returns how many buckets became launchable (in order)
// and writes the first of them to *first; -1 if idx was already marked this iteration.
int dcp_reducer_mark_ready(void* h, int idx, int* first);
// Number of buckets not yet launched this iteration.
int dcp_reducer_unlaunched(void* h);
int dcp_reducer_ready_order(void* h, int32_t* out, int cap);

/* ---- csrc/host/mesh.cpp ---- */
// Every rank listens on an ephemeral port and publishes "<prefix>/addr/<rank>" = "host:port".
// Rank i dials every j < i and accepts from every j > i; the dialler announces its rank.
void* dcp_mesh_create(void* store, int rank, int world, const char* my_host, const char* prefix, double timeout_s);
void dcp_mesh_destroy(void* h);
int dcp_mesh_send(void* h, int dst, const void* p, int64_t n, int tag);
int dcp_mesh_recv(void* h, int src, void* p, int64_t n, int tag);
// Full-duplex exchange: progress the send to `dst` and the receive from `src` together so that a
// ring of ranks that all send first cannot deadlock on full socket buffers.
int dcp_mesh_sendrecv(void* h, int dst, const void* sp, int64_t sn, int src, void* rp, int64_t rn, int tag);
int dcp_fd_listen(const char* name);
// Connect to `name` and send `nfds` descriptors tagged with the sender's rank.
int dcp_fd_send(const char* name, int my_rank, const int* fds, int nfds, double timeout_s);
// Accept one connection on `lfd`, receive its descriptors.  Returns the sender's rank (>=0) or <0.
int dcp_fd_recv(int lfd, int* out_fds, int max_fds, int* out_nfds, double timeout_s);
void dcp_fd_close(int fd);

/* ---- csrc/host/store.cpp ---- */
void* dcp_store_server_start(const char* host, int port, int* out_port);
void dcp_store_server_stop(void* h);
void* dcp_store_client_connect(const char* host, int port, double timeout_s);
void dcp_store_client_close(void* h);
int64_t dcp_store_set(void* h, const char* key, const void* val, int64_t len);
// Blocks until `key` exists or the timeout expires (-3).  The value stays inside the client until the
// next call; fetch it with dcp_store_copy_last.
int64_t dcp_store_get(void* h, const char* key, double timeout_s);
int64_t dcp_store_copy_last(void* h, void* buf, int64_t cap);
int64_t dcp_store_add(void* h, const char* key, int64_t delta);
int64_t dcp_store_check(void* h, const char* key);
int64_t dcp_store_delete(void* h, const char* key);
int64_t dcp_store_num_keys(void* h);

/* ---- csrc/ops/attention_tc.cu ---- */
// Debugging aid: device buffer (>= 256 long long) that one CTA fills with clock64 stamps, eight per key / query block (see
// ATTN_STAMP / ATTN_FSTAMP): CTA bx of the backward kernel, or - which = 1 - CTA (bx, 0) of the forward kernel.  nullptr = off.
void dcp_attn_set_trace(void* buf, int bx, int which);
// Causal self-attention forward for head dimension 64.  qkv: bf16 [B*T, 3*H*64] (the fused projection: q | k | v, heads inside
// each third); out: bf16 [B*T, H*64]; lse: fp32 [B*H, T], log2-domain log-sum-exp of the scaled scores (consumed by
// dcp_attn_bwd).  T must be a multiple of 128.  softmax scale = 1/sqrt(64).
int dcp_attn_fwd(const void* qkv, void* out, float* lse, int B, int T, int H, cudaStream_t s);
// Backward of dcp_attn_fwd.  dout / out: bf16 [B*T, H*64]; lse from the forward; dqkv: bf16 [B*T, 3*H*64] receives dq | dk | dv
// in the layout of qkv (what the projection's backward GEMMs consume - no split / merge copies).  Scratch from the caller:
// delta (B*H*T floats) and dq_acc (B*T*H*64 floats; zeroed here by the delta kernel).
int dcp_attn_bwd(const void* qkv, const void* out, const void* dout, const float* lse, void* dqkv, float* delta, float* dq_acc, int B, int T, int H, cudaStream_t s);

/* ---- csrc/ops/batchnorm.cu ---- */
// scratch: >= 4*C doubles.  Phase 1 of forward: zero + accumulate (sum, sumsq).
int dcp_bn_stats(const void* x, void* scratch, long M, int C, int dtype, cudaStream_t s);
// Phase 2 of forward.  training=0: uses running stats, scratch may be null.
int dcp_bn_apply(const void* x, void* y, const void* scratch, const float* gamma, const float* beta, float* run_mean, float* run_var, void* nbt, float* save_mean, float* save_rstd, long M, int C, double count, float eps, float momentum, int relu, int training, int dtype, cudaStream_t s);
// Backward phase 1: zero + accumulate (sum dy', sum dy'*xhat) into scratch[0..2C) and keep a local copy in [2C..4C).
int dcp_bn_bwd_reduce(const void* x, const void* y, const void* dy, const float* mean, const float* rstd, void* scratch, long M, int C, int relu, int dtype, cudaStream_t s);
// Backward phase 2.  count: rows behind scratch[0..2C) (global for SyncBN).
int dcp_bn_bwd_apply(const void* x, const void* y, const void* dy, void* dx, const float* mean, const float* rstd, const float* gamma, const void* scratch, float* dgamma, float* dbeta, long M, int C, double count, int relu, int dtype, cudaStream_t s);

/* ---- csrc/ops/conv_flat_tc.cu ---- */
int dcp_conv_flat_tc(const float* in, const float* wmat, float* out, const float* bias, const float* mask, long P, int Cin, int Cout, int taps, const int* off, int W, int HW, int Hv, int Wv, int mode, int num_ctas, cudaStream_t s);
// Convolution + bias + ReLU + 2x2 max-pool + channel dropout in ONE kernel (K3-K6 of the reference model, ref:main.py:34-37): the
// conv output never reaches memory.  pooled: [N][Cout][Hv/2][Wv/2] fp32 (the NCHW-flattened order fc1 consumes, ref:main.py:38-39),
// arg: one byte per pooled element = position of the maximum in its window (for the backward pass).  Dropout drops whole channels
// of an image with probability drop_p (scaled by 1/(1-p)), Philox stream `stream` + step counter, regenerated - not stored - in
// backward.  Needs Hv % 4 == 0, Wv even, 4 * W <= 128 flat pixels, Cout == 64.
int dcp_conv_flat_pool_tc(const float* in, const float* wmat, float* pooled, void* arg, const float* bias, long P, int Cin, int Cout, int taps, const int* off, int W, int HW, int Hv, int Wv, float drop_p, unsigned long long seed, unsigned long long stream, const void* step, int training, int num_ctas, cudaStream_t s);

/* ---- csrc/ops/conv_linear_simt.cu ---- */
int dcp_conv2d_fwd_simt(const void* x, const void* w, const void* bias, void* y, ConvP p, int relu, int dtype, cudaStream_t s);
int dcp_conv2d_dgrad_simt(const void* dy, const void* w, void* dx, ConvP p, int dtype, cudaStream_t s);
// dw32: float [K, R*S*C] accumulator (zeroed here).  db32: float [K] or null.
int dcp_conv2d_wgrad_simt(const void* dy, const void* x, float* dw32, float* db32, ConvP p, int dtype, cudaStream_t s);
// y[M,N] = x[M,K] . w[N,K]^T + b
int dcp_linear_fwd_simt(const void* x, const void* w, const void* bias, void* y, int M, int N, int K, int relu, int dtype, cudaStream_t s);
// dx[M,K] = dy[M,N] . w[N,K]
int dcp_linear_dgrad_simt(const void* dy, const void* w, void* dx, int M, int N, int K, int dtype, cudaStream_t s);
// dw32[N,K] = dy[M,N]^T . x[M,K] (float accumulator, zeroed here); db32[N] = column sums of dy
int dcp_linear_wgrad_simt(const void* dy, const void* x, float* dw32, float* db32, int M, int N, int K, int dtype, cudaStream_t s);
// float column sums of [M, C] (+= into zeroed out)
int dcp_colsum(const void* x, float* out, long M, int C, int dtype, cudaStream_t s);
int dcp_cast_from_f32(const float* in, void* out, long n, int dtype, cudaStream_t s);

/* ---- csrc/ops/conv_wgrad_flat_tc.cu ---- */
// x: [P, 32] fp32 flat pixels (pitch W); dy: [P, Cout] fp32 in the same frame; scratch: [num_ctas][Cout*288 + Cout] fp32 partial
// filters followed by partial bias gradients (finish with dcp_splitk_reduce2(scratch, dw, db, Cout*288, num_ctas, Cout*289)).
// 3x3 filters, Cin = 32.
// Exactly num_ctas partials are written (CTAs without pixels write zeros).
int dcp_conv_wgrad_flat_tc(const float* x, const float* dy, float* scratch, long P, int Cout, int W, int num_ctas, cudaStream_t s);

/* ---- csrc/ops/convnet_dp.cu ---- */
int dcp_dp_nvls_mode(int set);
// Payload path of the fused data-parallel tail: 1 = multimem.st through NVSwitch, 0 = one store per peer (default).
int dcp_dp_set_nvls(int on);
// Finish a split-K reduction (scratch: [nsplit][n] floats, n % 4 == 0) and push the n sums into floats
// [elem_off, elem_off + n) of this rank's slot in every rank's payload block; the last CTA signals (chan, blk).
// Nothing is written locally: dcp_dp_tail produces the averaged gradients.  ticket: one zeroed u32 in device memory
// owned by the caller.  step: device int64 counter that holds the same value on every rank when the kernel runs.
int dcp_splitk_reduce_push(void* h, const float* scratch, int nsplit, long n, void* payload, unsigned slot_elems, unsigned elem_off, int chan, int blk, const void* step, void* ticket, int pull, cudaStream_t s);
// Consumer of the pushed gradients: for the `ntensors` rows of the optimizer table (param, grad, state1, state2, numel; every
// grad a view into the bucket that starts at `bucket`, in bucket order = slot order) wait for the `nsig` producers
// (signal words blks[i] on `chan`) of every rank, average the W slots, store the averaged gradient and apply the update.
// row_sig[i] (i < ntensors <= 16) names the producer (index into blks) whose signal guards row i's gradients.
// kind: 0 Adadelta, 1 SGD, 2 Adam; hp: the optimizer's device hyper-parameter block (dcp_optim_advance'd by the caller).
int dcp_dp_tail(void* h, const void* table, int ntensors, long long max_numel, const float* hp, int kind, const void* bucket, void* payload, unsigned slot_elems, int chan, const int* blks, int nsig, const int* row_sig, const void* step, int pull, cudaStream_t s);

/* ---- csrc/ops/convnet_fused.cu ---- */
int dcp_convnet_conv1_fwd(const float* x, const float* w, const float* b, float* a1, int N, cudaStream_t s);
int dcp_convnet_pool_fwd(const float* a2, float* pooled, void* arg, int N, float p, unsigned long long seed, unsigned long long stream, const void* step, int training, cudaStream_t s);
int dcp_convnet_pool_bwd(const float* dpooled, const float* pooled, const void* arg, float* dy2, int N, float p, unsigned long long seed, unsigned long long stream, const void* step, int training, cudaStream_t s);
// dw1 / db1 are overwritten.  The partial-sum table is one per device: two of these launches running at the same time
// (different streams) would race on it.
int dcp_convnet_conv1_wgrad(const float* dz1, const float* x, float* dw1, float* db1, int N, cudaStream_t s);
int dcp_convnet_conv1_wgrad_allreduce(void* h, const float* dz1, const float* x, float* dw1, float* db1, int N, void* payload, int chan, int blk, cudaStream_t s);
// Push-only variant for the fused data-parallel tail (ops/dp_push.cuh, convnet_dp.cu): the 320 sums go into this rank's slot
// (floats [elem_off, elem_off + 320) in bucket order: dw1 then db1) of every rank's payload block, followed by one signal per
// rank on (chan, blk); dw1 / db1 are NOT written - dcp_dp_tail produces the averaged gradients.  ws: caller-owned scratch of
// 4 + 148 * 320 floats (ticket word + per-CTA partial sums), zero before the first call; one per concurrent user.
int dcp_convnet_conv1_wgrad_push(void* h, const float* dz1, const float* x, int N, void* payload, unsigned slot_elems, unsigned elem_off, int chan, int blk, const void* step, void* ws, cudaStream_t s);
// conv1's weight-gradient, its data-parallel average and its optimizer update in ONE kernel (struct C1LL above).  ll_block: a
// region of the symmetric heap holding 2 * world * 320 * 8 bytes, at the same offset on every rank, zeroed before first use and
// used by nothing else; rows: two rows of the optimizer's device pointer table - conv1.weight, conv1.bias - whose grad pointers
// receive the averaged gradient; hp / kind as for dcp_optim_step (dcp_optim_advance'd by the caller); step: device int64
// counter holding the same (non-zero, increasing) value on every rank when the kernel runs; ws: 4 + 148 * 320 floats of
// caller-owned scratch, zero before the first call.
int dcp_convnet_conv1_wgrad_ll(void* h, const float* dz1, const float* x, int N, void* ll_block, const void* rows, const float* hp, int kind, const void* step, void* ws, cudaStream_t s);
// Fused head (see above).  N <= 128.  logits_out (log-probabilities [N,10]) may be null; dbias1 ([128], may be null)
// receives the column sums of dh1, i.e. the bias gradient of the Linear layer in front.  All gradient outputs are
// fully overwritten.  Training mode only (batch statistics); p2 = dropout probability of the second dropout.
int dcp_convnet_head(const float* h1, const void* target, const float* gamma, const float* beta, float* run_mean, float* run_var, void* nbt, const float* w2, const float* b2, float* loss, float* logits_out, float* dh1, float* dgamma, float* dbeta, float* dw2, float* db2, float* dbias1, int N, float eps, float momentum, float p2, unsigned long long seed, unsigned long long stream_id, const void* step, int training, cudaStream_t s);
// The same kernel inside a data-parallel job: `bn_block` is a 1032-byte block of symmetric heap `h` laid out
// [running_mean 128 f32][running_var 128 f32][num_batches_tracked i64], allocated at the same offset on every rank.  Rank 0
// computes the new running statistics from its batch and writes them into EVERY rank's block (peer stores over NVLink from
// inside the kernel); the other ranks do not touch theirs.  That is what DistributedDataParallel's per-forward buffer
// broadcast from rank 0 amounts to - without a collective launch, a side stream or a join in front of the head.
// A rank may read the block after any later cross-rank synchronisation of the step (the gradient exchange is one).
int dcp_convnet_head_dp(void* h, const float* h1, const void* target, const float* gamma, const float* beta, void* bn_block, const float* w2, const float* b2, float* loss, float* logits_out, float* dh1, float* dgamma, float* dbeta, float* dw2, float* db2, float* dbias1, int N, float eps, float momentum, float p2, unsigned long long seed, unsigned long long stream_id, const void* step, int training, cudaStream_t s);

/* ---- csrc/ops/elementwise.cu ---- */
// dy: channels-last [N, Ho, Wo, K]; up: [N, Hup, Wup, K] (K * sizeof(element) must be a multiple of 16)
int dcp_upsample2_zero(const void* dy, void* up, int N, int Ho, int Wo, int Hup, int Wup, int K, int dtype, cudaStream_t s);
int dcp_im2col_small_c(const void* x, void* out, int N, int H, int W, int C, int R, int S, int stride, int pad, int Ho, int Wo, int Kpad, int dtype, cudaStream_t s);
int dcp_relu_fwd(const void* x, void* y, long n, int dtype, cudaStream_t s);
int dcp_relu_bwd(const void* y, const void* dy, void* dx, long n, int dtype, cudaStream_t s);
int dcp_add_relu_fwd(const void* a, const void* b, void* y, long n, int dtype, cudaStream_t s);
int dcp_dropout(const void* x, void* y, long n, float p, unsigned long long seed, unsigned long long stream_id, const void* step, int mode, long inner, long hw, int dtype, cudaStream_t s);
int dcp_maxpool_fwd(const void* x, void* y, void* arg, int N, int H, int W, int C, int Ho, int Wo, int k, int st, int pad, int out_nchw, int dtype, cudaStream_t s);
int dcp_maxpool_bwd(const void* dy, const void* arg, void* dx, int N, int H, int W, int C, int Ho, int Wo, int k, int st, int pad, int out_nchw, int dtype, cudaStream_t s);
int dcp_gap_fwd(const void* x, void* y, int N, int HW, int C, int dtype, cudaStream_t s);
int dcp_gap_bwd(const void* dy, void* dx, int N, int HW, int C, int dtype, cudaStream_t s);
int dcp_correct_count(const void* x, const void* target, void* count, int M, int C, int dtype, cudaStream_t s);

/* ---- csrc/ops/gemm_bf16_persistent.cu ---- */
// Hand the GEMM a scratch buffer (device memory, 16-byte aligned) for the tail split of two-wave problems; nullptr / 0 disables it.
// GEMMs that use it must not overlap on different streams.
int dcp_gemm_set_workspace(void* ws, size_t bytes);
// bf16 operands; see gemm_tc.cu::dcp_tc_gemm_ex for the meaning of the major flags.  ksplit > 1: `out` is a
// [ksplit][M * ldc] fp32 scratch that receives partial tiles (finish with dcp_splitk_reduce, which also adds the bias).
// K must be a multiple of 64; MN-major operands need their row count (M resp. N) to be a multiple of 8 (TMA stride rule).
int dcp_gemm_bf16_persistent(const void* A, const void* B, void* out, const void* bias, int M, int N, int K, long lda, long ldb, long ldc, int a_mn, int b_mn, int out_bf16, int bias_bf16, int relu, int ksplit, int num_ctas, cudaStream_t s);
// Implicit-GEMM convolution on the persistent kernel (bf16, stride 1 or 2): y[n,ho,wo,k] = sum x[n, ho*stride-pad+r, wo*stride-pad+s, c] * w[k][(r*S+s)*C + c]
//   x: channels-last [N,H,W,C] (C a multiple of 64); wmat: [Kout][R*S*C] row-major; y: channels-last [N,Ho,Wo,Kout] bf16 or fp32.
//   (Wb, Hb, Nb): box of output pixels per M tile, product 128.  The data-gradient is the same call on dy with flipped filters.
int dcp_conv_bf16_persistent(const void* x, const void* wmat, void* y, const void* bias, int N, int H, int W, int C, int Kout, int R, int S, int pad, int stride, int Ho, int Wo, int Wb, int Hb, int Nb, int out_bf16, int bias_bf16, int relu, cudaStream_t s);

/* ---- csrc/ops/gemm_tc.cu ---- */
// Host-side launch choice for wide, shallow fp32 / tf32 problems: 64-column tiles (twice the CTAs) when 1, 128-column when 0,
// the DCP_GEMM_NARROW default when -1.  Affects the launches issued after the call from this process.
void dcp_gemm_set_narrow(int v);
int dcp_tc_gemm(const void* A, const void* B, void* out, const void* bias, int M, int N, int K, long lda, long ldb, long ldc, int a_mn, int b_mn, int dtype, int out_bf16, int bias_bf16, int relu, int ksplit, cudaStream_t s);
int dcp_tc_gemm_ex(const void* A, const void* B, void* out, const void* bias, int M, int N, int K, long lda, long ldb, long ldc, int a_mn, int b_mn, int dtype, int out_bf16, int bias_bf16, int relu, int ksplit, int x3, cudaStream_t s);
int dcp_tc_conv(const void* x, const void* wmat, void* y, const void* bias, int N, int H, int W, int C, int Kout, int R, int S, int pad, int Ho, int Wo, int Wb, int Hb, int Nb, int dtype, int out_bf16, int bias_bf16, int relu, cudaStream_t s);
// Same with a convolution stride of 1 or 2: y[n,ho,wo,k] = sum x[n, ho*stride-pad+r, wo*stride-pad+s, c] * w[k][(r*S+s)*C + c].
// The strided input window is one TMA box with elementStrides = stride along W and H.
int dcp_tc_conv_strided(const void* x, const void* wmat, void* y, const void* bias, int N, int H, int W, int C, int Kout, int R, int S, int pad, int stride, int Ho, int Wo, int Wb, int Hb, int Nb, int dtype, int out_bf16, int bias_bf16, int relu, cudaStream_t s);
int dcp_tc_conv_wgrad(const void* dy, const void* x, float* dw32, int N, int H, int W, int C, int Kout, int R, int S, int pad, int Ho, int Wo, int Wb, int Hb, int Nb, int dtype, int zsplit, cudaStream_t s);
// Same for a convolution stride of 1 or 2 (x is read at ho*stride-pad+r, wo*stride-pad+s).
int dcp_tc_conv_wgrad_strided(const void* dy, const void* x, float* dw32, int N, int H, int W, int C, int Kout, int R, int S, int pad, int stride, int Ho, int Wo, int Wb, int Hb, int Nb, int dtype, int zsplit, cudaStream_t s);
// n and ncols must be multiples of 4.
int dcp_splitk_reduce(const float* scratch, float* out, const float* bias, int nsplit, long n, int ncols, cudaStream_t s);
// Same reduction with the summed vector going to two places: elements [0, n1) to `out`, [n1, n) to `out2` (n1 % 4 == 0).
int dcp_splitk_reduce2(const float* scratch, float* out, float* out2, long n1, int nsplit, long n, cudaStream_t s);
int dcp_flip_weights(const void* w, void* wt, int K, int R, int S, int C, int dtype, cudaStream_t s);

/* ---- csrc/ops/loss.cu ---- */
int dcp_log_softmax_fwd(const void* x, void* y, int M, int C, int dtype, cudaStream_t s);
int dcp_log_softmax_bwd(const void* y, const void* dy, void* dx, int M, int C, int dtype, cudaStream_t s);
int dcp_nll_fwd(const void* logp, const void* target, float* loss, int M, int C, float scale, int dtype, cudaStream_t s);
int dcp_nll_bwd(const void* target, const float* gout, void* dlogp, int M, int C, float scale, int dtype, cudaStream_t s);
int dcp_ce_fwd(const void* x, const void* target, float* loss, float* lse, int M, int C, float scale, int dtype, cudaStream_t s);
int dcp_ce_bwd(const void* x, const void* target, const float* lse, const float* gout, void* dx, int M, int C, float scale, int dtype, cudaStream_t s);

/* ---- csrc/ops/optim.cu ---- */
// Advance the device step counter (call once per optimizer.step(), before dcp_optim_step).
int dcp_optim_advance(float* hp, cudaStream_t s);
// kind: 0 Adadelta, 1 SGD, 2 Adam.  grad_dtype: DCP_F32 or DCP_BF16.  max_numel sizes the grid.
int dcp_optim_step(const void* table, int ntensors, long long max_numel, const float* hp, int kind, int grad_dtype, cudaStream_t s);
int dcp_counter_advance(void* counter, cudaStream_t s);

/* ---- csrc/ops/transformer.cu ---- */
int dcp_layernorm_fwd(const void* x, void* y, const float* gamma, const float* beta, float* mean, float* rstd, int rows, int C, float eps, int dtype, cudaStream_t s);
// dgamma / dbeta: fp32 [C], zeroed here.
int dcp_layernorm_bwd(const void* x, const void* dy, const float* gamma, const float* mean, const float* rstd, void* dx, float* dgamma, float* dbeta, int rows, int C, int dtype, cudaStream_t s);
// y = LayerNorm(x + res), sum_out = x + res (the residual stream after the add): one pass instead of an add kernel plus a
// LayerNorm.  Row length as in the warp-per-row path of dcp_layernorm_fwd (C a multiple of the 16-byte vector, <= 1024 bf16 /
// 512 fp32 elements), 16-byte aligned pointers; returns -1 otherwise (the caller falls back to add + layer norm).
int dcp_add_layernorm_fwd(const void* x, const void* res, void* sum_out, void* y, const float* gamma, const float* beta, float* mean, float* rstd, int rows, int C, float eps, int dtype, cudaStream_t s);
// Backward of dcp_add_layernorm_fwd: dx = LayerNorm-backward(dy) + dsum, where dsum is the gradient that reached the sum from
// the residual path (the same dx is the gradient of both x and res).  x = the saved sum.  dgamma / dbeta: fp32 [C], zeroed here.
// dxsum: optional fp32 [C] that receives the column sums of dx (zeroed here) - the bias gradient of the Linear that produced res.
int dcp_add_layernorm_bwd(const void* x, const void* dy, const void* dsum, const float* gamma, const float* mean, const float* rstd, void* dx, float* dgamma, float* dbeta, float* dxsum, int rows, int C, int dtype, cudaStream_t s);
// Token + position embedding of a transformer: out[r, :] = T(wte[idx[r], :] + wpe[r % Tlen, :]) for r < rows = B * Tlen.
// fp32 tables, f32 / bf16 output, C a multiple of 4.  (One gather kernel instead of two gathers, an add and a cast.)
int dcp_embed_fwd(const long* idx, const float* wte, const float* wpe, void* out, long rows, int Tlen, int C, int dtype, cudaStream_t s);
// Backward: dwte[idx[r], :] += dy[r, :] (vector reductions; dwte fp32 [V, C] zeroed by the caller - rows that are no token stay
// zero) and dwpe[t, :] = sum_b dy[b * Tlen + t, :] for t < Tlen (plain stores; the caller zeroes the rows beyond Tlen).
int dcp_embed_bwd(const long* idx, const void* dy, float* dwte, float* dwpe, long rows, int Tlen, int C, int dtype, cudaStream_t s);
// GELU backward over a [rows, C] matrix that also produces the column sums of dx (fp32 [C], zeroed here): C a multiple of the
// 16-byte vector, aligned pointers (-1 otherwise; the caller falls back to dcp_gelu_bwd).
int dcp_gelu_bwd_colsum(const void* x, const void* dy, void* dx, float* colsum, long rows, int C, int dtype, cudaStream_t s);
int dcp_gelu_fwd(const void* x, void* y, long n, int dtype, cudaStream_t s);
int dcp_gelu_bwd(const void* x, const void* dy, void* dx, long n, int dtype, cudaStream_t s);

/* ---- csrc/symm/symm.cu ---- */
const char* dcp_last_error();
void dcp_set_error(const char* fmt, ...);
// Whether device `device` supports multicast objects (NVLS).
int dcp_symm_multicast_supported(int device);
// Create this rank's heap.  mode: 0 = vmm (POSIX FD), 1 = ipc (cudaMalloc).  Returns handle or null.
void* dcp_symm_create(int rank, int world, int device, size_t bytes, int mode);
size_t dcp_symm_bytes(void* h);
void* dcp_symm_local_ptr(void* h);
void* dcp_symm_peer_ptr(void* h, int r);
void* dcp_symm_mc_ptr(void* h);
// vmm: export own heap as a POSIX FD (caller sends it to peers and closes it).
int dcp_symm_export_fd(void* h);
// vmm: map peer `r`'s heap from the FD it sent.
int dcp_symm_import_fd(void* h, int r, int fd);
// ipc: 64-byte handle of own heap / open a peer's.
int dcp_symm_ipc_handle(void* h, void* out64);
int dcp_symm_ipc_open(void* h, int r, const void* in64);
// NVLS step 1 (rank 0): create the multicast object, returns an FD to distribute (or -1).
int dcp_symm_mc_create(void* h);
// NVLS step 1' (other ranks): import the multicast object.
int dcp_symm_mc_import(void* h, int fd);
// NVLS step 2 (all ranks): add this device.  Every rank must finish this before anyone binds.
int dcp_symm_mc_add_device(void* h);
// NVLS step 3 (all ranks): bind own heap and map the multicast VA.
int dcp_symm_mc_bind(void* h);
// Freeze the device context after all peers are mapped.
int dcp_symm_finalize(void* h, double timeout_s, size_t p2p_slot_bytes, size_t staging_bytes, int allow_nvls);
size_t dcp_symm_user_offset(void* h);
size_t dcp_symm_staging_offset(void* h);
size_t dcp_symm_staging_bytes(void* h);
int dcp_symm_has_nvls(void* h);
const struct DcpDevCtx* dcp_symm_ctx(void* h);
// Non-zero once any kernel's bounded wait expired (a peer died or the ranks diverged).
int dcp_symm_error_flag(void* h);
void dcp_symm_clear_error(void* h);
void dcp_symm_destroy(void* h);
// Create `world` heaps of heap_bytes each on `device`, every one seeing the others as its peers (plain device pointers: same
// process, same GPU), no multicast.  Returns the group; dcp_emu_rank(group, r) is an ordinary heap handle for every other
// entry point of this library - except that collectives issued on it are deferred until dcp_emu_flush().
void* dcp_emu_create(int world, int device, size_t heap_bytes, double timeout_s, size_t p2p_slot_bytes, size_t staging_bytes);
void* dcp_emu_rank(void* group, int r);
int dcp_emu_pending(void* group);
// Launch everything the ranks have queued: slot i of every rank becomes ONE cooperative kernel of (grid, world) CTAs on
// `stream`.  The ranks must have queued the same collectives in the same order (same kernel, same grid) - as they would have
// to on N GPUs; point-to-point slots may be empty on ranks that take no part.  Returns -1 (and drops the queue) on a mismatch.
int dcp_emu_flush(void* group, cudaStream_t stream);
void dcp_emu_destroy(void* group);

/* ---- hand-maintained additions ---- */

/* Allocation-free BatchNorm forward.  This entry point has no counterpart in the reference.  It exists for hosts of THIS
 * library that cannot provide scratch memory - a C/C++ program linking libdcp_cuda.so without PyTorch's caching allocator,
 * or a process whose own arena already owns the whole GPU so that cudaMalloc would fail (and would stall the device anyway).
 * The per-channel accumulators live in a 32 KiB __device__ array inside the library (one copy per GPU, kept until
 * exit), so: C <= DCP_BN_NOALLOC_MAX_C (2048) or the call is refused; the array is shared by every call and
 * cleared at the start of each, hence calls that may overlap on the GPU race - keep them on one stream (the library
 * serialises the host side of concurrent callers, not the device side); the first call on a device resolves the array's
 * address (cudaGetSymbolAddress) and should come before any stream capture.  That first call is also the moment the CUDA
 * runtime loads the library's device image - array included - onto that GPU: a process that hands the whole GPU to its own
 * arena must make one call per device BEFORE it does so (or run with CUDA_MODULE_LOADING=EAGER).
 * Same results as dcp_bn_stats + dcp_bn_apply(training=1). */
#define DCP_BN_NOALLOC_MAX_C 2048
int dcp_bn_forward_noalloc(const void* x, void* y, const float* gamma, const float* beta, float* run_mean, float* run_var, void* nbt,
                           float* save_mean, float* save_rstd, long M, int C, float eps, float momentum, int relu, int dtype, cudaStream_t s);

#ifdef __cplusplus
}  /* extern "C" */
#endif
