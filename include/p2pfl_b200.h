/*
 * p2pfl_b200 - C ABI of the sm_100a kernel library (libp2pfl_b200.so).
 *
 * Every entry point returns 0 on success or a non-zero code; the text of the
 * last failure on the calling thread is available from p2pfl_last_error().
 * `stream` is a cudaStream_t passed as void*.  `device` is the CUDA ordinal the
 * pointers live on; entry points make it current before launching.
 * All kernels are asynchronous on `stream` unless stated otherwise.
 *
 * Reference citations name the site in Angel3245/p2pfl whose work the entry
 * point replaces (see SURVEY.md section 2.5 for the K / C numbering).
 */
#ifndef P2PFL_B200_H
#define P2PFL_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

const char* p2pfl_last_error(void);
int p2pfl_version(void);
/* Programmatic dependent launch for the train-step kernels (GEMM, loss, Adam): on by
 * default; returns the previous setting.  Off = plain stream-ordered launches. */
int p2pfl_set_pdl(int enabled);

/* ------------------------------------------------------------------ runtime
 * Device arenas that can be mapped into peer GPUs / peer processes.
 * (No counterpart in the reference: it keeps models in host state_dicts.) */
int p2pfl_rt_alloc(int device, uint64_t bytes, void** out_ptr);
int p2pfl_rt_free(int device, void* ptr);
int p2pfl_rt_ipc_export(int device, void* ptr, uint8_t out_handle[64]);
int p2pfl_rt_ipc_open(int device, const uint8_t handle[64], void** out_ptr);
int p2pfl_rt_ipc_close(int device, void* ptr);
int p2pfl_rt_enable_peer(int device, int peer_device);
int p2pfl_rt_can_access_peer(int device, int peer_device, int* out_yes);
int p2pfl_rt_memset32(int device, void* stream, void* ptr, uint32_t value, uint64_t count);
/* Puts the CALLING thread in CUDA's thread-local stream-capture mode: calls it makes (allocations,
 * device synchronisation) are no longer refused - nor do they invalidate the capture - while some
 * other thread of the process captures a graph.  Call once per thread that uses CUDA. */
int p2pfl_rt_thread_capture_mode_local(void);
/* a non-blocking stream owned by the caller (one per learner: never shared through a pool) */
int p2pfl_rt_stream_create(int device, void** out_stream);
int p2pfl_rt_stream_destroy(int device, void* stream);
/* Exchange watchdog (csrc/watchdog.cu): blocks the calling thread until `done_event` (recorded behind
 * an exchange kernel) has completed - polling in native code, so a Python caller holds no interpreter
 * lock meanwhile.  It returns early with
 *   P2PFL_WATCH_SLICE      after `slice_s` seconds (<= 0: never), so that the caller can run checks of
 *                          its own and call again;
 *   P2PFL_WATCH_CANCELLED  when `*cancel` (host memory, any thread may set it; may be NULL) is non-zero,
 *   P2PFL_WATCH_TIMEOUT    after `timeout_s` seconds (< 0: never);
 * in the last two cases it first writes 1 to `*abort_dev` (DEVICE memory: the abort word the kernel
 * was given; may be NULL) through a stream of its own and waits up to `grace_s` seconds (< 0: forever)
 * for the kernel to leave.  `*waited_s` (may be NULL) receives the seconds spent in this call.
 * Negative / other values: error, see p2pfl_last_error(). */
#define P2PFL_WATCH_DONE 0
#define P2PFL_WATCH_CANCELLED 1
#define P2PFL_WATCH_TIMEOUT 2
#define P2PFL_WATCH_SLICE 3
int p2pfl_rt_watch(int device, void* done_event, uint32_t* abort_dev, const volatile int32_t* cancel, double timeout_s,
                   double slice_s, double grace_s, double* waited_s);
/* peak SM clock in kHz (converts the device-side timeouts, which count SM cycles, from seconds) */
int p2pfl_rt_sm_clock_khz(int device, int* out_khz);
int p2pfl_rt_device_info(int device, int* sm_count, int* cc_major, int* cc_minor, uint64_t* total_mem);

/* One 2-D matrix inside the flat arena and where its bf16 transpose goes. */
typedef struct p2pfl_transpose_desc_s {
  uint64_t src_off; /* elements into src (fp32) */
  uint64_t dst_off; /* elements into dst (bf16) */
  int32_t rows, cols, ld_t, pad_;
} p2pfl_transpose_desc;

/* --------------------------------------------------------------- optimizers
 * K5: torch.optim.Adam.step (mlp.py:71-73, cnn.py:89-91) as ONE kernel over
 * the flat parameter arena; optionally also refreshes the bf16 working copy.
 * t = *step_ptr + step_offset is the 1-based Adam step (bias correction). */
int p2pfl_adam_step(int device, void* stream, float* p, const float* g, float* m, float* v,
                    void* p_bf16 /* nullable */, uint64_t n, float lr, float beta1, float beta2,
                    float eps, float weight_decay, const int32_t* step_ptr, int32_t step_offset);
/* Same, and additionally refreshes the bf16 TRANSPOSED copies of the listed weight
 * matrices (operands of the dgrad GEMMs) in the same pass - one launch less per step. */
int p2pfl_adam_step_t(int device, void* stream, float* p, const float* g, float* m, float* v,
                      void* p_bf16, uint64_t n, float lr, float beta1, float beta2, float eps,
                      float weight_decay, const int32_t* step_ptr, int32_t step_offset,
                      void* wt_bf16, const struct p2pfl_transpose_desc_s* descs, int count);
int p2pfl_sgd_step(int device, void* stream, float* p, const float* g, float* mom /* nullable */,
                   void* p_bf16 /* nullable */, uint64_t n, float lr, float momentum);
int p2pfl_add_i32(int device, void* stream, int32_t* ptr, int32_t delta);
/* fp32 -> bf16 cast of a flat vector */
int p2pfl_cast_f32_bf16(int device, void* stream, const float* src, void* dst, uint64_t n);
/* Batched 2-D transpose+cast: for each of `count` matrices (rows x cols fp32 at
 * src_off) write the bf16 transpose (cols x rows, leading dim ld_t) at dst_off. */
int p2pfl_transpose_cast(int device, void* stream, const float* src, void* dst,
                         const p2pfl_transpose_desc* descs, int count);

/* ------------------------------------------------------------------- losses
 * K3/K9: log_softmax + cross_entropy forward AND backward in one kernel
 * (mlp.py:68,78; cnn.py:73,96).  logits fp32 [B, ld_logits]; labels int32 [B].
 * Writes dlogits (already divided by B) as bf16 in both layouts the GEMMs
 * consume ([B, ld_d] and [C, ld_dt]), the last layer's bias gradient
 * (overwritten), and adds the summed loss to *loss_sum. */
int p2pfl_softmax_xent(int device, void* stream, const float* logits, int ld_logits,
                       const int32_t* labels, int B, int C, void* dlogits_bf16, int ld_d,
                       void* dlogits_t_bf16, int ld_dt, float* dbias /* nullable */,
                       float* loss_sum /* nullable */);
/* K10: evaluation reduce: adds sum of per-sample loss and number of correct
 * argmax predictions (mlp.py:96-99 torchmetrics Accuracy). */
int p2pfl_eval_reduce(int device, void* stream, const float* logits, int ld_logits,
                      const int32_t* labels, int B, int C, float* loss_sum, int32_t* correct);

/* ------------------------------------------------------------------- FedAvg
 * K12: fedavg.py:49-58 as one fused pass.  out = sum_i w[i] * model_i with the
 * (already normalised) weights; model_i is fp32 (is_bf16[i]==0) or bf16. */
#define P2PFL_MAX_MODELS 16
int p2pfl_fedavg(int device, void* stream, int n_models, const void* const* models,
                 const float* weights, const int32_t* is_bf16, float* out_f32 /* nullable */,
                 void* out_bf16 /* nullable */, uint64_t n);

/* ------------------------------------------------------------- gossip plane
 * C1/C3/C4 + K12: model push over NVLink fused with FedAvg on arrival
 * (gossiper.py:239 client.send + fedavg.py:49-58 + lightning_learner.py:152).
 *
 * An *inbox* lives in the receiver's arena: `slots` model-sized buffers in the
 * wire dtype plus one uint32 arrival flag per (slot, tile).  A sender stores
 * its tiles into slot[sender] of every peer and then publishes `seq` in the
 * tile's flag with release semantics at system scope; the receiver acquires the
 * flag before reading the tile.  `tile` is in elements. */
#define P2PFL_MAX_PEERS 16
typedef struct {
  void* inbox[P2PFL_MAX_PEERS];     /* peer-mapped base of slot[me] in peer j's arena */
  uint32_t* flags[P2PFL_MAX_PEERS]; /* peer-mapped base of flags[me] in peer j's arena */
  int32_t n_peers;
  int32_t wire_bf16; /* 1: bf16 wire, 0: fp32 wire */
} p2pfl_push_table;

typedef struct {
  const void* slot[P2PFL_MAX_PEERS];     /* local base of slot[j] */
  const uint32_t* flags[P2PFL_MAX_PEERS]; /* local base of flags[j] */
  float weight[P2PFL_MAX_PEERS];          /* normalised FedAvg weight of contributor j */
  int32_t n_slots;      /* contributors arriving through the inbox */
  int32_t self_pos;     /* position of the local model in the summation order, -1: not a contributor */
  float self_weight;
  int32_t wire_bf16;
  int32_t round_self;   /* 1: round the local model to the wire dtype first (bit-identical results on all nodes) */
  int32_t pad_;
} p2pfl_reduce_table;

/* status words written by the reduce kernels (device memory, int32[4]):
 * [0] = 1 if aborted/timeout, [1] = number of tiles finalised, [2] = bit j set when the
 * contributor at reduce-table position j was missing from at least one tile, [3] reserved */

/* G1/G3: push `src` (fp32, n elements) into every peer's inbox slot. */
int p2pfl_gossip_push(int device, void* stream, const float* src, uint64_t n, uint32_t tile,
                      uint32_t seq, const p2pfl_push_table* table);
/* G2: wait per tile for every expected slot, weighted-accumulate in fp32,
 * write the live parameters (fp32 and optional bf16 copy).  `abort_flag` is a
 * host-writable word polled while spinning; `timeout_cycles` bounds each wait. */
int p2pfl_gossip_reduce(int device, void* stream, float* params, void* params_bf16 /* nullable */,
                        uint64_t n, uint32_t tile, uint32_t seq, const p2pfl_reduce_table* table,
                        const uint32_t* abort_flag /* nullable */, uint64_t timeout_cycles,
                        int32_t* status /* nullable */);
/* G1+G2 in ONE kernel: push all local tiles, then reduce tiles as they land.
 * Uses device scratch: `status` (int32[4]). */
int p2pfl_gossip_fused(int device, void* stream, float* params, void* params_bf16 /* nullable */,
                       uint64_t n, uint32_t tile, uint32_t seq, const p2pfl_push_table* push,
                       const p2pfl_reduce_table* reduce, const uint32_t* abort_flag,
                       uint64_t timeout_cycles, int32_t* status);
/* Same, and every tile of the pre-exchange model is also copied to `snapshot` (fp32[n],
 * nullable) before the reduce overwrites it: a host that finds status[0] != 0 afterwards
 * can put the model back and exchange some other way instead of keeping a partly averaged one. */
int p2pfl_gossip_fused_snap(int device, void* stream, float* params, void* params_bf16 /* nullable */,
                            uint64_t n, uint32_t tile, uint32_t seq, const p2pfl_push_table* push,
                            const p2pfl_reduce_table* reduce, const uint32_t* abort_flag,
                            uint64_t timeout_cycles, int32_t* status, float* snapshot /* nullable */);
/* The exchange of SEVERAL federated nodes that live on one GPU (BASELINE config 4: simulated
 * nodes packed on a device) as ONE cooperative launch: member m runs the fused push+reduce
 * schedule with push_tables_dev[m] / reduce_tables_dev[m]; a table may name inboxes of the
 * other local members and of nodes on peer GPUs alike.  All three arrays are DEVICE memory.
 * (train_stage.py:115-177 for every co-located trainer at once.) */
typedef struct {
  float* params;              /* fp32[n] live model of the member, overwritten with the average */
  void* params_bf16;          /* nullable bf16 working copy, refreshed */
  float* snapshot;            /* nullable, see p2pfl_gossip_fused_snap */
  const uint32_t* abort_flag; /* nullable */
  int32_t* status;            /* int32[4] of the member */
} p2pfl_group_member;
int p2pfl_gossip_fused_group(int device, void* stream, int n_members,
                             const p2pfl_group_member* members_dev, uint64_t n, uint32_t tile,
                             uint32_t seq, const p2pfl_push_table* push_tables_dev,
                             const p2pfl_reduce_table* reduce_tables_dev, uint64_t timeout_cycles);
/* G1 / G3 with the copy engine (cp.async.bulk): every 16384-element tile of src[begin, end) is
 * fetched into shared memory by one bulk load, converted there when the table's wire is bf16, and
 * stored into every peer of the table by one bulk store each; then the tile's flag takes `seq`.
 * With `inbox` entries pointing at slot[me] this is the push of the fused exchange (mode B); with
 * entries pointing at the peers' PARAMETERS (fp32 wire) it is the model store of init / diffusion
 * (start_learning_stage.py:95-136, gossip_model_stage.py:85-132) - `begin` / `end` select a
 * sub-range of the arena (whole tiles; SURVEY appendix C rule 5).  The arenas' tile must be 16384. */
int p2pfl_gossip_push_bulk(int device, void* stream, const float* src, uint64_t n, uint64_t begin,
                           uint64_t end, uint32_t seq, const p2pfl_push_table* table);
/* G1 mode A: the sender scales its tiles by `weight` and the copy engine ADDS them into each
 * peer's fp32 accumulator (cp.reduce.async.bulk ... add.f32, performed in the peer's L2: no SM of
 * the receiver takes part); `inbox` entries are the accumulators, `flags` per-tile arrival counters.
 * The order of the additions is not fixed, so results are not bit-reproducible (fedavg.py:49-54). */
int p2pfl_gossip_push_reduce_bulk(int device, void* stream, const float* src, uint64_t n, float weight,
                                  const p2pfl_push_table* table);
/* Receiver side of mode A: per tile wait for `n_senders` arrivals, then params = acc + self_weight
 * * params (+ bf16 copy); acc and the counters are left zeroed for the next exchange. */
int p2pfl_gossip_finalize_acc(int device, void* stream, float* params, void* params_bf16 /* nullable */,
                              float* acc, uint32_t* counts, uint64_t n, int n_senders, float self_weight,
                              uint64_t timeout_cycles, int32_t* status /* nullable */);
/* Single-GPU emulation of R ranks exchanging through one cooperative launch
 * (test harness for machines with fewer GPUs than ranks). */
int p2pfl_gossip_fused_emulated(int device, void* stream, int n_ranks, float* const* params,
                                uint64_t n, uint32_t tile, uint32_t seq,
                                const p2pfl_push_table* push_tables,
                                const p2pfl_reduce_table* reduce_tables, uint64_t timeout_cycles,
                                int32_t* status /* int32[4 * n_ranks] */);

/* Allocation-free exchange among 2..8 nodes that share `device`.
 * (No counterpart in the reference: it has no device memory to manage.)
 *
 * For hosts that cannot ask CUDA for memory when a round's exchange is due - a
 * simulator that gave the whole GPU to its own pool at start-up, or a round loop
 * that cannot tolerate the device-wide synchronisation an allocation or a free
 * implies.  Same kernel and same arithmetic as p2pfl_gossip_fused_emulated (bf16
 * wire, contributors summed in rank order, every rank ends with identical bits),
 * but the inbox, the arrival flags, the kernel's tables and its status words live
 * in a 4200192-byte array inside the library image, so this call performs no
 * cudaMalloc / cudaFree and calls nothing that allocates.  Every GPU that loads
 * the library carries that array until the process exits.
 *
 *   params[r]  (host array of device pointers) fp32 model of rank r, n elements,
 *              16-byte aligned; overwritten with the weighted average
 *   weights[r] (host array) normalised FedAvg weight of rank r
 *   n          elements per model: a multiple of 8, at most 262144
 *   status_out (nullable; a DEVICE pointer, int32[4 * n_ranks] - it is filled by a
 *              device-to-device copy on `stream`, a host pointer would fault) per-rank
 *              status words, same meaning as for the reduce kernels
 *
 * Anything larger (more ranks, bigger models) is refused with a non-zero return
 * and a message in p2pfl_last_error(); use p2pfl_gossip_fused_emulated with inboxes
 * you own instead.
 *
 * The first call on a device looks up the array's device address
 * (cudaGetSymbolAddress); make that first call before starting any stream
 * capture on that device.  All calls share the one array and it keeps what the
 * previous call left (this entry point re-zeroes the flags and status itself):
 * two calls whose GPU work overlaps - different streams, different threads -
 * race on it.  Serialise them. */
int p2pfl_gossip_exchange_static(int device, void* stream, int n_ranks, float* const* params,
                                 const float* weights, uint64_t n, uint64_t timeout_cycles,
                                 int32_t* status_out);

/* --------------------------------------------------------------------- GEMM
 * K1/K2/K4/K8: D[M,N] = A[M,K] . B[N,K]^T on tcgen05 tensor cores, operands bf16
 * K-major staged by TMA (128B swizzle), fp32 accumulators in TMEM, fused
 * epilogue (mlp.py:63-67 Linear+ReLU forward, autograd dgrad/wgrad backward).
 * Epilogue, in order: + bias (per row if bias_per_row else per column), ReLU if
 * relu, zero where mask<=0 (mask bf16 [M, ld_mask], same layout as D), then any
 * of the four outputs (nullable), and row sums (rowsum[m] += sum_n D[m,n]; when
 * N <= 32 the row sum is stored, not accumulated, so no zero-init is needed). */
typedef struct {
  const void* A; int32_t lda;   /* bf16 [M, lda] */
  const void* B; int32_t ldb;   /* bf16 [N, ldb] */
  int32_t M, N, K;
  const float* bias; int32_t bias_per_row;
  int32_t relu;
  const void* mask; int32_t ld_mask;
  float* out_f32;   int32_t ld_f32;     /* [M, ld_f32] */
  void*  out_bf16;  int32_t ld_bf16;    /* [M, ld_bf16] */
  float* out_f32_t; int32_t ld_f32_t;   /* [N, ld_f32_t] (transposed) */
  void*  out_bf16_t; int32_t ld_bf16_t; /* [N, ld_bf16_t] (transposed) */
  float* rowsum;                        /* [M], atomically accumulated */
  float alpha;                          /* scales the accumulator before the epilogue */
  /* 0/1: one CTA per output tile.  -1 (pick) or n > 1: the K loop is cut in slices that
   * accumulate into out_f32 with atomics - the caller zero-fills out_f32, and bias / ReLU /
   * mask / the other outputs are not available.  For weight gradients (small D, long K). */
  int32_t split_k;
  /* bit 0: A is stored [K, lda] with M contiguous; bit 1: B is stored [K, ldb] with N
   * contiguous (MN-major tcgen05 operands), so D = A^T-stored . B-stored needs no transposed
   * copies: wgrad = dZ[pix, Cout]^T . col[pix, K], dgrad = dZ . W[Cout, K].
   * p2pfl_gemm_tn / _grouped only. */
  int32_t mn_major;
} p2pfl_gemm_args;
int p2pfl_gemm_tn(int device, void* stream, const p2pfl_gemm_args* args);
/* Up to 4 independent GEMMs in ONE launch (e.g. a layer's wgrad and dgrad, which
 * both depend only on that layer's dZ): fewer launches on the latency-bound step. */
int p2pfl_gemm_tn_grouped(int device, void* stream, const p2pfl_gemm_args* args, int count);

/* The dependent part of a small-batch train step as ONE launch: a single CTA walks a
 * list of links - GEMMs (same arguments as p2pfl_gemm_tn, K-major, no split-K; N tiled by 32) and at most
 * one softmax-cross-entropy link - each consuming what the previous one stored
 * (mlp.py:57-80 forward chain + loss + autograd dgrad chain). */
typedef struct {
  const float* logits; int32_t ld_logits;   /* fp32 [B, ld_logits] */
  const int32_t* labels; int32_t B, C;       /* C <= 256 */
  void* dlogits_bf16; int32_t ld_d;          /* bf16 [B, ld_d], already / B */
  void* dlogits_t_bf16; int32_t ld_dt;       /* bf16 [C, ld_dt] */
  float* dbias;                              /* [C], overwritten (nullable) */
  float* loss_sum;                           /* += sum of per-sample losses (nullable) */
} p2pfl_ce_args;
int p2pfl_gemm_chain(int device, void* stream, const p2pfl_gemm_args* gemms, int n_gemm,
                     const int32_t* link_kind /* 0 GEMM, 1 cross-entropy */, const int32_t* link_idx,
                     int n_links, const p2pfl_ce_args* ce);
/* The same chain as ONE thread-block cluster of `cluster_ctas` CTAs (1, 2, 4, 8; 16 where the
 * device allows it): the 128 x 32 tiles of a link are dealt round-robin to the CTAs and links
 * are separated by the hardware cluster barrier.  `phase_cnt[p]` > 1 makes link p a group of
 * independent GEMMs (gemms[idx] .. gemms[idx+cnt-1], e.g. every layer's weight gradient) whose
 * tiles share the link.  No counterpart in the reference. */
int p2pfl_gemm_chain_cluster(int device, void* stream, const p2pfl_gemm_args* gemms, int n_gemm,
                             const int32_t* phase_kind, const int32_t* phase_idx,
                             const int32_t* phase_cnt /* nullable */, int n_phases,
                             const p2pfl_ce_args* ce, int cluster_ctas);
/* Largest cluster_ctas (16, 8, 4, 2 or 1) the device can co-schedule for the chain kernel. */
int p2pfl_gemm_chain_max_cluster(int device);
/* Bring-up aid: while `device_buffer` (int64 [12][8], NULL = off) is set, every chain launch
 * makes CTA 0 record SM-clock stamps per link: 0 link start, 1 first operands landed, 2 MMAs
 * issued, 3 first accumulator complete, 4 epilogue stores issued, 5 fences done, 6 barrier passed. */
int p2pfl_gemm_chain_set_stamps(void* device_buffer);

/* A whole training EPOCH of the 784-256-128-10 MLP (mlp.py:53-80: forward, cross-entropy, backward,
 * Adam; batch 32) as ONE launch of one 16-CTA thread-block cluster: bf16 weights resident in shared
 * memory, layer-1 slices exchanged through distributed shared memory, weight gradients consumed by
 * Adam straight out of TMEM.  `flat` / `m` / `v` are the fp32 master and Adam moments in ParamLayout
 * order (l1.weight, l1.bias, l2.weight, l2.bias, l3.weight, l3.bias; slots 128-byte aligned), `w16`
 * the bf16 working copy of `flat`; all four are updated every step.  xs: bf16 [n_rows, 784], the
 * epoch's samples in training order; step s uses rows 32s .. 32s+31 (the last step may be partial).
 * Adam's step number is *step_ptr + s + 1. */
typedef struct {
  const void* xs;
  const int32_t* ys;      /* int32 [n_rows] */
  int32_t n_rows, n_steps; /* n_steps = ceil(n_rows / 32) */
  float* flat; float* m; float* v; void* w16;
  float* grads_out;       /* nullable fp32, same layout as flat: the LAST step's gradients (tests) */
  float* loss;            /* nullable [n_steps]: sum of the per-sample losses of each step */
  const int32_t* step_ptr; /* nullable */
  float lr, beta1, beta2, eps;
} p2pfl_mlp_epoch_args;
int p2pfl_mlp_epoch(int device, void* stream, const p2pfl_mlp_epoch_args* args);
/* 1 if the device can co-schedule the 16-CTA cluster (non-portable size) the epoch kernel needs */
int p2pfl_mlp_epoch_supported(int device);

/* ----------------------------------------------------------- convolution path
 * K6/K7 (cnn.py:78-81 Conv2d + ReLU + MaxPool2d through cuDNN/ATen) and the
 * ResNet-style blocks of BASELINE.json config 3.  Activations are NHWC bf16.
 * Convolutions are implicit GEMMs on p2pfl_gemm_tn: `col` is the im2col matrix
 * [B*Ho*Wo, ld_col] with K ordered (c, i, j) like a flattened PyTorch conv weight;
 * `col_t` (nullable) is its transpose [C*kh*kw, ld_t] for the weight-gradient GEMM. */
int p2pfl_im2col(int device, void* stream, const void* x_nhwc, int B, int H, int W, int C, int kh,
                 int kw, int stride, int pad, void* col, int ld_col, void* col_t, int64_t ld_t);
/* gradient of im2col: dx[b,h,w,c] gathers its kh*kw contributions (no atomics) */
int p2pfl_col2im(int device, void* stream, const void* dcol, int ld_col, int B, int H, int W, int C,
                 int kh, int kw, int stride, int pad, void* dx_nhwc);
/* 2x2/2 max pool; `arg` (uint8) keeps the winner; `flat` / `flat_t` (nullable) are the
 * output in NCHW-flatten order [B, C*Ho*Wo] and transposed, for a following Linear. */
int p2pfl_maxpool2_fwd(int device, void* stream, const void* x_nhwc, int B, int H, int W, int C,
                       void* y_nhwc /* nullable */, void* arg, void* flat, void* flat_t, int ld_flat_t);
/* backward of ReLU -> max pool in one pass: routes dy (NHWC or NCHW-flat, whichever is
 * given) to the winners, applies the ReLU mask of `act_nhwc`, writes [pixels, C] and
 * [C, ld_t], and accumulates the per-channel sum (conv bias gradient) into dbias. */
int p2pfl_maxpool2_relu_bwd(int device, void* stream, const void* dy_nhwc, const void* dy_flat,
                            const void* arg, const void* act_nhwc, int B, int H, int W, int C,
                            void* dx, void* dx_t /* nullable */, int64_t ld_t, float* dbias /* nullable */);
/* fp32 [rows, cols] -> bf16 [rows, ld] with zero padding (16-byte row pitch for TMA) */
int p2pfl_pad_cast(int device, void* stream, const float* src, int rows, int cols, void* dst_bf16, int ld);
int p2pfl_nchw_to_nhwc(int device, void* stream, const float* x, int B, int C, int H, int W, void* y_bf16);
/* training-mode BatchNorm2d over NHWC (+ optional residual add and ReLU), saving
 * mean / rstd for the backward pass and updating the running statistics */
int p2pfl_bn_fwd(int device, void* stream, const void* x_nhwc, int64_t pixels, int C,
                 const float* gamma, const float* beta, float* running_mean, float* running_var,
                 float momentum, float eps, int relu, const void* residual /* nullable */, void* y,
                 float* save_mean, float* save_rstd);
int p2pfl_bn_bwd(int device, void* stream, const void* x_nhwc, const void* y, const void* dy,
                 int64_t pixels, int C, const float* gamma, const float* save_mean,
                 const float* save_rstd, int relu, void* dx, void* d_residual /* nullable */,
                 float* dgamma, float* dbeta);
/* same, additionally writing dx transposed ([C, ld_t]) for the conv weight-gradient GEMM */
int p2pfl_bn_bwd_t(int device, void* stream, const void* x_nhwc, const void* y, const void* dy,
                   int64_t pixels, int C, const float* gamma, const float* save_mean,
                   const float* save_rstd, int relu, void* dx, void* dx_t /* nullable */, int64_t ld_t,
                   void* d_residual /* nullable */, float* dgamma, float* dbeta);
/* inference-mode BatchNorm (running statistics) (+ residual, + ReLU) */
int p2pfl_bn_infer(int device, void* stream, const void* x_nhwc, int64_t pixels, int C,
                   const float* gamma, const float* beta, const float* running_mean,
                   const float* running_var, float eps, int relu, const void* residual, void* y);
/* col2im that also adds `extra` (same shape as dx; e.g. the gradient of a residual branch) */
int p2pfl_col2im_add(int device, void* stream, const void* dcol, int ld_col, int B, int H, int W,
                     int C, int kh, int kw, int stride, int pad, void* dx_nhwc, const void* extra);
/* global average pool over HW: [B, HW, C] -> [B, C] (+ transpose), and its backward */
int p2pfl_avgpool_fwd(int device, void* stream, const void* x_nhwc, int B, int HW, int C, void* y,
                      void* y_t /* nullable */, int ld_t);
int p2pfl_avgpool_bwd(int device, void* stream, const void* dy, int B, int HW, int C, void* dx_nhwc);

/* Tap-major ("kc") convolution layout, for C % 8 == 0: the K axis of the implicit GEMM is
 * ordered (i, j, c) - channel fastest - so with NHWC activations im2col / col2im move
 * 16-byte channel vectors and the column matrix is one dense stream.
 *   col[(b,ho,wo), (i*kw + j)*C + c] = x[b, ho*s+i-p, wo*s+j-p, c]
 * The weights are repacked to the same K order (p2pfl_conv_pack, all convolutions of a
 * network in ONE launch, fp32 -> bf16 [cout, ldk]); the weight gradient the wgrad GEMM
 * accumulates in packed order is put back with p2pfl_conv_unpack.  `table_dev`: int64
 * [n_conv, 8] on the device, rows {flat offset, packed offset, cout, cin, taps, ldk,
 * tap_major (0 keeps PyTorch's (c,i,j) order: the C = 3 stem), 0}; `max_elems` = the
 * largest cout*cin*taps. */
int p2pfl_im2col_kc(int device, void* stream, const void* x_nhwc, int B, int H, int W, int C,
                    int kh, int kw, int stride, int pad, void* col, int ld_col);
int p2pfl_col2im_kc(int device, void* stream, const void* dcol, int ld_col, int B, int H, int W,
                    int C, int kh, int kw, int stride, int pad, void* dx_nhwc,
                    const void* extra /* nullable, added */);
int p2pfl_conv_pack(int device, void* stream, const float* flat, void* packed_bf16,
                    const int64_t* table_dev, int n_conv, int64_t max_elems);
int p2pfl_conv_unpack(int device, void* stream, const float* packed_f32, float* flat,
                      const int64_t* table_dev, int n_conv, int64_t max_elems);

/* ------------------------------------------------------------ data movement
 * K11/N8: device-resident dataset; gather a shuffled epoch in bf16, both as
 * [n, feat] and per-batch transposed [n/B][feat][B] for the wgrad GEMM. */
int p2pfl_gather_rows(int device, void* stream, const float* src, const int32_t* perm,
                      int n_rows, int feat, int batch, void* dst_bf16, void* dst_t_bf16 /* nullable */,
                      const int32_t* labels_src /* nullable */, int32_t* labels_dst /* nullable */);

#ifdef __cplusplus
}
#endif
#endif /* P2PFL_B200_H */
