/*
 * dls_b200 — C-ABI of the B200 (sm_100a) batched executor for the decentralized-learning
 * simulator's compute-graph solve (train / aggregate / test tasks).
 *
 * Boundary: the reference runs one task of one simulated node per call through
 *   Worker.start                      dasklearn/worker.py:21-38
 *   functions.train/aggregate/test    dasklearn/functions.py:28-77, 89-106, 109-133
 * This library executes a whole wavefront of same-kind tasks per call on rows of the stacked
 * parameter tensor `[S, ld]` (one fp32 row per node model, `parameters()` order).
 *
 * Conventions
 *   - plain C symbols, raw DEVICE pointers, explicit cudaStream_t (passed as void*);
 *   - the caller owns every buffer including workspaces; calls are asynchronous on `stream`;
 *   - return value: 0 = ok, negative = argument/shape error (DLS_E_*), positive = cudaError_t;
 *   - no CPU fallback: without a CUDA device every compute entry point returns an error.
 *   - rows must be 16-byte aligned: `ld % 4 == 0` and 16-byte aligned base pointers.
 */
#ifndef DLS_B200_H
#define DLS_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define DLS_E_ARG      (-1)   /* null pointer / negative size */
#define DLS_E_ALIGN    (-2)   /* pointer or leading dimension not 16-byte aligned */
#define DLS_E_SHAPE    (-3)   /* unsupported shape (batch too large, ld < P, ...) */
#define DLS_E_WS       (-4)   /* workspace too small */
#define DLS_E_TOO_BIG  (-5)   /* request exceeds a compiled-in maximum */

/* Model families (dasklearn/models/__init__.py:19-42): 0 = GNLeNet (cifar10), 1 = FEMNIST CNN */
#define DLS_ARCH_GNLENET 0
#define DLS_ARCH_FEMNIST 1

#define DLS_GNLENET_P   85354
#define DLS_FEMNIST_P 1690046
#define DLS_MAX_GROUP      64   /* max minibatch size (settings.learning.batch_size) */

/* Library / build information. */
int         dls_version(void);
const char *dls_build_info(void);
/* Number of kernels this library has launched in the calling process (all entry points). */
int64_t     dls_launch_count(void);
/* Diagnostic knobs for experiments (scripts/): key 0 = aggregate tile width in float4 columns, 1 = aggregate
 * pipeline stages, 2 = number of equal lanes a train wave is split into, 3 = 1: round-1 tensor-core convolution
 * kernels (one MMA per tap, fp32 activations converted by producer warps) instead of the plane / bulk-copy /
 * tap-pair kernels, 5 = 1: separate optimizer kernel for the whole row instead of the update fused into the
 * weight-gradient kernels, 6 = CTAs per SM of dls_push_rows, 7 = dls_push_rows kernel form (0: 16-byte accesses,
 * grid-stride; 1: 32-byte accesses; 2: persistent CTAs sweeping the transfer list front to back); value 0 restores the
 * default.  Process-wide, not thread-safe, never needed for correct results.  No reference counterpart. */
int         dls_tune(int32_t key, int32_t value);

/* ------------------------------------------------------------------------------------------
 * aggregate wave — replaces N calls of functions.aggregate -> ModelManager ->
 * FedAvg.aggregate (dasklearn/functions.py:89-106, dasklearn/model_manager.py:27-43,
 * dasklearn/gradient_aggregation/fedavg.py:12-26).
 *
 *   Y[out_rows[r], 0:P] = sum_{e in [rowptr[r], rowptr[r+1])} wval[e] * X[src_rows[colidx[e]], 0:P]
 *
 * evaluated per element in list order as acc = x_first*0; acc = fadd_rn(acc, fmul_rn(w, x))
 * (separate multiply and add, no FMA) — bit-identical to the reference's
 * `center.add_(w * p)` loop.  `src_rows` lists the n_src_rows distinct source rows of the wave
 * (NULL = identity: colidx indexes X directly); `out_rows == NULL` means row r; `nnz` =
 * rowptr[n_rows], the number of (colidx, wval) entries.  X and Y may be
 * the same allocation as long as no output row is also an input row of the same call.
 * `variant`: 0 = auto (gather), 1 = gather through L2, 2 = shared-memory staged column tiles,
 * 3 = staged + the caller asserts a uniform fan-in F = nnz / n_rows in [2, 16] (rowptr[r] == r*F),
 * 4 = like 3 and the caller also asserts that every entry carries the weight wval[0] (a D-PSGD wave without
 * explicit weights): the CSR is kept as 16-bit indices, so larger waves fit (1024 source rows x fan-in 11).
 * 2, 3 and 4 return DLS_E_SHAPE when the wave's source rows / CSR do not fit in shared memory.
 * ------------------------------------------------------------------------------------------ */
int dls_agg_spmm_f32(const float *X, int64_t ldx, float *Y, int64_t ldy,
                     const int32_t *rowptr, const int32_t *colidx, const float *wval,
                     const int32_t *src_rows, const int32_t *out_rows, int32_t n_rows,
                     int32_t n_src_rows, int32_t nnz, int64_t P, int32_t variant, void *stream);
/* 1 if `variant` (2, 3 or 4) accepts a wave of this shape on the current device, 0 if not (variant 1 always
 * does), negative on bad arguments: lets the caller choose a variant without a trial launch. */
int dls_agg_variant_fits(int32_t variant, int32_t n_rows, int32_t n_src_rows, int32_t nnz);

/* ------------------------------------------------------------------------------------------
 * reconstruct_from_chunks wave — replaces N calls of functions.reconstruct_from_chunks ->
 * ChunkManager.reconstruct_model (dasklearn/functions.py:142-146,
 * dasklearn/simulation/conflux/chunk_manager.py:34-53).  `functions.chunk` (functions.py:136-140,
 * chunk_manager.py:12-25) needs no kernel: a chunk is a column range of a row.
 *
 * Segment s (one chunk index of one reconstructed model):
 *   Y[seg_out_row[s], c0 : c0+len] = mean over e in [seg_ptr[s], seg_ptr[s+1]) of X[seg_src_rows[e], c0 : c0+len]
 * with c0 = seg_col0[s], len = seg_len[s], evaluated in the summation order of
 * torch.mean(torch.stack(pieces), dim=0) on the CPU (sequential over contributors with a 16-row cascade
 * for the columns below the last multiple of 32 of the piece, four interleaved partial sums for the rest;
 * sum divided by the contributor count) — bit-identical to the reference on the same pieces.
 * max_len = largest seg_len, max_contributors = largest contributor count (<= 1024, else DLS_E_TOO_BIG).
 * Output ranges of different segments must not overlap; X and Y may be the same allocation when no output
 * row is also a contributor row of the call.  No alignment requirement (chunk bounds are arbitrary).
 * ------------------------------------------------------------------------------------------ */
int dls_chunk_mean_f32(const float *X, int64_t ldx, float *Y, int64_t ldy, const int32_t *seg_ptr,
                       const int32_t *seg_src_rows, const int32_t *seg_out_row, const int64_t *seg_col0,
                       const int32_t *seg_len, int32_t n_segs, int32_t max_len, int32_t max_contributors,
                       void *stream);

/* Row plumbing on the stacked tensor (replaces the pickle round-trip copies of
 * serialize_model/unserialize_model, dasklearn/models/__init__.py:9-16). */
int dls_copy_rows(const float *X, int64_t ldx, float *Y, int64_t ldy, const int32_t *src_rows,
                  const int32_t *dst_rows, int32_t n_rows, int64_t P, void *stream);
int dls_broadcast_row(const float *src, float *Y, int64_t ldy, const int32_t *dst_rows,
                      int32_t n_rows, int64_t P, void *stream);

/* Multi-GPU halo push (replaces the broker-to-broker transfer of pickled models, dasklearn/broker.py:207-219):
 * transfer e copies X[src_rows[e], 0:P) to the row that starts at device address dst_ptrs[e], which may be
 * PEER memory mapped into this process (another GPU's halo rows).  Destinations must be 16-byte aligned and hold
 * at least ceil(P/4)*4 floats.  The kernel ends with a system-scope fence; the caller must run a cross-GPU
 * barrier on `stream` before any GPU reads the rows pushed to it.  At most 65535 transfers per call.  The kernel
 * works through the list roughly front to back, so its ORDER decides which peers are written at the same time:
 * interleave the destinations (round-robin starting at rank + 1) — sorted by destination, every GPU of a box writes
 * to the same peer at once and the exchange runs at half the bandwidth (345 vs 700 GB/s per GPU on 8 B200s). */
int dls_push_rows(const float *X, int64_t ldx, const int32_t *src_rows, const uint64_t *dst_ptrs,
                  int32_t n_transfers, int64_t P, void *stream);

/* ------------------------------------------------------------------------------------------
 * train wave — replaces N calls of functions.train -> ModelTrainer.train
 * (dasklearn/functions.py:28-77, dasklearn/model_trainer.py:60-131) + torch.optim.SGD
 * (dasklearn/optimizer/sgd.py:4-14).
 *
 * One call runs up to `n_steps` local SGD steps for `n_tasks` nodes, in place on their rows.
 *   X, ldx        stacked parameters; task t trains row task_row[t]
 *   M, ldm        momentum buffers, one row per NODE; task t uses row task_peer[t]
 *                 (momentum persists per node across rounds, model_trainer.py:89-92);
 *                 may be NULL when mu == 0
 *   mom_first[t]  1 if node task_peer[t] has no momentum buffer yet (buf = grad on its first step)
 *   images/labels the nodes' uint8 HWC partitions, image i at images + i*img_bytes
 *   mb_index      [n_steps][n_tasks][batch] GLOBAL image ids (-1 = empty slot; the empty slots of a task's
 *                 minibatch are its LAST ones: a partial batch fills a prefix)
 *   mb_count      [n_steps][n_tasks] delivered batch size (0 = task already finished its loader)
 *   loss_out      optional [n_steps][n_tasks] mean minibatch loss
 * Loss: NLL of log_softmax (GNLeNet) / cross entropy (FEMNIST), mean over the delivered batch.
 * Math: fp32 CUDA-core kernels when `precision == 0`; `precision == 1` uses bf16 tensor-core
 * operands with fp32 accumulation for the convolution GEMMs (GNLeNet: all three convolutions; FEMNIST: conv2 and the
 * 3136x512 dense layer, whose optimizer step then happens inside its weight-gradient kernel).
 * ------------------------------------------------------------------------------------------ */
int64_t dls_train_workspace_bytes(int32_t arch, int32_t n_tasks, int32_t batch);
int dls_train_wave(int32_t arch, float *X, int64_t ldx, float *M, int64_t ldm,
                   const int32_t *task_row, const int32_t *task_peer, const uint8_t *mom_first,
                   const uint8_t *images, const int64_t *labels,
                   const int32_t *mb_index, const int32_t *mb_count,
                   int32_t n_tasks, int32_t n_steps, int32_t batch,
                   float lr, float mu, float wd, float *loss_out, int32_t precision,
                   void *workspace, int64_t workspace_bytes, void *stream);

/* Gradient of one minibatch per task, written to G[t, 0:P] (no update).  Exposes the backward
 * pass for compute_gradient tasks (dasklearn/functions.py:80-82) and for parity tests. */
int dls_grad_wave(int32_t arch, const float *X, int64_t ldx, float *G, int64_t ldg,
                  const int32_t *task_row, const uint8_t *images, const int64_t *labels,
                  const int32_t *mb_index, const int32_t *mb_count, int32_t n_tasks, int32_t batch,
                  float *loss_out, int32_t precision, void *workspace, int64_t workspace_bytes,
                  void *stream);

/* gradient_update wave — replaces N calls of functions.gradient_update -> ModelTrainer.gradient_update
 * (dasklearn/functions.py:84-86, dasklearn/model_trainer.py:133-143): one step of a FRESH SGD optimizer
 * with a supplied gradient, X[out_rows[r]] = X[p_rows[r]] - lr * (X[g_rows[r]] + wd * X[p_rows[r]])
 * (single rounding, like torch's add_(buf, alpha=-lr)); gradients live in rows of X (as written by
 * dls_grad_wave followed by dls_copy_rows).  out_rows[r] may equal p_rows[r] (in place); rows of
 * different r must not alias.  The node's momentum rows are not touched.  X must be 16-byte aligned and
 * ldx a multiple of 4 (DLS_E_ALIGN otherwise): the kernel moves float4. */
int dls_gradient_update(float *X, int64_t ldx, const int32_t *p_rows, const int32_t *g_rows,
                        const int32_t *out_rows, int32_t n_rows, int64_t P, float lr, float wd, void *stream);

/* ------------------------------------------------------------------------------------------
 * test wave — replaces N calls of functions.test -> ModelEvaluator.evaluate_accuracy
 * (dasklearn/functions.py:109-133, dasklearn/model_evaluator.py:26-64).
 *
 * Evaluates `n_models` rows on the same `n_test` images.  Outputs per model:
 *   out_correct[m]   number of argmax hits (first maximum wins, like torch.argmax)
 *   out_loss_sum[m]  sum over samples of the per-sample loss (fp32 within each group of 32 samples,
 *                    fp64 across the groups, in sample order)
 * The caller derives accuracy = correct / n_test * 100 and loss = loss_sum / n_test.
 * ------------------------------------------------------------------------------------------ */
int64_t dls_eval_workspace_bytes(int32_t arch, int32_t n_models, int32_t n_test);
int dls_eval_wave(int32_t arch, const float *X, int64_t ldx, const int32_t *model_rows,
                  int32_t n_models, const uint8_t *test_images, const int64_t *test_labels,
                  int32_t n_test, int32_t *out_correct, double *out_loss_sum,
                  float *out_logits /* optional [n_models][n_test][classes] */, int32_t precision,
                  void *workspace, int64_t workspace_bytes, void *stream);

/* Diagnostic: run `n_mma` tcgen05.mma (kind::f16, bf16 operands, fp32 accumulate) instructions on
 * caller-provided shared-memory operand images and return the TMEM accumulator [128 lanes][ncols].
 * Used by the tests to pin the descriptor encodings of the tensor-core convolution kernels.
 * No reference counterpart. */
int dls_tc_probe(const uint8_t *a_img, int32_t a_bytes, const uint8_t *b_img, int32_t b_bytes,
                 const uint64_t *a_desc, const uint64_t *b_desc, int32_t n_mma, uint32_t idesc,
                 int32_t ncols, float *out, int32_t repeat /* re-issue the first 8 MMAs this many times */,
                 long long *cycles_out /* optional: SM cycles for the repeated MMAs */, void *stream);

/* Diagnostic: one CTA issues `n_mma` tcgen05.mma (M128 N96 K16, bf16) while four warps run `ld_loops` rounds of
 * TMEM loads (3 x 32 columns) and `shfl_per_round` warp shuffles; cycles[0] = SM cycles of the MMA stream,
 * cycles[1] = of the load loop.  Measures how the tensor pipe, TMEM reads and the shuffle path contend
 * (scripts/probe_contention.py).  No reference counterpart. */
int dls_tc_contention(int32_t n_mma, int32_t ld_loops, int32_t shfl_per_round, long long *cycles, float *sink,
                      void *stream);

/* ------------------------------------------------------------------------------------------
 * test wave without a workspace argument.  This entry point has NO counterpart in the reference;
 * it exists for callers that cannot allocate device memory (a foreign-function binding that only
 * forwards model / image / result buffers, a monitoring thread while the solver owns the training
 * workspace, a caller that must avoid cudaMalloc's device-wide synchronisation mid-round).
 *
 * GNLeNet only (arch must be DLS_ARCH_GNLENET; DLS_E_SHAPE for the FEMNIST CNN, whose fc1 activations
 * do not fit the fixed scratch).  Same outputs as dls_eval_wave (fp32 path, no logits).  The scratch memory is a fixed 32 MiB
 * array inside the library, present on every GPU that loads it until the process exits; it
 * limits a call to n_models <= 16 and n_test <= 4096, beyond which DLS_E_TOO_BIG is returned.
 * The call itself performs no allocation of any kind.  All calls on a device share that array
 * (it is re-initialised where needed at the start of each call), so calls that overlap on the
 * GPU race with each other: issue them one after another.  The first call on each device
 * resolves the array's address; make that first call outside of stream capture.
 * ------------------------------------------------------------------------------------------ */
int dls_eval_wave_noalloc(int32_t arch, const float *X, int64_t ldx, const int32_t *model_rows,
                          int32_t n_models, const uint8_t *test_images, const int64_t *test_labels,
                          int32_t n_test, int32_t *out_correct, double *out_loss_sum, void *stream);

#ifdef __cplusplus
}
#endif
#endif /* DLS_B200_H */
