/* per_b200.h — C ABI of the B200-native prioritized-replay library (libper_b200.so).
 *
 * This is the boundary a maintainer of Bing-Jing/Ape-X binds instead of the
 * pure-Python classes of the reference's memory.py (the reference has no FFI
 * of its own: it is Python only; the ctypes binding that replaces it is
 * ape-x_b200/per_native.py, and INTEGRATION.md shows the stub).  Every entry point
 * cites the reference function it replaces as  file:line  relative to the
 * reference repository root.
 *
 * Conventions
 *  - All pointers named d_* are DEVICE pointers on the handle's device; h_* are
 *    host pointers.  The caller owns every buffer it passes; the handle owns
 *    only the two trees, the transition ring and a small scratch area.
 *  - `stream` is a cudaStream_t passed as void*.  All work is enqueued on it and
 *    is stream-ordered; nothing synchronises unless the comment says so.
 *  - One host thread per handle at a time (the reference object is not
 *    thread-safe either; origin_repo/replay.py:92,121,142 serialises it).
 *  - Return value: 0 = PER_OK, negative = error; per_last_error() gives the text
 *    of the most recent failure on the calling thread.  Host-checkable
 *    preconditions the reference asserts (memory.py:34,224,284,313) return
 *    PER_ERR_ASSERT; data-dependent ones (priority <= 0, index out of range,
 *    memory.py:315-316,121) set bits in a device status word read by
 *    per_status() / seg_status().
 *  - Tree node arrays use the reference's heap layout (memory.py:36): 2*capacity
 *    float64, node 1 = root, children of i are 2i and 2i+1, leaf k at capacity+k,
 *    so a copy to the host compares element by element with `_value`.
 */
#ifndef PER_B200_H
#define PER_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define PER_OK            0
#define PER_ERR_ASSERT   -1   /* a precondition the reference asserts */
#define PER_ERR_CUDA     -2   /* a CUDA runtime call failed */
#define PER_ERR_ARG      -3   /* malformed argument (null handle, negative size, ...) */
#define PER_ERR_NOMEM    -4

/* bits of the device status word */
#define PER_STATUS_BAD_PRIORITY   1   /* priority <= 0 or NaN      (memory.py:315) */
#define PER_STATUS_BAD_INDEX      2   /* idx outside [0, len)      (memory.py:316, :90) */
#define PER_STATUS_BAD_PREFIXSUM  4   /* mass outside [0, sum+1e-5] (memory.py:121) */
#define PER_STATUS_NSTEP_EMPTY    8   /* done arrived on an empty n-step window (IndexError at memory.py:417) */
#define PER_STATUS_EXCHANGE_TIMEOUT 16 /* a peer's shard record did not arrive within 2 s (sharded replay) */

#define PER_TREE_SUM 0
#define PER_TREE_MIN 1

#define PER_PRIO_F64 0
#define PER_PRIO_F32 1

typedef struct seg_tree_s seg_tree;      /* one SumSegmentTree / MinSegmentTree          */
typedef struct per_buffer_s per_buffer;  /* one (Custom)PrioritizedReplayBuffer(_AQL)    */
typedef struct aql_plan_s aql_plan;      /* weights of one AQL Q_Network, prepared for the fused forward */

const char *per_last_error(void);
int per_version(void);

/* ---------------------------------------------------------------- segment trees */

/* SegmentTree.__init__ (memory.py:11-37): `capacity` must be a power of two
 * (else PER_ERR_ASSERT); kind PER_TREE_SUM (neutral 0.0) or PER_TREE_MIN (+inf). */
int seg_create(seg_tree **out, int device, int64_t capacity, int kind);
int seg_destroy(seg_tree *t);

/* Batched SegmentTree.__setitem__ (memory.py:76-87): equivalent to executing
 * t[d_idx[j]] = d_val[j] for j = 0..n-1 in order (a later duplicate wins). */
int seg_set(seg_tree *t, const int64_t *d_idx, const double *d_val, int64_t n, void *stream);

/* SegmentTree.reduce(start, end) (memory.py:39-74), half-open, same association as
 * the reference's recursion; result written to d_out[0].  end <= start is
 * PER_ERR_ASSERT (the reference recurses without bound there). */
int seg_reduce(seg_tree *t, int64_t start, int64_t end, double *d_out, void *stream);

/* Batched SumSegmentTree.find_prefixsum_idx (memory.py:106-129). */
int seg_find_prefixsum(seg_tree *t, const double *d_mass, int64_t n, int64_t *d_out_idx, void *stream);

/* Copies the 2*capacity node array (`_value`, memory.py:36) to d_out. */
int seg_read(seg_tree *t, double *d_out, void *stream);
/* SegmentTree.__getitem__ (memory.py:89-91): leaf idx to *h_out (8 bytes cross the bus, not the
 * tree).  idx outside [0, capacity) is PER_ERR_ASSERT (memory.py:90).  Synchronises `stream`. */
int seg_get(seg_tree *t, int64_t idx, double *h_out, void *stream);
int64_t seg_capacity(const seg_tree *t);
/* Reads and clears the status word.  Synchronises `stream`. */
int seg_status(seg_tree *t, int64_t *h_out, void *stream);

/* ---------------------------------------------------------------- AQL candidate Q + argmax
 * Q_Network.act (model.py:294-335): for each state b and candidate action k,
 *   e = action_out(a_bk); h = relu(advantage1(cat[e, f_b])); q = advantage2(h); argmax_k.
 * f [B,64] = q_feature(state) stays PyTorch (a per-state MLP); everything per candidate is
 * one fused kernel on the tensor cores (tcgen05, 3xTF32 for float32 parity).
 * continuous != 0: a_mu float32 [B,K,action_dim], action_out = Linear(A,128)-ReLU-Linear(128,64)-ReLU
 *                  (model.py:252-257); continuous == 0: a_mu int64 [B,K], action_out =
 *                  Linear(1,64)-ReLU (model.py:260-263). */
int aql_create(aql_plan **out, int device, int continuous, int action_dim);
int aql_destroy(aql_plan *p);
/* Device pointers to float32 weights in torch's [out, in] layout: action_out layers (d_wa2 /
 * d_ba2 NULL when discrete), then the EFFECTIVE advantage1 / advantage2 weights — mu, or
 * mu + sigma*epsilon in training mode (NoisyLinear.forward, model.py:132-143). */
int aql_set_weights(aql_plan *p, const float *d_wa1, const float *d_ba1, const float *d_wa2,
                    const float *d_ba2, const float *d_wadv1, const float *d_badv1,
                    const float *d_wadv2, const float *d_badv2, void *stream);
/* d_q float32 [B,K] out; d_argmax int64 [B] out or NULL (first maximum wins, torch.argmax). */
int aql_candidate_q(aql_plan *p, const float *d_wadv1, const float *d_badv1, const float *d_f,
                    const void *d_a_mu, int64_t B, int64_t K, float *d_q, int64_t *d_argmax,
                    void *stream);

/* Diagnostics: reads and clears the SM-clock cycles CTA 0 spent, summed over its tiles, in
 * {layer 1 + MMA issue, wait for layer-2 MMAs, layer-2 epilogue + issue, wait for layer-3 MMAs,
 * layer-3 epilogue}, the number of tiles it processed, and the cycles from kernel entry to its
 * first tile and to its exit (summed over launches).  The stamps are off until the first call
 * (they cost the traced CTA a global read-modify-write per phase).  Synchronises `stream`. */
int aql_trace(aql_plan *p, int64_t *h_out8, void *stream);

/* ---------------------------------------------------------------- n-step assembly
 * BatchStorage.add / multi_step_reward / compute_priorities (memory.py:415-440, 471-475,
 * 451-464) for n_streams actor streams at once.  Stream s is the steps
 * [d_offsets[s], d_offsets[s+1]) of the flat arrays d_reward (float64), d_action (int64),
 * d_done (uint8), d_q (float32 [T, n_actions], the actor's Q row at each step); it is
 * processed as by a BatchStorage created empty at its first step.  Emissions at steps before
 * d_emit_from[s] are suppressed (NULL: none are): pass a few earlier steps as pre-roll to
 * continue a stream whose windows are not empty.  h_gamma_pow: HOST table gamma**k for
 * k = 0..n_steps (CPython's pow, so returns are bit-identical).  Outputs, at slots
 * d_offsets[s] + e for the e-th emission of stream s (e < d_out_count[s]): step index of the
 * transition's state and of its bootstrap state, its action, n-step return (float64), done
 * (float32), actor priority (float64).  PER_STATUS_NSTEP_EMPTY is OR-ed into *d_status when a
 * done arrives on an empty window (the reference raises IndexError there). */
int nstep_assemble(int device, int64_t n_streams, const int64_t *d_offsets, const int64_t *d_emit_from,
                   const double *d_reward, const int64_t *d_action, const uint8_t *d_done,
                   const float *d_q, int64_t n_actions, int n_steps, const double *h_gamma_pow,
                   int64_t *d_out_first, int64_t *d_out_next, int64_t *d_out_action,
                   double *d_out_return, float *d_out_done, double *d_out_prio,
                   int64_t *d_out_count, uint64_t *d_status, void *stream);

/* Batched tree write for callers that cannot allocate device memory.  (No counterpart in
 * the reference, which has no device memory at all.)
 *
 * seg_set needs a handle from seg_create, and seg_create calls cudaMalloc for the node array
 * and for 8 bytes of arbitration stamps per leaf.  Some callers must not do that: a process
 * whose framework has already claimed the whole GPU for its own memory pool (a 2^21-row Atari
 * ring leaves little of the 180 GB), or a latency-critical loop that cannot afford the
 * device-wide synchronisation cudaMalloc/cudaFree imply.  This entry point allocates
 * nothing — no cudaMalloc, cudaFree or allocating library call on any path.  The caller
 * passes its own node arrays (heap layout of memory.py:36, 2*capacity float64 each, one of
 * the two may be NULL; initialise them to 0.0 / +inf) and the stamps live in a static
 * device array inside the library, SEG_NOALLOC_MAX_CAPACITY entries (8 MiB + 256 B on every
 * GPU that loads the library, held until the process exits).
 *
 *  - Semantics are seg_set's: for j in 0..n-1 in order, node[capacity + d_idx[j]] = d_val[j]
 *    in both trees, then every touched ancestor = op(left, right) (memory.py:76-87); a later
 *    duplicate wins.  Out-of-range indices are skipped and PER_STATUS_BAD_INDEX is OR-ed into
 *    *d_status (a caller-owned device word; may be NULL).
 *  - capacity > SEG_NOALLOC_MAX_CAPACITY is rejected with PER_ERR_ARG and a message naming
 *    the limit; use seg_create/seg_set there.
 *  - The first call on each device looks up the array's device address
 *    (cudaGetSymbolAddress, with that device current) and caches it; make that first call
 *    before any stream capture begins.
 *  - All calls on a device share the one array.  It must be all zero when a call starts and
 *    every call leaves it so: the item that wins a leaf puts that leaf's entry back to zero
 *    as it writes the leaf, so trees of different sizes can take turns and a call replayed
 *    from a CUDA graph starts clean.  Two calls whose kernels overlap on the GPU (different
 *    streams without ordering, or two host threads) race on it: order them. */
#define SEG_NOALLOC_MAX_CAPACITY (1 << 20)
int seg_set_noalloc(int device, double *d_sum_nodes, double *d_min_nodes, int64_t capacity,
                    const int64_t *d_idx, const double *d_val, int64_t n, uint64_t *d_status,
                    void *stream);

/* ---------------------------------------------------------------- prioritized replay */

/* PrioritizedReplayBuffer.__init__ (memory.py:209-233) + ReplayBuffer.__init__
 * (memory.py:147-157).  ring_size = the reference's `size` (the ring wraps there,
 * memory.py:169); tree capacity = next power of two (memory.py:227-229).
 * obs_bytes = bytes of one observation (28224 for uint8[4,84,84]); extra_bytes =
 * bytes of the per-row extra field (AQL's a_mu, memory.py:370), 0 if none.
 * alpha < 0 is PER_ERR_ASSERT (memory.py:224). */
int per_create(per_buffer **out, int device, int64_t ring_size, double alpha,
               int64_t obs_bytes, int64_t extra_bytes);
int per_destroy(per_buffer *b);

int64_t per_capacity(const per_buffer *b);   /* tree capacity (power of two)          */
int64_t per_len(const per_buffer *b);        /* len(_storage)      (memory.py:159-160) */
int64_t per_cursor(const per_buffer *b);     /* _next_idx          (memory.py:157,169) */
/* Restore support (the reference never saves its replay): after the rows
 * have been re-inserted with per_add_batch (saved leaves as d_leaf), puts `_next_idx` back
 * where it was and, if max_priority > 0, `_max_priority` too.  Synchronises `stream`. */
int per_set_cursor(per_buffer *b, int64_t cursor, double max_priority, void *stream);
/* _max_priority (memory.py:233,320).  Synchronises `stream`. */
int per_max_priority(per_buffer *b, double *h_out, void *stream);
int per_status(per_buffer *b, int64_t *h_out, void *stream);

/* n consecutive ReplayBuffer.add calls (memory.py:162-169) fused with their tree
 * writes.  Rows go to slots cursor, cursor+1, ... (mod ring_size); n <= ring_size.
 * Leaf values, first match wins:
 *   d_leaf != NULL : leaves = d_leaf[j] (float64, e.g. from per_pow_alpha_host: bit-exact
 *                    with the reference) and _max_priority is raised to batch_max_priority
 *   d_prio != NULL : leaves = d_prio[j] ** alpha on the device, running max raised
 *                    (CustomPrioritizedReplayBuffer.add, memory.py:334-346; the ingest
 *                    loop batchrecorder.py:133-137)
 *   both NULL      : leaves = _max_priority ** alpha   (memory.py:235-240, :368-378)
 * d_obs / d_next: n rows of obs_bytes; d_act int64[n]; d_rew, d_done float32[n];
 * d_extra: n rows of extra_bytes or NULL.  Row pointers may be NULL (tree-only use). */
int per_add_batch(per_buffer *b, const void *d_obs, const void *d_next, const int64_t *d_act,
                  const float *d_rew, const float *d_done, const void *d_extra,
                  const double *d_prio, const double *d_leaf, double batch_max_priority,
                  int64_t n, void *stream);

/* PrioritizedReplayBuffer._sample_proportional + the weight loop of sample()
 * (memory.py:242-250, 288-296).  d_u: batch uniforms in [0,1) standing for the
 * reference's random.random() draws, in order.  Outputs: d_out_idx int64[batch];
 * d_out_w64 float64[batch] and/or d_out_w32 float32[batch] (either may be NULL).
 * beta <= 0 is PER_ERR_ASSERT (memory.py:284); len < 2 is PER_ERR_ASSERT (the
 * reference's range sum over [0, len-2] does not terminate for len == 1). */
int per_sample_indices(per_buffer *b, const double *d_u, int64_t batch, double beta,
                       int64_t *d_out_idx, double *d_out_w64, float *d_out_w32, void *stream);

/* ReplayBuffer._encode_sample (memory.py:171-182, :348-362, :380-391): gathers the
 * rows d_idx[0..batch) into dense batch arrays.  Any output may be NULL. */
int per_gather(per_buffer *b, const int64_t *d_idx, int64_t batch,
               void *d_out_obs, void *d_out_next, int64_t *d_out_act, float *d_out_rew,
               float *d_out_done, void *d_out_extra, void *stream);

/* per_gather fused with the learner's widening of the frames (`torch.FloatTensor(states)`,
 * ApeX.py:50,53; origin_repo/learner.py:102-110): observation rows are written as
 * out_dtype 0 = float32, 1 = bfloat16, 2 = float16, value * scale (scale 1.0 reproduces the
 * reference's 0..255 floats; 1/255 normalises).  d_out_obs / d_out_next: batch rows of
 * obs_bytes ELEMENTS each (either may be NULL); the scalar columns as in per_gather.
 * Observation rows must be a multiple of 16 bytes (uint8[4,84,84] is). */
int per_gather_cast(per_buffer *b, const int64_t *d_idx, int64_t batch, int out_dtype, float scale,
                    void *d_out_obs, void *d_out_next, int64_t *d_out_act, float *d_out_rew,
                    float *d_out_done, void *stream);

/* The same with CHANNELS-LAST output rows (the layout cuDNN's convolutions prefer; the learner's first layer is one,
 * model.py:33): an observation is `channels` planes (uint8[4,84,84]: 4 planes of 7 056 bytes) and output row q holds, for
 * every pixel, its `channels` widened values next to each other — the memory of a torch tensor of shape
 * [batch, C, H, W] in torch.channels_last format.  channels must be 4 and the plane size a multiple of 4 bytes. */
int per_gather_cast_nhwc(per_buffer *b, const int64_t *d_idx, int64_t batch, int out_dtype, float scale, int channels,
                         void *d_out_obs, void *d_out_next, int64_t *d_out_act, float *d_out_rew, float *d_out_done,
                         void *stream);

/* PrioritizedReplayBuffer.update_priorities (memory.py:300-320), sequential
 * semantics (a later duplicate wins), running max raised.  Priorities are float64
 * (prio_dtype PER_PRIO_F64) or float32 (PER_PRIO_F32, what utils.py:77 hands back);
 * leaves are priority ** alpha evaluated on the device (CUDA pow, <= 2 ulp from
 * the C library's — for bit-identical leaves use per_update_leaves). */
int per_update_priorities(per_buffer *b, const int64_t *d_idx, const void *d_prio, int prio_dtype,
                          int64_t n, void *stream);

/* Same, but the caller supplies the leaves (priority ** alpha, float64) and the
 * raw priorities' maximum (for _max_priority, memory.py:320).  With leaves from
 * per_pow_alpha_host this reproduces the reference tree bit for bit. */
int per_update_leaves(per_buffer *b, const int64_t *d_idx, const double *d_leaf, double batch_max_priority,
                      int64_t n, void *stream);

/* ---- the same path with HOST buffers: what a binding that owns no device memory calls ----
 * (memory.py's sample() / update_priorities() mirrors go through these two; bench.py's `e2e`
 * number is measured on them.)
 *
 * PrioritizedReplayBuffer.sample (memory.py:252-298) in one call: h_u = `batch` uniforms in [0,1)
 * drawn by the caller (random.random(), memory.py:247); outputs are host pointers, NULL = not
 * wanted: frame batches h_obs / h_next (batch x obs_bytes, row-compact), h_act int64, h_rew /
 * h_done float32, h_extra (batch x extra_bytes), h_w64 float64 importance weights, h_idx int64.
 * Uploads the uniforms, runs the sampling and gather kernels into staging the handle keeps,
 * copies the results out (frames first: they are the PCIe time) and synchronises `stream`.
 * Pinned (page-locked) output buffers get the full PCIe rate; pageable ones work. */
int per_sample_host(per_buffer *b, const double *h_u, int64_t batch, double beta, void *h_obs,
                    void *h_next, int64_t *h_act, float *h_rew, float *h_done, void *h_extra,
                    double *h_w64, int64_t *h_idx, void *stream);

/* PrioritizedReplayBuffer.update_priorities (memory.py:300-320) with host arrays: checks the
 * reference's two assertions for every pair first (PER_ERR_ASSERT, nothing applied), raises the
 * priorities to alpha with the C library's pow (bit-identical leaves), uploads and launches the
 * update; stream-ordered, does not synchronise.  *h_batch_max (may be NULL) = max priority. */
int per_update_host(per_buffer *b, const int64_t *h_idx, const double *h_prio, int64_t n,
                    double *h_batch_max, void *stream);

/* Host helper: h_out[j] = pow(h_prio[j], alpha) with the C library's double pow —
 * the function CPython's `priority ** alpha` calls (memory.py:317).  No GPU work. */
int per_pow_alpha_host(const double *h_prio, double alpha, double *h_out, int64_t n);

/* The arithmetic of utils.compute_loss after the three forwards (utils.py:71-80,
 * identical in compute_loss_AQL utils.py:51-60) fused with update_priorities
 * (ApeX.py:58,63): double-DQN target, |td|, prios = 0.9*max_b(td)+0.1*td+1e-6,
 * Huber*weight loss, then leaves = prios**alpha written to both trees.
 * d_q, d_q_next (online net on s, s'), d_q_tgt (target net on s'): float32
 * [batch, n_actions] row-major.  gamma_n = gamma ** n_steps (float).  Outputs
 * (any may be NULL): d_out_prio float32[batch], d_out_td float32[batch],
 * d_out_loss float32[1].  If d_idx is NULL the trees are left alone. */
int per_td_priority_update(per_buffer *b, const float *d_q, const float *d_q_next, const float *d_q_tgt,
                           int64_t n_actions, const int64_t *d_act, const float *d_rew, const float *d_done,
                           const float *d_w, float gamma_n, const int64_t *d_idx, int64_t batch,
                           float *d_out_prio, float *d_out_td, float *d_out_loss, void *stream);

/* ---- sharded replay (the reference has one tree in one process) ------------
 * The replay of `global_size` transitions is cut into n_shards (power of two <= 16)
 * contiguous index ranges, one per GPU: with C = next_pow2(global_size) leaves in the
 * reference's single tree (memory.py:227-229), shard r owns leaves [r*C/n, (r+1)*C/n), i.e.
 * the subtree under the reference's node n + r, so the reference tree is reproduced exactly
 * by the n local trees plus the n-1 nodes above them.  A sharded handle interprets every
 * index it is given as GLOBAL (it subtracts r*C/n itself) and silently skips the value -1,
 * which per_sample_sharded writes for strata owned by other shards; per_gather,
 * per_update_*, per_td_priority_update can therefore be called on every rank with the same
 * full-batch arrays.  per_add_batch on a sharded handle takes the rows that fall into this
 * shard, in slot order (the caller routes them; ape-x_b200/sharded.py has the router). */
int per_create_shard(per_buffer **out, int device, int64_t global_size, int rank, int n_shards,
                     double alpha, int64_t obs_bytes, int64_t extra_bytes);

/* Writes this shard's 4-double record {root sum, root min, the shard-local part of the
 * reference's sum(0, len-1) chain (memory.py:244), local len} to d_out.  The caller
 * all-gathers the records of all shards in rank order (32 bytes per rank) and passes the
 * n_shards*4 doubles to per_sample_sharded.  global_len = len(_storage) of the whole replay. */
int per_shard_record(per_buffer *b, int64_t global_len, double *d_out, void *stream);

/* Optional: from now on every call that changes this shard's trees (per_update_*,
 * per_td_priority_update) also rewrites the record at d_record as its last step, computed
 * for a whole-replay len of global_len, so the learner loop (update -> sample, no inserts in
 * between) needs no per_shard_record launch.  Call again whenever the whole-replay len
 * changes; d_record == NULL detaches. */
int per_shard_attach_record(per_buffer *b, double *d_record, int64_t global_len);

/* memory.py:242-250 + 288-296 on the virtual global tree: every rank runs all `batch`
 * strata through the top log2(n_shards) levels (built from the gathered roots with the
 * reference's pairwise order), keeps the strata that land in its slice and finishes them on
 * its local tree.  d_out_idx[i] = global index, or -1 if stratum i belongs to another shard
 * (weights 0 there).  The union over ranks equals the single-tree result bit for bit.
 * The strata a shard owns are contiguous (masses ascend), and the kernel leaves their range
 * on the device: per_gather / per_update_* / per_td_priority_update calls that are handed the
 * SAME d_out_idx buffer and batch then visit only that range instead of skipping -1 entries one
 * by one, so per-rank work stays proportional to the rank's share of the batch.
 *
 * Binding between a sample and the calls that follow it (this entry point and per_sample_indices
 * alike): the library remembers the address and length of the index buffer a sample wrote.  A
 * per_gather / per_gather_cast / per_update_* / per_td_priority_update call that passes exactly
 * that buffer is taken to work on THAT sample: its indices ascend (so duplicates are adjacent and
 * the update needs no arbitration pass) and, sharded, only the owned strata range is visited.
 * The contents must therefore be unchanged since the sample wrote them, and the buffer must stay
 * allocated until the update has been issued (a freed buffer's address can come back from the
 * allocator holding other indices).  The binding ends with the first update that uses it, with
 * per_add_batch, and with the next sample; any other index buffer takes the general path.  An
 * index buffer that is found not to ascend on the bound path sets PER_STATUS_BAD_INDEX. */
int per_sample_sharded(per_buffer *b, const double *d_records, const double *d_u, int64_t batch,
                       double beta, int64_t global_len, int64_t *d_out_idx, double *d_out_w64,
                       float *d_out_w32, void *stream);

/* Record exchange WITHOUT a collective.  Each rank owns a mailbox of 2 * n_shards * 64 bytes in
 * memory its peers can write (per_ipc_alloc in one process, per_ipc_open in the others; NVLink
 * peer access).  After per_shard_attach_exchange, per_sample_sharded may be called with
 * d_records == NULL: the sampling kernel's first warp then computes this shard's record, stores
 * it into every rank's mailbox (eight aligned 8-byte stores to peer-mapped addresses, each carrying 32 bits of the record and
 * the sample's epoch: no flag and no fence — a word that shows the awaited epoch is whole, in whatever order the words
 * arrive), and one warp per CTA polls until every word of all n_shards slots of this sample shows its epoch before the CTA descends
 * — compute and exchange in ONE kernel, no NCCL launch, capturable in a CUDA graph.  Slots are
 * double-buffered by sample parity (a rank can be at most one sample ahead of its slowest peer,
 * because its next sample needs that peer's next record).  A peer that never publishes does not
 * hang the kernel: after 2 s PER_STATUS_EXCHANGE_TIMEOUT is set and the sample proceeds.
 * h_peer_mailboxes: host array of n_shards device pointers, entry r = rank r's mailbox as mapped
 * in THIS process (entry `rank` = d_mailbox).  Every rank must call per_sample_sharded the same
 * number of times.  per_shard_publish publishes the record for the NEXT sample by itself (used to
 * emulate several ranks on one GPU, where one rank's kernel cannot wait for another's). */
int per_ipc_alloc(int device, int64_t bytes, void **d_ptr, uint8_t *h_handle64);
/* Learner-ward delivery (sampled transitions move peer-to-peer to the learner GPU; the reference
 * ships them as pickles over ZeroMQ, origin_repo/replay.py:38-45, origin_repo/learner.py:71-131).  The row columns get
 * there by passing peer-mapped output pointers to per_gather; this call does the same for the sample's scalars: for
 * every stratum this shard owns (d_idx[q] != -1) it stores d_idx[q] / d_w32[q] into the learner's full-batch arrays
 * d_dst_idx / d_dst_w32 (peer-mapped; either may be NULL).  Returned priorities need no call of their own: hand the
 * learner's (peer-mapped) priority array to per_update_priorities together with the sample's index buffer and the
 * update reads just the owned entries over NVLink. */
int per_shard_deliver(per_buffer *b, const int64_t *d_idx, const float *d_w32, int64_t batch, int64_t *d_dst_idx,
                      float *d_dst_w32, void *stream);
int per_ipc_open(int device, const uint8_t *h_handle64, void **d_ptr);
int per_ipc_close(int device, void *d_ptr);
int per_ipc_free(int device, void *d_ptr);
int per_shard_attach_exchange(per_buffer *b, double *d_mailbox, void *const *h_peer_mailboxes);
int per_shard_publish(per_buffer *b, int64_t global_len, void *stream);
/* Reads and clears {exchanges, total ns, max ns} that the sampling kernel's first CTA spent
 * between its start and the arrival of the last peer record (publish + NVLink + skew between
 * ranks).  Synchronises `stream`.  Diagnostics for profiles/. */
int per_shard_exchange_stats(per_buffer *b, int64_t *h_out3, void *stream);

/* Launch-shape knobs of the gather kernel, for the sweeps under profiles/ (not needed in
 * normal use; defaults are the measured best).  Keys: "gather_mode" (0 TMA bulk copies
 * through shared memory, 1 register-staged LDG/STG), "gather_warps", "gather_stages",
 * "gather_piece" (bytes, multiple of 16), "gather_ctas_per_sm"; "update_trace" (0/1, see
 * per_update_trace); "pdl" (bit mask, default 3, 7 on a shard of a sharded replay: 1 = sampling kernels, 2 = gather, 4 = the one-launch
 * update are launched with programmatic stream serialization and begin with
 * griddepcontrol.wait, so their launch and prologue overlap the tail of the kernel in front;
 * measured on B200 the cluster launch of the update loses from it, profiles/r01_notes.md);
 * "update_ctas" (cap on the thread-block cluster of the one-launch update: 0 = the library's choice, 8 CTAs, or 16 for
 * sorted batches of 1 025 .. 2 048 items; 16 needs the device's non-portable cluster size);
 * "sorted_update" (default 1: an update of exactly the index buffer the last sample wrote takes
 * the sorted fast path of update_kernels.cuh; 0 = always the level-synchronous rounds);
 * "dense_update_lmax" (default 0: such an update takes the dense streaming form only when it is too large for the
 * per-CTA climb; L = 12..26 sends every such update of a tree of up to 2^L leaves there, whatever its size);
 * "early_publish" (default 0; needs per_shard_attach_exchange): every priority update ends by
 * storing the refreshed shard record, tagged with the next exchange epoch, into all ranks'
 * mailboxes, so the record is on its way over NVLink while the update kernel retires and the next
 * per_sample_sharded only has to wait for it.  Contract while it is on, on EVERY rank alike:
 * between an update and the next sample nothing else may modify the tree or the replay length
 * (per_add_batch, a second update, per_shard_attach_record with another len fail loudly); ingest
 * goes between a sample and its update, or switch the knob off around it.
 * No reference counterpart. */
int per_tune(per_buffer *b, const char *key, int64_t value);

/* Diagnostics: after per_tune(b, "update_trace", 1) thread 0 of CTA 0 of the one-launch update kernel stores the GPU
 * global timer (ns) at its phase boundaries; this reads the 16 stamps of the last launch (0 where a phase did not
 * run): [0] past the launch wait, [1] TD done, [2] arbitration done, [3] leaves written / cluster barrier passed,
 * [4] sparse rounds done, [5] top rebuilt, and finer ones: [6] TD inputs staged, [7] CTA maximum known,
 * [8] cluster barrier of the TD phase passed, [9] batch maximum known, [10] sorted path: leaves and siblings in,
 * [11] sorted path: climb done, [12] dense top loaded, [13] dense top written out.  When the update runs as the TD
 * launch + k_upd_dense_sorted, the second kernel stamps [14] item runs found, [1] priorities ready, [10] leaves written,
 * [11] chunk lifted, [3] fenced (all CTA 0), then the CTA that finishes last: [12] elected, [5] top lifted, [13] end.
 * Synchronises. */
int per_update_trace(per_buffer *b, int64_t *h_out16, void *stream);
/* Diagnostics: after per_tune(b, "timeline", entries) thread 0 of CTA 0 of every sample / gather /
 * update launch stores, into the six entries that launch owns (plain stores, nothing waits for
 * them; a launch captured in a CUDA graph keeps its entries across replays, so after a run the
 * list holds the latest execution of each), (tag << 56) | GPU-global-timer ns stamps (tags: 1 sample entered, 10 shard
 * records in, 2 sample past its programmatic-launch wait, 11 owned strata known, 3 sample CTA 0
 * done; 4 gather entered, 5 past its wait, 6 CTA 0 done, 13 last issuing lane of the grid done; 7 update entered, 8 past its wait, 12 TD
 * arithmetic done, 9 CTA 0 done).  Copies up to max_entries non-empty entries to h_out (unordered:
 * sort by time), stores the number in h_count and clears the list.  Works inside CUDA-graph replays.  Synchronises `stream`. */
int per_timeline(per_buffer *b, int64_t *h_out, int64_t max_entries, int64_t *h_count, void *stream);

/* Copies one tree's node array (`_it_sum._value` / `_it_min._value`) to d_out[2*capacity]. */
int per_tree_read(per_buffer *b, int which, double *d_out, void *stream);
/* `_it_sum[idx]` / `_it_min[idx]` (memory.py:89-91, used at :293) and `_it_sum.sum(start, end)` /
 * `_it_min.min(start, end)` (memory.py:54-74, used at :244,289) on the buffer's own trees: one leaf
 * to *h_out (synchronises `stream`), or the range reduction with the reference's association to
 * d_out[0] (stream-ordered).  Same argument rules as seg_get / seg_reduce. */
int per_tree_get(per_buffer *b, int which, int64_t idx, double *h_out, void *stream);
int per_tree_reduce(per_buffer *b, int which, int64_t start, int64_t end, double *d_out, void *stream);

#ifdef __cplusplus
}
#endif
#endif /* PER_B200_H */
