/* b2c — B200-native gradient synchronisation: C-ABI boundary.
 *
 * This header is the drop-in boundary for the ONE data-parallel hot path of
 * ray_lightning (per-step gradient allreduce + initial parameter broadcast).
 * The reference owns no native code: each entry point below replaces a
 * *library call* the reference reaches through torch c10d / Horovod.  Citations
 * are `path:line` into the reference tree (ray_lightning/...) or, where the
 * arithmetic lives upstream, `torch:<file>:<line>` into the installed torch
 * 2.11 (see SURVEY.md §2a for the call-site inventory C1..C9).
 *
 * Conventions
 *   - plain C, no torch / C++ types; loadable with ctypes / cgo / JNI.
 *   - return 0 on success, negative B2C_E* library code, or a positive
 *     cudaError_t.  Never throws, never exits.  b2c_last_error() returns a
 *     thread-local human-readable string for the last failing call.
 *   - data calls (allreduce / broadcast / barrier) only ENQUEUE on the given
 *     stream; no host synchronisation.  create / import / destroy are
 *     synchronous and must be called by every rank of the group.
 *   - all ranks of a comm must issue the same sequence of data calls with the
 *     same sizes and parameters (the cross-GPU epoch counter is implicit).
 *   - the caller owns `buf` and keeps it alive until the stream passes the op;
 *     the library owns arena, flags and peer mappings.
 *   - calls on one comm must be stream-ordered (one stream at a time).
 */
#ifndef B2C_H_
#define B2C_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B2C_VERSION 100          /* 0.1.0 */
#define B2C_MAX_WORLD 8          /* one NVSwitch domain */
#define B2C_MAX_BLOCKS 256       /* independent flag channels per comm */
#define B2C_EXPORT_BLOB_BYTES 256

typedef struct b2c_comm b2c_comm;
typedef void* b2c_stream_t;      /* a cudaStream_t; NULL = legacy default stream */

/* element types (dtype = what lives in the caller's buffer,
 * wire_dtype = what crosses NVLink / sits in the arena) */
enum { B2C_F32 = 0, B2C_BF16 = 1, B2C_U8 = 2 };

/* algorithm selector for b2c_allreduce */
enum {
  B2C_ALGO_AUTO = 0,          /* one-shot below the crossover, two-shot above */
  B2C_ALGO_ONESHOT = 1,       /* every rank pulls all W copies and reduces locally */
  B2C_ALGO_TWOSHOT = 2,       /* reduce-scatter then all-gather (variant by param) */
  B2C_ALGO_NVLS = 3,          /* in-switch reduction: multimem.ld_reduce + multimem.st on the
                                 multicast arena (needs b2c_nvls_* set-up) */
};

/* keys for b2c_set_param / b2c_get_param */
enum {
  B2C_PARAM_CROSSOVER_BYTES = 0,  /* AUTO: wire bytes above which two-shot is used */
  B2C_PARAM_MAX_BLOCKS_ONESHOT = 1,
  B2C_PARAM_MAX_BLOCKS_TWOSHOT = 2,
  B2C_PARAM_THREADS = 3,          /* threads per CTA: 128..1024, multiple of 32 */
  B2C_PARAM_TIMEOUT_MS = 4,       /* bounded spin in the cross-GPU barrier */
  B2C_PARAM_TWOSHOT_VARIANT = 5,  /* 0 = pull-reduce + push (default), 1 = pull + pull */
  B2C_PARAM_MAX_BLOCKS_BCAST = 6,
  B2C_PARAM_ZEROCOPY = 7,         /* 1 (default): two-shot runs in place on registered / local-group
                                     buffers with no staging; 0: always stage through the arena */
  B2C_PARAM_ZEROCOPY_MIN_BYTES = 8, /* AUTO: at or above this size a registered buffer uses the
                                     zero-copy two-shot even below the one-shot crossover */
  B2C_PARAM_NVLS_MIN_BYTES = 9,   /* AUTO: wire bytes at or above which the NVLS path is used when
                                     the multicast arena is up (0 = never pick it automatically) */
  B2C_PARAM_MAX_BLOCKS_NVLS = 10,
  B2C_PARAM_UNROLL_NVLS = 11,     /* multimem.ld_reduce vectors in flight per thread: 2, 4, 8 or 16 */
  B2C_PARAM_UNROLL_ZC = 12,       /* zero-copy two-shot, W = 8 only: 0 (default 2), 1, 2 or 4 vectors per peer in flight */
  B2C_PARAM_PDL = 13,             /* 1 (default): the lean NVLS kernel is launched with programmatic stream
                                     serialization, so a train of bucket allreduces overlaps launch with drain */
  B2C_PARAM_THREADS_NVLS = 14,    /* lean NVLS kernel: 512 (default) or 256 threads per CTA (the 256-thread
                                     build stays within 64 registers and fits beside backward kernels) */
  B2C_PARAM_NVLS_STAGE_BYTES = 15, /* size of one of the two multicast staging slots, carved out of the arena
                                     on first use (tensors outside the arena, multicast broadcast) */
  B2C_PARAM_NVLS_LEAN = 16,       /* 1 (default): in-place NVLS uses the lean kernel (multicast flag barrier);
                                     0: the round-1 kernel (unicast flags), kept for A/B measurements */
  B2C_PARAM_BCAST_MC = 17,        /* 1 (default): with the multicast arena up b2c_broadcast goes through the
                                     switch (root sends its bytes once); 0: every rank pulls from the root */
  B2C_PARAM_MC_ACQ_POLL = 18,     /* multicast flag barrier: 1 (default) spin with ld.acquire.sys, 0 spin relaxed
                                     and fence once at the end (a measurement knob) */
  B2C_PARAM_BCAST_INPLACE_MIN_BYTES = 19, /* multicast broadcast of a tensor inside the arena: in place (two barriers,
                                     no copy) from this size up, through a staging slot (one barrier) below; default 1 MiB */
  B2C_PARAM__COUNT = 20
};

/* error codes */
enum {
  B2C_OK = 0,
  B2C_EINVAL = -1,     /* bad argument */
  B2C_ESTATE = -2,     /* call not valid in this state (e.g. data call before import) */
  B2C_ETOOBIG = -3,    /* request exceeds a stated capacity */
  B2C_ETIMEOUT = -4,   /* a peer did not arrive at the flag barrier in time (sticky) */
  B2C_ENOPEER = -5,    /* peer memory cannot be mapped (no P2P path / different node) */
  B2C_EUNSUPPORTED = -6
};

typedef struct b2c_stats {
  uint64_t calls_allreduce;
  uint64_t calls_broadcast;
  uint64_t calls_barrier;
  uint64_t launches;            /* kernels launched by this comm */
  uint64_t bytes_allreduce;     /* user bytes reduced */
  uint64_t bytes_broadcast;
  uint64_t oneshot_launches;
  uint64_t twoshot_launches;
  uint64_t zerocopy_launches;
  uint64_t nvls_launches;
  uint32_t last_algo;           /* B2C_ALGO_* actually used by the last allreduce */
  uint32_t last_blocks;         /* grid.x of the last launch */
  uint32_t last_threads;
  uint32_t device_status;       /* 0, or the sticky error code the kernels reported */
  uint32_t device_status_detail[3]; /* epoch, block, peer at failure */
  uint32_t reserved;
  uint64_t bcast_mc_launches;   /* broadcasts that went through the switch (multimem.st) */
  uint64_t noop_calls;          /* allreduce calls answered without a launch (world 1, scale 1) */
  uint64_t fused_steps;         /* b2c_fused_step calls */
} b2c_stats;

/* ---- library ---------------------------------------------------------- */

int b2c_version(void);
const char* b2c_last_error(void);

/* ---- communicator life cycle -------------------------------------------
 * Replaces the NCCL communicator that torch builds lazily behind
 * ray_lightning/ray_ddp.py:232-237 (init_process_group("nccl")) for the
 * gradient path only; the torch process group stays for control collectives.
 *
 * b2c_comm_create: cudaMalloc one symmetric region on `device`
 *   (flag page + 2 arena slots, `arena_bytes` total for the two slots).
 * b2c_comm_export: fill `blob` (B2C_EXPORT_BLOB_BYTES) with the CUDA IPC handle,
 *   GPU UUID, pid and geometry; `*len` is in/out.
 * b2c_comm_import: `blobs` = world blobs concatenated in rank order (own blob
 *   included); maps every peer with cudaIpcOpenMemHandle (or direct peer access
 *   when the peer lives in this process).  After import the comm is ready.
 * b2c_comm_destroy: closes peer mappings and frees the region.  Replaces what
 *   ray_ddp.py:109-121 (teardown) has to do before the actors are killed. */
int b2c_comm_create(int rank, int world, int device, size_t arena_bytes, b2c_comm** out);
int b2c_comm_export(b2c_comm* comm, void* blob, size_t* len);
int b2c_comm_import(b2c_comm* comm, const void* blobs, size_t len_each);
int b2c_comm_destroy(b2c_comm* comm);

/* Single-device group: `world` virtual ranks on ONE GPU, peers addressed
 * directly, every data call is ONE cooperative launch that runs all ranks'
 * CTAs side by side (grid.y = world).  Used to exercise the complete
 * cross-rank protocol (flags, epochs, double buffering) on a one-GPU box and
 * for the HBM-bound roofline at N=1.  No reference counterpart. */
int b2c_group_create_local(int world, int device, size_t arena_bytes, b2c_comm** out_comms);
int b2c_group_destroy_local(b2c_comm** comms, int world);

/* ---- data path ---------------------------------------------------------
 * b2c_allreduce: in place, buf[i] = (sum over ranks r=0..W-1, in rank order,
 *   fp32 accumulate) * scale, rounded once to `dtype`.
 *   Replaces C3 — the per-bucket allreduce the c10d Reducer issues behind
 *   ray_ddp.py:218-219 (torch:distributed/algorithms/ddp_comm_hooks/
 *   default_hooks.py:18-32, div by W then all_reduce SUM) — and C6, Horovod's
 *   allreduce_async_(op=Average) behind ray_horovod.py:159.
 *   (dtype, wire_dtype) in {(F32,F32), (BF16,BF16), (F32,BF16)}; the last is the
 *   bf16_compress_hook analogue (default_hooks.py:57-93): cast fused into the
 *   staging pass, result is bf16-representable and written back as fp32.
 * b2c_broadcast: bit-exact copy of root's `bytes` to every rank.  Replaces C2
 *   (_sync_module_states, torch:distributed/utils.py:289-328, reached from the
 *   DDP constructor behind ray_ddp.py:218-219), C4 (_sync_buffers) and C5
 *   (hvd.broadcast_parameters behind ray_horovod.py:158).
 * b2c_barrier: cross-GPU flag barrier on the stream.  Replaces C7 (hvd.join). */
int b2c_allreduce(b2c_comm* comm, void* buf, size_t count, int dtype, int wire_dtype,
                  float scale, int algo, b2c_stream_t stream);
int b2c_broadcast(b2c_comm* comm, void* buf, size_t bytes, int root, b2c_stream_t stream);
int b2c_barrier(b2c_comm* comm, b2c_stream_t stream);

/* the same three on a local (single-device) group: bufs[r] is rank r's buffer */
int b2c_group_allreduce(b2c_comm** comms, void** bufs, size_t count, int dtype, int wire_dtype,
                        float scale, int algo, b2c_stream_t stream);
int b2c_group_broadcast(b2c_comm** comms, void** bufs, size_t bytes, int root, b2c_stream_t stream);
int b2c_group_barrier(b2c_comm** comms, b2c_stream_t stream);

/* ---- registered (zero-copy) buffers --------------------------------------
 * A gradient bucket that lives in plain cudaMalloc memory (torch's caching allocator
 * without expandable_segments) can be mapped into every peer once; b2c_allreduce then
 * reduces it IN PLACE over NVLink with no staging pass through the arena.  Collective:
 * every rank exports its buffer, the blobs are swapped (rank order) and imported.
 * After that any b2c_allreduce whose [buf, buf+bytes) lies inside a registration — at the
 * SAME offset on every rank — takes the zero-copy two-shot path when it is 16-byte
 * aligned, (dtype == wire_dtype) and two-shot is selected.  No reference counterpart (NCCL
 * user-buffer registration is the closest upstream analogue). */
#define B2C_REG_BLOB_BYTES 128
int b2c_register_export(b2c_comm* comm, void* ptr, size_t bytes, void* blob, size_t* len);
int b2c_register_import(b2c_comm* comm, void* ptr, size_t bytes, const void* blobs, size_t len_each);
int b2c_deregister(b2c_comm* comm, void* ptr);

/* ---- NVLS (NVLink SHARP) multicast arena ---------------------------------
 * A second symmetric arena, allocated with cuMemCreate and bound to an NVLink multicast
 * object, so that the switch does the reduction (multimem.ld_reduce) and the replication
 * (multimem.st): each GPU moves ~S bytes per direction instead of 2(W-1)/W*S, which is the
 * only way past the peer-to-peer ceiling for the large DDP buckets.  Optional: without it
 * every other path keeps working.  Collective, three phases with a host barrier between
 * import and bind (every device must have been added before any memory is bound):
 *   b2c_nvls_export  allocate local memory; rank 0 creates the multicast object; fill blob
 *   b2c_nvls_import  blobs in rank order: adopt rank 0's multicast handle, add this device
 *   -- host barrier --
 *   b2c_nvls_bind    bind + map (unicast and multicast views); -- host barrier -- ready
 * b2c_nvls_arena returns the local unicast view: tensors placed inside it are reduced in
 * place (no staging) when passed to b2c_allreduce with algo NVLS/AUTO at the same offset on
 * every rank (16-byte aligned; a ragged tail of less than one 16-byte vector is reduced by a
 * second, staged launch so that bytes after the tensor are never touched).  Arithmetic: the switch adds in fp32 in its own order and the result is scaled
 * afterwards (sum * scale): within the stated tolerances of the reference path, bitwise
 * identical across replicas, but NOT bit-identical to the rank-order contract of the other
 * algorithms.  No reference counterpart (NCCL's NVLS algorithm is the upstream analogue). */
#define B2C_NVLS_BLOB_BYTES 64
int b2c_nvls_export(b2c_comm* comm, size_t bytes, void* blob, size_t* len);
int b2c_nvls_import(b2c_comm* comm, const void* blobs, size_t len_each);
int b2c_nvls_bind(b2c_comm* comm);
int b2c_nvls_arena(b2c_comm* comm, void** base, size_t* bytes);
/* Allocation inside the arena: first fit with coalescing, 256-byte aligned, deterministic — ranks that
 * issue the same sequence of b2c_nvls_alloc / b2c_nvls_free calls get the same offsets, which is what
 * the in-place kernels need.  The first 64 KiB of the arena hold the multicast flag counters.
 * b2c_nvls_pool_bind makes this comm's arena the backing store of b2c_pool_alloc / b2c_pool_free,
 * which have the signature torch.cuda.memory.CUDAPluggableAllocator expects: tensors allocated inside a
 * torch.cuda.MemPool on that allocator live INSIDE the multicast arena, so the per-step allreduce is
 * the in-place NVLS kernel with no staging at all. */
int b2c_nvls_alloc(b2c_comm* comm, size_t bytes, void** ptr);
int b2c_nvls_free(b2c_comm* comm, void* ptr);
int b2c_nvls_pool_bind(b2c_comm* comm);
void* b2c_pool_alloc(ptrdiff_t size, int device, b2c_stream_t stream);
void b2c_pool_free(void* ptr, ptrdiff_t size, int device, b2c_stream_t stream);

/* ---- fused optimizer step -------------------------------------------------
 * the step on the far side of the path.  What the reference reaches after the gradient
 * allreduce — optimizer.step() inside the PTL train loop entered at ray_ddp.py:218-219 / ray_horovod.py:159 —
 * is folded into the reduction: the rank that reduces shard r of a gradient bucket applies the optimizer to
 * shard r (fp32 state for the owned shard only: each rank holds 1/W of the momentum / Adam moments) and the
 * updated WEIGHTS are written into every rank's parameter bucket.  Arithmetic follows torch.optim.SGD / Adam /
 * AdamW operation by operation in fp32 (the gradient average is the allreduce contract above).
 *   grad / param : flat fp32 buckets of `count` elements with the same layout, both inside the multicast arena
 *                  (in-switch path) or both registered with every peer (peer-to-peer path), 16-byte aligned,
 *                  count % 4 == 0 when world > 1.  `grad` is left untouched; `param` is updated on every rank.
 *   state1/2     : fp32, b2c_fused_shard() elements each, this rank's shard only, zero-initialised by the caller
 *                  (SGD: momentum buffer / unused; Adam: exp_avg / exp_avg_sq).
 *   bcast_*      : optional C4 fold (multicast arena only): `bcast_bytes` (multiple of 16) at `bcast_buf` inside
 *                  the arena are replicated from rank `bcast_root` in the same launch — the per-forward module
 *                  buffer broadcast (_sync_buffers) without a launch of its own.
 * No reference counterpart as an entry point (the reference calls the optimizer in Python). */
enum { B2C_OPT_KIND_SGD = 1, B2C_OPT_KIND_ADAM = 2, B2C_OPT_KIND_ADAMW = 3 };
typedef struct b2c_opt {
  int kind;                 /* B2C_OPT_KIND_* */
  int nesterov;             /* SGD */
  int first_step;           /* SGD with momentum: 1 on the first step (buf = grad, as torch creates it) */
  /* hyper-parameters in double, as the Python optimizers hold them: every derived scalar (1 - beta2, lr /
   * bias_correction1, ...) is formed in double and rounded to float once, like torch's own kernels get them */
  double lr, momentum, dampening, weight_decay;
  double beta1, beta2, eps; /* Adam / AdamW */
  double bias_correction1;  /* 1 - beta1^t */
  double bias_correction2;  /* 1 - beta2^t */
} b2c_opt;
int b2c_fused_shard(size_t count, int world, int rank, size_t* first_elem, size_t* n_elems);
int b2c_fused_step(b2c_comm* comm, void* grad, void* param, size_t count, int dtype, float scale, const b2c_opt* opt,
                   void* state1, void* state2, void* bcast_buf, size_t bcast_bytes, int bcast_root, b2c_stream_t stream);
int b2c_group_fused_step(b2c_comm** comms, void** grads, void** params, size_t count, int dtype, float scale,
                         const b2c_opt* opt, void** state1, void** state2, b2c_stream_t stream);

/* ---- allocation-free variant ----------------------------------------------
 * No counterpart in the reference.
 * b2c_allreduce_local_noalloc: the one-shot allreduce of a single-device group
 * (b2c_group_allreduce with B2C_ALGO_ONESHOT), for callers that must not allocate device
 * memory where they need the reduction: code that runs after the training framework's
 * caching allocator has claimed the GPU (a cudaMalloc there can fail, or make the framework
 * flush its cache), and code on the autograd hook thread in the middle of backward, where a
 * cudaMalloc / cudaFree would synchronise the whole device and stall the overlapped step.
 * It needs no b2c_comm: flags and staging slots live in a file-scope device array inside
 * libb2c.so (about 2.1 MiB, held by every GPU that loads the library until the process
 * exits).  bufs[r] (r < world <= 8) are `count` elements of `dtype` (B2C_F32 or B2C_BF16) on
 * the CURRENT device; each ends up holding sum_r bufs[r] * scale, same arithmetic contract
 * as b2c_allreduce.  Limit: count * element size <= B2C_NOALLOC_MAX_BYTES per buffer;
 * anything larger is refused with B2C_ETOOBIG.  The call enqueues a few small memsets (the
 * flag / counter words are cleared every time, because the array keeps whatever the last
 * call left) and one cooperative kernel on `stream`, and calls no allocating routine.
 * The first call on a device looks up the array's device address; make that first call
 * before starting any stream capture.  All calls on a device share the one array: two calls
 * that can overlap on the GPU (different streams, no dependency) race with each other. */
#define B2C_NOALLOC_MAX_BYTES (256 * 1024)
int b2c_allreduce_local_noalloc(void* const* bufs, int world, size_t count, int dtype, float scale,
                                b2c_stream_t stream);
/* The call above only enqueues, so it cannot itself report that its kernel gave up waiting in the flag
 * barrier (2 s; with a cooperative launch this means the GPU was taken away mid-kernel).  The kernel
 * records that in four sticky words next to the scratch array — error code, epoch, channel, virtual
 * rank waited for — which no later call wipes.  b2c_noalloc_status waits for the current device, copies
 * them to out[4] (may be NULL), returns B2C_OK when they are clear and B2C_ETIMEOUT otherwise (the
 * buffers of the failing call were left unreduced); reset != 0 clears them.  Allocates nothing.
 * No counterpart in the reference. */
int b2c_noalloc_status(uint32_t out[4], int reset);

/* ---- knobs, counters, introspection ------------------------------------ */
int b2c_set_param(b2c_comm* comm, int key, long value);
int b2c_get_param(b2c_comm* comm, int key, long* value);
int b2c_get_stats(b2c_comm* comm, b2c_stats* out);
/* base address / size of one arena slot (device pointer, local rank) */
int b2c_comm_arena(b2c_comm* comm, void** base, size_t* slot_bytes);
int b2c_comm_info(b2c_comm* comm, int* rank, int* world, int* device);
/* Device-side phase clock of the LAST kernel of this comm (channel 0, %globaltimer ns):
 * out[0] kernel start, out[1] passed barrier A, out[2] passed barrier B, out[3] end.
 * Synchronises the device; a profiling aid, not for the hot path. */
int b2c_get_phase_ns(b2c_comm* comm, uint64_t out[4]);
/* The same four stamps for the last (up to 32) kernels of this comm, oldest first, 4 x u64 each, so a
 * whole training step can be attributed bucket by bucket afterwards: start->A is the wait for the
 * slowest rank, A->B the data phase.  Returns the number of entries written (<= max_entries) or a
 * negative error; *total = kernels stamped since the comm was created.  Synchronises the device. */
int b2c_get_phase_ring(b2c_comm* comm, uint64_t* out, int max_entries, uint32_t* total);

#ifdef __cplusplus
}
#endif
#endif /* B2C_H_ */
