// b2c host side: the C-ABI declared in include/b2c.h.
// Plain CUDA runtime only — no torch, no NCCL, no C++ types across the boundary.
#include <stdio.h>
#include <stdarg.h>
#include <string.h>
#include <stdlib.h>
#include <unistd.h>
#include <sys/syscall.h>
#include <errno.h>
#include <pthread.h>
#include <math.h>
#include <cuda.h>   // types only: the driver API is reached through cudaGetDriverEntryPoint (no -lcuda)

#include "b2c_kernels.cuh"

// ---------------------------------------------------------------- errors
static thread_local char g_err[512] = "";

static int fail(int code, const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
  return code;
}
static int cuda_fail(cudaError_t e, const char* what) {
  snprintf(g_err, sizeof(g_err), "%s: %s (%s)", what, cudaGetErrorName(e), cudaGetErrorString(e));
  return (int)e;
}
#define CU(call)                                         \
  do {                                                   \
    cudaError_t e_ = (call);                             \
    if (e_ != cudaSuccess) return cuda_fail(e_, #call);  \
  } while (0)

struct DeviceGuard {
  int prev = -1;
  bool ok = true;
  explicit DeviceGuard(int dev) {
    if (cudaGetDevice(&prev) != cudaSuccess) prev = -1;
    if (prev != dev) ok = (cudaSetDevice(dev) == cudaSuccess);
  }
  ~DeviceGuard() {
    int cur = -1;
    if (prev >= 0 && cudaGetDevice(&cur) == cudaSuccess && cur != prev) cudaSetDevice(prev);
  }
};

// ---------------------------------------------------------------- comm
#define B2C_MAGIC 0x62326331u /* "b2c1" */

struct ExportBlob {
  uint32_t magic;
  uint32_t version;
  int32_t rank;
  int32_t world;
  int64_t pid;
  int32_t device;       // CUDA ordinal in the exporting process
  int32_t pad0;
  uint64_t region_bytes;
  uint64_t slot_bytes;
  uint64_t raw_ptr;     // valid only inside the exporting process
  uint64_t host_id;     // gethostid ^ boot-id hash: IPC handles do not cross nodes
  unsigned char uuid[16];
  cudaIpcMemHandle_t handle;  // 64 bytes
};
static_assert(sizeof(ExportBlob) <= B2C_EXPORT_BLOB_BYTES, "blob too large");

struct b2c_comm {
  int rank, world, device;
  size_t arena_bytes, slot_bytes, region_bytes;
  char* region;                      // local allocation
  char* peer_region[B2C_MAX_WORLD];  // every rank's region as mapped here ([rank] == region)
  bool peer_ipc[B2C_MAX_WORLD];      // opened with cudaIpcOpenMemHandle (must be closed)
  uint32_t* status_h;                // pinned, mapped: sticky device-side error words
  uint32_t* status_d;
  bool ready;
  bool local_group;
  b2c_comm** group;                  // local group: all members (owned by member 0)
  long params[B2C_PARAM__COUNT];
  b2c_stats stats;
  struct Registration* regs;         // registered caller buffers (zero-copy path)
  int nregs, regs_cap;
  struct OpenedHandle* opened;       // IPC base allocations already mapped (one open per handle)
  int nopened, opened_cap;
  struct Nvls* nvls;                 // multicast arena (optional)
};

#define B2C_NVLS_MAX_BLOCKS 512
struct ArenaBlock { size_t off, size; bool used; };
struct Nvls {
  int phase;                              // 1 exported, 2 imported, 3 bound (usable)
  size_t bytes;
  size_t slot_bytes, slot_off;            // staging slots (2 x slot_bytes at slot_off), carved out on first use
  CUmemGenericAllocationHandle mem, mc;
  int mc_fd;                              // rank 0: exported fd of the multicast object
  CUdeviceptr uc, mcva;
  bool uc_mapped, mc_mapped, bound;
  // first-fit allocator with coalescing over [B2C_MC_HEAD_BYTES, bytes): address ordered block list.
  // Deterministic: ranks that issue the same alloc/free sequence get the same offsets.
  ArenaBlock blocks[B2C_NVLS_MAX_BLOCKS];
  int nblocks;
};
struct Registration {
  char* local;                       // caller pointer on this rank
  size_t bytes;
  char* peer[B2C_MAX_WORLD];         // the same buffer of rank q, mapped here ([rank] == local)
};
struct OpenedHandle {
  cudaIpcMemHandle_t handle;
  char* base;
  int refs;
};
struct RegBlob {
  uint32_t magic;
  int32_t rank;
  int64_t pid;
  int32_t device;
  int32_t pad0;
  uint64_t raw_ptr;
  uint64_t offset;                   // ptr - base of its cudaMalloc allocation
  uint64_t bytes;
  cudaIpcMemHandle_t handle;         // of the base allocation
};
static_assert(sizeof(RegBlob) <= B2C_REG_BLOB_BYTES, "reg blob too large");

static void nvls_release(b2c_comm* c);
static int nvls_stage_slots(b2c_comm* c);
static pthread_mutex_t g_pool_lock = PTHREAD_MUTEX_INITIALIZER;
static b2c_comm* g_pool_comm = nullptr;  // whose multicast arena backs b2c_pool_alloc / b2c_pool_free

static uint64_t host_identity() {
  uint64_t h = (uint64_t)gethostid();
  FILE* f = fopen("/proc/sys/kernel/random/boot_id", "r");
  if (f) {
    char buf[64] = {0};
    if (fgets(buf, sizeof(buf), f))
      for (const char* p = buf; *p; ++p) h = h * 1099511628211ull ^ (unsigned char)*p;
    fclose(f);
  }
  return h;
}

static void default_params(b2c_comm* c) {
  c->params[B2C_PARAM_CROSSOVER_BYTES] = 512 * 1024;
  c->params[B2C_PARAM_MAX_BLOCKS_ONESHOT] = 32;
  c->params[B2C_PARAM_MAX_BLOCKS_TWOSHOT] = 64;
  c->params[B2C_PARAM_THREADS] = 512;
  c->params[B2C_PARAM_TIMEOUT_MS] = 10000;
  c->params[B2C_PARAM_TWOSHOT_VARIANT] = 0;
  c->params[B2C_PARAM_MAX_BLOCKS_BCAST] = 32;
  c->params[B2C_PARAM_ZEROCOPY] = 1;
  c->params[B2C_PARAM_ZEROCOPY_MIN_BYTES] = 256 * 1024;
  c->params[B2C_PARAM_NVLS_MIN_BYTES] = 512 * 1024;  // below this one barrier (staged one-shot) beats two
  c->params[B2C_PARAM_MAX_BLOCKS_NVLS] = 8;   // the switch saturates with 4-8 CTAs; more only adds congestion
  c->params[B2C_PARAM_UNROLL_NVLS] = 8;
  c->params[B2C_PARAM_UNROLL_ZC] = 0;
  c->params[B2C_PARAM_PDL] = 1;
  c->params[B2C_PARAM_THREADS_NVLS] = 512;
  c->params[B2C_PARAM_NVLS_LEAN] = 1;
  c->params[B2C_PARAM_BCAST_MC] = 1;
  c->params[B2C_PARAM_MC_ACQ_POLL] = 1;
  c->params[B2C_PARAM_BCAST_INPLACE_MIN_BYTES] = 1l << 20;
  c->params[B2C_PARAM_NVLS_STAGE_BYTES] = 32l << 20;
}

static int comm_alloc(int rank, int world, int device, size_t arena_bytes, b2c_comm** out) {
  if (!out) return fail(B2C_EINVAL, "b2c_comm_create: out is NULL");
  *out = nullptr;
  if (world < 1 || world > B2C_MAX_WORLD)
    return fail(B2C_EINVAL, "b2c_comm_create: world=%d outside 1..%d (one NVSwitch domain)", world, B2C_MAX_WORLD);
  if (rank < 0 || rank >= world) return fail(B2C_EINVAL, "b2c_comm_create: rank=%d outside 0..%d", rank, world - 1);
  if (arena_bytes < 2 * 4096) return fail(B2C_EINVAL, "b2c_comm_create: arena_bytes=%zu too small (min 8192)", arena_bytes);
  DeviceGuard g(device);
  if (!g.ok) return fail(B2C_EINVAL, "b2c_comm_create: cannot select CUDA device %d", device);
  b2c_comm* c = (b2c_comm*)calloc(1, sizeof(b2c_comm));
  if (!c) return fail(B2C_EINVAL, "out of host memory");
  c->rank = rank; c->world = world; c->device = device;
  c->slot_bytes = (arena_bytes / 2) & ~(size_t)255;
  c->arena_bytes = 2 * c->slot_bytes;
  c->region_bytes = B2C_HEADER_BYTES + c->arena_bytes;
  default_params(c);
  cudaError_t e = cudaMalloc((void**)&c->region, c->region_bytes);
  if (e != cudaSuccess) { free(c); return cuda_fail(e, "cudaMalloc(region)"); }
  e = cudaMemset(c->region, 0, B2C_HEADER_BYTES);
  if (e == cudaSuccess) e = cudaHostAlloc((void**)&c->status_h, 64, cudaHostAllocMapped);
  if (e == cudaSuccess) { memset(c->status_h, 0, 64); e = cudaHostGetDevicePointer((void**)&c->status_d, c->status_h, 0); }
  if (e == cudaSuccess) e = cudaDeviceSynchronize();
  if (e != cudaSuccess) {
    cudaFree(c->region);
    if (c->status_h) cudaFreeHost(c->status_h);
    free(c);
    return cuda_fail(e, "b2c_comm_create: region set-up");
  }
  c->peer_region[rank] = c->region;
  *out = c;
  return B2C_OK;
}

extern "C" int b2c_version(void) { return B2C_VERSION; }
extern "C" const char* b2c_last_error(void) { return g_err; }

extern "C" int b2c_comm_create(int rank, int world, int device, size_t arena_bytes, b2c_comm** out) {
  int rc = comm_alloc(rank, world, device, arena_bytes, out);
  if (rc == B2C_OK && world == 1) (*out)->ready = true;  // nothing to import
  return rc;
}

extern "C" int b2c_comm_export(b2c_comm* c, void* blob, size_t* len) {
  if (!c || !blob || !len) return fail(B2C_EINVAL, "b2c_comm_export: NULL argument");
  if (*len < B2C_EXPORT_BLOB_BYTES) return fail(B2C_EINVAL, "b2c_comm_export: blob needs %d bytes, got %zu", B2C_EXPORT_BLOB_BYTES, *len);
  if (c->local_group) return fail(B2C_ESTATE, "b2c_comm_export: local groups are not exportable");
  DeviceGuard g(c->device);
  ExportBlob b;
  memset(&b, 0, sizeof(b));
  b.magic = B2C_MAGIC; b.version = B2C_VERSION;
  b.rank = c->rank; b.world = c->world; b.pid = (int64_t)getpid(); b.device = c->device;
  b.region_bytes = c->region_bytes; b.slot_bytes = c->slot_bytes;
  b.raw_ptr = (uint64_t)(uintptr_t)c->region;
  b.host_id = host_identity();
  cudaDeviceProp prop;
  CU(cudaGetDeviceProperties(&prop, c->device));
  memcpy(b.uuid, &prop.uuid, 16);
  CU(cudaIpcGetMemHandle(&b.handle, c->region));
  memset(blob, 0, B2C_EXPORT_BLOB_BYTES);
  memcpy(blob, &b, sizeof(b));
  *len = B2C_EXPORT_BLOB_BYTES;
  return B2C_OK;
}

extern "C" int b2c_comm_import(b2c_comm* c, const void* blobs, size_t len_each) {
  if (!c || !blobs) return fail(B2C_EINVAL, "b2c_comm_import: NULL argument");
  if (len_each < sizeof(ExportBlob)) return fail(B2C_EINVAL, "b2c_comm_import: len_each=%zu too small", len_each);
  if (c->local_group) return fail(B2C_ESTATE, "b2c_comm_import: local groups need no import");
  if (c->ready && c->world > 1) return fail(B2C_ESTATE, "b2c_comm_import: already imported");
  DeviceGuard g(c->device);
  const uint64_t me_host = host_identity();
  for (int q = 0; q < c->world; ++q) {
    ExportBlob b;
    memcpy(&b, (const char*)blobs + (size_t)q * len_each, sizeof(b));
    if (b.magic != B2C_MAGIC || b.version != B2C_VERSION)
      return fail(B2C_EINVAL, "b2c_comm_import: blob %d has bad magic/version", q);
    if (b.rank != q || b.world != c->world)
      return fail(B2C_EINVAL, "b2c_comm_import: blob %d is rank %d of %d; blobs must be in rank order", q, b.rank, b.world);
    if (b.slot_bytes != c->slot_bytes)
      return fail(B2C_EINVAL, "b2c_comm_import: rank %d arena slot %llu B != ours %zu B (arena must be symmetric)", q, (unsigned long long)b.slot_bytes, c->slot_bytes);
    if (q == c->rank) continue;
    if (b.host_id != me_host)
      return fail(B2C_ENOPEER, "b2c_comm_import: rank %d is on another node; b2c is single NVSwitch domain only", q);
    if (b.pid == (int64_t)getpid()) {
      // peer lives in this process (one process driving several GPUs): direct peer access
      if (b.device != c->device) {
        int can = 0;
        CU(cudaDeviceCanAccessPeer(&can, c->device, b.device));
        if (!can) return fail(B2C_ENOPEER, "b2c_comm_import: device %d cannot access peer device %d", c->device, b.device);
        cudaError_t e = cudaDeviceEnablePeerAccess(b.device, 0);
        if (e == cudaErrorPeerAccessAlreadyEnabled) cudaGetLastError();
        else if (e != cudaSuccess) return cuda_fail(e, "cudaDeviceEnablePeerAccess");
      }
      c->peer_region[q] = (char*)(uintptr_t)b.raw_ptr;
      c->peer_ipc[q] = false;
    } else {
      void* p = nullptr;
      cudaError_t e = cudaIpcOpenMemHandle(&p, b.handle, cudaIpcMemLazyEnablePeerAccess);
      if (e != cudaSuccess) {
        cudaGetLastError();
        snprintf(g_err, sizeof(g_err), "b2c_comm_import: cudaIpcOpenMemHandle(rank %d): %s (%s)", q, cudaGetErrorName(e), cudaGetErrorString(e));
        return (int)e;
      }
      c->peer_region[q] = (char*)p;
      c->peer_ipc[q] = true;
    }
  }
  c->ready = true;
  return B2C_OK;
}

extern "C" int b2c_comm_destroy(b2c_comm* c) {
  if (!c) return B2C_OK;
  if (c->local_group) return fail(B2C_ESTATE, "b2c_comm_destroy: member of a local group; use b2c_group_destroy_local");
  DeviceGuard g(c->device);
  cudaDeviceSynchronize();
  int rc = B2C_OK;
  for (int q = 0; q < c->world; ++q)
    if (q != c->rank && c->peer_ipc[q] && c->peer_region[q]) {
      cudaError_t e = cudaIpcCloseMemHandle(c->peer_region[q]);
      if (e != cudaSuccess && rc == B2C_OK) rc = cuda_fail(e, "cudaIpcCloseMemHandle");
    }
  nvls_release(c);
  for (int i = 0; i < c->nopened; ++i)
    if (c->opened[i].base) cudaIpcCloseMemHandle(c->opened[i].base);
  free(c->opened);
  free(c->regs);
  if (c->region) cudaFree(c->region);
  if (c->status_h) cudaFreeHost(c->status_h);
  free(c);
  return rc;
}

extern "C" int b2c_group_create_local(int world, int device, size_t arena_bytes, b2c_comm** out) {
  if (!out) return fail(B2C_EINVAL, "b2c_group_create_local: out is NULL");
  if (world < 1 || world > B2C_MAX_WORLD) return fail(B2C_EINVAL, "b2c_group_create_local: world=%d outside 1..%d", world, B2C_MAX_WORLD);
  b2c_comm** members = (b2c_comm**)calloc(world, sizeof(b2c_comm*));
  for (int r = 0; r < world; ++r) {
    int rc = comm_alloc(r, world, device, arena_bytes, &members[r]);
    if (rc != B2C_OK) {
      for (int k = 0; k < r; ++k) { members[k]->local_group = false; b2c_comm_destroy(members[k]); }
      free(members);
      return rc;
    }
  }
  for (int r = 0; r < world; ++r) {
    for (int q = 0; q < world; ++q) members[r]->peer_region[q] = members[q]->region;
    members[r]->local_group = true;
    members[r]->group = members;
    members[r]->ready = true;
    out[r] = members[r];
  }
  return B2C_OK;
}

extern "C" int b2c_group_destroy_local(b2c_comm** comms, int world) {
  if (!comms || world < 1) return B2C_OK;
  b2c_comm** members = comms[0] ? comms[0]->group : nullptr;
  for (int r = 0; r < world; ++r)
    if (comms[r]) { comms[r]->local_group = false; b2c_comm_destroy(comms[r]); comms[r] = nullptr; }
  free(members);
  return B2C_OK;
}

// ---------------------------------------------------------------- launch helpers
template <int MODE> static const void* oneshot_fn(int W) {
  switch (W) {
    case 1: return (const void*)k_allreduce_oneshot<MODE, 1>;
    case 2: return (const void*)k_allreduce_oneshot<MODE, 2>;
    case 3: return (const void*)k_allreduce_oneshot<MODE, 3>;
    case 4: return (const void*)k_allreduce_oneshot<MODE, 4>;
    case 5: return (const void*)k_allreduce_oneshot<MODE, 5>;
    case 6: return (const void*)k_allreduce_oneshot<MODE, 6>;
    case 7: return (const void*)k_allreduce_oneshot<MODE, 7>;
    default: return (const void*)k_allreduce_oneshot<MODE, 8>;
  }
}
template <int MODE> static const void* twoshot_fn(int W) {
  switch (W) {
    case 1: return (const void*)k_allreduce_twoshot<MODE, 1>;
    case 2: return (const void*)k_allreduce_twoshot<MODE, 2>;
    case 3: return (const void*)k_allreduce_twoshot<MODE, 3>;
    case 4: return (const void*)k_allreduce_twoshot<MODE, 4>;
    case 5: return (const void*)k_allreduce_twoshot<MODE, 5>;
    case 6: return (const void*)k_allreduce_twoshot<MODE, 6>;
    case 7: return (const void*)k_allreduce_twoshot<MODE, 7>;
    default: return (const void*)k_allreduce_twoshot<MODE, 8>;
  }
}

template <int MODE> static const void* twoshot_zc_fn(int W) {
  switch (W) {
    case 1: return (const void*)k_allreduce_twoshot_zc<MODE, 1>;
    case 2: return (const void*)k_allreduce_twoshot_zc<MODE, 2>;
    case 3: return (const void*)k_allreduce_twoshot_zc<MODE, 3>;
    case 4: return (const void*)k_allreduce_twoshot_zc<MODE, 4>;
    case 5: return (const void*)k_allreduce_twoshot_zc<MODE, 5>;
    case 6: return (const void*)k_allreduce_twoshot_zc<MODE, 6>;
    case 7: return (const void*)k_allreduce_twoshot_zc<MODE, 7>;
    default: return (const void*)k_allreduce_twoshot_zc<MODE, 8>;
  }
}

static const void* nvls_fn(int mode, int unroll) {
  if (mode == MODE_F32) {
    switch (unroll) {
      case 2: return (const void*)k_allreduce_nvls<MODE_F32, 2>;
      case 4: return (const void*)k_allreduce_nvls<MODE_F32, 4>;
      case 16: return (const void*)k_allreduce_nvls<MODE_F32, 16>;
      default: return (const void*)k_allreduce_nvls<MODE_F32, 8>;
    }
  }
  switch (unroll) {
    case 2: return (const void*)k_allreduce_nvls<MODE_BF16, 2>;
    case 4: return (const void*)k_allreduce_nvls<MODE_BF16, 4>;
    case 16: return (const void*)k_allreduce_nvls<MODE_BF16, 16>;
    default: return (const void*)k_allreduce_nvls<MODE_BF16, 8>;
  }
}
// lean in-place NVLS kernel: (threads, vectors in flight per thread).  256-thread variants are built for
// <= 64 registers so that a CTA fits on an SM beside the backward kernels of a training step.
static const void* nvls_lean_fn(int mode, int threads, int unroll) {
  if (mode == MODE_F32) {
    if (threads <= 256) return unroll <= 4 ? (const void*)k_nvls_inplace<MODE_F32, 4, 256, 4> : (const void*)k_nvls_inplace<MODE_F32, 8, 256, 4>;
    return unroll <= 4 ? (const void*)k_nvls_inplace<MODE_F32, 4, 512, 1> : (const void*)k_nvls_inplace<MODE_F32, 8, 512, 1>;
  }
  if (threads <= 256) return unroll <= 4 ? (const void*)k_nvls_inplace<MODE_BF16, 4, 256, 4> : (const void*)k_nvls_inplace<MODE_BF16, 8, 256, 4>;
  return unroll <= 4 ? (const void*)k_nvls_inplace<MODE_BF16, 4, 512, 1> : (const void*)k_nvls_inplace<MODE_BF16, 8, 512, 1>;
}

static const void* zc8_fn(int mode, int unroll) {
  if (mode == MODE_F32) {
    switch (unroll) {
      case 1: return (const void*)k_allreduce_twoshot_zc<MODE_F32, 8, 1>;
      case 4: return (const void*)k_allreduce_twoshot_zc<MODE_F32, 8, 4>;
      default: return (const void*)k_allreduce_twoshot_zc<MODE_F32, 8, 2>;
    }
  }
  switch (unroll) {
    case 1: return (const void*)k_allreduce_twoshot_zc<MODE_BF16, 8, 1>;
    case 4: return (const void*)k_allreduce_twoshot_zc<MODE_BF16, 8, 4>;
    default: return (const void*)k_allreduce_twoshot_zc<MODE_BF16, 8, 2>;
  }
}

// registration covering [p, p+bytes), or NULL
// An exact (pointer, size) match wins; otherwise the NEWEST registration that contains the range (an older
// one may describe memory that has since been freed and handed out again).
static Registration* find_reg(b2c_comm* c, const void* p, size_t bytes) {
  const char* q = (const char*)p;
  for (int i = c->nregs - 1; i >= 0; --i)
    if (q == c->regs[i].local && bytes == c->regs[i].bytes) return &c->regs[i];
  for (int i = c->nregs - 1; i >= 0; --i)
    if (q >= c->regs[i].local && q + bytes <= c->regs[i].local + c->regs[i].bytes) return &c->regs[i];
  return nullptr;
}

static int check_sticky(b2c_comm* c) {
  if (c->status_h && c->status_h[0] != 0)
    return fail(-(int)c->status_h[0], "b2c: comm is dead: flag barrier timed out earlier (epoch %u, channel %u, waiting for rank %u)",
                c->status_h[1], c->status_h[2], c->status_h[3]);
  return B2C_OK;
}

// geometry: identical on every rank because it depends on (size, params) only
static void geometry(const b2c_comm* c, size_t nvec_span, long max_blocks, int* nblocks, int* threads, uint32_t* chunk_vecs) {
  long th = c->params[B2C_PARAM_THREADS];
  if (max_blocks < 1) max_blocks = 1;
  if (max_blocks > B2C_MAX_BLOCKS) max_blocks = B2C_MAX_BLOCKS;
  size_t per_block_min = (size_t)th;
  size_t cu = (nvec_span + per_block_min * max_blocks - 1) / (per_block_min * max_blocks);
  if (cu < 1) cu = 1;
  if (cu > 8) cu = 8;
  size_t cv = per_block_min * cu;
  size_t nb = (nvec_span + cv - 1) / cv;
  if (nb < 1) nb = 1;
  if (nb > (size_t)max_blocks) nb = max_blocks;
  *nblocks = (int)nb; *threads = (int)th; *chunk_vecs = (uint32_t)cv;
}

// W == 1 only: not limited by the flag-channel count
static void geometry_wide(const b2c_comm* c, size_t nvec, long max_blocks, int* nblocks, int* threads, uint32_t* chunk_vecs) {
  const size_t th = (size_t)c->params[B2C_PARAM_THREADS];
  size_t cu = (nvec + th * max_blocks - 1) / (th * max_blocks);
  if (cu < 1) cu = 1;
  if (cu > 8) cu = 8;
  size_t nb = (nvec + th * cu - 1) / (th * cu);
  if (nb < 1) nb = 1;
  if (nb > (size_t)max_blocks) nb = max_blocks;
  *nblocks = (int)nb; *threads = (int)th; *chunk_vecs = (uint32_t)(th * cu);
}

// local group: all W virtual ranks' CTAs must be co-resident (cooperative launch), so the
// per-rank channel count is capped at (resident CTAs on the device) / W
static long group_block_cap(const b2c_comm* c, const void* fn, long want) {
  // tiny cache: the occupancy query costs microseconds of host time per call
  static thread_local struct { const void* fn; int threads, device, resident; } cache[8];
  static thread_local int cache_n = 0;
  const int threads = (int)c->params[B2C_PARAM_THREADS];
  int resident = -1;
  for (int i = 0; i < cache_n; ++i)
    if (cache[i].fn == fn && cache[i].threads == threads && cache[i].device == c->device) resident = cache[i].resident;
  if (resident < 0) {
    int per_sm = 0, sms = 0;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, fn, threads, 0) != cudaSuccess ||
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, c->device) != cudaSuccess) {
      cudaGetLastError();
      return want;
    }
    resident = per_sm * sms;
    cache[cache_n % 8] = {fn, threads, c->device, resident};
    if (cache_n < 8) cache_n++;
  }
  long cap = (long)resident / c->world;
  if (cap < 1) cap = 1;
  return want < cap ? want : cap;
}

struct Launch {
  const void* fn;
  B2cKArgs args;
  int nblocks, threads;
};

static int do_launch(b2c_comm* lead, Launch& L, int grid_y, cudaStream_t stream) {
  void* params[] = {(void*)&L.args};
  dim3 grid(L.nblocks, grid_y), block(L.threads);
  cudaError_t e;
  if (grid_y > 1) {
    // all virtual ranks' CTAs must be co-resident: they wait on one another
    e = cudaLaunchCooperativeKernel(L.fn, grid, block, params, 0, stream);
  } else {
    e = cudaLaunchKernel(L.fn, grid, block, params, 0, stream);
  }
  if (e != cudaSuccess) { cudaGetLastError(); return cuda_fail(e, "kernel launch"); }
  lead->stats.launches++;
  lead->stats.last_blocks = L.nblocks;
  lead->stats.last_threads = L.threads;
  return B2C_OK;
}

// Launch with programmatic stream serialization: the grid may become resident while the previous kernel
// on the stream is still draining (it waits in griddepcontrol.wait); only for kernels that execute that wait.
static int launch_pdl(b2c_comm* c, const void* fn, void* args, int nblocks, int threads, cudaStream_t stream) {
  cudaLaunchConfig_t cfg;
  memset(&cfg, 0, sizeof(cfg));
  cfg.gridDim = dim3(nblocks);
  cfg.blockDim = dim3(threads);
  cfg.stream = stream;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = at;
  cfg.numAttrs = c->params[B2C_PARAM_PDL] ? 1 : 0;
  void* params[] = {args};
  cudaError_t e = cudaLaunchKernelExC(&cfg, fn, params);
  if (e != cudaSuccess) { cudaGetLastError(); return cuda_fail(e, "kernel launch"); }
  c->stats.launches++;
  c->stats.last_blocks = nblocks;
  c->stats.last_threads = threads;
  return B2C_OK;
}

static void fill_common(b2c_comm* c, B2cKArgs& a, bool group) {
  memset(&a, 0, sizeof(a));
  for (int q = 0; q < c->world; ++q) a.region[q] = c->peer_region[q];
  if (group) {
    for (int q = 0; q < c->world; ++q) a.status[q] = c->group[q]->status_d;
    a.rank = -1;
  } else {
    a.status[c->rank] = c->status_d;
    a.rank = c->rank;
  }
  a.world = c->world;
  a.slot_bytes = c->slot_bytes;
  a.timeout_ns = (unsigned long long)c->params[B2C_PARAM_TIMEOUT_MS] * 1000000ull;
  a.variant = (int)c->params[B2C_PARAM_TWOSHOT_VARIANT];
}

static int mode_of(int dtype, int wire, int* mode, int* vw, int* ub) {
  if (dtype == B2C_F32 && wire == B2C_F32) { *mode = MODE_F32; *vw = 4; *ub = 4; return 0; }
  if (dtype == B2C_BF16 && wire == B2C_BF16) { *mode = MODE_BF16; *vw = 8; *ub = 2; return 0; }
  if (dtype == B2C_F32 && wire == B2C_BF16) { *mode = MODE_F32_BF16; *vw = 8; *ub = 4; return 0; }
  return fail(B2C_EUNSUPPORTED, "b2c_allreduce: (dtype=%d, wire_dtype=%d) unsupported; use (F32,F32), (BF16,BF16) or (F32,BF16)", dtype, wire);
}

// one allreduce, possibly split into several launches when it exceeds an arena slot.
// bufs: group ? per-rank buffers : only bufs[c->rank] is used
static int allreduce_impl(b2c_comm* c, void* const* bufs, bool group, size_t count, int dtype, int wire,
                          float scale, int algo, cudaStream_t stream) {
  if (!c) return fail(B2C_EINVAL, "b2c_allreduce: comm is NULL");
  if (!c->ready) return fail(B2C_ESTATE, "b2c_allreduce: comm not imported yet (call b2c_comm_import on every rank)");
  int rc = check_sticky(c);
  if (rc) return rc;
  int mode, vw, ub;
  rc = mode_of(dtype, wire, &mode, &vw, &ub);
  if (rc) return rc;
  if (algo < B2C_ALGO_AUTO || algo > B2C_ALGO_NVLS) return fail(B2C_EINVAL, "b2c_allreduce: algo=%d", algo);
  c->stats.calls_allreduce++;
  c->stats.bytes_allreduce += (uint64_t)count * ub;
  if (count == 0) return B2C_OK;
  const int W = c->world;
  for (int r = 0; r < W; ++r) {
    if (!group && r != c->rank) continue;
    if (!bufs[r]) return fail(B2C_EINVAL, "b2c_allreduce: buf is NULL");
    if ((uintptr_t)bufs[r] % ub) return fail(B2C_EINVAL, "b2c_allreduce: buf %p is not aligned to its element size %d", bufs[r], ub);
  }
  DeviceGuard g(c->device);

  // ---- NVLS: in-switch reduction on the multicast arena
  const bool nvls_up = !group && c->nvls && c->nvls->phase == 3;
  if (algo == B2C_ALGO_NVLS && !nvls_up)
    return fail(B2C_ESTATE, "b2c_allreduce: algo NVLS requested but the multicast arena is not set up (b2c_nvls_export/import/bind)");
  bool nvls_inside = false;
  if (nvls_up) {
    char* p0 = (char*)bufs[c->rank];
    nvls_inside = p0 >= (char*)c->nvls->uc && p0 + count * ub <= (char*)c->nvls->uc + c->nvls->bytes && (((uintptr_t)p0 & 15u) == 0);
  }
  // AUTO takes the switch only for tensors that live in the arena (in place); staging through the
  // arena loses to the zero-copy peer path and is only run on explicit request
  if (nvls_up && W > 1 && mode != MODE_F32_BF16 &&
      (algo == B2C_ALGO_NVLS || (algo == B2C_ALGO_AUTO && nvls_inside && c->params[B2C_PARAM_NVLS_MIN_BYTES] > 0 &&
                                 count * (size_t)ub >= (size_t)c->params[B2C_PARAM_NVLS_MIN_BYTES]))) {
    Nvls* nv = c->nvls;
    char* p = (char*)bufs[c->rank];
    const bool inside = nvls_inside;
    // In place the kernel works on whole 16-byte vectors of the caller's tensor.  A ragged tail (fewer
    // than one vector) must not drag its neighbours in the arena into the reduction, so it goes through
    // the staged one-shot kernel as a second, tiny launch.
    const size_t tail = inside ? count % (size_t)vw : 0;
    const size_t body = count - tail;
    if (inside && body && c->params[B2C_PARAM_NVLS_LEAN]) {
      // the hot path of a training step: one lean launch, multicast flag barrier, no staging
      B2cNvlsArgs a;
      memset(&a, 0, sizeof(a));
      a.uc = (char*)nv->uc; a.mc = (char*)nv->mcva;
      a.ctl = (uint32_t*)c->region; a.status = c->status_d;
      a.off = (size_t)(p - (char*)nv->uc);
      a.nvec = body / vw;
      a.timeout_ns = (unsigned long long)c->params[B2C_PARAM_TIMEOUT_MS] * 1000000ull;
      a.scale = scale; a.rank = c->rank; a.world = W;
      a.acq_poll = (int)c->params[B2C_PARAM_MC_ACQ_POLL];
      const int threads = c->params[B2C_PARAM_THREADS_NVLS] <= 256 ? 256 : 512;
      const int unroll = c->params[B2C_PARAM_UNROLL_NVLS] <= 4 ? 4 : 8;
      const size_t nvs = (a.nvec + W - 1) / W;
      const size_t per = (size_t)threads * unroll;                 // vectors one CTA moves per iteration
      long mb = c->params[B2C_PARAM_MAX_BLOCKS_NVLS];
      if (threads <= 256) mb *= 2;                                 // same bytes in flight with half-size CTAs
      if (mb > B2C_MAX_BLOCKS) mb = B2C_MAX_BLOCKS;
      size_t nb = (nvs + per - 1) / per;
      if (nb < 1) nb = 1;
      if (nb > (size_t)mb) nb = mb;
      a.chunk_vecs = (uint32_t)per;
      c->stats.nvls_launches++;
      c->stats.last_algo = B2C_ALGO_NVLS;
      rc = launch_pdl(c, nvls_lean_fn(mode, threads, unroll), &a, (int)nb, threads, stream);
      if (rc) return rc;
    } else {
      if (!inside) {
        rc = nvls_stage_slots(c);
        if (rc) return rc;
      }
      const size_t piece = inside ? (body ? body : 1) : (nv->slot_bytes / 16) * vw;
      for (size_t off = 0; off < body; off += piece) {
        const size_t n = (body - off < piece) ? body - off : piece;
        Launch L;
        fill_common(c, L.args, false);
        L.args.buf[c->rank] = p + off * ub;
        L.args.count = n;
        L.args.scale = scale;
        L.args.nvls_uc = (char*)nv->uc;
        L.args.nvls_mc = (char*)nv->mcva;
        L.args.nvls_slot_bytes = nv->slot_bytes;
        L.args.nvls_inplace = inside ? 1 : 0;
        L.args.nvls_off = inside ? (size_t)(p - (char*)nv->uc) : nv->slot_off;
        const size_t nvec = (n + vw - 1) / vw;
        L.fn = nvls_fn(mode, (int)c->params[B2C_PARAM_UNROLL_NVLS]);
        geometry(c, (nvec + W - 1) / W, c->params[B2C_PARAM_MAX_BLOCKS_NVLS], &L.nblocks, &L.threads, &L.args.chunk_vecs);
        c->stats.nvls_launches++;
        c->stats.last_algo = B2C_ALGO_NVLS;
        rc = do_launch(c, L, 1, stream);
        if (rc) return rc;
      }
    }
    if (tail) {
      void* tb[B2C_MAX_WORLD] = {0};
      tb[c->rank] = p + body * ub;
      c->stats.calls_allreduce--;                       // one user call, two launches
      c->stats.bytes_allreduce -= (uint64_t)tail * ub;
      const long keep = c->params[B2C_PARAM_ZEROCOPY];
      c->params[B2C_PARAM_ZEROCOPY] = 0;
      rc = allreduce_impl(c, tb, false, tail, dtype, wire, scale, B2C_ALGO_ONESHOT, stream);
      c->params[B2C_PARAM_ZEROCOPY] = keep;
      if (rc) return rc;
      c->stats.last_algo = B2C_ALGO_NVLS;
    }
    return B2C_OK;
  }

  // ---- zero-copy two-shot: the buffers themselves are peer-visible (local group, or registered)
  if (W > 1 && c->params[B2C_PARAM_ZEROCOPY] && mode != MODE_F32_BF16 && algo != B2C_ALGO_ONESHOT &&
      (algo == B2C_ALGO_TWOSHOT || count * (size_t)ub >= (size_t)c->params[B2C_PARAM_ZEROCOPY_MIN_BYTES])) {
    Registration* reg = group ? nullptr : find_reg(c, bufs[c->rank], count * (size_t)ub);
    bool ok = group || reg != nullptr;
    for (int r = 0; ok && r < W; ++r)
      if ((group || r == c->rank) && ((uintptr_t)bufs[r] & 15u)) ok = false;
    if (ok) {
      Launch L;
      fill_common(c, L.args, group);
      if (group) {
        for (int r = 0; r < W; ++r) L.args.buf[r] = bufs[r];
      } else {
        const size_t roff = (char*)bufs[c->rank] - reg->local;  // same offset on every rank (contract)
        for (int r = 0; r < W; ++r) L.args.buf[r] = reg->peer[r] + roff;
      }
      L.args.count = count;
      L.args.scale = scale;
      const size_t nvs = (count / vw + W - 1) / W;
      L.fn = mode == MODE_F32 ? twoshot_zc_fn<MODE_F32>(W) : twoshot_zc_fn<MODE_BF16>(W);
      if (W == 8 && c->params[B2C_PARAM_UNROLL_ZC] > 0) L.fn = zc8_fn(mode, (int)c->params[B2C_PARAM_UNROLL_ZC]);
      long mb = c->params[B2C_PARAM_MAX_BLOCKS_TWOSHOT];
      if (group) mb = group_block_cap(c, L.fn, mb);
      geometry(c, nvs, mb, &L.nblocks, &L.threads, &L.args.chunk_vecs);
      c->stats.twoshot_launches++;
      c->stats.zerocopy_launches++;
      c->stats.last_algo = B2C_ALGO_TWOSHOT;
      return do_launch(c, L, group ? W : 1, stream);
    }
  }

  // ---- a single rank averaging over itself: the result is the input, bit for bit (x * 1.0f == x, NaN
  // payloads included); a bf16 wire would still round, so that mode goes on
  if (W == 1 && scale == 1.0f && mode != MODE_F32_BF16) {
    c->stats.last_algo = B2C_ALGO_ONESHOT;
    c->stats.noop_calls++;
    return B2C_OK;
  }
  // ---- a single rank: streaming scale pass, no arena, no flags
  if (W == 1 && mode != MODE_F32_BF16 && (((uintptr_t)bufs[c->rank < 0 ? 0 : c->rank]) & 15u) == 0) {
    void* b = bufs[c->rank];
    const size_t nvec = count / vw;
    size_t blocks = (nvec + 256 * 4 - 1) / (256 * 4);
    if (blocks < 1) blocks = 1;
    if (blocks > 148 * 8) blocks = 148 * 8;
    if (mode == MODE_F32) k_scale_inplace<MODE_F32><<<(unsigned)blocks, 256, 0, stream>>>(b, count, scale);
    else k_scale_inplace<MODE_BF16><<<(unsigned)blocks, 256, 0, stream>>>(b, count, scale);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return cuda_fail(e, "kernel launch (scale pass)");
    c->stats.launches++;
    c->stats.oneshot_launches++;
    c->stats.last_algo = B2C_ALGO_ONESHOT;
    c->stats.last_blocks = (uint32_t)blocks;
    c->stats.last_threads = 256;
    return B2C_OK;
  }

  const size_t slot_vecs = c->slot_bytes / 16;
  // pieces keep 16-byte alignment relative to buf; a single rank never touches the arena, so no split
  const size_t piece_elems = (W == 1) ? (count > 0 ? count : 1) : slot_vecs * vw;
  for (size_t off = 0; off < count; off += piece_elems) {
    const size_t n = (count - off < piece_elems) ? count - off : piece_elems;
    const size_t nvec = (n + vw - 1) / vw;
    int use = algo;
    if (W == 1) use = B2C_ALGO_ONESHOT;
    else if (use == B2C_ALGO_AUTO) use = (nvec * 16 <= (size_t)c->params[B2C_PARAM_CROSSOVER_BYTES]) ? B2C_ALGO_ONESHOT : B2C_ALGO_TWOSHOT;
    Launch L;
    fill_common(c, L.args, group);
    for (int r = 0; r < W; ++r)
      if (group || r == c->rank) L.args.buf[r] = (char*)bufs[r] + off * ub;
    L.args.count = n;
    L.args.scale = scale;
    if (use == B2C_ALGO_ONESHOT) {
      L.fn = mode == MODE_F32 ? oneshot_fn<MODE_F32>(W) : mode == MODE_BF16 ? oneshot_fn<MODE_BF16>(W) : oneshot_fn<MODE_F32_BF16>(W);
      long mb = c->params[B2C_PARAM_MAX_BLOCKS_ONESHOT];
      if (group) mb = group_block_cap(c, L.fn, mb);
      if (W == 1) {
        // no peer, no flags: a plain HBM-bound scale pass, so fill the whole chip (several CTAs per SM)
        geometry_wide(c, nvec, 148 * 4, &L.nblocks, &L.threads, &L.args.chunk_vecs);
      } else
      geometry(c, nvec, mb, &L.nblocks, &L.threads, &L.args.chunk_vecs);
      c->stats.oneshot_launches++;
    } else {
      const size_t nvs = (nvec + W - 1) / W;
      L.fn = mode == MODE_F32 ? twoshot_fn<MODE_F32>(W) : mode == MODE_BF16 ? twoshot_fn<MODE_BF16>(W) : twoshot_fn<MODE_F32_BF16>(W);
      long mb = c->params[B2C_PARAM_MAX_BLOCKS_TWOSHOT];
      if (group) mb = group_block_cap(c, L.fn, mb);
      geometry(c, nvs, mb, &L.nblocks, &L.threads, &L.args.chunk_vecs);
      c->stats.twoshot_launches++;
    }
    c->stats.last_algo = use;
    rc = do_launch(c, L, group ? W : 1, stream);
    if (rc) return rc;
  }
  return B2C_OK;
}

static int broadcast_impl(b2c_comm* c, void* const* bufs, bool group, size_t bytes, int root, cudaStream_t stream) {
  if (!c) return fail(B2C_EINVAL, "b2c_broadcast: comm is NULL");
  if (!c->ready) return fail(B2C_ESTATE, "b2c_broadcast: comm not imported yet");
  if (root < 0 || root >= c->world) return fail(B2C_EINVAL, "b2c_broadcast: root=%d outside 0..%d", root, c->world - 1);
  int rc = check_sticky(c);
  if (rc) return rc;
  c->stats.calls_broadcast++;
  c->stats.bytes_broadcast += bytes;
  if (bytes == 0 || c->world == 1) return B2C_OK;
  for (int r = 0; r < c->world; ++r)
    if ((group || r == c->rank) && !bufs[r]) return fail(B2C_EINVAL, "b2c_broadcast: buf is NULL");
  DeviceGuard g(c->device);
  // ---- through the switch: root sends S once (multimem.st), everyone else copies locally
  const bool nvls_up = !group && c->nvls && c->nvls->phase == 3;
  if (nvls_up && c->params[B2C_PARAM_BCAST_MC]) {
    Nvls* nv = c->nvls;
    char* p = (char*)bufs[c->rank];
    // in place: the tensor itself lives in the arena (same offset on every rank), whole vectors only
    // (in place needs an entry barrier on top of the exit barrier — the root writes straight into the peers' tensor — so
    // below 1 MiB, where the barriers are the whole cost, the staged form with its single barrier is the faster one)
    const bool inside = p >= (char*)nv->uc && p + bytes <= (char*)nv->uc + nv->bytes && (((uintptr_t)p & 15u) == 0) && bytes % 16 == 0 &&
                        bytes >= (size_t)c->params[B2C_PARAM_BCAST_INPLACE_MIN_BYTES];
    if (!inside) {
      rc = nvls_stage_slots(c);
      if (rc) return rc;
    }
    const size_t piece = inside ? bytes : nv->slot_bytes;
    for (size_t off = 0; off < bytes; off += piece) {
      const size_t n = (bytes - off < piece) ? bytes - off : piece;
      Launch L;
      fill_common(c, L.args, false);
      L.args.buf[c->rank] = p + off;
      L.args.count = n;
      L.args.root = root;
      L.args.nvls_uc = (char*)nv->uc;
      L.args.nvls_mc = (char*)nv->mcva;
      L.args.nvls_slot_bytes = nv->slot_bytes;
      L.args.nvls_inplace = inside ? 1 : 0;
      L.args.nvls_off = inside ? (size_t)(p - (char*)nv->uc) : nv->slot_off;
      L.fn = (const void*)k_broadcast_mc;
      geometry(c, (n + 15) / 16, c->params[B2C_PARAM_MAX_BLOCKS_BCAST], &L.nblocks, &L.threads, &L.args.chunk_vecs);
      c->stats.bcast_mc_launches++;
      rc = do_launch(c, L, 1, stream);
      if (rc) return rc;
    }
    return B2C_OK;
  }
  const size_t piece = c->slot_bytes;
  for (size_t off = 0; off < bytes; off += piece) {
    const size_t n = (bytes - off < piece) ? bytes - off : piece;
    Launch L;
    fill_common(c, L.args, group);
    for (int r = 0; r < c->world; ++r)
      if (group || r == c->rank) L.args.buf[r] = (char*)bufs[r] + off;
    L.args.count = n;
    L.args.root = root;
    L.fn = (const void*)k_broadcast;
    long mb = c->params[B2C_PARAM_MAX_BLOCKS_BCAST];
    if (group) mb = group_block_cap(c, L.fn, mb);
    geometry(c, (n + 15) / 16, mb, &L.nblocks, &L.threads, &L.args.chunk_vecs);
    rc = do_launch(c, L, group ? c->world : 1, stream);
    if (rc) return rc;
  }
  return B2C_OK;
}

static int barrier_impl(b2c_comm* c, bool group, cudaStream_t stream) {
  if (!c) return fail(B2C_EINVAL, "b2c_barrier: comm is NULL");
  if (!c->ready) return fail(B2C_ESTATE, "b2c_barrier: comm not imported yet");
  int rc = check_sticky(c);
  if (rc) return rc;
  c->stats.calls_barrier++;
  DeviceGuard g(c->device);
  Launch L;
  fill_common(c, L.args, group);
  L.nblocks = 1; L.threads = 32; L.args.chunk_vecs = 32;
  L.fn = (const void*)k_barrier;
  return do_launch(c, L, group ? c->world : 1, stream);
}

// ---------------------------------------------------------------- C-ABI data path
extern "C" int b2c_allreduce(b2c_comm* c, void* buf, size_t count, int dtype, int wire_dtype, float scale,
                             int algo, b2c_stream_t stream) {
  if (c && c->local_group) return fail(B2C_ESTATE, "b2c_allreduce: comm belongs to a local group; use b2c_group_allreduce");
  void* bufs[B2C_MAX_WORLD] = {0};
  if (c && c->rank >= 0 && c->rank < B2C_MAX_WORLD) bufs[c->rank] = buf;
  return allreduce_impl(c, bufs, false, count, dtype, wire_dtype, scale, algo, (cudaStream_t)stream);
}
extern "C" int b2c_broadcast(b2c_comm* c, void* buf, size_t bytes, int root, b2c_stream_t stream) {
  if (c && c->local_group) return fail(B2C_ESTATE, "b2c_broadcast: comm belongs to a local group; use b2c_group_broadcast");
  void* bufs[B2C_MAX_WORLD] = {0};
  if (c && c->rank >= 0 && c->rank < B2C_MAX_WORLD) bufs[c->rank] = buf;
  return broadcast_impl(c, bufs, false, bytes, root, (cudaStream_t)stream);
}
extern "C" int b2c_barrier(b2c_comm* c, b2c_stream_t stream) {
  if (c && c->local_group) return fail(B2C_ESTATE, "b2c_barrier: comm belongs to a local group; use b2c_group_barrier");
  return barrier_impl(c, false, (cudaStream_t)stream);
}

static int group_check(b2c_comm** comms, const char* who) {
  if (!comms || !comms[0]) return fail(B2C_EINVAL, "%s: comms is NULL", who);
  if (!comms[0]->local_group) return fail(B2C_ESTATE, "%s: not a local group", who);
  return B2C_OK;
}
extern "C" int b2c_group_allreduce(b2c_comm** comms, void** bufs, size_t count, int dtype, int wire_dtype,
                                   float scale, int algo, b2c_stream_t stream) {
  int rc = group_check(comms, "b2c_group_allreduce");
  if (rc) return rc;
  if (!bufs) return fail(B2C_EINVAL, "b2c_group_allreduce: bufs is NULL");
  return allreduce_impl(comms[0], bufs, true, count, dtype, wire_dtype, scale, algo, (cudaStream_t)stream);
}
extern "C" int b2c_group_broadcast(b2c_comm** comms, void** bufs, size_t bytes, int root, b2c_stream_t stream) {
  int rc = group_check(comms, "b2c_group_broadcast");
  if (rc) return rc;
  if (!bufs) return fail(B2C_EINVAL, "b2c_group_broadcast: bufs is NULL");
  return broadcast_impl(comms[0], bufs, true, bytes, root, (cudaStream_t)stream);
}
extern "C" int b2c_group_barrier(b2c_comm** comms, b2c_stream_t stream) {
  int rc = group_check(comms, "b2c_group_barrier");
  if (rc) return rc;
  return barrier_impl(comms[0], true, (cudaStream_t)stream);
}

// ---------------------------------------------------------------- NVLS multicast arena
struct NvlsBlob {
  uint32_t magic;
  int32_t rank;
  int64_t pid;
  int32_t mc_fd;
  int32_t pad0;
  uint64_t bytes;
};
static_assert(sizeof(NvlsBlob) <= B2C_NVLS_BLOB_BYTES, "nvls blob too large");

struct Drv {
  decltype(&cuDeviceGet) DeviceGet;
  decltype(&cuMemCreate) MemCreate;
  decltype(&cuMemRelease) MemRelease;
  decltype(&cuMemGetAllocationGranularity) MemGetAllocationGranularity;
  decltype(&cuMemAddressReserve) MemAddressReserve;
  decltype(&cuMemAddressFree) MemAddressFree;
  decltype(&cuMemMap) MemMap;
  decltype(&cuMemUnmap) MemUnmap;
  decltype(&cuMemSetAccess) MemSetAccess;
  decltype(&cuMemExportToShareableHandle) MemExportToShareableHandle;
  decltype(&cuMemImportFromShareableHandle) MemImportFromShareableHandle;
  decltype(&cuMulticastCreate) MulticastCreate;
  decltype(&cuMulticastAddDevice) MulticastAddDevice;
  decltype(&cuMulticastBindMem) MulticastBindMem;
  decltype(&cuMulticastGetGranularity) MulticastGetGranularity;
  decltype(&cuGetErrorString) GetErrorString;
  bool ok;
};
static Drv* drv() {
  static Drv d;
  static bool tried = false;
  if (tried) return d.ok ? &d : nullptr;
  tried = true;
  cudaDriverEntryPointQueryResult qr;
  bool ok = true;
#define B2C_SYM(field, name) ok = ok && (cudaGetDriverEntryPoint(name, (void**)&d.field, cudaEnableDefault, &qr) == cudaSuccess) && d.field
  B2C_SYM(DeviceGet, "cuDeviceGet");
  B2C_SYM(MemCreate, "cuMemCreate");
  B2C_SYM(MemRelease, "cuMemRelease");
  B2C_SYM(MemGetAllocationGranularity, "cuMemGetAllocationGranularity");
  B2C_SYM(MemAddressReserve, "cuMemAddressReserve");
  B2C_SYM(MemAddressFree, "cuMemAddressFree");
  B2C_SYM(MemMap, "cuMemMap");
  B2C_SYM(MemUnmap, "cuMemUnmap");
  B2C_SYM(MemSetAccess, "cuMemSetAccess");
  B2C_SYM(MemExportToShareableHandle, "cuMemExportToShareableHandle");
  B2C_SYM(MemImportFromShareableHandle, "cuMemImportFromShareableHandle");
  B2C_SYM(MulticastCreate, "cuMulticastCreate");
  B2C_SYM(MulticastAddDevice, "cuMulticastAddDevice");
  B2C_SYM(MulticastBindMem, "cuMulticastBindMem");
  B2C_SYM(MulticastGetGranularity, "cuMulticastGetGranularity");
  B2C_SYM(GetErrorString, "cuGetErrorString");
#undef B2C_SYM
  cudaGetLastError();
  d.ok = ok;
  return ok ? &d : nullptr;
}
static int drv_fail(Drv* d, CUresult r, const char* what) {
  const char* msg = nullptr;
  if (d) d->GetErrorString(r, &msg);
  return fail(B2C_EUNSUPPORTED, "b2c_nvls: %s failed: driver error %d (%s)", what, (int)r, msg ? msg : "?");
}
#define DRV(call, what)                                  \
  do {                                                   \
    CUresult r_ = (call);                                \
    if (r_ != CUDA_SUCCESS) return drv_fail(d, r_, what); \
  } while (0)

static void nvls_release(b2c_comm* c) {
  pthread_mutex_lock(&g_pool_lock);
  if (g_pool_comm == c) g_pool_comm = nullptr;
  pthread_mutex_unlock(&g_pool_lock);
  Nvls* n = c->nvls;
  Drv* d = drv();
  if (!n) return;
  if (d) {
    if (n->mc_mapped) { d->MemUnmap(n->mcva, n->bytes); d->MemAddressFree(n->mcva, n->bytes); }
    if (n->uc_mapped) { d->MemUnmap(n->uc, n->bytes); d->MemAddressFree(n->uc, n->bytes); }
    if (n->mc) d->MemRelease(n->mc);
    if (n->mem) d->MemRelease(n->mem);
  }
  if (n->mc_fd > 0) close(n->mc_fd);
  free(n);
  c->nvls = nullptr;
}

extern "C" int b2c_nvls_export(b2c_comm* c, size_t bytes, void* blob, size_t* len) {
  if (!c || !blob || !len) return fail(B2C_EINVAL, "b2c_nvls_export: NULL argument");
  if (*len < B2C_NVLS_BLOB_BYTES) return fail(B2C_EINVAL, "b2c_nvls_export: blob needs %d bytes", B2C_NVLS_BLOB_BYTES);
  if (c->local_group || c->world < 2) return fail(B2C_ESTATE, "b2c_nvls_export: needs a cross-process comm with world >= 2");
  if (!c->ready) return fail(B2C_ESTATE, "b2c_nvls_export: comm not imported yet");
  if (c->nvls) return fail(B2C_ESTATE, "b2c_nvls_export: already set up");
  DeviceGuard g(c->device);
  cudaFree(0);
  Drv* d = drv();
  if (!d) return fail(B2C_EUNSUPPORTED, "b2c_nvls_export: this driver has no multicast entry points");
  CUdevice dev;
  DRV(d->DeviceGet(&dev, c->device), "cuDeviceGet");
  CUmulticastObjectProp mp;
  memset(&mp, 0, sizeof(mp));
  mp.numDevices = (unsigned)c->world;
  mp.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  mp.size = bytes;
  size_t mgran = 0, agran = 0;
  DRV(d->MulticastGetGranularity(&mgran, &mp, CU_MULTICAST_GRANULARITY_MINIMUM), "cuMulticastGetGranularity");
  CUmemAllocationProp ap;
  memset(&ap, 0, sizeof(ap));
  ap.type = CU_MEM_ALLOCATION_TYPE_PINNED;
  ap.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  ap.location.id = c->device;
  ap.requestedHandleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  DRV(d->MemGetAllocationGranularity(&agran, &ap, CU_MEM_ALLOC_GRANULARITY_MINIMUM), "cuMemGetAllocationGranularity");
  size_t gran = mgran > agran ? mgran : agran;
  size_t size = (bytes + gran - 1) / gran * gran;
  Nvls* n = (Nvls*)calloc(1, sizeof(Nvls));
  n->bytes = size;
  n->mc_fd = -1;
  c->nvls = n;
  CUresult r = d->MemCreate(&n->mem, size, &ap, 0);
  if (r != CUDA_SUCCESS) { nvls_release(c); return drv_fail(d, r, "cuMemCreate"); }
  if (c->rank == 0) {
    mp.size = size;
    r = d->MulticastCreate(&n->mc, &mp);
    if (r != CUDA_SUCCESS) { nvls_release(c); return drv_fail(d, r, "cuMulticastCreate"); }
    r = d->MemExportToShareableHandle(&n->mc_fd, n->mc, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0);
    if (r != CUDA_SUCCESS) { nvls_release(c); return drv_fail(d, r, "cuMemExportToShareableHandle(multicast)"); }
  }
  NvlsBlob b;
  memset(&b, 0, sizeof(b));
  b.magic = B2C_MAGIC ^ 0x33333333u;
  b.rank = c->rank; b.pid = (int64_t)getpid(); b.mc_fd = n->mc_fd; b.bytes = size;
  memset(blob, 0, B2C_NVLS_BLOB_BYTES);
  memcpy(blob, &b, sizeof(b));
  *len = B2C_NVLS_BLOB_BYTES;
  n->phase = 1;
  return B2C_OK;
}

extern "C" int b2c_nvls_import(b2c_comm* c, const void* blobs, size_t len_each) {
  if (!c || !blobs) return fail(B2C_EINVAL, "b2c_nvls_import: NULL argument");
  if (!c->nvls || c->nvls->phase != 1) return fail(B2C_ESTATE, "b2c_nvls_import: call b2c_nvls_export first");
  if (len_each < sizeof(NvlsBlob)) return fail(B2C_EINVAL, "b2c_nvls_import: len_each too small");
  DeviceGuard g(c->device);
  Drv* d = drv();
  Nvls* n = c->nvls;
  NvlsBlob b0;
  memcpy(&b0, blobs, sizeof(b0));
  for (int q = 0; q < c->world; ++q) {
    NvlsBlob b;
    memcpy(&b, (const char*)blobs + (size_t)q * len_each, sizeof(b));
    if (b.magic != (B2C_MAGIC ^ 0x33333333u) || b.rank != q) return fail(B2C_EINVAL, "b2c_nvls_import: blob %d malformed / out of order", q);
    if (b.bytes != n->bytes) return fail(B2C_EINVAL, "b2c_nvls_import: rank %d arena %llu B != ours %zu B", q, (unsigned long long)b.bytes, n->bytes);
  }
  if (c->rank != 0) {
    // fetch rank 0's file descriptor: pidfd_open + pidfd_getfd (Linux >= 5.6, same user / CAP_SYS_PTRACE)
#if defined(SYS_pidfd_open) && defined(SYS_pidfd_getfd)
    int pidfd = (int)syscall(SYS_pidfd_open, (pid_t)b0.pid, 0);
    if (pidfd < 0) return fail(B2C_EUNSUPPORTED, "b2c_nvls_import: pidfd_open(%lld) failed (errno %d)", (long long)b0.pid, errno);
    int fd = (int)syscall(SYS_pidfd_getfd, pidfd, b0.mc_fd, 0);
    int err = errno;
    close(pidfd);
    if (fd < 0) return fail(B2C_EUNSUPPORTED, "b2c_nvls_import: pidfd_getfd failed (errno %d); ranks must share a user / ptrace scope", err);
    CUresult r = d->MemImportFromShareableHandle(&n->mc, (void*)(uintptr_t)fd, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR);
    close(fd);
    if (r != CUDA_SUCCESS) return drv_fail(d, r, "cuMemImportFromShareableHandle(multicast)");
#else
    return fail(B2C_EUNSUPPORTED, "b2c_nvls_import: built without pidfd syscalls");
#endif
  }
  CUdevice dev;
  DRV(d->DeviceGet(&dev, c->device), "cuDeviceGet");
  DRV(d->MulticastAddDevice(n->mc, dev), "cuMulticastAddDevice");
  n->phase = 2;
  return B2C_OK;
}

extern "C" int b2c_nvls_bind(b2c_comm* c) {
  if (!c) return fail(B2C_EINVAL, "b2c_nvls_bind: comm is NULL");
  if (!c->nvls || c->nvls->phase != 2) return fail(B2C_ESTATE, "b2c_nvls_bind: call b2c_nvls_import (on every rank, then a barrier) first");
  DeviceGuard g(c->device);
  Drv* d = drv();
  Nvls* n = c->nvls;
  DRV(d->MulticastBindMem(n->mc, 0, n->mem, 0, n->bytes, 0), "cuMulticastBindMem");
  n->bound = true;
  CUmemAccessDesc acc;
  memset(&acc, 0, sizeof(acc));
  acc.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  acc.location.id = c->device;
  acc.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
  DRV(d->MemAddressReserve(&n->uc, n->bytes, 0, 0, 0), "cuMemAddressReserve(uc)");
  DRV(d->MemMap(n->uc, n->bytes, 0, n->mem, 0), "cuMemMap(uc)");
  n->uc_mapped = true;
  DRV(d->MemSetAccess(n->uc, n->bytes, &acc, 1), "cuMemSetAccess(uc)");
  DRV(d->MemAddressReserve(&n->mcva, n->bytes, 0, 0, 0), "cuMemAddressReserve(mc)");
  DRV(d->MemMap(n->mcva, n->bytes, 0, n->mc, 0), "cuMemMap(mc)");
  n->mc_mapped = true;
  DRV(d->MemSetAccess(n->mcva, n->bytes, &acc, 1), "cuMemSetAccess(mc)");
  CU(cudaMemset((void*)n->uc, 0, n->bytes));
  CU(cudaDeviceSynchronize());
  if (n->bytes <= B2C_MC_HEAD_BYTES + 4096) return fail(B2C_EINVAL, "b2c_nvls_bind: arena of %zu bytes is too small", n->bytes);
  n->blocks[0].off = B2C_MC_HEAD_BYTES;   // [0, head): multicast flag counters
  n->blocks[0].size = n->bytes - B2C_MC_HEAD_BYTES;
  n->blocks[0].used = false;
  n->nblocks = 1;
  n->phase = 3;
  return B2C_OK;
}

extern "C" int b2c_nvls_arena(b2c_comm* c, void** base, size_t* bytes) {
  if (!c) return fail(B2C_EINVAL, "b2c_nvls_arena: comm is NULL");
  if (!c->nvls || c->nvls->phase != 3) return fail(B2C_ESTATE, "b2c_nvls_arena: multicast arena is not set up");
  if (base) *base = (void*)c->nvls->uc;
  if (bytes) *bytes = c->nvls->bytes;
  return B2C_OK;
}

// first fit over the address-ordered block list; caller holds g_pool_lock
static int arena_alloc_locked(Nvls* n, size_t bytes, size_t* off) {
  const size_t need = ((bytes ? bytes : 1) + 255) & ~(size_t)255;
  for (int i = 0; i < n->nblocks; ++i) {
    ArenaBlock& b = n->blocks[i];
    if (b.used || b.size < need) continue;
    if (b.size > need) {
      if (n->nblocks == B2C_NVLS_MAX_BLOCKS) return fail(B2C_ETOOBIG, "b2c_nvls_alloc: more than %d live blocks", B2C_NVLS_MAX_BLOCKS);
      memmove(&n->blocks[i + 2], &n->blocks[i + 1], (size_t)(n->nblocks - i - 1) * sizeof(ArenaBlock));
      n->blocks[i + 1].off = b.off + need;
      n->blocks[i + 1].size = b.size - need;
      n->blocks[i + 1].used = false;
      n->nblocks++;
      n->blocks[i].size = need;
    }
    n->blocks[i].used = true;
    *off = n->blocks[i].off;
    return B2C_OK;
  }
  size_t used = 0;
  for (int i = 0; i < n->nblocks; ++i) if (n->blocks[i].used) used += n->blocks[i].size;
  return fail(B2C_ETOOBIG, "b2c_nvls_alloc: %zu bytes do not fit (arena %zu, in use %zu, %d blocks)", bytes, n->bytes, used, n->nblocks);
}
static int arena_free_locked(Nvls* n, size_t off) {
  for (int i = 0; i < n->nblocks; ++i) {
    if (n->blocks[i].off != off || !n->blocks[i].used) continue;
    n->blocks[i].used = false;
    if (i + 1 < n->nblocks && !n->blocks[i + 1].used) {  // merge with the right neighbour
      n->blocks[i].size += n->blocks[i + 1].size;
      memmove(&n->blocks[i + 1], &n->blocks[i + 2], (size_t)(n->nblocks - i - 2) * sizeof(ArenaBlock));
      n->nblocks--;
    }
    if (i > 0 && !n->blocks[i - 1].used) {               // and with the left one
      n->blocks[i - 1].size += n->blocks[i].size;
      memmove(&n->blocks[i], &n->blocks[i + 1], (size_t)(n->nblocks - i - 1) * sizeof(ArenaBlock));
      n->nblocks--;
    }
    return B2C_OK;
  }
  return fail(B2C_EINVAL, "b2c_nvls_free: offset %zu is not a live allocation", off);
}

extern "C" int b2c_nvls_alloc(b2c_comm* c, size_t bytes, void** ptr) {
  if (!c || !ptr) return fail(B2C_EINVAL, "b2c_nvls_alloc: NULL argument");
  if (!c->nvls || c->nvls->phase != 3) return fail(B2C_ESTATE, "b2c_nvls_alloc: multicast arena is not set up");
  Nvls* n = c->nvls;
  size_t off = 0;
  pthread_mutex_lock(&g_pool_lock);
  int rc = arena_alloc_locked(n, bytes, &off);
  pthread_mutex_unlock(&g_pool_lock);
  if (rc == B2C_OK) *ptr = (void*)(n->uc + off);
  return rc;
}
extern "C" int b2c_nvls_free(b2c_comm* c, void* ptr) {
  if (!c) return fail(B2C_EINVAL, "b2c_nvls_free: comm is NULL");
  if (!c->nvls || c->nvls->phase != 3) return fail(B2C_ESTATE, "b2c_nvls_free: multicast arena is not set up");
  Nvls* n = c->nvls;
  if ((char*)ptr < (char*)n->uc || (char*)ptr >= (char*)n->uc + n->bytes) return fail(B2C_EINVAL, "b2c_nvls_free: %p is outside the arena", ptr);
  pthread_mutex_lock(&g_pool_lock);
  int rc = arena_free_locked(n, (size_t)((char*)ptr - (char*)n->uc));
  pthread_mutex_unlock(&g_pool_lock);
  return rc;
}
// staging slots for tensors that live OUTSIDE the arena (explicit B2C_ALGO_NVLS requests, the multicast
// broadcast): 2 x slot, carved out once; collective by construction (every rank makes the same first call)
static int nvls_stage_slots(b2c_comm* c) {
  Nvls* n = c->nvls;
  if (n->slot_bytes) return B2C_OK;
  size_t slot = (size_t)c->params[B2C_PARAM_NVLS_STAGE_BYTES] & ~(size_t)255;
  size_t off = 0;
  int rc = B2C_ETOOBIG;
  pthread_mutex_lock(&g_pool_lock);
  for (; slot >= (64u << 10); slot /= 2) {   // same arena state on every rank => same slot size is found
    rc = arena_alloc_locked(n, 2 * slot, &off);
    if (rc == B2C_OK) break;
  }
  pthread_mutex_unlock(&g_pool_lock);
  if (rc != B2C_OK) return fail(B2C_ETOOBIG, "b2c: no room left in the multicast arena for staging slots (enlarge B2C_NVLS_ARENA_BYTES)");
  n->slot_bytes = slot;
  n->slot_off = off;
  return B2C_OK;
}
extern "C" int b2c_nvls_pool_bind(b2c_comm* c) {
  if (c && (!c->nvls || c->nvls->phase != 3)) return fail(B2C_ESTATE, "b2c_nvls_pool_bind: multicast arena is not set up");
  pthread_mutex_lock(&g_pool_lock);
  g_pool_comm = c;
  pthread_mutex_unlock(&g_pool_lock);
  return B2C_OK;
}
extern "C" void* b2c_pool_alloc(ptrdiff_t size, int device, b2c_stream_t stream) {
  (void)device; (void)stream;
  void* p = nullptr;
  if (!g_pool_comm || size < 0 || b2c_nvls_alloc(g_pool_comm, (size_t)size, &p) != B2C_OK) return nullptr;
  return p;
}
extern "C" void b2c_pool_free(void* ptr, ptrdiff_t size, int device, b2c_stream_t stream) {
  (void)size; (void)device; (void)stream;
  b2c_comm* c = g_pool_comm;
  if (c && c->nvls) b2c_nvls_free(c, ptr);  // after the comm is gone the arena is unmapped as a whole
}


// ---------------------------------------------------------------- fused optimizer step (§8(f)4)
static int opt_to_k(const b2c_opt* o, B2cOptK* k) {
  if (!o) return fail(B2C_EINVAL, "b2c_fused_step: opt is NULL");
  if (o->kind != B2C_OPT_SGD && o->kind != B2C_OPT_ADAM && o->kind != B2C_OPT_ADAMW)
    return fail(B2C_EUNSUPPORTED, "b2c_fused_step: optimizer kind %d (1 = SGD, 2 = Adam, 3 = AdamW)", o->kind);
  memset(k, 0, sizeof(*k));
  k->kind = o->kind; k->nesterov = o->nesterov; k->first = o->first_step;
  k->lr = (float)o->lr; k->momentum = (float)o->momentum; k->one_m_dampening = (float)(1.0 - o->dampening);
  k->weight_decay = (float)o->weight_decay;
  k->one_m_beta1 = (float)(1.0 - o->beta1); k->beta2 = (float)o->beta2; k->one_m_beta2 = (float)(1.0 - o->beta2);
  k->eps = (float)o->eps; k->adamw_keep = (float)(1.0 - o->lr * o->weight_decay);
  if (o->kind != B2C_OPT_SGD) {
    if (!(o->bias_correction1 > 0.0) || !(o->bias_correction2 > 0.0))
      return fail(B2C_EINVAL, "b2c_fused_step: Adam needs bias_correction1/2 = 1 - beta^t > 0");
    k->step_size = (float)(o->lr / o->bias_correction1);
    k->bc2_sqrt = (float)sqrt(o->bias_correction2);
  }
  return B2C_OK;
}

extern "C" int b2c_fused_shard(size_t count, int world, int rank, size_t* first_elem, size_t* n_elems) {
  if (world < 1 || world > B2C_MAX_WORLD || rank < 0 || rank >= world) return fail(B2C_EINVAL, "b2c_fused_shard: rank %d of %d", rank, world);
  if (world == 1) {
    if (first_elem) *first_elem = 0;
    if (n_elems) *n_elems = count;
    return B2C_OK;
  }
  const size_t nvec = count / 4, nvs = (nvec + world - 1) / world;
  const size_t s0 = (size_t)rank * nvs;
  const size_t sl = s0 >= nvec ? 0 : (s0 + nvs <= nvec ? nvs : nvec - s0);
  if (first_elem) *first_elem = s0 * 4;
  if (n_elems) *n_elems = sl * 4;
  return B2C_OK;
}

template <int W> static const void* fused_zc_fn_w(int kind) {
  switch (kind) {
    case B2C_OPT_SGD: return (const void*)k_fused_step_zc<B2C_OPT_SGD, W>;
    case B2C_OPT_ADAM: return (const void*)k_fused_step_zc<B2C_OPT_ADAM, W>;
    default: return (const void*)k_fused_step_zc<B2C_OPT_ADAMW, W>;
  }
}
static const void* fused_zc_fn(int kind, int W) {
  switch (W) {
    case 2: return fused_zc_fn_w<2>(kind);
    case 3: return fused_zc_fn_w<3>(kind);
    case 4: return fused_zc_fn_w<4>(kind);
    case 5: return fused_zc_fn_w<5>(kind);
    case 6: return fused_zc_fn_w<6>(kind);
    case 7: return fused_zc_fn_w<7>(kind);
    default: return fused_zc_fn_w<8>(kind);
  }
}

// grads / params: group ? per-rank pointers : only [c->rank]
static int fused_impl(b2c_comm* c, void* const* grads, void* const* params, bool group, size_t count, int dtype, float scale,
                      const b2c_opt* opt, void* const* st1, void* const* st2, void* bcast_buf, size_t bcast_bytes, int bcast_root,
                      cudaStream_t stream) {
  if (!c) return fail(B2C_EINVAL, "b2c_fused_step: comm is NULL");
  if (!c->ready) return fail(B2C_ESTATE, "b2c_fused_step: comm not imported yet");
  if (dtype != B2C_F32) return fail(B2C_EUNSUPPORTED, "b2c_fused_step: fp32 gradients and parameters only (dtype=%d)", dtype);
  int rc = check_sticky(c);
  if (rc) return rc;
  B2cOptK ok;
  rc = opt_to_k(opt, &ok);
  if (rc) return rc;
  const int W = c->world;
  const bool need1 = ok.kind != B2C_OPT_SGD || ok.momentum != 0.f, need2 = ok.kind != B2C_OPT_SGD;
  for (int r = 0; r < W; ++r) {
    if (!group && r != c->rank) continue;
    if (!grads[r] || !params[r]) return fail(B2C_EINVAL, "b2c_fused_step: NULL gradient / parameter bucket");
    if (((uintptr_t)grads[r] | (uintptr_t)params[r]) & (W == 1 ? 3u : 15u))
      return fail(B2C_EINVAL, "b2c_fused_step: buckets must be 16-byte aligned");
    if ((need1 && !st1[r]) || (need2 && !st2[r])) return fail(B2C_EINVAL, "b2c_fused_step: optimizer state buffer missing");
    if (W > 1 && ((need1 && ((uintptr_t)st1[r] & 15u)) || (need2 && ((uintptr_t)st2[r] & 15u))))
      return fail(B2C_EINVAL, "b2c_fused_step: optimizer state buffers must be 16-byte aligned");
  }
  if (W > 1 && count % 4) return fail(B2C_EINVAL, "b2c_fused_step: count=%zu must be a multiple of 4 (buckets are padded to 16 bytes)", count);
  if (bcast_bytes && (bcast_root < 0 || bcast_root >= W)) return fail(B2C_EINVAL, "b2c_fused_step: bcast_root=%d", bcast_root);
  c->stats.fused_steps++;
  if (count == 0 && !bcast_bytes) return B2C_OK;
  DeviceGuard g(c->device);

  if (W == 1) {   // everything is local: one streaming kernel
    if (count == 0) return B2C_OK;
    size_t blocks = (count / 4 + 255) / 256;
    if (blocks < 1) blocks = 1;
    if (blocks > 148 * 8) blocks = 148 * 8;
    const float* gp = (const float*)grads[c->rank];
    float* wp = (float*)params[c->rank];
    float* a1 = (float*)st1[c->rank];
    float* a2 = (float*)st2[c->rank];
    if (((uintptr_t)gp | (uintptr_t)wp | (uintptr_t)a1 | (uintptr_t)a2) & 15u) return fail(B2C_EINVAL, "b2c_fused_step: buffers must be 16-byte aligned");
    switch (ok.kind) {
      case B2C_OPT_SGD: k_fused_step_local<B2C_OPT_SGD><<<(unsigned)blocks, 256, 0, stream>>>(gp, wp, a1, a2, count, scale, ok); break;
      case B2C_OPT_ADAM: k_fused_step_local<B2C_OPT_ADAM><<<(unsigned)blocks, 256, 0, stream>>>(gp, wp, a1, a2, count, scale, ok); break;
      default: k_fused_step_local<B2C_OPT_ADAMW><<<(unsigned)blocks, 256, 0, stream>>>(gp, wp, a1, a2, count, scale, ok); break;
    }
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return cuda_fail(e, "kernel launch (fused step, local)");
    c->stats.launches++;
    c->stats.last_blocks = (uint32_t)blocks;
    c->stats.last_threads = 256;
    return B2C_OK;
  }

  // ---- through the switch: both buckets (and the buffers to broadcast) live in the multicast arena
  Nvls* nv = (!group && c->nvls && c->nvls->phase == 3) ? c->nvls : nullptr;
  auto inside = [&](const void* p, size_t bytes) {
    return nv && (const char*)p >= (const char*)nv->uc && (const char*)p + bytes <= (const char*)nv->uc + nv->bytes;
  };
  if (nv && inside(grads[c->rank], count * 4) && inside(params[c->rank], count * 4) &&
      (!bcast_bytes || (inside(bcast_buf, bcast_bytes) && bcast_bytes % 16 == 0 && ((uintptr_t)bcast_buf & 15u) == 0))) {
    B2cNvlsFusedArgs fa;
    memset(&fa, 0, sizeof(fa));
    B2cNvlsArgs& a = fa.n;
    a.uc = (char*)nv->uc; a.mc = (char*)nv->mcva;
    a.ctl = (uint32_t*)c->region; a.status = c->status_d;
    a.off = (size_t)((char*)grads[c->rank] - (char*)nv->uc);
    a.nvec = count / 4;
    a.timeout_ns = (unsigned long long)c->params[B2C_PARAM_TIMEOUT_MS] * 1000000ull;
    a.scale = scale; a.rank = c->rank; a.world = W; a.acq_poll = 1;
    fa.param_off = (size_t)((char*)params[c->rank] - (char*)nv->uc);
    fa.st1 = (float*)st1[c->rank]; fa.st2 = (float*)st2[c->rank];
    if (bcast_bytes) {
      fa.bcast_off = (size_t)((char*)bcast_buf - (char*)nv->uc);
      fa.bcast_vecs = bcast_bytes / 16;
      fa.bcast_root = bcast_root;
    }
    fa.opt = ok;
    const int threads = 512, unroll = 4;
    const size_t nvs = (a.nvec + W - 1) / W;
    const size_t per = (size_t)threads * unroll;
    long mb = c->params[B2C_PARAM_MAX_BLOCKS_NVLS] * 2;   // half the vectors in flight per thread of the plain kernel
    if (mb > B2C_MAX_BLOCKS) mb = B2C_MAX_BLOCKS;
    size_t nb = (nvs + per - 1) / per;
    if (nb < 1) nb = 1;
    if (nb > (size_t)mb) nb = mb;
    a.chunk_vecs = (uint32_t)per;
    const void* fn = ok.kind == B2C_OPT_SGD ? (const void*)k_fused_step_nvls<B2C_OPT_SGD, 4>
                   : ok.kind == B2C_OPT_ADAM ? (const void*)k_fused_step_nvls<B2C_OPT_ADAM, 4>
                                             : (const void*)k_fused_step_nvls<B2C_OPT_ADAMW, 4>;
    c->stats.nvls_launches++;
    c->stats.last_algo = B2C_ALGO_NVLS;
    return launch_pdl(c, fn, &fa, (int)nb, threads, stream);
  }

  // ---- peer to peer: registered buckets (or a local group); the C4 fold is not available here
  if (bcast_bytes) return fail(B2C_ESTATE, "b2c_fused_step: folding the buffer broadcast into the step needs the multicast arena");
  Registration* rg = group ? nullptr : find_reg(c, grads[c->rank], count * 4);
  Registration* rp = group ? nullptr : find_reg(c, params[c->rank], count * 4);
  if (!group && (!rg || !rp))
    return fail(B2C_ESTATE, "b2c_fused_step: gradient and parameter buckets must live in the multicast arena or be registered "
                            "with every peer (b2c_register_*); there is no staged form of the fused step");
  B2cFusedArgs fa;
  memset(&fa, 0, sizeof(fa));
  fill_common(c, fa.k, group);
  for (int r = 0; r < W; ++r) {
    if (group) {
      fa.k.buf[r] = grads[r]; fa.pbuf[r] = params[r];
      fa.st1[r] = (float*)st1[r]; fa.st2[r] = (float*)st2[r];
    } else {
      fa.k.buf[r] = rg->peer[r] + ((char*)grads[c->rank] - rg->local);
      fa.pbuf[r] = rp->peer[r] + ((char*)params[c->rank] - rp->local);
    }
  }
  if (!group) { fa.st1[c->rank] = (float*)st1[c->rank]; fa.st2[c->rank] = (float*)st2[c->rank]; }
  fa.k.count = count;
  fa.k.scale = scale;
  fa.opt = ok;
  Launch L;
  L.fn = fused_zc_fn(ok.kind, W);
  long mb = c->params[B2C_PARAM_MAX_BLOCKS_TWOSHOT];
  if (group) mb = group_block_cap(c, L.fn, mb);
  const size_t nvs = (count / 4 + W - 1) / W;
  geometry(c, nvs, mb, &L.nblocks, &L.threads, &fa.k.chunk_vecs);
  void* kp[] = {(void*)&fa};
  dim3 grid(L.nblocks, group ? W : 1), block(L.threads);
  cudaError_t e = group ? cudaLaunchCooperativeKernel(L.fn, grid, block, kp, 0, stream) : cudaLaunchKernel(L.fn, grid, block, kp, 0, stream);
  if (e != cudaSuccess) { cudaGetLastError(); return cuda_fail(e, "kernel launch (fused step)"); }
  c->stats.launches++;
  c->stats.twoshot_launches++;
  c->stats.zerocopy_launches++;
  c->stats.last_algo = B2C_ALGO_TWOSHOT;
  c->stats.last_blocks = L.nblocks;
  c->stats.last_threads = L.threads;
  return B2C_OK;
}

extern "C" int b2c_fused_step(b2c_comm* c, void* grad, void* param, size_t count, int dtype, float scale, const b2c_opt* opt,
                              void* state1, void* state2, void* bcast_buf, size_t bcast_bytes, int bcast_root, b2c_stream_t stream) {
  if (c && c->local_group) return fail(B2C_ESTATE, "b2c_fused_step: comm belongs to a local group; use b2c_group_fused_step");
  void* g[B2C_MAX_WORLD] = {0}; void* p[B2C_MAX_WORLD] = {0}; void* a1[B2C_MAX_WORLD] = {0}; void* a2[B2C_MAX_WORLD] = {0};
  if (c && c->rank >= 0 && c->rank < B2C_MAX_WORLD) { g[c->rank] = grad; p[c->rank] = param; a1[c->rank] = state1; a2[c->rank] = state2; }
  return fused_impl(c, g, p, false, count, dtype, scale, opt, a1, a2, bcast_buf, bcast_bytes, bcast_root, (cudaStream_t)stream);
}
extern "C" int b2c_group_fused_step(b2c_comm** comms, void** grads, void** params, size_t count, int dtype, float scale,
                                    const b2c_opt* opt, void** state1, void** state2, b2c_stream_t stream) {
  int rc = group_check(comms, "b2c_group_fused_step");
  if (rc) return rc;
  if (!grads || !params || !state1 || !state2) return fail(B2C_EINVAL, "b2c_group_fused_step: NULL argument");
  return fused_impl(comms[0], grads, params, true, count, dtype, scale, opt, state1, state2, nullptr, 0, 0, (cudaStream_t)stream);
}

// ---------------------------------------------------------------- allocation-free local allreduce
// For callers that cannot allocate where they need the reduction (see include/b2c.h): the
// scratch a local group normally gets from cudaMalloc — control words, flag page and one staging
// slot per virtual rank — lives in this file-scope device array instead.  Every GPU that loads
// the library carries it (B2C_MAX_WORLD x (16 KiB header + 256 KiB slot) = 2,228,224 bytes).
#define B2C_NOALLOC_REGION_BYTES (B2C_HEADER_BYTES + B2C_NOALLOC_MAX_BYTES)
static __device__ __align__(256) unsigned char g_noalloc_scratch[B2C_MAX_WORLD * B2C_NOALLOC_REGION_BYTES];
// sticky error words of the allocation-free entry point (code, epoch, channel, peer).  Kept apart from the
// scratch above because that is wiped at the start of every call; these survive until read and reset.
static __device__ uint32_t g_noalloc_status[4];

#define B2C_NOALLOC_MAX_DEVICES 64
static pthread_mutex_t g_noalloc_lock = PTHREAD_MUTEX_INITIALIZER;
static char* g_noalloc_addr[B2C_NOALLOC_MAX_DEVICES];   // per device: where the array lives ON that device
static uint32_t* g_noalloc_status_addr[B2C_NOALLOC_MAX_DEVICES];
static int g_noalloc_blocks[B2C_NOALLOC_MAX_DEVICES];   // per device: co-resident CTAs of the widest kernel

// Once per device, with that device current.  In host code `g_noalloc_scratch` is only a handle for
// the runtime, not a device address: the address has to be asked for, per device.
static int noalloc_setup(int dev, char** base, int* resident, uint32_t** status = nullptr) {
  if (dev < 0 || dev >= B2C_NOALLOC_MAX_DEVICES) return fail(B2C_EINVAL, "b2c_allreduce_local_noalloc: device %d out of range", dev);
  pthread_mutex_lock(&g_noalloc_lock);
  int rc = B2C_OK;
  if (!g_noalloc_addr[dev]) {
    void* p = nullptr;
    void* ps = nullptr;
    cudaError_t e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
    if (e == cudaSuccess) e = cudaGetSymbolAddress(&ps, g_noalloc_status);
    int per_sm = 0, sms = 0;
    if (e == cudaSuccess) e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_allreduce_oneshot<MODE_F32, B2C_MAX_WORLD>, 512, 0);
    if (e == cudaSuccess) e = cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    if (e != cudaSuccess || !p || !ps) {
      cudaGetLastError();
      rc = cuda_fail(e, "b2c_allreduce_local_noalloc: set-up (cudaGetSymbolAddress / occupancy)");
    } else {
      g_noalloc_blocks[dev] = per_sm * sms;
      g_noalloc_status_addr[dev] = (uint32_t*)ps;
      g_noalloc_addr[dev] = (char*)p;
    }
  }
  *base = g_noalloc_addr[dev];
  *resident = g_noalloc_blocks[dev];
  if (status) *status = g_noalloc_status_addr[dev];
  pthread_mutex_unlock(&g_noalloc_lock);
  return rc;
}

extern "C" int b2c_allreduce_local_noalloc(void* const* bufs, int world, size_t count, int dtype, float scale,
                                           b2c_stream_t stream) {
  if (!bufs) return fail(B2C_EINVAL, "b2c_allreduce_local_noalloc: bufs is NULL");
  if (world < 1 || world > B2C_MAX_WORLD) return fail(B2C_EINVAL, "b2c_allreduce_local_noalloc: world=%d outside 1..%d", world, B2C_MAX_WORLD);
  int mode, vw, ub;
  if (dtype != B2C_F32 && dtype != B2C_BF16) return fail(B2C_EUNSUPPORTED, "b2c_allreduce_local_noalloc: dtype=%d (B2C_F32 or B2C_BF16)", dtype);
  int rc = mode_of(dtype, dtype, &mode, &vw, &ub);
  if (rc) return rc;
  if (count > (size_t)B2C_NOALLOC_MAX_BYTES / ub)
    return fail(B2C_ETOOBIG, "b2c_allreduce_local_noalloc: %zu elements x %d bytes exceed the built-in scratch of %d bytes per buffer; "
                "use b2c_group_allreduce (which allocates an arena) for larger buffers", count, ub, (int)B2C_NOALLOC_MAX_BYTES);
  for (int r = 0; r < world; ++r) {
    if (!bufs[r]) return fail(B2C_EINVAL, "b2c_allreduce_local_noalloc: bufs[%d] is NULL", r);
    if ((uintptr_t)bufs[r] % ub) return fail(B2C_EINVAL, "b2c_allreduce_local_noalloc: bufs[%d] is not aligned to its element size", r);
  }
  if (count == 0) return B2C_OK;
  int dev = -1;
  CU(cudaGetDevice(&dev));
  char* base = nullptr;
  uint32_t* status = nullptr;
  int resident = 0;
  rc = noalloc_setup(dev, &base, &resident, &status);
  if (rc) return rc;
  cudaStream_t st = (cudaStream_t)stream;

  Launch L;
  memset(&L.args, 0, sizeof(L.args));
  for (int q = 0; q < world; ++q) {
    char* region = base + (size_t)q * B2C_NOALLOC_REGION_BYTES;
    // the array remembers the previous call (which may have had another world size): epochs,
    // call counter, exit ticket, error words and flags all have to start from zero
    CU(cudaMemsetAsync(region, 0, B2C_CTL_BYTES + B2C_FLAG_BYTES, st));
    L.args.region[q] = region;
    L.args.status[q] = status;  // shared by the virtual ranks, NOT wiped per call: see b2c_noalloc_status
    L.args.buf[q] = bufs[q];
  }
  L.args.count = count;
  L.args.slot_bytes = B2C_NOALLOC_MAX_BYTES;
  L.args.timeout_ns = 2000ull * 1000000ull;
  L.args.scale = scale;
  L.args.rank = -1;  // one cooperative launch runs every virtual rank (grid.y = world)
  L.args.world = world;
  L.fn = mode == MODE_F32 ? oneshot_fn<MODE_F32>(world) : oneshot_fn<MODE_BF16>(world);
  const size_t nvec = (count + vw - 1) / vw;
  long cap = resident / world;
  if (cap > 32) cap = 32;
  if (cap < 1) cap = 1;
  size_t nb = (nvec + 511) / 512;
  if (nb > (size_t)cap) nb = cap;
  if (nb < 1) nb = 1;
  size_t cu = (nvec + 512 * nb - 1) / (512 * nb);  // <= 4 here: 16384 vectors at most
  L.nblocks = (int)nb;
  L.threads = 512;
  L.args.chunk_vecs = (uint32_t)(512 * (cu < 1 ? 1 : cu));
  void* params[] = {(void*)&L.args};
  dim3 grid(L.nblocks, world), block(L.threads);
  cudaError_t e = (world > 1) ? cudaLaunchCooperativeKernel(L.fn, grid, block, params, 0, st)
                              : cudaLaunchKernel(L.fn, grid, block, params, 0, st);
  if (e != cudaSuccess) { cudaGetLastError(); return cuda_fail(e, "b2c_allreduce_local_noalloc: kernel launch"); }
  return B2C_OK;
}

// A flag-barrier timeout inside b2c_allreduce_local_noalloc cannot be returned by the call that hit it (the
// call only enqueues).  This reads the sticky words back; it waits for the device, allocates nothing, and
// with `reset` clears them so the entry point can be used again.
extern "C" int b2c_noalloc_status(uint32_t out[4], int reset) {
  int dev = -1;
  CU(cudaGetDevice(&dev));
  char* base = nullptr;
  uint32_t* status = nullptr;
  int resident = 0;
  int rc = noalloc_setup(dev, &base, &resident, &status);
  if (rc) return rc;
  uint32_t st[4] = {0, 0, 0, 0};
  CU(cudaDeviceSynchronize());
  CU(cudaMemcpy(st, status, sizeof(st), cudaMemcpyDeviceToHost));
  if (out) memcpy(out, st, sizeof(st));
  if (reset && st[0]) CU(cudaMemset(status, 0, sizeof(st)));
  if (st[0])
    return fail(-(int)st[0], "b2c_allreduce_local_noalloc: an earlier call timed out in its flag barrier (epoch %u, channel %u, "
                "waiting for virtual rank %u); the buffers of that call were left unreduced", st[1], st[2], st[3]);
  return B2C_OK;
}

// ---------------------------------------------------------------- registered buffers
typedef int (*cuMemGetAddressRange_t)(unsigned long long* pbase, size_t* psize, unsigned long long dptr);

static int alloc_base_of(void* ptr, char** base, size_t* size) {
  static cuMemGetAddressRange_t fn = nullptr;
  if (!fn) {
    void* sym = nullptr;
    cudaDriverEntryPointQueryResult qr;
    cudaError_t e = cudaGetDriverEntryPoint("cuMemGetAddressRange", &sym, cudaEnableDefault, &qr);
    if (e != cudaSuccess || !sym) { cudaGetLastError(); return fail(B2C_EUNSUPPORTED, "cuMemGetAddressRange is not available from this driver"); }
    fn = (cuMemGetAddressRange_t)sym;
  }
  unsigned long long b = 0;
  size_t sz = 0;
  int rc = fn(&b, &sz, (unsigned long long)(uintptr_t)ptr);
  if (rc != 0) return fail(B2C_EINVAL, "b2c_register: %p is not inside a CUDA allocation (driver error %d)", ptr, rc);
  *base = (char*)(uintptr_t)b;
  *size = sz;
  return B2C_OK;
}

extern "C" int b2c_register_export(b2c_comm* c, void* ptr, size_t bytes, void* blob, size_t* len) {
  if (!c || !ptr || !blob || !len) return fail(B2C_EINVAL, "b2c_register_export: NULL argument");
  if (*len < B2C_REG_BLOB_BYTES) return fail(B2C_EINVAL, "b2c_register_export: blob needs %d bytes", B2C_REG_BLOB_BYTES);
  if (c->local_group) return fail(B2C_ESTATE, "b2c_register_export: local groups are zero-copy without registration");
  DeviceGuard g(c->device);
  char* base = nullptr;
  size_t size = 0;
  int rc = alloc_base_of(ptr, &base, &size);
  if (rc) return rc;
  if ((char*)ptr + bytes > base + size) return fail(B2C_EINVAL, "b2c_register_export: [%p, +%zu) crosses the end of its allocation", ptr, bytes);
  RegBlob b;
  memset(&b, 0, sizeof(b));
  b.magic = B2C_MAGIC ^ 0x5a5a5a5au;
  b.rank = c->rank; b.pid = (int64_t)getpid(); b.device = c->device;
  b.raw_ptr = (uint64_t)(uintptr_t)ptr;
  b.offset = (uint64_t)((char*)ptr - base);
  b.bytes = bytes;
  cudaError_t e = cudaIpcGetMemHandle(&b.handle, base);
  if (e != cudaSuccess) {
    cudaGetLastError();
    return fail((int)e, "b2c_register_export: cudaIpcGetMemHandle(%p): %s — memory from cudaMallocAsync / expandable_segments "
                "cannot be exported; the staged path stays available", base, cudaGetErrorName(e));
  }
  memset(blob, 0, B2C_REG_BLOB_BYTES);
  memcpy(blob, &b, sizeof(b));
  *len = B2C_REG_BLOB_BYTES;
  return B2C_OK;
}

static int open_base(b2c_comm* c, const cudaIpcMemHandle_t& h, char** base) {
  for (int i = 0; i < c->nopened; ++i)
    if (memcmp(&c->opened[i].handle, &h, sizeof(h)) == 0) { c->opened[i].refs++; *base = c->opened[i].base; return B2C_OK; }
  void* p = nullptr;
  cudaError_t e = cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess);
  if (e != cudaSuccess) { cudaGetLastError(); return cuda_fail(e, "b2c_register_import: cudaIpcOpenMemHandle"); }
  if (c->nopened == c->opened_cap) {
    c->opened_cap = c->opened_cap ? 2 * c->opened_cap : 16;
    c->opened = (OpenedHandle*)realloc(c->opened, c->opened_cap * sizeof(OpenedHandle));
  }
  c->opened[c->nopened].handle = h;
  c->opened[c->nopened].base = (char*)p;
  c->opened[c->nopened].refs = 1;
  c->nopened++;
  *base = (char*)p;
  return B2C_OK;
}

extern "C" int b2c_register_import(b2c_comm* c, void* ptr, size_t bytes, const void* blobs, size_t len_each) {
  if (!c || !ptr || !blobs) return fail(B2C_EINVAL, "b2c_register_import: NULL argument");
  if (len_each < sizeof(RegBlob)) return fail(B2C_EINVAL, "b2c_register_import: len_each too small");
  if (!c->ready) return fail(B2C_ESTATE, "b2c_register_import: comm not imported yet");
  DeviceGuard g(c->device);
  Registration r;
  memset(&r, 0, sizeof(r));
  r.local = (char*)ptr;
  r.bytes = bytes;
  for (int q = 0; q < c->world; ++q) {
    RegBlob b;
    memcpy(&b, (const char*)blobs + (size_t)q * len_each, sizeof(b));
    if (b.magic != (B2C_MAGIC ^ 0x5a5a5a5au) || b.rank != q)
      return fail(B2C_EINVAL, "b2c_register_import: blob %d is malformed or out of rank order", q);
    if (b.bytes != bytes) return fail(B2C_EINVAL, "b2c_register_import: rank %d registered %llu bytes, we registered %zu", q, (unsigned long long)b.bytes, bytes);
    if (q == c->rank) { r.peer[q] = (char*)ptr; continue; }
    if (b.pid == (int64_t)getpid()) {
      r.peer[q] = (char*)(uintptr_t)b.raw_ptr;  // same process: peer access was enabled at comm import
    } else {
      char* base = nullptr;
      int rc = open_base(c, b.handle, &base);
      if (rc) return rc;
      r.peer[q] = base + b.offset;
    }
  }
  if (c->nregs == c->regs_cap) {
    c->regs_cap = c->regs_cap ? 2 * c->regs_cap : 16;
    c->regs = (Registration*)realloc(c->regs, c->regs_cap * sizeof(Registration));
  }
  c->regs[c->nregs++] = r;
  return B2C_OK;
}

extern "C" int b2c_deregister(b2c_comm* c, void* ptr) {
  if (!c) return fail(B2C_EINVAL, "b2c_deregister: comm is NULL");
  for (int i = 0; i < c->nregs; ++i)
    if (c->regs[i].local == (char*)ptr) {
      memmove(&c->regs[i], &c->regs[i + 1], (size_t)(c->nregs - i - 1) * sizeof(Registration));  // keeps age order
      c->nregs--;                        // peer mappings stay open until destroy (shared per allocation)
      return B2C_OK;
    }
  return fail(B2C_EINVAL, "b2c_deregister: %p is not registered", ptr);
}

// ---------------------------------------------------------------- knobs
extern "C" int b2c_set_param(b2c_comm* c, int key, long v) {
  if (!c) return fail(B2C_EINVAL, "b2c_set_param: comm is NULL");
  if (key < 0 || key >= B2C_PARAM__COUNT) return fail(B2C_EINVAL, "b2c_set_param: unknown key %d", key);
  switch (key) {
    case B2C_PARAM_THREADS:
      if (v < 128 || v > 512 || v % 32) return fail(B2C_EINVAL, "b2c_set_param: threads=%ld must be 128..512, multiple of 32", v);
      break;
    case B2C_PARAM_MAX_BLOCKS_ONESHOT:
    case B2C_PARAM_MAX_BLOCKS_TWOSHOT:
    case B2C_PARAM_MAX_BLOCKS_BCAST:
    case B2C_PARAM_MAX_BLOCKS_NVLS:
      if (v < 1 || v > B2C_MAX_BLOCKS) return fail(B2C_EINVAL, "b2c_set_param: blocks=%ld must be 1..%d", v, B2C_MAX_BLOCKS);
      break;
    case B2C_PARAM_TIMEOUT_MS:
      if (v < 1) return fail(B2C_EINVAL, "b2c_set_param: timeout_ms=%ld must be >= 1", v);
      break;
    case B2C_PARAM_TWOSHOT_VARIANT:
      if (v != 0 && v != 1) return fail(B2C_EINVAL, "b2c_set_param: twoshot variant %ld (0 = push, 1 = pull)", v);
      break;
    case B2C_PARAM_UNROLL_NVLS:
      if (v != 2 && v != 4 && v != 8 && v != 16) return fail(B2C_EINVAL, "b2c_set_param: nvls unroll %ld (2, 4, 8 or 16)", v);
      break;
    case B2C_PARAM_UNROLL_ZC:
      if (v != 0 && v != 1 && v != 2 && v != 4) return fail(B2C_EINVAL, "b2c_set_param: zero-copy unroll %ld (0, 1, 2 or 4)", v);
      break;
    case B2C_PARAM_ZEROCOPY:
    case B2C_PARAM_PDL:
    case B2C_PARAM_NVLS_LEAN:
    case B2C_PARAM_BCAST_MC:
    case B2C_PARAM_MC_ACQ_POLL:
      if (v != 0 && v != 1) return fail(B2C_EINVAL, "b2c_set_param: key %d takes 0 or 1, got %ld", key, v);
      break;
    case B2C_PARAM_THREADS_NVLS:
      if (v != 256 && v != 512) return fail(B2C_EINVAL, "b2c_set_param: nvls threads %ld (256 or 512)", v);
      break;
    case B2C_PARAM_NVLS_STAGE_BYTES:
      if (v < (64 << 10)) return fail(B2C_EINVAL, "b2c_set_param: staging slot %ld bytes (>= 64 KiB)", v);
      break;
    case B2C_PARAM_ZEROCOPY_MIN_BYTES:
    case B2C_PARAM_NVLS_MIN_BYTES:
    case B2C_PARAM_BCAST_INPLACE_MIN_BYTES:
    case B2C_PARAM_CROSSOVER_BYTES:
      if (v < 0) return fail(B2C_EINVAL, "b2c_set_param: crossover=%ld", v);
      break;
  }
  if (c->local_group && c->group) {
    for (int r = 0; r < c->world; ++r) c->group[r]->params[key] = v;
  } else {
    c->params[key] = v;
  }
  return B2C_OK;
}
extern "C" int b2c_get_param(b2c_comm* c, int key, long* v) {
  if (!c || !v) return fail(B2C_EINVAL, "b2c_get_param: NULL argument");
  if (key < 0 || key >= B2C_PARAM__COUNT) return fail(B2C_EINVAL, "b2c_get_param: unknown key %d", key);
  *v = c->params[key];
  return B2C_OK;
}
extern "C" int b2c_get_stats(b2c_comm* c, b2c_stats* out) {
  if (!c || !out) return fail(B2C_EINVAL, "b2c_get_stats: NULL argument");
  *out = c->stats;
  uint32_t st[4] = {0, 0, 0, 0};
  if (c->local_group && c->group) {
    for (int r = 0; r < c->world; ++r)
      if (c->group[r]->status_h[0]) { memcpy(st, c->group[r]->status_h, sizeof(st)); break; }
  } else if (c->status_h) {
    memcpy(st, c->status_h, sizeof(st));
  }
  out->device_status = st[0];
  out->device_status_detail[0] = st[1];
  out->device_status_detail[1] = st[2];
  out->device_status_detail[2] = st[3];
  return B2C_OK;
}
extern "C" int b2c_comm_arena(b2c_comm* c, void** base, size_t* slot_bytes) {
  if (!c) return fail(B2C_EINVAL, "b2c_comm_arena: comm is NULL");
  if (base) *base = c->region + B2C_HEADER_BYTES;
  if (slot_bytes) *slot_bytes = c->slot_bytes;
  return B2C_OK;
}
extern "C" int b2c_get_phase_ns(b2c_comm* c, uint64_t out[4]) {
  if (!c || !out) return fail(B2C_EINVAL, "b2c_get_phase_ns: NULL argument");
  DeviceGuard g(c->device);
  CU(cudaDeviceSynchronize());
  CU(cudaMemcpy(out, c->region + B2C_CTL_PHASE_U64 * 8, 4 * sizeof(uint64_t), cudaMemcpyDeviceToHost));
  return B2C_OK;
}
extern "C" int b2c_get_phase_ring(b2c_comm* c, uint64_t* out, int max_entries, uint32_t* total) {
  if (!c || !out || !total) return fail(B2C_EINVAL, "b2c_get_phase_ring: NULL argument");
  DeviceGuard g(c->device);
  CU(cudaDeviceSynchronize());
  uint32_t seq = 0;
  CU(cudaMemcpy(&seq, c->region + B2C_CTL_PHASE_SEQ * 4, 4, cudaMemcpyDeviceToHost));
  uint64_t ring[B2C_PHASE_RING * 4];
  CU(cudaMemcpy(ring, c->region + B2C_CTL_RING_BYTE, sizeof(ring), cudaMemcpyDeviceToHost));
  int n = (int)(seq < (uint32_t)B2C_PHASE_RING ? seq : (uint32_t)B2C_PHASE_RING);
  if (n > max_entries) n = max_entries;
  for (int i = 0; i < n; ++i) {   // oldest first
    const uint32_t k = (seq - (uint32_t)n + (uint32_t)i) % B2C_PHASE_RING;
    memcpy(out + (size_t)i * 4, ring + (size_t)k * 4, 32);
  }
  *total = seq;
  return n;
}
extern "C" int b2c_comm_info(b2c_comm* c, int* rank, int* world, int* device) {
  if (!c) return fail(B2C_EINVAL, "b2c_comm_info: comm is NULL");
  if (rank) *rank = c->rank;
  if (world) *world = c->world;
  if (device) *device = c->device;
  return B2C_OK;
}
