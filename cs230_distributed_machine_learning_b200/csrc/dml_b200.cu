// C ABI + orchestration of the sweep (see include/dml_b200.h).
#include "../../include/dml_b200.h"
#include "dml_common.cuh"
#include "dml_gemm.cuh"
#include "dml_lbfgs.cuh"
#include "dml_tc.cuh"
#include "dml_ridge.cuh"

#include <dlfcn.h>
#include <stdarg.h>
#include <string.h>
#include <stdlib.h>
#include <algorithm>
#include <atomic>
#include <mutex>
#include <vector>

using namespace dml;

// ---- minimal NCCL surface, resolved at run time from the libnccl the process already has ----
typedef struct ncclComm* ncclComm_t;
typedef struct { char internal[128]; } ncclUniqueId_t;
enum { kNcclSum = 0, kNcclInt64 = 4, kNcclUint64 = 5, kNcclFloat64 = 8 };
struct NcclApi {
  void* lib = nullptr;
  int (*GetUniqueId)(ncclUniqueId_t*) = nullptr;
  int (*CommInitRank)(ncclComm_t*, int, ncclUniqueId_t, int) = nullptr;
  int (*CommDestroy)(ncclComm_t) = nullptr;
  int (*AllReduce)(const void*, void*, size_t, int, int, ncclComm_t, cudaStream_t) = nullptr;
  const char* (*GetErrorString)(int) = nullptr;
  bool load() {
    if (lib) return true;
    const char* names[] = {"libnccl.so.2", "libnccl.so"};
    for (const char* nm : names) { lib = dlopen(nm, RTLD_NOW | RTLD_GLOBAL); if (lib) break; }
    if (!lib) return false;
    GetUniqueId = (decltype(GetUniqueId))dlsym(lib, "ncclGetUniqueId");
    CommInitRank = (decltype(CommInitRank))dlsym(lib, "ncclCommInitRank");
    CommDestroy = (decltype(CommDestroy))dlsym(lib, "ncclCommDestroy");
    AllReduce = (decltype(AllReduce))dlsym(lib, "ncclAllReduce");
    GetErrorString = (decltype(GetErrorString))dlsym(lib, "ncclGetErrorString");
    return GetUniqueId && CommInitRank && CommDestroy && AllReduce;
  }
};
static NcclApi g_nccl;
static char g_create_err[512] = "";

// cudaFuncSetAttribute acts on the current device's copy of a kernel, and one process may hold a ctx on
// every GPU (worker.py runs one Worker thread per GPU): the opt-ins for more than 48 KB of dynamic shared
// memory are therefore remembered per device.  Setting an attribute twice is harmless, so a flag that
// two threads race to set needs no lock.
constexpr int kMaxDevices = 64;
struct DeviceOnce {
  std::atomic<bool> done[kMaxDevices];
  DeviceOnce() { for (auto& d : done) d.store(false); }
  bool test(int dev) const { return dev >= 0 && dev < kMaxDevices && done[dev].load(std::memory_order_acquire); }
  void set(int dev) { if (dev >= 0 && dev < kMaxDevices) done[dev].store(true, std::memory_order_release); }
};

struct dml_ctx {
  int device = 0;
  cudaStream_t stream = nullptr;
  char* ws = nullptr; size_t ws_bytes = 0;
  int* h_scalar = nullptr;  // pinned
  int n_sms = 148;
  char err[512] = "";
  ncclComm_t comm = nullptr; int n_ranks = 1, rank = 0;
  std::vector<cudaEvent_t> ev_comm; size_t ev_comm_used = 0;   // pairs around each ncclAllReduce of the last call
  float comm_ms = 0.f; int comm_calls = 0; double comm_bytes = 0.0;
  // timing of the last sweep
  std::vector<cudaEvent_t> ev;  // pairs
  size_t ev_used = 0;
  cudaEvent_t ev_total[2] = {nullptr, nullptr};
  cudaEvent_t ev_round[2] = {nullptr, nullptr};   // read-back of a round's list counts (the host runs one round ahead)
  int round_ahead = 1;                             // DML_ROUND_AHEAD=0: wait for every round's counts before sizing the next
  float total_ms = 0.f, contraction_ms = 0.f; int contraction_launches = 0, total_launches = 0;
  size_t r_budget = (size_t)4 << 30;  // bytes of residual scratch per row chunk
  bool r_budget_known = false;
  size_t first_pairs = 0; float first_ms = 0.f; int64_t first_cols = 0;   // round 0 when it ran once per row mask
  // the gradient kernel can skip 64-column segments too (DML_GRAD_SKIP=1).  Measured on config 2: per launch
  // 0.85 ms against 0.76 ms (N = 64 MMAs cost more than half an N = 128 one) but a third less HBM traffic and
  // tensor work, and the same sustained sweep time at the power cap -- so it is off by default.
  int grad_skip = 0;
  int score_by_group = 1;     // scoring pass: a one-fold row tile only visits its own fold's column blocks (DML_SCORE_BY_GROUP=0: every block)
  int fold_skip = 1;          // fold-major column groups, so that row tiles of one fold can skip that fold's columns
  // gradient kernel: cross terms of the 3xTF32 product as bf16 MMAs over a K-concatenated operand
  // (DML_CROSS_BF16=1).  Measured on config 2: same accuracy (G within 1e-6 of fp64), tensor pipe 60 % -> 43 %
  // busy, launch time unchanged (the kernel is bound by the MMA warp's issue path, not by the pipe): off.
  int cross_bf16 = 0;
  int grad_zero_fill = 1;     // hole chunks of R: not stored by the forward kernel, zero-filled (not read) by the gradient kernel
  int gb_in_grad = 1;         // sum r per column in the gradient kernel instead of the forward epilogue (DML_GB_IN_GRAD=0: forward)
  int first_round_once = 1;   // evaluate round 0 once per row mask instead of once per candidate
  int tc_rps = 8192;          // most rows per gradient split (DML_TC_RPS, multiple of 128 in 2048..8192)
  int tc_prefetch = 0, tc_prefetch_grad = 0;  // L2 prefetch distance of the forward X tiles in jobs (0 = off; measured: no gain, the ring is not latency-bound)
  int tc_cluster = 2;                 // CTAs per cluster in the forward kernel (X tiles multicast): 1 or 2
  int tc_mask_hi = 0;                 // 1 = write the masked hi operand back before the MMAs read it (not needed:
                                      // kind::tf32 ignores the 13 low mantissa bits, measured identical results)
  long long slot_rounds = 0; int rounds = 0;  // sum over rounds of evaluated columns
  int last_cluster = 1;                       // what fwd_cluster decided for the last call's X (dml_last_cluster)
  // debug: scratch of the last evaluation (dml_debug_fetch)
  const void* dbg_ptr[4] = {nullptr, nullptr, nullptr, nullptr}; size_t dbg_bytes[4] = {0, 0, 0, 0};
};

static int fail(dml_ctx* c, int code, const char* fmt, ...) {
  va_list ap; va_start(ap, fmt);
  vsnprintf(c ? c->err : g_create_err, 512, fmt, ap);
  va_end(ap);
  return code;
}
#define CU(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) \
  return fail(ctx, DML_ERR_CUDA, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e_), __FILE__, __LINE__); } while (0)
#define KCHECK() CU(cudaGetLastError())

// bump allocator over the ctx arena: first pass measures, second pass hands out
struct Arena {
  char* base; size_t off = 0;
  explicit Arena(char* b) : base(b) {}
  template <class T> T* take(size_t count) {
    off = (off + 255) & ~(size_t)255;
    T* p = base ? reinterpret_cast<T*>(base + off) : nullptr;
    off += count * sizeof(T);
    return p;
  }
};
static int ensure_ws(dml_ctx* ctx, size_t bytes) {
  if (bytes <= ctx->ws_bytes) return DML_OK;
  if (ctx->ws) { CU(cudaFree(ctx->ws)); ctx->ws = nullptr; ctx->ws_bytes = 0; }
  CU(cudaMalloc(&ctx->ws, bytes));
  ctx->ws_bytes = bytes;
  return DML_OK;
}
static inline int64_t cdiv(int64_t a, int64_t b) { return (a + b - 1) / b; }
static inline int64_t rup(int64_t a, int64_t b) { return cdiv(a, b) * b; }
static inline bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }

// timing helpers.  A timing call that fails must neither fail the sweep nor leave its error behind for the
// caller's next CUDA call to trip over (the runtime keeps the last error per host thread): EVC clears it and
// says so on stderr.
#define EVC(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) { cudaGetLastError(); \
  fprintf(stderr, "[dml_b200] timing call %s failed: %s (%s:%d)\n", #call, cudaGetErrorString(e_), __FILE__, __LINE__); } } while (0)
static void tick(dml_ctx* ctx) {
  if (ctx->ev_used + 2 > ctx->ev.size()) return;
  EVC(cudaEventRecord(ctx->ev[ctx->ev_used], ctx->stream));
}
static void tock(dml_ctx* ctx, int launches) {
  ctx->contraction_launches += launches;
  if (ctx->ev_used + 2 > ctx->ev.size()) return;
  EVC(cudaEventRecord(ctx->ev[ctx->ev_used + 1], ctx->stream));
  ctx->ev_used += 2;
}
static void timing_begin(dml_ctx* ctx) {
  ctx->ev_used = 0; ctx->contraction_launches = 0; ctx->total_launches = 0;
  ctx->slot_rounds = 0; ctx->rounds = 0;
  ctx->total_ms = ctx->contraction_ms = 0.f;
  ctx->first_pairs = 0; ctx->first_ms = 0.f; ctx->first_cols = 0;
  ctx->ev_comm_used = 0; ctx->comm_ms = 0.f; ctx->comm_calls = 0; ctx->comm_bytes = 0.0;
  EVC(cudaEventRecord(ctx->ev_total[0], ctx->stream));
}
static void timing_end(dml_ctx* ctx) {
  EVC(cudaEventRecord(ctx->ev_total[1], ctx->stream));
  EVC(cudaEventSynchronize(ctx->ev_total[1]));
  EVC(cudaEventElapsedTime(&ctx->total_ms, ctx->ev_total[0], ctx->ev_total[1]));
  float acc = 0.f;
  for (size_t i = 0; i + 1 < ctx->ev_used; i += 2) {
    float ms = 0.f;
    const cudaError_t e = cudaEventElapsedTime(&ms, ctx->ev[i], ctx->ev[i + 1]);
    if (e == cudaSuccess) {
      acc += ms;
      if (i / 2 < ctx->first_pairs) ctx->first_ms += ms;
    } else {
      cudaGetLastError();
      fprintf(stderr, "[dml_b200] elapsed time of event pair %zu of %zu failed: %s\n", i / 2, ctx->ev_used / 2, cudaGetErrorString(e));
    }
  }
  ctx->contraction_ms = acc;
  for (size_t i = 0; i + 1 < ctx->ev_comm_used; i += 2) {
    float ms = 0.f;
    if (cudaEventElapsedTime(&ms, ctx->ev_comm[i], ctx->ev_comm[i + 1]) == cudaSuccess) ctx->comm_ms += ms;
    else cudaGetLastError();
  }
  // nothing of this call may stay behind in the thread's error state
  const cudaError_t stray = cudaGetLastError();
  if (stray != cudaSuccess) fprintf(stderr, "[dml_b200] a CUDA error was pending at the end of the call: %s\n", cudaGetErrorString(stray));
}

static int allreduce(dml_ctx* ctx, void* buf, size_t count, int dtype) {
  if (!ctx->comm) return DML_OK;
  const bool timed = ctx->ev_comm_used + 2 <= ctx->ev_comm.size();
  if (timed) EVC(cudaEventRecord(ctx->ev_comm[ctx->ev_comm_used], ctx->stream));
  const int rc = g_nccl.AllReduce(buf, buf, count, dtype, kNcclSum, ctx->comm, ctx->stream);
  if (rc != 0) return fail(ctx, DML_ERR_NCCL, "ncclAllReduce failed: %s", g_nccl.GetErrorString ? g_nccl.GetErrorString(rc) : "?");
  if (timed) { EVC(cudaEventRecord(ctx->ev_comm[ctx->ev_comm_used + 1], ctx->stream)); ctx->ev_comm_used += 2; }
  ctx->comm_calls += 1; ctx->comm_bytes += (double)count * 8.0;   // every reduced type here is 8 bytes wide
  return DML_OK;
}

extern "C" int dml_version(void) { return 100; }

extern "C" const char* dml_last_error(dml_ctx* ctx) { return ctx ? ctx->err : g_create_err; }

extern "C" int dml_ctx_create(int device, void* stream, dml_ctx** out) {
  dml_ctx* ctx = nullptr;
  if (!out) return fail(nullptr, DML_ERR_ARG, "dml_ctx_create: out is NULL");
  int count = 0;
  cudaError_t e = cudaGetDeviceCount(&count);
  if (e != cudaSuccess || device < 0 || device >= count)
    return fail(nullptr, DML_ERR_CUDA, "dml_ctx_create: no CUDA device %d (%s)", device,
                e != cudaSuccess ? cudaGetErrorString(e) : "index out of range");
  cudaDeviceProp prop;
  cudaGetDeviceProperties(&prop, device);
  if (prop.major != 10)
    return fail(nullptr, DML_ERR_CUDA, "dml_ctx_create: device %d is sm_%d%d; this library is built for sm_100a only",
                device, prop.major, prop.minor);
  ctx = new dml_ctx();
  ctx->device = device; ctx->stream = (cudaStream_t)stream; ctx->n_sms = prop.multiProcessorCount;
  if (const char* e = getenv("DML_TC_MASK_HI")) ctx->tc_mask_hi = atoi(e);
  if (const char* e = getenv("DML_TC_PREFETCH")) ctx->tc_prefetch = atoi(e);
  if (const char* e = getenv("DML_TC_PREFETCH_GRAD")) ctx->tc_prefetch_grad = atoi(e);
  if (const char* e = getenv("DML_FIRST_ROUND_ONCE")) ctx->first_round_once = atoi(e);
  if (const char* e = getenv("DML_FOLD_SKIP")) ctx->fold_skip = atoi(e);
  if (const char* e = getenv("DML_SCORE_BY_GROUP")) ctx->score_by_group = atoi(e);
  if (const char* e = getenv("DML_GRAD_SKIP")) ctx->grad_skip = atoi(e);
  if (const char* e = getenv("DML_CROSS_BF16")) ctx->cross_bf16 = atoi(e);
  if (const char* e = getenv("DML_ROUND_AHEAD")) ctx->round_ahead = atoi(e);
  if (const char* e = getenv("DML_GB_IN_GRAD")) ctx->gb_in_grad = atoi(e);
  if (const char* e = getenv("DML_GRAD_ZERO_FILL")) ctx->grad_zero_fill = atoi(e);
  if (const char* e = getenv("DML_TC_RPS")) { const int v = atoi(e); if (v >= 2048 && v <= 8192 && v % 128 == 0) ctx->tc_rps = v; }
  if (const char* e = getenv("DML_TC_CLUSTER")) ctx->tc_cluster = atoi(e) == 1 ? 1 : 2;
  CU(cudaSetDevice(device));
  CU(cudaHostAlloc(&ctx->h_scalar, 64, cudaHostAllocDefault));
  ctx->ev.resize(2 * 2048);
  for (auto& ev : ctx->ev) CU(cudaEventCreate(&ev));
  CU(cudaEventCreate(&ctx->ev_total[0]));
  CU(cudaEventCreate(&ctx->ev_total[1]));
  CU(cudaEventCreateWithFlags(&ctx->ev_round[0], cudaEventDisableTiming));
  CU(cudaEventCreateWithFlags(&ctx->ev_round[1], cudaEventDisableTiming));
  CU(cudaFuncSetAttribute(fwd_kernel<FWD_LOGREG>, cudaFuncAttributeMaxDynamicSharedMemorySize, kFwdSmemBytes));
  CU(cudaFuncSetAttribute(fwd_kernel<FWD_SCORE_CLS>, cudaFuncAttributeMaxDynamicSharedMemorySize, kFwdSmemBytes));
  CU(cudaFuncSetAttribute(fwd_kernel<FWD_SCORE_REG>, cudaFuncAttributeMaxDynamicSharedMemorySize, kFwdSmemBytes));
  *out = ctx;
  return DML_OK;
}

extern "C" void dml_ctx_destroy(dml_ctx* ctx) {
  if (!ctx) return;
  cudaSetDevice(ctx->device);
  cudaStreamSynchronize(ctx->stream);
  if (ctx->comm && g_nccl.CommDestroy) g_nccl.CommDestroy(ctx->comm);
  if (ctx->ws) cudaFree(ctx->ws);
  if (ctx->h_scalar) cudaFreeHost(ctx->h_scalar);
  for (auto& ev : ctx->ev) cudaEventDestroy(ev);
  for (auto& ev : ctx->ev_comm) cudaEventDestroy(ev);
  cudaEventDestroy(ctx->ev_total[0]); cudaEventDestroy(ctx->ev_total[1]);
  if (ctx->ev_round[0]) cudaEventDestroy(ctx->ev_round[0]);
  if (ctx->ev_round[1]) cudaEventDestroy(ctx->ev_round[1]);
  delete ctx;
}

extern "C" int dml_nccl_unique_id(void* id128_host) {
  if (!g_nccl.load()) return fail(nullptr, DML_ERR_NCCL, "libnccl.so.2 not found");
  ncclUniqueId_t id;
  const int rc = g_nccl.GetUniqueId(&id);
  if (rc != 0) return fail(nullptr, DML_ERR_NCCL, "ncclGetUniqueId failed (%d)", rc);
  memcpy(id128_host, &id, 128);
  return DML_OK;
}

extern "C" int dml_ctx_init_comm(dml_ctx* ctx, int n_ranks, int rank, const void* id128_host) {
  if (!ctx) return DML_ERR_ARG;
  if (!g_nccl.load()) return fail(ctx, DML_ERR_NCCL, "libnccl.so.2 not found");
  CU(cudaSetDevice(ctx->device));
  ncclUniqueId_t id; memcpy(&id, id128_host, 128);
  const int rc = g_nccl.CommInitRank(&ctx->comm, n_ranks, id, rank);
  if (rc != 0) return fail(ctx, DML_ERR_NCCL, "ncclCommInitRank failed: %s", g_nccl.GetErrorString ? g_nccl.GetErrorString(rc) : "?");
  ctx->n_ranks = n_ranks; ctx->rank = rank;
  ctx->ev_comm.resize(2 * 512);
  for (auto& ev : ctx->ev_comm) CU(cudaEventCreate(&ev));
  return DML_OK;
}

extern "C" int dml_last_comm(dml_ctx* ctx, float* allreduce_ms, int32_t* allreduce_calls, double* allreduce_bytes) {
  if (!ctx) return DML_ERR_ARG;
  if (allreduce_ms) *allreduce_ms = ctx->comm_ms;
  if (allreduce_calls) *allreduce_calls = ctx->comm_calls;
  if (allreduce_bytes) *allreduce_bytes = ctx->comm_bytes;
  return DML_OK;
}

extern "C" int dml_last_timing(dml_ctx* ctx, float* total_ms, float* contraction_ms,
                               int32_t* contraction_launches, int32_t* total_launches) {
  if (!ctx) return DML_ERR_ARG;
  if (total_ms) *total_ms = ctx->total_ms;
  if (contraction_ms) *contraction_ms = ctx->contraction_ms;
  if (contraction_launches) *contraction_launches = ctx->contraction_launches;
  if (total_launches) *total_launches = ctx->total_launches;
  return DML_OK;
}

// Debug aid: copy scratch buffer `which` (0 R (chunk-major on the tcgen05 path), 2 Gpart, 3 W_hi) of the last evaluation to the host.
extern "C" int dml_debug_fetch(dml_ctx* ctx, int which, void* dst_host, size_t bytes, size_t* avail) {
  if (!ctx || which < 0 || which > 3) return DML_ERR_ARG;
  if (avail) *avail = ctx->dbg_bytes[which];
  if (dst_host && ctx->dbg_ptr[which]) {
    CU(cudaMemcpy(dst_host, ctx->dbg_ptr[which], std::min(bytes, ctx->dbg_bytes[which]), cudaMemcpyDeviceToHost));
  }
  return DML_OK;
}

extern "C" int dml_last_first_round(dml_ctx* ctx, float* contraction_ms, int64_t* column_rounds) {
  if (!ctx) return DML_ERR_ARG;
  if (contraction_ms) *contraction_ms = ctx->first_ms;
  if (column_rounds) *column_rounds = ctx->first_cols;
  return DML_OK;
}

extern "C" int dml_last_cluster(dml_ctx* ctx) { return ctx ? ctx->last_cluster : 0; }

extern "C" int dml_last_counters(dml_ctx* ctx, int64_t* column_rounds, int32_t* rounds) {
  if (!ctx) return DML_ERR_ARG;
  if (column_rounds) *column_rounds = ctx->slot_rounds;
  if (rounds) *rounds = ctx->rounds;
  return DML_OK;
}


// ---- tcgen05 path: tensor maps + dispatch ---------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static EncodeTiledFn g_encode = nullptr;
static int make_map_2d(dml_ctx* ctx, CUtensorMap* m, const void* base, uint64_t inner, uint64_t outer,
                       uint64_t stride_bytes, uint32_t box_inner, uint32_t box_outer, bool mn_major = false) {
  if (!g_encode) {
    void* fn = nullptr;
    cudaDriverEntryPointQueryResult q;
    CU(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q));
    if (!fn || q != cudaDriverEntryPointSuccess) return fail(ctx, DML_ERR_CUDA, "cuTensorMapEncodeTiled not available");
    g_encode = (EncodeTiledFn)fn;
  }
  cuuint64_t gdim[2] = {inner, outer};
  cuuint64_t gstr[1] = {stride_bytes};
  cuuint32_t box[2] = {box_inner, box_outer};
  cuuint32_t est[2] = {1, 1};
  const CUresult r = g_encode(m, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<void*>(base), gdim, gstr, box, est,
                              CU_TENSOR_MAP_INTERLEAVE_NONE,
                              // MN-major tf32 operands need the 128B swizzle with 32-byte atoms
                              // (UMMA layout SWIZZLE_128B_BASE32B); K-major ones the plain 128B swizzle
                              mn_major ? CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B : CU_TENSOR_MAP_SWIZZLE_128B,
                              CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return fail(ctx, DML_ERR_CUDA, "cuTensorMapEncodeTiled failed (%d) inner=%llu outer=%llu stride=%llu box=%ux%u",
                                     (int)r, (unsigned long long)inner, (unsigned long long)outer,
                                     (unsigned long long)stride_bytes, box_inner, box_outer);
  return DML_OK;
}
// 3-D view (32 columns, row, 32-column chunk) of the residual scratch, which is stored chunk-major
// ([chunk][row][32 floats]) so that a box of 32 rows is one contiguous 4 KB run in HBM (row-major R made
// every box 32 separate 128-byte pieces).  A box (32, box_rows, box_chunks) lands in shared memory as
// [chunk][row][128 B], the MN-major operand layout of the gradient kernels.  Chunks past `cols` read as zero.
static int make_map_chunked(dml_ctx* ctx, CUtensorMap* m, const void* base, uint64_t cols, uint64_t rows,
                            uint32_t box_rows, uint32_t box_chunks) {
  if (!g_encode) { CUtensorMap tmp; int rc = make_map_2d(ctx, &tmp, base, 32, 1, 128, 32, 1); (void)rc; }
  if (!g_encode) return fail(ctx, DML_ERR_CUDA, "cuTensorMapEncodeTiled not available");
  // memory is [chunk][row][32 floats]: rows of a chunk 128 B apart, chunks rows * 128 B apart
  cuuint64_t gdim[3] = {32, rows, (cols + 31) / 32};
  cuuint64_t gstr[2] = {128, rows * 128};
  cuuint32_t box[3] = {32, box_rows, box_chunks};
  cuuint32_t est[3] = {1, 1, 1};
  const CUresult r = g_encode(m, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, const_cast<void*>(base), gdim, gstr, box, est,
                              CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B,
                              CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return fail(ctx, DML_ERR_CUDA, "cuTensorMapEncodeTiled(3d) failed (%d)", (int)r);
  return DML_OK;
}
// column-block width: the candidate with the least padding, larger on ties
// (a fit's T columns never straddle a block, so the width must be a multiple of T); 0 = none fits
static int pick_nc(int64_t n_slots, int T = 1) {
  const int cand[5] = {160, 128, 96, 64, 32};
  int best = 0; int64_t best_pad = INT64_MAX;
  for (int c : cand) {
    if (c % T || (T > 1 && c / T > 32)) continue;   // at most 32 fits per column block (epilogue registers)
    const int64_t pad = cdiv(n_slots, c) * c;
    if (pad < best_pad) { best_pad = pad; best = c; }
  }
  return best;
}
// tcgen05 path: TMA needs 16-byte aligned rows; multinomial fits need a block width divisible by T <= 16
static bool tc_eligible(const float* X, int64_t ldx, int T) {
  return ldx % 4 == 0 && aligned16(X) && (T == 1 || (T >= 3 && T <= 16 && pick_nc(1, T) != 0));
}

// CTAs per cluster of the forward kernel.  With 2 the X tiles are loaded with TMA multicast, and
// TMA multicast loads here read only this GPU's memory, since on B200 a TMA multicast load from
// another GPU's memory hung the GPU.  X is the caller's pointer: unless the driver says that the whole
// range the tensor map covers is device memory of the context's own GPU, the kernel runs unclustered
// and every CTA loads its own tiles (plain TMA loads).
// The tensor map covers [X, X + bytes), so the proof has to hold for every byte of that range, not
// for the first one: the range must lie inside ONE allocation made with cudaMalloc / cuMemAlloc on
// the context's GPU (such an allocation is device memory of a single GPU from end to end).  Memory
// mapped with cuMemMap (a virtual range can stitch together handles of several GPUs), pool
// allocations, IPC imports from a peer, managed and host memory, a range that straddles two
// allocations, or any failed query all take the unclustered kernel.
typedef CUresult (*PtrGetAttrFn)(void*, CUpointer_attribute, CUdeviceptr);
typedef CUresult (*MemGetRangeFn)(CUdeviceptr*, size_t*, CUdeviceptr);
static bool range_is_own_device_allocation(dml_ctx* ctx, const void* p, size_t bytes) {
  static PtrGetAttrFn get_attr = nullptr;
  static MemGetRangeFn get_range = nullptr;
  if (!get_attr || !get_range) {
    void* f1 = nullptr; void* f2 = nullptr;
    cudaDriverEntryPointQueryResult q1, q2;
    if (cudaGetDriverEntryPoint("cuPointerGetAttribute", &f1, cudaEnableDefault, &q1) != cudaSuccess ||
        cudaGetDriverEntryPoint("cuMemGetAddressRange", &f2, cudaEnableDefault, &q2) != cudaSuccess ||
        q1 != cudaDriverEntryPointSuccess || q2 != cudaDriverEntryPointSuccess || !f1 || !f2) {
      cudaGetLastError();
      return false;
    }
    get_attr = (PtrGetAttrFn)f1; get_range = (MemGetRangeFn)f2;
  }
  if (!p || bytes == 0) return false;
  const CUdeviceptr first = (CUdeviceptr)(uintptr_t)p, last = first + bytes - 1;
  if (last < first) return false;
  for (CUdeviceptr q : {first, last}) {
    cudaPointerAttributes at{};
    if (cudaPointerGetAttributes(&at, (const void*)(uintptr_t)q) != cudaSuccess) { cudaGetLastError(); return false; }
    if (at.type != cudaMemoryTypeDevice || at.device != ctx->device) return false;
    unsigned int mtype = 0; int ordinal = -1; int legacy_ipc = 0;
    if (get_attr(&mtype, CU_POINTER_ATTRIBUTE_MEMORY_TYPE, q) != CUDA_SUCCESS || mtype != CU_MEMORYTYPE_DEVICE) return false;
    if (get_attr(&ordinal, CU_POINTER_ATTRIBUTE_DEVICE_ORDINAL, q) != CUDA_SUCCESS || ordinal != ctx->device) return false;
    // true only for cuMemAlloc / cudaMalloc allocations: not for cuMemMap'd ranges, not for pool memory
    if (get_attr(&legacy_ipc, CU_POINTER_ATTRIBUTE_IS_LEGACY_CUDA_IPC_CAPABLE, q) != CUDA_SUCCESS || !legacy_ipc) return false;
  }
  CUdeviceptr base = 0, base2 = 0; size_t size = 0, size2 = 0;
  if (get_range(&base, &size, first) != CUDA_SUCCESS || get_range(&base2, &size2, last) != CUDA_SUCCESS) return false;
  if (base != base2 || size != size2) return false;              // first and last byte in the same allocation
  return first >= base && last - base < size;
}
// `bytes` = extent of the X tensor map: ((n - 1) * ldx + d) floats.
static int fwd_cluster(dml_ctx* ctx, const void* X, size_t bytes) {
  ctx->last_cluster = (ctx->tc_cluster == 2 && range_is_own_device_allocation(ctx, X, bytes)) ? 2 : 1;
  return ctx->last_cluster;
}
static inline size_t x_extent_bytes(int64_t n, int d, int64_t ldx) { return ((size_t)(n - 1) * (size_t)ldx + (size_t)d) * sizeof(float); }

template <int NC, int MODE, int CL>
static int launch_tc_fwd_nc(dml_ctx* ctx, int grid, const CUtensorMap& mx, const CUtensorMap& mwh, const CUtensorMap& mwl,
                            const tc::TcFwdParams& p) {
  static DeviceOnce attr_set;
  if (!attr_set.test(ctx->device)) {
    CU(cudaFuncSetAttribute(tc::tc_fwd_kernel<NC, MODE, CL>, cudaFuncAttributeMaxDynamicSharedMemorySize, tc::FwdSmem<NC>::kBytes));
    attr_set.set(ctx->device);
  }
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3((unsigned)grid); cfg.blockDim = dim3(tc::kFwdThreads);
  cfg.dynamicSmemBytes = tc::FwdSmem<NC>::kBytes; cfg.stream = ctx->stream;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeClusterDimension;
  at[0].val.clusterDim.x = CL; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
  cfg.attrs = at; cfg.numAttrs = 1;
  CU(cudaLaunchKernelEx(&cfg, tc::tc_fwd_kernel<NC, MODE, CL>, mx, mwh, mwl, p));
  return DML_OK;
}
// `cl` = CTAs per cluster (1 or 2); the W maps must have been built with box rows NC / cl
template <int MODE>
static int launch_tc_fwd(dml_ctx* ctx, int NC, int cl, int grid, const CUtensorMap& mx, const CUtensorMap& mwh,
                         const CUtensorMap& mwl, const tc::TcFwdParams& p) {
#define DML_FWD_CASE(N) case N: return cl == 2 ? launch_tc_fwd_nc<N, MODE, 2>(ctx, grid, mx, mwh, mwl, p) \
                                               : launch_tc_fwd_nc<N, MODE, 1>(ctx, grid, mx, mwh, mwl, p);
  switch (NC) {
    DML_FWD_CASE(160) DML_FWD_CASE(128) DML_FWD_CASE(96) DML_FWD_CASE(64)
    default: return cl == 2 ? launch_tc_fwd_nc<32, MODE, 2>(ctx, grid, mx, mwh, mwl, p)
                            : launch_tc_fwd_nc<32, MODE, 1>(ctx, grid, mx, mwh, mwl, p);
  }
#undef DML_FWD_CASE
}
// gradient kernel on CTA pairs: cluster of two along the feature-block axis (an odd block count leaves
// the last pair's second CTA on zero-filled features), R map with a one-chunk box
static int launch_tc_grad2(dml_ctx* ctx, dim3 grid, const CUtensorMap& mx, const CUtensorMap& mr, const tc::TcGradParams& p) {
  static DeviceOnce attr_set;
  if (!attr_set.test(ctx->device)) {
    CU(cudaFuncSetAttribute(tc::tc_grad2_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, tc::Grad2Smem::kBytes));
    attr_set.set(ctx->device);
  }
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = grid; cfg.blockDim = dim3(tc::kGrad2Threads);
  cfg.dynamicSmemBytes = tc::Grad2Smem::kBytes; cfg.stream = ctx->stream;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeClusterDimension;
  at[0].val.clusterDim.x = 2; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
  cfg.attrs = at; cfg.numAttrs = 1;
  CU(cudaLaunchKernelEx(&cfg, tc::tc_grad2_kernel, mx, mr, p));
  return DML_OK;
}

// ---- shared: one scoring pass over all rows for `n_slots` columns -------------------------
struct ScoreBufs { int32_t* cntPart; float* ssePart; };
static int run_score(dml_ctx* ctx, int mode, const float* X, int64_t n, int d, int64_t ldx,
                     const int32_t* ycls, const float* yreg, const int8_t* fold_id, const int8_t* hold_id,
                     const float* Wt, int ldw, const float* bt, const int32_t* sel, const int32_t* val,
                     int n_slots, int T, void* part /*[tiles][n_slots] 4-byte*/) {
  FwdParams fp{};
  fp.X = X; fp.ldx = ldx; fp.row0 = 0; fp.rows = n; fp.d = d;
  fp.xvec = (ldx % 4 == 0) && aligned16(X);
  fp.Wt = Wt; fp.ldw = ldw; fp.wvec = (ldw % 4 == 0) && aligned16(Wt); fp.bt = bt;
  fp.n_slots = n_slots; fp.T = T; fp.tile_n = (BN / T) * T;
  fp.ycls = ycls; fp.yreg = yreg; fp.fold_id = fold_id; fp.hold_id = hold_id;
  fp.slot_sel = sel; fp.slot_val = val;
  fp.cntPart = (int32_t*)part; fp.ssePart = (float*)part;
  dim3 grid((unsigned)cdiv(n, BM), (unsigned)cdiv(n_slots, fp.tile_n));
  if (grid.y > 65535) return fail(ctx, DML_ERR_LIMIT, "too many columns for one scoring launch (%d)", n_slots);
  tick(ctx);
  if (mode == FWD_SCORE_CLS) fwd_kernel<FWD_SCORE_CLS><<<grid, NT, kFwdSmemBytes, ctx->stream>>>(fp);
  else fwd_kernel<FWD_SCORE_REG><<<grid, NT, kFwdSmemBytes, ctx->stream>>>(fp);
  tock(ctx, 1);
  ctx->total_launches += 1;
  KCHECK();
  return DML_OK;
}

// mixed[0] = number of 128-row tiles whose rows do not share one fold, mixed[1 ..] = their indices in ascending
// order (one block; the order fixes which CTA scores which tile, so sums stay reproducible)
__global__ void mixed_tiles_kernel(const uint8_t* __restrict__ tile_fold, int64_t n_tiles, int n_groups, int* __restrict__ mixed) {
  __shared__ int base, warp_tot[32];
  if (threadIdx.x == 0) base = 0;
  __syncthreads();
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5, nw = blockDim.x >> 5;
  for (int64_t t0 = 0; t0 < n_tiles; t0 += blockDim.x) {
    const int64_t t = t0 + threadIdx.x;
    const int flag = (t < n_tiles) ? ((int)tile_fold[t] >= n_groups) : 0;
    const unsigned bal = __ballot_sync(0xffffffffu, flag);
    if (lane == 0) warp_tot[w] = __popc(bal);
    __syncthreads();
    int off = base;
    for (int i = 0; i < w; ++i) off += warp_tot[i];
    if (flag) mixed[1 + off + __popc(bal & ((1u << lane) - 1))] = (int)t;
    __syncthreads();
    if (threadIdx.x == 0) { int tt = 0; for (int i = 0; i < nw; ++i) tt += warp_tot[i]; base += tt; }
    __syncthreads();
  }
  if (threadIdx.x == 0) mixed[0] = base;
}

// ---- scoring pass on the tensor cores (classifier counts or squared errors) -----------------
// Launches the scoring kernel(s) for prepared parameters `fp` (labels, masks, bias, groups): builds the tensor
// maps, zeroes `part` ([<= 2 * n_sms * 4][n_slots] entries of `esz` bytes) and returns the number of partial rows.
// With fold-major column groups of width `group_width` (> 0, a multiple of 32) the pass runs by group: a row tile
// of one fold only meets its own group's blocks (tc_fwd_kernel, score_bpg), X is read once; a second small launch
// covers the tiles that mix folds.
static int score_launches(dml_ctx* ctx, int mode, int T, const float* X, int64_t n, int d, int64_t ldx,
                          const float* Whi, const float* Wlo, int ldw, int n_slots, tc::TcFwdParams fp,
                          int group_width, void* part, size_t esz, int64_t* n_parts, int* mixed = nullptr /*[tiles + 1] scratch*/) {
  cudaStream_t st = ctx->stream;
  int NC = pick_nc(n_slots, T);
  if (NC == 0) return fail(ctx, DML_ERR_ARG, "scoring: no column-block width for T=%d", T);
  const bool by_group = ctx->score_by_group && mixed != nullptr && T == 1 && group_width > 0 && fp.gstart != nullptr && fp.tile_fold != nullptr && mode != tc::FWD_SCORE_MULTI;
  if (by_group) { const int cand[5] = {160, 128, 96, 64, 32}; for (int c : cand) if (group_width % c == 0) { NC = c; break; } }
  CUtensorMap mX, mX2, mWh, mWl;
  int rc;
  if ((rc = make_map_2d(ctx, &mX, X, d, n, ldx * 4, 32, 128))) return rc;
  if ((rc = make_map_2d(ctx, &mX2, X, d, n, ldx * 4, 32, 64))) return rc;
  if ((rc = make_map_2d(ctx, &mWh, Whi, d, n_slots, (uint64_t)ldw * 4, 32, NC))) return rc;
  if ((rc = make_map_2d(ctx, &mWl, Wlo, d, n_slots, (uint64_t)ldw * 4, 32, NC))) return rc;
  fp.n_rows_total = n; fp.row0 = 0; fp.rows = n; fp.d_chunks = (int)cdiv(d, 32);
  fp.n_slots = n_slots; fp.n_col_blocks = (int)cdiv(n_slots, NC); fp.T = T; fp.loss_shift = 0.f;
  fp.mask_hi = ctx->tc_mask_hi; fp.prefetch_jobs = 0;
  const int ncb = fp.n_col_blocks;
  const int64_t tiles = cdiv(n, 128);
  auto launch = [&](int clr, int grid, const tc::TcFwdParams& q) -> int {
    const CUtensorMap& mx = clr == 2 ? mX2 : mX;
    if (mode == FWD_SCORE_REG) return launch_tc_fwd<FWD_SCORE_REG>(ctx, NC, clr, grid, mx, mWh, mWl, q);
    if (T == 1) return launch_tc_fwd<FWD_SCORE_CLS>(ctx, NC, clr, grid, mx, mWh, mWl, q);
    return launch_tc_fwd<tc::FWD_SCORE_MULTI>(ctx, NC, clr, grid, mx, mWh, mWl, q);
  };
  if (by_group) {
    fp.score_bpg = group_width / NC;
    const int64_t jobs0 = tiles * fp.score_bpg * (1 + (fp.hold_group ? 1 : 0)), jobs1 = tiles * ncb;
    const int grid0 = (int)std::min<int64_t>(jobs0, ctx->n_sms), grid1 = (int)std::min<int64_t>(jobs1, ctx->n_sms);
    CU(cudaMemsetAsync(part, 0, esz * (size_t)(grid0 + grid1) * 4 * n_slots, st));
    tick(ctx);
    fp.score_phase = 0;
    rc = launch(1, grid0, fp);
    if (!rc) {
      mixed_tiles_kernel<<<1, 1024, 0, st>>>(fp.tile_fold, tiles, fp.n_groups, mixed);
      tc::TcFwdParams q = fp;
      q.score_phase = 1; q.mixed_tiles = mixed;
      if (q.cntPart) q.cntPart += (size_t)grid0 * 4 * n_slots;
      if (q.ssePart) q.ssePart += (size_t)grid0 * 4 * n_slots;
      rc = launch(1, grid1, q);
    }
    tock(ctx, 2);
    ctx->total_launches += 2;
    if (rc) return rc;
    KCHECK();
    *n_parts = (int64_t)(grid0 + grid1) * 4;
    return DML_OK;
  }
  fp.score_bpg = 0; fp.score_phase = 0;
  const int clr = (fwd_cluster(ctx, X, x_extent_bytes(n, d, ldx)) == 2 && ncb >= 2 && (ncb % 2 == 0 || ncb >= 5)) ? 2 : 1;
  const int64_t jobs = tiles * cdiv(ncb, clr);
  const int grid = clr * (int)std::min<int64_t>(jobs, ctx->n_sms / clr);
  CU(cudaMemsetAsync(part, 0, esz * (size_t)grid * 4 * n_slots, st));
  tick(ctx);
  rc = launch(clr, grid, fp);
  tock(ctx, 1);
  ctx->total_launches += 1;
  if (rc) return rc;
  KCHECK();
  *n_parts = (int64_t)grid * 4;
  return DML_OK;
}
// Wt [n_slots][ldw] fp32 coefficients; Whi/Wlo: scratch of the same shape; part: [<= 2 * n_sms * 4][n_slots]
// int32 (classifier) or double (regressor) partials, summed by reduce_counts / reduce_sse_parts.
static int run_score_tc(dml_ctx* ctx, int mode, const float* X, int64_t n, int d, int64_t ldx,
                        const int32_t* ycls, const float* yreg, const int8_t* fold_id, const int8_t* hold_id,
                        const float* Wt, int ldw, const float* bt, const int32_t* sel, const int32_t* val,
                        int n_slots, int T, float* Whi, float* Wlo, void* part, int64_t* n_parts,
                        const uint8_t* tile_fold = nullptr, const int* gstart = nullptr, int n_groups = 0, int hold_group = 0,
                        int group_width = 0, int* mixed = nullptr) {
  cudaStream_t st = ctx->stream;
  tc::split_weights_kernel<<<(unsigned)cdiv((int64_t)n_slots * ldw, 256), 256, 0, st>>>(Wt, Whi, Wlo, (int64_t)n_slots * ldw);
  ctx->total_launches += 1;
  tc::TcFwdParams fp{};
  fp.bt = bt; fp.ycls = ycls; fp.yreg = yreg;
  fp.fold_id = fold_id; fp.hold_id = hold_id; fp.slot_sel = sel; fp.slot_val = val;
  fp.cntPart = (int32_t*)part; fp.ssePart = (double*)part;
  fp.tile_fold = tile_fold; fp.gstart = gstart; fp.n_groups = n_groups; fp.hold_group = hold_group;
  const int m = (mode == FWD_SCORE_REG) ? FWD_SCORE_REG : (T == 1 ? FWD_SCORE_CLS : tc::FWD_SCORE_MULTI);
  return score_launches(ctx, m, T, X, n, d, ldx, Whi, Wlo, ldw, n_slots, fp, group_width, part,
                        mode == FWD_SCORE_REG ? sizeof(double) : sizeof(int32_t), n_parts, mixed);
}
// sse[fit] = sum of the partials of the fit's score column (n_cols columns per partial row; see score_slot)
__global__ void reduce_sse_parts_kernel(const double* part, int64_t n_parts, int n_cols, double* sse,
                                        int n_fits, int fpc, int npad) {
  const int fit = blockIdx.x * blockDim.x + threadIdx.x;
  if (fit >= n_fits) return;
  const int col = score_slot(fit, fpc, npad);
  double a = 0.0;
  for (int64_t t = 0; t < n_parts; ++t) a += part[t * n_cols + col];
  sse[fit] = a;
}

// Forward pass of round 0 without a GEMM (binary problems, at most 32 columns): every coefficient is zero,
// so z = 0, p = 1/2 and the residual of a training row is 1/2 - y -- exactly what the tcgen05 kernel
// computes there (ex2(0) = 1, rcp(2) = 1/2, lg2(2) = 1, so loss - ln 2 = 0).  Writes the chunk-major residual
// scratch (one 32-column chunk) and per-block column sums of r.
struct ZeroRoundParams {
  int64_t row0, rows; const int32_t* y; const int8_t* fold_id; const int8_t* hold_id;
  const int32_t* sel; const int32_t* val; int n_slots;
  float* R; double* gbPart;   // R: [rows][32] (chunk 0); gbPart: [gridDim.x][n_slots]
};
__global__ void __launch_bounds__(256) zero_round_kernel(const ZeroRoundParams p) {
  __shared__ int s_sel[32], s_val[32];
  __shared__ float s_red[8][32];
  if (threadIdx.x < 32) {
    s_sel[threadIdx.x] = threadIdx.x < p.n_slots ? p.sel[threadIdx.x] : -1;
    s_val[threadIdx.x] = threadIdx.x < p.n_slots ? (p.val[threadIdx.x] & 0xff) : 0;
  }
  __syncthreads();
  float part[32];
#pragma unroll
  for (int c = 0; c < 32; ++c) part[c] = 0.f;
  // rows up to the next multiple of 128 are written too (zeros): the gradient kernel reads whole 32-row
  // stages and sums the residuals it reads
  const int64_t rows_pad = (p.rows + 127) / 128 * 128;
  for (int64_t row = (int64_t)blockIdx.x * 256 + threadIdx.x; row < rows_pad; row += (int64_t)gridDim.x * 256) {
    const bool live = row < p.rows;
    const int64_t g = p.row0 + (live ? row : 0);
    const float base = 0.5f - (float)p.y[g];
    const int fv = (int)(uint8_t)p.fold_id[g], hv = p.hold_id ? ((int)(uint8_t)p.hold_id[g] & 1) : 0;
    float out[32];
#pragma unroll
    for (int c = 0; c < 32; ++c) {
      const int sl = s_sel[c];
      const bool held = !live || sl < 0 || (sl == 0 ? fv == s_val[c] : hv == s_val[c]);
      out[c] = held ? 0.f : base;
      part[c] += out[c];            // multiples of 1/2: exact
    }
    float4* dst = reinterpret_cast<float4*>(p.R + row * 32);
#pragma unroll
    for (int q = 0; q < 8; ++q) dst[q] = make_float4(out[4 * q], out[4 * q + 1], out[4 * q + 2], out[4 * q + 3]);
  }
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
#pragma unroll
  for (int c = 0; c < 32; ++c) {
    float v = part[c];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if (lane == 0) s_red[w][c] = v;
  }
  __syncthreads();
  if (threadIdx.x < p.n_slots) {
    double a = 0.0;
    for (int i = 0; i < 8; ++i) a += (double)s_red[i][threadIdx.x];
    p.gbPart[(int64_t)blockIdx.x * p.n_slots + threadIdx.x] = a;
  }
}

// =============================================================================================
// One batched loss/gradient evaluation for `n_slots` columns: the buffers and the two code paths
// (tcgen05 3xTF32, or FFMA for shapes the tensor-core path does not take).
struct EvalCtx {
  // problem
  const float* X = nullptr; int64_t n = 0, ldx = 0; int d = 0, T = 1;
  const int32_t* y = nullptr; const int8_t* fold_id = nullptr; const int8_t* hold_id = nullptr;
  // plan
  bool use_tc = false, xvec = false; int cl = 2; int NC = 160, nt = 1, tile_n = BN, ldw = 0, ldg = 0, ldgr = 0, splits = 1;
  int64_t S_max = 0, ldr_max = 0, chunk_rows = 0, rps = 0, tc_rps = 2048, tc_rps_max = 8192, tiles_total = 0; int ctx_sms = 148;
  // buffers
  float *Wt = nullptr, *bt = nullptr, *R = nullptr, *Whi = nullptr, *Wlo = nullptr, *Gpart = nullptr;
  double *lossPart = nullptr, *gbPart = nullptr, *Gred = nullptr, *gb2 = nullptr; int32_t *sel = nullptr, *val = nullptr, *cntPart = nullptr;
  CUtensorMap mXf, mXf2, mXg, mWh, mWl, mR1;   // mR1: residuals, one 32-column chunk x 32 rows per box
  const int* gstart = nullptr; int n_groups = 0; uint8_t* tile_fold = nullptr; int* mixed = nullptr;   // mixed: list of row tiles that mix folds
  int hold_group = 0;                // scoring: column group n_groups belongs to the hold-out fits
  const int* n_list_dev = nullptr;   // device count of live list entries (n_slots passed to evaluate is then an upper bound)   // fold-major column groups of the current list (device), or null   // mXf2: forward X box of a 2-CTA cluster member (64 rows)
  unsigned long long* hist_local = nullptr;   // [512] label histogram of this rank's rows
  bool weighted = false; float* cwS = nullptr; double* swS = nullptr;   // class weights per column (write_trial_kernel)
  // the forward kernel sums (loss - shift) with shift = loss of the zero model: ln 2, or ln T (multinomial)
  float loss_shift() const { return T == 1 ? 0.693147182f : logf((float)T); }

  void plan(dml_ctx* ctx, int64_t n_groups_max) {
    ctx_sms = ctx->n_sms;
    if (!ctx->r_budget_known) {
      // residual scratch per row chunk: a third of what is free (arena included), at most 16 GB.  Asked once
      // per context: cudaMemGetInfo goes to the kernel driver and takes milliseconds on some hosts.
      size_t free_b = 0, total_b = 0;
      if (cudaMemGetInfo(&free_b, &total_b) == cudaSuccess)
        ctx->r_budget = std::max<size_t>((size_t)1 << 30, std::min<size_t>((size_t)16 << 30, (free_b + ctx->ws_bytes) / 3));
      ctx->r_budget_known = true;
    }
    cl = fwd_cluster(ctx, X, x_extent_bytes(n, d, ldx)); tc_rps_max = ctx->tc_rps;
    ldw = (int)rup(d, 4); ldg = (int)rup(d + 1, 4); ldgr = d + 2;
    tiles_total = cdiv(n, BM); tile_n = (BN / T) * T;
    NC = pick_nc(S_max, T); if (NC == 0) NC = 160;
    ldr_max = use_tc ? cdiv(S_max, NC) * NC : rup(S_max, 4);   // tcgen05: whole column blocks, so 32-column chunks stay inside a row
    // row chunk so that the residual scratch stays within budget
    chunk_rows = std::max<int64_t>(BM, (int64_t)(ctx->r_budget / (ldr_max * 4)) / BM * BM);
    chunk_rows = std::min<int64_t>(chunk_rows, rup(n, BM));
    nt = (int)cdiv(d + 1, BN);
    const int mt_max = (int)cdiv(S_max, BM);
    // short fp32 accumulation chains (<= 2048 rows), fp64 across splits: keeps the gradient noise
    // near 1e-7 relative (DESIGN.md "precision"); at least enough splits to fill the machine
    splits = (int)std::max<int64_t>(cdiv(chunk_rows, 2048),
                                    std::min<int64_t>(cdiv(2 * ctx->n_sms, (int64_t)mt_max * nt), cdiv(chunk_rows, 256)));
    rps = rup(cdiv(chunk_rows, splits), 16);
    splits = (int)cdiv(chunk_rows, rps);
    if (use_tc) splits = std::max(splits, (int)cdiv(chunk_rows, tc_rps));
    (void)n_groups_max;
  }
  void take(Arena& a, int64_t n_groups_max) {
    Gred = a.take<double>((size_t)S_max * ldgr);
    tile_fold = a.take<uint8_t>((size_t)cdiv(n, 128) + 16);
    mixed = a.take<int>((size_t)cdiv(n, 128) + 8);
    hist_local = a.take<unsigned long long>(512);
    Wt = a.take<float>((size_t)S_max * ldw); bt = a.take<float>(S_max);
    sel = a.take<int32_t>(S_max); val = a.take<int32_t>(S_max);
    if (weighted) { cwS = a.take<float>(2 * (size_t)S_max); swS = a.take<double>(S_max); }
    R = a.take<float>((size_t)chunk_rows * ldr_max);
    // FFMA path: one partial per row tile; tcgen05 path: one per (CTA, epilogue warp)
    const size_t tc_parts = (size_t)ctx_sms * 4;
    lossPart = a.take<double>(use_tc ? tc_parts * S_max : (size_t)tiles_total * n_groups_max);
    cntPart = a.take<int32_t>(use_tc ? 2 * tc_parts * S_max : (size_t)tiles_total * S_max);   // scoring may run as two launches
    Gpart = a.take<float>((size_t)splits * S_max * ldg);
    if (use_tc) {
      gbPart = a.take<double>(tc_parts * S_max);
      gb2 = a.take<double>((size_t)splits * S_max);
      Whi = a.take<float>((size_t)S_max * ldw); Wlo = a.take<float>((size_t)S_max * ldw);
    }
  }
  int make_maps(dml_ctx* ctx) {
    if (!use_tc) return DML_OK;
    int rc;
    // forward X box: 128 rows (or a cluster member's share of them) x 32 features
    if ((rc = make_map_2d(ctx, &mXf, X, d, n, ldx * 4, 32, 128))) return rc;
    if ((rc = make_map_2d(ctx, &mXf2, X, d, n, ldx * 4, 32, 64))) return rc;   // forward: 128 rows x 32 features
    if ((rc = make_map_2d(ctx, &mXg, X, d, n, ldx * 4, 32, 32, true))) return rc;    // gradient: 32 rows x 32 features
    if ((rc = make_map_2d(ctx, &mWh, Whi, d, S_max, (uint64_t)ldw * 4, 32, NC))) return rc;
    if ((rc = make_map_2d(ctx, &mWl, Wlo, d, S_max, (uint64_t)ldw * 4, 32, NC))) return rc;
    if ((rc = make_map_chunked(ctx, &mR1, R, ldr_max, chunk_rows, 32, 1))) return rc;
    return DML_OK;
  }
  // Gred[slot] = [sum r x (d), sum r, sum loss] over the training rows of each column
  // zero_w: every coefficient of the evaluated columns is zero (round 0)
  int evaluate(dml_ctx* ctx, int n_slots, bool zero_w = false) {
    cudaStream_t st = ctx->stream;
    int launches = 0;
    if (use_tc) {
      tc::split_weights_kernel<<<(unsigned)cdiv((int64_t)n_slots * ldw, 256), 256, 0, st>>>(Wt, Whi, Wlo, (int64_t)n_slots * ldw);
      ctx->total_launches += 1;
    }
    tick(ctx);
    for (int64_t r0 = 0, ci = 0; r0 < n; r0 += chunk_rows, ++ci) {
      const int64_t rows = std::min(chunk_rows, n - r0);
      if (use_tc) {
        const int ncb = (int)cdiv(n_slots, NC);
        tc::TcFwdParams fp{};
        fp.n_rows_total = n; fp.row0 = r0; fp.rows = rows; fp.d_chunks = (int)cdiv(d, 32);
        fp.n_slots = n_slots; fp.n_col_blocks = ncb; fp.bt = bt; fp.ycls = y; fp.yreg = nullptr;
        fp.fold_id = fold_id; fp.hold_id = hold_id; fp.slot_sel = sel; fp.slot_val = val;
        fp.R = R; fp.ldr = (int)ldr_max; fp.r_chunk_stride = chunk_rows * 32; fp.lossPart = lossPart; fp.gbPart = gbPart;
        fp.mask_hi = ctx->tc_mask_hi; fp.prefetch_jobs = ctx->tc_prefetch; fp.T = T; fp.loss_shift = loss_shift();
        fp.tile_fold = tile_fold; fp.gstart = gstart; fp.n_groups = n_groups;
        fp.grad_skips = ctx->grad_skip ? 1 : (ctx->grad_zero_fill ? 2 : 0);
        fp.n_list_dev = n_list_dev;
        fp.cw = weighted ? cwS : nullptr;
        // sum r per column: by the gradient kernel's split warps for binary problems (not with segment skipping or
        // bf16 cross terms, whose tile layouts differ), else by the forward epilogue
        const bool gb_grad = T == 1 && !ctx->grad_skip && !ctx->cross_bf16 && ctx->gb_in_grad;
        fp.gb_in_grad = gb_grad ? 1 : 0;
        // clusters of 2 share each X tile between two column blocks; with an odd number of blocks one
        // CTA of the last pair would run on empty columns, so small / odd cases run unclustered
        const int clr = (cl == 2 && ncb >= 2 && (ncb % 2 == 0 || ncb >= 5)) ? 2 : 1;
        const int64_t jobs = cdiv(rows, 128) * cdiv(ncb, clr);     // per cluster
        const int grid = clr * (int)std::min<int64_t>(jobs, ctx->n_sms / clr);
        CU(cudaMemsetAsync(lossPart, 0, sizeof(double) * (size_t)grid * 4 * n_slots, st));
        CU(cudaMemsetAsync(gbPart, 0, sizeof(double) * (size_t)grid * 4 * n_slots, st));
        int rc = DML_OK;
        int64_t n_parts = (int64_t)grid * 4;
        if (zero_w && T == 1 && n_slots <= 32 && ctx->first_round_once && !weighted) {
          const int zgrid = (int)std::min<int64_t>(cdiv(rows, 256), 4 * ctx->n_sms);
          ZeroRoundParams zp{r0, rows, y, fold_id, hold_id, sel, val, n_slots, R, gbPart};
          zero_round_kernel<<<zgrid, 256, 0, st>>>(zp);
          n_parts = zgrid;
        } else {
          rc = T == 1 ? launch_tc_fwd<FWD_LOGREG>(ctx, NC, clr, grid, clr == 2 ? mXf2 : mXf, mWh, mWl, fp)
                      : launch_tc_fwd<tc::FWD_LOGREG_MULTI>(ctx, NC, clr, grid, clr == 2 ? mXf2 : mXf, mWh, mWl, fp);
        }
        if (rc) return rc;
        // rows per split: the scratch is planned for 2048 (tc_rps); longer splits mean fewer CTAs (each pays
        // its pipeline fill and drain) and fewer partial gradients to reduce, as long as two waves of CTA
        // pairs remain.  Measured on config 2: 2048 / 4096 / 8192 rows -> 0.81 / 0.73 / 0.69 ms per launch.
        const int64_t pair_blocks = cdiv(n_slots, tc::kGrad2NC) * cdiv(cdiv(d, 128), 2);
        const int64_t min_splits = cdiv((int64_t)ctx->n_sms, pair_blocks);            // 2 waves of n_sms / 2 pairs
        const int64_t rps_here = std::min<int64_t>(tc_rps_max, std::max<int64_t>(tc_rps, (rows / min_splits) / 128 * 128));
        tc::TcGradParams gp{r0, rows, rps_here, d, n_slots, Gpart, ldg, ctx->tc_mask_hi, ctx->tc_prefetch_grad,
                            tile_fold, (ctx->grad_skip || ctx->grad_zero_fill) ? gstart : nullptr, n_groups, ctx->grad_skip,
                            ctx->cross_bf16, n_list_dev, T,
                            gb_grad ? gb2 : nullptr};
        const int splits_here = (int)cdiv(rows, rps_here);
        rc = launch_tc_grad2(ctx, dim3((unsigned)(2 * cdiv(cdiv(d, 128), 2)), (unsigned)cdiv(n_slots, tc::kGrad2NC), (unsigned)splits_here),
                             mXg, mR1, gp);
        if (rc) return rc;
        ReduceParams rp{Gpart, ldg, splits_here, n_slots, d, T, d, nullptr, 0, 0, Gred, ldgr, ci > 0 ? 1 : 0, n_list_dev};
        reduce_grad_kernel<<<n_slots, 256, 0, st>>>(rp);
        tc::TcReduceParams tr{lossPart, gbPart, n_parts, n_slots, T, d, Gred, ldgr, ci > 0 ? 1 : 0,
                              hist_local, sel, val, n, (double)loss_shift(), n_list_dev, gb_grad ? gb2 : nullptr, splits_here,
                              weighted ? swS : nullptr};
        tc::tc_reduce_tiles_kernel<<<(unsigned)cdiv(n_slots, 32), dim3(32, 32), 0, st>>>(tr);
      } else {
        const int ldr = (int)rup(n_slots, 4);
        FwdParams fp{};
        fp.X = X; fp.ldx = ldx; fp.row0 = r0; fp.rows = rows; fp.d = d; fp.xvec = xvec;
        fp.Wt = Wt; fp.ldw = ldw; fp.wvec = true; fp.bt = bt;
        fp.n_slots = n_slots; fp.T = T; fp.tile_n = tile_n;
        fp.ycls = y; fp.yreg = nullptr; fp.fold_id = fold_id; fp.hold_id = hold_id;
        fp.slot_sel = sel; fp.slot_val = val;
        fp.R = R; fp.ldr = ldr; fp.lossPart = lossPart;
        fp.cw = weighted ? cwS : nullptr;
        dim3 g1((unsigned)cdiv(rows, BM), (unsigned)cdiv(n_slots, tile_n));
        fwd_kernel<FWD_LOGREG><<<g1, NT, kFwdSmemBytes, st>>>(fp);
        GradParams gp{X, ldx, r0, rows, d, xvec, R, ldr, n_slots, rps, Gpart, ldg};
        const int splits_here = (int)cdiv(rows, rps);
        dim3 g2((unsigned)cdiv(n_slots, BM), (unsigned)nt, (unsigned)splits_here);
        grad_kernel<<<g2, NT, kGemmSmemBytes, st>>>(gp);
        ReduceParams rp{Gpart, ldg, splits_here, n_slots, d, T, d + 1, lossPart, r0 / BM, cdiv(rows, BM), Gred, ldgr, ci > 0 ? 1 : 0, nullptr};
        reduce_grad_kernel<<<n_slots, 256, 0, st>>>(rp);
        reduce_tiles_kernel<<<(unsigned)cdiv(n_slots / T, 32), dim3(32, 8), 0, st>>>(rp);
      }
      launches += 2;
      ctx->total_launches += 4;
    }
    tock(ctx, launches);
    KCHECK();
    ctx->dbg_ptr[0] = R; ctx->dbg_bytes[0] = (size_t)chunk_rows * ldr_max * 4;
    ctx->dbg_ptr[1] = nullptr; ctx->dbg_bytes[1] = 0;
    ctx->dbg_ptr[2] = Gpart; ctx->dbg_bytes[2] = (size_t)splits * S_max * ldg * 4;
    ctx->dbg_ptr[3] = Whi; ctx->dbg_bytes[3] = Whi ? (size_t)S_max * ldw * 4 : 0;
    return DML_OK;
  }
  // correct-prediction counts of `n_slots` columns (groups of T) on their held-out rows -> cntPart
  int score_cls(dml_ctx* ctx, int n_slots, int64_t* n_parts, int group_width = 0) {
    cudaStream_t st = ctx->stream;
    if (use_tc) {
      tc::split_weights_kernel<<<(unsigned)cdiv((int64_t)n_slots * ldw, 256), 256, 0, st>>>(Wt, Whi, Wlo, (int64_t)n_slots * ldw);
      ctx->total_launches += 1;
      tc::TcFwdParams fp{};
      fp.bt = bt; fp.ycls = y;
      fp.fold_id = fold_id; fp.hold_id = hold_id; fp.slot_sel = sel; fp.slot_val = val;
      fp.cntPart = cntPart;
      fp.tile_fold = tile_fold; fp.gstart = gstart; fp.n_groups = n_groups; fp.hold_group = hold_group;
      return score_launches(ctx, T == 1 ? (int)FWD_SCORE_CLS : (int)tc::FWD_SCORE_MULTI, T, X, n, d, ldx, Whi, Wlo, ldw, n_slots, fp,
                            group_width, cntPart, sizeof(int32_t), n_parts, mixed);
    }
    *n_parts = tiles_total;
    return run_score(ctx, FWD_SCORE_CLS, X, n, d, ldx, y, nullptr, fold_id, hold_id, Wt, ldw, bt, sel, val, n_slots, T, cntPart);
  }
};

// Round 0 starts every fit at w = 0, where the loss and gradient sums depend on the row mask only:
// `Gu` holds the sums of candidate 0's fits (fit k, class t at row k*T + t); every list entry gets the
// row of its own fit number within the candidate.
__global__ void broadcast_first_round_kernel(double* __restrict__ Gred, const double* __restrict__ Gu, int ldgr,
                                             const int* __restrict__ active_fits, int n_list, int fpc, int T) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t total = (int64_t)n_list * T * ldgr;
  if (i >= total) return;
  const int64_t slot = i / ldgr; const int j = (int)(i % ldgr);
  const int fit = active_fits[slot / T];
  if (fit < 0) return;
  Gred[i] = Gu[(int64_t)((fit % fpc) * T + (int)(slot % T)) * ldgr + j];
}
__global__ void set_int_kernel(int* p, int v) { *p = v; }
// column groups of the fold-major score layout: group k = [k * npad, (k + 1) * npad)
__global__ void score_groups_kernel(int* gstart, int n_groups, int npad) {
  if ((int)threadIdx.x <= n_groups) gstart[threadIdx.x] = (int)threadIdx.x * npad;
}
// tile_fold[t] = fold id shared by all rows of the 128-row tile t, or 255 when the tile mixes folds
__global__ void tile_fold_kernel(const int8_t* __restrict__ fold_id, int64_t n, uint8_t* __restrict__ tile_fold) {
  const int64_t t = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  if (t * 128 >= n) return;
  const int first = (int)(uint8_t)fold_id[t * 128];
  bool same = true;
  for (int i = lane; i < 128; i += 32) {
    const int64_t r = t * 128 + i;
    if (r < n && (int)(uint8_t)fold_id[r] != first) same = false;
  }
  same = __all_sync(0xffffffffu, same);
  if (lane == 0) tile_fold[t] = same ? (uint8_t)first : (uint8_t)255;
}

__global__ void fill_n_test_kernel(const unsigned long long* __restrict__ hist, int F, int fpc, int n_folds,
                                   int64_t* __restrict__ n_test) {
  const int f = blockIdx.x * blockDim.x + threadIdx.x;
  if (f >= F) return;
  const int k = f % fpc;
  n_test[f] = (int64_t)(k < n_folds ? hist[k] : hist[256 + 1]);
}

static int logreg_sweep_impl(dml_ctx* ctx, const float* X, int64_t n, int d, int64_t ldx,
                             const int32_t* y, int n_classes,
                             const int8_t* fold_id, int n_folds, const int8_t* hold_id,
                             const double* C, const double* tol, const uint8_t* fit_intercept,
                             const int32_t* max_iter, int n_cand, const double* class_weight,
                             float* coef_out, float* intercept_out,
                             int32_t* n_iter_out, int32_t* status_out, int32_t* nfev_out,
                             int64_t* correct_out, int64_t* n_test_out,
                             int32_t* rounds_out, int use_tensor_cores) {
  if (!ctx) return DML_ERR_ARG;
  if (!X || !y || !fold_id || !C || !tol || !fit_intercept || !max_iter || !coef_out || !intercept_out)
    return fail(ctx, DML_ERR_ARG, "dml_logreg_sweep: NULL argument");
  if (n <= 0 || d <= 0 || ldx < d || n_classes < 2 || n_cand <= 0 || n_folds < 0 || n_folds > 127)
    return fail(ctx, DML_ERR_ARG, "dml_logreg_sweep: bad shape (n=%lld d=%d ldx=%lld classes=%d cand=%d folds=%d)",
                (long long)n, d, (long long)ldx, n_classes, n_cand, n_folds);
  CU(cudaSetDevice(ctx->device));
  cudaGetLastError();      // an error another library left pending on this thread is not this call's
  cudaStream_t st = ctx->stream;
  const int T = n_classes == 2 ? 1 : n_classes;
  if (T > BN) return fail(ctx, DML_ERR_LIMIT, "n_classes %d > %d not supported", n_classes, BN);
  const int fpc = n_folds + (hold_id ? 1 : 0);
  if (fpc == 0) return fail(ctx, DML_ERR_ARG, "dml_logreg_sweep: no fits requested");
  const int F = n_cand * fpc, P = T * (d + 1);
  EvalCtx ev;
  ev.X = X; ev.n = n; ev.ldx = ldx; ev.d = d; ev.T = T; ev.y = y; ev.fold_id = fold_id; ev.hold_id = hold_id;
  ev.use_tc = use_tensor_cores != 0 && tc_eligible(X, ldx, T);
  ev.weighted = class_weight != nullptr;
  // fold-major lists pad every fold's group of live fits to 32 columns (binary problems on the tcgen05 path)
  const int pad_to = (ev.use_tc && T == 1 && ctx->fold_skip) ? 32 : 0;
  const int F_pad = F + (pad_to ? fpc * pad_to : 0);
  ev.S_max = (int64_t)F_pad * T;
  ev.use_tc = use_tensor_cores != 0 && tc_eligible(X, ldx, T);
  ev.xvec = (ldx % 4 == 0) && aligned16(X);
  if (use_tensor_cores == 1 && !ev.use_tc)
    return fail(ctx, DML_ERR_ARG, "tensor-core path needs a binary problem, ldx %% 4 == 0 and 16-byte aligned X");
  ev.plan(ctx, F);
  if (cdiv(ev.S_max, ev.tile_n) > 65535) return fail(ctx, DML_ERR_LIMIT, "too many columns (%lld)", (long long)ev.S_max);

  FitState* d_st; double *d_x, *d_g, *d_dir, *d_gt, *d_S, *d_Y, *d_ys;
  int *d_active, *d_nactive, *d_gstart; double* d_Gu; unsigned long long* d_hist; long long* d_correct; int32_t* d_i32;
  unsigned long long *d_hfc_loc = nullptr, *d_hfc = nullptr; float* d_cwfit = nullptr; double *d_swloc = nullptr, *d_swglob = nullptr;
  const int KC = n_classes;
  for (int pass = 0; pass < 2; ++pass) {
    Arena a(pass ? ctx->ws : nullptr);
    d_st = a.take<FitState>(F); d_i32 = a.take<int32_t>(3 * (size_t)F);
    d_x = a.take<double>((size_t)F * P); d_g = a.take<double>((size_t)F * P);
    d_dir = a.take<double>((size_t)F * P); d_gt = a.take<double>((size_t)F * P);
    d_S = a.take<double>((size_t)F * kMaxCor * P); d_Y = a.take<double>((size_t)F * kMaxCor * P);
    d_ys = a.take<double>((size_t)F * kMaxCor);
    d_active = a.take<int>(F_pad); d_nactive = a.take<int>(8); d_gstart = a.take<int>(fpc + 1);
    d_Gu = a.take<double>((size_t)fpc * T * (d + 2));
    d_hist = a.take<unsigned long long>(512); d_correct = a.take<long long>(F);
    if (class_weight) {
      d_hfc_loc = a.take<unsigned long long>(512 * (size_t)KC); d_hfc = a.take<unsigned long long>(512 * (size_t)KC);
      d_cwfit = a.take<float>((size_t)F * KC); d_swloc = a.take<double>(F); d_swglob = a.take<double>(F);
    }
    ev.take(a, F);
    if (!pass) { int rc = ensure_ws(ctx, a.off + 256); if (rc) return rc; }
  }
  { int rc = ev.make_maps(ctx); if (rc) return rc; }
  timing_begin(ctx);
  CU(cudaMemsetAsync(d_x, 0, sizeof(double) * (size_t)F * P, st));
  CU(cudaMemsetAsync(d_g, 0, sizeof(double) * (size_t)F * P, st));
  CU(cudaMemsetAsync(d_dir, 0, sizeof(double) * (size_t)F * P, st));
  CU(cudaMemsetAsync(ev.hist_local, 0, sizeof(unsigned long long) * 512, st));
  label_hist_kernel<<<(unsigned)std::min<int64_t>(cdiv(n, 256), 4 * ctx->n_sms), 256, 0, st>>>(fold_id, hold_id, n, ev.hist_local);
  KCHECK();
  CU(cudaMemcpyAsync(d_hist, ev.hist_local, sizeof(unsigned long long) * 512, cudaMemcpyDeviceToDevice, st));
  if (pad_to) {
    tile_fold_kernel<<<(unsigned)cdiv(cdiv(n, 128), 8), 256, 0, st>>>(fold_id, n, ev.tile_fold);
    KCHECK();
    ctx->total_launches += 1;
  }
  int64_t n_total = n;
  if (ctx->comm) {
    int rc = allreduce(ctx, d_hist, 512, kNcclUint64); if (rc) return rc;
    std::vector<unsigned long long> h(256);   // total rows = sum of the fold histogram
    CU(cudaMemcpyAsync(h.data(), d_hist, 256 * sizeof(unsigned long long), cudaMemcpyDeviceToHost, st));
    CU(cudaStreamSynchronize(st));
    n_total = 0; for (auto v : h) n_total += (int64_t)v;
  }
  if (class_weight) {      // weighted training-row counts per fit, on this rank (loss shift) and over all ranks (1 / sum s_i)
    CU(cudaMemsetAsync(d_hfc_loc, 0, sizeof(unsigned long long) * 512 * (size_t)KC, st));
    label_class_hist_kernel<<<(unsigned)std::min<int64_t>(cdiv(n, 256), 2 * ctx->n_sms), 256, 0, st>>>(fold_id, hold_id, y, n, KC, d_hfc_loc);
    KCHECK();
    CU(cudaMemcpyAsync(d_hfc, d_hfc_loc, sizeof(unsigned long long) * 512 * (size_t)KC, cudaMemcpyDeviceToDevice, st));
    if (ctx->comm) { int rc = allreduce(ctx, d_hfc, 512 * (size_t)KC, kNcclUint64); if (rc) return rc; }
    fit_weights_kernel<<<(unsigned)cdiv(F, 128), 128, 0, st>>>(class_weight, F, fpc, n_folds, KC, d_hfc_loc, d_hfc, d_cwfit, d_swloc, d_swglob);
    KCHECK();
    ctx->total_launches += 2;
  }
  CU(cudaMemsetAsync(d_nactive, 0, 8 * sizeof(int), st));
  InitParams ip{F, n_folds, hold_id ? 1 : 0, fpc, n_total, d_hist, C, tol, fit_intercept, max_iter, d_st, d_active, d_nactive,
                class_weight ? d_swglob : nullptr};
  init_fits_kernel<<<(unsigned)cdiv(F, 128), 128, 0, st>>>(ip);
  KCHECK();
  TrialParams tp{d, T, P, ev.ldw, d_st, d_x, d_dir, d_active, d_nactive, ev.Wt, ev.bt, ev.sel, ev.val,
                 class_weight ? d_cwfit : nullptr, KC, d_swloc, ev.cwS, ev.swS};
  CompactParams cp{d_st, F, fpc, pad_to, d_active, d_nactive, d_gstart, d_nactive + 4};
  LbfgsParams lp{d, T, P, ev.ldgr, d_st, d_x, d_g, d_dir, d_gt, d_S, d_Y, d_ys, d_active, d_nactive, ev.Gred};
  ctx->total_launches += 1;
  int n_list = 0, n_real = 0, rounds = 0;
  auto next_list = [&]() -> int {      // live fits -> list of the next round; trial weights in that layout
    compact_groups_kernel<<<1, 1024, 0, st>>>(cp);
    write_trial_kernel<<<F_pad, 128, 0, st>>>(tp);
    KCHECK();
    CU(cudaMemcpyAsync(ctx->h_scalar, d_nactive + 4, 3 * sizeof(int), cudaMemcpyDeviceToHost, st));
    CU(cudaStreamSynchronize(st));
    n_list = ctx->h_scalar[0]; n_real = ctx->h_scalar[1];
    ev.gstart = ctx->h_scalar[2] ? d_gstart : nullptr; ev.n_groups = n_folds;   // only CV folds own row tiles (group n_folds = hold-out split)
    ctx->total_launches += 2;
    return DML_OK;
  };
  auto evaluate_list = [&]() -> int {
    ctx->slot_rounds += (int64_t)n_real * T; ctx->rounds += 1; ++rounds;
    int rc = ev.evaluate(ctx, n_list * T);
    if (rc) return rc;
    if (ctx->comm) rc = allreduce(ctx, ev.Gred, (size_t)n_list * T * ev.ldgr, kNcclFloat64);
    return rc;
  };
  // round 0: all coefficients are zero (_logistic.py:439,457), so z = 0 for every column and the sums
  // differ only by row mask -- evaluate candidate 0's fits (fits 0 .. fpc-1) and copy the sums
  if (ctx->first_round_once && n_cand > 1 && !class_weight) {
    set_int_kernel<<<1, 1, 0, st>>>(d_nactive, fpc);
    write_trial_kernel<<<fpc, 128, 0, st>>>(tp);
    KCHECK();
    ev.gstart = nullptr;
    ctx->slot_rounds += (int64_t)fpc * T; ctx->rounds += 1; ++rounds;
    int rc = ev.evaluate(ctx, fpc * T, true);
    if (rc) return rc;
    if (ctx->comm) { rc = allreduce(ctx, ev.Gred, (size_t)fpc * T * ev.ldgr, kNcclFloat64); if (rc) return rc; }
    ctx->first_pairs = ctx->ev_used / 2; ctx->first_cols = (int64_t)fpc * T;
    CU(cudaMemcpyAsync(d_Gu, ev.Gred, sizeof(double) * (size_t)fpc * T * ev.ldgr, cudaMemcpyDeviceToDevice, st));
    rc = next_list();
    if (rc) return rc;
    const int64_t total = (int64_t)n_list * T * ev.ldgr;
    broadcast_first_round_kernel<<<(unsigned)cdiv(total, 256), 256, 0, st>>>(ev.Gred, d_Gu, ev.ldgr, d_active, n_list, fpc, T);
    KCHECK();
    ctx->total_launches += 4;
  } else {
    int rc = next_list();
    if (rc) return rc;
    rc = evaluate_list();
    if (rc) return rc;
  }
  if (!(ev.use_tc && ctx->round_ahead)) {
    while (true) {
      lbfgs_update_kernel<<<n_list, 256, 0, st>>>(lp);
      ctx->total_launches += 1;
      int rc = next_list();
      if (rc) return rc;
      if (n_real == 0) break;
      rc = evaluate_list();
      if (rc) return rc;
      if (rounds > 200000) return fail(ctx, DML_ERR_LIMIT, "dml_logreg_sweep: round cap hit");
    }
  } else {
    // The host runs one round ahead of the GPU.  List k+1 (made by the k-th compaction) is evaluated with grids
    // sized from the length of list k -- the newest length whose read-back is certain to have arrived without
    // waiting -- and the kernels take the true length from device memory (EvalCtx::n_list_dev).  A list never
    // grows (compact_groups_kernel keeps a dense list dense), so that length is an upper bound.  The sweep ends
    // one read-back late, after one round of kernels that find an empty list.
    ev.n_list_dev = d_nactive;
    ev.gstart = pad_to ? d_gstart : nullptr; ev.n_groups = n_folds;   // the device table says whether the list is grouped
    int bound = n_list;
    for (int k = 0;; ++k) {
      lbfgs_update_kernel<<<bound, 256, 0, st>>>(lp);
      compact_groups_kernel<<<1, 1024, 0, st>>>(cp);
      write_trial_kernel<<<F_pad, 128, 0, st>>>(tp);
      KCHECK();
      int* h_next = ctx->h_scalar + 4 * ((k + 1) & 1);
      CU(cudaMemcpyAsync(h_next, d_nactive + 4, 3 * sizeof(int), cudaMemcpyDeviceToHost, st));
      CU(cudaEventRecord(ctx->ev_round[(k + 1) & 1], st));
      ctx->total_launches += 3;
      if (k >= 1) {            // counts of list k: the list whose evaluation was queued in the previous iteration
        CU(cudaEventSynchronize(ctx->ev_round[k & 1]));
        const int* h = ctx->h_scalar + 4 * (k & 1);
        if (h[1] == 0) break;
        ctx->slot_rounds += (int64_t)h[1] * T; ctx->rounds += 1; ++rounds;
        bound = h[0];
      }
      int rc = ev.evaluate(ctx, bound * T);
      if (rc) return rc;
      if (ctx->comm) { rc = allreduce(ctx, ev.Gred, (size_t)bound * T * ev.ldgr, kNcclFloat64); if (rc) return rc; }
      if (rounds > 200000) return fail(ctx, DML_ERR_LIMIT, "dml_logreg_sweep: round cap hit");
    }
    ev.n_list_dev = nullptr;
  }
  // results + score pass on every fit's held-out rows.  With fold-major score columns (binary problems on
  // the tcgen05 path, at least two folds) a row tile of fold k only multiplies and counts group k's columns
  // and the hold-out group's: a fifth of the work at cv = 5.
  const int npad = (pad_to && n_folds >= 2) ? (int)rup(n_cand, 32) : 0;
  const int n_score = npad ? fpc * npad : F * T;
  if (npad) {      // padding columns: zero weights, mask -1 (no row is scored on them)
    CU(cudaMemsetAsync(ev.Wt, 0, sizeof(float) * (size_t)n_score * ev.ldw, st));
    CU(cudaMemsetAsync(ev.bt, 0, sizeof(float) * (size_t)n_score, st));
    CU(cudaMemsetAsync(ev.sel, 0xff, sizeof(int32_t) * (size_t)n_score, st));
    CU(cudaMemsetAsync(ev.val, 0, sizeof(int32_t) * (size_t)n_score, st));
    score_groups_kernel<<<1, 160, 0, st>>>(d_gstart, fpc, npad);
  }
  FinalParams fin{F, d, T, P, ev.ldw, d_st, d_x, coef_out, intercept_out, n_iter_out ? n_iter_out : d_i32,
                  status_out ? status_out : d_i32 + F, nfev_out ? nfev_out : d_i32 + 2 * F,
                  ev.Wt, ev.bt, ev.sel, ev.val, fpc, npad};
  finalize_kernel<<<F, 128, 0, st>>>(fin);
  KCHECK();
  ctx->total_launches += 1;
  if (correct_out || n_test_out) {
    int64_t n_parts = 0;
    ev.gstart = npad ? d_gstart : nullptr; ev.n_groups = n_folds; ev.hold_group = hold_id ? 1 : 0;
    int rc = ev.score_cls(ctx, n_score, &n_parts, npad);
    ev.gstart = nullptr;
    if (rc) return rc;
    reduce_counts_kernel<<<F, 256, 0, st>>>(ev.cntPart, n_parts, n_score, T, d_correct, fpc, npad);
    KCHECK();
    if (ctx->comm) { rc = allreduce(ctx, d_correct, F, kNcclInt64); if (rc) return rc; }
    if (correct_out) CU(cudaMemcpyAsync(correct_out, d_correct, sizeof(int64_t) * F, cudaMemcpyDeviceToDevice, st));
    if (n_test_out) {   // n_test of fit (cand, k) = histogram entry of its mask
      fill_n_test_kernel<<<(unsigned)cdiv(F, 256), 256, 0, st>>>(d_hist, F, fpc, n_folds, n_test_out);
      KCHECK();
    }
    ctx->total_launches += 1;
  }
  timing_end(ctx);
  CU(cudaStreamSynchronize(st));
  if (rounds_out) *rounds_out = rounds;
  return DML_OK;
}

extern "C" int dml_logreg_sweep(dml_ctx* ctx, const float* X, int64_t n, int d, int64_t ldx,
                                const int32_t* y, int n_classes,
                                const int8_t* fold_id, int n_folds, const int8_t* hold_id,
                                const double* C, const double* tol, const uint8_t* fit_intercept,
                                const int32_t* max_iter, int n_cand,
                                float* coef_out, float* intercept_out,
                                int32_t* n_iter_out, int32_t* status_out, int32_t* nfev_out,
                                int64_t* correct_out, int64_t* n_test_out,
                                int32_t* rounds_out, int use_tensor_cores) {
  return logreg_sweep_impl(ctx, X, n, d, ldx, y, n_classes, fold_id, n_folds, hold_id, C, tol, fit_intercept, max_iter, n_cand,
                           nullptr, coef_out, intercept_out, n_iter_out, status_out, nfev_out, correct_out, n_test_out,
                           rounds_out, use_tensor_cores);
}

extern "C" int dml_logreg_sweep_weighted(dml_ctx* ctx, const float* X, int64_t n, int d, int64_t ldx,
                                         const int32_t* y, int n_classes,
                                         const int8_t* fold_id, int n_folds, const int8_t* hold_id,
                                         const double* C, const double* tol, const uint8_t* fit_intercept,
                                         const int32_t* max_iter, int n_cand, const double* class_weight,
                                         float* coef_out, float* intercept_out,
                                         int32_t* n_iter_out, int32_t* status_out, int32_t* nfev_out,
                                         int64_t* correct_out, int64_t* n_test_out,
                                         int32_t* rounds_out, int use_tensor_cores) {
  if (ctx && !class_weight) return fail(ctx, DML_ERR_ARG, "dml_logreg_sweep_weighted: class_weight is NULL");
  return logreg_sweep_impl(ctx, X, n, d, ldx, y, n_classes, fold_id, n_folds, hold_id, C, tol, fit_intercept, max_iter, n_cand,
                           class_weight, coef_out, intercept_out, n_iter_out, status_out, nfev_out, correct_out, n_test_out,
                           rounds_out, use_tensor_cores);
}

// Loss and gradient sums of given coefficient columns (binary logistic): one batched evaluation.
extern "C" int dml_logreg_eval(dml_ctx* ctx, const float* X, int64_t n, int d, int64_t ldx, const int32_t* y,
                               const int8_t* fold_id, const int8_t* hold_id,
                               const float* W, const float* b, const int32_t* mask_sel, const int32_t* mask_val,
                               int n_slots, double* out, int use_tensor_cores) {
  if (!ctx) return DML_ERR_ARG;
  if (!X || !y || !fold_id || !W || !b || !mask_sel || !mask_val || !out || n <= 0 || d <= 0 || ldx < d || n_slots <= 0)
    return fail(ctx, DML_ERR_ARG, "dml_logreg_eval: bad argument");
  CU(cudaSetDevice(ctx->device));
  cudaStream_t st = ctx->stream;
  EvalCtx ev;
  ev.X = X; ev.n = n; ev.ldx = ldx; ev.d = d; ev.T = 1; ev.y = y; ev.fold_id = fold_id; ev.hold_id = hold_id;
  ev.S_max = n_slots;
  ev.use_tc = use_tensor_cores != 0 && tc_eligible(X, ldx, 1);
  ev.xvec = (ldx % 4 == 0) && aligned16(X);
  if (use_tensor_cores == 1 && !ev.use_tc) return fail(ctx, DML_ERR_ARG, "tensor-core path needs ldx %% 4 == 0 and aligned X");
  ev.plan(ctx, n_slots);
  for (int pass = 0; pass < 2; ++pass) {
    Arena a(pass ? ctx->ws : nullptr);
    ev.take(a, n_slots);
    if (!pass) { int rc = ensure_ws(ctx, a.off + 256); if (rc) return rc; }
  }
  { int rc = ev.make_maps(ctx); if (rc) return rc; }
  timing_begin(ctx);
  CU(cudaMemcpy2DAsync(ev.Wt, (size_t)ev.ldw * 4, W, (size_t)d * 4, (size_t)d * 4, n_slots, cudaMemcpyDeviceToDevice, st));
  if (ev.ldw > d) {
    // pad columns are never read by the FFMA path; the TMA path reads them (zero them)
    CU(cudaMemset2DAsync(ev.Wt + d, (size_t)ev.ldw * 4, 0, (size_t)(ev.ldw - d) * 4, n_slots, st));
  }
  CU(cudaMemcpyAsync(ev.bt, b, sizeof(float) * n_slots, cudaMemcpyDeviceToDevice, st));
  CU(cudaMemcpyAsync(ev.sel, mask_sel, sizeof(int32_t) * n_slots, cudaMemcpyDeviceToDevice, st));
  CU(cudaMemcpyAsync(ev.val, mask_val, sizeof(int32_t) * n_slots, cudaMemcpyDeviceToDevice, st));
  CU(cudaMemsetAsync(ev.hist_local, 0, sizeof(unsigned long long) * 512, st));
  label_hist_kernel<<<(unsigned)std::min<int64_t>(cdiv(n, 256), 4 * ctx->n_sms), 256, 0, st>>>(fold_id, hold_id, n, ev.hist_local);
  int rc = ev.evaluate(ctx, n_slots);
  if (rc) return rc;
  if (ctx->comm) { rc = allreduce(ctx, ev.Gred, (size_t)n_slots * ev.ldgr, kNcclFloat64); if (rc) return rc; }
  CU(cudaMemcpyAsync(out, ev.Gred, sizeof(double) * (size_t)n_slots * ev.ldgr, cudaMemcpyDeviceToDevice, st));
  timing_end(ctx);
  CU(cudaStreamSynchronize(st));
  return DML_OK;
}

// =============================================================================================
// per-fit statistics of the held-out targets, fp64, deterministic: kStatSlices blocks per label value, each over a
// contiguous slice of the rows, partial sums combined in slice order
constexpr int kStatSlices = 64;
__global__ void target_stats_kernel(const float* y, const int8_t* label, int64_t n, double* part /*[labels][slices][2]*/) {
  __shared__ double red[33];
  const int want = blockIdx.x, sl = blockIdx.y;
  const int64_t per = (n + kStatSlices - 1) / kStatSlices, r0 = sl * per, r1 = min(n, r0 + per);
  double a = 0.0, b = 0.0;
  for (int64_t r = r0 + threadIdx.x; r < r1; r += blockDim.x)
    if (label[r] == want) { const double v = (double)y[r]; a += v; b += v * v; }
  a = block_reduce<false>(a, red);
  b = block_reduce<false>(b, red);
  if (threadIdx.x == 0) { part[((int64_t)want * kStatSlices + sl) * 2] = a; part[((int64_t)want * kStatSlices + sl) * 2 + 1] = b; }
}
__global__ void target_stats_sum_kernel(const double* part, double* sy, double* syy) {
  const int want = blockIdx.x;
  if (threadIdx.x != 0) return;
  double a = 0.0, b = 0.0;
  for (int sl = 0; sl < kStatSlices; ++sl) { a += part[((int64_t)want * kStatSlices + sl) * 2]; b += part[((int64_t)want * kStatSlices + sl) * 2 + 1]; }
  sy[want] = a; syy[want] = b;
}
__global__ void expand_stats_kernel(const double* sy_l, const double* syy_l, const double* sy_h, const double* syy_h,
                                    int n_fits, int fpc, int n_folds, double* sy, double* syy) {
  const int f = blockIdx.x * blockDim.x + threadIdx.x;
  if (f >= n_fits) return;
  const int k = f % fpc;
  if (sy) sy[f] = k < n_folds ? sy_l[k] : sy_h[1];
  if (syy) syy[f] = k < n_folds ? syy_l[k] : syy_h[1];
}

extern "C" int dml_ridge_sweep(dml_ctx* ctx, const float* X, int64_t n, int d, int64_t ldx,
                               const float* y_reg, const int32_t* y_cls, int n_classes, int is_classifier,
                               const int8_t* fold_id, int n_folds, const int8_t* hold_id,
                               const double* alpha, const uint8_t* fit_intercept, int n_cand,
                               float* coef_out, float* intercept_out,
                               int64_t* correct_out, int64_t* n_test_out,
                               double* sse_out, double* sy_out, double* syy_out,
                               int32_t* info_out, int use_tensor_cores) {
  if (!ctx) return DML_ERR_ARG;
  if (!X || !fold_id || !alpha || !fit_intercept || !coef_out || !intercept_out || (is_classifier ? !y_cls : !y_reg))
    return fail(ctx, DML_ERR_ARG, "dml_ridge_sweep: NULL argument");
  if (n <= 0 || d <= 0 || ldx < d || n_cand <= 0 || n_folds < 0 || n_folds > 127 || (is_classifier && n_classes < 2))
    return fail(ctx, DML_ERR_ARG, "dml_ridge_sweep: bad shape");
  CU(cudaSetDevice(ctx->device));
  cudaGetLastError();
  cudaStream_t st = ctx->stream;
  const int T = is_classifier ? (n_classes == 2 ? 1 : n_classes) : 1;
  if (T > BN) return fail(ctx, DML_ERR_LIMIT, "n_classes %d > %d not supported", n_classes, BN);
  const int has_hold = hold_id ? 1 : 0, fpc = n_folds + has_hold;
  if (fpc == 0) return fail(ctx, DML_ERR_ARG, "dml_ridge_sweep: no fits requested");
  const int F = n_cand * fpc, D = d + T + 1, ldm = (int)rup(D, 4), ldw = (int)rup(d, 4);
  const int64_t S_all = (int64_t)F * T, tiles_total = cdiv(n, BM);
  const int tile_n = (BN / T) * T;
  if (cdiv(S_all, tile_n) > 65535) return fail(ctx, DML_ERR_LIMIT, "too many columns (%lld)", (long long)S_all);
  // tcgen05 Gram: 128 x 128 feature blocks (upper block triangle), long row splits (the drain warps
  // keep round-to-nearest fp32 sums over 128-row tensor-core chains), first-order moments by a
  // separate HBM-bound kernel
  const bool use_tc = use_tensor_cores != 0 && ldx % 4 == 0 && aligned16(X) && n_folds <= 16 && d <= 1024 && T <= 16;
  if (use_tensor_cores == 1 && !use_tc)
    return fail(ctx, DML_ERR_ARG, "tensor-core Gram needs ldx %% 4 == 0, aligned X, n_folds <= 16, d <= 1024, <= 16 targets");
  const int nb = (int)cdiv(d, 128), ldm_tc = nb * 128;
  const int64_t rps = use_tc ? 32768 : 16384;
  const int splits = (int)cdiv(n, rps);
  const int64_t xty_rows = 4096; const int xty_chunks = (int)cdiv(n, xty_rows);
  const int gfolds = std::max(n_folds, 2);  // the hold-out pass uses labels {0,1}
  // candidates per solve batch: keep the fp64 systems under ~2 GB
  int cand_chunk = (int)std::max<int64_t>(1, ((int64_t)2 << 30) / ((int64_t)fpc * d * d * 16));   // A and L^T
  cand_chunk = std::min(cand_chunk, n_cand);

  float *d_Mpart, *d_Wt, *d_bt, *d_xty = nullptr, *d_Whi = nullptr, *d_Wlo = nullptr; void* d_part; uint16_t* d_bmask;
  double* d_part_tc = nullptr;
  // scoring on the tensor cores needs a column-block width that holds whole fits
  const bool score_tc = use_tc && (T == 1 || (T >= 3 && pick_nc(1, T) != 0));
  // fold-major score columns (see dml_logreg_sweep): group k = the fits that hold out fold k, padded to 32
  const int npad = (score_tc && T == 1 && ctx->fold_skip && n_folds >= 2) ? (int)rup(n_cand, 32) : 0;
  const int64_t S_score = npad ? (int64_t)fpc * npad : S_all;
  uint8_t* d_tfold = nullptr; int* d_gs = nullptr; int* d_mixed = nullptr;
  double *d_Mfold, *d_Mtot, *d_A, *d_AT, *d_B, *d_mu, *d_stats, *d_sse, *d_statpart; int32_t *d_sel, *d_val, *d_info; long long* d_cnt;
  for (int pass = 0; pass < 2; ++pass) {
    Arena a(pass ? ctx->ws : nullptr);
    d_Mpart = a.take<float>(use_tc ? (size_t)gfolds * splits * ldm_tc * ldm_tc : (size_t)gfolds * splits * D * ldm);
    if (use_tc) d_xty = a.take<float>((size_t)gfolds * xty_chunks * (d + 2) * (1 + T));
    d_Mfold = a.take<double>((size_t)(n_folds + 2) * D * D);
    d_Mtot = a.take<double>((size_t)D * D);
    d_bmask = a.take<uint16_t>((size_t)cdiv(n, 16) + 16);
    d_A = a.take<double>((size_t)cand_chunk * fpc * d * d);
    d_AT = a.take<double>((size_t)cand_chunk * fpc * d * d);
    d_B = a.take<double>((size_t)cand_chunk * fpc * T * d);
    d_mu = a.take<double>((size_t)cand_chunk * fpc * (d + T));
    d_Wt = a.take<float>((size_t)S_score * ldw); d_bt = a.take<float>(S_score);
    d_sel = a.take<int32_t>(S_score); d_val = a.take<int32_t>(S_score);
    d_tfold = a.take<uint8_t>((size_t)cdiv(n, 128) + 16); d_gs = a.take<int>(fpc + 1);
    d_mixed = a.take<int>((size_t)cdiv(n, 128) + 8);
    d_info = a.take<int32_t>(F); d_cnt = a.take<long long>(F);
    d_stats = a.take<double>(4 * 256); d_sse = a.take<double>(S_all);
    d_statpart = a.take<double>(256 * (size_t)kStatSlices * 2);
    d_part = a.take<float>((size_t)tiles_total * S_all);
    if (score_tc) {
      d_Whi = a.take<float>((size_t)S_score * ldw); d_Wlo = a.take<float>((size_t)S_score * ldw);
      d_part_tc = a.take<double>((size_t)ctx->n_sms * 8 * S_score);
    }
    if (!pass) { int rc = ensure_ws(ctx, a.off + 256); if (rc) return rc; }
  }
  CUtensorMap mXg;
  if (use_tc) { int rc = make_map_2d(ctx, &mXg, X, d, n, ldx * 4, 32, 32, true); if (rc) return rc; }
  static DeviceOnce gram_attr;
  if (use_tc && !gram_attr.test(ctx->device)) {
    CU(cudaFuncSetAttribute(tc::tc_gram_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, tc::GramSmem::kBytes));
    gram_attr.set(ctx->device);
  }
  timing_begin(ctx);
  const bool xvec = (ldx % 4 == 0) && aligned16(X);
  CU(cudaMemsetAsync(d_info, 0, sizeof(int32_t) * F, st));
  // ---- moments: one pass over X for the CV folds (+ one for the hold-out split) ----
  for (int pass = 0; pass < 1 + has_hold; ++pass) {
    const int8_t* lab = pass == 0 ? fold_id : hold_id;
    const int nf = pass == 0 ? n_folds : 2;
    if (nf == 0) continue;
    const bool use_mask = nf <= 16;
    if (use_mask) { blockmask_kernel<<<(unsigned)cdiv(cdiv(n, 16), 256), 256, 0, st>>>(lab, n, d_bmask); KCHECK(); }
    double* dst = pass == 0 ? d_Mfold : d_Mfold + (size_t)n_folds * D * D;  // hold pass writes [n_folds], [n_folds+1]
    if (use_tc) {
      CU(cudaMemsetAsync(d_Mpart, 0, sizeof(float) * (size_t)nf * splits * ldm_tc * ldm_tc, st));
      tc::TcGramParams gp{n, rps, d, nb, lab, d_bmask, splits, d_Mpart, ldm_tc};
      tick(ctx);
      tc::tc_gram_kernel<<<dim3((unsigned)(nb * (nb + 1) / 2), 1, (unsigned)(nf * splits)), tc::kGramThreads,
                           tc::GramSmem::kBytes, st>>>(mXg, gp);
      tock(ctx, 1);
      KCHECK();
      tc::XtyParams xp{X, ldx, n, d, T, lab, d_bmask, is_classifier ? nullptr : y_reg, is_classifier ? y_cls : nullptr,
                       xty_rows, xty_chunks, d_xty};
      {
        const dim3 xg((unsigned)xty_chunks, (unsigned)nf);
        if (d <= 512 && T <= 2) tc::xty_kernel<1, 2><<<xg, 256, 0, st>>>(xp);
        else if (T <= 2) tc::xty_kernel<2, 2><<<xg, 256, 0, st>>>(xp);
        else if (d <= 512) tc::xty_scalar_kernel<2, 16><<<xg, 256, 0, st>>>(xp);
        else tc::xty_scalar_kernel<4, 16><<<xg, 256, 0, st>>>(xp);
      }
      KCHECK();
      tc::GramReduceParams rp{d_Mpart, splits, nb, ldm_tc, d_xty, xty_chunks, d, T, dst};
      tc::gram_reduce_kernel<<<dim3((unsigned)cdiv(D, 128), D, nf), 128, 0, st>>>(rp);
      KCHECK();
      ctx->total_launches += 5;
      continue;
    }
    GramParams gp{X, ldx, n, d, xvec, lab, use_mask ? d_bmask : nullptr, T, is_classifier ? nullptr : y_reg,
                  is_classifier ? y_cls : nullptr, rps, splits, d_Mpart, ldm};
    dim3 grid((unsigned)cdiv(D, BM), (unsigned)cdiv(D, BN), (unsigned)(nf * splits));
    tick(ctx);
    gram_kernel<<<grid, NT, kGemmSmemBytes, st>>>(gp);
    tock(ctx, 1);
    KCHECK();
    reduce_moments_kernel<<<dim3((unsigned)cdiv(D, 128), D, nf), 128, 0, st>>>(d_Mpart, splits, D, ldm, dst);
    KCHECK();
    ctx->total_launches += 3;
  }
  if (has_hold) {
    // keep only the label-1 moments of the hold-out pass, in slot n_folds
    CU(cudaMemcpyAsync(d_Mfold + (size_t)n_folds * D * D, d_Mfold + (size_t)(n_folds + 1) * D * D,
                       sizeof(double) * D * D, cudaMemcpyDeviceToDevice, st));
  }
  if (ctx->comm) { int rc = allreduce(ctx, d_Mfold, (size_t)fpc * D * D, kNcclFloat64); if (rc) return rc; }
  if (n_folds > 0) {
    total_moments_kernel<<<(unsigned)cdiv((int64_t)D * D, 256), 256, 0, st>>>(d_Mfold, n_folds, D, d_Mtot);
  } else {
    // only a hold-out split: total = label0 + label1 moments of that pass (label 0 was overwritten above) -> recompute
    return fail(ctx, DML_ERR_ARG, "dml_ridge_sweep: n_folds must be >= 2");
  }
  KCHECK();
  if (npad) {      // score layout: padding columns (zero weights, mask -1), group starts, fold id of every row tile
    CU(cudaMemsetAsync(d_Wt, 0, sizeof(float) * (size_t)S_score * ldw, st));
    CU(cudaMemsetAsync(d_bt, 0, sizeof(float) * (size_t)S_score, st));
    CU(cudaMemsetAsync(d_sel, 0xff, sizeof(int32_t) * (size_t)S_score, st));
    CU(cudaMemsetAsync(d_val, 0, sizeof(int32_t) * (size_t)S_score, st));
    score_groups_kernel<<<1, 160, 0, st>>>(d_gs, fpc, npad);
    tile_fold_kernel<<<(unsigned)cdiv(cdiv(n, 128), 8), 256, 0, st>>>(fold_id, n, d_tfold);
    KCHECK();
  }
  // ---- batched solves ----
  for (int c0 = 0; c0 < n_cand; c0 += cand_chunk) {
    const int nc = std::min(cand_chunk, n_cand - c0), nsys = nc * fpc;
    RidgeSysParams sp{d, T, D, n_folds, has_hold, c0, nc, d_Mfold, d_Mtot, alpha, fit_intercept, d_A, d_B, d_mu};
    if (d <= kCholMaxD) {      // the blocked kernel forms the systems itself from the moments (RidgeFused)
      const size_t chol_doubles = (size_t)kCholMaxD * (kCholNB + 1) + 1 + (size_t)kCholNB * (kCholNB + 2) + kCholMaxD + BN + 2 + 16 * (size_t)kCholMaxD;
      const size_t chol_smem = ((size_t)d * (kCholNB + 1) + 1 + (size_t)kCholNB * (kCholNB + 2) + d + T + 2 + 16 * (size_t)d) * sizeof(double);   // + ring of 4 x 4 rows
      static DeviceOnce chol_attr;
      if (!chol_attr.test(ctx->device)) {
        CU(cudaFuncSetAttribute(ridge_cholesky_blocked_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                (int)(chol_doubles * sizeof(double))));
        chol_attr.set(ctx->device);
      }
      const int threads = (int)std::min<int64_t>(512, rup(d, 32));
      RidgeFused rf{d_Mfold, d_Mtot, alpha, fit_intercept, D, n_folds, has_hold, c0, d_mu};
      long long* d_prof = getenv("DML_CHOL_PROF") ? (long long*)d_statpart : nullptr;   // (scratch that is free at this point)
      ridge_cholesky_blocked_kernel<<<nsys, threads, chol_smem, st>>>(d_A, d_AT, d_B, d, T, d_info + c0 * fpc, rf, 1, d_prof);
      if (d_prof) {
        long long h[6];
        CU(cudaMemcpyAsync(h, d_prof, sizeof(h), cudaMemcpyDeviceToHost, st));
        CU(cudaStreamSynchronize(st));
        fprintf(stderr, "chol block 0 clocks: update %lld form %lld diag %lld rows %lld writeback %lld solves %lld\n", h[0], h[1], h[2], h[3], h[4], h[5]);
      }
    } else {
      ridge_assemble_kernel<<<dim3(32, nsys), 256, 0, st>>>(sp);
      ridge_cholesky_kernel<<<nsys, 512, d * sizeof(double), st>>>(d_A, d_B, d, T, d_info + c0 * fpc);
    }
    RidgeOutParams op{d, T, n_folds, has_hold, c0, ldw, d_B, d_mu, coef_out, intercept_out, d_Wt, d_bt, d_sel, d_val, npad};
    ridge_output_kernel<<<dim3(nsys, T), 128, 0, st>>>(op);
    KCHECK();
    ctx->total_launches += 3;
  }
  if (info_out) CU(cudaMemcpyAsync(info_out, d_info, sizeof(int32_t) * F, cudaMemcpyDeviceToDevice, st));
  // ---- score every fit on its held-out rows ----
  int rc; int64_t n_parts = tiles_total; const void* parts = d_part;
  if (score_tc) {
    rc = run_score_tc(ctx, is_classifier ? FWD_SCORE_CLS : FWD_SCORE_REG, X, n, d, ldx, y_cls, y_reg, fold_id, hold_id,
                      d_Wt, ldw, d_bt, d_sel, d_val, (int)S_score, T, d_Whi, d_Wlo, d_part_tc, &n_parts,
                      npad ? d_tfold : nullptr, npad ? d_gs : nullptr, n_folds, has_hold, npad, d_mixed);
    parts = d_part_tc;
  } else {
    rc = run_score(ctx, is_classifier ? FWD_SCORE_CLS : FWD_SCORE_REG, X, n, d, ldx, y_cls, y_reg, fold_id, hold_id,
                   d_Wt, ldw, d_bt, d_sel, d_val, (int)S_all, T, d_part);
  }
  if (rc) return rc;
  if (is_classifier) {
    reduce_counts_kernel<<<F, 256, 0, st>>>((const int32_t*)parts, n_parts, (int)(score_tc ? S_score : S_all), T, d_cnt, fpc, npad);
    KCHECK();
    if (ctx->comm) { rc = allreduce(ctx, d_cnt, F, kNcclInt64); if (rc) return rc; }
    if (correct_out) CU(cudaMemcpyAsync(correct_out, d_cnt, sizeof(int64_t) * F, cudaMemcpyDeviceToDevice, st));
  } else {
    if (score_tc) reduce_sse_parts_kernel<<<(unsigned)cdiv(S_all, 128), 128, 0, st>>>((const double*)parts, n_parts, (int)S_score, d_sse, (int)S_all, fpc, npad);
    else reduce_sse_kernel<<<(unsigned)S_all, 256, 0, st>>>((const float*)d_part, tiles_total, (int)S_all, d_sse);
    KCHECK();
    if (ctx->comm) { rc = allreduce(ctx, d_sse, S_all, kNcclFloat64); if (rc) return rc; }
    if (sse_out) CU(cudaMemcpyAsync(sse_out, d_sse, sizeof(double) * S_all, cudaMemcpyDeviceToDevice, st));
    if (sy_out || syy_out) {
      target_stats_kernel<<<dim3(n_folds, kStatSlices), 256, 0, st>>>(y_reg, fold_id, n, d_statpart);
      target_stats_sum_kernel<<<n_folds, 32, 0, st>>>(d_statpart, d_stats, d_stats + 256);
      if (has_hold) {
        target_stats_kernel<<<dim3(2, kStatSlices), 256, 0, st>>>(y_reg, hold_id, n, d_statpart);
        target_stats_sum_kernel<<<2, 32, 0, st>>>(d_statpart, d_stats + 512, d_stats + 768);
      }
      KCHECK();
      if (ctx->comm) { rc = allreduce(ctx, d_stats, 1024, kNcclFloat64); if (rc) return rc; }
      expand_stats_kernel<<<(unsigned)cdiv(F, 128), 128, 0, st>>>(d_stats, d_stats + 256, d_stats + 512, d_stats + 768,
                                                                 F, fpc, n_folds, sy_out, syy_out);
      KCHECK();
    }
  }
  if (n_test_out) {
    ridge_teststats_kernel<<<(unsigned)cdiv(F, 128), 128, 0, st>>>(d_Mfold, D, d, F, fpc, n_folds,
                                                                  (long long*)n_test_out, nullptr, nullptr);
    KCHECK();
  }
  ctx->total_launches += 3;
  timing_end(ctx);
  CU(cudaStreamSynchronize(st));
  return DML_OK;
}

// =============================================================================================
extern "C" int dml_score(dml_ctx* ctx, const float* X, int64_t n, int d, int64_t ldx,
                         const int32_t* y_cls, const float* y_reg,
                         const int8_t* fold_id, const int8_t* hold_id,
                         const float* W, const float* b, const int32_t* mask_sel, const int32_t* mask_val,
                         int n_slots, int T, int64_t* correct_out, double* sse_out) {
  if (!ctx) return DML_ERR_ARG;
  if (!X || !fold_id || !W || !b || !mask_sel || !mask_val || n <= 0 || d <= 0 || n_slots <= 0 || T <= 0 ||
      n_slots % T != 0 || T > BN || (!y_cls && !y_reg))
    return fail(ctx, DML_ERR_ARG, "dml_score: bad argument");
  CU(cudaSetDevice(ctx->device));
  cudaGetLastError();
  cudaStream_t st = ctx->stream;
  const int64_t tiles_total = cdiv(n, BM);
  const bool cls = y_cls != nullptr && correct_out != nullptr;
  if (!cls && (!sse_out || !y_reg)) return fail(ctx, DML_ERR_ARG, "dml_score: regression scoring needs y_reg and sse_out");
  // tensor-core scoring (the kernel the sweeps use) whenever the shape allows: TMA needs 16-byte aligned
  // rows of X; the coefficients are copied into a padded scratch copy, so W may have any d
  const bool use_tc = tc_eligible(X, ldx, T) && (cls || T == 1);
  const int ldw = use_tc ? (int)rup(d, 4) : d;
  void* d_part; long long* d_cnt; double* d_sse;
  float *d_Wt = nullptr, *d_Whi = nullptr, *d_Wlo = nullptr;
  for (int pass = 0; pass < 2; ++pass) {
    Arena a(pass ? ctx->ws : nullptr);
    if (use_tc) {
      d_part = a.take<double>((size_t)ctx->n_sms * 8 * n_slots);
      d_Wt = a.take<float>((size_t)n_slots * ldw); d_Whi = a.take<float>((size_t)n_slots * ldw); d_Wlo = a.take<float>((size_t)n_slots * ldw);
    } else {
      d_part = a.take<float>((size_t)tiles_total * n_slots);
    }
    d_cnt = a.take<long long>(n_slots); d_sse = a.take<double>(n_slots);
    if (!pass) { int rc = ensure_ws(ctx, a.off + 256); if (rc) return rc; }
  }
  timing_begin(ctx);
  int rc; int64_t n_parts = tiles_total;
  if (use_tc) {
    if (ldw > d) CU(cudaMemsetAsync(d_Wt, 0, sizeof(float) * (size_t)n_slots * ldw, st));
    CU(cudaMemcpy2DAsync(d_Wt, (size_t)ldw * 4, W, (size_t)d * 4, (size_t)d * 4, n_slots, cudaMemcpyDeviceToDevice, st));
    rc = run_score_tc(ctx, cls ? FWD_SCORE_CLS : FWD_SCORE_REG, X, n, d, ldx, y_cls, y_reg, fold_id, hold_id,
                      d_Wt, ldw, b, mask_sel, mask_val, n_slots, T, d_Whi, d_Wlo, d_part, &n_parts);
  } else {
    rc = run_score(ctx, cls ? FWD_SCORE_CLS : FWD_SCORE_REG, X, n, d, ldx, y_cls, y_reg, fold_id, hold_id,
                   W, d, b, mask_sel, mask_val, n_slots, T, d_part);
  }
  if (rc) return rc;
  if (cls) {
    reduce_counts_kernel<<<n_slots / T, 256, 0, st>>>((const int32_t*)d_part, n_parts, n_slots, T, d_cnt);
    KCHECK();
    if (ctx->comm) { rc = allreduce(ctx, d_cnt, n_slots / T, kNcclInt64); if (rc) return rc; }
    CU(cudaMemcpyAsync(correct_out, d_cnt, sizeof(int64_t) * (n_slots / T), cudaMemcpyDeviceToDevice, st));
  } else {
    if (use_tc) reduce_sse_parts_kernel<<<(unsigned)cdiv(n_slots, 128), 128, 0, st>>>((const double*)d_part, n_parts, n_slots, d_sse, n_slots, 1, 0);
    else reduce_sse_kernel<<<n_slots, 256, 0, st>>>((const float*)d_part, tiles_total, n_slots, d_sse);
    KCHECK();
    if (ctx->comm) { rc = allreduce(ctx, d_sse, n_slots, kNcclFloat64); if (rc) return rc; }
    CU(cudaMemcpyAsync(sse_out, d_sse, sizeof(double) * n_slots, cudaMemcpyDeviceToDevice, st));
  }
  timing_end(ctx);
  CU(cudaStreamSynchronize(st));
  return DML_OK;
}

// =============================================================================================
// Scoring without any allocation.
//
// Some callers of this library must not allocate device memory while they run: a worker that
// was admitted with a fixed device-memory reservation and has handed all of it out already, or
// a scoring call issued on a side stream while a sweep owns the GPU -- cudaMalloc / cudaFree
// serialise the whole device and would stall that sweep (and fail outright when the reservation
// is spent).  dml_score reaches into the ctx arena, which grows with cudaMalloc on demand; this
// variant keeps its only scratch (per-row-tile partial counts) in the fixed array below, so the
// call launches kernels and nothing else.  The array is resident on every GPU that loads the
// library (4 MiB) until the process exits.
constexpr size_t kStaticScoreInts = (size_t)1 << 20;   // 1 Mi int32 = 4 MiB
static __device__ __align__(256) int32_t g_score_scratch[kStaticScoreInts];
// Host code must not hand `g_score_scratch` itself to a kernel: on the host that name is only a
// stand-in for the symbol.  The device address is looked up once per device, below.
static int32_t* g_score_scratch_dev[kMaxDevices] = {nullptr};
static std::mutex g_score_scratch_mu;   // two threads may make the first call on a device at the same time

extern "C" int dml_score_static(int device, void* stream_v, const float* X, int64_t n, int d, int64_t ldx,
                                const int32_t* y_cls, const int8_t* fold_id, const int8_t* hold_id,
                                const float* W, const float* b, const int32_t* mask_sel, const int32_t* mask_val,
                                int n_slots, int T, int64_t* correct_out) {
  dml_ctx* ctx = nullptr;   // messages go to the context-free slot of dml_last_error(NULL)
  if (device < 0 || device >= kMaxDevices) return fail(ctx, DML_ERR_ARG, "dml_score_static: device %d out of range", device);
  if (!X || !y_cls || !fold_id || !W || !b || !mask_sel || !mask_val || !correct_out || n <= 0 || d <= 0 || ldx < d ||
      n_slots <= 0 || T <= 0 || T > BN || n_slots % T != 0)
    return fail(ctx, DML_ERR_ARG, "dml_score_static: bad argument");
  const int64_t tiles = cdiv(n, BM);
  const int tile_n = (BN / T) * T;
  if ((size_t)tiles * (size_t)n_slots > kStaticScoreInts)
    return fail(ctx, DML_ERR_LIMIT,
                "dml_score_static: ceil(n/128) * n_slots = %lld exceeds the fixed scratch of %llu entries "
                "(4 MiB); score fewer columns per call or use dml_score",
                (long long)(tiles * n_slots), (unsigned long long)kStaticScoreInts);
  if (cdiv(n_slots, tile_n) > 65535) return fail(ctx, DML_ERR_LIMIT, "dml_score_static: too many columns (%d)", n_slots);
  CU(cudaSetDevice(device));
  int32_t* part = nullptr;
  {
    std::lock_guard<std::mutex> guard(g_score_scratch_mu);
    if (!g_score_scratch_dev[device]) {
      // one-time set-up on this device (no allocation: address lookup + a function attribute)
      int32_t* ptr = nullptr;
      CU(cudaGetSymbolAddress((void**)&ptr, g_score_scratch));
      CU(cudaFuncSetAttribute(fwd_kernel<FWD_SCORE_CLS>, cudaFuncAttributeMaxDynamicSharedMemorySize, kFwdSmemBytes));
      g_score_scratch_dev[device] = ptr;
    }
    part = g_score_scratch_dev[device];
  }
  cudaStream_t st = (cudaStream_t)stream_v;
  // The array still holds the previous call's partial counts.  Every entry this call reads is
  // written by it first, but the used part is zeroed anyway so that stale counts can never leak.
  CU(cudaMemsetAsync(part, 0, sizeof(int32_t) * (size_t)tiles * n_slots, st));
  FwdParams fp{};
  fp.X = X; fp.ldx = ldx; fp.row0 = 0; fp.rows = n; fp.d = d;
  fp.xvec = (ldx % 4 == 0) && aligned16(X);
  fp.Wt = W; fp.ldw = d; fp.wvec = (d % 4 == 0) && aligned16(W); fp.bt = b;
  fp.n_slots = n_slots; fp.T = T; fp.tile_n = tile_n;
  fp.ycls = y_cls; fp.yreg = nullptr; fp.fold_id = fold_id; fp.hold_id = hold_id;
  fp.slot_sel = mask_sel; fp.slot_val = mask_val;
  fp.cntPart = part; fp.ssePart = nullptr;
  dim3 grid((unsigned)tiles, (unsigned)cdiv(n_slots, tile_n));
  fwd_kernel<FWD_SCORE_CLS><<<grid, NT, kFwdSmemBytes, st>>>(fp);
  KCHECK();
  reduce_counts_kernel<<<n_slots / T, 256, 0, st>>>(part, tiles, n_slots, T, (long long*)correct_out);
  KCHECK();
  CU(cudaStreamSynchronize(st));
  return DML_OK;
}
