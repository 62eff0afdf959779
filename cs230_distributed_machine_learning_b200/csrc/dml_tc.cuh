// tcgen05 (5th-gen tensor core) path of the sweep: 3xTF32 split-precision contractions
// with TMA-staged operands and TMEM accumulators, sm_100a only.
//
//   tc_fwd_kernel  : Z = X W^T for a 128-row tile x NC columns, epilogue in registers:
//                    loss / residual (one fp32 array R) or counts
//   tc_grad2_kernel: G = X^T R for two 128-feature blocks x 128 columns over a row split, on CTA pairs
//   tc_gram_kernel : per-fold X^T X blocks for Ridge
//
// Split precision: a = a_hi + a_lo with a_hi = a & 0xffffe000 (exactly a TF32 number);
// a.b ~= a_hi.b_hi + a_hi.b_lo + a_lo.b_hi, three kind::tf32 MMAs into one fp32
// TMEM accumulator.  The operand derived from X is split inside the kernel and staged
// in tensor memory; W arrives pre-split (split_weights_kernel), R is split inside the
// gradient kernel (the split is elementwise, so it works on the swizzled tile as it lies).
//
// Reference call sites these stand in for: the X @ w and X.T @ g products inside
// every L-BFGS function evaluation of every fit (sklearn/linear_model/_linear_loss.py:216-222,
// 355-368, reached from aws-prod/worker/worker.py:315,326).
#pragma once
#include <cuda.h>
#include "dml_common.cuh"
#include "dml_gemm.cuh"

namespace dml {
namespace tc {

// ---------------------------------------------------------------- PTX wrappers
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
// suspend-time hint of try_wait: the warp may sleep that long inside the instruction instead of coming back
// to spin (without it a quarter of the forward kernel's issued instructions were try_wait / branch / yield
// of waiting warps, in a kernel that is bound by instruction issue)
constexpr uint32_t kWaitHintNs = 20000;
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  const uint32_t addr = smem_u32(bar);
  uint32_t done;
  do {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2, %3;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(done) : "r"(addr), "r"(parity), "r"(kWaitHintNs) : "memory");
  } while (!done);
}
__device__ __forceinline__ void fence_barrier_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

__device__ __forceinline__ void tma_load_2d(const CUtensorMap* map, uint64_t* bar, void* dst, int c0, int c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
      ::"r"(smem_u32(dst)), "l"(map), "r"(smem_u32(bar)), "r"(c0), "r"(c1) : "memory");
}
__device__ __forceinline__ void tma_load_3d(const CUtensorMap* map, uint64_t* bar, void* dst, int c0, int c1, int c2) {
  asm volatile(
      "cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];"
      ::"r"(smem_u32(dst)), "l"(map), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2) : "memory");
}
__device__ __forceinline__ void tma_prefetch_2d(const CUtensorMap* map, int c0, int c1) {
  asm volatile("cp.async.bulk.prefetch.tensor.2d.L2.global.tile [%0, {%1, %2}];" ::"l"(map), "r"(c0), "r"(c1) : "memory");
}
__device__ __forceinline__ void tma_load_2d_mcast(const CUtensorMap* map, uint64_t* bar, void* dst, int c0, int c1, uint16_t mask) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes.multicast::cluster [%0], [%1, {%3, %4}], [%2], %5;"
      ::"r"(smem_u32(dst)), "l"(map), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "h"(mask) : "memory");
}
__device__ __forceinline__ void umma_commit_mcast(uint64_t* bar, uint16_t mask) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
               ::"r"(smem_u32(bar)), "h"(mask) : "memory");
}
__device__ __forceinline__ uint32_t cluster_ctarank() { uint32_t r; asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r)); return r; }
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
  asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// ---- CTA-pair (cta_group::2) variants: one MMA spans the two SMs of a cluster of two
__device__ __forceinline__ uint32_t mapa_rank(uint32_t saddr, uint32_t rank) {   // same offset in CTA `rank`
  uint32_t r; asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(saddr), "r"(rank)); return r;
}
__device__ __forceinline__ void mbar_arrive_cluster(uint32_t cluster_addr) {
  // default semantics on purpose: the data published here lives in tensor memory and is ordered by the
  // tcgen05 fences; ".release.cluster" compiles to MEMBAR.ALL.GPU per arrive (measured: the staging warps
  // then spend a quarter of their time in it)
  asm volatile("mbarrier.arrive.shared::cluster.b64 _, [%0];" ::"r"(cluster_addr) : "memory");
}
__device__ __forceinline__ void mbar_wait_cluster(uint64_t* bar, uint32_t parity) {   // arrivals may come from the peer CTA
  const uint32_t addr = smem_u32(bar);
  uint32_t done;
  do {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2, %3;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(done) : "r"(addr), "r"(parity), "r"(kWaitHintNs) : "memory");
  } while (!done);
}
__device__ __forceinline__ void tmem_alloc_pair(uint32_t* slot, uint32_t cols) {
  asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(slot)), "r"(cols) : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc_pair(uint32_t addr, uint32_t cols) {
  asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(addr), "r"(cols) : "memory");
}
// D[256 x N] += A[256 x 8] B[8 x N]: rows 0-127 of A and D in the leader's tensor memory, 128-255 in the
// peer's (same addresses); columns 0..N/2-1 of B in the leader's shared memory, the rest in the peer's
__device__ __forceinline__ void umma_tf32_ts_pair(uint32_t tmem_c, uint32_t tmem_a, uint64_t db, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::2.kind::tf32 [%0], [%1], %2, %3, p;\n\t}"
      ::"r"(tmem_c), "r"(tmem_a), "l"(db), "r"(idesc), "r"(accumulate) : "memory");
}
// bf16 operands (A: tensor memory, two K-adjacent elements per 32-bit column; B: shared memory), fp32 accumulate
__device__ __forceinline__ void umma_bf16_ts_pair(uint32_t tmem_c, uint32_t tmem_a, uint64_t db, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::2.kind::f16 [%0], [%1], %2, %3, p;\n\t}"
      ::"r"(tmem_c), "r"(tmem_a), "l"(db), "r"(idesc), "r"(accumulate) : "memory");
}
__device__ __forceinline__ void umma_commit_pair(uint64_t* bar) {   // arrives on `bar` in both CTAs of the pair
  asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
               ::"r"(smem_u32(bar)), "h"((uint16_t)3) : "memory");
}
__device__ __forceinline__ void tmem_alloc(uint32_t* slot, uint32_t cols) {
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(slot)), "r"(cols) : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t addr, uint32_t cols) {
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(addr), "r"(cols) : "memory");
}
__device__ __forceinline__ void umma_tf32(uint32_t tmem_c, uint64_t da, uint64_t db, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}"
      ::"r"(tmem_c), "l"(da), "l"(db), "r"(idesc), "r"(accumulate) : "memory");
}
// A operand in tensor memory ("TS" form): A[128 x 8] = lanes 0..127, 8 consecutive 32-bit columns
__device__ __forceinline__ void umma_tf32_ts(uint32_t tmem_c, uint32_t tmem_a, uint64_t db, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::tf32 [%0], [%1], %2, %3, p;\n\t}"
      ::"r"(tmem_c), "r"(tmem_a), "l"(db), "r"(idesc), "r"(accumulate) : "memory");
}
__device__ __forceinline__ void tmem_st32(uint32_t taddr, const uint32_t (&v)[32]) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,"
      "%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31,%32};"
      ::"r"(taddr), "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7]),
        "r"(v[8]), "r"(v[9]), "r"(v[10]), "r"(v[11]), "r"(v[12]), "r"(v[13]), "r"(v[14]), "r"(v[15]),
        "r"(v[16]), "r"(v[17]), "r"(v[18]), "r"(v[19]), "r"(v[20]), "r"(v[21]), "r"(v[22]), "r"(v[23]),
        "r"(v[24]), "r"(v[25]), "r"(v[26]), "r"(v[27]), "r"(v[28]), "r"(v[29]), "r"(v[30]), "r"(v[31]) : "memory");
}
__device__ __forceinline__ void tmem_st_wait() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ uint32_t lds32u(uint32_t a) { uint32_t v; asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(a)); return v; }
__device__ __forceinline__ void umma_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, uint32_t (&v)[16]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
        "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
      : "r"(taddr));
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ bool elect_one() {
  uint32_t pred;
  asm volatile("{\n\t.reg .pred p;\n\telect.sync _|p, 0xffffffff;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(pred));
  return pred != 0;
}

// ---------------------------------------------------------------- descriptors
// Shared-memory matrix descriptor (sm_100 "version 1").
//   K-major  (layout 2 = SWIZZLE_128B): rows of 128 B (32 tf32 along K), 8-row groups SBO = 1024 B
//             apart; the K offset inside the swizzle atom is added to the start address.
//   MN-major (layout 1 = SWIZZLE_128B_BASE32B, the only layout tf32 accepts when transposed): rows
//             of 128 B (32 tf32 along M/N) stacked along K, 32-byte chunks XOR-ed with (row & 3);
//             4-row atoms SBO = 512 B apart, the next 32 elements along M/N LBO bytes away.
//             TMA writes this pattern with CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B.
constexpr uint32_t kLayoutSw128 = 2, kLayoutSw128Base32 = 1;
__device__ __forceinline__ uint64_t smem_desc(uint32_t saddr, uint32_t lbo_bytes, uint32_t sbo_bytes,
                                              uint32_t layout = kLayoutSw128) {
  uint64_t d = 0;
  d |= (uint64_t)((saddr >> 4) & 0x3fff);
  d |= (uint64_t)((lbo_bytes >> 4) & 0x3fff) << 16;
  d |= (uint64_t)((sbo_bytes >> 4) & 0x3fff) << 32;
  d |= (uint64_t)1 << 46;   // descriptor version (Blackwell)
  d |= (uint64_t)layout << 61;
  return d;
}
// Instruction descriptor: D fp32, A/B tf32, M x N, transposes (MN-major) as flags.
__host__ __device__ constexpr uint32_t instr_desc_tf32(int M, int N, int a_mn_major, int b_mn_major) {
  return (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)a_mn_major << 15) | ((uint32_t)b_mn_major << 16) |
         ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}

// D fp32, A/B bf16
__host__ __device__ constexpr uint32_t instr_desc_bf16(int M, int N, int a_mn_major, int b_mn_major) {
  return (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)a_mn_major << 15) | ((uint32_t)b_mn_major << 16) |
         ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}
__device__ __forceinline__ uint32_t pack_bf16(float lo_k, float hi_k) {   // element k in the low half, k + 1 in the high half
  uint32_t r;
  asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(hi_k), "f"(lo_k));
  return r;
}

constexpr uint32_t kHiMask = 0xffffe000u;
// class weights in the forward epilogues; -DDML_NO_CW compiles them out (A/B builds only)
#ifdef DML_NO_CW
#define DML_HAS_CW(p) false
#else
#define DML_HAS_CW(p) ((p).cw != nullptr)
#endif
__device__ __forceinline__ float tf32_hi(float x) { return __uint_as_float(__float_as_uint(x) & kHiMask); }

__device__ __forceinline__ float4 lds128(uint32_t a) {
  float4 v;
  asm volatile("ld.shared.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(a));
  return v;
}
__device__ __forceinline__ void sts128(uint32_t a, const float4& v) {
  asm volatile("st.shared.v4.f32 [%0], {%1,%2,%3,%4};" ::"r"(a), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}
// e^x for x <= 0 as one FMUL and one MUFU: __expf without -ftz wraps the MUFU in a range test and two predicated
// multiplies (3 of the ~16 instructions per element of the logistic epilogue) to keep results below 2^-126, which
// no consumer here can see (1 + e, e / (1 + e) - y and log(1 + e) are the same fp32 numbers with e flushed to 0)
__device__ __forceinline__ float exp_ftz(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x * 1.4426950408889634f));
  return y;
}
__device__ __forceinline__ float lds32f(uint32_t a) { float v; asm volatile("ld.shared.f32 %0, [%1];" : "=f"(v) : "r"(a)); return v; }
__device__ __forceinline__ double lds64d(uint32_t a) { double v; asm volatile("ld.shared.f64 %0, [%1];" : "=d"(v) : "r"(a)); return v; }
__device__ __forceinline__ void sts64d(uint32_t a, double v) { asm volatile("st.shared.f64 [%0], %1;" ::"r"(a), "d"(v) : "memory"); }
__device__ __forceinline__ void stg64(float* a, float x, float y) { asm volatile("st.global.v2.f32 [%0], {%1, %2};" ::"l"(a), "f"(x), "f"(y) : "memory"); }
__device__ __forceinline__ float2 lds64f(uint32_t a) { float2 v; asm volatile("ld.shared.v2.f32 {%0, %1}, [%2];" : "=f"(v.x), "=f"(v.y) : "r"(a)); return v; }
__device__ __forceinline__ void sts64f(uint32_t a, float x, float y) { asm volatile("st.shared.v2.f32 [%0], {%1, %2};" ::"r"(a), "f"(x), "f"(y) : "memory"); }
__device__ __forceinline__ uint32_t lds16u(uint32_t a) { uint16_t v; asm volatile("ld.shared.u16 %0, [%1];" : "=h"(v) : "r"(a)); return v; }

// split a stage buffer: lo <- x - (x & mask) (and optionally hi <- x & mask).  `bytes` multiple of
// 2048; `hi`, `lo` are shared-space addresses.  128 threads, 4 x 16 bytes in flight per thread.
__device__ __forceinline__ void split_stage(uint32_t hi, uint32_t lo, int bytes, int tid, int nthreads, bool mask_hi) {
  const int n4 = bytes >> 4;
  for (int i0 = tid; i0 < n4; i0 += 4 * nthreads) {
    float4 x[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) if (i0 + u * nthreads < n4) x[u] = lds128(hi + (i0 + u * nthreads) * 16);
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int i = i0 + u * nthreads;
      if (i < n4) {
        const float4 h = make_float4(tf32_hi(x[u].x), tf32_hi(x[u].y), tf32_hi(x[u].z), tf32_hi(x[u].w));
        sts128(lo + i * 16, make_float4(x[u].x - h.x, x[u].y - h.y, x[u].z - h.z, x[u].w - h.w));
        if (mask_hi) sts128(hi + i * 16, h);
      }
    }
  }
}

// ---------------------------------------------------------------- forward kernel
constexpr int kFwdStages = 3;
constexpr int kFwdEpiGroups = 4;    // epilogue warps per TMEM lane quadrant (FwdSmem::kMultiAcc is sized for 4)
// warps 0-3 operand staging, 4 .. 4+4G-1 epilogue, then the TMA producer and the MMA issuer: the SM's warp
// arbiter favours the highest warp id, and the two single-thread roles must never wait behind the math warps
constexpr int kFwdThreads = 192 + 128 * kFwdEpiGroups;
constexpr int kFwdTmaWarp = 4 + 4 * kFwdEpiGroups, kFwdMmaWarp = kFwdTmaWarp + 1;
constexpr int kMaxLabel = 128;      // fold ids 0..126 (+ hold flag) in the held-column tables

struct TcFwdParams {
  int64_t n_rows_total;      // rows of X visible through the tensor map (for masking the last tile)
  int64_t row0, rows;        // chunk
  int d_chunks;              // ceil(d / 32)
  int n_slots, n_col_blocks;
  const float* bt;           // [n_slots]
  const int32_t* ycls; const float* yreg;
  const int8_t* fold_id; const int8_t* hold_id;
  const int32_t* slot_sel; const int32_t* slot_val;
  // fold-aware skipping (binary training mode): tile_fold[t] = the fold id shared by all rows of global row
  // tile t (255: mixed); gstart[k] .. gstart[k+1] = the columns of fold k's fits (multiples of 32).  A tile of
  // fold k trains none of them: no MMAs, no epilogue math for that column range.  gstart == nullptr: off.
  const uint8_t* tile_fold; const int* gstart; int n_groups;
  int hold_group;                        // scoring mode: group n_groups holds the hold-out fits (scored on rows of every tile)
  // Scoring with fold-major column groups of equal width that is a multiple of the block width NC
  // (score_bpg = blocks per group > 0; clusters of one): a row tile whose rows all sit in fold k only visits
  // the score_bpg blocks of group k (and those of the hold-out group), so X is streamed ONCE for all
  // columns instead of once per pair of column blocks.  score_phase 0: those tiles; score_phase 1: the few
  // tiles that mix folds, against every block.  Both launches skip the other's tiles.
  int score_bpg, score_phase;
  const int* mixed_tiles;                // score_phase 1: [0] = number of row tiles that mix folds, [1 ..] = their indices, ascending
  int grad_skips;                        // what the gradient kernel leaves unread of a hole, so nothing is stored there:
                                         // 1 = whole 64-column segments, 2 = 32-column chunks (TcGradParams::seg_skip)
  // n_slots is the host's upper bound (and the pitch of the partial-sum arrays); when n_list_dev is set the
  // live columns are *n_list_dev * T <= n_slots, read on the device, so that the host can size a round's grids
  // before it has read back the list length the previous round produced
  const int* n_list_dev;
  int gb_in_grad;                        // the gradient kernel sums r per column (TcGradParams::gb2): no butterfly for it here
  float* R; int ldr;                     // fp32 residuals, [ldr / 32 column chunks][chunk rows][32]: a chunk's rows are contiguous,
  int64_t r_chunk_stride;                // so a gradient box (32 rows x 32 columns) is one 4 KB run; stride in floats
  double* lossPart; double* gbPart;      // [gridDim.x * 4][n_slots], zeroed by the caller
  int32_t* cntPart; double* ssePart;     // score modes, same shape
  int mask_hi;
  int prefetch_jobs;                     // L2 prefetch distance of the X tiles, in jobs of this CTA (0 = off)
  int T;                                 // columns per fit (1 = binary; multinomial modes: classes, <= 16)
  float loss_shift;                      // per-row losses are summed as (loss - loss_shift): ln 2, or ln T
  // class weights (training modes; null = unweighted), [n_slots][2] as in FwdParams::cw: a row's (loss - shift)
  // and residual are multiplied by the weight of its class
  const float* cw;
};
enum { FWD_LOGREG_MULTI = 3, FWD_SCORE_MULTI = 4 };   // softmax / arg-max over T consecutive columns

template <int NC>
struct FwdSmem {
  static constexpr int kA = 128 * 128;          // 128 rows x 32 tf32 (TMA landing zone of X; A_hi/A_lo go to TMEM)
  static constexpr int kB = NC * 128;
  static constexpr int kStage = kA + 2 * kB;
  // per column: bias; per label value: 16-bit held-column masks, one per 16-column chunk
  // multinomial epilogue: fp64 running sums of every epilogue warp, [warp][fit of the warp][17] (16 class columns + the loss)
  static_assert(kFwdEpiGroups == 4, "kMultiAcc / kMultiStage below are laid out for 16 epilogue warps");
  static constexpr int kMultiAcc = 16 * (32 / 4) * 17 * 8;
  // multinomial epilogue: [32 rows][T <= 16] fp32 residuals of one fit per epilogue warp, turned row-major -> line-major before they are stored
  static constexpr int kMultiStage = 16 * 32 * 16 * 4;
  static constexpr int kMeta = NC * 4 + (kMaxLabel + 2) * (NC / 16) * 2 + (kMaxLabel + 2) * 4 + 64 + NC * 8 + kMultiAcc + kMultiStage;   // + column-group starts, class weights
  static constexpr int kBytes = kFwdStages * kStage + 1024 /*align*/ + 256 /*barriers*/ + kMeta;
  static_assert(kBytes <= 227 * 1024, "forward kernel: shared memory of one CTA");
};

// Sum v[c] over the 32 lanes of the warp for 16 columns with 16 shuffles (recursive halving):
// lane l ends up with the total of column bfly_col(l); lanes l and l^1 hold the same column.
__device__ __forceinline__ int bfly_col(int lane) { return ((lane >> 4) & 1) * 8 + ((lane >> 3) & 1) * 4 + ((lane >> 2) & 1) * 2 + ((lane >> 1) & 1); }
template <class T>
__device__ __forceinline__ T bfly16(const T (&v)[16], int lane) {
  T a[8], b[4], c[2], d;
  bool up = lane & 16;
#pragma unroll
  for (int i = 0; i < 8; ++i) { const T keep = up ? v[8 + i] : v[i], send = up ? v[i] : v[8 + i]; a[i] = keep + __shfl_xor_sync(0xffffffffu, send, 16); }
  up = lane & 8;
#pragma unroll
  for (int i = 0; i < 4; ++i) { const T keep = up ? a[4 + i] : a[i], send = up ? a[i] : a[4 + i]; b[i] = keep + __shfl_xor_sync(0xffffffffu, send, 8); }
  up = lane & 4;
#pragma unroll
  for (int i = 0; i < 2; ++i) { const T keep = up ? b[2 + i] : b[i], send = up ? b[i] : b[2 + i]; c[i] = keep + __shfl_xor_sync(0xffffffffu, send, 4); }
  up = lane & 2;
  { const T keep = up ? c[1] : c[0], send = up ? c[0] : c[1]; d = keep + __shfl_xor_sync(0xffffffffu, send, 2); }
  return d + __shfl_xor_sync(0xffffffffu, d, 1);
}

// arr[i] += val for a run-time i < n with arr kept in registers: n predicated adds with compile-time indices
#define DML_ACC_N(arr, n, val) do { const auto v_ = (val); _Pragma("unroll") for (int k_ = 0; k_ < (n); ++k_) if (k_ == i) arr[k_] += v_; } while (0)

// Jobs are ordered column block first: every CTA walks the row tiles of one column block (W stays
// hot in L2, per-column running sums stay in registers) and changes block at most a few times.
// CL = 2: the kernel runs as clusters of two CTAs that work on the SAME row tile and on neighbouring
// column blocks (2k, 2k+1); each CTA fetches half of every X chunk and TMA-multicasts it to both, so
// X crosses HBM / L2 once per row tile instead of once per (row tile, column block).
template <int NC, int MODE, int CL>
// (22 warps = 6 on the fullest of the four SM sub-partitions x 16 384 registers: 80 per thread is the ceiling at this block size)
__global__ void __launch_bounds__(kFwdThreads, 1)
tc_fwd_kernel(const __grid_constant__ CUtensorMap mapX, const __grid_constant__ CUtensorMap mapWhi,
              const __grid_constant__ CUtensorMap mapWlo, const TcFwdParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  using L = FwdSmem<NC>;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + kFwdStages * L::kStage);
  uint64_t* full = bars;                    // [S] TMA landed
  uint64_t* split = bars + kFwdStages;      // [S] lo operand written
  uint64_t* empty = bars + 2 * kFwdStages;  // [S] MMAs that read the stage retired
  uint64_t* tfull = bars + 3 * kFwdStages;  // [2] accumulator ready
  uint64_t* tempty = tfull + 2;             // [2] accumulator drained
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty + 2);
  float* colBias = reinterpret_cast<float*>(smem + kFwdStages * L::kStage + 256);   // [NC]
  // heldTab[v][ci] bit j: column ci*16+j holds out the rows whose label equals v (v < kMaxLabel: fold
  // id; v = kMaxLabel + h: hold flag h), or does not exist.  A row's held columns = OR of two entries.
  uint16_t* heldTab = reinterpret_cast<uint16_t*>(colBias + NC);                     // [kMaxLabel + 2][NC / 16]
  int* gsTab = reinterpret_cast<int*>(heldTab + (kMaxLabel + 2) * (NC / 16));        // [n_groups + 1] copy of p.gstart
  float* colCw = reinterpret_cast<float*>((reinterpret_cast<uintptr_t>(gsTab + kMaxLabel + 2) + 7) & ~(uintptr_t)7);   // [NC][2] class weights
  double* accS = reinterpret_cast<double*>(colCw + 2 * NC);                          // multinomial running sums, FwdSmem::kMultiAcc bytes
  float* stageS = reinterpret_cast<float*>(reinterpret_cast<uint8_t*>(accS) + L::kMultiAcc);   // multinomial residual staging, FwdSmem::kMultiStage bytes

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  // tensor memory: two accumulators of NC columns, then per stage 32 columns of X_hi and 32 of X_lo
  constexpr uint32_t kTmemUse = 2 * NC + kFwdStages * 64;
  constexpr uint32_t kTmemCols = kTmemUse <= 256 ? 256 : 512;
  constexpr uint32_t kTmemA = 2 * NC;
  static_assert(kTmemUse <= 512 && NC % 16 == 0, "NC");

  if (warp == kFwdTmaWarp && lane == 0) {
    for (int s = 0; s < kFwdStages; ++s) { mbar_init(full + s, 1); mbar_init(split + s, 4); mbar_init(empty + s, CL); }
    for (int a = 0; a < 2; ++a) { mbar_init(tfull + a, 1); mbar_init(tempty + a, 4 * kFwdEpiGroups); }
    fence_barrier_init();
  }
  if (warp == kFwdMmaWarp) tmem_alloc(tmem_slot, kTmemCols);
  constexpr bool kScoreBin = MODE == FWD_SCORE_CLS || MODE == FWD_SCORE_REG;
  const bool holes = (MODE == FWD_LOGREG || kScoreBin) && p.gstart != nullptr;
  if (holes) for (int i = threadIdx.x; i <= p.n_groups + (kScoreBin ? p.hold_group : 0); i += blockDim.x) gsTab[i] = p.gstart[i];
  tc_fence_before();
  __syncthreads();
  if (CL > 1) cluster_sync_all();           // peer barriers are initialised before anyone signals them
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  // job list of this CTA: (column block, row tile); a cluster takes CL neighbouring row tiles
  const int64_t n_row_tiles = (p.rows + 127) / 128;
  const int crank = CL > 1 ? (int)cluster_ctarank() : 0;
  const int n_live = p.n_list_dev ? min(p.n_slots, *p.n_list_dev * p.T) : p.n_slots;
  const int ncb_live = p.n_list_dev ? (n_live + NC - 1) / NC : p.n_col_blocks;
  const int64_t cb_groups = (ncb_live + CL - 1) / CL;            // column blocks are dealt out CL at a time
  constexpr bool kScoreAny = MODE == FWD_SCORE_CLS || MODE == FWD_SCORE_REG;
  const bool by_group = kScoreAny && CL == 1 && p.score_bpg > 0 && p.gstart != nullptr;
  // second scoring launch: only the listed tiles (looking every tile's fold up again cost one dependent global load
  // per job and role: 0.47 ms at config 3 for four tiles of real work)
  const bool by_list = by_group && p.score_phase == 1;
  const int64_t n_mixed = by_list ? p.mixed_tiles[0] : 1;      // (1: never used as a count, only keeps the divisions below defined)
  const int64_t n_jobs = by_list ? n_mixed * ncb_live
                       : by_group ? n_row_tiles * (p.score_bpg * (1 + (p.hold_group ? 1 : 0)))
                                  : n_row_tiles * cb_groups;     // jobs per cluster (column-block group major)
  const int64_t job0 = blockIdx.x / CL, job_step = gridDim.x / CL;
  auto job_rt = [&](int64_t job) -> int64_t { return by_list ? (int64_t)p.mixed_tiles[1 + job % n_mixed] : job % n_row_tiles; };
  // column block of a job; -1 = nothing to do for this job in this launch (every role skips it)
  auto job_cb = [&](int64_t job) -> int {
    if (by_list) return (int)(job / n_mixed);
    if (!by_group) return (int)(job / n_row_tiles) * CL + crank;                     // may be >= n_col_blocks (no columns)
    const int j = (int)(job / n_row_tiles);
    const int k = (int)p.tile_fold[p.row0 / 128 + job % n_row_tiles];
    if (k >= p.n_groups) return -1;                                                  // a tile that mixes folds: second launch
    return j < p.score_bpg ? k * p.score_bpg + j : p.n_groups * p.score_bpg + (j - p.score_bpg);
  };
  // block-local column range [sa, sb) that a tile whose rows all sit in fold k does not train
  // (empty: sa == sb == 0); k = 255: the tile mixes folds
  auto tile_k = [&](int64_t rt) -> int { return holes ? (int)p.tile_fold[p.row0 / 128 + rt] : 255; };
  // The two block-local column ranges [a0, b0) and [a1, b1) a tile needs when all its rows sit in fold k:
  // training mode: everything but fold k's group (the rows train none of its columns);
  // scoring mode : fold k's group (only its fits are scored on these rows) and the hold-out group.
  // Any other tile, or a list without groups: [0, 0) and [0, NC).
  auto ranges_of = [&](int cb, int k, int& a0, int& b0, int& a1, int& b1) {
    a0 = 0; b0 = 0; a1 = 0; b1 = NC;
    if (!holes || k >= p.n_groups) return;
    auto clip = [&](int g, int& lo, int& hi) {
      lo = max(gsTab[g], cb * NC) - cb * NC; hi = min(gsTab[g + 1], cb * NC + NC) - cb * NC;
      if (hi <= lo) { lo = 0; hi = 0; }
    };
    int lo, hi;
    clip(k, lo, hi);
    if (MODE == FWD_LOGREG) {
      if (hi > lo) { b0 = lo; a1 = hi; }
    } else {
      a0 = lo; b0 = hi; a1 = 0; b1 = 0;
      if (p.hold_group) clip(p.n_groups, a1, b1);
    }
  };

  if (warp == kFwdTmaWarp) {
    // ===== TMA producer =====
    if (lane == 0) {
      uint32_t it = 0;
      for (int64_t job = job0; job < n_jobs; job += job_step) {
        const int cb = job_cb(job); const int64_t rt = job_rt(job);
        if (cb < 0) continue;
        const int row = (int)(p.row0 + rt * 128), slot = cb * NC;
        // L2 prefetch distance: X comes from HBM (latency the 3-stage ring cannot cover by itself),
        // so the tile this CTA will need `kPf` jobs from now is pulled into L2 chunk by chunk
        const int64_t pf_job = job + p.prefetch_jobs * job_step;
        const bool pf = p.prefetch_jobs > 0 && pf_job < n_jobs;
        const int pf_row = pf ? (int)(p.row0 + job_rt(pf_job) * 128) : 0;
        for (int c = 0; c < p.d_chunks; ++c, ++it) {
          const int s = it % kFwdStages; const uint32_t ph = (it / kFwdStages) & 1;
          if (pf) tma_prefetch_2d(&mapX, c * 32, pf_row);
          mbar_wait(empty + s, ph ^ 1);
          uint8_t* st = smem + s * L::kStage;
          mbar_expect_tx(full + s, L::kA + 2 * L::kB);
          if (CL == 1) tma_load_2d(&mapX, full + s, st, c * 32, row);
          else         // this CTA's share of the X chunk (128 / CL rows), delivered to every CTA of the cluster
            tma_load_2d_mcast(&mapX, full + s, st + crank * (128 / CL) * 128, c * 32, row + crank * (128 / CL), (1 << CL) - 1);
          tma_load_2d(&mapWhi, full + s, st + L::kA, c * 32, slot);
          tma_load_2d(&mapWlo, full + s, st + L::kA + L::kB, c * 32, slot);
        }
      }
    }
  } else if (warp == kFwdMmaWarp) {
    // ===== MMA issuer =====
    uint32_t it = 0, tile = 0;
    int k_next = job0 < n_jobs ? tile_k(job_rt(job0)) : 255;
    for (int64_t job = job0; job < n_jobs; job += job_step) {
      const int cb_j = job_cb(job);
      const int k_here = k_next;
      if (job + job_step < n_jobs) k_next = tile_k(job_rt(job + job_step));   // in flight while this tile's MMAs are issued
      if (cb_j < 0) continue;
      const int acc = tile & 1; const uint32_t aph = (tile >> 1) & 1;
      ++tile;
      int a0, b0, a1, b1;
      ranges_of(cb_j, k_here, a0, b0, a1, b1);
      const uint32_t idesc_a = instr_desc_tf32(128, b0 - a0, 0, 0), idesc_b = instr_desc_tf32(128, b1 - a1, 0, 0);
      mbar_wait(tempty + acc, aph ^ 1);
      tc_fence_after();
      const uint32_t tacc = tmem_base + acc * NC;
      for (int c = 0; c < p.d_chunks; ++c, ++it) {
        const int s = it % kFwdStages; const uint32_t ph = (it / kFwdStages) & 1;
        mbar_wait(full + s, ph);
        mbar_wait(split + s, ph);
        tc_fence_after();
        if (elect_one()) {
          const uint32_t b_hi = smem_u32(smem + s * L::kStage) + L::kA, b_lo = b_hi + L::kB;
          const uint32_t ta_hi = tmem_base + kTmemA + s * 64, ta_lo = ta_hi + 32;
          if (b0 > a0) {
            const uint32_t off = (uint32_t)a0 * 128;       // W rows are 128 B apart (8-row swizzle atoms: a0 % 8 == 0)
#pragma unroll
            for (int k = 0; k < 4; ++k) {
              const uint64_t dbh = smem_desc(b_hi + off + k * 32, 16, 1024), dbl = smem_desc(b_lo + off + k * 32, 16, 1024);
              umma_tf32_ts(tacc + a0, ta_lo + k * 8, dbh, idesc_a, (c | k) != 0);   // small terms first
              umma_tf32_ts(tacc + a0, ta_hi + k * 8, dbl, idesc_a, 1);
              umma_tf32_ts(tacc + a0, ta_hi + k * 8, dbh, idesc_a, 1);
            }
          }
          if (b1 > a1) {
            const uint32_t off = (uint32_t)a1 * 128;
#pragma unroll
            for (int k = 0; k < 4; ++k) {
              const uint64_t dbh = smem_desc(b_hi + off + k * 32, 16, 1024), dbl = smem_desc(b_lo + off + k * 32, 16, 1024);
              umma_tf32_ts(tacc + a1, ta_lo + k * 8, dbh, idesc_b, (c | k) != 0);
              umma_tf32_ts(tacc + a1, ta_hi + k * 8, dbl, idesc_b, 1);
              umma_tf32_ts(tacc + a1, ta_hi + k * 8, dbh, idesc_b, 1);
            }
          }
          if (CL == 1) umma_commit(empty + s);
          else umma_commit_mcast(empty + s, (1 << CL) - 1);   // the stage is refilled by both CTAs
          if (c == p.d_chunks - 1) umma_commit(tfull + acc);
        }
        __syncwarp();
      }
    }
  } else if (warp < 4) {
    // ===== operand staging: thread = row of the tile = TMEM lane.  Reads its 32 features of the chunk
    // from the swizzled TMA tile and writes X_hi (the raw words: the tensor core ignores the 13 low
    // mantissa bits) and X_lo = X - tf32(X) into tensor memory, where all three MMAs read A from. =====
    const int rq = (warp & 3) * 32 + lane;      // a warp may only touch the 32 TMEM lanes of its quadrant
    uint32_t it = 0;
    for (int64_t job = job0; job < n_jobs; job += job_step) {
      if (job_cb(job) < 0) continue;
      for (int c = 0; c < p.d_chunks; ++c, ++it) {
        const int s = it % kFwdStages; const uint32_t ph = (it / kFwdStages) & 1;
        mbar_wait(full + s, ph);
        const uint32_t row_s = smem_u32(smem + s * L::kStage) + rq * 128;
        uint32_t hi[32], lo[32];
#pragma unroll
        for (int ch = 0; ch < 8; ++ch) {            // 128-byte swizzle: 16-byte chunk ch sits at ch ^ (row & 7)
          const float4 x = lds128(row_s + ((ch ^ (rq & 7)) << 4));
          hi[4 * ch + 0] = __float_as_uint(x.x); hi[4 * ch + 1] = __float_as_uint(x.y);
          hi[4 * ch + 2] = __float_as_uint(x.z); hi[4 * ch + 3] = __float_as_uint(x.w);
        }
#pragma unroll
        for (int e = 0; e < 32; ++e) { const float x = __uint_as_float(hi[e]); lo[e] = __float_as_uint(x - tf32_hi(x)); }
        const uint32_t ta = tmem_base + kTmemA + s * 64 + ((uint32_t)((warp & 3) * 32) << 16);
        tmem_st32(ta, hi);
        tmem_st32(ta + 32, lo);
        tmem_st_wait();
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(split + s);
      }
    }
  } else {
    if constexpr (MODE == FWD_LOGREG_MULTI || MODE == FWD_SCORE_MULTI) {
      // ===== multinomial epilogue: a fit owns T consecutive columns; warp group g takes the fits
      // g, g + G, ... of the column block; thread = row.  softmax / first arg-max over the T columns. =====
      constexpr int G = kFwdEpiGroups;
      constexpr int kMyF = 32 / G;                          // fits per warp: the host only picks NC / T <= 32
      const int q = warp & 3, grp = (warp - 4) >> 2;
      const int qrow = q * 32 + lane;
      const int et = threadIdx.x - 128;
      const int mycol = bfly_col(lane);
      // Running sums.  Training: fp64 in shared memory, one row of 17 per (warp, fit of the warp): entry c < T is the
      // intercept gradient of class c, entry 15 the loss (T = 16: entry 16, summed on its own).  The lane that
      // bfly16 leaves column c on (the even lane with bfly_col == c) owns the entry; nothing else touches it, so
      // no barrier is needed, and the rolled fit loop needs no predicated walk over register arrays.
      // Scoring: one count per fit in registers.
      int accC[kMyF];
#pragma unroll
      for (int i = 0; i < kMyF; ++i) accC[i] = 0;
      const uint32_t acc_s = smem_u32(accS) + (uint32_t)(((warp - 4) * kMyF) * 17 * 8);
      const bool own_col = !(lane & 1);
      if (MODE == FWD_LOGREG_MULTI) {
#pragma unroll
        for (int i = 0; i < kMyF; ++i) {
          if (own_col) sts64d(acc_s + (i * 17 + mycol) * 8, 0.0);
          if (lane == 0) sts64d(acc_s + (i * 17 + 16) * 8, 0.0);
        }
      }
      int cur_cb = -1;
      const int64_t part_row = ((int64_t)blockIdx.x * 4 + q) * p.n_slots;
      const uint32_t bias_s = smem_u32(colBias), held_s = smem_u32(heldTab);
      constexpr int NCH = NC / 16;
      const int T = p.T, nfit = NC / T;
      auto flush = [&](int cb) {
        if (cb < 0) return;
#pragma unroll
        for (int i = 0; i < kMyF; ++i) {
          const int f = G * i + grp;
          const int slot = cb * NC + f * T;
          const bool live = f < nfit && slot < p.n_slots;
          if (MODE == FWD_LOGREG_MULTI) {
            const uint32_t mine = acc_s + (i * 17 + mycol) * 8, loss_a = acc_s + (i * 17 + 16) * 8;
            if (own_col) {
              const double a = lds64d(mine);
              if (live && mycol < T) p.gbPart[part_row + slot + mycol] = a;
              if (live && mycol == 15 && T < 16) p.lossPart[part_row + slot] = a;
              sts64d(mine, 0.0);
            }
            if (lane == 0) {
              if (live && T == 16) p.lossPart[part_row + slot] = lds64d(loss_a);
              sts64d(loss_a, 0.0);
            }
          } else {
            if (live && lane == 0) p.cntPart[part_row + slot] = accC[i];
            accC[i] = 0;
          }
        }
      };
      int last_f = -1;
      for (int i = 0; i < kMyF; ++i) if (G * i + grp < nfit) last_f = G * i + grp;
      const bool even_T = !(T & 1);
      // Residual stores, T even.  A thread holds one ROW of a fit (T floats of that row's 128-byte line of R), so a
      // store instruction straight from registers touches 32 lines for 8 bytes each, and the L1 spends one tag
      // cycle per line: 160 per fit at T = 10, which is what bounded this epilogue (ncu: warps waiting to reuse
      // the address registers of stores still queued).  The fit's 32 x T block goes through shared memory instead
      // and leaves line-major: pair number lane + 32 k of the block (T / 2 pairs per row), 32 / (T / 2) rows and
      // about 7 lines per instruction.
      const int H = T >> 1;
      const uint32_t stage_s = smem_u32(stageS) + (uint32_t)((warp - 4) * (32 * 16 * 4));
      const int pr0 = even_T ? lane / H : 0, pc0 = even_T ? lane % H : 0, pdq = even_T ? 32 / H : 0, pdr = even_T ? 32 % H : 0;
      uint32_t tile = 0;
      for (int64_t job = job0; job < n_jobs; job += job_step, ++tile) {
        const int cb = job_cb(job); const int64_t rt = job_rt(job);
        const int slot0 = cb * NC;
        if (cb != cur_cb) {
          flush(cur_cb);
          cur_cb = cb;
          asm volatile("bar.sync 1, %0;" ::"n"(128 * G) : "memory");
          for (int c = et; c < NC; c += 128 * G) colBias[c] = slot0 + c < n_live ? __ldg(p.bt + slot0 + c) : 0.f;
          if (DML_HAS_CW(p)) for (int c = et; c < NC; c += 128 * G) colCw[2 * c] = slot0 + c < n_live ? __ldg(p.cw + (int64_t)(slot0 + c) * 2) : 1.f;
          for (int e = et; e < (kMaxLabel + 2) * NCH; e += 128 * G) {
            const int v = e / NCH, ci = e % NCH;
            uint32_t bits = 0;
            for (int j = 0; j < 16; ++j) {
              const int slot = slot0 + ci * 16 + j;
              bool held = true;
              if (slot < n_live) {
                const int sel = __ldg(p.slot_sel + slot), val = __ldg(p.slot_val + slot) & 0xff;
                held = sel < 0 || (v < kMaxLabel ? (sel == 0 && val == v) : (sel == 1 && val == v - kMaxLabel));   // mask -1: padding column
              }
              bits |= (held ? 1u : 0u) << j;
            }
            heldTab[e] = (uint16_t)bits;
          }
          asm volatile("bar.sync 1, %0;" ::"n"(128 * G) : "memory");
        }
        const int64_t grow = p.row0 + rt * 128 + qrow;
        const int64_t lrow = rt * 128 + qrow;
        const bool valid = rt < n_row_tiles && grow < p.row0 + p.rows && grow < p.n_rows_total;
        int yv = 0, fold_v = 0, hold_v = 0;
        if (valid) {
          yv = p.ycls[grow];
          fold_v = (int)(uint8_t)p.fold_id[grow] & (kMaxLabel - 1);
          hold_v = p.hold_id ? ((int)(uint8_t)p.hold_id[grow] & 1) : 0;
        }
        const uint32_t held_f = held_s + fold_v * NCH * 2, held_h = held_s + (kMaxLabel + hold_v) * NCH * 2;
        const int acc = tile & 1; const uint32_t aph = (tile >> 1) & 1;
        float* const r_row = MODE == FWD_LOGREG_MULTI ? p.R + lrow * 32 : nullptr;      // this row inside a 32-column chunk of R
        mbar_wait(tfull + acc, aph);
        tc_fence_after();
        const uint32_t tacc = tmem_base + acc * NC + ((uint32_t)(q * 32) << 16);
        if (last_f < 0) { tc_fence_before(); __syncwarp(); if (lane == 0) mbar_arrive(tempty + acc); }
#pragma unroll 1      // rolled for the same reason as the binary epilogue's chunk loop below (8 fits x 16 classes unrolled: 178 KB of SASS)
        for (int i = 0; i < kMyF; ++i) {
          const int f = G * i + grp;
          if (f >= nfit) break;
          const int c0 = f * T;
          uint32_t v[16];
          tmem_ld16(tacc + c0, v);           // T <= 16 columns of this fit (the tail of the vector is ignored)
          tmem_ld_wait();
          if (f == last_f) { tc_fence_before(); __syncwarp(); if (lane == 0) mbar_arrive(tempty + acc); }
          const uint32_t hb = valid ? (lds16u(held_f + (c0 >> 4) * 2) | lds16u(held_h + (c0 >> 4) * 2)) : 0xffffu;
          const bool held = (hb >> (c0 & 15)) & 1u;
          const bool exists = slot0 + c0 < n_live;
          // The T <= 16 classes are walked in quarters of four under a test that is uniform for the launch
          // (4 * q < T), so that T = 10 costs 12 element slots instead of 16 (the epilogue is what bounds this
          // kernel: ncu at config 4 has the tensor pipe 35 % busy).
          float z[16];
          float m = -INFINITY; int am = 0;
#pragma unroll
          for (int q4 = 0; q4 < 4; ++q4) {
            if (4 * q4 < T) {
              float b4[4];
              if (even_T) {                        // c0 even: the intercepts of a pair of classes are one 8-byte word
                const float2 b01 = lds64f(bias_s + (c0 + 4 * q4) * 4);
                b4[0] = b01.x; b4[1] = b01.y;
                if (4 * q4 + 2 < T) { const float2 b23 = lds64f(bias_s + (c0 + 4 * q4 + 2) * 4); b4[2] = b23.x; b4[3] = b23.y; }
                else { b4[2] = 0.f; b4[3] = 0.f; }
              } else {
#pragma unroll
                for (int cc = 0; cc < 4; ++cc) b4[cc] = 4 * q4 + cc < T ? lds32f(bias_s + (c0 + 4 * q4 + cc) * 4) : 0.f;
              }
#pragma unroll
              for (int cc = 0; cc < 4; ++cc) {
                const int c = 4 * q4 + cc;
                z[c] = c < T ? __uint_as_float(v[c]) + b4[cc] : -INFINITY;
                if (MODE == FWD_SCORE_MULTI) { if (z[c] > m) { m = z[c]; am = c; } }      // first arg-max
                else m = fmaxf(m, z[c]);
              }
            } else {
#pragma unroll
              for (int cc = 0; cc < 4; ++cc) z[4 * q4 + cc] = -INFINITY;
            }
          }
          if (MODE == FWD_SCORE_MULTI) {
            const int ok = (valid && held && exists && am == yv) ? 1 : 0;
            DML_ACC_N(accC, kMyF, __popc(__ballot_sync(0xffffffffu, ok)));
          } else {
            const bool train = valid && !held && exists;
            float ssum = 0.f, zy = 0.f, rs[16];
#pragma unroll
            for (int q4 = 0; q4 < 4; ++q4) {
              if (4 * q4 < T) {
#pragma unroll
                for (int cc = 0; cc < 4; ++cc) {
                  const int c = 4 * q4 + cc;
                  rs[c] = c < T ? exp_ftz(z[c] - m) : 0.f; ssum += rs[c]; if (c == yv) zy = z[c];
                }
              } else {
#pragma unroll
                for (int cc = 0; cc < 4; ++cc) rs[4 * q4 + cc] = 0.f;
              }
            }
            const float sw = (DML_HAS_CW(p) && train) ? colCw[2 * (c0 + yv)] : 1.f;      // weight of the row's class for this fit
            const float scale = train ? sw * __fdividef(1.f, ssum) : 0.f, hot = train ? sw : 0.f;
            const float loss = train ? sw * ((__logf(ssum) + m - zy) - p.loss_shift) : 0.f;
#pragma unroll
            for (int q4 = 0; q4 < 4; ++q4)
              if (4 * q4 < T) {
#pragma unroll
                for (int cc = 0; cc < 4; ++cc) { const int c = 4 * q4 + cc; rs[c] = rs[c] * scale - (c == yv ? hot : 0.f); }
              }
            if (rt < n_row_tiles) {
              // A fit's T columns sit in one 32-column chunk of R or straddle into the next one: two row
              // pointers, picked per element.  T even: the fit starts on an even column, a pair of classes
              // never straddles (chunks are 32 wide) and goes out as one 8-byte store.
              const int colb = slot0 + c0, cross = 32 - (colb & 31);
              float* const pa = r_row + ((int64_t)(colb >> 5) * p.r_chunk_stride + (colb & 31));
              float* const pb = pa + (p.r_chunk_stride - 32);
              if (even_T && colb + T <= p.ldr) {
                const uint32_t mine = stage_s + (uint32_t)(lane * T * 4);
#pragma unroll
                for (int q4 = 0; q4 < 4; ++q4)
                  if (4 * q4 < T) {
#pragma unroll
                    for (int cc = 0; cc < 4; cc += 2) {
                      const int c = 4 * q4 + cc;
                      if (c < T) sts64f(mine + c * 4, rs[c], rs[c + 1]);
                    }
                  }
                __syncwarp();
                float* const g0 = pa - lane * 32;                 // the warp's first row
                const int64_t hop = p.r_chunk_stride - 32;
                int prow = pr0, pcol = pc0;
#pragma unroll
                for (int k = 0; k < 8; ++k) {
                  if (k < H) {
                    const float2 v2 = lds64f(stage_s + (uint32_t)((lane + 32 * k) * 8));
                    const int c = 2 * pcol;
                    stg64(g0 + (prow * 32 + c) + (c >= cross ? hop : 0), v2.x, v2.y);
                    prow += pdq; pcol += pdr;
                    if (pcol >= H) { pcol -= H; ++prow; }
                  }
                }
                __syncwarp();                                     // the next fit overwrites the block
              } else {
#pragma unroll
                for (int q4 = 0; q4 < 4; ++q4)
                  if (4 * q4 < T) {
#pragma unroll
                    for (int cc = 0; cc < 4; ++cc) {
                      const int c = 4 * q4 + cc;
                      if (c < T && colb + c < p.ldr) (c >= cross ? pb : pa)[c] = rs[c];
                    }
                  }
              }
            }
            // column sums over the warp's 32 rows in fp32, then into the fp64 running sums; the loss rides along
            // as column 15 of the same butterfly when that slot is free
            if (T < 16) rs[15] = loss;                       // slot 15 held a zero
            const float colsum = bfly16(rs, lane);
            if (own_col && (mycol < T || mycol == 15)) {
              const uint32_t mine = acc_s + (i * 17 + mycol) * 8;
              sts64d(mine, lds64d(mine) + (double)colsum);
            }
            if (T == 16) {
              const float ls = warp_sum(loss);
              if (lane == 0) { const uint32_t la = acc_s + (i * 17 + 16) * 8; sts64d(la, lds64d(la) + (double)ls); }
            }
          }
        }
      }
      flush(cur_cb);
    } else {
    // ===== epilogue: thread = row of the tile (TMEM lane); kFwdEpiGroups warps per lane quadrant,
    // warp group g taking the 16-column chunks g, g + G, g + 2G, ...  Column sums over the 32 rows of
    // a warp are formed in fp32 (the loss shifted by ln 2 so the partial sums stay O(1)) and
    // accumulated in fp64 across tiles; see DESIGN.md "precision" for why the loss needs this care. =====
    constexpr int G = kFwdEpiGroups, NCH = NC / 16;
    const int q = warp & 3;                 // TMEM lane quadrant this warp may read ((warp - 6) % 4 == (warp + 2) % 4)
    const int grp = (warp - 4) >> 2;        // chunk residue of this warp
    const int qrow = (warp & 3) * 32 + lane; // row within the tile = TMEM lane
    const int et = threadIdx.x - 128;       // 0 .. 128*G-1 among the epilogue threads
    const int mycol = bfly_col(lane);
    constexpr int kMy = (NCH + G - 1) / G;  // chunks per warp (upper bound)
    double accA[kMy], accB[kMy]; int accC[kMy];
#pragma unroll
    for (int i = 0; i < kMy; ++i) { accA[i] = 0.0; accB[i] = 0.0; accC[i] = 0; }
    int cur_cb = -1;
    const int64_t part_row = ((int64_t)blockIdx.x * 4 + q) * p.n_slots;
    const uint32_t bias_s = smem_u32(colBias), held_s = smem_u32(heldTab);
    auto flush = [&](int cb) {
      if (cb < 0 || (lane & 1)) return;
#pragma unroll
      for (int i = 0; i < kMy; ++i) {
        const int ci = G * i + grp;
        const int slot = cb * NC + ci * 16 + mycol;
        if (ci < NCH && slot < p.n_slots) {
          if (MODE == FWD_LOGREG) { p.lossPart[part_row + slot] = accA[i]; p.gbPart[part_row + slot] = accB[i]; }
          else if (MODE == FWD_SCORE_CLS) p.cntPart[part_row + slot] = accC[i];
          else p.ssePart[part_row + slot] = accA[i];
        }
        accA[i] = 0.0; accB[i] = 0.0; accC[i] = 0;
      }
    };
    // last chunk this warp reads of every accumulator (-1: none, arrive straight away)
    int last_ci = -1;
#pragma unroll
    for (int i = 0; i < kMy; ++i) if (G * i + grp < NCH) last_ci = G * i + grp;
    uint32_t tile = 0;
    for (int64_t job = job0; job < n_jobs; job += job_step) {
      const int cb = job_cb(job); const int64_t rt = job_rt(job);
      if (cb < 0) continue;
      const uint32_t tile_now = tile++;
      const int slot0 = cb * NC;
      if (cb != cur_cb) {
        flush(cur_cb);
        cur_cb = cb;
        asm volatile("bar.sync 1, %0;" ::"n"(128 * G) : "memory");   // everyone is done with the old column tables
        for (int c = et; c < NC; c += 128 * G) colBias[c] = slot0 + c < n_live ? __ldg(p.bt + slot0 + c) : 0.f;
        if (MODE == FWD_LOGREG && DML_HAS_CW(p))
          for (int c = et; c < 2 * NC; c += 128 * G) colCw[c] = slot0 + (c >> 1) < n_live ? __ldg(p.cw + (int64_t)slot0 * 2 + c) : 1.f;
        for (int e = et; e < (kMaxLabel + 2) * NCH; e += 128 * G) {
          const int v = e / NCH, ci = e % NCH;
          uint32_t bits = 0;
          for (int j = 0; j < 16; ++j) {
            const int slot = slot0 + ci * 16 + j;
            bool held = true;                                        // a column that does not exist holds out every row
            if (slot < n_live) {
              const int sel = __ldg(p.slot_sel + slot), val = __ldg(p.slot_val + slot) & 0xff;
              held = sel < 0 || (v < kMaxLabel ? (sel == 0 && val == v) : (sel == 1 && val == v - kMaxLabel));   // mask -1: padding column
            }
            bits |= (held ? 1u : 0u) << j;
          }
          heldTab[e] = (uint16_t)bits;
        }
        asm volatile("bar.sync 1, %0;" ::"n"(128 * G) : "memory");
      }
      const int64_t grow = p.row0 + rt * 128 + qrow;       // global row
      const bool valid = rt < n_row_tiles && grow < p.row0 + p.rows && grow < p.n_rows_total;
      int yv = 0, fold_v = 0, hold_v = 0; float yf = 0.f;
      const int kf = tile_k(rt);                           // independent of the row loads below: one latency
      if (valid) {
        if (p.ycls) yv = p.ycls[grow];
        if (p.yreg) yf = p.yreg[grow];
        fold_v = (int)(uint8_t)p.fold_id[grow] & (kMaxLabel - 1);
        hold_v = p.hold_id ? ((int)(uint8_t)p.hold_id[grow] & 1) : 0;
      }
      const float yy = (float)yv;
      const uint32_t held_f = held_s + fold_v * NCH * 2, held_h = held_s + (kMaxLabel + hold_v) * NCH * 2;
      const int acc = tile_now & 1; const uint32_t aph = (tile_now >> 1) & 1;
      // Training mode: bit ci of hole_mask = no row of this tile trains the columns of chunk ci (no math, residual
      // zero); bit ci + 16 = the gradient kernel skips the chunk's 64-column segment as well, so nothing is stored.
      // Scoring mode: bit ci = none of the chunk's columns is scored on this tile's rows (nothing to count).
      uint32_t hole_mask = 0;
      {
        int a0, b0, a1, b1;
        ranges_of(cb, kf, a0, b0, a1, b1);
#pragma unroll
        for (int ci = 0; ci < NCH; ++ci) {
          const int c = ci * 16;
          if (MODE == FWD_LOGREG) {
            if (c >= b0 && c + 16 <= a1) {          // inside the hole [b0, a1)
              const int seg = ((slot0 + c) & ~63) - slot0;       // block-local start of the 64-column segment
              const int ch = ((slot0 + c) & ~31) - slot0;        // ... and of its 32-column chunk
              const bool unread = p.grad_skips == 1 ? (seg >= b0 && seg + 64 <= a1)
                                : p.grad_skips == 2 ? (ch >= b0 && ch + 32 <= a1) : false;
              hole_mask |= unread ? (0x10001u << ci) : (1u << ci);
            }
          } else if (kScoreBin) {
            if (!((c >= a0 && c + 16 <= b0) || (c >= a1 && c + 16 <= b1))) hole_mask |= 1u << ci;
          }
        }
      }
      mbar_wait(tfull + acc, aph);
      tc_fence_after();
      const uint32_t tacc = tmem_base + acc * NC + ((uint32_t)(q * 32) << 16);
      if (last_ci < 0) { tc_fence_before(); __syncwarp(); if (lane == 0) mbar_arrive(tempty + acc); }
      // The chunk loop stays rolled: unrolled, the body of one tile was 80 KB of SASS (3 chunks x two 16-element
      // paths), the epilogue warps missed the instruction cache on every pass ("no instruction" was the second
      // stall reason) and no chunk overlaps the next anyway (each starts with a tcgen05.ld wait).  The per-chunk
      // accumulators stay in registers: DML_ACC adds under a predicate with a compile-time index.
#define DML_ACC(arr, val) DML_ACC_N(arr, kMy, val)
#pragma unroll 1
      for (int i = 0; i < kMy; ++i) {
        const int ci = G * i + grp;
        if (ci >= NCH) break;
        const int c0 = ci * 16;
        const bool in_hole = (MODE == FWD_LOGREG || kScoreBin) && ((hole_mask >> ci) & 1u);   // the MMAs
        uint32_t v[16];                                                         // never wrote these TMEM columns
        if (!in_hole) {
          tmem_ld16(tacc + c0, v);
          tmem_ld_wait();
        }
        if (ci == last_ci) {                 // this warp's last read of the accumulator
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(tempty + acc);
        }
        if (MODE == FWD_LOGREG && ((hole_mask >> (ci + 16)) & 1u)) continue;
        if (kScoreBin && ((hole_mask >> ci) & 1u)) continue;
        // bit j set: column c0 + j does not train on (LOGREG) / is scored on (SCORE) this row
        const uint32_t heldbits = valid ? (lds16u(held_f + ci * 2) | lds16u(held_h + ci * 2)) : 0xffffu;
        if (MODE == FWD_LOGREG) {
          float ls[16], rs[16];
          if (in_hole) {
#pragma unroll
            for (int j = 0; j < 16; ++j) rs[j] = 0.f;
          } else {
            // most (tile, chunk) pairs hold out nothing: every row of a one-fold tile trains every column
            // of the other folds -- that case runs without the per-element selects
            if (__all_sync(0xffffffffu, heldbits == 0u)) {
#pragma unroll
              for (int j = 0; j < 16; ++j) {
                const float z = __uint_as_float(v[j]) + lds32f(bias_s + (c0 + j) * 4);
                const float e = exp_ftz(-fabsf(z));
                const float inv = __fdividef(1.f, 1.f + e);
                rs[j] = (z >= 0.f ? inv : e * inv) - yy;
                ls[j] = (fmaxf(z, 0.f) - yy * z) + (__logf(1.f + e) - p.loss_shift);
              }
            } else {
#pragma unroll
              for (int j = 0; j < 16; ++j) {
                const bool train = !((heldbits >> j) & 1u);
                const float z = __uint_as_float(v[j]) + lds32f(bias_s + (c0 + j) * 4);
                const float e = exp_ftz(-fabsf(z));
                const float inv = __fdividef(1.f, 1.f + e);
                const float pr = z >= 0.f ? inv : e * inv;
                rs[j] = train ? pr - yy : 0.f;
                ls[j] = train ? (fmaxf(z, 0.f) - yy * z) + (__logf(1.f + e) - p.loss_shift) : 0.f;
              }
            }
            if (DML_HAS_CW(p)) {             // class weights: the row's class picks one of the column's two weights
              const uint32_t cw_s = smem_u32(colCw) + (uint32_t)(c0 * 8 + yv * 4);
#pragma unroll
              for (int j = 0; j < 16; ++j) { const float sw = lds32f(cw_s + j * 8); rs[j] *= sw; ls[j] *= sw; }
            }
            DML_ACC(accA, (double)bfly16(ls, lane));
            if (!p.gb_in_grad) DML_ACC(accB, (double)bfly16(rs, lane));
            // 4x4 transpose of float4 inside each group of 4 lanes (rows 4g..4g+3): afterwards lane t
            // holds float4 #t of rows 4g..4g+3, so the 4 lanes of a group store 64 contiguous bytes
            // of one row per instruction instead of 16 bytes of four different rows.
            const bool o1 = lane & 1, o2 = lane & 2;
#pragma unroll
            for (int c1 = 0; c1 < 2; ++c1)
#pragma unroll
              for (int e = 0; e < 4; ++e) {
                const float snd = o1 ? rs[8 * c1 + e] : rs[8 * c1 + 4 + e];
                const float rcv = __shfl_xor_sync(0xffffffffu, snd, 1);
                if (o1) rs[8 * c1 + e] = rcv; else rs[8 * c1 + 4 + e] = rcv;
              }
#pragma unroll
            for (int cc = 0; cc < 2; ++cc)
#pragma unroll
              for (int e = 0; e < 4; ++e) {
                const float snd = o2 ? rs[4 * cc + e] : rs[4 * cc + 8 + e];
                const float rcv = __shfl_xor_sync(0xffffffffu, snd, 2);
                if (o2) rs[4 * cc + e] = rcv; else rs[4 * cc + 8 + e] = rcv;
              }
          }
          const int tq = lane & 3;                                // which float4 of the 16-column chunk
          const int64_t lrow4 = rt * 128 + q * 32 + (lane & ~3);   // first row of this lane group
          if (rt < n_row_tiles && slot0 + c0 + 4 * tq < p.ldr) {
#pragma unroll
            for (int j = 0; j < 4; ++j) {                         // j = row inside the group
              const int col = slot0 + c0 + 4 * tq;
              const int64_t o = (int64_t)(col >> 5) * p.r_chunk_stride + (lrow4 + j) * 32 + (col & 31);
              *reinterpret_cast<float4*>(p.R + o) = make_float4(rs[4 * j], rs[4 * j + 1], rs[4 * j + 2], rs[4 * j + 3]);
            }
          }
        } else if (MODE == FWD_SCORE_CLS) {
          int ok[16];
#pragma unroll
          for (int j = 0; j < 16; ++j) {
            const bool held = valid && ((heldbits >> j) & 1u) && slot0 + c0 + j < p.n_slots;
            const float z = __uint_as_float(v[j]) + lds32f(bias_s + (c0 + j) * 4);
            ok[j] = (held && (((z > 0.f) ? 1 : 0) == yv)) ? 1 : 0;
          }
          DML_ACC(accC, bfly16(ok, lane));
        } else {
          float se[16];
#pragma unroll
          for (int j = 0; j < 16; ++j) {
            const bool held = valid && ((heldbits >> j) & 1u) && slot0 + c0 + j < p.n_slots;
            const float e = yf - (__uint_as_float(v[j]) + lds32f(bias_s + (c0 + j) * 4));
            se[j] = held ? e * e : 0.f;
          }
          DML_ACC(accA, (double)bfly16(se, lane));
        }
      }
    }
#undef DML_ACC
#undef DML_ACC_N
    flush(cur_cb);
    }
  }
  tc_fence_before();
  __syncthreads();
  if (CL > 1) cluster_sync_all();           // no CTA leaves while its peer may still write into it
  if (warp == kFwdMmaWarp) tmem_dealloc(tmem_base, kTmemCols);
}

// ---------------------------------------------------------------- gradient kernel
// G[j][slot] = sum over the split's rows of X[row][j] * R[row][slot]: A = X^T (features = TMEM lanes,
// gathered column-wise from the TMA tile and written to tensor memory as hi / lo), B = R (MN-major,
// UMMA layout SWIZZLE_128B_BASE32B + TMA SWIZZLE_128B_ATOM_32B, the only layout tf32 accepts
// transposed).  Two-level accumulation: the tensor core's fp32 accumulator truncates, so chains stay
// short (kChain stages = 128 rows, measured 1.3e-5 relative after 2048 rows in one chain) and the drain
// warps add each finished TMEM tile into fp32 registers with round-to-nearest while the MMAs fill the
// other accumulator.
constexpr int kChain = 4;           // stages per TMEM accumulation chain
struct TcGradParams {
  int64_t row0_x;            // first row of the chunk inside X
  int64_t rows;              // rows in the chunk (R is chunk-local)
  int64_t rows_per_split;    // multiple of 32
  int d, n_slots;
  float* Gpart; int ldg;     // [splits][n_slots][ldg]
  int mask_hi;
  int prefetch_stages;       // L2 prefetch distance of the operand boxes, in 32-row stages (0 = off)
  // fold-aware skipping (see TcFwdParams): a 128-row chain whose rows all sit in fold k contributes
  // nothing to fold k's columns; 64-column segments inside that group are not loaded or multiplied
  const uint8_t* tile_fold; const int* gstart; int n_groups;
  // with gstart set: seg_skip = 1 skips whole 64-column segments as described above (N = 64 MMAs); 0 keeps the
  // N = ncols MMAs and only leaves the 32-column chunks of the hole unread -- their TMA load is aimed past the
  // end of the residual array, which fills the shared-memory tile with zeros without touching HBM, and the
  // forward kernel does not store them (TcFwdParams::grad_skips = 2)
  int seg_skip;
  // cross terms x_lo r_hi + x_hi r_lo as bf16 MMAs over a K-concatenated operand (128-column blocks only):
  // 8 MMAs per 32-row stage instead of 12; a cross term is 2^-11 of the product and bf16 keeps 8 bits of it
  int cross_bf16;
  const int* n_list_dev; int list_T;   // live columns = *n_list_dev * list_T (see TcFwdParams); n_slots stays the pitch
  // column sums of R over the split's rows, [splits][n_slots] doubles (sum r = the intercept's gradient): the
  // split warps read every residual anyway, which saves the forward epilogue one 16-column butterfly per chunk.
  // Written by the first feature-block pair only; null = off (the forward kernel sums r itself).
  double* gb2;
};

// ---------------------------------------------------------------- gradient kernel, CTA pair
// Same contraction with cta_group::2 MMAs (M = 256): the two CTAs of a cluster take feature blocks
// 2k and 2k+1 and share the R operand -- each loads and holds half of the block's columns, so the
// shared-memory traffic of the B operand is halved per SM.
//   leader (rank 0): issues every MMA; its fullB / split / tempty barriers also count the peer's
//                    TMA bytes, staging warps and drain warps (remote arrives)
//   both CTAs      : TMA producer (own X boxes -> own fullA; own R half -> the leader's fullB),
//                    staging warps (X^T hi/lo into own tensor memory), drain warps (own features)
//   tcgen05.commit multicasts "stage free" and "accumulator full" to both CTAs.
// Column blocks are 128 wide, the last one 64 when that covers the rest (N/2 must be whole 32-column
// chunks of the MN-major swizzled layout).  Used when d spans an even number of 128-feature blocks.
// Three rings so that the look-ahead is set by shared memory, not by tensor memory: X boxes (freed as
// soon as the staging warps have read them), R halves (freed by the MMAs) and the A slots in tensor
// memory (four, freed by the MMAs).  Measured with a single 4-stage ring: the MMA warp waited for data
// a third of the time (HBM latency under load exceeds 4 x 768 clocks).
constexpr int kGrad2SX = 6, kGrad2SR = 6, kGrad2TA = 4;
constexpr int kGrad2NC = 128;
// warps 0-3 X staging, 4-7 drain, 8-9 R_lo split, 10 TMA producer, 11 MMA issuer
constexpr int kGrad2Threads = 384, kGrad2SplitWarp = 8, kGrad2TmaWarp = 10, kGrad2MmaWarp = 11;
struct Grad2Smem {
  static constexpr int kA = 4 * 32 * 128;                    // own X boxes of one stage
  static constexpr int kBh = (kGrad2NC / 2 / 32) * 32 * 128; // own half of R (fp32; read as R_hi) and of R_lo
  static constexpr int kB = 2 * kBh;
  static constexpr int kBytes = kGrad2SX * kA + kGrad2SR * kB + 1024 + 512 + (kMaxLabel + 2) * 4 + 64 + 4096 + 16;   // + column-group starts, chain fold ids, column-sum staging
};

// R is the fp32 residual as the forward kernel left it: the MMAs read it as R_hi (kind::tf32 ignores the
// 13 low mantissa bits) and two warps per CTA write R_lo = R - tf32(R) next to it in shared memory, so
// the residual crosses HBM once in each direction instead of twice.
// A chain = 4 stages = one 128-row tile of X.  Normally the leader holds the first half of the block's
// columns, the peer the second half, and one N = ncols MMA covers the block.  A column block is also two
// 64-column segments: when all rows of the chain sit in fold k and a segment lies inside fold k's column
// group, that segment is skipped for the chain by every role (no TMA, no split, no MMA, nothing added by
// the drain warps); the other segment is then loaded interleaved (leader: its even 32-column chunk, peer:
// the odd one) and multiplied by N = 64 MMAs, which keeps the accumulator columns in natural order.
__global__ void __launch_bounds__(kGrad2Threads, 1)
tc_grad2_kernel(const __grid_constant__ CUtensorMap mapX, const __grid_constant__ CUtensorMap mapR,
                const TcGradParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  using L = Grad2Smem;
  constexpr int SX = kGrad2SX, SR = kGrad2SR, TA = kGrad2TA;
  uint8_t* smX = smem;
  uint8_t* smR = smem + SX * L::kA;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smR + SR * L::kB);
  uint64_t* fullA = bars; uint64_t* emptyA = fullA + SX;
  uint64_t* fullR = emptyA + SX;      // own R half landed (local)
  uint64_t* bsplit = fullR + SR;      // leader: R_lo of both CTAs written (4 remote/local arrives)
  uint64_t* emptyB = bsplit + SR;
  uint64_t* split = emptyB + SR; uint64_t* aempty = split + TA;
  uint64_t* tfull = aempty + TA;      // [2]
  uint64_t* tempty = tfull + 2;       // [2]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty + 2);
  int* gsTab = reinterpret_cast<int*>(tmem_slot + 2);        // [n_groups + 1] copy of p.gstart
  uint8_t* tfTab = reinterpret_cast<uint8_t*>(gsTab + kMaxLabel + 2);   // [<= 64] fold id of each chain's row tile
  double* gsm = reinterpret_cast<double*>((reinterpret_cast<uintptr_t>(tfTab + 64) + 15) & ~(uintptr_t)15);   // [2][32][8] column-sum staging
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank();
  const bool leader = rank == 0;
  constexpr uint32_t kTmemCols = 512, kTmemA = 2 * kGrad2NC;   // two accumulators, then 64 columns of A per slot
  static_assert(2 * kGrad2NC + TA * 64 <= 512, "tensor memory");
  // a column block past the live columns (the host sized the grid from an upper bound): both CTAs of the
  // pair share blockIdx.y and leave before anything is set up
  if (p.n_list_dev && (int)blockIdx.y * kGrad2NC >= min(p.n_slots, *p.n_list_dev * p.list_T)) return;

  if (warp == kGrad2TmaWarp && lane == 0) {
    for (int s = 0; s < SX; ++s) { mbar_init(fullA + s, 1); mbar_init(emptyA + s, 4); }
    for (int s = 0; s < SR; ++s) { mbar_init(fullR + s, 1); mbar_init(bsplit + s, 4); mbar_init(emptyB + s, 1); }
    for (int s = 0; s < TA; ++s) { mbar_init(split + s, 8); mbar_init(aempty + s, 1); }
    for (int a = 0; a < 2; ++a) { mbar_init(tfull + a, 1); mbar_init(tempty + a, 8); }
    fence_barrier_init();
  }
  if (warp == kGrad2MmaWarp) tmem_alloc_pair(tmem_slot, kTmemCols);
  const int jb = blockIdx.x, cb = blockIdx.y, sp = blockIdx.z;
  const int col0 = cb * kGrad2NC;
  const int n_live = p.n_list_dev ? min(p.n_slots, *p.n_list_dev * p.list_T) : p.n_slots;
  const int ncols = min(kGrad2NC, max(0, ((n_live - col0 + 63) / 64) * 64));   // columns of this pair's block
  const int nseg = ncols / 64;                                            // 64-column segments = chunks per CTA
  const int64_t kb = (int64_t)sp * p.rows_per_split;
  const int64_t ke = min(p.rows, kb + p.rows_per_split);
  const int n_stages = (int)((ke - kb + 31) / 32);
  const int n_groups = (n_stages + kChain - 1) / kChain;
  const bool holes = p.gstart != nullptr && n_groups <= 64;
  if (holes) {              // the tables every role consults once per chain live in shared memory
    for (int i = threadIdx.x; i <= p.n_groups; i += blockDim.x) gsTab[i] = p.gstart[i];
    for (int i = threadIdx.x; i < n_groups; i += blockDim.x) tfTab[i] = p.tile_fold[(p.row0_x + kb) / 128 + i];
  }
  tc_fence_before();
  cluster_sync_all();       // barriers of both CTAs initialised before any remote arrive or multicast
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  // bit s set: segment s is skipped for chain g (rows_per_split and row0_x are multiples of 128)
  auto skip_of = [&](int g) -> uint32_t {
    if (!holes || !p.seg_skip) return 0u;
    const int k = tfTab[g];
    if (k >= p.n_groups) return 0u;
    uint32_t m = 0;
    for (int sg = 0; sg < nseg; ++sg) {
      const int c = col0 + sg * 64;
      if (c >= gsTab[k] && c + 64 <= gsTab[k + 1]) m |= 1u << sg;
    }
    return m;
  };

  // bit c set: chunk c of this CTA's half lies inside the column group of chain g's fold (zero-filled, not read)
  auto zero_of = [&](int g) -> uint32_t {
    if (!holes || p.seg_skip) return 0u;
    const int k = tfTab[g];
    if (k >= p.n_groups) return 0u;
    uint32_t m = 0;
    for (int c = 0; c < nseg; ++c) {
      const int col = col0 + ((int)rank * nseg + c) * 32;
      if (col >= gsTab[k] && col + 32 <= gsTab[k + 1]) m |= 1u << c;
    }
    return m;
  };
  constexpr int kPastEnd = 1 << 20;   // chunk coordinate beyond every residual array

  const bool use_bf = p.cross_bf16 != 0 && nseg == 2;   // only chains without a skipped segment take this form

  if (warp == kGrad2TmaWarp) {
    if (lane == 0) {
      uint32_t skip = 0, zfill = 0;
      for (int it = 0; it < n_stages; ++it) {
        if (it % kChain == 0) skip = skip_of(it / kChain);
        const int sx = it % SX, sr = it % SR;
        const int lr = (int)(kb + (int64_t)it * 32);
        const int xr = (int)(p.row0_x + lr);
        mbar_wait(emptyA + sx, ((it / SX) & 1) ^ 1);
        uint8_t* ax = smX + sx * L::kA;
        mbar_expect_tx(fullA + sx, L::kA);
#pragma unroll
        for (int b = 0; b < 4; ++b) tma_load_2d(&mapX, fullA + sx, ax + b * 4096, jb * 128 + b * 32, xr);
        mbar_wait(emptyB + sr, ((it / SR) & 1) ^ 1);
        uint8_t* br = smR + sr * L::kB;
        if (skip == 0) {           // whole block: this CTA's half = nseg consecutive chunks
          if (it % kChain == 0) zfill = zero_of(it / kChain);
          mbar_expect_tx(fullR + sr, (uint32_t)(nseg * 4096));
          for (int c = 0; c < nseg; ++c)
            tma_load_3d(&mapR, fullR + sr, br + c * 4096, 0, lr, ((zfill >> c) & 1u) ? kPastEnd : col0 / 32 + (int)rank * nseg + c);
        } else {                   // segments: chunk (2 sg + rank) of the block, at position sg
          int n_load = 0;
          for (int sg = 0; sg < nseg; ++sg) if (!((skip >> sg) & 1u)) ++n_load;
          if (n_load == 0) mbar_arrive(fullR + sr);
          else {
            mbar_expect_tx(fullR + sr, (uint32_t)(n_load * 4096));
            for (int sg = 0; sg < nseg; ++sg)
              if (!((skip >> sg) & 1u)) tma_load_3d(&mapR, fullR + sr, br + sg * 4096, 0, lr, col0 / 32 + 2 * sg + (int)rank);
          }
        }
      }
    }
  } else if (warp == kGrad2MmaWarp) {
    // The whole warp runs the loop and one elected lane issues: with a single-lane branch around the loop
    // the compiler cannot keep descriptors in uniform registers and the launch takes 1.16 ms instead of 0.78.
    if (leader) {
      constexpr uint32_t idesc_seg = instr_desc_tf32(256, 64, 0, 1);
      const uint32_t idesc_all = instr_desc_tf32(256, ncols, 0, 1), idesc_bf = instr_desc_bf16(256, ncols, 0, 1);
      // descriptors of stage slot 0; a slot / K step / lo buffer adds a constant to the address field
      const uint64_t dR0 = smem_desc(smem_u32(smR), 4096, 512, kLayoutSw128Base32);
      const uint64_t dB0 = smem_desc(smem_u32(smR) + L::kBh, 8192, 1024, kLayoutSw128);
      for (int g = 0, it = 0; g < n_groups; ++g) {
        const int acc = g & 1; const uint32_t aph = (g >> 1) & 1;
        const uint32_t skip = skip_of(g);
        mbar_wait_cluster(tempty + acc, aph ^ 1);
        tc_fence_after();
        const uint32_t tacc = tmem_base + acc * kGrad2NC;
        const int g_end = min(n_stages, (g + 1) * kChain);
        for (uint32_t first = 0; it < g_end; ++it, first = 1) {
          const int sr = it % SR, ta = it % TA;
          mbar_wait_cluster(bsplit + sr, (it / SR) & 1);
          mbar_wait_cluster(split + ta, (it / TA) & 1);
          tc_fence_after();
          if (elect_one()) {
            const uint32_t ta_hi = tmem_base + kTmemA + ta * 64, ta_lo = ta_hi + 32;
            const uint64_t dh = dR0 + (uint64_t)((sr * L::kB) >> 4), dl = dh + (L::kBh >> 4);
            if (skip == 0 && use_bf) {
              // B' = [bf16(R) rows 0-31 ; bf16(R_lo) rows 32-63], 64 columns per CTA, MN-major SWIZZLE_128B;
              // A' = [x_lo | x_hi] as bf16 pairs in columns 32..63 of the slot: 16 K values per MMA
              const uint64_t db = dB0 + (uint64_t)((sr * L::kB) >> 4);
#pragma unroll
              for (int j = 0; j < 4; ++j) umma_bf16_ts_pair(tacc, ta_lo + j * 8, db + j * (2048 >> 4), idesc_bf, j == 0 ? first : 1u);
#pragma unroll
              for (int k = 0; k < 4; ++k) umma_tf32_ts_pair(tacc, ta_hi + k * 8, dh + k * (1024 >> 4), idesc_all, 1);
            } else if (skip == 0) {
#pragma unroll
              for (int k = 0; k < 4; ++k) {
                umma_tf32_ts_pair(tacc, ta_lo + k * 8, dh + k * (1024 >> 4), idesc_all, k == 0 ? first : 1u);
                umma_tf32_ts_pair(tacc, ta_hi + k * 8, dl + k * (1024 >> 4), idesc_all, 1);
                umma_tf32_ts_pair(tacc, ta_hi + k * 8, dh + k * (1024 >> 4), idesc_all, 1);
              }
            } else {
              for (int sg = 0; sg < nseg; ++sg) {
                if ((skip >> sg) & 1u) continue;
                const uint64_t sh = dh + sg * (4096 >> 4), sl = dl + sg * (4096 >> 4);
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                  umma_tf32_ts_pair(tacc + sg * 64, ta_lo + k * 8, sh + k * (1024 >> 4), idesc_seg, k == 0 ? first : 1u);
                  umma_tf32_ts_pair(tacc + sg * 64, ta_hi + k * 8, sl + k * (1024 >> 4), idesc_seg, 1);
                  umma_tf32_ts_pair(tacc + sg * 64, ta_hi + k * 8, sh + k * (1024 >> 4), idesc_seg, 1);
                }
              }
            }
            umma_commit_pair(emptyB + sr);
            umma_commit_pair(aempty + ta);
            if (it == g_end - 1) umma_commit_pair(tfull + acc);
          }
          __syncwarp();
        }
      }
    }
  } else if (warp >= kGrad2SplitWarp && warp < kGrad2TmaWarp) {
    // R_lo = R - tf32(R), elementwise on the swizzled tile (position-preserving): each of the two warps
    // takes half of every loaded 4 KB chunk; then the buffer is handed to the async proxy (the MMAs)
    const int half = warp - kGrad2SplitWarp;
    uint32_t skip = 0;
    // column sums of R (p.gb2): a lane always meets the same four columns of a chunk -- its 16-byte slot is
    // lane & 7 and the 32-byte swizzle of the tile depends on row & 3 = (lane >> 3) & 3 for all of its rows
    const bool want_gb = p.gb2 != nullptr && blockIdx.x < 2;
    double gs[2][4] = {{0.0, 0.0, 0.0, 0.0}, {0.0, 0.0, 0.0, 0.0}};
    float fs[2][4] = {{0.f, 0.f, 0.f, 0.f}, {0.f, 0.f, 0.f, 0.f}};   // fp32 over one chain (16 values), fp64 across chains
    for (int it = 0; it < n_stages; ++it) {
      if (it % kChain == 0) {
        skip = skip_of(it / kChain);
#pragma unroll
        for (int q = 0; q < 8; ++q) { gs[q >> 2][q & 3] += (double)fs[q >> 2][q & 3]; fs[q >> 2][q & 3] = 0.f; }
      }
      const int sr = it % SR;
      mbar_wait(fullR + sr, (it / SR) & 1);
      const uint32_t hi_s = smem_u32(smR + sr * L::kB), lo_s = hi_s + L::kBh;
      if (use_bf && skip == 0) {
        // unit = (row k, 8 columns): 32 source bytes (one swizzled sub-chunk of the TMA tile) -> 16 bytes of
        // bf16(R) in row k and 16 bytes of bf16(R - tf32(R)) in row 32 + k of the bf16 tile
#pragma unroll
        for (int u4 = 0; u4 < 4; ++u4) {
          const int u = u4 * 64 + half * 32 + lane, k = u >> 3, c8 = u & 7;
          const uint32_t src = hi_s + (c8 >> 2) * 4096 + k * 128 + (((c8 & 3) ^ (k & 3)) << 5);
          const float4 a = lds128(src), b = lds128(src + 16);
          float4 h, l;
          h.x = __uint_as_float(pack_bf16(a.x, a.y)); h.y = __uint_as_float(pack_bf16(a.z, a.w));
          h.z = __uint_as_float(pack_bf16(b.x, b.y)); h.w = __uint_as_float(pack_bf16(b.z, b.w));
          l.x = __uint_as_float(pack_bf16(a.x - tf32_hi(a.x), a.y - tf32_hi(a.y)));
          l.y = __uint_as_float(pack_bf16(a.z - tf32_hi(a.z), a.w - tf32_hi(a.w)));
          l.z = __uint_as_float(pack_bf16(b.x - tf32_hi(b.x), b.y - tf32_hi(b.y)));
          l.w = __uint_as_float(pack_bf16(b.z - tf32_hi(b.z), b.w - tf32_hi(b.w)));
          const uint32_t dst = lo_s + k * 128 + ((c8 ^ (k & 7)) << 4);
          sts128(dst, h);
          sts128(dst + 32 * 128, l);
        }
      } else
      for (int sg = 0; sg < nseg; ++sg) {
        if ((skip >> sg) & 1u) continue;
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          const int i = sg * 256 + half * 128 + u * 32 + lane;       // 16-byte piece of the chunk
          const float4 v = lds128(hi_s + i * 16);
          sts128(lo_s + i * 16, make_float4(v.x - tf32_hi(v.x), v.y - tf32_hi(v.y), v.z - tf32_hi(v.z), v.w - tf32_hi(v.w)));
          if (want_gb) {
            if (sg == 0) { fs[0][0] += v.x; fs[0][1] += v.y; fs[0][2] += v.z; fs[0][3] += v.w; }
            else         { fs[1][0] += v.x; fs[1][1] += v.y; fs[1][2] += v.z; fs[1][3] += v.w; }
          }
        }
      }
      fence_proxy_async();
      __syncwarp();
      if (lane == 0) mbar_arrive_cluster(mapa_rank(smem_u32(bsplit + sr), 0));
    }
    if (want_gb) {      // 8 lanes (4 per warp) hold partial sums of each column: add them in a fixed order
#pragma unroll
      for (int sg = 0; sg < 2; ++sg)
#pragma unroll
        for (int c = 0; c < 4; ++c) gsm[(half * 32 + lane) * 8 + sg * 4 + c] = gs[sg][c] + (double)fs[sg][c];
      asm volatile("bar.sync 2, 64;" ::: "memory");
      const int x = half * 32 + lane;                  // column of this CTA's half: chunk x / 32, column x % 32
      if (x < nseg * 32) {
        const int sg = x >> 5, cg = (x & 31) >> 2, cc = x & 3;
        double a = 0.0;
        for (int h = 0; h < 2; ++h)
          for (int l = 0; l < 32; ++l)
            if ((((((l & 7) >> 1) ^ ((l >> 3) & 3)) << 1) | (l & 1)) == cg) a += gsm[(h * 32 + l) * 8 + sg * 4 + cc];
        const int slot = col0 + ((int)rank * nseg + sg) * 32 + (x & 31);     // the CTA holds chunks rank * nseg + sg
        if (slot < p.n_slots) p.gb2[(int64_t)sp * p.n_slots + slot] = a;
      }
    }
  } else if (warp < 4) {
    const int qd = warp & 3;
    uint32_t sk = 0;
    for (int it = 0; it < n_stages; ++it) {
      const int sx = it % SX, ta = it % TA;
      mbar_wait(fullA + sx, (it / SX) & 1);
      const uint32_t box = smem_u32(smX + sx * L::kA) + qd * 4096;
      uint32_t hi[32], lo[32];
#pragma unroll
      for (int k = 0; k < 32; ++k)
        hi[k] = lds32u(box + k * 128 + ((((lane >> 3) ^ (k & 3)) << 5) | ((lane & 7) << 2)));
#pragma unroll
      for (int e = 0; e < 32; ++e) { const float x = __uint_as_float(hi[e]); lo[e] = __float_as_uint(x - tf32_hi(x)); }
      __syncwarp();
      if (lane == 0) mbar_arrive(emptyA + sx);              // the box is in registers: the producer may refill it
      mbar_wait(aempty + ta, ((it / TA) & 1) ^ 1);          // the MMAs that read this A slot are done
      tc_fence_after();
      const uint32_t taddr = tmem_base + kTmemA + ta * 64 + ((uint32_t)(qd * 32) << 16);
      tmem_st32(taddr, hi);
      if (use_bf && (it % kChain == 0 ? (sk = skip_of(it / kChain)) : sk) == 0) {
        uint32_t pk[32];       // columns 32..47: x_lo as bf16 pairs (K order), 48..63: x as bf16 pairs
#pragma unroll
        for (int e = 0; e < 16; ++e) {
          pk[e] = pack_bf16(__uint_as_float(lo[2 * e]), __uint_as_float(lo[2 * e + 1]));
          pk[16 + e] = pack_bf16(__uint_as_float(hi[2 * e]), __uint_as_float(hi[2 * e + 1]));
        }
        tmem_st32(taddr + 32, pk);
      } else {
        tmem_st32(taddr + 32, lo);
      }
      tmem_st_wait();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive_cluster(mapa_rank(smem_u32(split + ta), 0));
    }
  } else {
    const int q = warp & 3;
    const int j = jb * 128 + q * 32 + lane;
    float sum[kGrad2NC];
#pragma unroll
    for (int c = 0; c < kGrad2NC; ++c) sum[c] = 0.f;
    for (int g = 0; g < n_groups; ++g) {
      const int acc = g & 1; const uint32_t aph = (g >> 1) & 1;
      const uint32_t skip = skip_of(g);
      mbar_wait(tfull + acc, aph);
      tc_fence_after();
      const uint32_t tacc = tmem_base + acc * kGrad2NC + ((uint32_t)(q * 32) << 16);
#pragma unroll
      for (int c0 = 0; c0 < kGrad2NC; c0 += 16) {
        if (c0 < ncols && !((skip >> (c0 / 64)) & 1u)) {
          uint32_t v[16];
          tmem_ld16(tacc + c0, v);
          tmem_ld_wait();
#pragma unroll
          for (int c = 0; c < 16; ++c) sum[c0 + c] += __uint_as_float(v[c]);
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive_cluster(mapa_rank(smem_u32(tempty + acc), 0));
    }
    float* out = p.Gpart + (int64_t)sp * p.n_slots * p.ldg;
#pragma unroll
    for (int c = 0; c < kGrad2NC; ++c) {
      const int slot = col0 + c;
      if (c < ncols && slot < n_live && j < p.d) out[(int64_t)slot * p.ldg + j] = sum[c];
    }
  }
  tc_fence_before();
  cluster_sync_all();       // the peer's shared and tensor memory stay alive until the leader's MMAs are done
  if (warp == kGrad2MmaWarp) tmem_dealloc_pair(tmem_base, kTmemCols);
}

// ---------------------------------------------------------------- Gram kernel (Ridge)
// Mpart[fold][split][i][c] = sum over the split's rows of that fold of X[row][i] * X[row][c]
// for a 128 x 128 block (ib, cb), cb >= ib (the moment matrix is symmetric; the reducer mirrors).
//   A = X^T block ib: staged into tensor memory (hi, lo) with the rows of other folds zeroed
//   B = X block cb  : four TMA boxes [32 rows x 32 features]; B_lo written to shared memory by
//                     four extra warps (elementwise on the swizzled buffer)
// 32-row stages without a row of the fold are skipped by every role (per-16-row presence bits).
// Replaces the sgemm in sklearn's _solve_cholesky (sklearn/linear_model/_ridge.py:220) for all folds
// of cross_val_score (worker.py:326/341) in one pass over X.
constexpr int kGramStages = 3;
constexpr int kGramThreads = 448;   // warps 0-3 A staging, 4-7 B split, 8-11 drain, 12 TMA, 13 MMA
constexpr int kGramNC = 128;

struct TcGramParams {
  int64_t n; int64_t rows_per_split; int d; int n_blocks;   // n_blocks = ceil(d / 128)
  const int8_t* label; const uint16_t* blockmask;            // presence bits per 16 rows (bit f: fold f occurs)
  int splits;
  float* Mpart; int ldm;                                     // [folds][splits][n_blocks*128][ldm], zero-initialised
};
struct GramSmem {
  static constexpr int kA = 4 * 4096, kB = 4 * 4096;
  static constexpr int kStage = kA + 2 * kB;
  static constexpr int kBytes = kGramStages * kStage + 1024 + 256;
};

__global__ void __launch_bounds__(kGramThreads, 1)
tc_gram_kernel(const __grid_constant__ CUtensorMap mapX, const TcGramParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  using L = GramSmem;
  constexpr int NC = kGramNC;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + kGramStages * L::kStage);
  uint64_t* full = bars; uint64_t* split = bars + kGramStages; uint64_t* empty = bars + 2 * kGramStages;
  uint64_t* tfull = bars + 3 * kGramStages; uint64_t* tempty = tfull + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty + 2);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  constexpr uint32_t kTmemCols = 512, kTmemA = 2 * NC;       // 2 x 128 accumulator + 3 x 64 operand columns

  // block pair (ib <= cb) from the linear index
  int ib = 0, cb = 0;
  { int t = blockIdx.x; for (ib = 0; ib < p.n_blocks; ++ib) { const int len = p.n_blocks - ib; if (t < len) { cb = ib + t; break; } t -= len; } }
  const int fold = blockIdx.z / p.splits, sp = blockIdx.z % p.splits;
  const int64_t kb = (int64_t)sp * p.rows_per_split;
  const int64_t ke = min(p.n, kb + p.rows_per_split);
  const int n_stages = (int)((ke - kb + 31) / 32);
  const uint16_t* bm = p.blockmask + (kb >> 4);
  const int64_t bm_len = ((p.n + 15) >> 4) - (kb >> 4);
  auto stage_active = [&](int it) -> bool {
    unsigned m = bm[2 * it];
    if (2 * it + 1 < bm_len) m |= bm[2 * it + 1];
    return (m >> fold) & 1u;
  };
  int n_act = 0;
  for (int it = 0; it < n_stages; ++it) n_act += stage_active(it) ? 1 : 0;
  if (n_act == 0) return;                                   // nothing of this fold in the split (uniform exit)
  const int n_groups = (n_act + kChain - 1) / kChain;

  if (warp == 12 && lane == 0) {
    for (int s = 0; s < kGramStages; ++s) { mbar_init(full + s, 1); mbar_init(split + s, 8); mbar_init(empty + s, 1); }
    for (int a = 0; a < 2; ++a) { mbar_init(tfull + a, 1); mbar_init(tempty + a, 4); }
    fence_barrier_init();
  }
  if (warp == 13) tmem_alloc(tmem_slot, kTmemCols);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 12) {
    if (lane == 0) {
      for (int it = 0, ai = 0; it < n_stages; ++it) {
        if (!stage_active(it)) continue;
        const int s = ai % kGramStages; const uint32_t ph = (ai / kGramStages) & 1; ++ai;
        mbar_wait(empty + s, ph ^ 1);
        uint8_t* st = smem + s * L::kStage;
        mbar_expect_tx(full + s, L::kA + L::kB);
        const int xr = (int)(kb + (int64_t)it * 32);
#pragma unroll
        for (int b = 0; b < 4; ++b) {
          tma_load_2d(&mapX, full + s, st + b * 4096, ib * 128 + b * 32, xr);
          tma_load_2d(&mapX, full + s, st + L::kA + b * 4096, cb * 128 + b * 32, xr);
        }
      }
    }
  } else if (warp == 13) {
    constexpr uint32_t idesc = instr_desc_tf32(128, NC, 0, 1);
    for (int g = 0, ai = 0; g < n_groups; ++g) {
      const int acc = g & 1; const uint32_t aph = (g >> 1) & 1;
      mbar_wait(tempty + acc, aph ^ 1);
      tc_fence_after();
      const uint32_t tacc = tmem_base + acc * NC;
      const int g_end = min(n_act, (g + 1) * kChain);
      for (int first = 1; ai < g_end; ++ai) {
        const int s = ai % kGramStages; const uint32_t ph = (ai / kGramStages) & 1;
        mbar_wait(full + s, ph);
        mbar_wait(split + s, ph);
        tc_fence_after();
        if (elect_one()) {
          const uint32_t b_hi = smem_u32(smem + s * L::kStage) + L::kA, b_lo = b_hi + L::kB;
          const uint32_t ta_hi = tmem_base + kTmemA + s * 64, ta_lo = ta_hi + 32;
#pragma unroll
          for (int k = 0; k < 4; ++k) {
            const uint64_t dbh = smem_desc(b_hi + k * 1024, 4096, 512, kLayoutSw128Base32);
            const uint64_t dbl = smem_desc(b_lo + k * 1024, 4096, 512, kLayoutSw128Base32);
            umma_tf32_ts(tacc, ta_lo + k * 8, dbh, idesc, (first && k == 0) ? 0u : 1u);
            umma_tf32_ts(tacc, ta_hi + k * 8, dbl, idesc, 1);
            umma_tf32_ts(tacc, ta_hi + k * 8, dbh, idesc, 1);
          }
          umma_commit(empty + s);
          if (ai == g_end - 1) umma_commit(tfull + acc);
        }
        __syncwarp();
        first = 0;
      }
    }
  } else if (warp < 4) {
    // A staging: lane = feature of block ib; rows of other folds (and rows past n) contribute zero
    const int qd = warp & 3;
    for (int it = 0, ai = 0; it < n_stages; ++it) {
      if (!stage_active(it)) continue;
      const int s = ai % kGramStages; const uint32_t ph = (ai / kGramStages) & 1; ++ai;
      const int64_t row0 = kb + (int64_t)it * 32;
      uint32_t keep = 0;                                      // bit k: row k of the stage belongs to the fold
      { const int64_t r = row0 + lane; const int ok = (r < p.n) && (p.label[r] == fold); keep = __ballot_sync(0xffffffffu, ok); }
      mbar_wait(full + s, ph);
      const uint32_t box = smem_u32(smem + s * L::kStage) + qd * 4096;
      uint32_t hi[32], lo[32];
#pragma unroll
      for (int k = 0; k < 32; ++k) {
        const uint32_t w = lds32u(box + k * 128 + ((((lane >> 3) ^ (k & 3)) << 5) | ((lane & 7) << 2)));
        hi[k] = ((keep >> k) & 1u) ? w : 0u;
      }
#pragma unroll
      for (int e = 0; e < 32; ++e) { const float x = __uint_as_float(hi[e]); lo[e] = __float_as_uint(x - tf32_hi(x)); }
      const uint32_t ta = tmem_base + kTmemA + s * 64 + ((uint32_t)(qd * 32) << 16);
      tmem_st32(ta, hi);
      tmem_st32(ta + 32, lo);
      tmem_st_wait();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(split + s);
    }
  } else if (warp < 8) {
    // B split: lo = x - tf32(x), elementwise on the swizzled tile
    const int tid = threadIdx.x - 128;
    for (int it = 0, ai = 0; it < n_stages; ++it) {
      if (!stage_active(it)) continue;
      const int s = ai % kGramStages; const uint32_t ph = (ai / kGramStages) & 1; ++ai;
      mbar_wait(full + s, ph);
      const uint32_t hi = smem_u32(smem + s * L::kStage) + L::kA;
      split_stage(hi, hi + L::kB, L::kB, tid, 128, false);
      fence_proxy_async();
      __syncwarp();
      if (lane == 0) mbar_arrive(split + s);
    }
  } else if (warp < 12) {
    const int q = warp & 3;
    const int i = ib * 128 + q * 32 + lane;
    float sum[NC];
#pragma unroll
    for (int c = 0; c < NC; ++c) sum[c] = 0.f;
    for (int g = 0; g < n_groups; ++g) {
      const int acc = g & 1; const uint32_t aph = (g >> 1) & 1;
      mbar_wait(tfull + acc, aph);
      tc_fence_after();
      const uint32_t tacc = tmem_base + acc * NC + ((uint32_t)(q * 32) << 16);
#pragma unroll
      for (int c0 = 0; c0 < NC; c0 += 16) {
        uint32_t v[16];
        tmem_ld16(tacc + c0, v);
        tmem_ld_wait();
#pragma unroll
        for (int c = 0; c < 16; ++c) sum[c0 + c] += __uint_as_float(v[c]);
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(tempty + acc);
    }
    float* out = p.Mpart + (((int64_t)fold * p.splits + sp) * (p.n_blocks * 128) + i) * p.ldm + cb * 128;
#pragma unroll
    for (int c = 0; c < NC; ++c) out[c] = sum[c];
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 13) tmem_dealloc(tmem_base, kTmemCols);
}

// First-order moments and targets, per fold, HBM-bound: for a chunk of rows and one fold,
// part[j][0] = sum x_j, part[j][1+t] = sum x_j y_t (j < d); row d: [count, sum y_t ...]; row d+1: [0, sum y_t^2 ...].
// Y: regression target, or +-1 one-vs-rest columns of the class index (T = 1: class 1 positive).
struct XtyParams {
  const float* X; int64_t ldx; int64_t n; int d; int T;
  const int8_t* label; const uint16_t* blockmask;
  const float* yreg; const int32_t* ycls;
  int64_t rows_per_chunk; int chunks;
  float* part;   // [folds][chunks][d + 2][1 + T]
};
// one feature per thread and row: the general form (many targets)
template <int JMAX, int TMAX>   // features per thread (d <= 256 * JMAX), targets (T <= TMAX)
__global__ void __launch_bounds__(256) xty_scalar_kernel(const XtyParams p) {
  const int fold = blockIdx.y, ch = blockIdx.x;
  const int64_t r0 = (int64_t)ch * p.rows_per_chunk, r1 = min(p.n, r0 + p.rows_per_chunk);
  const int T = p.T, W = 1 + T;
  float ax[JMAX], axy[JMAX][TMAX];
#pragma unroll
  for (int a = 0; a < JMAX; ++a) { ax[a] = 0.f;
#pragma unroll
    for (int t = 0; t < TMAX; ++t) axy[a][t] = 0.f; }
  float cnt = 0.f, sy[TMAX], syy[TMAX];
#pragma unroll
  for (int t = 0; t < TMAX; ++t) { sy[t] = 0.f; syy[t] = 0.f; }
  constexpr int U = 8;                                  // rows in flight per thread (the pass is latency-bound otherwise)
  for (int64_t r = r0; r < r1; r += U) {                // r0 and the chunk length are multiples of 16
    if (!((p.blockmask[r >> 4] >> fold) & 1u)) continue;
    bool in[U]; float xv[U][JMAX];
#pragma unroll
    for (int u = 0; u < U; ++u) in[u] = (r + u < r1) && (p.label[r + u] == fold);
#pragma unroll
    for (int u = 0; u < U; ++u)
#pragma unroll
      for (int a = 0; a < JMAX; ++a) {
        const int j = threadIdx.x + a * 256;
        xv[u][a] = (in[u] && j < p.d) ? __ldg(p.X + (r + u) * p.ldx + j) : 0.f;
      }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      if (!in[u]) continue;
      float yv[TMAX];
#pragma unroll
      for (int t = 0; t < TMAX; ++t) {
        if (t < T) {
          if (p.yreg) yv[t] = p.yreg[r + u];
          else { const int c = p.ycls[r + u]; yv[t] = (T == 1 ? (c == 1) : (c == t)) ? 1.f : -1.f; }
        } else yv[t] = 0.f;
      }
#pragma unroll
      for (int a = 0; a < JMAX; ++a) {
        ax[a] += xv[u][a];
#pragma unroll
        for (int t = 0; t < TMAX; ++t) if (t < T) axy[a][t] = fmaf(xv[u][a], yv[t], axy[a][t]);
      }
      if (threadIdx.x == 0) {
        cnt += 1.f;
#pragma unroll
        for (int t = 0; t < TMAX; ++t) if (t < T) { sy[t] += yv[t]; syy[t] = fmaf(yv[t], yv[t], syy[t]); }
      }
    }
  }
  float* out = p.part + ((int64_t)fold * p.chunks + ch) * (p.d + 2) * W;
#pragma unroll
  for (int a = 0; a < JMAX; ++a) {
    const int j = threadIdx.x + a * 256;
    if (j < p.d) {
      out[(int64_t)j * W] = ax[a];
#pragma unroll
      for (int t = 0; t < TMAX; ++t) if (t < T) out[(int64_t)j * W + 1 + t] = axy[a][t];
    }
  }
  if (threadIdx.x == 0) {
    out[(int64_t)p.d * W] = cnt; out[(int64_t)(p.d + 1) * W] = 0.f;
#pragma unroll
    for (int t = 0; t < TMAX; ++t) if (t < T) { out[(int64_t)p.d * W + 1 + t] = sy[t]; out[(int64_t)(p.d + 1) * W + 1 + t] = syy[t]; }
  }
}

// Thread = four consecutive features (one 16-byte load per row; the tcgen05 path guarantees 16-byte aligned rows)
// times JV groups 512 features apart, and one of two row phases (even / odd rows of the chunk); the two phases are
// added through shared memory at the end.  The scalar one-feature-per-thread form this replaces was bound by
// instruction issue (ncu, config 3: 74 % of the issue slots at 3.2 TB/s).
template <int JV, int TMAX>   // feature groups per thread (d <= 512 * JV), targets (T <= TMAX)
__global__ void __launch_bounds__(256) xty_kernel(const XtyParams p) {
  const int fold = blockIdx.y, ch = blockIdx.x;
  const int64_t r0 = (int64_t)ch * p.rows_per_chunk, r1 = min(p.n, r0 + p.rows_per_chunk);
  const int T = p.T, W = 1 + T;
  const int g = threadIdx.x & 127, ph = threadIdx.x >> 7;
  float ax[JV][4], axy[JV][4][TMAX];
#pragma unroll
  for (int a = 0; a < JV; ++a)
#pragma unroll
    for (int e = 0; e < 4; ++e) { ax[a][e] = 0.f;
#pragma unroll
      for (int t = 0; t < TMAX; ++t) axy[a][e][t] = 0.f; }
  float cnt = 0.f, sy[TMAX], syy[TMAX];
#pragma unroll
  for (int t = 0; t < TMAX; ++t) { sy[t] = 0.f; syy[t] = 0.f; }
  constexpr int U = 4;                                  // rows in flight per thread
  for (int64_t rb = r0; rb < r1; rb += 2 * U) {         // r0 and the chunk length are multiples of 16; 2 U = 8 rows share a mask word
    if (!((p.blockmask[rb >> 4] >> fold) & 1u)) continue;
    bool in[U]; float4 xv[U][JV];
#pragma unroll
    for (int u = 0; u < U; ++u) { const int64_t r = rb + 2 * u + ph; in[u] = (r < r1) && (p.label[r] == fold); }
#pragma unroll
    for (int u = 0; u < U; ++u)
#pragma unroll
      for (int a = 0; a < JV; ++a) {
        const int j = 4 * (g + 128 * a);
        const int64_t r = rb + 2 * u + ph;
        float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
        if (in[u] && j < p.d) {
          const float* src = p.X + r * p.ldx + j;
          if (j + 3 < p.d) v = __ldg(reinterpret_cast<const float4*>(src));
          else { v.x = __ldg(src); if (j + 1 < p.d) v.y = __ldg(src + 1); if (j + 2 < p.d) v.z = __ldg(src + 2); }
        }
        xv[u][a] = v;
      }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      if (!in[u]) continue;
      const int64_t r = rb + 2 * u + ph;
      float yv[TMAX];
#pragma unroll
      for (int t = 0; t < TMAX; ++t) {
        if (t < T) {
          if (p.yreg) yv[t] = p.yreg[r];
          else { const int c = p.ycls[r]; yv[t] = (T == 1 ? (c == 1) : (c == t)) ? 1.f : -1.f; }
        } else yv[t] = 0.f;
      }
#pragma unroll
      for (int a = 0; a < JV; ++a) {
        const float x4[4] = {xv[u][a].x, xv[u][a].y, xv[u][a].z, xv[u][a].w};
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          ax[a][e] += x4[e];
#pragma unroll
          for (int t = 0; t < TMAX; ++t) if (t < T) axy[a][e][t] = fmaf(x4[e], yv[t], axy[a][e][t]);
        }
      }
      if (g == 0) {
        cnt += 1.f;
#pragma unroll
        for (int t = 0; t < TMAX; ++t) if (t < T) { sy[t] += yv[t]; syy[t] = fmaf(yv[t], yv[t], syy[t]); }
      }
    }
  }
  // odd-row phase -> shared memory -> even-row phase adds and writes
  __shared__ float sh[128 * JV * 4 * (1 + TMAX) + 1 + 2 * TMAX];
  float* shx = sh;
  if (ph == 1) {
#pragma unroll
    for (int a = 0; a < JV; ++a)
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        float* q = shx + (((g * JV + a) * 4 + e) * (1 + TMAX));
        q[0] = ax[a][e];
#pragma unroll
        for (int t = 0; t < TMAX; ++t) q[1 + t] = axy[a][e][t];
      }
    if (g == 0) {
      float* q = sh + 128 * JV * 4 * (1 + TMAX);
      q[0] = cnt;
#pragma unroll
      for (int t = 0; t < TMAX; ++t) { q[1 + t] = sy[t]; q[1 + TMAX + t] = syy[t]; }
    }
  }
  __syncthreads();
  if (ph == 1) return;
  float* out = p.part + ((int64_t)fold * p.chunks + ch) * (p.d + 2) * W;
#pragma unroll
  for (int a = 0; a < JV; ++a)
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      const int j = 4 * (g + 128 * a) + e;
      if (j < p.d) {
        const float* q = shx + (((g * JV + a) * 4 + e) * (1 + TMAX));
        out[(int64_t)j * W] = ax[a][e] + q[0];
#pragma unroll
        for (int t = 0; t < TMAX; ++t) if (t < T) out[(int64_t)j * W + 1 + t] = axy[a][e][t] + q[1 + t];
      }
    }
  if (g == 0) {
    const float* q = sh + 128 * JV * 4 * (1 + TMAX);
    out[(int64_t)p.d * W] = cnt + q[0]; out[(int64_t)(p.d + 1) * W] = 0.f;
#pragma unroll
    for (int t = 0; t < TMAX; ++t) if (t < T) {
      out[(int64_t)p.d * W + 1 + t] = sy[t] + q[1 + t]; out[(int64_t)(p.d + 1) * W + 1 + t] = syy[t] + q[1 + TMAX + t];
    }
  }
}

// Partials -> Mfold[fold][D][D] (fp64, D = d + T + 1), fixed summation order.
//   (i, c) with i, c < d   : Gram blocks, mirrored from the computed upper block triangle
//   (i, d + t), (d + t, i) : X^T Y ; (i, D-1), (D-1, i): column sums; (d+t, d+t): sum y^2; (d+t, D-1): sum y; (D-1, D-1): count
struct GramReduceParams {
  const float* Mpart; int splits; int n_blocks; int ldm;
  const float* xty; int chunks; int d; int T;
  double* Mfold;
};
__global__ void gram_reduce_kernel(const GramReduceParams p) {
  const int fold = blockIdx.z, i = blockIdx.y;
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  const int D = p.d + p.T + 1, W = 1 + p.T;
  if (c >= D || i >= D) return;
  double acc = 0.0;
  if (i < p.d && c < p.d) {
    const bool upper = (c >> 7) >= (i >> 7);
    const int64_t rr = upper ? i : c, cc = upper ? c : i;
    const int64_t rows = (int64_t)p.n_blocks * 128;
    for (int s = 0; s < p.splits; ++s) acc += (double)p.Mpart[(((int64_t)fold * p.splits + s) * rows + rr) * p.ldm + cc];
  } else {
    // which entry of the xty partial record?
    int j, k;                                   // record row j (feature, or d = counts/sums, d+1 = squares), column k
    const int hi = max(i, c), lo = min(i, c);
    if (lo < p.d) { j = lo; k = (hi == D - 1) ? 0 : 1 + (hi - p.d); }          // X^T 1 or X^T Y
    else if (hi == D - 1) { j = p.d; k = (lo == D - 1) ? 0 : 1 + (lo - p.d); }  // count or sum y_t
    else if (hi == lo) { j = p.d + 1; k = 1 + (lo - p.d); }                     // sum y_t^2
    else { j = -1; k = 0; }                                                      // y_t y_t' cross terms: not used
    if (j >= 0)
      for (int s = 0; s < p.chunks; ++s) acc += (double)p.xty[(((int64_t)fold * p.chunks + s) * (p.d + 2) + j) * W + k];
  }
  p.Mfold[((int64_t)fold * D + i) * D + c] = acc;
}

// ---------------------------------------------------------------- small helpers
// trial weights -> hi / lo copies for the forward B operand
__global__ void split_weights_kernel(const float* W, float* Whi, float* Wlo, int64_t n) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) { const float x = W[i], h = tf32_hi(x); Whi[i] = h; Wlo[i] = x - h; }
}

// per-(CTA, warp) fp64 partials [n_parts][n_slots] -> Gred columns d (sum of residuals) and d+1 (loss)
struct TcReduceParams {
  const double* lossPart; const double* gbPart; int64_t n_parts; int n_slots, T, d;
  double* Gred; int ldgr; int accumulate;
  // the forward kernel sums (loss - ln 2): add ln 2 * (training rows of the column) back once
  const unsigned long long* hist_local; const int32_t* sel; const int32_t* val; int64_t n_local; double loss_shift;
  const int* n_list_dev;   // live columns = *n_list_dev * T when set (n_slots: upper bound and pitch)
  const double* gb2; int gb2_splits;   // sum r per (gradient split, column) from tc_grad2_kernel, or null (then gbPart)
  const double* sw_slot;               // class weights: weighted count of this rank's training rows per column, or null
};
__global__ void tc_reduce_tiles_kernel(const TcReduceParams p) {   // block (32 slots, 32 partial lanes)
  const int n_live = p.n_list_dev ? min(p.n_slots, *p.n_list_dev * p.T) : p.n_slots;
  __shared__ double sh[2][32][33];
  const int slot = blockIdx.x * 32 + threadIdx.x;
  double a = 0.0, b = 0.0;
  if (slot < n_live) {
    for (int64_t t = threadIdx.y; t < p.n_parts; t += 32) {
      a += p.lossPart[t * p.n_slots + slot];
      if (!p.gb2) b += p.gbPart[t * p.n_slots + slot];
    }
    if (p.gb2) for (int t = threadIdx.y; t < p.gb2_splits; t += 32) b += p.gb2[(int64_t)t * p.n_slots + slot];
  }
  sh[0][threadIdx.y][threadIdx.x] = a; sh[1][threadIdx.y][threadIdx.x] = b;
  __syncthreads();
  if (threadIdx.y == 0 && slot < n_live) {
    double l = 0.0, g = 0.0;
    for (int i = 0; i < 32; ++i) { l += sh[0][i][threadIdx.x]; g += sh[1][i][threadIdx.x]; }   // fixed order
    double* q = p.Gred + (int64_t)slot * p.ldgr;
    q[p.d] = p.accumulate ? q[p.d] + g : g;
    if (!p.accumulate && p.sel[slot] >= 0)     // (a padding column has mask -1 and no rows)
      l += p.loss_shift * (p.sw_slot ? p.sw_slot[slot]
                                     : (double)(p.n_local - (int64_t)p.hist_local[p.sel[slot] * 256 + (p.val[slot] & 0xff)]));
    if (slot % p.T == 0) q[p.d + 1] = p.accumulate ? q[p.d + 1] + l : l;
  }
}

}  // namespace tc
}  // namespace dml
