// Ridge / RidgeClassifier: batched normal-equation solves from per-fold moments.
//
// scikit-learn path being reproduced (reference call sites worker.py:315,326 with
// the registry extended to Ridge/RidgeClassifier):
//   centre X and y on the training fold        sklearn/linear_model/_base.py:189-199
//   A = Xc^T Xc + alpha I ; w = solve(A, Xc^T yc, assume_a="pos")   _ridge.py:215-234
//   intercept = ybar - xbar . w                 _base.py:318-335
// Instead of copying and centring each training fold, the centred system is formed
// algebraically from the uncentred moments (fp64):
//   Xc^T Xc = (M_all - M_fold)[x,x] - n_tr * mu mu^T,  mu = (s_all - s_fold) / n_tr.
#pragma once
#include "dml_common.cuh"
#include "dml_tc.cuh"      // mbarrier helpers

namespace dml {

// 1-D bulk copy global -> shared (TMA, no tensor map): `bytes` and both addresses multiples of 16
__device__ __forceinline__ void bulk_load_1d(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               ::"r"(tc::smem_u32(dst)), "l"(src), "r"(bytes), "r"(tc::smem_u32(bar)) : "memory");
}

// Mpart[fold][split][D][ldm] (fp32) -> Mfold[fold][D][D] (fp64), fixed order over splits
__global__ void reduce_moments_kernel(const float* Mpart, int splits, int D, int ldm, double* Mfold) {
  const int fold = blockIdx.z, r = blockIdx.y;
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= D) return;
  double acc = 0.0;
  for (int s = 0; s < splits; ++s)
    acc += (double)Mpart[(((int64_t)fold * splits + s) * D + r) * ldm + c];
  Mfold[((int64_t)fold * D + r) * D + c] = acc;
}

// Mtot = sum over folds
__global__ void total_moments_kernel(const double* Mfold, int n_folds, int D, double* Mtot) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= (int64_t)D * D) return;
  double acc = 0.0;
  for (int f = 0; f < n_folds; ++f) acc += Mfold[(int64_t)f * D * D + i];
  Mtot[i] = acc;
}

struct RidgeSysParams {
  int d, T, D;            // D = d + T + 1
  int n_folds, has_hold;  // systems per candidate = n_folds (+1: hold-out split, moments in slot n_folds)
  int cand0, n_cand_chunk;
  const double* Mfold;    // [n_folds (+1)][D][D]; entry n_folds = rows with hold_id == 1
  const double* Mtot;
  const double* alpha; const uint8_t* fit_icpt;  // per candidate
  double* A;              // [systems][d][d]
  double* B;              // [systems][T][d]  (rhs, overwritten by the solution)
  double* mu;             // [systems][d + T]
};

// system index = (cand - cand0) * fits_per_cand + k
__global__ void ridge_assemble_kernel(const RidgeSysParams p) {
  const int fits_per_cand = p.n_folds + p.has_hold;
  const int sys = blockIdx.y, cand = p.cand0 + sys / fits_per_cand, k = sys % fits_per_cand;
  const int d = p.d, T = p.T, D = p.D;
  const double* Mf = p.Mfold + (int64_t)k * D * D;
  const double n_tr = p.Mtot[(int64_t)(D - 1) * D + (D - 1)] - Mf[(int64_t)(D - 1) * D + (D - 1)];
  const bool ic = p.fit_icpt[cand] != 0;
  const double alpha = p.alpha[cand];
  double* A = p.A + (int64_t)sys * d * d;
  double* B = p.B + (int64_t)sys * T * d;
  double* mu = p.mu + (int64_t)sys * (d + T);
  const int64_t total = (int64_t)d * (d + T);
  for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += (int64_t)gridDim.x * blockDim.x) {
    const int i = (int)(e / (d + T)), c = (int)(e % (d + T));
    const double mi = (p.Mtot[(int64_t)i * D + (D - 1)] - Mf[(int64_t)i * D + (D - 1)]) / n_tr;
    const double mc = (p.Mtot[(int64_t)c * D + (D - 1)] - Mf[(int64_t)c * D + (D - 1)]) / n_tr;
    double v = p.Mtot[(int64_t)i * D + c] - Mf[(int64_t)i * D + c];
    if (ic) v -= n_tr * mi * mc;
    if (c < d) A[(int64_t)i * d + c] = v + (i == c ? alpha : 0.0);
    else B[(int64_t)(c - d) * d + i] = v;
    if (i == 0) mu[c] = ic ? mc : 0.0;
  }
}

// In-place lower Cholesky of A (row-major, only the lower triangle is read/written),
// then L L^T W = B for T right-hand sides.  One block per system.
// A pivot (squared diagonal entry of L before the square root) below kRidgePivotTol times the largest
// diagonal entry of A marks the system as numerically rank-deficient: the Gram matrix is accumulated at
// fp32 / 3xTF32 accuracy (about 1e-7 of its diagonal), so such a pivot is mostly noise and the solution
// is not the one LAPACK gives scikit-learn in float64 (which, for LinearRegression, is the minimum-norm
// lstsq solution, and for Ridge falls back to SVD on LinAlgError, _ridge.py:753-761).  Reported as
// info = -(j + 1); the worker re-runs such candidates on the scikit-learn executor.
constexpr double kRidgePivotTol = 1e-4;
__device__ __forceinline__ double max_diag(const double* A, int d, double* red) {
  double m = 0.0;
  for (int j = threadIdx.x; j < d; j += blockDim.x) m = fmax(m, fabs(A[(int64_t)j * d + j]));
  return block_reduce<true>(m, red);
}
__global__ void __launch_bounds__(512) ridge_cholesky_kernel(double* Aall, double* Ball, int d, int T, int* info) {
  extern __shared__ double colj[];  // [d]
  __shared__ double piv;
  __shared__ double red[33];
  double* A = Aall + (int64_t)blockIdx.x * d * d;
  double* B = Ball + (int64_t)blockIdx.x * T * d;
  const int nw = blockDim.x >> 5, w = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const double small = kRidgePivotTol * max_diag(A, d, red);
  int weak = 0;
  for (int j = 0; j < d; ++j) {
    if (threadIdx.x == 0) {
      const double a = A[(int64_t)j * d + j];
      if (!(a > 0.0)) atomicExch(info + blockIdx.x, j + 1);  // not positive definite
      else if (a < small && !weak) weak = j + 1;
      piv = sqrt(a > 0.0 ? a : 1.0);
    }
    __syncthreads();
    const double pj = piv;
    for (int i = j + threadIdx.x; i < d; i += blockDim.x) {
      const double v = (i == j) ? pj : A[(int64_t)i * d + j] / pj;
      A[(int64_t)i * d + j] = v;
      colj[i] = v;
    }
    __syncthreads();
    // trailing update, one warp per row: A[i][k] -= L[i][j] L[k][j], j < k <= i
    for (int i = j + 1 + w; i < d; i += nw) {
      const double lij = colj[i];
      double* Ai = A + (int64_t)i * d;
      for (int k = j + 1 + lane; k <= i; k += 32) Ai[k] -= lij * colj[k];
    }
    __syncthreads();
  }
  if (threadIdx.x == 0 && weak && info[blockIdx.x] == 0) info[blockIdx.x] = -weak;
  // forward then backward substitution, one right-hand side at a time
  for (int t = 0; t < T; ++t) {
    double* b = B + (int64_t)t * d;
    for (int j = 0; j < d; ++j) {
      if (threadIdx.x == 0) { piv = b[j] / A[(int64_t)j * d + j]; b[j] = piv; }
      __syncthreads();
      const double zj = piv;
      for (int i = j + 1 + threadIdx.x; i < d; i += blockDim.x) b[i] -= A[(int64_t)i * d + j] * zj;
      __syncthreads();
    }
    for (int j = d - 1; j >= 0; --j) {
      if (threadIdx.x == 0) { piv = b[j] / A[(int64_t)j * d + j]; b[j] = piv; }
      __syncthreads();
      const double wj = piv;
      for (int i = threadIdx.x; i < j; i += blockDim.x) b[i] -= A[(int64_t)j * d + i] * wj;
      __syncthreads();
    }
  }
}

// Blocked (left-looking) variant for d <= kCholMaxD: a 32-column panel lives in shared memory, is
// updated with all previously factored columns (each thread owns one row and streams it from L2),
// factored, and written back once.  Same result as the column-by-column kernel above up to fp64
// rounding; ~100x less memory traffic (the rank-1 version rewrites the trailing matrix d times).
constexpr int kCholNB = 32, kCholMaxD = 512;
// `ATall` receives L^T (row j of AT = column j of L): with one thread per matrix row, the panel update
// and the forward substitution walk L by columns, which is contiguous (coalesced) only in L^T.
// With `f.Mtot` set the kernel forms the centred normal equations itself while it reads them (the entries of A and
// of the right-hand sides are each read once), straight from the fold moments that all systems of a fold share and
// that stay in L2, instead of reading a matrix ridge_assemble_kernel wrote to HBM for every system: same numbers
// (same expression, fp64), one kernel and 3 MB of traffic per system less.
struct RidgeFused {
  const double* Mfold; const double* Mtot; const double* alpha; const uint8_t* fit_icpt;
  int D, n_folds, has_hold, cand0;
  double* mu;             // [systems][d + T], written here (ridge_output_kernel reads it)
};
__global__ void __launch_bounds__(512) ridge_cholesky_blocked_kernel(double* Aall, double* ATall, double* Ball, int d, int T, int* info,
                                                                     const RidgeFused f, const int bulk_smem, long long* prof) {
  // optional phase profile of block 0 (compile with -DDML_CHOL_PROFILE and run with DML_CHOL_PROF=1): clocks spent in
  // 0 panel update, 1 panel formation, 2 diagonal block, 3 (unused), 4 rows below the block + write-back, 5 solves
#ifdef DML_CHOL_PROFILE
  long long pc[6] = {0, 0, 0, 0, 0, 0};
  long long tprev = clock64();
#define DML_PROF(i) do { if (prof) { const long long now_ = clock64(); pc[i] += now_ - tprev; tprev = now_; } } while (0)
#else
#define DML_PROF(i) do { } while (0)
#endif
  extern __shared__ double sm[];
  constexpr int NB = kCholNB, PP = NB + 1;
  double* P = sm;                     // [d][PP] panel rows k0..d-1
  double* Lk = sm + (((size_t)d * PP + 1) & ~(size_t)1);   // [NB][NB+2] staged rows of L (16-byte aligned): Lk[jj][c] = L[k0 + c][j0 + jj]
  double* mus = Lk + NB * (NB + 2);   // [d + T] training-fold means (fused form)
  constexpr int kRing = 4, kRows = 4;   // bulk-copy ring of the panel update: kRing stages of kRows rows of L^T (d doubles each)
  double* stage = mus + (((size_t)d + T + 1) & ~(size_t)1);   // [kRing][kRows][d], 16-byte aligned
  __shared__ __align__(8) uint64_t full[kRing];
  uint32_t issued = 0, consumed = 0;    // chunks issued (thread 0) / consumed (all threads) since the kernel began
  // bulk copies need 16-byte aligned rows: even d (k0 and nr are then even too).  `bulk_smem` says the launch
  // reserved the ring.
  const bool bulk = bulk_smem && (d % 2 == 0);
  __shared__ int bad, weak;
  __shared__ double red[33];
  __shared__ double invD[kCholNB];
  __shared__ double colj[kCholNB];
  double* A = Aall + (int64_t)blockIdx.x * d * d;
  double* AT = ATall + (int64_t)blockIdx.x * d * d;
  double* B = Ball + (int64_t)blockIdx.x * T * d;
  const int t = threadIdx.x;
  if (t == 0) {
    bad = 0; weak = 0;
    for (int i = 0; i < kRing; ++i) tc::mbar_init(full + i, 1);
    tc::fence_barrier_init();
  }
  __syncthreads();
  const bool fused = f.Mtot != nullptr;
  const int D = f.D;
  const double* Mf = nullptr; double n_tr = 0.0, alpha = 0.0; bool ic = false;
  if (fused) {
    const int fpc = f.n_folds + f.has_hold, cand = f.cand0 + blockIdx.x / fpc, k = blockIdx.x % fpc;
    Mf = f.Mfold + (int64_t)k * D * D;
    n_tr = f.Mtot[(int64_t)(D - 1) * D + (D - 1)] - Mf[(int64_t)(D - 1) * D + (D - 1)];
    alpha = f.alpha[cand]; ic = f.fit_icpt[cand] != 0;
    for (int c = t; c < d + T; c += blockDim.x) {
      const double mc = (f.Mtot[(int64_t)c * D + (D - 1)] - Mf[(int64_t)c * D + (D - 1)]) / n_tr;
      mus[c] = mc;
      f.mu[(int64_t)blockIdx.x * (d + T) + c] = ic ? mc : 0.0;
    }
    __syncthreads();
  }
  // entry (i, c) of [A | B]: c < d is A, c = d + target is a right-hand side (ridge_assemble_kernel's expression)
  auto entry = [&](int i, int c) -> double {
    double v = f.Mtot[(int64_t)i * D + c] - Mf[(int64_t)i * D + c];
    if (ic) v -= n_tr * mus[i] * mus[c];
    return (i == c) ? v + alpha : v;
  };
  double md = 0.0;
  if (fused) {
    for (int j = t; j < d; j += blockDim.x) md = fmax(md, fabs(entry(j, j)));
    md = block_reduce<true>(md, red);
  } else md = max_diag(A, d, red);
  const double small = kRidgePivotTol * md;
  for (int k0 = 0; k0 < d; k0 += NB) {
    const int nc = min(NB, d - k0), nr = d - k0;
    const int r = k0 + t;                              // this thread's row (valid when t < nr)
    // Panel update P = A[k0:, k0:k0+nc] - L[k0:, :k0] L[k0:k0+nc, :k0]^T with a 4-row x 8-column register tile per
    // thread (row group t / 4, column group t % 4).  One row per thread with all 32 columns re-read the whole
    // staged 32 x 32 block of L from shared memory for every row: 16 LDS.128 per 32 DFMA, which bound the kernel
    // by shared-memory return bandwidth at a tenth of the fp64 rate (ncu, config 3); the tile needs 4.
    const int rg = t >> 2, cg = t & 3;
    const int r0 = k0 + 4 * rg;                        // first of this thread's four rows
    const bool tile_on = 4 * rg < nr;
    double acc[4][8];
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
      for (int c = 0; c < 8; ++c) acc[i][c] = 0.0;
    // rows past the matrix read row d - 1 again (their results are dropped)
    const int ri0 = min(r0, d - 1), ri1 = min(r0 + 1, d - 1), ri2 = min(r0 + 2, d - 1), ri3 = min(r0 + 3, d - 1);
    // the 32 x 32 block of L each step needs is fetched one step ahead (registers), so that its L2 / DRAM latency
    // runs under the previous step's FMAs instead of between two barriers (ncu: a fifth of the kernel sat there)
    constexpr int kStage = 4;                          // elements per thread: covers NB * NB with >= 256 threads
    double nxt[kStage];
    auto fetch = [&](int j0) {
#pragma unroll
      for (int u = 0; u < kStage; ++u) {
        const int e = t + u * blockDim.x, c = e / NB, jj = e % NB;      // coalesced along jj
        nxt[u] = (e < NB * NB && c < nc) ? A[(int64_t)(k0 + c) * d + j0 + jj] : 0.0;
      }
    };
    const bool ahead = blockDim.x * kStage >= NB * NB;       // small systems (< 256 threads) stage the block directly
    if (bulk) {
      // Rows j (< k0) of L^T, elements k0 .. d-1, stream through a ring of kRing stages of kRows rows each by bulk
      // async copies (one thread issues, an mbarrier counts the bytes): a stage's first nc entries of a row are the
      // block L[k0 + c][j] the tile multiplies with, its entries 4 rg .. 4 rg + 3 the thread's own rows.  The copies
      // of the next stages run under the FMAs of this one; nothing waits on L2 / DRAM between barriers.
      const int n_chunks = k0 / kRows;
      const uint32_t row_bytes = (uint32_t)nr * 8u;
      auto issue = [&](int c, uint32_t g) {
        const uint32_t st = g % kRing;
        tc::mbar_expect_tx(full + st, kRows * row_bytes);
#pragma unroll
        for (int q = 0; q < kRows; ++q)
          bulk_load_1d(stage + ((size_t)st * kRows + q) * d, AT + (int64_t)(c * kRows + q) * d + k0, row_bytes, full + st);
      };
      if (t == 0) for (int c = 0; c < min(kRing - 1, n_chunks); ++c) issue(c, issued++);
      for (int c = 0; c < n_chunks; ++c) {
        if (t == 0 && c + kRing - 1 < n_chunks) issue(c + kRing - 1, issued++);
        const uint32_t st = consumed % kRing;
        DML_PROF(0);
        tc::mbar_wait(full + st, (consumed / kRing) & 1u);
        ++consumed;
        DML_PROF(3);   // waiting for the stage's bytes
        if (tile_on) {
#pragma unroll
          for (int q = 0; q < kRows; ++q) {
            const double* row = stage + ((size_t)st * kRows + q) * d;
            const double2 a01 = *reinterpret_cast<const double2*>(row + 4 * rg), a23 = *reinterpret_cast<const double2*>(row + 4 * rg + 2);
            const double2* lk = reinterpret_cast<const double2*>(row + 8 * cg);
            const double2 l01 = lk[0], l23 = lk[1], l45 = lk[2], l67 = lk[3];
            const double l[8] = {l01.x, l01.y, l23.x, l23.y, l45.x, l45.y, l67.x, l67.y};
#pragma unroll
            for (int cc = 0; cc < 8; ++cc) {
              acc[0][cc] = fma(a01.x, l[cc], acc[0][cc]); acc[1][cc] = fma(a01.y, l[cc], acc[1][cc]);
              acc[2][cc] = fma(a23.x, l[cc], acc[2][cc]); acc[3][cc] = fma(a23.y, l[cc], acc[3][cc]);
            }
          }
        }
        __syncthreads();                             // the stage may be refilled
      }
    } else {
    if (ahead && k0 > 0) fetch(0);
    for (int j0 = 0; j0 < k0; j0 += NB) {
      __syncthreads();
      if (ahead) {
#pragma unroll
        for (int u = 0; u < kStage; ++u) {
          const int e = t + u * blockDim.x;
          if (e < NB * NB) Lk[(e % NB) * (NB + 2) + e / NB] = nxt[u];
        }
      } else {
        for (int e = t; e < NB * NB; e += blockDim.x) {
          const int c = e / NB, jj = e % NB;
          Lk[jj * (NB + 2) + c] = (c < nc) ? A[(int64_t)(k0 + c) * d + j0 + jj] : 0.0;
        }
      }
      __syncthreads();
      if (ahead && j0 + NB < k0) fetch(j0 + NB);
      if (tile_on) {
        const double* at = AT + (int64_t)j0 * d;       // L[r][j0 + jj] = AT[j0 + jj][r]: consecutive rows are contiguous
#pragma unroll 4
        for (int jj = 0; jj < NB; ++jj) {
          const double* col = at + (int64_t)jj * d;
          const double a0 = col[ri0], a1 = col[ri1], a2 = col[ri2], a3 = col[ri3];
          const double2* lk = reinterpret_cast<const double2*>(Lk + jj * (NB + 2) + 8 * cg);
          const double2 l01 = lk[0], l23 = lk[1], l45 = lk[2], l67 = lk[3];
          const double l[8] = {l01.x, l01.y, l23.x, l23.y, l45.x, l45.y, l67.x, l67.y};
#pragma unroll
          for (int c = 0; c < 8; ++c) {
            acc[0][c] = fma(a0, l[c], acc[0][c]); acc[1][c] = fma(a1, l[c], acc[1][c]);
            acc[2][c] = fma(a2, l[c], acc[2][c]); acc[3][c] = fma(a3, l[c], acc[3][c]);
          }
        }
      }
    }
    }
    __syncthreads();
    DML_PROF(0);   // update loop
    if (tile_on) {
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const int row = r0 + i;
        if (row < d) {
#pragma unroll
          for (int c = 0; c < 8; ++c) {
            const int cc = 8 * cg + c;
            if (cc < nc) P[(size_t)(row - k0) * PP + cc] = (fused ? entry(row, k0 + cc) : A[(int64_t)row * d + k0 + cc]) - acc[i][c];
          }
        }
      }
    }
    __syncthreads();
    DML_PROF(1);   // panel formed
    // Factor the panel.  (A) warp 0 factors the nc x nc diagonal block with one row per lane in registers and
    // lane-to-lane shuffles; (B) every other row is an independent forward substitution against that block,
    // done by the row's thread in registers.  The column-by-column form this replaces (two block barriers and a
    // serial shared-memory update loop per column) kept 14 of 16 warps at a barrier (ncu: 57 % of the samples).
    if (t < 32) {
      double a[NB];
#pragma unroll
      for (int c = 0; c < NB; ++c) a[c] = (t < nc && c <= t) ? P[(size_t)t * PP + c] : 0.0;
#pragma unroll
      for (int j = 0; j < NB; ++j) {
        const double ajj = __shfl_sync(0xffffffffu, a[j], j);
        if (j < nc) {
          if (t == 0) {
            if (!(ajj > 0.0)) bad = k0 + j + 1;
            else if (ajj < small && !weak) weak = k0 + j + 1;
          }
          // one reciprocal square root instead of a square root and a division on the critical path of the step
          const double rs = rsqrt(ajj > 0.0 ? ajj : 1.0);
          a[j] = (t == j) ? (ajj > 0.0 ? ajj * rs : 1.0) : (t > j ? a[j] * rs : 0.0);
          colj[t] = a[j];                                              // column j of L, for every lane to read
          __syncwarp();
#pragma unroll
          // no "c <= t" test: entries above the diagonal (c > t) only ever hold garbage nobody reads, and 31 predicates
          // per step made the compiler pack and unpack predicate masks (300 instructions per step, 2 000 clocks)
          for (int c = j + 1; c < NB; ++c) a[c] = fma(-a[j], colj[c], a[c]);
          __syncwarp();
        }
      }
      if (t < nc) {
#pragma unroll
        for (int c = 0; c < NB; ++c) if (c <= t) P[(size_t)t * PP + c] = a[c];
#pragma unroll
        for (int c = 0; c < NB; ++c) if (c == t) invD[t] = 1.0 / a[c];
      }
    }
    __syncthreads();
    DML_PROF(2);   // (A) diagonal block
    if (t >= nc && t < nr) {
      double x[NB];
#pragma unroll
      for (int c = 0; c < NB; ++c) x[c] = (c < nc) ? P[(size_t)t * PP + c] : 0.0;
#pragma unroll
      for (int j = 0; j < NB; ++j) {
        if (j < nc) {
          double sacc = x[j];
#pragma unroll
          for (int c = 0; c < j; ++c) sacc = fma(-x[c], P[(size_t)j * PP + c], sacc);   // L_d[j][c]: one address per warp
          x[j] = sacc * invD[j];
        }
      }
#pragma unroll
      for (int c = 0; c < NB; ++c) if (c < nc) P[(size_t)t * PP + c] = x[c];
    }
    __syncthreads();
    DML_PROF(3);   // (B) rows below the diagonal block
    if (t < nr) {
      for (int c = 0; c < nc; ++c)                     // L^T: thread = row, coalesced over threads
        AT[(int64_t)(k0 + c) * d + r] = (k0 + c <= r) ? P[(size_t)t * PP + c] : 0.0;
    }
    {                                                  // L: lane = column of the panel, a warp writes whole 256-byte rows
      const int lane = t & 31, w = t >> 5, nw = blockDim.x >> 5;
      if (lane < nc)
        for (int row = w; row < nr; row += nw)
          if (lane <= row) A[(int64_t)(k0 + row) * d + k0 + lane] = P[(size_t)row * PP + lane];
    }
    if (bulk) asm volatile("fence.proxy.async;" ::: "memory");   // these rows are read back by bulk async copies
    __syncthreads();
    DML_PROF(4);   // write-back of the panel
  }
  if (t == 0 && (bad || weak)) atomicExch(info + blockIdx.x, bad ? bad : -weak);
  // Blocked substitution, one right-hand side at a time: the vector lives in shared memory, a 32 x 32
  // diagonal block is solved by one warp with shuffles, and every thread then updates its own row with the
  // 32 new unknowns (coalesced reads of L^T going forward, of L going backward).  Two block barriers per 32
  // unknowns instead of two per unknown (the column-by-column version spent about half of the kernel here).
  __syncthreads();
  double* bs = sm;                      // [d]      (the panel storage is free now)
  double* Ld = sm + ((d + 1) & ~1);     // [32][33]
  const int lane = t & 31;
  for (int tt = 0; tt < T; ++tt) {
    double* b = B + (int64_t)tt * d;
    for (int i = t; i < d; i += blockDim.x) bs[i] = fused ? entry(i, d + tt) : b[i];
    for (int k0 = 0; k0 < d; k0 += NB) {            // forward: L z = b
      const int nc = min(NB, d - k0);
      __syncthreads();
      for (int e = t; e < NB * NB; e += blockDim.x) {
        const int i = e / NB, j = e % NB;
        Ld[i * 33 + j] = (i < nc && j <= i) ? A[(int64_t)(k0 + i) * d + k0 + j] : (i == j ? 1.0 : 0.0);
      }
      __syncthreads();
      if (t < 32) {
        double bi = lane < nc ? bs[k0 + lane] : 0.0;
        for (int j = 0; j < NB; ++j) {
          const double zj = __shfl_sync(0xffffffffu, bi, j) / Ld[j * 33 + j];
          if (lane == j) bi = zj; else if (lane > j) bi -= Ld[lane * 33 + j] * zj;
        }
        if (lane < nc) bs[k0 + lane] = bi;
      }
      __syncthreads();
      const int i = k0 + NB + t;                    // rows below the block: one per thread, strided if d > blockDim
      for (int r = i; r < d; r += blockDim.x) {
        double acc = 0.0;
#pragma unroll 8
        for (int j = 0; j < nc; ++j) acc = fma(AT[(int64_t)(k0 + j) * d + r], bs[k0 + j], acc);
        bs[r] -= acc;
      }
    }
    for (int k0 = ((d - 1) / NB) * NB; k0 >= 0; k0 -= NB) {   // backward: L^T w = z
      const int nc = min(NB, d - k0);
      __syncthreads();
      for (int e = t; e < NB * NB; e += blockDim.x) {
        const int i = e / NB, j = e % NB;
        Ld[i * 33 + j] = (i < nc && j <= i) ? A[(int64_t)(k0 + i) * d + k0 + j] : (i == j ? 1.0 : 0.0);
      }
      __syncthreads();
      if (t < 32) {
        double bi = lane < nc ? bs[k0 + lane] : 0.0;
        for (int j = NB - 1; j >= 0; --j) {
          const double wj = __shfl_sync(0xffffffffu, bi, j) / Ld[j * 33 + j];
          if (lane == j) bi = wj; else if (lane < j) bi -= Ld[j * 33 + lane] * wj;
        }
        if (lane < nc) bs[k0 + lane] = bi;
      }
      __syncthreads();
      for (int r = t; r < k0; r += blockDim.x) {     // rows above the block
        double acc = 0.0;
#pragma unroll 8
        for (int j = 0; j < nc; ++j) acc = fma(A[(int64_t)(k0 + j) * d + r], bs[k0 + j], acc);
        bs[r] -= acc;
      }
    }
    __syncthreads();
    for (int i = t; i < d; i += blockDim.x) b[i] = bs[i];
    __syncthreads();
  }
  DML_PROF(5);
#ifdef DML_CHOL_PROFILE
  if (prof && blockIdx.x == 0 && t == 0) for (int i = 0; i < 6; ++i) prof[i] = pc[i];
#endif
}

#undef DML_PROF
// solution -> fp32 coef/intercept outputs and the score-pass operands
struct RidgeOutParams {
  int d, T, n_folds, has_hold, cand0, ldw;
  const double* B; const double* mu;
  float* coef; float* icpt;        // [n_cand][fits_per_cand][T][d], [..][T]
  float* Wt; float* bt; int32_t* slot_sel; int32_t* slot_val;   // all fits
  int npad;                        // score-pass column of a fit: see score_slot (0 = fit-major)
};
__global__ void ridge_output_kernel(const RidgeOutParams p) {
  __shared__ double red[33];
  const int fits_per_cand = p.n_folds + p.has_hold;
  const int sys = blockIdx.x, t = blockIdx.y, k = sys % fits_per_cand;
  const int fit = p.cand0 * fits_per_cand + sys, slot = fit * p.T + t;
  const int ss = score_slot(fit, fits_per_cand, p.npad) * p.T + t;
  const double* w = p.B + ((int64_t)sys * p.T + t) * p.d;
  const double* mu = p.mu + (int64_t)sys * (p.d + p.T);
  double dot = 0.0;
  for (int j = threadIdx.x; j < p.d; j += blockDim.x) {
    const float v = (float)w[j];   // coef_ keeps X.dtype
    p.coef[(int64_t)slot * p.d + j] = v;
    p.Wt[(int64_t)ss * p.ldw + j] = v;
    dot += mu[j] * w[j];
  }
  dot = block_reduce<false>(dot, red);
  if (threadIdx.x == 0) {
    const float b = (float)(mu[p.d + t] - dot);
    p.icpt[slot] = b; p.bt[ss] = b;
    p.slot_sel[ss] = (k < p.n_folds) ? 0 : 1;
    p.slot_val[ss] = (k < p.n_folds) ? k : 1;
  }
}

// per-fit test-fold statistics out of the moments: n_test, sum y, sum y^2 (regression)
__global__ void ridge_teststats_kernel(const double* Mfold, int D, int d, int n_fits, int fits_per_cand, int n_folds,
                                       long long* n_test, double* sy, double* syy) {
  const int f = blockIdx.x * blockDim.x + threadIdx.x;
  if (f >= n_fits) return;
  const int k = f % fits_per_cand;
  const double* Mf = Mfold + (int64_t)k * D * D;
  n_test[f] = (long long)(Mf[(int64_t)(D - 1) * D + (D - 1)] + 0.5);
  if (sy) sy[f] = Mf[(int64_t)d * D + (D - 1)];
  if (syy) syy[f] = Mf[(int64_t)d * D + d];
}

}  // namespace dml
