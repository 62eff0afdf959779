// CUDA-core (FFMA) tiled contractions with fused sweep epilogues.
//
// These kernels are the shape-general path of the sweep: any d, any number of
// columns, fp32 operands, fp32 accumulation per 128-row tile / per row split,
// fp64 across tiles.  They implement, for every (candidate x fold) column at
// once, what the reference worker does one fit at a time through scikit-learn
// (aws-prod/worker/worker.py:315,322,326):
//   fwd   : Z = X W^T + b, then per-row loss / residual (sklearn
//           _linear_loss.py:216-222 + _loss.pyx.tp:686-751,1300-1335), or the
//           prediction + integer correct count / squared error
//           (_base.py:391,416-418; worker.py:322-323,336-338)
//   grad  : G = R^T X                      (_linear_loss.py:355-368)
//   gram  : per-fold X^T [X | Y | 1]       (_ridge.py:220-221, _base.py:193-199)
// Fold membership is a per-row int8 label compared with a per-column value;
// no row of X is ever copied.
#pragma once
#include "dml_common.cuh"

namespace dml {

constexpr int BM = 128, BN = 128, BK = 16, LDS = 132, NT = 256;
constexpr int ZP = 129;  // pitch of the epilogue tile (odd => conflict-free both ways)
constexpr int kGemmSmemBytes = 2 * 2 * BK * LDS * 4;                 // A,B double-buffered
constexpr int kEpiSmemBytes = BM * ZP * 4 + BM * 16 + 8192;          // Z tile + row meta + red
constexpr int kFwdSmemBytes = kEpiSmemBytes > kGemmSmemBytes ? kEpiSmemBytes : kGemmSmemBytes;

// ---- operand loaders ----------------------------------------------------------------
// K-contiguous source: elem(m, k) = base[m*ld + k]   (rows of X, rows of W^T)
struct LoadK {
  const float* base; int64_t ld; int mcount; int kcount; bool vec;
  __device__ __forceinline__ void load(float4 (&r)[2], int k0) const {
#pragma unroll
    for (int i = 0; i < 2; ++i) {
      const int idx = threadIdx.x + i * NT, m = idx >> 2, k = k0 + (idx & 3) * 4;
      float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
      if (m < mcount) {
        const float* p = base + (int64_t)m * ld + k;
        if (vec && k + 3 < kcount) {
          v = __ldg(reinterpret_cast<const float4*>(p));
        } else {
          if (k + 0 < kcount) v.x = __ldg(p + 0);
          if (k + 1 < kcount) v.y = __ldg(p + 1);
          if (k + 2 < kcount) v.z = __ldg(p + 2);
          if (k + 3 < kcount) v.w = __ldg(p + 3);
        }
      }
      r[i] = v;
    }
  }
  __device__ __forceinline__ void store(float* S, const float4 (&r)[2]) const {
#pragma unroll
    for (int i = 0; i < 2; ++i) {
      const int idx = threadIdx.x + i * NT, m = idx >> 2, kq = (idx & 3) * 4;
      S[(kq + 0) * LDS + m] = r[i].x;
      S[(kq + 1) * LDS + m] = r[i].y;
      S[(kq + 2) * LDS + m] = r[i].z;
      S[(kq + 3) * LDS + m] = r[i].w;
    }
  }
};

// MN-contiguous source: elem(m, k) = base[k*ld + m]  (R rows, X rows as X^T).
// masked   : row k contributes only when label[k] == want (fold membership)
// augmented: columns past the dX features are the targets and a column of ones,
//            i.e. the operand is [X | Y | 1] (ridge moment matrix)
struct LoadMN {
  const float* base; int64_t ld; int mcount; int kcount; bool vec;
  bool masked; const int8_t* label; int want;
  bool augmented; int m0; int dX; int T; const float* yreg; const int32_t* ycls;
  __device__ __forceinline__ float aug(int k, int m) const {  // m = absolute column >= dX
    const int t = m - dX;
    if (t == T) return 1.f;
    if (t > T) return 0.f;
    if (yreg) return __ldg(yreg + k);
    const int c = __ldg(ycls + k);
    return (T == 1 ? (c == 1) : (c == t)) ? 1.f : -1.f;
  }
  __device__ __forceinline__ void load(float4 (&r)[2], int k0) const {
#pragma unroll
    for (int i = 0; i < 2; ++i) {
      const int idx = threadIdx.x + i * NT, k = k0 + (idx >> 5), m = (idx & 31) * 4;
      float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
      if (k < kcount && !(masked && __ldg(label + k) != want)) {
        const float* p = base + (int64_t)k * ld + m;
        if (vec && m + 3 < mcount) {
          v = __ldg(reinterpret_cast<const float4*>(p));
        } else {
          float e[4];
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            e[j] = 0.f;
            if (m + j < mcount) e[j] = __ldg(p + j);
            else if (augmented) e[j] = aug(k, m0 + m + j);
          }
          v = make_float4(e[0], e[1], e[2], e[3]);
        }
      }
      r[i] = v;
    }
  }
  __device__ __forceinline__ void store(float* S, const float4 (&r)[2]) const {
#pragma unroll
    for (int i = 0; i < 2; ++i) {
      const int idx = threadIdx.x + i * NT;
      *reinterpret_cast<float4*>(S + (idx >> 5) * LDS + (idx & 31) * 4) = r[i];
    }
  }
};

__device__ __forceinline__ void mma_tile(float (&acc)[8][8], const float* As, const float* Bs) {
  const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
#pragma unroll
  for (int k = 0; k < BK; ++k) {
    const float4 a0 = *reinterpret_cast<const float4*>(As + k * LDS + ty * 4);
    const float4 a1 = *reinterpret_cast<const float4*>(As + k * LDS + 64 + ty * 4);
    const float4 b0 = *reinterpret_cast<const float4*>(Bs + k * LDS + tx * 4);
    const float4 b1 = *reinterpret_cast<const float4*>(Bs + k * LDS + 64 + tx * 4);
    const float a[8] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w};
    const float b[8] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w};
#pragma unroll
    for (int i = 0; i < 8; ++i)
#pragma unroll
      for (int j = 0; j < 8; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
  }
}

// acc += A(:, k range) B(k range, :) ; kskip (optional) = per-16-row bitmask, bit `want`
template <class LA, class LB>
__device__ __forceinline__ void gemm_mainloop(float (&acc)[8][8], const LA& la, const LB& lb,
                                              int kBegin, int kEnd, float* smem,
                                              const uint16_t* kskip, int want) {
  float* As[2] = {smem, smem + BK * LDS};
  float* Bs[2] = {smem + 2 * BK * LDS, smem + 3 * BK * LDS};
  float4 ra[2], rb[2];
  int k0 = kBegin;
  if (kskip) while (k0 < kEnd && !((kskip[k0 >> 4] >> want) & 1)) k0 += BK;
  if (k0 >= kEnd) return;
  la.load(ra, k0); lb.load(rb, k0);
  la.store(As[0], ra); lb.store(Bs[0], rb);
  __syncthreads();
  int buf = 0;
  while (true) {
    int k1 = k0 + BK;
    if (kskip) while (k1 < kEnd && !((kskip[k1 >> 4] >> want) & 1)) k1 += BK;
    const bool more = k1 < kEnd;
    if (more) { la.load(ra, k1); lb.load(rb, k1); }
    mma_tile(acc, As[buf], Bs[buf]);
    if (!more) break;
    la.store(As[buf ^ 1], ra); lb.store(Bs[buf ^ 1], rb);
    __syncthreads();
    buf ^= 1; k0 = k1;
  }
}

// row / col of accumulator element (i, j) inside the 128x128 tile
__device__ __forceinline__ int acc_row(int i) { return (i < 4 ? 0 : 64) + (threadIdx.x >> 4) * 4 + (i & 3); }
__device__ __forceinline__ int acc_col(int j) { return (j < 4 ? 0 : 64) + (threadIdx.x & 15) * 4 + (j & 3); }

// ---- forward kernel -----------------------------------------------------------------
enum FwdMode { FWD_LOGREG = 0, FWD_SCORE_CLS = 1, FWD_SCORE_REG = 2 };

struct FwdParams {
  const float* X; int64_t ldx; int64_t row0; int64_t rows; int d; bool xvec;
  const float* Wt; int ldw; bool wvec; const float* bt;
  int n_slots; int T; int tile_n;
  const int32_t* ycls; const float* yreg;
  const int8_t* fold_id; const int8_t* hold_id;   // row labels (hold_id may be null)
  const int32_t* slot_sel; const int32_t* slot_val; // per slot: which label array, which value is held out
  float* R; int ldr;                                // [rows][ldr], chunk local (FWD_LOGREG)
  double* lossPart;                                 // [row tiles][n_slots / T]: fp64, see DESIGN.md "precision"
  int32_t* cntPart; float* ssePart;                 // [row tiles][n_slots] (score modes)
  // class weights (FWD_LOGREG; null = unweighted): cw[slot][2]; binary: the weights of class 0 and class 1 for
  // the column's fit; multinomial: cw[slot][0] = weight of class (slot % T) of the fit that owns the column.
  // A row's loss and residual are multiplied by the weight of its class (sklearn _logistic.py:
  // sample_weight = class_weight_[y], _linear_loss.py:291-379).
  const float* cw;
};

__device__ __forceinline__ float softplus_f(float z) {  // log(1 + e^z), stable
  return fmaxf(z, 0.f) + log1pf(expf(-fabsf(z)));
}
__device__ __forceinline__ float sigmoid_f(float z) {
  const float e = expf(-fabsf(z));
  return z >= 0.f ? 1.f / (1.f + e) : e / (1.f + e);
}

template <int MODE>
__global__ void __launch_bounds__(NT) fwd_kernel(const FwdParams p) {
  extern __shared__ __align__(16) float smem[];
  const int64_t tile_row0 = p.row0 + (int64_t)blockIdx.x * BM;
  const int rows_here = (int)min((int64_t)BM, p.row0 + p.rows - tile_row0);
  const int slot0 = blockIdx.y * p.tile_n;
  const int slots_here = min(p.tile_n, p.n_slots - slot0);
  const int64_t tile_id = tile_row0 / BM;

  float acc[8][8];
#pragma unroll
  for (int i = 0; i < 8; ++i)
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[i][j] = 0.f;
  LoadK la{p.X + tile_row0 * p.ldx, p.ldx, rows_here, p.d, p.xvec};
  LoadK lb{p.Wt + (int64_t)slot0 * p.ldw, p.ldw, slots_here, p.d, p.wvec};
  gemm_mainloop(acc, la, lb, 0, p.d, smem, nullptr, 0);
  __syncthreads();

  // ---- epilogue: Z tile -> smem, row metadata -> smem ----
  float* Zs = smem;                                        // [BM][ZP]
  int32_t* rowY = reinterpret_cast<int32_t*>(smem + BM * ZP);
  float* rowYf = reinterpret_cast<float*>(rowY + BM);
  int32_t* rowLab = rowY + 2 * BM;                         // fold | hold<<8
  double* redd = reinterpret_cast<double*>(rowY + 4 * BM); // [512] doubles
  int* redi = reinterpret_cast<int*>(redd + 512);          // [256] ints
#pragma unroll
  for (int i = 0; i < 8; ++i)
#pragma unroll
    for (int j = 0; j < 8; ++j) Zs[acc_row(i) * ZP + acc_col(j)] = acc[i][j];
  if (threadIdx.x < BM) {
    const int r = threadIdx.x;
    int yv = 0, lab = 0xffff; float yf = 0.f;
    if (r < rows_here) {
      const int64_t g = tile_row0 + r;
      if (p.ycls) yv = p.ycls[g];
      if (p.yreg) yf = p.yreg[g];
      lab = (int)(uint8_t)p.fold_id[g] | ((p.hold_id ? (int)(uint8_t)p.hold_id[g] : 0) << 8);
    }
    rowY[r] = yv; rowYf[r] = yf; rowLab[r] = lab;
  }
  __syncthreads();

  const int T = p.T;
  if (T == 1) {
    // thread -> (slot column, half of the rows); running sums stay in registers.  The loss is
    // summed in fp64: More'-Thuente interpolates on differences of f, and fp32 summation noise
    // (1e-8 relative) is enough to move the accepted step (DESIGN.md "precision").
    const int c = threadIdx.x & 127, half = threadIdx.x >> 7;
    double s0 = 0.0; int cnt = 0;
    if (c < slots_here) {
      const int sel = p.slot_sel[slot0 + c], val = p.slot_val[slot0 + c];
      const float b = p.bt[slot0 + c];
      for (int r = half * 64; r < half * 64 + 64; ++r) {
        const int lab = rowLab[r];
        if (lab == 0xffff) { if (MODE == FWD_LOGREG) Zs[r * ZP + c] = 0.f; continue; }
        const bool held = ((lab >> (8 * sel)) & 0xff) == val;
        const float z = Zs[r * ZP + c] + b;
        if (MODE == FWD_LOGREG) {
          float res = 0.f;
          if (!held) {
            const float yv = (float)rowY[r];
            const float sw = p.cw ? p.cw[(int64_t)(slot0 + c) * 2 + rowY[r]] : 1.f;
            s0 += (double)(sw * (softplus_f(z) - yv * z));
            res = sw * (sigmoid_f(z) - yv);
          }
          Zs[r * ZP + c] = res;
        } else if (MODE == FWD_SCORE_CLS) {
          if (held) cnt += ((z > 0.f) ? 1 : 0) == rowY[r];
        } else {
          if (held) { const float e = rowYf[r] - z; s0 += (double)(e * e); }
        }
      }
    } else if (MODE == FWD_LOGREG) {
      for (int r = half * 64; r < half * 64 + 64; ++r) Zs[r * ZP + c] = 0.f;
    }
    redd[threadIdx.x] = s0;
    redi[threadIdx.x] = cnt;
    __syncthreads();
    if (threadIdx.x < slots_here) {
      const int64_t o = tile_id * p.n_slots + slot0 + threadIdx.x;
      if (MODE == FWD_LOGREG) p.lossPart[o] = redd[threadIdx.x] + redd[128 + threadIdx.x];
      else if (MODE == FWD_SCORE_CLS) p.cntPart[o] = redi[threadIdx.x] + redi[128 + threadIdx.x];
      else p.ssePart[o] = (float)(redd[threadIdx.x] + redd[128 + threadIdx.x]);
    }
  } else {
    // thread -> (row, group of T consecutive slots = one fit); a warp covers 32 rows of one group
    const int groups = slots_here / T;
    for (int it = threadIdx.x; it < groups * BM; it += NT) {
      const int r = it & 127, g = it >> 7, c0 = g * T;
      const int lab = rowLab[r];
      const int sel = p.slot_sel[slot0 + c0], val = p.slot_val[slot0 + c0];
      const bool valid = lab != 0xffff;
      const bool held = valid && (((lab >> (8 * sel)) & 0xff) == val);
      float* zr = Zs + r * ZP + c0;
      float loss = 0.f; int ok = 0;
      if (MODE == FWD_LOGREG) {
        if (valid && !held) {
          float m = -INFINITY;
          for (int c = 0; c < T; ++c) { const float z = zr[c] + p.bt[slot0 + c0 + c]; zr[c] = z; m = fmaxf(m, z); }
          float s = 0.f;
          for (int c = 0; c < T; ++c) s += expf(zr[c] - m);
          const int yv = rowY[r];
          const float sw = p.cw ? p.cw[(int64_t)(slot0 + c0 + yv) * 2] : 1.f;
          loss = sw * (logf(s) + m - zr[yv]);
          const float inv = 1.f / s;
          for (int c = 0; c < T; ++c) zr[c] = sw * (expf(zr[c] - m) * inv - (c == yv ? 1.f : 0.f));
        } else {
          for (int c = 0; c < T; ++c) zr[c] = 0.f;
        }
      } else if (MODE == FWD_SCORE_CLS) {
        if (held) {
          float m = -INFINITY; int am = 0;
          for (int c = 0; c < T; ++c) { const float z = zr[c] + p.bt[slot0 + c0 + c]; if (z > m) { m = z; am = c; } }
          ok = (am == rowY[r]);
        }
      }
      // one (group, 32-row block) per warp iteration
      const int wslot = it >> 5;  // = g*4 + row block
      if (MODE == FWD_LOGREG) {
        const double l = warp_sum((double)loss);
        if ((threadIdx.x & 31) == 0) redd[wslot] = l;
      } else {
        const unsigned bal = __ballot_sync(0xffffffffu, ok);
        if ((threadIdx.x & 31) == 0) redi[wslot] = __popc(bal);
      }
    }
    __syncthreads();
    if (threadIdx.x < groups) {
      if (MODE == FWD_LOGREG) {
        const double* q = redd + threadIdx.x * 4;
        p.lossPart[tile_id * (p.n_slots / T) + slot0 / T + threadIdx.x] = (q[0] + q[1]) + (q[2] + q[3]);
      } else {
        const int* q = redi + threadIdx.x * 4;
        p.cntPart[tile_id * p.n_slots + slot0 + threadIdx.x * T] = q[0] + q[1] + q[2] + q[3];
      }
    }
  }
  if (MODE == FWD_LOGREG) {
    __syncthreads();
    // coalesced store of the residual tile; pad columns [n_slots, ldr) are written as zeros
    const int64_t lrow0 = tile_row0 - p.row0;
    const int ncols = min(BN, p.ldr - slot0);
    const bool last_tile = (slot0 + p.tile_n >= p.n_slots);
    const int wcols = last_tile ? ncols : slots_here;
    for (int it = threadIdx.x; it < rows_here * BN; it += NT) {
      const int r = it >> 7, c = it & 127;
      if (c < wcols) p.R[(lrow0 + r) * p.ldr + slot0 + c] = (c < slots_here) ? Zs[r * ZP + c] : 0.f;
    }
  }
}

// ---- gradient kernel: Gpart[split][slot][j] = sum_rows R[row][slot] [X | 1][row][j] -------
struct GradParams {
  const float* X; int64_t ldx; int64_t row0; int64_t rows; int d; bool xvec;
  const float* R; int ldr; int n_slots;
  int64_t rows_per_split;
  float* Gpart; int ldg;  // [splits][n_slots][ldg], ldg >= d + 1 (column d = sum of residuals)
};

__global__ void __launch_bounds__(NT) grad_kernel(const GradParams p) {
  extern __shared__ __align__(16) float smem[];
  const int slot0 = blockIdx.x * BM, j0 = blockIdx.y * BN, split = blockIdx.z;
  const int64_t kb = (int64_t)split * p.rows_per_split;
  const int64_t ke = min(p.rows, kb + p.rows_per_split);
  float acc[8][8];
#pragma unroll
  for (int i = 0; i < 8; ++i)
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[i][j] = 0.f;
  if (kb < ke) {
    LoadMN la{p.R + kb * p.ldr + slot0, p.ldr, min(BM, p.ldr - slot0), (int)(ke - kb), true,
              false, nullptr, 0, false, 0, 0, 0, nullptr, nullptr};
    // columns: the d features, then a column of ones (intercept gradient = sum of residuals)
    LoadMN lb{p.X + (p.row0 + kb) * p.ldx + j0, p.ldx, max(0, min(BN, p.d - j0)), (int)(ke - kb), p.xvec,
              false, nullptr, 0, true, j0, p.d, 0, nullptr, nullptr};
    gemm_mainloop(acc, la, lb, 0, (int)(ke - kb), smem, nullptr, 0);
  }
  float* out = p.Gpart + (int64_t)split * p.n_slots * p.ldg;
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const int s = slot0 + acc_row(i);
    if (s >= p.n_slots) continue;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const int c = j0 + acc_col(j);
      if (c <= p.d) out[(int64_t)s * p.ldg + c] = acc[i][j];
    }
  }
}

// ---- ridge moments: with Z = [X | Y | 1] (n x (d+T+1)),
//      Mpart[fold][split] = sum over the split's rows of that fold of z z^T.
// One pass over X gives X^T X, X^T Y, column sums, Y^T Y, sum Y and the row count of
// every fold; the training-fold moments are (sum over folds) - (own fold).
struct GramParams {
  const float* X; int64_t ldx; int64_t n; int d; bool xvec;
  const int8_t* label; const uint16_t* blockmask;  // per-16-row presence bits (may be null)
  int T; const float* yreg; const int32_t* ycls;
  int64_t rows_per_split; int splits;
  float* Mpart; int ldm;  // [folds][splits][d+T+1][ldm], ldm >= d + T + 1
};

__global__ void __launch_bounds__(NT) gram_kernel(const GramParams p) {
  extern __shared__ __align__(16) float smem[];
  const int i0 = blockIdx.x * BM, c0 = blockIdx.y * BN;
  const int fold = blockIdx.z / p.splits, split = blockIdx.z % p.splits;
  const int64_t kb = (int64_t)split * p.rows_per_split;
  const int64_t ke = min(p.n, kb + p.rows_per_split);
  const int ncols = p.d + p.T + 1;
  float acc[8][8];
#pragma unroll
  for (int i = 0; i < 8; ++i)
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[i][j] = 0.f;
  if (kb < ke) {
    const float* yr = p.yreg ? p.yreg + kb : nullptr;
    const int32_t* yc = p.ycls ? p.ycls + kb : nullptr;
    LoadMN la{p.X + kb * p.ldx + i0, p.ldx, max(0, min(BM, p.d - i0)), (int)(ke - kb), p.xvec,
              true, p.label + kb, fold, true, i0, p.d, p.T, yr, yc};
    LoadMN lb{p.X + kb * p.ldx + c0, p.ldx, max(0, min(BN, p.d - c0)), (int)(ke - kb), p.xvec,
              false, nullptr, 0, true, c0, p.d, p.T, yr, yc};
    gemm_mainloop(acc, la, lb, 0, (int)(ke - kb), smem, p.blockmask ? p.blockmask + (kb >> 4) : nullptr, fold);
  }
  float* out = p.Mpart + ((int64_t)fold * p.splits + split) * ncols * p.ldm;
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const int r = i0 + acc_row(i);
    if (r >= ncols) continue;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const int c = c0 + acc_col(j);
      if (c < ncols) out[(int64_t)r * p.ldm + c] = acc[i][j];
    }
  }
}

// per-16-row block: bit f set when some row of the block has label f (f < 16)
__global__ void blockmask_kernel(const int8_t* label, int64_t n, uint16_t* mask) {
  const int64_t b = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (b * 16 >= n) return;
  unsigned m = 0;
  for (int64_t r = b * 16; r < min(n, b * 16 + 16); ++r) m |= 1u << (label[r] & 15);
  mask[b] = (uint16_t)m;
}

}  // namespace dml
