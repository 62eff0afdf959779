// Batched L-BFGS-B (unbounded case) for every (candidate x fold) fit at once.
//
// One thread block owns one fit; every fit sits at its own point of its own
// More'-Thuente line search, so one batched loss/gradient evaluation ("round")
// serves all of them.  The rule set is the one scikit-learn's
// LogisticRegression(solver="lbfgs") reaches through scipy
// (sklearn/linear_model/_logistic.py:580-598 -> scipy/optimize/_lbfgsb_py.py:393-461,
// L-BFGS-B 3.0 mainlb/lnsrlb + MINPACK-2 dcsrch/dcstep); reference call sites
// aws-prod/worker/worker.py:315,326.  Scalars and optimiser state are fp64, the
// trial point handed to the contraction kernels is rounded to fp32 exactly where
// sklearn casts it (_linear_loss.py:216-217).
#pragma once
#include "dml_common.cuh"

namespace dml {

struct FitState {
  // constants of the fit
  double l2, tol, inv_ntr;
  int max_iter, fit_icpt, mask_sel, mask_val;
  // L-BFGS-B
  double f, f_old, gd, gd_old, stp, theta;
  int nit, nfev, col, head, status, active, phase, ifun;
  // dcsrch
  int brackt, stage;
  double finit, ginit, gtest, width, width1, stx, fx, gx, sty, fy, gy, stmin, stmax;
};

enum { TASK_FG = 0, TASK_CONV = 1, TASK_WARN = 2 };

__device__ inline void dcstep(double& stx, double& fx, double& dx, double& sty, double& fy, double& dy,
                              double& stp, double fp, double dp, int& brackt, double stpmin, double stpmax) {
  const double sgnd = dp * (dx / fabs(dx));
  double stpf;
  if (fp > fx) {
    const double theta = 3.0 * (fx - fp) / (stp - stx) + dx + dp;
    const double s = fmax(fabs(theta), fmax(fabs(dx), fabs(dp)));
    double gamma = s * sqrt((theta / s) * (theta / s) - (dx / s) * (dp / s));
    if (stp < stx) gamma = -gamma;
    const double p = (gamma - dx) + theta, q = ((gamma - dx) + gamma) + dp, r = p / q;
    const double stpc = stx + r * (stp - stx);
    const double stpq = stx + ((dx / ((fx - fp) / (stp - stx) + dx)) / 2.0) * (stp - stx);
    stpf = (fabs(stpc - stx) < fabs(stpq - stx)) ? stpc : stpc + (stpq - stpc) / 2.0;
    brackt = 1;
  } else if (sgnd < 0.0) {
    const double theta = 3.0 * (fx - fp) / (stp - stx) + dx + dp;
    const double s = fmax(fabs(theta), fmax(fabs(dx), fabs(dp)));
    double gamma = s * sqrt((theta / s) * (theta / s) - (dx / s) * (dp / s));
    if (stp > stx) gamma = -gamma;
    const double p = (gamma - dp) + theta, q = ((gamma - dp) + gamma) + dx, r = p / q;
    const double stpc = stp + r * (stx - stp);
    const double stpq = stp + (dp / (dp - dx)) * (stx - stp);
    stpf = (fabs(stpc - stp) > fabs(stpq - stp)) ? stpc : stpq;
    brackt = 1;
  } else if (fabs(dp) < fabs(dx)) {
    const double theta = 3.0 * (fx - fp) / (stp - stx) + dx + dp;
    const double s = fmax(fabs(theta), fmax(fabs(dx), fabs(dp)));
    double gamma = s * sqrt(fmax(0.0, (theta / s) * (theta / s) - (dx / s) * (dp / s)));
    if (stp > stx) gamma = -gamma;
    const double p = (gamma - dp) + theta, q = (gamma + (dx - dp)) + gamma, r = p / q;
    double stpc;
    if (r < 0.0 && gamma != 0.0) stpc = stp + r * (stx - stp);
    else if (stp > stx) stpc = stpmax;
    else stpc = stpmin;
    const double stpq = stp + (dp / (dp - dx)) * (stx - stp);
    if (brackt) {
      stpf = (fabs(stpc - stp) < fabs(stpq - stp)) ? stpc : stpq;
      if (stp > stx) stpf = fmin(stp + 0.66 * (sty - stp), stpf);
      else stpf = fmax(stp + 0.66 * (sty - stp), stpf);
    } else {
      stpf = (fabs(stpc - stp) > fabs(stpq - stp)) ? stpc : stpq;
      stpf = fmin(stpmax, stpf);
      stpf = fmax(stpmin, stpf);
    }
  } else {
    if (brackt) {
      const double theta = 3.0 * (fp - fy) / (sty - stp) + dy + dp;
      const double s = fmax(fabs(theta), fmax(fabs(dy), fabs(dp)));
      double gamma = s * sqrt((theta / s) * (theta / s) - (dy / s) * (dp / s));
      if (stp > sty) gamma = -gamma;
      const double p = (gamma - dp) + theta, q = ((gamma - dp) + gamma) + dy, r = p / q;
      stpf = stp + r * (sty - stp);
    } else if (stp > stx) stpf = stpmax;
    else stpf = stpmin;
  }
  if (fp > fx) { sty = stp; fy = fp; dy = dp; }
  else {
    if (sgnd < 0.0) { sty = stx; fy = fx; dy = dx; }
    stx = stp; fx = fp; dx = dp;
  }
  stp = stpf;
}

__device__ inline void dcsrch_start(FitState& s, double stp, double f, double g) {
  s.brackt = 0; s.stage = 1;
  s.finit = f; s.ginit = g; s.gtest = 1e-3 * g;
  s.width = kStpMax - 0.0; s.width1 = s.width / 0.5;
  s.stx = 0.0; s.fx = f; s.gx = g;
  s.sty = 0.0; s.fy = f; s.gy = g;
  s.stmin = 0.0; s.stmax = stp + 4.0 * stp;
}

// ftol 1e-3, gtol 0.9, xtol 0.1, stpmin 0, stpmax 1e10 (L-BFGS-B lnsrlb constants)
__device__ inline int dcsrch_step(FitState& s, double& stp, double f, double g) {
  const double stpmin = 0.0, stpmax = kStpMax;
  const double ftest = s.finit + stp * s.gtest;
  if (s.stage == 1 && f <= ftest && g >= 0.0) s.stage = 2;
  int task = TASK_FG;
  if (s.brackt && (stp <= s.stmin || stp >= s.stmax)) task = TASK_WARN;
  if (s.brackt && s.stmax - s.stmin <= 0.1 * s.stmax) task = TASK_WARN;
  if (stp == stpmax && f <= ftest && g <= s.gtest) task = TASK_WARN;
  if (stp == stpmin && (f > ftest || g >= s.gtest)) task = TASK_WARN;
  if (f <= ftest && fabs(g) <= 0.9 * (-s.ginit)) task = TASK_CONV;
  if (task != TASK_FG) return task;
  if (s.stage == 1 && f <= s.fx && f > ftest) {
    const double fm = f - stp * s.gtest;
    double fxm = s.fx - s.stx * s.gtest, fym = s.fy - s.sty * s.gtest;
    const double gm = g - s.gtest;
    double gxm = s.gx - s.gtest, gym = s.gy - s.gtest;
    dcstep(s.stx, fxm, gxm, s.sty, fym, gym, stp, fm, gm, s.brackt, s.stmin, s.stmax);
    s.fx = fxm + s.stx * s.gtest; s.fy = fym + s.sty * s.gtest;
    s.gx = gxm + s.gtest; s.gy = gym + s.gtest;
  } else {
    dcstep(s.stx, s.fx, s.gx, s.sty, s.fy, s.gy, stp, f, g, s.brackt, s.stmin, s.stmax);
  }
  if (s.brackt) {
    if (fabs(s.sty - s.stx) >= 0.66 * s.width1) stp = s.stx + 0.5 * (s.sty - s.stx);
    s.width1 = s.width;
    s.width = fabs(s.sty - s.stx);
  }
  if (s.brackt) { s.stmin = fmin(s.stx, s.sty); s.stmax = fmax(s.stx, s.sty); }
  else { s.stmin = stp + 1.1 * (stp - s.stx); s.stmax = stp + 4.0 * (stp - s.stx); }
  stp = fmin(fmax(stp, stpmin), stpmax);
  if ((s.brackt && (stp <= s.stmin || stp >= s.stmax)) ||
      (s.brackt && s.stmax - s.stmin <= 0.1 * s.stmax)) stp = s.stx;
  return TASK_FG;
}

struct LbfgsParams {
  int d, T, P;              // P = T*(d+1)
  int ldgr;                 // pitch of Gred rows: [slot][ldgr] doubles = d weights, then gb, then loss
  FitState* st;
  double *x, *g, *dir, *gt; // [F][P]
  double *S, *Y;            // [F][kMaxCor][P]
  double *ys;               // [F][kMaxCor]
  const int* active_fits;   // compaction used by the evaluation that just finished
  const int* n_active;
  const double* Gred;       // reduced sums for the current slots
};

// One block per active fit.  All threads take the same branches (scalars come from
// block reductions or from the state record), thread 0 writes the record back.
__global__ void __launch_bounds__(256) lbfgs_update_kernel(const LbfgsParams p) {
  __shared__ double red[33];
  const int a = blockIdx.x;
  if (a >= *p.n_active) return;
  const int fit = p.active_fits[a];
  if (fit < 0) return;                 // padding entry of a fold-major list (compact_groups_kernel)
  FitState s = p.st[fit];
  const int d = p.d, T = p.T, P = p.P, D1 = d + 1;
  double* x = p.x + (int64_t)fit * P;
  double* g = p.g + (int64_t)fit * P;
  double* dir = p.dir + (int64_t)fit * P;
  double* gt = p.gt + (int64_t)fit * P;
  double* S = p.S + (int64_t)fit * kMaxCor * P;
  double* Y = p.Y + (int64_t)fit * kMaxCor * P;
  double* ys = p.ys + (int64_t)fit * kMaxCor;
  const double* Gr = p.Gred + (int64_t)a * T * p.ldgr;

  // ---- objective and gradient at the trial point x + stp*dir ----
  const double stp_in = s.phase == 0 ? 0.0 : s.stp;
  double pen = 0.0, gd_t = 0.0, gmax = 0.0;
  for (int idx = threadIdx.x; idx < P; idx += blockDim.x) {
    const int t = idx / D1, j = idx - t * D1;
    const double xt = x[idx] + stp_in * dir[idx];
    double gn;
    if (j < d) { gn = Gr[(int64_t)t * p.ldgr + j] * s.inv_ntr + s.l2 * xt; pen += xt * xt; }
    else gn = s.fit_icpt ? Gr[(int64_t)t * p.ldgr + d] * s.inv_ntr : 0.0;
    gt[idx] = gn;
    gd_t += gn * dir[idx];
    gmax = fmax(gmax, fabs(gn));
  }
  pen = block_reduce<false>(pen, red);
  gd_t = block_reduce<false>(gd_t, red);
  gmax = block_reduce<true>(gmax, red);
  const double f_t = Gr[d + 1] * s.inv_ntr + 0.5 * s.l2 * pen;

  bool need_dir = false;  // compute a fresh direction and open a line search
  if (s.phase == 0) {
    s.f = f_t; s.nfev = 1; s.nit = 0; s.col = 0; s.head = 0; s.theta = 1.0;
    for (int idx = threadIdx.x; idx < P; idx += blockDim.x) g[idx] = gt[idx];
    if (gmax <= s.tol) { s.active = 0; s.status = 0; }
    else { s.phase = 1; need_dir = true; }
  } else {
    double stp = s.stp;
    const int task = dcsrch_step(s, stp, f_t, gd_t);
    if (task == TASK_FG) {
      if (s.ifun >= kMaxLs) {          // mainlb: iback >= maxls -> drop the memory or give up
        if (s.col == 0) { s.active = 0; s.status = 2; }
        else { s.col = 0; s.head = 0; s.theta = 1.0; need_dir = true; }
      } else { s.ifun += 1; s.nfev += 1; s.stp = stp; }
    } else {
      // ---- step accepted: x <- x + stp*dir, g <- g(x) ----
      const double dr = (gd_t - s.gd_old) * s.stp, ddum = -s.gd_old * s.stp;
      const bool upd = dr > kEps * ddum;
      int pos = 0;
      if (upd) {
        if (s.col < kMaxCor) { pos = (s.head + s.col) % kMaxCor; }
        else { pos = s.head; }
      }
      double yy = 0.0;
      for (int idx = threadIdx.x; idx < P; idx += blockDim.x) {
        const double sv = s.stp * dir[idx], yv = gt[idx] - g[idx];
        x[idx] += sv; g[idx] = gt[idx];
        if (upd) { S[(int64_t)pos * P + idx] = sv; Y[(int64_t)pos * P + idx] = yv; yy += yv * yv; }
      }
      yy = block_reduce<false>(yy, red);
      s.f_old = s.f; s.f = f_t; s.nit += 1;
      if (s.nit >= s.max_iter) { s.active = 0; s.status = 1; }
      else if (gmax <= s.tol) { s.active = 0; s.status = 0; }
      else if ((s.f_old - s.f) <= kEps * kFactr * fmax(fmax(fabs(s.f_old), fabs(s.f)), 1.0)) { s.active = 0; s.status = 0; }
      else {
        if (upd) {
          if (s.col < kMaxCor) s.col += 1; else s.head = (s.head + 1) % kMaxCor;
          if (threadIdx.x == 0) ys[pos] = dr;
          s.theta = yy / dr;
        }
        need_dir = true;
      }
    }
  }
  __syncthreads();

  while (need_dir) {
    // two-loop recursion with H0 = I/theta (q lives in dir)
    double alpha[kMaxCor];
    for (int idx = threadIdx.x; idx < P; idx += blockDim.x) dir[idx] = g[idx];
    for (int i = s.col - 1; i >= 0; --i) {
      const int pos = (s.head + i) % kMaxCor;
      double sq = 0.0;
      for (int idx = threadIdx.x; idx < P; idx += blockDim.x) sq += S[(int64_t)pos * P + idx] * dir[idx];
      sq = block_reduce<false>(sq, red);
      alpha[i] = sq / ys[pos];
      for (int idx = threadIdx.x; idx < P; idx += blockDim.x) dir[idx] -= alpha[i] * Y[(int64_t)pos * P + idx];
    }
    const double inv_theta = 1.0 / s.theta;
    for (int idx = threadIdx.x; idx < P; idx += blockDim.x) dir[idx] *= inv_theta;
    for (int i = 0; i < s.col; ++i) {
      const int pos = (s.head + i) % kMaxCor;
      double yq = 0.0;
      for (int idx = threadIdx.x; idx < P; idx += blockDim.x) yq += Y[(int64_t)pos * P + idx] * dir[idx];
      yq = block_reduce<false>(yq, red);
      const double c = alpha[i] - yq / ys[pos];
      for (int idx = threadIdx.x; idx < P; idx += blockDim.x) dir[idx] += c * S[(int64_t)pos * P + idx];
    }
    double gd = 0.0, dd = 0.0;
    for (int idx = threadIdx.x; idx < P; idx += blockDim.x) {
      const double v = -dir[idx];
      dir[idx] = v; gd += g[idx] * v; dd += v * v;
    }
    gd = block_reduce<false>(gd, red);
    dd = block_reduce<false>(dd, red);
    if (gd >= 0.0) {                   // lnsrlb info = -4: not a descent direction
      if (s.col == 0) { s.active = 0; s.status = 2; break; }
      s.col = 0; s.head = 0; s.theta = 1.0;
      continue;
    }
    const double stp0 = (s.nit == 0) ? fmin(1.0 / sqrt(dd), kStpMax) : 1.0;
    dcsrch_start(s, stp0, s.f, gd);
    s.gd = gd; s.gd_old = gd; s.stp = stp0; s.ifun = 1; s.nfev += 1;
    break;
  }
  if (threadIdx.x == 0) p.st[fit] = s;
}

// fits still running -> dense list; one block
__global__ void compact_kernel(const FitState* st, int n_fits, int* active_fits, int* n_active, int* n_active_host) {
  __shared__ int base, warp_tot[32];
  if (threadIdx.x == 0) base = 0;
  __syncthreads();
  for (int f0 = 0; f0 < n_fits; f0 += blockDim.x) {
    const int f = f0 + threadIdx.x;
    const int flag = (f < n_fits) ? (st[f].active != 0) : 0;
    const unsigned bal = __ballot_sync(0xffffffffu, flag);
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    if (lane == 0) warp_tot[w] = __popc(bal);
    __syncthreads();
    int off = base;
    for (int i = 0; i < w; ++i) off += warp_tot[i];
    if (flag) active_fits[off + __popc(bal & ((1u << lane) - 1))] = f;
    __syncthreads();
    if (threadIdx.x == 0) { int t = 0; for (int i = 0; i < (int)(blockDim.x >> 5); ++i) t += warp_tot[i]; base += t; }
    __syncthreads();
  }
  if (threadIdx.x == 0) { *n_active = base; *n_active_host = base; }
}

// Live fits -> list for the next round.  With `pad_to` > 0 the list is fold-major: the fits of fold
// (or hold-out split) k, in candidate order, then -1 entries up to a multiple of `pad_to`; group k
// occupies list positions [gstart[k], gstart[k+1]).  A row tile whose rows all sit in fold k trains
// none of group k's columns, so the contraction kernels skip that column range for it.  The padded
// form is used while it costs at most an eighth more columns than the dense list; otherwise (and
// with pad_to == 0) the list is dense in fit order and gstart[0] = -1 (no column ranges).  One block.
struct CompactParams {
  const FitState* st; int n_fits, fpc, pad_to;
  int* active_fits;
  int* counts;        // [0] list length (n_active of the next round), [1] live fits, [2] 1 = fold-major padded,
                      // [3] the list has been dense once (zeroed by the caller at the start of a sweep)
  int* gstart;        // [fpc + 1]
  int* counts_host;   // copy of counts[0..2] the host reads
};
__global__ void compact_groups_kernel(const CompactParams p) {
  __shared__ int cnt[128], start[129], warp_tot[32], base, grouped;
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5, nw = blockDim.x >> 5;
  for (int k = threadIdx.x; k < p.fpc; k += blockDim.x) cnt[k] = 0;
  __syncthreads();
  for (int f = threadIdx.x; f < p.n_fits; f += blockDim.x)
    if (p.st[f].active != 0) atomicAdd(&cnt[f % p.fpc], 1);
  __syncthreads();
  if (threadIdx.x == 0) {
    int dense = 0, padded = 0;
    start[0] = 0;
    for (int k = 0; k < p.fpc; ++k) {
      dense += cnt[k];
      if (p.pad_to > 0) padded += (cnt[k] + p.pad_to - 1) / p.pad_to * p.pad_to;
      start[k + 1] = padded;
    }
    grouped = (p.pad_to > 0 && dense > 0 && padded * 8 <= ((dense + 31) / 32 * 32) * 9) ? 1 : 0;
    // once dense, always dense: a list must never get longer from one round to the next (the host sizes
    // a round's grids from the previous list's length); counts[3] remembers the switch for the sweep
    if (p.counts[3]) grouped = 0;
    if (!grouped) p.counts[3] = 1;
    base = 0;
    p.counts[0] = grouped ? padded : dense; p.counts[1] = dense; p.counts[2] = grouped;
    p.counts_host[0] = p.counts[0]; p.counts_host[1] = dense; p.counts_host[2] = grouped;
    for (int k = 0; k <= p.fpc; ++k) p.gstart[k] = grouped ? start[k] : -1;
  }
  __syncthreads();
  if (!grouped) {                       // dense, fit order
    for (int f0 = 0; f0 < p.n_fits; f0 += blockDim.x) {
      const int f = f0 + threadIdx.x;
      const int flag = (f < p.n_fits) ? (p.st[f].active != 0) : 0;
      const unsigned bal = __ballot_sync(0xffffffffu, flag);
      if (lane == 0) warp_tot[w] = __popc(bal);
      __syncthreads();
      int off = base;
      for (int i = 0; i < w; ++i) off += warp_tot[i];
      if (flag) p.active_fits[off + __popc(bal & ((1u << lane) - 1))] = f;
      __syncthreads();
      if (threadIdx.x == 0) { int t = 0; for (int i = 0; i < nw; ++i) t += warp_tot[i]; base += t; }
      __syncthreads();
    }
    return;
  }
  const int n_cand = p.n_fits / p.fpc;
  for (int k = 0; k < p.fpc; ++k) {
    if (threadIdx.x == 0) base = start[k];
    __syncthreads();
    for (int c0 = 0; c0 < n_cand; c0 += blockDim.x) {
      const int c = c0 + threadIdx.x;
      const int f = c * p.fpc + k;
      const int flag = (c < n_cand) ? (p.st[f].active != 0) : 0;
      const unsigned bal = __ballot_sync(0xffffffffu, flag);
      if (lane == 0) warp_tot[w] = __popc(bal);
      __syncthreads();
      int off = base;
      for (int i = 0; i < w; ++i) off += warp_tot[i];
      if (flag) p.active_fits[off + __popc(bal & ((1u << lane) - 1))] = f;
      __syncthreads();
      if (threadIdx.x == 0) { int t = 0; for (int i = 0; i < nw; ++i) t += warp_tot[i]; base += t; }
      __syncthreads();
    }
    for (int i = start[k] + cnt[k] + threadIdx.x; i < start[k + 1]; i += blockDim.x) p.active_fits[i] = -1;
    __syncthreads();
  }
}

// trial point of every active fit, rounded to fp32, in the layout the contraction kernels read
struct TrialParams {
  int d, T, P, ldw;
  const FitState* st; const double *x, *dir;
  const int* active_fits; const int* n_active;
  float* Wt; float* bt; int32_t* slot_sel; int32_t* slot_val;
  // class weights (null = unweighted): cw_fit [n_fits][K] -> per-slot table cwS [slots][2] the forward kernels
  // read (binary: weights of class 0 / 1; multinomial: [slot][0] = weight of class slot % T), and
  // swS[slot] = weighted count of this rank's training rows of the fit (sw_loc), for the loss shift
  const float* cw_fit; int K; const double* sw_loc; float* cwS; double* swS;
};
__global__ void write_trial_kernel(const TrialParams p) {
  const int a = blockIdx.x;
  if (a >= *p.n_active) return;
  const int fit = p.active_fits[a], D1 = p.d + 1;
  if (p.cw_fit && threadIdx.x < p.T) {
    const int slot = a * p.T + threadIdx.x;
    if (fit < 0) { p.cwS[2 * slot] = 1.f; p.cwS[2 * slot + 1] = 1.f; p.swS[slot] = 0.0; }
    else {
      const float* w = p.cw_fit + (int64_t)fit * p.K;
      if (p.T == 1) { p.cwS[2 * slot] = w[0]; p.cwS[2 * slot + 1] = w[1]; }
      else { p.cwS[2 * slot] = w[threadIdx.x]; p.cwS[2 * slot + 1] = 1.f; }
      p.swS[slot] = p.sw_loc[fit];
    }
  }
  if (fit < 0) {                       // padding entry: zero weights, mask -1 = "holds out every row"
    for (int idx = threadIdx.x; idx < p.P; idx += blockDim.x) {
      const int t = idx / D1, j = idx - t * D1, slot = a * p.T + t;
      if (j < p.d) p.Wt[(int64_t)slot * p.ldw + j] = 0.f;
      else { p.bt[slot] = 0.f; p.slot_sel[slot] = -1; p.slot_val[slot] = 0; }
    }
    return;
  }
  const FitState& s = p.st[fit];
  const double stp = s.phase == 0 ? 0.0 : s.stp;
  const double* x = p.x + (int64_t)fit * p.P;
  const double* dir = p.dir + (int64_t)fit * p.P;
  for (int idx = threadIdx.x; idx < p.P; idx += blockDim.x) {
    const int t = idx / D1, j = idx - t * D1, slot = a * p.T + t;
    const float v = (float)(x[idx] + stp * dir[idx]);
    if (j < p.d) p.Wt[(int64_t)slot * p.ldw + j] = v;
    else { p.bt[slot] = v; p.slot_sel[slot] = s.mask_sel; p.slot_val[slot] = s.mask_val; }
  }
}

// partial sums -> fp64 per-slot record [G(0..d-1), gb, loss]; fixed summation order.
// `accumulate` adds to the record (row chunks after the first).
struct ReduceParams {
  const float* Gpart; int ldg; int splits; int n_slots; int d; int T; int jcount;  // jcount = d+1 (FFMA path: column d = sum r) or d
  const double* lossPart; int64_t tile0, n_tiles;
  double* Gred; int ldgr; int accumulate;
  const int* n_list_dev;   // live columns = *n_list_dev * T when set (n_slots: upper bound and pitch)
};
__global__ void reduce_grad_kernel(const ReduceParams p) {
  const int slot = blockIdx.x;
  if (p.n_list_dev && slot >= *p.n_list_dev * p.T) return;
  const int64_t stride = (int64_t)p.n_slots * p.ldg;
  for (int j = threadIdx.x; j < p.jcount; j += blockDim.x) {
    const float* src = p.Gpart + (int64_t)slot * p.ldg + j;
    double a[8] = {0.0, 0.0, 0.0, 0.0, 0.0, 0.0, 0.0, 0.0};   // 8 independent chains; fixed association: deterministic
    int s = 0;
    for (; s + 8 <= p.splits; s += 8) {
      float v[8];
#pragma unroll
      for (int u = 0; u < 8; ++u) v[u] = src[(int64_t)(s + u) * stride];
#pragma unroll
      for (int u = 0; u < 8; ++u) a[u] += (double)v[u];
    }
    for (; s < p.splits; ++s) a[0] += (double)src[(int64_t)s * stride];
    const double acc = ((a[0] + a[1]) + (a[2] + a[3])) + ((a[4] + a[5]) + (a[6] + a[7]));
    double* q = p.Gred + (int64_t)slot * p.ldgr + j;
    *q = p.accumulate ? *q + acc : acc;
  }
}
// loss of group g (fit) = sum over this chunk's row tiles; block = (32 groups, 8 tile lanes)
__global__ void reduce_tiles_kernel(const ReduceParams p) {
  __shared__ double sh[8][33];
  const int n_groups = p.n_slots / p.T;
  const int g = blockIdx.x * 32 + threadIdx.x;
  double a = 0.0;
  if (g < n_groups)
    for (int64_t t = p.tile0 + threadIdx.y; t < p.tile0 + p.n_tiles; t += 8) a += p.lossPart[t * n_groups + g];
  sh[threadIdx.y][threadIdx.x] = a;
  __syncthreads();
  if (threadIdx.y == 0 && g < n_groups) {
    double l = 0.0;
    for (int i = 0; i < 8; ++i) l += sh[i][threadIdx.x];
    double* q = p.Gred + (int64_t)g * p.T * p.ldgr + p.d + 1;
    *q = p.accumulate ? *q + l : l;
  }
}

// labels -> histogram (256 bins per label array); n_test of a fit = hist[sel][val]
__global__ void label_hist_kernel(const int8_t* fold_id, const int8_t* hold_id, int64_t n, unsigned long long* hist) {
  __shared__ unsigned int sh[512];
  for (int i = threadIdx.x; i < 512; i += blockDim.x) sh[i] = 0;
  __syncthreads();
  for (int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; r < n; r += (int64_t)gridDim.x * blockDim.x) {
    atomicAdd(&sh[(uint8_t)fold_id[r]], 1u);
    if (hold_id) atomicAdd(&sh[256 + (uint8_t)hold_id[r]], 1u);
  }
  __syncthreads();
  for (int i = threadIdx.x; i < 512; i += blockDim.x) if (sh[i]) atomicAdd(&hist[i], (unsigned long long)sh[i]);
}

// (label value, class) histogram for class weights: hist[(sel * 256 + value) * K + class]
__global__ void label_class_hist_kernel(const int8_t* fold_id, const int8_t* hold_id, const int32_t* y, int64_t n, int K,
                                        unsigned long long* hist) {
  for (int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; r < n; r += (int64_t)gridDim.x * blockDim.x) {
    const int c = y[r];
    atomicAdd(&hist[(int64_t)(uint8_t)fold_id[r] * K + c], 1ull);
    if (hold_id) atomicAdd(&hist[(int64_t)(256 + (uint8_t)hold_id[r]) * K + c], 1ull);
  }
}
// sw[fit] = sum over classes of class_weight[fit][class] * (training rows of that class), from a (label value,
// class) histogram: this rank's rows (sw_loc) or all ranks' (sw_glob).  Also narrows the weights to fp32 (cw_fit).
// One thread per fit.
__global__ void fit_weights_kernel(const double* class_weight, int n_fits, int fpc, int n_folds, int K,
                                   const unsigned long long* hist_loc, const unsigned long long* hist_glob,
                                   float* cw_fit, double* sw_loc, double* sw_glob) {
  const int f = blockIdx.x * blockDim.x + threadIdx.x;
  if (f >= n_fits) return;
  const int k = f % fpc;
  const int64_t held = (k < n_folds) ? (int64_t)k * K : (int64_t)(256 + 1) * K;
  double a = 0.0, b = 0.0;
  for (int c = 0; c < K; ++c) {
    const double w = class_weight[(int64_t)f * K + c];
    cw_fit[(int64_t)f * K + c] = (float)w;
    unsigned long long tl = 0, tg = 0;
    for (int v = 0; v < 256; ++v) { tl += hist_loc[(int64_t)v * K + c]; tg += hist_glob[(int64_t)v * K + c]; }
    // the weight the kernels apply is the fp32 one
    a += (double)(float)w * (double)(tl - hist_loc[held + c]);
    b += (double)(float)w * (double)(tg - hist_glob[held + c]);
  }
  sw_loc[f] = a; sw_glob[f] = b;
}

struct InitParams {
  int n_fits, n_folds, has_hold, fits_per_cand;
  int64_t n_total;                 // rows over all ranks
  const unsigned long long* hist;  // summed over ranks
  const double* C; const double* tol; const uint8_t* fit_icpt; const int32_t* max_iter;  // per candidate
  FitState* st; int* active_fits; int* n_active;
  const double* sw_glob;           // class weights: weighted training-row count per fit (null = unweighted)
};
// fit index = cand*fits_per_cand + k ; k < n_folds: CV fold k held out ; k == n_folds: hold-out split
__global__ void init_fits_kernel(const InitParams p) {
  const int f = blockIdx.x * blockDim.x + threadIdx.x;
  if (f == 0) *p.n_active = p.n_fits;
  if (f >= p.n_fits) return;
  const int cand = f / p.fits_per_cand, k = f % p.fits_per_cand;
  FitState s;
  memset(&s, 0, sizeof(s));
  s.mask_sel = (k < p.n_folds) ? 0 : 1;
  s.mask_val = (k < p.n_folds) ? k : 1;
  const double n_te = (double)p.hist[s.mask_sel * 256 + s.mask_val];
  // with sample weights the objective is sum(s_i loss_i) / sum(s_i) + l2/2 |w|^2, l2 = 1 / (C sum(s_i))
  // (_linear_loss.py:340-361, _logistic.py:580 with sw_sum)
  const double n_tr = p.sw_glob ? p.sw_glob[f] : (double)p.n_total - n_te;
  s.inv_ntr = 1.0 / n_tr;
  s.l2 = 1.0 / (p.C[cand] * n_tr);   // _logistic.py:580
  s.tol = p.tol[cand];
  s.max_iter = p.max_iter[cand];
  s.fit_icpt = p.fit_icpt[cand];
  s.theta = 1.0; s.active = 1; s.phase = 0;
  p.st[f] = s;
  p.active_fits[f] = f;
}

struct FinalParams {
  int n_fits, d, T, P, ldw;
  const FitState* st; const double* x;
  float* coef; float* icpt; int32_t* n_iter; int32_t* status; int32_t* nfev;
  float* Wt; float* bt; int32_t* slot_sel; int32_t* slot_val;  // score-pass operands (all fits)
  // score-pass column of fit (cand, k): fit-major cand * fpc + k when npad == 0, else fold-major
  // k * npad + cand (npad = candidates rounded up to 32; the caller presets the padding columns), so that
  // a row tile of fold k only needs the column range of group k
  int fpc, npad;
};
__device__ __forceinline__ int score_slot(int fit, int fpc, int npad) {
  return npad ? (fit % fpc) * npad + fit / fpc : fit;
}
__global__ void finalize_kernel(const FinalParams p) {
  const int fit = blockIdx.x, D1 = p.d + 1;
  const FitState& s = p.st[fit];
  const double* x = p.x + (int64_t)fit * p.P;
  for (int idx = threadIdx.x; idx < p.P; idx += blockDim.x) {
    const int t = idx / D1, j = idx - t * D1, slot = fit * p.T + t;
    const int ss = score_slot(fit, p.fpc, p.npad) * p.T + t;
    const float v = (float)x[idx];   // coef_ is cast to X.dtype (_logistic.py:702-712)
    if (j < p.d) { p.coef[(int64_t)slot * p.d + j] = v; p.Wt[(int64_t)ss * p.ldw + j] = v; }
    else { p.icpt[slot] = v; p.bt[ss] = v; p.slot_sel[ss] = s.mask_sel; p.slot_val[ss] = s.mask_val; }
  }
  if (threadIdx.x == 0) {
    p.n_iter[fit] = min(s.nit, s.max_iter);   // sklearn/utils/optimize.py:405-460
    p.status[fit] = s.status;
    p.nfev[fit] = s.nfev;
  }
}

// score partials -> totals
__global__ void reduce_counts_kernel(const int32_t* cntPart, int64_t n_tiles, int n_slots, int T, long long* correct,
                                     int fpc = 1, int npad = 0) {
  __shared__ long long sh[256];
  const int fit = blockIdx.x;
  const int col = score_slot(fit, fpc, npad) * T;
  long long a = 0;
  for (int64_t t = threadIdx.x; t < n_tiles; t += blockDim.x) a += cntPart[t * n_slots + col];
  sh[threadIdx.x] = a;
  __syncthreads();
  for (int o = blockDim.x / 2; o > 0; o >>= 1) { if (threadIdx.x < o) sh[threadIdx.x] += sh[threadIdx.x + o]; __syncthreads(); }
  if (threadIdx.x == 0) correct[fit] = sh[0];
}
__global__ void reduce_sse_kernel(const float* ssePart, int64_t n_tiles, int n_slots, double* sse) {
  __shared__ double red[33];
  const int slot = blockIdx.x;
  double a = 0.0;
  for (int64_t t = threadIdx.x; t < n_tiles; t += blockDim.x) a += (double)ssePart[t * n_slots + slot];
  a = block_reduce<false>(a, red);
  if (threadIdx.x == 0) sse[slot] = a;
}

}  // namespace dml
