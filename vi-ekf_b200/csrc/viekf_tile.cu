// viekf_tile.cu -- register-tile tape executor of the VI-EKF batch engine (sm_100a, FP64).
//
// Same contract as viekf_kernel.cu (one thread group owns one filter for a whole tape slice; events in order; HBM layout of
// x / packed P unchanged), different placement of the covariance on chip:
//
//   P_ff (feature-feature, the O(N^2) part)  lives in REGISTERS: the lower block-triangle of 2x2-feature tiles (6x6 doubles),
//                                            one tile per thread of the TW "tile warps"; diagonal tiles keep a mirrored copy
//   P_bb (16x16, full symmetric), P_fb (3N x 16)  live in shared memory and are owned by the last ("base") warp
//
// so the two O(N^2) loops of the filter -- the structured A_d P A_d^T of propagate_state (vi_ekf.cpp:302-304) and the
// Joseph / partial covariance update (vi_ekf_meas.cpp:254-265) -- run as fully unrolled FP64 FMA streams on register
// operands; shared memory only supplies the O(N) row / column vectors, mostly as warp broadcasts.  Feature removal
// (clear_feature / keep_only_features, vi_ekf_feat.cpp:50-142) reshuffles the tiles through the filter's own packed P in
// global memory (L2 round trip), which is also the load / store path at the ends of a slice.
//
// Padding: with `len` live features the tile grid covers lenp = len rounded up to even; every quantity of the padding
// feature (A_d rows, G_d rows, Y rows, P_fb rows, update vectors, tile entries) is an exact zero, so no per-element guards
// are needed inside the unrolled loops.
#include <stdlib.h>
#include "viekf_dev.cuh"

namespace vk {

constexpr int PBS = 18;  // row stride of P_bb: even (16-byte rows for vector loads), 2-way conflicts at worst for row-per-lane access

struct TLayout {
  int x, Pbb, Pfb, lam, dx, Abd, AbT, T, Ac, Gb, Afd, Gd, Y, upd, small, ids, smap, evb, fixk, fixv, zvr, total, TR;
};
__host__ __device__ inline TLayout make_tlayout(int N) {
  TLayout L;
  L.TR = (N + 1) / 2;
  const int NP = 2 * L.TR;
  const int nmaxp = 16 + 3 * NP;
  int o = 0;
  L.x = o;     o += ev2(17 + 5 * N);
  L.Pbb = o;   o += 16 * PBS;           // full symmetric 16 x 16, row stride PBS
  L.Pfb = o;   o += ev2(3 * NP * 17);   // [feature row][17]: base columns 0..15, stride 17 (conflict-free row-per-lane access)
  L.lam = o;   o += ev2(nmaxp);
  L.dx = o;    o += ev2(nmaxp);
  L.Abd = o;   o += 256;
  L.AbT = o;   o += 256;
  L.T = o;     o += 16 * 17;             // A_bb P_bb, row stride 17 (conflict-free row-per-lane reads); also the packed P_bb copy
  L.Ac = o;    o += 144;
  L.Gb = o;    o += 54 + 2;
  L.Afd = o;   o += ev2(3 * NP * 19);
  L.Gd = o;    o += ev2(nmaxp * 6);
  L.Y = o;     o += ev2(3 * NP * 17 > 27 * N ? 3 * NP * 17 : 27 * N);
  L.upd = L.Afd;
  const int upd_end = L.upd + ev2(10 * nmaxp);
  if (upd_end > o) o = upd_end;
  L.small = o; o += 80;
  L.ids = o;   o += ev2(N + 12);
  L.smap = o;  o += ev2(NP / 2 + 2);
  L.evb = o;   o += 64;                  // two staged viekf_event records (31 doubles each): the current and the next event
  L.fixk = o;  o += ev2(NP / 2 + 1);     // pending fix_depth covariance corrections per feature: kind (int32) ...
  L.fixv = o;  o += NP;                  // ... and value
  L.zvr = o;   o += 64;                  // ring of 8 resolved value blocks (8 doubles each): event r reads block r & 7
  L.total = ev2(o);
  return L;
}

size_t tile_smem_bytes(int N) { return (size_t)make_tlayout(N).total * sizeof(double); }

// Optional phase timing (build with -DVIEKF_PHASE_TIMING): thread 0 of every CTA adds the clocks between consecutive marks to
// g_phase_clk[k]; since the marks sit right after barriers this is the elapsed time of each phase on the CTA's critical path.
#ifdef VIEKF_PHASE_TIMING
__device__ unsigned long long g_phase_clk[4 * 64];
__device__ unsigned long long g_phase_cnt[4 * 64];
#define PH_DECL long long ph_last;
#define PH_INIT(c) (c).ph_last = clock64();
// slot k: work of the phase (mark before the barrier); slot k + 32: wait at the barrier (mark after it); one row per warp
#define PH(c, k)                                                                                  \
  if ((threadIdx.x & 31) == 0) {                                                                  \
    const long long ph_now = clock64();                                                           \
    atomicAdd(&g_phase_clk[64 * ((threadIdx.x >> 5) & 3) + (k)], (unsigned long long)(ph_now - (c).ph_last)); \
    atomicAdd(&g_phase_cnt[64 * ((threadIdx.x >> 5) & 3) + (k)], 1ull);                                 \
    (c).ph_last = ph_now;                                                                         \
  }
#define PHS(c, k) \
  PH(c, k)        \
  (c).sync();     \
  PH(c, (k) + 32)
#else
#define PH_DECL
#define PH_INIT(c)
#define PH(c, k)
#define PHS(c, k) (c).sync();
#endif

template <int TW>
struct TCtx : Ctx<32 * (TW + 1)> {
  double *Pbb, *Pfb, *fixv;
  int *smap, *fixk;
  int ti, tj;   // tile coordinates in units of 2 features, ti >= tj
  bool tile;    // this thread owns a tile of the (capacity) grid
  bool basew;   // this thread belongs to the base warp
  int lane;
  // per-instance values (Philox noise: a ~400-instruction dependent chain per value) are resolved four events at a time --
  // every fourth event, one warp resolves events e+1..e+4 on 4 x 8 lanes -- by a warp that has slack in that event: the base
  // warp in propagate phase D2, warp 1 in the rank-update phase; otherwise by warp 1 up front
  int64_t next_e;    // first event to resolve in this event's slot, -1 = none
  int64_t ginst;
  double* zvr;       // ring of 8 value blocks
  bool resolved;
  PH_DECL
};

template <int TW>
__device__ __forceinline__ void resolve_next(TCtx<TW>& c, const KParams& p, int first_tid) {
  if (c.next_e >= 0 && c.tid >= first_tid && c.tid < first_tid + 32) {
    const int l = c.tid - first_tid;
    const int64_t ev = c.next_e + (l >> 3);
    if (ev < p.ev_first + p.ev_count) resolve_values(p, ev, c.b, c.ginst, l & 7, c.zvr + 8 * (int)((ev - p.ev_first) & 7));
  }
  c.resolved = true;
}

#define PT(r, c) Pt[(r) * 6 + (c)]

#ifndef VIEKF_LOCKSTEP
#define VIEKF_LOCKSTEP 1
#endif
constexpr int LOCKSTEP = VIEKF_LOCKSTEP;  // events between CTA-wide barriers (power of two)


// Base-warp threads own no tile: their Pt is all zeros for the whole launch.  Re-zeroing it at the end of a base-warp-only
// section makes the 72 tile registers dead inside that section, so its loops get the full register budget (loads batched
// ahead of the FMAs instead of one load-use round trip per FMA pair).
__device__ __forceinline__ void tile_dead(double (&Pt)[36]) {
#pragma unroll
  for (int i = 0; i < 36; i++) Pt[i] = 0.0;
}

template <int TW>
__device__ __forceinline__ void tile_mirror(const TCtx<TW>& c, double (&Pt)[36]) {
  if (c.tile && c.ti == c.tj) {
#pragma unroll
    for (int r = 0; r < 6; r++)
#pragma unroll
      for (int cc = r + 1; cc < 6; cc++) PT(r, cc) = PT(cc, r);
  }
}

// tiles <- packed P in global memory; new feature f comes from old feature smap[f] (identity when !use_map); dead blocks = 0
template <int TW>
__device__ __forceinline__ void tile_load(const TCtx<TW>& c, const double* gP, double (&Pt)[36], int len, bool use_map) {
#pragma unroll
  for (int bi = 0; bi < 2; bi++)
#pragma unroll
    for (int bj = 0; bj < 2; bj++) {
      const int fi = 2 * c.ti + bi, fj = 2 * c.tj + bj;
      const bool live = c.tile && fi < len && fj < len;
      int oi = fi, oj = fj;
      if (live && use_map) {
        oi = c.smap[fi];
        oj = c.smap[fj];
      }
#pragma unroll
      for (int rr = 0; rr < 3; rr++)
#pragma unroll
        for (int cc = 0; cc < 3; cc++) {
          double v = 0.0;
          if (live) v = __ldcg(gP + psym(16 + 3 * oi + rr, 16 + 3 * oj + cc));
          PT(3 * bi + rr, 3 * bj + cc) = v;
        }
    }
}

// packed P in global memory <- tiles (lower triangle of the live part only)
template <int TW>
__device__ __forceinline__ void tile_store(const TCtx<TW>& c, double* gP, const double (&Pt)[36], int len) {
#pragma unroll
  for (int bi = 0; bi < 2; bi++)
#pragma unroll
    for (int bj = 0; bj < 2; bj++) {
      const int fi = 2 * c.ti + bi, fj = 2 * c.tj + bj;
      const bool live = c.tile && fi < len && fj < len && fi >= fj;
      if (live) {
#pragma unroll
        for (int rr = 0; rr < 3; rr++)
#pragma unroll
          for (int cc = 0; cc < 3; cc++)
            if (fi > fj || rr >= cc) gP[plo(16 + 3 * fi + rr, 16 + 3 * fj + cc)] = PT(3 * bi + rr, 3 * bj + cc);
      }
    }
}

// fix_depth (vi_ekf_helper.cpp:128-156), split so that it needs no barrier of its own: the thread that just box-plussed
// feature i checks its inverse depth, resets it and records the covariance correction (P_rr += err^2, or P_rr = P0_feat);
// the owner of the diagonal tile applies the recorded correction after the barrier that ends the event, before anything
// reads P again.
template <int TW>
__device__ __forceinline__ void tile_fix_x(const TCtx<TW>& c, const KParams& p, int i) {
  double* rho = c.x + 17 + 5 * i + 4;
  const double reset = p.depth_reset;
  double r = *rho;
  uint32_t fl = 0;
  int nfix = 0;
  if (r != r) {
    r = reset;
    fl |= VIEKF_FLAG_DEPTH_NAN_FIXED;
    nfix++;
  }
  if (r < 0.0) {
    const double err = reset - r;
    c.fixv[i] = err * err;
    c.fixk[i] = 1;
    r = reset;
    fl |= VIEKF_FLAG_DEPTH_NEG_FIXED;
    nfix++;
  } else if (r > 1e2) {
    c.fixk[i] = 2;
    r = reset;
    fl |= VIEKF_FLAG_DEPTH_BIG_FIXED;
    nfix++;
  }
  if (nfix) {
    *rho = r;
    atomicOr((unsigned*)&c.sc[SC_FLAGS], fl);
    atomicAdd(&c.sc[SC_NFIX], nfix);
  }
}
template <int TW>
__device__ __forceinline__ void tile_apply_fix(const TCtx<TW>& c, const KParams& p, double (&Pt)[36]) {
  if (!(c.tile && c.ti == c.tj)) return;
#pragma unroll
  for (int bi = 0; bi < 2; bi++) {
    const int i = 2 * c.ti + bi;
    const int k = c.fixk[i];
    if (k) {
      if (k == 1)
        PT(3 * bi + 2, 3 * bi + 2) += c.fixv[i];
      else
        PT(3 * bi + 2, 3 * bi + 2) = p.P0f[2];
      c.fixk[i] = 0;
    }
  }
}

// acc[0..W) += sum_{k<K} a[k * AS] * M[k * MS + 0..W), software-pipelined by hand: the 16 operands of row k+1 are loaded
// before the 16 FMAs of row k are issued (ptxas otherwise pairs every load with its two consumers and exposes one
// shared-memory round trip, ~29 clk, per FMA pair).  VEC: rows are 16-byte aligned (8 x LDS.128), else 16 x LDS.64.
template <int W, bool VEC>
__device__ __forceinline__ void load_row(double (&v)[W], const double* M) {
  if (VEC) {
    const double2* r = (const double2*)M;
#pragma unroll
    for (int q = 0; q < W / 2; q++) {
      const double2 t = r[q];
      v[2 * q] = t.x;
      v[2 * q + 1] = t.y;
    }
  } else {
#pragma unroll
    for (int q = 0; q < W; q++) v[q] = M[q];
  }
}
template <int W>
__device__ __forceinline__ void fma_row(double (&acc)[W], double a, const double (&v)[W]) {
#pragma unroll
  for (int q = 0; q < W; q++) acc[q] = fma(a, v[q], acc[q]);
}
// W = 16 on threads with a free register file (base warp), W = 8 (two column halves) on threads that hold a tile
template <int W, int K, int AS, int MS, bool VEC>
__device__ __forceinline__ void rows(double (&acc)[W], const double* a, const double* M) {
  double v[W], w[W];
  load_row<W, VEC>(v, M);
  double av = a[0], aw = 0.0;
#pragma unroll 1
  for (int k = 0; k + 2 <= K; k += 2) {
    load_row<W, VEC>(w, M + (k + 1) * MS);
    aw = a[(k + 1) * AS];
    fma_row<W>(acc, av, v);
    if (k + 2 < K) {
      load_row<W, VEC>(v, M + (k + 2) * MS);
      av = a[(k + 2) * AS];
    }
    fma_row<W>(acc, aw, w);
  }
  if (K & 1) fma_row<W>(acc, av, v);
}

// Base-warp loops live in out-of-line functions: the caller's 72 tile registers are dead there (tile_dead), so these get
// their own register allocation and a latency-tolerant schedule (loads batched ahead of the FMAs) instead of the
// pressure-minimising one the 128-register kernel body forces.
//
// P'_fb = Y A_bb^T + G_d,f Qu G_d,b^T: one row per lane, 16 independent accumulators, operands as warp broadcasts.
// GqT[q][cc] = Qu[q] G_d(cc, q) goes to the T buffer (idle since D1).
static __device__ __noinline__ void base_pfb(const double* Y, const double* AbT, const double* Gd, double* GqT, double* Pfb, const double* Qu, int nfr,
                                             int lane) {
  __builtin_assume(__isShared(Y));
  __builtin_assume(__isShared(AbT));
  __builtin_assume(__isShared(Gd));
  __builtin_assume(__isShared(GqT));
  __builtin_assume(__isShared(Pfb));
  for (int e = lane; e < 96; e += 32) {
    const int q = e >> 4, cc = e & 15;
    GqT[e] = Gd[cc * 6 + q] * Qu[q];
  }
  __syncwarp();
  for (int fr = lane; fr < nfr; fr += 32) {
    const double* Yr = Y + fr * 17;
    const double* Gr = Gd + (16 + fr) * 6;
    double acc[16];
#pragma unroll
    for (int q = 0; q < 16; q++) acc[q] = 0.0;
    rows<16, 16, 1, 16, true>(acc, Yr, AbT);
    rows<16, 6, 1, 16, true>(acc, Gr, GqT);
    double* Pr = Pfb + fr * 17;
#pragma unroll
    for (int q = 0; q < 16; q++) Pr[q] = acc[q];
  }
}

// ------------------------------------------------------------------------------------------------------------------
// PROP: propagate_state (vi_ekf.cpp:262-318) with the structured covariance product
// ------------------------------------------------------------------------------------------------------------------
template <int TW>
__device__ __forceinline__ void tile_propagate(TCtx<TW>& c, const KParams& p, const viekf_event& E, double (&Pt)[36]) {
  constexpr int G = 32 * (TW + 1);
  double dt;
  int len;
  if (!prop_discretise(c, p, E, dt, len, 1)) return;
  PH(c, 1)
  const int lenp = (len + 1) & ~1;
  const int nfr = 3 * lenp;
  // ---- phase C: one row of Y = A_fb P_bb + A_ff P_fb (3 lenp x 16) or of T = A_bb P_bb per thread ----
  // (Y rows on the tile warps, T rows on the base warp: no divergent warp)
  if (!c.basew) {
    for (int fr = c.tid; fr < nfr; fr += 32 * TW) {
      const int i = fr / 3;
      const double* Ad = c.Afd + fr * 19;
      double* Yr = c.Y + fr * 17;
#pragma unroll 1
      for (int h = 0; h < 16; h += 8) {  // two column halves: 8 accumulators + two operand rows in flight next to the 72 tile registers
        double acc[8];
#pragma unroll
        for (int q = 0; q < 8; q++) acc[q] = 0.0;
        rows<8, 13, 1, PBS, true>(acc, Ad + 3, c.Pbb + 3 * PBS + h);
        rows<8, 3, 1, 17, false>(acc, Ad + 16, c.Pfb + 3 * i * 17 + h);
#pragma unroll
        for (int q = 0; q < 8; q++) Yr[h + q] = acc[q];
      }
    }
  } else if (c.lane < 16) {
    const int r = c.lane;
    double acc[16];
#pragma unroll
    for (int q = 0; q < 16; q++) acc[q] = 0.0;
    rows<16, 16, 1, PBS, true>(acc, c.Abd + r * 16, c.Pbb);
#pragma unroll
    for (int q = 0; q < 16; q++) c.T[r * 17 + q] = acc[q];
    tile_dead(Pt);
  }
  PHS(c, 2)
  // ---- phase D1: tiles: Z = A_ff,i P_ij + A_fb,i P_bf,j (last readers of the old P_fb); base warp: P'_bb in place (T holds
  //      A_bb P_bb) and the state boxplus ----
  const bool live_tile = c.tile && 2 * c.ti < len;
  const int R = 6 * c.ti, C = 6 * c.tj;
  if (c.tid == G - 1 || c.tid < len) boxplus_inplace(c, c.dx, dt, len);  // only the warps with a share make the out-of-line call
  if (c.tid < len) tile_fix_x(c, p, c.tid);  // phase F of the other executor, by the thread that owns the feature's state
  if (c.tid == 0) *c.tcur = E.t;
  if (live_tile) {
    // Z = A_ff,i P_ij  (block rows), in place
#pragma unroll
    for (int bi = 0; bi < 2; bi++) {
      double Af[9];
#pragma unroll
      for (int a = 0; a < 3; a++)
#pragma unroll
        for (int k = 0; k < 3; k++) Af[a * 3 + k] = c.Afd[(R + 3 * bi + a) * 19 + 16 + k];
#pragma unroll
      for (int cc = 0; cc < 6; cc++) {
        const double t0 = PT(3 * bi, cc), t1 = PT(3 * bi + 1, cc), t2 = PT(3 * bi + 2, cc);
#pragma unroll
        for (int a = 0; a < 3; a++) PT(3 * bi + a, cc) = fma(Af[a * 3 + 2], t2, fma(Af[a * 3 + 1], t1, Af[a * 3] * t0));
      }
    }
    // Z += A_fb,i P_bf,j   (base columns VEL..MU: 3..15)
#pragma unroll 1
    for (int k = 3; k < 16; k++) {
      double ar[6], bc[6];
#pragma unroll
      for (int r = 0; r < 6; r++) ar[r] = c.Afd[(R + r) * 19 + k];
#pragma unroll
      for (int cc = 0; cc < 6; cc++) bc[cc] = c.Pfb[(C + cc) * 17 + k];
#pragma unroll
      for (int r = 0; r < 6; r++)
#pragma unroll
        for (int cc = 0; cc < 6; cc++) PT(r, cc) = fma(ar[r], bc[cc], PT(r, cc));
    }
  }
  if (c.basew) {
    // P'_bb = T A_bb^T + G_d Qu G_d^T + Qx: lower triangle, mirrored
    for (int e = c.lane; e < 136; e += 32) {
      int r = (int)((sqrtf(8.0f * (float)e + 1.0f) - 1.0f) * 0.5f);
      while (((r + 1) * (r + 2) >> 1) <= e) r++;
      while (((r * (r + 1)) >> 1) > e) r--;
      const int cc = e - ((r * (r + 1)) >> 1);
      double s = 0;
#pragma unroll
      for (int k = 0; k < 16; k++) s = fma(c.T[r * 17 + k], c.AbT[k * 16 + cc], s);
#pragma unroll
      for (int q = 0; q < 6; q++) s = fma(c.Gd[r * 6 + q] * p.Qu[q], c.Gd[cc * 6 + q], s);
      if (r == cc) s += p.Qx[r];
      c.Pbb[r * PBS + cc] = s;
      c.Pbb[cc * PBS + r] = s;
    }
    tile_dead(Pt);
  }
  PHS(c, 3)
  // ---- phase D2: tiles: P'_ij = Z A_ff,j^T + Y_i A_fb,j^T + G_d,i Qu G_d,j^T; base warp: P'_fb = Y A_bb^T + ... in place ----
  if (live_tile) {
    // O = Z A_ff,j^T  (block columns), in place
#pragma unroll
    for (int bj = 0; bj < 2; bj++) {
      double Af[9];
#pragma unroll
      for (int b = 0; b < 3; b++)
#pragma unroll
        for (int k = 0; k < 3; k++) Af[b * 3 + k] = c.Afd[(C + 3 * bj + b) * 19 + 16 + k];
#pragma unroll
      for (int r = 0; r < 6; r++) {
        const double t0 = PT(r, 3 * bj), t1 = PT(r, 3 * bj + 1), t2 = PT(r, 3 * bj + 2);
#pragma unroll
        for (int b = 0; b < 3; b++) PT(r, 3 * bj + b) = fma(t2, Af[b * 3 + 2], fma(t1, Af[b * 3 + 1], t0 * Af[b * 3]));
      }
    }
    // O += Y_i A_fb,j^T
#pragma unroll 1
    for (int k = 3; k < 16; k++) {
      double yr[6], ac[6];
#pragma unroll
      for (int r = 0; r < 6; r++) yr[r] = c.Y[(R + r) * 17 + k];
#pragma unroll
      for (int cc = 0; cc < 6; cc++) ac[cc] = c.Afd[(C + cc) * 19 + k];
#pragma unroll
      for (int r = 0; r < 6; r++)
#pragma unroll
        for (int cc = 0; cc < 6; cc++) PT(r, cc) = fma(yr[r], ac[cc], PT(r, cc));
    }
    // O += G_d,i Qu G_d,j^T
#pragma unroll 1
    for (int q = 0; q < 6; q++) {
      double gr[6], gc[6];
      const double qu = p.Qu[q];
#pragma unroll
      for (int r = 0; r < 6; r++) gr[r] = c.Gd[(16 + R + r) * 6 + q] * qu;
#pragma unroll
      for (int cc = 0; cc < 6; cc++) gc[cc] = c.Gd[(16 + C + cc) * 6 + q];
#pragma unroll
      for (int r = 0; r < 6; r++)
#pragma unroll
        for (int cc = 0; cc < 6; cc++) PT(r, cc) = fma(gr[r], gc[cc], PT(r, cc));
    }
    if (c.ti == c.tj) {
#pragma unroll
      for (int bi = 0; bi < 2; bi++)
        if (2 * c.ti + bi < len) {
#pragma unroll
          for (int a = 0; a < 3; a++) PT(3 * bi + a, 3 * bi + a) += p.Qxf[a];
        }
    }
    tile_mirror(c, Pt);
  }
  if (c.basew) {
    base_pfb(c.Y, c.AbT, c.Gd, c.T, c.Pfb, p.Qu, nfr, c.lane);
    resolve_next(c, p, 32 * TW);
    tile_dead(Pt);
  } else {
    c.resolved = true;
  }
  PH(c, 4)
  // no barrier here: the one that ends the event publishes P'_fb, the resolved values and the pending depth corrections
}

// ------------------------------------------------------------------------------------------------------------------
// UPD: VIEKF::update (vi_ekf_meas.cpp:196-278) as a symmetric rank update
//   P_ij += (1 - a_i a_j) ( W_i.K_j - K_i.H_j ),   H = P H^T,  K = H S^-1,  W = K S - H,  a = 1 - lambda (partial update) or 0
// which is the Joseph form (I-KH) P (I-KH)^T + K R K^T - P = K S K^T - K (HP) - (HP)^T K^T with the two terms that share
// K_j merged.  Vectors live in the update workspace (padding rows are exact zeros).
// ------------------------------------------------------------------------------------------------------------------
struct TUpdVec {
  double2 *Kv, *Hv, *Wv;
  double *K2, *H2, *W2, *av;
  __device__ __forceinline__ TUpdVec(double* w, int nmaxp) {
    Kv = (double2*)w;
    Hv = (double2*)(w + 2 * nmaxp);
    Wv = (double2*)(w + 4 * nmaxp);
    K2 = w + 6 * nmaxp;
    H2 = w + 7 * nmaxp;
    W2 = w + 8 * nmaxp;
    av = w + 9 * nmaxp;
  }
};

template <int RD>
__device__ __forceinline__ double rank_delta(double2 wr, double2 kr, double w2r, double k2r, double2 kc, double2 hc, double k2c, double h2c) {
  double d = wr.x * kc.x;
  d = fma(-kr.x, hc.x, d);
  if (RD >= 2) {
    d = fma(wr.y, kc.y, d);
    d = fma(-kr.y, hc.y, d);
  }
  if (RD == 3) {
    d = fma(w2r, k2c, d);
    d = fma(-k2r, h2c, d);
  }
  return d;
}

// rows of [P_bb; P_fb]: one row per lane, three rows per pass; column vectors are warp broadcasts (out of line: see base_pfb)
template <int RD>
static __device__ __noinline__ void base_rank(TUpdVec v, double* Pbb, double* Pfb, int npad, int lane) {
  constexpr int NR = 3;
  // every pointer is into shared memory: lets the compiler emit LDS / STS instead of generic loads and stores
  __builtin_assume(__isShared(Pbb));
  __builtin_assume(__isShared(Pfb));
  __builtin_assume(__isShared(v.Kv));
  __builtin_assume(__isShared(v.Hv));
  __builtin_assume(__isShared(v.Wv));
  __builtin_assume(__isShared(v.K2));
  __builtin_assume(__isShared(v.H2));
  __builtin_assume(__isShared(v.W2));
  __builtin_assume(__isShared(v.av));
  for (int ib = lane; ib < npad; ib += 32 * NR) {
    double2 wr[NR], kr[NR];
    double ar[NR], w2r[NR], k2r[NR];
    double* prow[NR];
    int jmax[NR], ii[NR];
#pragma unroll
    for (int m = 0; m < NR; m++) {
      const int i = ib + 32 * m;
      const bool valid = i < npad;
      ii[m] = valid ? i : 0;
      wr[m] = v.Wv[ii[m]];
      kr[m] = v.Kv[ii[m]];
      ar[m] = v.av[ii[m]];
      w2r[m] = (RD == 3) ? v.W2[ii[m]] : 0.0;
      k2r[m] = (RD == 3) ? v.K2[ii[m]] : 0.0;
      prow[m] = (ii[m] < 16) ? Pbb + ii[m] * PBS : Pfb + (ii[m] - 16) * 17;
      jmax[m] = valid ? (ii[m] < 16 ? ii[m] : 15) : -1;
    }
    // the three rows go to registers first (immediate-offset loads, no per-column address arithmetic), then 16 fully
    // unrolled columns: column vectors are warp broadcasts, the three rows give independent FMA chains
    double pr[NR][16];
#pragma unroll
    for (int m = 0; m < NR; m++)
#pragma unroll
      for (int j = 0; j < 16; j++) pr[m][j] = prow[m][j];
#pragma unroll
    for (int j = 0; j < 16; j++) {
      const double2 kj = v.Kv[j], hj = v.Hv[j];
      const double aj = v.av[j];
      const double k2j = (RD == 3) ? v.K2[j] : 0.0, h2j = (RD == 3) ? v.H2[j] : 0.0;
#pragma unroll
      for (int m = 0; m < NR; m++) {
        const double d = rank_delta<RD>(wr[m], kr[m], w2r[m], k2r[m], kj, hj, k2j, h2j);
        const double u = fma(-ar[m], aj, 1.0);
        pr[m][j] = fma(d, u, pr[m][j]);
      }
    }
#pragma unroll
    for (int m = 0; m < NR; m++)
#pragma unroll
      for (int j = 0; j < 16; j++) {
        const bool ok = j <= jmax[m];
        if (ok) prow[m][j] = pr[m][j];
        if (ok && ii[m] < 16) Pbb[j * PBS + ii[m]] = pr[m][j];
      }
  }
}

template <int TW, int RD>
__device__ __forceinline__ int tile_update_core(TCtx<TW>& c, const KParams& p, const viekf_event& E, int n, int len, const HReg& H,
                                                double (&Pt)[36]) {
  constexpr int G = 32 * (TW + 1);
  constexpr int MC = 6;
  int* cols = c.sc + SC_COUNT;
  const int lenp = (len + 1) & ~1;
  const int npad = 16 + 3 * lenp;
  TUpdVec v(c.upd, 16 + 3 * ((p.N + 1) & ~1));
  PH(c, 8)
  const bool featcols = (H.m > 0) && (H.col[0] >= 16);  // every H touches either base columns or the columns of one feature
  // ---- P H^T ----
  if (!featcols) {
    for (int i = c.tid; i < npad; i += G) {
      double s[3] = {0, 0, 0};
#pragma unroll
      for (int k = 0; k < MC; k++)
        if (k < H.m) {
          const double pv = (i < 16) ? c.Pbb[i * PBS + H.col[k]] : c.Pfb[(i - 16) * 17 + H.col[k]];
#pragma unroll
          for (int a = 0; a < RD; a++) s[a] = fma(pv, H.h[a * MC + k], s[a]);
        }
      v.Hv[i] = make_double2(s[0], s[1]);
      if (RD == 3) v.H2[i] = s[2];
    }
  } else {
    const int fi = (H.col[0] - 16) / 3, tb = fi >> 1, sb = fi & 1;
    // effective 3-column Jacobian of the feature block: he[a][kk] = sum over the k with local column kk
    double he[RD][3];
#pragma unroll
    for (int a = 0; a < RD; a++)
#pragma unroll
      for (int kk = 0; kk < 3; kk++) he[a][kk] = 0.0;
#pragma unroll
    for (int k = 0; k < 2; k++)  // feature-column models have at most two live columns (FEAT, QZETA: 2; DEPTH, INV_DEPTH: 1)
      if (k < H.m) {
        const int lk = H.col[k] - 16 - 3 * fi;
#pragma unroll
        for (int a = 0; a < RD; a++)
#pragma unroll
          for (int kk = 0; kk < 3; kk++)
            if (lk == kk) he[a][kk] += H.h[a * MC + k];
      }
    if (c.basew && c.lane < 16) {
      const int i = c.lane;
      double s[3] = {0, 0, 0};
#pragma unroll
      for (int kk = 0; kk < 3; kk++) {
        const double pv = c.Pfb[(3 * fi + kk) * 17 + i];
#pragma unroll
        for (int a = 0; a < RD; a++) s[a] = fma(pv, he[a][kk], s[a]);
      }
      v.Hv[i] = make_double2(s[0], s[1]);
      if (RD == 3) v.H2[i] = s[2];
    }
    if (c.tile && 2 * c.ti < len) {
      if (c.tj == tb) {  // rows of this tile against the columns of feature fi
#pragma unroll
        for (int r = 0; r < 6; r++) {
          double s[3] = {0, 0, 0};
#pragma unroll
          for (int kk = 0; kk < 3; kk++) {
            const double pv = sb ? PT(r, 3 + kk) : PT(r, kk);
#pragma unroll
            for (int a = 0; a < RD; a++) s[a] = fma(pv, he[a][kk], s[a]);
          }
          v.Hv[16 + 6 * c.ti + r] = make_double2(s[0], s[1]);
          if (RD == 3) v.H2[16 + 6 * c.ti + r] = s[2];
        }
      } else if (c.ti == tb) {  // columns of this tile against the rows of feature fi (symmetry)
#pragma unroll
        for (int cc = 0; cc < 6; cc++) {
          double s[3] = {0, 0, 0};
#pragma unroll
          for (int kk = 0; kk < 3; kk++) {
            const double pv = sb ? PT(3 + kk, cc) : PT(kk, cc);
#pragma unroll
            for (int a = 0; a < RD; a++) s[a] = fma(pv, he[a][kk], s[a]);
          }
          v.Hv[16 + 6 * c.tj + cc] = make_double2(s[0], s[1]);
          if (RD == 3) v.H2[16 + 6 * c.tj + cc] = s[2];
        }
      }
    }
  }
  if (c.tid == 0) cols[8] = H.hnan ? 1 : 0;
  PHS(c, 9)
  // ---- S = H P H^T + R, S^-1, Mahalanobis gate: every thread evaluates the same registers ----
  double S[3][3], Si[3][3];
  {
    double ph[MC][3];
#pragma unroll
    for (int k = 0; k < MC; k++) {
      const int ck = (k < H.m) ? H.col[k] : 0;
      const double2 t = v.Hv[ck];
      ph[k][0] = t.x;
      ph[k][1] = t.y;
      ph[k][2] = (RD == 3) ? v.H2[ck] : 0.0;
    }
#pragma unroll
    for (int a = 0; a < RD; a++)
#pragma unroll
      for (int b = 0; b < RD; b++) {
        double s = 0;
#pragma unroll
        for (int k = 0; k < MC; k++)
          if (k < H.m) s += H.h[a * MC + k] * ph[k][b];
        S[a][b] = s + E.R[a + b * RD];
      }
  }
  const double mahal = inverse_and_mahal<RD>(S, Si, H.res);
  if (mahal > 9.0) {
    if (c.tid == 0) c.sc[SC_NGATED] += 1;
    return VIEKF_MEAS_GATED;  // early return: no fix_depth (vi_ekf_meas.cpp:238)
  }
  // ---- K = P H^T S^-1, W = K S - P H^T, state correction lambda o (K r) ----
  bool bad = false;
  for (int i = c.tid; i < npad; i += G) {
    const double2 h01 = v.Hv[i];
    double ph[3] = {h01.x, h01.y, 0.0};
    if (RD == 3) ph[2] = v.H2[i];
    double k3[3] = {0, 0, 0}, w3[3] = {0, 0, 0};
#pragma unroll
    for (int b = 0; b < RD; b++) {
      double s = 0;
#pragma unroll
      for (int a = 0; a < RD; a++) s += ph[a] * Si[a][b];
      k3[b] = s;
      bad |= (s != s);
    }
    double d = 0;
#pragma unroll
    for (int a = 0; a < RD; a++) d += k3[a] * H.res[a];
#pragma unroll
    for (int a = 0; a < RD; a++) {
      double s = 0;
#pragma unroll
      for (int b = 0; b < RD; b++) s += k3[b] * S[b][a];
      w3[a] = s - ph[a];
    }
    v.Kv[i] = make_double2(k3[0], k3[1]);
    v.Wv[i] = make_double2(w3[0], w3[1]);
    if (RD == 3) {
      v.K2[i] = k3[2];
      v.W2[i] = w3[2];
    }
    const double li = c.lam[i];
    v.av[i] = p.use_partial ? 1.0 - li : 0.0;
    c.dx[i] = p.use_partial ? li * d : d;
  }
  if (bad) atomicOr((unsigned*)&cols[8], 1u);
  PHS(c, 10)
  if (!cols[8]) {
    // base part on the spare lane of the last tile warp (the base warp has the longer rank update); warps without a share of the
    // boxplus skip the out-of-line call altogether
    if (c.tid == 32 * TW - 1 || c.tid < len) boxplus_inplace(c, c.dx, 1.0, len, 32 * TW - 1);
    PH(c, 12)
    if (c.tile && 2 * c.ti < len) {
      const int R = 16 + 6 * c.ti, C = 16 + 6 * c.tj;
      // three columns at a time: 72 tile registers + 3 column vectors + 1 row vector stay inside the register budget
#pragma unroll
      for (int ch = 0; ch < 2; ch++) {
        double2 kc[3], hc[3];
        double ac[3], k2c[3], h2c[3];
#pragma unroll
        for (int cc = 0; cc < 3; cc++) {
          kc[cc] = v.Kv[C + 3 * ch + cc];
          hc[cc] = v.Hv[C + 3 * ch + cc];
          ac[cc] = v.av[C + 3 * ch + cc];
          k2c[cc] = (RD == 3) ? v.K2[C + 3 * ch + cc] : 0.0;
          h2c[cc] = (RD == 3) ? v.H2[C + 3 * ch + cc] : 0.0;
        }
#pragma unroll
        for (int r = 0; r < 6; r++) {
          const double2 wr = v.Wv[R + r], kr = v.Kv[R + r];
          const double ar = v.av[R + r];
          const double w2r = (RD == 3) ? v.W2[R + r] : 0.0, k2r = (RD == 3) ? v.K2[R + r] : 0.0;
#pragma unroll
          for (int cc = 0; cc < 3; cc++) {
            const double d = rank_delta<RD>(wr, kr, w2r, k2r, kc[cc], hc[cc], k2c[cc], h2c[cc]);
            const double u = fma(-ar, ac[cc], 1.0);
            PT(r, 3 * ch + cc) = fma(d, u, PT(r, 3 * ch + cc));
          }
        }
      }
      tile_mirror(c, Pt);
    }
    PH(c, 13)
    resolve_next(c, p, 32);
    if (c.basew) {
      base_rank<RD>(v, c.Pbb, c.Pfb, npad, c.lane);
      PH(c, 14)
      tile_dead(Pt);
    }
  } else if (c.tid == 0) {
    c.sc[SC_FLAGS] |= VIEKF_FLAG_NAN_GAIN;
    c.sc[SC_NNAN] += 1;
  }
  if (c.tid < len) tile_fix_x(c, p, c.tid);
  PH(c, 11)
  return VIEKF_MEAS_SUCCESS;
}

// ------------------------------------------------------------------------------------------------------------------
// feature management
// ------------------------------------------------------------------------------------------------------------------
// init_feature (vi_ekf_feat.cpp:6-47).  Returns (uniformly) whether a slot was available.
template <int TW>
__device__ __forceinline__ bool tile_init_feature(TCtx<TW>& c, const KParams& p, double l0, double l1, double depth, double (&Pt)[36]) {
  constexpr int G = 32 * (TW + 1);
  const int len = c.sc[SC_LEN];
  if (len >= p.N) return false;
  c.sync();
  if (c.tid == 0) {
    const double lc0 = l0 - p.cc[0], lc1 = l1 - p.cc[1];
    V3 z = mk3(lc0, lc1 * (p.fl[1] / p.fl[0]), p.fl[0]);
    const double nz = norm(z);
    z = mk3(z.x / nz, z.y / nz, z.z / nz);
    Q4 q = from_two_unit_vectors(mk3(0, 0, 1), z);
    double d = depth;
    if (d != d) d = 2.0 * p.min_depth;
    double* xf = c.x + 17 + 5 * len;
    stq(xf, q);
    xf[4] = 1.0 / d;
    c.ids[len] = c.sc[SC_NEXT];
    c.sc[SC_NEXT] += 1;
    c.sc[SC_LEN] = len + 1;
  }
  // covariance: rows / columns of the new feature (and of the padding feature that shares its tile row) zero, P0_feat on the diagonal
  const int tb = len >> 1, sb = len & 1;
  if (c.tile && c.ti == tb) {
#pragma unroll
    for (int r = 0; r < 6; r++)
#pragma unroll
      for (int cc = 0; cc < 6; cc++)
        if (sb == 0 || r >= 3 || (c.tj == tb && cc >= 3)) PT(r, cc) = 0.0;
    if (c.tj == tb) {
#pragma unroll
      for (int k = 0; k < 3; k++) {
        if (sb)
          PT(3 + k, 3 + k) = p.P0f[k];
        else
          PT(k, k) = p.P0f[k];
      }
    }
  }
  {
    const int r0 = 3 * len, r1 = 3 * ((len + 2) & ~1);  // through the end of the tile row
    for (int e = c.tid; e < (r1 - r0) * 16; e += G) c.Pfb[(r0 + (e >> 4)) * 17 + (e & 15)] = 0.0;
  }
  c.sync();
  return true;
}

// Order-preserving removal of features: new feature f <- old feature smap[f], f < newlen (clear_feature, vi_ekf_feat.cpp:50-73,
// applied once for all removals).  Tiles and P_fb go through the filter's packed P in global memory; x and ids through registers.  Precondition: smap[] published by a barrier.
template <int TW>
__device__ __forceinline__ void tile_compact(TCtx<TW>& c, const KParams& p, int len, int newlen, double (&Pt)[36]) {
  constexpr int G = 32 * (TW + 1);
  double* gP = p.P + c.b * p.ps;
  tile_store(c, gP, Pt, len);
  double xf[5];
  int idv = 0;
  if (c.tid < newlen) {
    const int src = c.smap[c.tid];
#pragma unroll
    for (int k = 0; k < 5; k++) xf[k] = c.x[17 + 5 * src + k];
    idv = c.ids[src];
  }
  for (int e = c.tid; e < 3 * len * 16; e += G) {
    const int fr = e >> 4, cc = e & 15;
    gP[plo(16 + fr, cc)] = c.Pfb[fr * 17 + cc];
  }
  c.sync();
  const int newp = (newlen + 1) & ~1;
  for (int e = c.tid; e < 3 * newp * 16; e += G) {
    const int fr = e >> 4, cc = e & 15, f = fr / 3;
    c.Pfb[fr * 17 + cc] = (f < newlen) ? __ldcg(gP + plo(16 + 3 * c.smap[f] + (fr - 3 * f), cc)) : 0.0;
  }
  if (c.tid < newlen) {
#pragma unroll
    for (int k = 0; k < 5; k++) c.x[17 + 5 * c.tid + k] = xf[k];
    c.ids[c.tid] = idv;
  }
  tile_load(c, gP, Pt, newlen, true);
  if (c.tid == 0) c.sc[SC_LEN] = newlen;
  c.sync();
}

template <int TW>
__device__ __forceinline__ void tile_clear_feature(TCtx<TW>& c, const KParams& p, int local, double (&Pt)[36]) {
  constexpr int G = 32 * (TW + 1);
  const int len = c.sc[SC_LEN];
  if (local < 0 || local >= len) return;
  c.sync();
  for (int f = c.tid; f < len - 1; f += G) c.smap[f] = f + (f >= local ? 1 : 0);
  c.sync();
  tile_compact(c, p, len, len - 1, Pt);
}

// packed lower triangle of P_bb into the T buffer (idle outside propagate): the statistics and the keyframe-reset pose code
// read the base block through c.P
template <int TW>
__device__ __forceinline__ void tile_pack_base(TCtx<TW>& c) {
  constexpr int G = 32 * (TW + 1);
  for (int e = c.tid; e < 256; e += G) {
    const int r = e >> 4, cc = e & 15;
    if (cc <= r) c.T[plo(r, cc)] = c.Pbb[r * PBS + cc];
  }
  c.P = c.T;
}

// keyframe_reset (vi_ekf_kfr.cpp:56-157): P = N P N^T, N = I except N(POS,POS) = 0 and the 3x3 attitude block Nb.
// P_ff is untouched; P_fb gets N^T from the right, P_bb both sides (base warp), then its upper half is re-mirrored.
template <int TW>
__device__ __forceinline__ void tile_keyframe_reset(TCtx<TW>& c, const KParams& p) {
  constexpr int G = 32 * (TW + 1);
  const int len = c.sc[SC_LEN];
  c.sync();
  tile_pack_base(c);
  c.sync();
  double* Nb = c.sm + SM_HC;  // 9 doubles, row-major
  if (c.tid == 0) kfreset_pose(c, p, Nb);
  c.sync();
  for (int fr = c.tid; fr < 3 * len; fr += G) {
    double* row = c.Pfb + fr * 17;
    double v[3];
#pragma unroll
    for (int a = 0; a < 3; a++) {
      row[a] = 0.0;
      v[a] = row[6 + a];
    }
#pragma unroll
    for (int a = 0; a < 3; a++) row[6 + a] = Nb[a * 3] * v[0] + Nb[a * 3 + 1] * v[1] + Nb[a * 3 + 2] * v[2];
  }
  if (c.basew) {
    const int l = c.lane;
    if (l < 16) {  // right factor N^T, one row per lane
      double* row = c.Pbb + l * PBS;
      double v[3];
#pragma unroll
      for (int a = 0; a < 3; a++) {
        row[a] = 0.0;
        v[a] = row[6 + a];
      }
#pragma unroll
      for (int a = 0; a < 3; a++) row[6 + a] = Nb[a * 3] * v[0] + Nb[a * 3 + 1] * v[1] + Nb[a * 3 + 2] * v[2];
    }
    __syncwarp();
    if (l < 16) {  // left factor N, one column per lane
      double v[3];
#pragma unroll
      for (int a = 0; a < 3; a++) {
        c.Pbb[a * PBS + l] = 0.0;
        v[a] = c.Pbb[(6 + a) * PBS + l];
      }
#pragma unroll
      for (int a = 0; a < 3; a++) c.Pbb[(6 + a) * PBS + l] = Nb[a * 3] * v[0] + Nb[a * 3 + 1] * v[1] + Nb[a * 3 + 2] * v[2];
    }
    __syncwarp();
    if (l < 16)
      for (int cc = 0; cc < l; cc++) c.Pbb[cc * PBS + l] = c.Pbb[l * PBS + cc];
  }
  c.sync();
}

// keep_only_features (vi_ekf_feat.cpp:81-142)
template <int TW>
__device__ __forceinline__ void tile_keep_only(TCtx<TW>& c, const KParams& p, const int32_t* feats, int nf, double (&Pt)[36]) {
  constexpr int G = 32 * (TW + 1);
  c.sync();
  int* keepf = (int*)c.upd;  // [N] flags, update workspace is idle here
  if (c.tid == 0) c.sc[SC_TMP1] = 0;
  c.sync();
  const int len = c.sc[SC_LEN];
  const int nkf = c.sc[SC_NKF];
  {
    const int32_t* kf = p.kf_ids + c.b * KF_CAP;
    for (int l = c.tid; l < len; l += G) {
      const int id = c.ids[l];
      bool keep = false;
      for (int i = 0; i < nf; i++) keep |= (feats[i] == id);
      if (keep && p.use_kfreset) {
        bool ov = false;
        for (int j = 0; j < nkf; j++) ov |= (kf[j] == id);
        if (ov) atomicAdd(&c.sc[SC_TMP1], 1);
      }
      keepf[l] = keep ? 1 : 0;
    }
  }
  c.sync();
  if (c.tid == 0) {
    int m = 0;
    for (int l = 0; l < len; l++)
      if (keepf[l]) c.smap[m++] = l;
    c.sc[SC_TMP0] = m;
  }
  c.sync();
  const int newlen = c.sc[SC_TMP0];
  const int overlap = c.sc[SC_TMP1];
  if (newlen < len) tile_compact(c, p, len, newlen, Pt);
  const bool reset = p.use_kfreset && nkf > 0 && ((double)overlap / (double)nkf < p.kf_thresh);
  if (reset) tile_keyframe_reset(c, p);
  c.sync();  // every thread has taken the (uniform) reset decision from the old list size before thread 0 replaces it
  if (p.use_kfreset && (reset || nkf == 0)) {
    int32_t* kf = p.kf_ids + c.b * KF_CAP;
    const int m = nf < KF_CAP ? nf : KF_CAP;
    for (int i = c.tid; i < m; i += G) kf[i] = feats[i];
    if (c.tid == 0) c.sc[SC_NKF] = m;
  }
  c.sync();
}

// NaNsInTheHouse (vi_ekf_error.cpp:6-19), this thread's share: live x, P_bb, live rows of P_fb, and the owner's tile (entries
// of a padding feature are exact zeros)
template <int TW>
__device__ __forceinline__ bool tile_nan_scan(const TCtx<TW>& c, const double (&Pt)[36]) {
  constexpr int G = 32 * (TW + 1);
  const int len = c.sc[SC_LEN];
  bool bad = false;
  for (int i = c.tid; i < 17 + 5 * len; i += G) bad |= (c.x[i] != c.x[i]);
  for (int e = c.tid; e < 256; e += G) {
    const double v = c.Pbb[(e >> 4) * PBS + (e & 15)];
    bad |= (v != v);
  }
  for (int e = c.tid; e < 3 * len * 16; e += G) {
    const double v = c.Pfb[(e >> 4) * 17 + (e & 15)];
    bad |= (v != v);
  }
  if (c.tile && 2 * c.ti < len) {
#pragma unroll
    for (int i = 0; i < 36; i++) bad |= (Pt[i] != Pt[i]);
  }
  return bad;
}

template <int TW>
struct TileCore {
  TCtx<TW>& c;
  const KParams& p;
  double (&Pt)[36];
  template <int RD>
  __device__ __forceinline__ int run(const viekf_event& E, int n, int len, const HReg& H) {
    return tile_update_core<TW, RD>(c, p, E, n, len, H, Pt);
  }
  __device__ __forceinline__ bool init_feature(double l0, double l1, double depth) { return tile_init_feature(c, p, l0, l1, depth, Pt); }
  __device__ __forceinline__ void fix_depth(int len) {
    if (c.tid < len) tile_fix_x(c, p, c.tid);
  }
};

// ------------------------------------------------------------------------------------------------------------------
// the persistent tape executor, register-tile variant: a group of 32 (TW + 1) threads per filter, FPC groups per CTA.
// The groups of a CTA synchronise among themselves with named barriers inside an event and all together at the end of
// every event: the FPC filters stay within one event of each other, so the warps of an SM run the same few thousand
// instructions at any time (the executor is instruction-fetch bound when its CTAs drift apart over the ~0.5 MB of code).
// ------------------------------------------------------------------------------------------------------------------
template <int TW, int FPC>
__global__ void __launch_bounds__(32 * (TW + 1) * FPC, (32 * (TW + 1) * FPC <= 128 ? 4 : (32 * (TW + 1) * FPC <= 256 ? 2 : 1))) viekf_tile_kernel(const __grid_constant__ KParams p) {
  extern __shared__ __align__(16) double smem[];
  constexpr int G = 32 * (TW + 1);
  const TLayout L = make_tlayout(p.N);
  TCtx<TW> c;
  c.gi = threadIdx.x / G;
#ifdef VIEKF_ROLE_ROTATE
  // warp w of the CTA issues on SM sub-partition w % 4: with G = 128 the base warps of all filters of a CTA would share one
  // sub-partition (and its 16 FP64 lanes); rotating the roles by the filter index spreads them
  c.tid = (int)((threadIdx.x % G) + 32 * c.gi) % G;
#else
  c.tid = threadIdx.x % G;
#endif
  c.lane = c.tid & 31;
  c.basew = (c.tid >> 5) == TW;
  {
    const int t = c.tid;
    int ti = (int)((sqrtf(8.0f * (float)t + 1.0f) - 1.0f) * 0.5f);
    while (((ti + 1) * (ti + 2) >> 1) <= t) ti++;
    while (((ti * (ti + 1)) >> 1) > t) ti--;
    c.ti = ti;
    c.tj = t - ((ti * (ti + 1)) >> 1);
    c.tile = !c.basew && ti < L.TR;
  }
  double* base = smem + (size_t)c.gi * L.total;
  c.x = base + L.x; c.P = base + L.T; c.lam = base + L.lam; c.dx = base + L.dx; c.Abd = base + L.Abd; c.AbT = base + L.AbT; c.T = base + L.T;
  c.Ac = base + L.Ac; c.Gb = base + L.Gb; c.Afd = base + L.Afd; c.Gd = base + L.Gd; c.Y = base + L.Y; c.upd = base + L.upd;
  c.Pbb = base + L.Pbb; c.Pfb = base + L.Pfb;
  c.sm = base + L.small;
  c.ids = (int*)(base + L.ids);
  c.rm = c.ids + p.N;
  c.sc = c.rm + p.N;
  c.smap = (int*)(base + L.smap);
  c.fixk = (int*)(base + L.fixk);
  c.fixv = base + L.fixv;
  c.zvr = base + L.zvr;
  c.tcur = c.sm + SM_T;
  double* evb = base + L.evb;
  constexpr int EVW = (int)(sizeof(viekf_event) / sizeof(double));
  static_assert(sizeof(viekf_event) % sizeof(double) == 0 && EVW <= 32, "viekf_event must fit one warp of doubles");
  constexpr int EVT0 = (G >= 96) ? 64 : 32;  // first thread of the warp that stages the next event record (warp 2; the base warp when TW = 1)
  const int nmaxp = 16 + 6 * L.TR;
  for (int i = c.tid; i < nmaxp; i += G) c.lam[i] = (i < 16) ? p.lam[i] : p.lamf[(i - 16) % 3];
  for (int i = c.tid; i < 2 * L.TR; i += G) c.fixk[i] = 0;
  double Pt[36];
  PH_INIT(c)

  for (int64_t b0 = (int64_t)blockIdx.x * FPC; b0 < p.B; b0 += (int64_t)gridDim.x * FPC) {
    const bool active = b0 + c.gi < p.B;  // a group without a filter still takes part in the CTA-wide event barriers
    const int64_t b = active ? b0 + c.gi : 0;
    c.b = b;
    const int64_t ginst = p.inst_off + b;
    c.ginst = ginst;
    FilterMeta& M = p.meta[b];
    double* gP = p.P + b * p.ps;
    // ---- load the filter: meta, ids, live part of x; P_bb / P_fb into shared memory, P_ff into the register tiles ----
    c.sync();
    const int len0 = M.len;
    if (c.tid == 0) {
      c.sc[SC_LEN] = len0; c.sc[SC_NEXT] = M.next_id; c.sc[SC_STARTED] = M.started; c.sc[SC_FLAGS] = (int)M.flags;
      c.sc[SC_NGATED] = (int)M.n_gated; c.sc[SC_NFIX] = (int)M.n_fix; c.sc[SC_NNAN] = (int)M.n_nan_gain; c.sc[SC_NKF] = M.n_kf;
      *c.tcur = M.t;
      for (int k = 0; k < 6; k++) c.sm[SM_UB + k] = M.ub[k];
    }
    for (int i = c.tid; i < len0; i += G) c.ids[i] = p.ids[b * (p.N > 0 ? p.N : 1) + i];
    {
      const double* gx = p.x + b * p.xs;
      const int nx = 17 + 5 * len0;
      for (int i = c.tid; i < nx; i += G) c.x[i] = gx[i];
      for (int e = c.tid; e < 256; e += G) {
        const int r = e >> 4, cc = e & 15;
        c.Pbb[r * PBS + cc] = gP[psym(r, cc)];
      }
      const int nfr0 = 3 * ((len0 + 1) & ~1);
      for (int e = c.tid; e < nfr0 * 16; e += G) {
        const int fr = e >> 4, cc = e & 15;
        c.Pfb[fr * 17 + cc] = (fr < 3 * len0) ? gP[plo(16 + fr, cc)] : 0.0;
      }
      tile_load(c, gP, Pt, len0, false);
    }
    if (p.ev_count > 0 && c.tid < 8) resolve_values(p, p.ev_first, b, ginst, c.tid, c.zvr);
    if (p.ev_count > 0 && c.tid >= EVT0 && c.tid < EVT0 + EVW) evb[c.tid - EVT0] = ((const double*)(p.ev + p.ev_first))[c.tid - EVT0];
    PHS(c, 0)
    // ---- run the tape slice ----
    for (int64_t e = p.ev_first; e < p.ev_first + p.ev_count; e++) {
      // the event record is read from shared memory; the next one is fetched now (one double per lane of warp 2) and
      // staged just before the barrier that ends this event
      const viekf_event& E = *(const viekf_event*)(evb + 32 * (int)((e - p.ev_first) & 1));
      tile_apply_fix(c, p, Pt);  // depth corrections recorded by the previous event
      double ev_next = 0.0;
      const bool stage = (e + 1 < p.ev_first + p.ev_count) && c.tid >= EVT0 && c.tid < EVT0 + EVW;
      if (stage) ev_next = __ldg((const double*)(p.ev + e + 1) + (c.tid - EVT0));
      const int kind = E.kind;
      c.zv = c.zvr + 8 * (int)((e - p.ev_first) & 7);
      c.next_e = (((e - p.ev_first) & 3) == 0 && e + 1 < p.ev_first + p.ev_count) ? e + 1 : -1;
      c.resolved = false;
      int result = VIEKF_MEAS_NONE;
      if (active) {
      if (kind != VIEKF_EV_PROP && kind != VIEKF_EV_UPD) resolve_next(c, p, 32);
      switch (kind) {
        case VIEKF_EV_PROP:
          tile_propagate(c, p, E, Pt);
          break;
        case VIEKF_EV_UPD: {
          TileCore<TW> core{c, p, Pt};
          result = ev_update(c, p, E, core);
          break;
        }
        case VIEKF_EV_INIT:
          result = tile_init_feature(c, p, c.zv[0], c.zv[1], E.z[4], Pt) ? 1 : 0;
          break;
        case VIEKF_EV_CLEAR:
          tile_clear_feature(c, p, find_feature(c, E.id), Pt);
          break;
        case VIEKF_EV_KEEP:
          tile_keep_only(c, p, p.id_pool + E.ids_off, E.n_ids, Pt);
          break;
        case VIEKF_EV_KFRESET:
          tile_keyframe_reset(c, p);
          break;
        case VIEKF_EV_STAT:
          c.sync();
          tile_pack_base(c);
          c.sync();
          ev_stat(c, p, E, ginst, [&]() { return tile_nan_scan(c, Pt); });
          break;
        default:
          break;
      }
      if (!c.resolved) resolve_next(c, p, 32);  // the handler returned before its deferred slot (uniform)
      }
      if (stage) evb[32 * (int)((e + 1 - p.ev_first) & 1) + (c.tid - EVT0)] = ev_next;
      PH(c, 16 + kind)
      // the filters of a CTA re-align every LOCKSTEP events (all start the next event together); in between only the group
      // synchronises, so a filter that finishes an event early (gated update, fewer live features) is not held back each time
      if (FPC > 1 && ((int)(e - p.ev_first) & (LOCKSTEP - 1)) == LOCKSTEP - 1)
        __syncthreads();
      else
        c.sync();
      PH(c, 16 + kind + 32)
      if (p.results && active && c.tid == 0 && b >= p.res_first && b < p.res_first + p.res_count)
        p.results[(b - p.res_first) * p.n_events + e] = (uint8_t)result;
    }
    // ---- store the filter ----
    c.sync();
    if (!active) continue;
    tile_apply_fix(c, p, Pt);
    latch_nan_state(c, tile_nan_scan(c, Pt));  // NaNsInTheHouse at the end of the slice
    const int len1 = c.sc[SC_LEN];
    if (c.tid == 0) {
      M.len = len1; M.next_id = c.sc[SC_NEXT]; M.started = c.sc[SC_STARTED]; M.flags = (uint32_t)c.sc[SC_FLAGS];
      M.n_gated = (uint32_t)c.sc[SC_NGATED]; M.n_fix = (uint32_t)c.sc[SC_NFIX]; M.n_nan_gain = (uint32_t)c.sc[SC_NNAN]; M.n_kf = c.sc[SC_NKF];
      M.t = *c.tcur;
      for (int k = 0; k < 6; k++) M.ub[k] = c.sm[SM_UB + k];
    }
    for (int i = c.tid; i < len1; i += G) p.ids[b * (p.N > 0 ? p.N : 1) + i] = c.ids[i];
    {
      double* gx = p.x + b * p.xs;
      const int nx = 17 + 5 * len1;
      for (int i = c.tid; i < nx; i += G) gx[i] = c.x[i];
      for (int e = c.tid; e < 256; e += G) {
        const int r = e >> 4, cc = e & 15;
        if (cc <= r) gP[plo(r, cc)] = c.Pbb[r * PBS + cc];
      }
      for (int e = c.tid; e < 3 * len1 * 16; e += G) {
        const int fr = e >> 4, cc = e & 15;
        gP[plo(16 + fr, cc)] = c.Pfb[fr * 17 + cc];
      }
      tile_store(c, gP, Pt, len1);
    }
    PH(c, 15)
  }
}

#ifdef VIEKF_PHASE_TIMING
extern "C" int viekf_debug_phase_clocks(unsigned long long* clk, unsigned long long* cnt, int reset) {
  cudaDeviceSynchronize();
  cudaMemcpyFromSymbol(clk, g_phase_clk, sizeof(g_phase_clk));
  cudaMemcpyFromSymbol(cnt, g_phase_cnt, sizeof(g_phase_cnt));
  if (reset) {
    static unsigned long long z[4 * 64];
    cudaMemcpyToSymbol(g_phase_clk, z, sizeof(z));
    cudaMemcpyToSymbol(g_phase_cnt, z, sizeof(z));
  }
  return 0;
}
#endif

// 28 tiles (N <= 14) fit one tile warp, 91 tiles (N <= 26) three, 171 tiles (N <= 36) six.  VIEKF_TILE_TW1=0 keeps N <= 14 on three
// tile warps (A/B switch for the one-warp mapping).
int tile_warps_for(int N) {
  if (N >= 7 && N <= 14) {
    const char* s = getenv("VIEKF_TILE_TW1");
    return (s && atoi(s) == 0) ? 3 : 1;
  }
  return (N >= 15 && N <= 26) ? 3 : ((N >= 27 && N <= 36) ? 6 : 0);
}

template <int TW, int FPC>
static cudaError_t launch_tile_fpc(const KParams& p, int sm_count, cudaStream_t stream) {
  constexpr int CTA = 32 * (TW + 1) * FPC;
  size_t smem = tile_smem_bytes(p.N) * FPC;
  if (const char* s = getenv("VIEKF_TILE_MAXOCC")) {  // tuning aid: pad the shared-memory request so that at most k CTAs fit on an SM
    const int k = atoi(s);
    if (k >= 1 && k <= 8) {
      const size_t pad = (size_t)(227 * 1024) / k - 2048;
      if (pad > smem) smem = pad;
    }
  }
  cudaError_t e = cudaFuncSetAttribute(viekf_tile_kernel<TW, FPC>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) return e;
  e = cudaFuncSetAttribute(viekf_tile_kernel<TW, FPC>, cudaFuncAttributePreferredSharedMemoryCarveout, (int)cudaSharedmemCarveoutMaxShared);
  if (e != cudaSuccess) return e;
  int occ = 0;
  e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, viekf_tile_kernel<TW, FPC>, CTA, smem);
  if (e != cudaSuccess) return e;
  if (occ < 1) return cudaErrorInvalidConfiguration;
  int64_t grid = (int64_t)sm_count * occ;  // persistent: a whole number of waves of the SM count
  const int64_t need = (p.B + FPC - 1) / FPC;
  if (need < grid) grid = need;
  if (grid < 1) grid = 1;
  viekf_tile_kernel<TW, FPC><<<(unsigned)grid, CTA, smem, stream>>>(p);
  return cudaGetLastError();
}

cudaError_t launch_tile(const KParams& p, int sm_count, cudaStream_t stream) {
  const int tw = tile_warps_for(p.N);
  if (tw == 0) return cudaErrorInvalidValue;
  // filters per CTA in event lock-step.  TW = 3: 3 (384 threads, 168 registers each: the 72-register tile plus a schedule with
  // loads in flight; 4 x 128 registers spills on the hot path).  TW = 6: 2 (448 threads, 144 registers).  Small batches spread
  // over the SMs first.
  // TW = 1 (N <= 14, config #3 at N = 8): 6 filters of 64 threads per CTA.
  const int full = (tw == 1) ? 6 : ((tw == 3) ? 3 : 2);
  int fpc = (p.B < full * (int64_t)sm_count) ? 1 : full;
  if (const char* s = getenv("VIEKF_TILE_FPC")) fpc = atoi(s);  // test / tuning override; a value without an instantiation is an error
#ifdef VIEKF_DEV_FAST  // development build: only the configs[3] instantiations (each one costs ~25 s of ptxas)
  if (tw == 3 && fpc == 1) return launch_tile_fpc<3, 1>(p, sm_count, stream);
  if (tw == 3 && fpc == 3) return launch_tile_fpc<3, 3>(p, sm_count, stream);
  return cudaErrorInvalidValue;
#else
  if (tw == 1) {
    switch (fpc) {
      case 1: return launch_tile_fpc<1, 1>(p, sm_count, stream);
      case 2: return launch_tile_fpc<1, 2>(p, sm_count, stream);
      case 4: return launch_tile_fpc<1, 4>(p, sm_count, stream);
      case 6: return launch_tile_fpc<1, 6>(p, sm_count, stream);
      default: return cudaErrorInvalidValue;
    }
  }
  if (tw == 3) {
    switch (fpc) {
      case 1: return launch_tile_fpc<3, 1>(p, sm_count, stream);
      case 2: return launch_tile_fpc<3, 2>(p, sm_count, stream);
      case 3: return launch_tile_fpc<3, 3>(p, sm_count, stream);
      case 4: return launch_tile_fpc<3, 4>(p, sm_count, stream);
      default: return cudaErrorInvalidValue;
    }
  }
  switch (fpc) {
    case 1: return launch_tile_fpc<6, 1>(p, sm_count, stream);
    case 2: return launch_tile_fpc<6, 2>(p, sm_count, stream);
    default: return cudaErrorInvalidValue;
  }
#endif
}

}  // namespace vk
