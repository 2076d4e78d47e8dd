// viekf_tile.cu -- register-tile tape executor of the VI-EKF batch engine (sm_100a, FP64).
//
// Same contract as viekf_kernel.cu (one thread group owns one filter for a whole tape slice; events in order; HBM layout of
// x / packed P unchanged), different placement of the covariance on chip and a warp-specialised software pipeline with two
// roles per filter:
//
//   P warps (TW of them): the COVARIANCE CHAIN.  P_ff (feature-feature, the O(N^2) part) lives in REGISTERS: the lower
//     block-triangle of 2x2-feature tiles (6x6 doubles), one tile per thread; diagonal tiles keep a mirrored copy.  P_bb (16x16,
//     full symmetric) and P_fb (3N x 16) live in shared memory, one row per P thread.  These warps run P H^T, K, the Joseph /
//     partial rank update (vi_ekf_meas.cpp:254-265) and the structured A_d P A_d^T of propagate_state (vi_ekf.cpp:302-304) as
//     unrolled FP64 FMA streams on register operands; shared memory only supplies the O(N) row / column vectors.
//   x warp (one): the STATE CHAIN -- everything that depends on the state x only: boxplus, fix_depth, the next event's
//     measurement model h / H / residual or the next propagate's continuous dynamics, per-instance noise resolution, and S^-1 with
//     the Mahalanobis gate.  It runs one event ahead of the P warps: while they apply the rank update of event e, the x warp
//     applies the state correction of e and prepares the model of e + 1, so the serial scalar work (dependent FP64 chains,
//     sincos, divisions) is off the critical path.  Its code is one out-of-line function (x_role) with its own register
//     allocation; the kernel body proper is the P role, whose live ranges are the tile and the operands of the running loop.
//
// Feature removal (clear_feature / keep_only_features, vi_ekf_feat.cpp:50-142) reshuffles the tiles through the filter's own
// packed P in global memory (L2 round trip), which is also the load / store path at the ends of a slice.
//
// Padding: with `len` live features the tile grid covers lenp = len rounded up to even; every quantity of the padding
// feature (A_d rows, G_d rows, Y rows, P_fb rows, update vectors, tile entries) is an exact zero, so no per-element guards
// are needed inside the unrolled loops.
#include <stdlib.h>
#include "viekf_dev.cuh"

namespace vk {

constexpr int PBS = 18;  // row stride of P_bb: even (16-byte rows for vector loads), 2-way conflicts at worst for row-per-lane access

struct TLayout {
  int x, Pbb, Pfb, lam, dx, Abd, AbT, T, GqT, Ac, Gb, Afd, Gd, Y, upd, small, ids, smap, evb, fixk, fixv, zvr, model, gate, total, TR;
};
// prepared model of an event (written by the x warp one event ahead, double buffered by event parity), in doubles
enum { MD_H = 0 /*18*/, MD_RES = 18 /*3*/, MD_DT = 21, MD_INT = 24 /*int32 part*/, MD_SIZE = 32 };
enum { MI_COL = 0 /*6*/, MI_M = 6, MI_KIND = 7, MI_FI = 8, MI_HNAN = 9, MI_LEN = 10, MI_PREP = 11 };
enum { MK_NONE = 0, MK_NAN, MK_INVALID, MK_NEWFEAT, MK_INACTIVE, MK_RUN, MK_PROP_LATCH, MK_PROP_NOP, MK_PROP_BACK, MK_PROP_RUN };
// S^-1 / gate block of the running update (written by the x warp between the two barriers of the gain computation)
enum { GD_SI = 0 /*9*/, GD_S = 9 /*9*/, GD_INT = 18, GD_SIZE = 24 };
enum { GI_GATED = 0, GI_NAN = 1 };
constexpr int EVSLOTS = 4;  // staged viekf_event records: events e .. e+1 are resident while e runs, e+2 is in flight

// The layout is a function of the capacity CLASS of the executor (14 / 26 / 36 features for 1 / 3 / 6 tile warps), not of the
// handle's N: every offset is then a compile-time constant, shared-memory accesses use immediate offsets off one base register,
// and the ~30 per-array pointers need no registers (they were the bulk of the kernel's local-memory traffic).
__host__ __device__ constexpr TLayout make_tlayout(int N) {
  TLayout L{};
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
  L.GqT = o;   o += 96;                  // Qu[q] G_d(cc, q), [q][cc]: the base factor of G_d Qu G_d^T for the P'_fb rows
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
  L.evb = o;   o += 32 * EVSLOTS;        // staged viekf_event records (31 doubles each)
  L.fixk = o;  o += ev2(NP / 2 + 1);     // pending fix_depth covariance corrections per feature: kind (int32) ...
  L.fixv = o;  o += NP;                  // ... and value
  L.zvr = o;   o += 64;                  // ring of 8 resolved value blocks (8 doubles each): event r reads block r & 7
  L.model = o; o += 2 * MD_SIZE;
  L.gate = o;  o += GD_SIZE;
  L.total = ev2(o);
  return L;
}

template <int TW>
__host__ __device__ constexpr int tile_cap() {
  return TW == 1 ? 14 : (TW == 3 ? 26 : 36);
}
size_t tile_smem_bytes(int N) {
  const int tw = tile_warps_for(N);
  return (size_t)make_tlayout(tw == 1 ? 14 : (tw == 3 ? 26 : 36)).total * sizeof(double);
}

// Optional phase timing (build with -DVIEKF_PHASE_TIMING): lane 0 of every warp adds the clocks between consecutive marks to
// g_phase_clk[role][k]; since the marks sit right after barriers this is the elapsed time of each phase on that warp.
#ifdef VIEKF_PHASE_TIMING
__device__ unsigned long long g_phase_clk[8 * 64];
__device__ unsigned long long g_phase_cnt[8 * 64];
#define PH_DECL long long ph_last;
#define PH_INIT(c) (c).ph_last = clock64();
// slot k: work of the phase (mark before the barrier); slot k + 32: wait at the barrier (mark after it); one row per warp role
#define PH(c, k)                                                                                  \
  if (((c).tid & 31) == 0) {                                                                      \
    const long long ph_now = clock64();                                                           \
    atomicAdd(&g_phase_clk[64 * (((c).tid >> 5) & 7) + (k)], (unsigned long long)(ph_now - (c).ph_last)); \
    atomicAdd(&g_phase_cnt[64 * (((c).tid >> 5) & 7) + (k)], 1ull);                                 \
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

// G = 32 (TW + 1) threads per filter: TW P warps, then the x warp
template <int TW>
struct TCtx : Ctx<32 * (TW + 1)> {
  double *Pbb, *Pfb, *GqT, *fixv, *zvr, *model, *gate, *evb;
  int *smap, *fixk;
  int ti, tj;   // tile coordinates in units of 2 features, ti >= tj
  bool tile;    // this thread owns a tile of the (capacity) grid
  bool xw;      // this thread belongs to the x warp
  int lane;
  int64_t ginst;
  PH_DECL
  // barrier among the P warps of the group only (ids 9.. next to the group barriers 1..8)
  __device__ __forceinline__ void psync() const { asm volatile("bar.sync %0, %1;" ::"r"(this->gi + 9), "r"(32 * TW) : "memory"); }
};

#define PT(r, c) Pt[(r) * 6 + (c)]

#ifndef VIEKF_LOCKSTEP
#define VIEKF_LOCKSTEP 1
#endif
constexpr int LOCKSTEP = VIEKF_LOCKSTEP;  // events between CTA-wide barriers (power of two)

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

// ------------------------------------------------------------------------------------------------------------------
// the x warp: state chain
// ------------------------------------------------------------------------------------------------------------------
template <int TW>
__device__ __forceinline__ void init_tctx(TCtx<TW>& c, double* base) {
  constexpr TLayout L = make_tlayout(tile_cap<TW>());
  constexpr int N = tile_cap<TW>();
  c.x = base + L.x; c.P = base + L.T; c.lam = base + L.lam; c.dx = base + L.dx; c.Abd = base + L.Abd; c.AbT = base + L.AbT; c.T = base + L.T;
  c.Ac = base + L.Ac; c.Gb = base + L.Gb; c.Afd = base + L.Afd; c.Gd = base + L.Gd; c.Y = base + L.Y; c.upd = base + L.upd;
  c.Pbb = base + L.Pbb; c.Pfb = base + L.Pfb; c.GqT = base + L.GqT;
  c.sm = base + L.small;
  c.ids = (int*)(base + L.ids);
  c.rm = c.ids + N;
  c.sc = c.rm + N;
  c.smap = (int*)(base + L.smap);
  c.fixk = (int*)(base + L.fixk);
  c.fixv = base + L.fixv;
  c.zvr = base + L.zvr;
  c.model = base + L.model;
  c.gate = base + L.gate;
  c.evb = base + L.evb;
  c.tcur = c.sm + SM_T;
  c.zv = c.zvr;
}

// x <- x [+] d s (vi_ekf_helper.cpp:88-98) on the x warp: item i < len is feature i, item len is the base state; one item per lane,
// all in one pass for len <= 31.  Both kinds of item share the expensive part -- q' = exp(v) (x) q or q (x) exp(v), with sincos,
// a square root and a division inside exp -- so the base state does not cost a second pass through it.
static __device__ __noinline__ void xw_boxplus(double* __restrict__ x, const double* __restrict__ d, double s, int len, int lane) {
  // the 13 additive base components (POS, VEL | B_A, B_G, MU) one per lane: a single lane ran them as 13 dependent
  // load / FMA / store round trips in front of the quaternion, on the critical path of every event
  if (lane < 13) {
    const int xi = lane < 6 ? lane : lane + 4, di = lane < 6 ? lane : lane + 3;
    x[xi] = fma(d[di], s, x[xi]);
  }
  for (int it = lane; it <= len; it += 32) {
    const bool isb = (it == len);
    double* xq = isb ? x + 6 : x + 17 + 5 * it;
    const Q4 q = ldq(xq);
    V3 dv;
    if (isb) {
      dv = mk3(d[6] * s, d[7] * s, d[8] * s);
    } else {
      const double* df = d + 16 + 3 * it;
      const double d0 = df[0] * s, d1 = df[1] * s;
      const V3 T0 = rota_ex(q), T1 = rota_ey(q);
      dv = mk3(T0.x * d0 + T1.x * d1, T0.y * d0 + T1.y * d1, T0.z * d0 + T1.z * d1);
      xq[4] = fma(df[2], s, xq[4]);
    }
    const Q4 e = qexp(dv);
    const Q4 a = isb ? q : e, b = isb ? e : q;  // base: q (x) exp(v); feature: exp(T_zeta d) (x) q
    stq(xq, otimes(a, b));
  }
}

// per-instance values (Philox noise: a ~400-instruction dependent chain per value) of events rel .. rel+3, 4 x 8 lanes, into the ring
template <int TW>
__device__ __forceinline__ void xw_resolve(const TCtx<TW>& c, const KParams& p, int64_t rel) {
  const int64_t r = rel + (c.lane >> 3);
  if (r < p.ev_count) resolve_values(p, p.ev_first + r, c.b, c.ginst, c.lane & 7, c.zvr + 8 * (int)(r & 7));
}

// continuous dynamics and Jacobian blocks of propagate_state (phase A, vi_ekf_dyn.cpp:14-135) on the 32 lanes of the x warp
template <int TW>
__device__ __forceinline__ void xw_phase_a(const TCtx<TW>& c, const KParams& p, const double* ub, int len) {
  base_dynamics(c, p, ub);  // zero fill by the warp, entries by lane 31
  for (int i = c.lane; i < len; i += 32) feature_dynamics(c, p, ub, i);
}

// Prepares event `rel` (relative to ev_first) while the covariance warps are still busy with the event before it (kind
// `cur_kind`, -1 = none): everything that depends on the state x only.
//   UPD:  add_measurement's bookkeeping (NaN reject, feature lookup, new feature / invalid id, vi_ekf_meas.cpp:130-147), the
//         measurement model h, H and the residual (vi_ekf_meas.cpp:203-228,281-395)
//   PROP: the body-frame input, the guards of propagate_state (vi_ekf.cpp:265-290) and the continuous dynamics (phase A)
// Precondition: x, ids, len, started, time are those event `rel` will see (the x warp itself wrote them, or a group barrier
// followed the handler that did).
template <int TW>
static __device__ __noinline__ void xw_prepare(const KParams& p, double* base, int gi, int lane, int64_t b, int rel, int cur_kind) {
  constexpr int G = 32 * (TW + 1);
  TCtx<TW> c;
  init_tctx(c, base);
  c.gi = gi;
  c.tid = G - 32 + lane;
  c.lane = lane;
  c.b = b;
  const viekf_event& E = *(const viekf_event*)(c.evb + 32 * (rel & (EVSLOTS - 1)));
  double* M = c.model + MD_SIZE * (rel & 1);
  int* Mi = (int*)(M + MD_INT);
  const double* zvs = c.zvr + 8 * (rel & 7);
  const int kind = E.kind;
  if (kind == VIEKF_EV_PROP) {
    double ub[6];
    {
      Q4 qbu = ldq(p.qbu);
      V3 a = rota(qbu, mk3(zvs[0], zvs[1], zvs[2])), g = rota(qbu, mk3(zvs[3], zvs[4], zvs[5]));
      ub[0] = a.x; ub[1] = a.y; ub[2] = a.z; ub[3] = g.x; ub[4] = g.y; ub[5] = g.z;
    }
    // the body-frame input of the latest propagate_state call is kept for UPD events with aux == 1 (vi_ekf.cpp:265-267); any
    // such event before this PROP has already taken its values
    if (lane < 6) c.sm[SM_UB + lane] = ub[lane];
    const int started = c.sc[SC_STARTED];
    const double dt = E.t - *c.tcur;
    int mk = MK_PROP_RUN;
    if (!started) mk = MK_PROP_LATCH;
    else if (fabs(dt) < 1e-6) mk = MK_PROP_NOP;
    else if (dt < 0) mk = MK_PROP_BACK;
    const int len = c.sc[SC_LEN];
    // phase A writes the continuous feature blocks into the Y buffer, which a running propagate still reads: deferred then
    const bool defer = (cur_kind == VIEKF_EV_PROP);
    if (mk == MK_PROP_RUN && !defer) xw_phase_a(c, p, ub, len);
    if (lane == 0) {
      Mi[MI_KIND] = mk;
      Mi[MI_LEN] = len;
      Mi[MI_PREP] = (mk == MK_PROP_RUN && !defer) ? 1 : 0;
      M[MD_DT] = dt;
    }
  } else if (kind == VIEKF_EV_UPD) {
    const double* zv = (E.aux == 1) ? (c.sm + SM_UB) : zvs;
    const int type = E.type, zdim = E.zdim, rd = E.rdim;
    int mk = MK_RUN, fi = -1;
    bool zn = false;
    for (int k = 0; k < zdim; k++) zn |= (zv[k] != zv[k]);
    if (zn) {
      mk = MK_NAN;
    } else {
      const bool feat_type = (type == VIEKF_FEAT || type == VIEKF_QZETA || type == VIEKF_DEPTH || type == VIEKF_INV_DEPTH);
      if (feat_type) {
        fi = (E.id >= 0) ? find_feature(c, E.id) : -1;
        if (fi < 0)  // aux == 2: update() replayed by handle_measurements, which never initialises
          mk = (type == VIEKF_FEAT && E.id >= 0 && E.aux != 2) ? MK_NEWFEAT : MK_INVALID;
      }
      if (mk == MK_RUN && !E.active) mk = MK_INACTIVE;  // logged only: no gate, no correction, but fix_depth still runs
    }
    if (mk == MK_RUN) {
      HReg H;
      if (fast_model(c.x, p, type, rd, zv, fi, H)) {
        if (lane == 0) {
#pragma unroll
          for (int k = 0; k < 18; k++) M[MD_H + k] = H.h[k];
#pragma unroll
          for (int k = 0; k < 3; k++) M[MD_RES + k] = H.res[k];
#pragma unroll
          for (int k = 0; k < 6; k++) Mi[MI_COL + k] = H.col[k];
          Mi[MI_M] = H.m;
          Mi[MI_HNAN] = H.hnan ? 1 : 0;
        }
      } else if (lane == 0) {
        int cols[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};
        double res[4];
        generic_model(c.x, p, type, zdim, zv, fi, M + MD_H, res, cols);
        for (int k = 0; k < 3; k++) M[MD_RES + k] = res[k];
        for (int k = 0; k < 6; k++) Mi[MI_COL + k] = cols[k];
        Mi[MI_M] = cols[6];
        Mi[MI_HNAN] = cols[8];
      }
    }
    if (lane == 0) {
      Mi[MI_KIND] = mk;
      Mi[MI_FI] = fi;
      Mi[MI_LEN] = c.sc[SC_LEN];
    }
  }
  __syncwarp();
}

// Pt(r, cc) += sum_k a[r * ars + k * aks] * b[cc * brs + k * bks], k < K: the rank-K correction of a 6x6 tile from two 6xK operand
// blocks in shared memory.  Software-pipelined by hand like rows<>: the twelve operands of step k+1 are loaded into their own
// registers before the 36 FMAs of step k are issued (left to itself ptxas keeps ONE operand register and exposes a
// shared-memory round trip, ~30 clk, per group of six FMAs).
__device__ __forceinline__ void ld6(double (&v)[6], const double* s, int stride) {
#pragma unroll
  for (int r = 0; r < 6; r++) v[r] = s[r * stride];
}
__device__ __forceinline__ void fma66(double (&Pt)[36], const double (&a)[6], const double (&b)[6]) {
#pragma unroll
  for (int r = 0; r < 6; r++)
#pragma unroll
    for (int cc = 0; cc < 6; cc++) PT(r, cc) = fma(a[r], b[cc], PT(r, cc));
}
template <int K>
__device__ __forceinline__ void outer_k(double (&Pt)[36], const double* a, int ars, int aks, const double* b, int brs, int bks) {
  double a0[6], b0[6], a1[6], b1[6];
  ld6(a0, a, ars);
  ld6(b0, b, brs);
#pragma unroll 1
  for (int k = 0; k + 2 <= K; k += 2) {
    ld6(a1, a + (k + 1) * aks, ars);
    ld6(b1, b + (k + 1) * bks, brs);
    fma66(Pt, a0, b0);
    if (k + 2 < K) {
      ld6(a0, a + (k + 2) * aks, ars);
      ld6(b0, b + (k + 2) * bks, brs);
    }
    fma66(Pt, a1, b1);
  }
  if (K & 1) fma66(Pt, a0, b0);
}
// Pt(r, cc) += sum_q (gr[r * 6 + q] * Qu[q]) * gc[cc * 6 + q], q < 6   (G_d,i Qu G_d,j^T)
__device__ __forceinline__ void outer_gq(double (&Pt)[36], const double* gr, const double* gc, const double* Qu) {
  double a0[6], b0[6], a1[6], b1[6];
  ld6(a0, gr, 6);
  ld6(b0, gc, 6);
#pragma unroll 1
  for (int q = 0; q < 6; q += 2) {
    ld6(a1, gr + q + 1, 6);
    ld6(b1, gc + q + 1, 6);
    {
      const double qu = Qu[q];
#pragma unroll
      for (int r = 0; r < 6; r++) a0[r] *= qu;
    }
    fma66(Pt, a0, b0);
    if (q + 2 < 6) {
      ld6(a0, gr + q + 2, 6);
      ld6(b0, gc + q + 2, 6);
    }
    {
      const double qu = Qu[q + 1];
#pragma unroll
      for (int r = 0; r < 6; r++) a1[r] *= qu;
    }
    fma66(Pt, a1, b1);
  }
}

// ------------------------------------------------------------------------------------------------------------------
// PROP: propagate_state (vi_ekf.cpp:262-318) with the structured covariance product
//   model (x warp, one event ahead): guards, body-frame input, continuous dynamics (phase A)
//   B   P warps: discretised rows [A_d | G_d], one row per thread                    || x warp: noise values four events ahead
//   C+D1  P warps: own row of T = A_bb P_bb / Y = A_fb P_bb + A_ff P_fb (old P), then the tile Z = A_ff,i P_ij + A_fb,i P_bf,j
//                                                                                      || x warp: state boxplus, fix_depth, time
//   D2  P warps: P'_bb, own row of P'_fb = Y A_bb^T + G_d Qu G_d^T (in place), tile P'_ij = Z A_ff,j^T + Y_i A_fb,j^T + G_d,i Qu G_d,j^T
//                                                                                      || x warp: the next event's model
// ------------------------------------------------------------------------------------------------------------------
template <int TW>
__device__ __forceinline__ void p_propagate(TCtx<TW>& c, const KParams& p, const double* M, double (&Pt)[36]) {
  constexpr int PTH = 32 * TW;
  const int* Mi = (const int*)(M + MD_INT);
  if (Mi[MI_KIND] != MK_PROP_RUN) return;  // the call does not advance the filter (vi_ekf.cpp:274-290)
  const double dt = M[MD_DT];
  const int len = Mi[MI_LEN];
  const int n = 16 + 3 * len;
  const int lenp = (len + 1) & ~1;
  const int nfr = 3 * lenp;
  const int npad = 16 + nfr;
  if (!Mi[MI_PREP]) c.sync();  // phase A could not run ahead (the previous event was a propagate that still read the Y buffer)
  PH(c, 1)
  // ---- phase B ----
  discretise_rows(c, dt, n, npad, c.tid, PTH);
  PHS(c, 2)
  // ---- phase C: own rows of T and Y (two column halves: 8 accumulators + two operand rows in flight next to the 72 tile
  //      registers), and GqT[q][cc] = Qu[q] G_d(cc, q) ----
  for (int e = c.tid; e < 96; e += PTH) {
    const int q = e >> 4, cc = e & 15;
    c.GqT[e] = c.Gd[cc * 6 + q] * p.Qu[q];
  }
  for (int r = c.tid; r < npad; r += PTH) {
    if (r < 16) {
#pragma unroll 1
      for (int h = 0; h < 16; h += 8) {
        double acc[8];
#pragma unroll
        for (int q = 0; q < 8; q++) acc[q] = 0.0;
        rows<8, 16, 1, PBS, true>(acc, c.Abd + r * 16, c.Pbb + h);
#pragma unroll
        for (int q = 0; q < 8; q++) c.T[r * 17 + h + q] = acc[q];
      }
    } else {
      const int fr = r - 16, i = fr / 3;
      const double* Ad = c.Afd + fr * 19;
      double* Yr = c.Y + fr * 17;
#pragma unroll 1
      for (int h = 0; h < 16; h += 8) {
        double acc[8];
#pragma unroll
        for (int q = 0; q < 8; q++) acc[q] = 0.0;
        rows<8, 13, 1, PBS, true>(acc, Ad + 3, c.Pbb + 3 * PBS + h);
        rows<8, 3, 1, 17, false>(acc, Ad + 16, c.Pfb + 3 * i * 17 + h);
#pragma unroll
        for (int q = 0; q < 8; q++) Yr[h + q] = acc[q];
      }
    }
  }
  // ---- phase D1: Z = A_ff,i P_ij + A_fb,i P_bf,j (last readers of the old P_fb) ----
  const bool live_tile = c.tile && 2 * c.ti < len;
  const int R = 6 * c.ti, C = 6 * c.tj;
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
    // Z += A_fb,i P_bf,j   (base columns VEL..MU: 3..15), operands of column k+1 in flight while the FMAs of column k issue
    outer_k<13>(Pt, c.Afd + R * 19 + 3, 19, 1, c.Pfb + C * 17 + 3, 17, 1);
  }
  PHS(c, 3)
  // ---- phase D2 ----
  // P'_bb = T A_bb^T + G_d Qu G_d^T + Qx: lower triangle, mirrored
  for (int e = c.tid; e < 136; e += PTH) {
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
  // own row of P'_fb = Y A_bb^T + G_d,f Qu G_d,b^T, in place (D1 was the last reader of the old P_fb)
  for (int r = 16 + c.tid; r < npad; r += PTH) {
    const int fr = r - 16;
    const double* Yr = c.Y + fr * 17;
    const double* Gr = c.Gd + r * 6;
    double* Pr = c.Pfb + fr * 17;
#pragma unroll 1
    for (int h = 0; h < 16; h += 8) {
      double acc[8];
#pragma unroll
      for (int q = 0; q < 8; q++) acc[q] = 0.0;
      rows<8, 16, 1, 16, true>(acc, Yr, c.AbT + h);
      rows<8, 6, 1, 16, true>(acc, Gr, c.GqT + h);
#pragma unroll
      for (int q = 0; q < 8; q++) Pr[h + q] = acc[q];
    }
  }
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
    outer_k<13>(Pt, c.Y + R * 17 + 3, 17, 1, c.Afd + C * 19 + 3, 19, 1);
    // O += (G_d,i Qu) G_d,j^T: the row factor scaled by Qu[q] first, as G_d Qu G_d^T is evaluated everywhere
    outer_gq(Pt, c.Gd + (16 + R) * 6, c.Gd + (16 + C) * 6, p.Qu);
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
  PH(c, 4)
  // no barrier here: the one that ends the event publishes P'_bb, P'_fb and the pending depth corrections
}

// the x warp's part of a propagate
template <int TW>
__device__ __forceinline__ void x_propagate(TCtx<TW>& c, const KParams& p, const viekf_event& E, const double* M, int64_t rel, bool& resolved) {
  const int* Mi = (const int*)(M + MD_INT);
  const int mk = Mi[MI_KIND];
  if (mk != MK_PROP_RUN) {  // the call does not advance the filter (vi_ekf.cpp:274-290)
    if (c.lane == 0) {
      if (mk == MK_PROP_LATCH) {
        c.sc[SC_STARTED] = 1;
        *c.tcur = E.t;
      } else if (mk == MK_PROP_BACK) {
        c.sc[SC_FLAGS] |= VIEKF_FLAG_BACKWARD_DT;
      }
    }
    __syncwarp();
    return;
  }
  const double dt = M[MD_DT];
  const int len = Mi[MI_LEN];
  if (!Mi[MI_PREP]) {
    xw_phase_a(c, p, c.sm + SM_UB, len);
    c.sync();
  }
  PH(c, 1)
  if ((rel & 3) == 0) xw_resolve(c, p, rel + 4);  // idle while the P warps discretise
  resolved = true;
  PHS(c, 2)
  // the state: x <- x [+] xdot dt, fix_depth, time (phases E / F of the other executor)
  xw_boxplus(c.x, c.dx, dt, len, c.lane);
  for (int i = c.lane; i < len; i += 32) tile_fix_x(c, p, i);
  if (c.lane == 0) *c.tcur = E.t;
  PHS(c, 3)
}

// ------------------------------------------------------------------------------------------------------------------
// UPD: VIEKF::update (vi_ekf_meas.cpp:196-278) as a symmetric rank update
//   P_ij += (1 - a_i a_j) ( W_i.K_j - K_i.H_j ),   H = P H^T,  K = H S^-1,  W = K S - H,  a = 1 - lambda (partial update) or 0
// which is the Joseph form (I-KH) P (I-KH)^T + K R K^T - P = K S K^T - K (HP) - (HP)^T K^T with the two terms that share
// K_j merged.  Vectors live in the update workspace (padding rows are exact zeros).
//   model (x warp, one event ahead): bookkeeping, h, H, residual
//   P1  tile + base warps: P H^T rows                      ||  x warp: value resolution
//   P2  x warp: S = H P H^T + R, S^-1, Mahalanobis gate
//   P3  tile + base warps: K, W rows, lambda o (K r)
//   P4  tile warps: rank update of the tiles  ||  base warp: rank update of [P_bb; P_fb]  ||  x warp: state boxplus, fix_depth,
//       then (in the event loop) the next event's model
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

// own row r of [P_bb; P_fb] (16 columns): the row goes to registers, the column vectors are warp broadcasts
template <int RD>
__device__ __forceinline__ void p_row_rank(const TUpdVec& v, double* Pbb, double* Pfb, int r) {
  const double2 wr = v.Wv[r], kr = v.Kv[r];
  const double ar = v.av[r];
  const double w2r = (RD == 3) ? v.W2[r] : 0.0, k2r = (RD == 3) ? v.K2[r] : 0.0;
  double* prow = (r < 16) ? Pbb + r * PBS : Pfb + (r - 16) * 17;
  const int jmax = (r < 16) ? r : 15;
  double pr[16];
#pragma unroll
  for (int j = 0; j < 16; j++) pr[j] = prow[j];
#pragma unroll
  for (int j = 0; j < 16; j++) {
    const double2 kj = v.Kv[j], hj = v.Hv[j];
    const double aj = v.av[j];
    const double k2j = (RD == 3) ? v.K2[j] : 0.0, h2j = (RD == 3) ? v.H2[j] : 0.0;
    const double d = rank_delta<RD>(wr, kr, w2r, k2r, kj, hj, k2j, h2j);
    const double u = fma(-ar, aj, 1.0);
    pr[j] = fma(d, u, pr[j]);
  }
  // P_bb keeps both halves: the owner of row r writes (r, j) and its mirror (j, r) for j <= r; what it read above the
  // diagonal is recomputed but never stored
#pragma unroll
  for (int j = 0; j < 16; j++) {
    const bool ok = j <= jmax;
    if (ok) prow[j] = pr[j];
    if (ok && r < 16) Pbb[j * PBS + r] = pr[j];
  }
}

template <int TW, int RD>
__device__ __forceinline__ int p_update(TCtx<TW>& c, const KParams& p, const double* M, double (&Pt)[36]) {
  constexpr int PTH = 32 * TW;  // threads that own covariance rows
  constexpr int MC = 6;
  const int* Mi = (const int*)(M + MD_INT);
  int* Gi = (int*)(c.gate + GD_INT);
  const int len = Mi[MI_LEN];
  const int lenp = (len + 1) & ~1;
  const int npad = 16 + 3 * lenp;
  const int m = Mi[MI_M];
  TUpdVec v(c.upd, 16 + 6 * ((tile_cap<TW>() + 1) / 2));
  PH(c, 8)
  // ---- P1: P H^T ----
  {
    double h[RD][MC];
    int col[MC];
#pragma unroll
    for (int k = 0; k < MC; k++) {
      col[k] = Mi[MI_COL + k];
#pragma unroll
      for (int a = 0; a < RD; a++) h[a][k] = M[MD_H + a * MC + k];
    }
    const bool featcols = (m > 0) && (col[0] >= 16);  // every H touches either base columns or the columns of one feature
    if (!featcols) {
      for (int i = c.tid; i < npad; i += PTH) {
        double s[3] = {0, 0, 0};
#pragma unroll
        for (int k = 0; k < MC; k++)
          if (k < m) {
            const double pv = (i < 16) ? c.Pbb[i * PBS + col[k]] : c.Pfb[(i - 16) * 17 + col[k]];
#pragma unroll
            for (int a = 0; a < RD; a++) s[a] = fma(pv, h[a][k], s[a]);
          }
        v.Hv[i] = make_double2(s[0], s[1]);
        if (RD == 3) v.H2[i] = s[2];
      }
    } else {
      const int fi = (col[0] - 16) / 3, tb = fi >> 1, sb = fi & 1;
      // effective 3-column Jacobian of the feature block: he[a][kk] = sum over the k with local column kk
      double he[RD][3];
#pragma unroll
      for (int a = 0; a < RD; a++)
#pragma unroll
        for (int kk = 0; kk < 3; kk++) he[a][kk] = 0.0;
#pragma unroll
      for (int k = 0; k < 2; k++)  // feature-column models have at most two live columns (FEAT, QZETA: 2; DEPTH, INV_DEPTH: 1)
        if (k < m) {
          const int lk = col[k] - 16 - 3 * fi;
#pragma unroll
          for (int a = 0; a < RD; a++)
#pragma unroll
            for (int kk = 0; kk < 3; kk++)
              if (lk == kk) he[a][kk] += h[a][k];
        }
      if (c.tid < 16) {
        const int i = c.tid;
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
  }
  PHS(c, 9)
  PHS(c, 10)  // the x warp evaluates S^-1 and the gate between these two barriers
  if (Gi[GI_GATED]) return VIEKF_MEAS_GATED;  // early return: no fix_depth (vi_ekf_meas.cpp:238)
  // ---- P3: K = P H^T S^-1, W = K S - P H^T, state correction lambda o (K r) ----
  {
    double S[3][3], Si[3][3];
#pragma unroll
    for (int a = 0; a < RD; a++) {
#pragma unroll
      for (int b = 0; b < RD; b++) {
        Si[a][b] = c.gate[GD_SI + a * 3 + b];
        S[a][b] = c.gate[GD_S + a * 3 + b];
      }
    }
    bool bad = false;
    for (int i = c.tid; i < npad; i += PTH) {
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
      v.av[i] = p.use_partial ? 1.0 - c.lam[i] : 0.0;
    }
    if (bad) atomicOr((unsigned*)&Gi[GI_NAN], 1u);
  }
  PH(c, 11)
  c.psync();  // the x warp forms the state correction lambda o (K r) itself and is already applying it
  PH(c, 11 + 32)
  // ---- P4: rank update of the own tile and the own row(s) of [P_bb; P_fb] ----
  if (!Gi[GI_NAN]) {
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
    for (int r = c.tid; r < npad; r += PTH) p_row_rank<RD>(v, c.Pbb, c.Pfb, r);
    PH(c, 14)
  }
  return VIEKF_MEAS_SUCCESS;
}

// the x warp's part of an update: S^-1 and the gate between the gain barriers, then the state correction
template <int TW, int RD>
__device__ __forceinline__ int x_update(TCtx<TW>& c, const KParams& p, const viekf_event& E, const double* M, int64_t rel, bool& resolved) {
  constexpr int MC = 6;
  const int* Mi = (const int*)(M + MD_INT);
  int* Gi = (int*)(c.gate + GD_INT);
  const int len = Mi[MI_LEN];
  const int m = Mi[MI_M];
  TUpdVec v(c.upd, 16 + 6 * ((tile_cap<TW>() + 1) / 2));
  PH(c, 8)
  if ((rel & 3) == 0) xw_resolve(c, p, rel + 4);  // idle while the P warps form P H^T
  resolved = true;
  PHS(c, 9)
  // ---- P2: S = H P H^T + R, S^-1 (partial-pivot LU), Mahalanobis gate (vi_ekf_meas.cpp:232-239) ----
  double S[3][3], Si[3][3], res[3];
  bool gated;
  {
    {
      double ph[MC][3];
#pragma unroll
      for (int k = 0; k < MC; k++) {
        const int ck = (k < m) ? Mi[MI_COL + k] : 0;
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
            if (k < m) s += M[MD_H + a * MC + k] * ph[k][b];
          S[a][b] = s + E.R[a + b * RD];
        }
    }
#pragma unroll
    for (int a = 0; a < 3; a++) res[a] = M[MD_RES + a];
    const double mahal = inverse_and_mahal<RD>(S, Si, res);
    gated = mahal > 9.0;
    if (c.lane == 0) {
#pragma unroll
      for (int a = 0; a < RD; a++)
#pragma unroll
        for (int b = 0; b < RD; b++) {
          c.gate[GD_SI + a * 3 + b] = Si[a][b];
          c.gate[GD_S + a * 3 + b] = S[a][b];
        }
      Gi[GI_GATED] = gated ? 1 : 0;
      Gi[GI_NAN] = Mi[MI_HNAN];
      if (gated) c.sc[SC_NGATED] += 1;
    }
  }
  PHS(c, 10)
  if (gated) return VIEKF_MEAS_GATED;
  // ---- the state correction lambda o (K r), K = P H^T S^-1 row by row (the same arithmetic as the P warps' K rows): the x warp
  //      does not wait for them ----
  bool bad = Mi[MI_HNAN] != 0;
  {
    const int npad = 16 + 3 * ((len + 1) & ~1);
    // the rows of all passes at once (loads first): three dependent shared-memory round trips in a row cost the x warp ~0.5 k cycles
    constexpr int NIT = (16 + 3 * ((tile_cap<TW>() + 1) & ~1) + 31) / 32;
    double2 h01[NIT];
    double h2[NIT], lm[NIT];
#pragma unroll
    for (int u = 0; u < NIT; u++) {
      const int i = c.lane + 32 * u;
      const bool in = i < npad;
      h01[u] = in ? v.Hv[i] : make_double2(0.0, 0.0);
      h2[u] = (RD == 3 && in) ? v.H2[i] : 0.0;
      lm[u] = (p.use_partial && in) ? c.lam[i] : 1.0;
    }
#pragma unroll
    for (int u = 0; u < NIT; u++) {
      const int i = c.lane + 32 * u;
      const double ph[3] = {h01[u].x, h01[u].y, h2[u]};
      double k3[3] = {0, 0, 0};
#pragma unroll
      for (int b = 0; b < RD; b++) {
        double s = 0;
#pragma unroll
        for (int a = 0; a < RD; a++) s += ph[a] * Si[a][b];
        k3[b] = s;
        bad |= (i < npad) && (s != s);
      }
      double d = 0;
#pragma unroll
      for (int a = 0; a < RD; a++) d += k3[a] * res[a];
      if (i < npad) c.dx[i] = p.use_partial ? lm[u] * d : d;
    }
  }
  bad = __any_sync(0xffffffffu, bad);
  __syncwarp();  // the dx rows of all lanes are visible to the boxplus below
  PH(c, 11)
  if (!bad) {
    xw_boxplus(c.x, c.dx, 1.0, len, c.lane);
  } else if (c.lane == 0) {  // NO_NANS(K) && NO_NANS(H) failed: the correction is skipped (vi_ekf_meas.cpp:247)
    c.sc[SC_FLAGS] |= VIEKF_FLAG_NAN_GAIN;
    c.sc[SC_NNAN] += 1;
  }
  __syncwarp();
  for (int i = c.lane; i < len; i += 32) tile_fix_x(c, p, i);
  PH(c, 12)
  return VIEKF_MEAS_SUCCESS;
}

// ------------------------------------------------------------------------------------------------------------------
// feature management
// ------------------------------------------------------------------------------------------------------------------
// init_feature (vi_ekf_feat.cpp:6-47).  Returns (uniformly) whether a slot was available.
// TILE = false: the x warp's instantiation of a handler (it takes part in the barriers and in the shared-memory work, owns no tile)
template <int TW, bool TILE>
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
  if (TILE && c.tile && c.ti == tb) {
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
template <int TW, bool TILE>
__device__ __forceinline__ void tile_compact(TCtx<TW>& c, const KParams& p, int len, int newlen, double (&Pt)[36]) {
  constexpr int G = 32 * (TW + 1);
  double* gP = p.P + c.b * p.ps;
  if (TILE) tile_store(c, gP, Pt, len);
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
  if (TILE) tile_load(c, gP, Pt, newlen, true);
  if (c.tid == 0) c.sc[SC_LEN] = newlen;
  c.sync();
}

template <int TW, bool TILE>
__device__ __forceinline__ void tile_clear_feature(TCtx<TW>& c, const KParams& p, int local, double (&Pt)[36]) {
  constexpr int G = 32 * (TW + 1);
  const int len = c.sc[SC_LEN];
  if (local < 0 || local >= len) return;
  c.sync();
  for (int f = c.tid; f < len - 1; f += G) c.smap[f] = f + (f >= local ? 1 : 0);
  c.sync();
  tile_compact<TW, TILE>(c, p, len, len - 1, Pt);
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
  if ((c.tid >> 5) == 0) {  // warp 0: P_bb, one row / column per lane
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
template <int TW, bool TILE>
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
  if (newlen < len) tile_compact<TW, TILE>(c, p, len, newlen, Pt);
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
template <int TW, bool TILE>
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
  if (TILE && c.tile && 2 * c.ti < len) {
#pragma unroll
    for (int i = 0; i < 36; i++) bad |= (Pt[i] != Pt[i]);
  }
  return bad;
}

// ------------------------------------------------------------------------------------------------------------------
// The event loop.  One handler per event kind and role; `TILE` selects the P role (owns a tile, Pt) or the x role.  Both roles
// execute the same sequence of group barriers for every event.
// ------------------------------------------------------------------------------------------------------------------
// the handlers that reshuffle the filter (all G threads cooperate; rare next to PROP / UPD)
template <int TW, bool TILE>
__device__ __forceinline__ int generic_event(TCtx<TW>& c, const KParams& p, const viekf_event& E, int kind, int64_t ginst, double (&Pt)[36]) {
  int result = VIEKF_MEAS_NONE;
  switch (kind) {
    case VIEKF_EV_INIT:
      result = tile_init_feature<TW, TILE>(c, p, c.zv[0], c.zv[1], E.z[4], Pt) ? 1 : 0;
      break;
    case VIEKF_EV_CLEAR:
      tile_clear_feature<TW, TILE>(c, p, find_feature(c, E.id), Pt);
      break;
    case VIEKF_EV_KEEP:
      tile_keep_only<TW, TILE>(c, p, p.id_pool + E.ids_off, E.n_ids, Pt);
      break;
    case VIEKF_EV_KFRESET:
      tile_keyframe_reset(c, p);
      break;
    case VIEKF_EV_STAT:
      c.sync();
      tile_pack_base(c);
      c.sync();
      ev_stat(c, p, E, ginst, [&]() { return tile_nan_scan<TW, TILE>(c, Pt); });
      break;
    default:
      break;
  }
  return result;
}

// The x role: the state chain of one filter slot of the CTA, for every filter that slot processes.  Out of line: its registers
// (no tile) are allocated separately from the kernel body.
template <int TW, int FPC>
static __device__ __noinline__ void x_role(const KParams& p, double* base, int gi, int lane) {
  constexpr int G = 32 * (TW + 1);
  constexpr int EVW = (int)(sizeof(viekf_event) / sizeof(double));
  static_assert(sizeof(viekf_event) % sizeof(double) == 0 && EVW <= 32, "viekf_event must fit one warp of doubles");
  TCtx<TW> c;
  init_tctx(c, base);
  c.gi = gi;
  c.tid = G - 32 + lane;
  c.lane = lane;
  c.tile = false;
  c.xw = true;
  c.ti = c.tj = 0;
  double Pz[36];  // never touched: every TILE = false handler skips the tile parts
  PH_INIT(c)
  for (int64_t b0 = (int64_t)blockIdx.x * FPC; b0 < p.B; b0 += (int64_t)gridDim.x * FPC) {
    const bool active = b0 + gi < p.B;
    const int64_t b = active ? b0 + gi : 0;
    c.b = b;
    const int64_t ginst = p.inst_off + b;
    c.ginst = ginst;
    c.sync();  // the previous filter of this slot has been stored
    // while the P warps load the filter: stage the first two event records, resolve the per-instance values of the first four events
    for (int r = 0; r < 2; r++)
      if (r < p.ev_count && lane < EVW) c.evb[32 * r + lane] = __ldg((const double*)(p.ev + p.ev_first + r) + lane);
    xw_resolve(c, p, 0);
    PHS(c, 0)
    if (active && p.ev_count > 0) xw_prepare<TW>(p, base, gi, lane, b, 0, -1);
    c.sync();
    for (int64_t rel = 0; rel < p.ev_count; rel++) {
      const int64_t e = p.ev_first + rel;
      const viekf_event& E = *(const viekf_event*)(c.evb + 32 * (int)(rel & (EVSLOTS - 1)));
      const double* Mod = c.model + MD_SIZE * (int)(rel & 1);
      // the record two events ahead is fetched now (one double per lane) and staged before the barrier that ends this event
      double ev_next = 0.0;
      const bool stage = (rel + 2 < p.ev_count) && lane < EVW;
      if (stage) ev_next = __ldg((const double*)(p.ev + e + 2) + lane);
      const int kind = E.kind;
      c.zv = c.zvr + 8 * (int)(rel & 7);
      bool resolved = false;
      if (active) {
        if (kind == VIEKF_EV_PROP) {
          x_propagate(c, p, E, Mod, rel, resolved);
        } else if (kind == VIEKF_EV_UPD) {
          const int mk = ((const int*)(Mod + MD_INT))[MI_KIND];
          if (mk == MK_NEWFEAT) {
            tile_init_feature<TW, false>(c, p, c.zv[0], c.zv[1], E.z[4], Pz);
          } else if (mk == MK_INACTIVE) {
            const int len = c.sc[SC_LEN];
            for (int i = lane; i < len; i += 32) tile_fix_x(c, p, i);
          } else if (mk == MK_RUN) {
            const int rd = E.rdim;  // one call site per residual dimension
            if (rd == 1) x_update<TW, 1>(c, p, E, Mod, rel, resolved);
            else if (rd == 2) x_update<TW, 2>(c, p, E, Mod, rel, resolved);
            else x_update<TW, 3>(c, p, E, Mod, rel, resolved);
          }
        } else {
          generic_event<TW, false>(c, p, E, kind, ginst, Pz);
        }
        // noise values four events ahead (if the handler had no idle slot for it), then the next event's model, while the P warps
        // finish this event.  Every handler that changes x / ids / len on other warps ends with a group barrier.
        __syncwarp();
        if (!resolved && (rel & 3) == 0) xw_resolve(c, p, rel + 4);
        if (rel + 1 < p.ev_count) xw_prepare<TW>(p, base, gi, lane, b, (int)(rel + 1), kind);
      }
      if (stage) c.evb[32 * (int)((rel + 2) & (EVSLOTS - 1)) + lane] = ev_next;
      PH(c, 16 + kind)
      if (FPC > 1 && ((int)rel & (LOCKSTEP - 1)) == LOCKSTEP - 1)
        __syncthreads();
      else
        c.sync();
      PH(c, 16 + kind + 32)
    }
    c.sync();
    if (!active) continue;
    latch_nan_state(c, tile_nan_scan<TW, false>(c, Pz));  // NaNsInTheHouse at the end of the slice (with the P warps)
  }
}

// ------------------------------------------------------------------------------------------------------------------
// the persistent tape executor, register-tile variant: a group of 32 (TW + 1) threads per filter, FPC groups per CTA.
// The groups of a CTA synchronise among themselves with named barriers inside an event and all together at the end of
// every event: the FPC filters stay within one event of each other, so the warps of an SM run the same few thousand
// instructions at any time (the executor is instruction-fetch bound when its CTAs drift apart over the ~0.5 MB of code).
// ------------------------------------------------------------------------------------------------------------------
template <int TW, int FPC>
__global__ void __launch_bounds__(32 * (TW + 1) * FPC, (384 / (32 * (TW + 1) * FPC) >= 1 ? 384 / (32 * (TW + 1) * FPC) : 1))
    viekf_tile_kernel(const __grid_constant__ KParams p) {
  extern __shared__ __align__(16) double smem[];
  constexpr int G = 32 * (TW + 1);
  constexpr int PTH = 32 * TW;
  constexpr TLayout L = make_tlayout(tile_cap<TW>());
  const int gi = threadIdx.x / G;
  // warp w of the CTA issues on SM sub-partition w % 4: when G is a multiple of 128 the x warps of all filters of the CTA would
  // share one sub-partition; rotating the roles by the filter index spreads them
#ifndef VIEKF_TILE_ROT
#define VIEKF_TILE_ROT 1
#endif
  const int tid = (G % 128 == 0 && VIEKF_TILE_ROT) ? (int)((threadIdx.x % G) + 32 * gi) % G : (int)(threadIdx.x % G);
  double* base = smem + gi * L.total;
  if ((tid >> 5) == TW) {
    x_role<TW, FPC>(p, base, gi, tid & 31);
    return;
  }
  // ---- the P role ----
  TCtx<TW> c;
  init_tctx(c, base);
  c.gi = gi;
  c.tid = tid;
  c.lane = tid & 31;
  c.xw = false;
  {
    const int t = tid;
    int ti = (int)((sqrtf(8.0f * (float)t + 1.0f) - 1.0f) * 0.5f);
    while (((ti + 1) * (ti + 2) >> 1) <= t) ti++;
    while (((ti * (ti + 1)) >> 1) > t) ti--;
    c.ti = ti;
    c.tj = t - ((ti * (ti + 1)) >> 1);
    c.tile = ti < L.TR;
  }
  constexpr int nmaxp = 16 + 6 * L.TR;
  for (int i = tid; i < nmaxp; i += PTH) c.lam[i] = (i < 16) ? p.lam[i] : p.lamf[(i - 16) % 3];
  for (int i = tid; i < 2 * L.TR; i += PTH) c.fixk[i] = 0;
  double Pt[36];
  PH_INIT(c)

  for (int64_t b0 = (int64_t)blockIdx.x * FPC; b0 < p.B; b0 += (int64_t)gridDim.x * FPC) {
    const bool active = b0 + gi < p.B;  // a group without a filter still takes part in the CTA-wide event barriers
    const int64_t b = active ? b0 + gi : 0;
    c.b = b;
    const int64_t ginst = p.inst_off + b;
    c.ginst = ginst;
    FilterMeta& M = p.meta[b];
    double* gP = p.P + b * p.ps;
    // ---- load the filter: meta, ids, live part of x; P_bb / P_fb into shared memory, P_ff into the register tiles ----
    c.sync();
    const int len0 = M.len;
    if (tid == 0) {
      c.sc[SC_LEN] = len0; c.sc[SC_NEXT] = M.next_id; c.sc[SC_STARTED] = M.started; c.sc[SC_FLAGS] = (int)M.flags;
      c.sc[SC_NGATED] = (int)M.n_gated; c.sc[SC_NFIX] = (int)M.n_fix; c.sc[SC_NNAN] = (int)M.n_nan_gain; c.sc[SC_NKF] = M.n_kf;
      *c.tcur = M.t;
      for (int k = 0; k < 6; k++) c.sm[SM_UB + k] = M.ub[k];
    }
    for (int i = tid; i < len0; i += PTH) c.ids[i] = p.ids[b * (p.N > 0 ? p.N : 1) + i];
    {
      const double* gx = p.x + b * p.xs;
      const int nx = 17 + 5 * len0;
      for (int i = tid; i < nx; i += PTH) c.x[i] = gx[i];
      for (int e = tid; e < 256; e += PTH) {
        const int r = e >> 4, cc = e & 15;
        c.Pbb[r * PBS + cc] = gP[psym(r, cc)];
      }
      const int nfr0 = 3 * ((len0 + 1) & ~1);
      for (int e = tid; e < nfr0 * 16; e += PTH) {
        const int fr = e >> 4, cc = e & 15;
        c.Pfb[fr * 17 + cc] = (fr < 3 * len0) ? gP[plo(16 + fr, cc)] : 0.0;
      }
      tile_load(c, gP, Pt, len0, false);
    }
    PHS(c, 0)
    c.sync();  // the x warp has prepared the first event
    // ---- run the tape slice ----
    for (int64_t rel = 0; rel < p.ev_count; rel++) {
      const int64_t e = p.ev_first + rel;
      // the event record and its prepared model are read from shared memory (staged / written by the x warp)
      const viekf_event& E = *(const viekf_event*)(c.evb + 32 * (int)(rel & (EVSLOTS - 1)));
      const double* Mod = c.model + MD_SIZE * (int)(rel & 1);
      tile_apply_fix(c, p, Pt);  // depth corrections recorded by the previous event
      const int kind = E.kind;
      c.zv = c.zvr + 8 * (int)(rel & 7);
      int result = VIEKF_MEAS_NONE;
      if (active) {
        if (kind == VIEKF_EV_PROP) {
          p_propagate(c, p, Mod, Pt);
        } else if (kind == VIEKF_EV_UPD) {
          const int mk = ((const int*)(Mod + MD_INT))[MI_KIND];
          if (mk == MK_NAN) {
            result = VIEKF_MEAS_NAN;
          } else if (mk == MK_INVALID) {
            result = VIEKF_MEAS_INVALID;  // the reference would index the state with -1 here (UB)
          } else if (mk == MK_NEWFEAT) {
            tile_init_feature<TW, true>(c, p, c.zv[0], c.zv[1], E.z[4], Pt);
            result = VIEKF_MEAS_NEW_FEATURE;
          } else if (mk == MK_INACTIVE) {
            result = VIEKF_MEAS_SUCCESS;  // logged only: the x warp runs fix_depth (vi_ekf_meas.cpp:230,271)
          } else {
            const int rd = E.rdim;  // one call site per residual dimension
            if (rd == 1) result = p_update<TW, 1>(c, p, Mod, Pt);
            else if (rd == 2) result = p_update<TW, 2>(c, p, Mod, Pt);
            else result = p_update<TW, 3>(c, p, Mod, Pt);
          }
        } else {
          result = generic_event<TW, true>(c, p, E, kind, ginst, Pt);
        }
      }
      PH(c, 16 + kind)
      // the filters of a CTA re-align every LOCKSTEP events (all start the next event together); in between only the group
      // synchronises, so a filter that finishes an event early (gated update, fewer live features) is not held back each time
      if (FPC > 1 && ((int)rel & (LOCKSTEP - 1)) == LOCKSTEP - 1)
        __syncthreads();
      else
        c.sync();
      PH(c, 16 + kind + 32)
      if (p.results && active && tid == 0 && b >= p.res_first && b < p.res_first + p.res_count)
        p.results[(b - p.res_first) * p.n_events + e] = (uint8_t)result;
    }
    // ---- store the filter ----
    c.sync();
    if (!active) continue;
    tile_apply_fix(c, p, Pt);
    latch_nan_state(c, tile_nan_scan<TW, true>(c, Pt));  // NaNsInTheHouse at the end of the slice
    const int len1 = c.sc[SC_LEN];
    if (tid == 0) {
      M.len = len1; M.next_id = c.sc[SC_NEXT]; M.started = c.sc[SC_STARTED]; M.flags = (uint32_t)c.sc[SC_FLAGS];
      M.n_gated = (uint32_t)c.sc[SC_NGATED]; M.n_fix = (uint32_t)c.sc[SC_NFIX]; M.n_nan_gain = (uint32_t)c.sc[SC_NNAN]; M.n_kf = c.sc[SC_NKF];
      M.t = *c.tcur;
      for (int k = 0; k < 6; k++) M.ub[k] = c.sm[SM_UB + k];
    }
    for (int i = tid; i < len1; i += PTH) p.ids[b * (p.N > 0 ? p.N : 1) + i] = c.ids[i];
    {
      double* gx = p.x + b * p.xs;
      const int nx = 17 + 5 * len1;
      for (int i = tid; i < nx; i += PTH) gx[i] = c.x[i];
      for (int e = tid; e < 256; e += PTH) {
        const int r = e >> 4, cc = e & 15;
        if (cc <= r) gP[plo(r, cc)] = c.Pbb[r * PBS + cc];
      }
      for (int e = tid; e < 3 * len1 * 16; e += PTH) {
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
    static unsigned long long z[8 * 64];
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
  // filters per CTA in event lock-step: 384 threads (168 registers each) per CTA wherever the shared memory allows --
  // TW = 1 (N <= 14, config #3 at N = 8): 6 filters of 64 threads; TW = 3 (N <= 26, config #4): 3 filters of 128 threads;
  // TW = 6 (N <= 36, config #5): 2 filters of 224 threads (144 registers).  Small batches spread over the SMs first.
  const int full = (tw == 1) ? 6 : ((tw == 3) ? 3 : 2);
  int fpc = (p.B < full * (int64_t)sm_count) ? 1 : full;
  if (const char* s = getenv("VIEKF_TILE_FPC")) fpc = atoi(s);  // test / tuning override; a value without an instantiation is an error
#ifdef VIEKF_DEV_FAST  // development build: only the configs[3] instantiations (each one costs ~25 s of ptxas)
  if (tw == 3 && fpc == 1) return launch_tile_fpc<3, 1>(p, sm_count, stream);
  if (tw == 3 && fpc == 3) return launch_tile_fpc<3, 3>(p, sm_count, stream);
  if (tw == 3 && fpc == 2) return launch_tile_fpc<3, 2>(p, sm_count, stream);
  return cudaErrorInvalidValue;
#else
  if (tw == 1) {
    switch (fpc) {
      case 1: return launch_tile_fpc<1, 1>(p, sm_count, stream);
      case 2: return launch_tile_fpc<1, 2>(p, sm_count, stream);
      case 6: return launch_tile_fpc<1, 6>(p, sm_count, stream);
      default: return cudaErrorInvalidValue;
    }
  }
  if (tw == 3) {
    switch (fpc) {
      case 1: return launch_tile_fpc<3, 1>(p, sm_count, stream);
      case 2: return launch_tile_fpc<3, 2>(p, sm_count, stream);
      case 3: return launch_tile_fpc<3, 3>(p, sm_count, stream);
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
