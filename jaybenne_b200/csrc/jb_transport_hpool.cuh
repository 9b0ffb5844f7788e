// jb_transport_hpool.cuh — the hybrid IMC / DDMC history kernel on warp-private photon pools (v12).
// Included by jb_kernels.cu after jb_transport_pool.cuh and jb_transport_ddmc.cuh (shares WarpPool's idea,
// pool_pick, Rng2, rng2_draw1, imc_step_lockstep, load_props, ld_once, bface_sample, apply_swarm_bc).
//
// Reference behaviour: TransportPhotons_DDMC, src/jaybenne/transport_ddmc.cpp:28-237, with ptcl_transport_step /
// ptcl_ddmc_step / ptcl_ddmc_albedo (transport_utils.hpp:111-397), scatter (scattering.hpp:21-29),
// SampleDDMCBlockFace for on-rank block changes (sample_ddmc_bface.cpp), the swarm boundary conditions
// (boundaries.hpp:24-84) and the hand-over (jaybenne.cpp:26-61). Same per-photon Philox streams, draw order and
// arithmetic as k_transport_hybrid (jb_transport_ddmc.cuh), so results are bit-identical; only the order in which
// a warp visits its photons differs.
//
// Why: k_transport_hybrid keeps one photon per lane in 96-128 registers (20 warps per SM) and every DDMC segment
// ends in one of seven outcomes, the expensive ones — census resampling (3 Philox blocks, sqrt, sincos), a leak's
// or an albedo rejection's cosine-law direction (Philox, 2 sqrt, sincos), the IMC scatter — each taken by part of
// the warp (profiles/r01_v4_ncu_transport_ddmc_cells64.md: 9.8 of 32 lanes). Here, as in k_transport_pool, the
// photons live in shared memory and a warp runs one PHASE per iteration, each (nearly) full width:
//   * retire + refill: absorbed photons deposit, census photons go back to the front of their chunk (the pool
//     stays ordered, see jb_transport_pool.cuh), freed slots take the next photons of the chunk;
//   * deferred work: everything that ends in "draw (mu, phi), build a direction" — IMC scatter, DDMC leak, albedo
//     rejection, DDMC census resampling (which first draws x, y) — shares ONE sqrt + sincos evaluation, selected by
//     a per-slot mode; a photon that left its meshblock (1-2 % of the segments) is handled here too, after its
//     direction is known (swarm BCs, neighbour lookup, block-face resampling, outbox record);
//   * segment: one iteration of the while (t < t_end) loop up to the point where the outcome is known — the IMC
//     step, or the DDMC albedo test + leak / absorb / census selection.
#ifndef JAYBENNE_B200_JB_TRANSPORT_HPOOL_CUH_
#define JAYBENNE_B200_JB_TRANSPORT_HPOOL_CUH_

// deferred-work modes (low 3 bits of HPool::mode; bits 3-4 = axis, bit 5 = upper face)
enum { HM_SCAT = 0, HM_LEAK = 1, HM_REJECT = 2, HM_CENSUS = 3, HM_EXIT = 4 };
// census resamplings wait in a state of their own: they are half of a DDMC run's deferred work and take one draw
// pair and an Xtoijk more than the others, so serving the two kinds in separate rounds keeps both full width
// (mixed: 19 of 32 lanes, profiles/r02_v12_ncu_transport_hpool_cells64.md)
enum { PS_PEND2 = 5 };

template <int S>
struct __align__(16) HPool {
  double x[S], y[S], z[S], t[S], vx[S], vy[S], vz[S];
  double w[S];       // weight
  double e[S];       // photon frequency (swarm field `energy`)
  double bpref[S];   // B_nu prefactor 2 h nu^3 / c^2 (photon_freq)
  uint2 ida[S];      // (birth index, birth gid)
  uint2 idb[S];      // (source call, draws consumed)
  uint2 carry[S];    // words 2, 3 of Philox block n >> 1 while n is odd (Rng2)
  int2 blk[S];       // (local block, gid)
  unsigned idx[S];   // swarm slot the photon was loaded from, relative to `first` (tells its chunk)
  int cell[S];       // (ip + 1) | (jp + 1) << 10 | (kp + 1) << 20: one cell beyond the block fits
  unsigned char st[64];
  unsigned char mode[64];
  unsigned char list[64];
};

__device__ __forceinline__ int hp_pack(int ip, int jp, int kp) { return (ip + 1) | ((jp + 1) << 10) | ((kp + 1) << 20); }
__device__ __forceinline__ void hp_unpack(int c, int &ip, int &jp, int &kp) {
  ip = (c & 1023) - 1;
  jp = ((c >> 10) & 1023) - 1;
  kp = ((c >> 20) & 1023) - 1;
}

// ptcl_ddmc_albedo (transport_utils.hpp:279-397) up to the decision: the arithmetic and the single draw of
// ddmc_albedo2; a rejected photon is put eps_imc outside the face, its direction is sampled later (HM_REJECT).
// Returns -1 (not at a face, or admitted) or the face 0..5 it was rejected at.
template <int NDIM>
__device__ __forceinline__ int hp_albedo(Rng2 &rng, const PKeys &K, const CellEnv &c, const double vv, Photon &p) {
  constexpr bool multi_d = NDIM >= 2, three_d = NDIM == 3;
  const double dx = c.xu - c.xl, dy = c.yu - c.yl, dz = c.zu - c.zl;
  const double tol = 2.5 * EPS_IMC;
  const int face = fuzzy_eq(p.x, c.xl, dx, tol) ? 0
                 : fuzzy_eq(p.x, c.xu, dx, tol) ? 1
                 : (fuzzy_eq(p.y, c.yl, dy, tol) && multi_d) ? 2
                 : (fuzzy_eq(p.y, c.yu, dy, tol) && multi_d) ? 3
                 : (fuzzy_eq(p.z, c.zl, dz, tol) && three_d) ? 4
                 : (fuzzy_eq(p.z, c.zu, dz, tol) && three_d) ? 5 : -1;
  int rejected = -1;
  if (face >= 0) {
    const int axis = face >> 1;
    const bool upper = (face & 1) != 0;
    const double dcell = axis == 0 ? dx : (axis == 1 ? dy : dz);
    const double vn = axis == 0 ? p.vx : (axis == 1 ? p.vy : p.vz);
    const double sign_in = upper ? -1.0 : 1.0;
    const double Pf = jbm::div_normal(2.0 / 3.0, (c.aa + c.ss) * dcell + 2.0 * LAM_EXT);  // denominator >= 1.42
    const double P = 2.0 * Pf * (1.0 + jbm::div_normal(sign_in * 1.5 * vn, vv));
    if (rng2_draw1(rng, K) > P) {
      const double lo = axis == 0 ? c.xl : (axis == 1 ? c.yl : c.zl);
      const double hi = axis == 0 ? c.xu : (axis == 1 ? c.yu : c.zu);
      const double out = upper ? hi + EPS_IMC * dcell : lo - EPS_IMC * dcell;
      p.x = axis == 0 ? out : p.x;
      p.y = axis == 1 ? out : p.y;
      p.z = axis == 2 ? out : p.z;
      rejected = face;
    }
  }
  if (rejected < 0) {
    p.x = 0.5 * (c.xl + c.xu);
    p.y = 0.5 * (c.yl + c.yu);
    p.z = 0.5 * (c.zl + c.zu);
  }
  return rejected;
}

// ptcl_ddmc_step (transport_utils.hpp:163-277) up to the point where the outcome is known: the arithmetic and the
// first draw pair of ddmc_step2. Outcomes: 0 nothing further (no event selected), 1 absorbed, 2 leak through
// `face` (position and cell index moved; direction later: HM_LEAK), 3 census (t and z set; x, y and the direction
// later: HM_CENSUS).
template <int NDIM>
__device__ __forceinline__ int hp_ddmc_core(Rng2 &rng, const PKeys &K, const double t_end, const CellEnv &c,
                                            const double vv, const double P[6], Photon &p, int &face) {
  constexpr int multi_d = NDIM >= 2 ? 1 : 0, three_d = NDIM == 3 ? 1 : 0;
  const double rmin = DBL_MIN;
  const double eps = EPS_DDMC;
  const double dx = c.xu - c.xl, dy = c.yu - c.yl, dz = c.zu - c.zl;
  const double leakx_l = jbm::div_normal(P[0], dx), leakx_u = jbm::div_normal(P[3], dx);
  const double leaky_l = jbm::div_normal(P[1], dy), leaky_u = jbm::div_normal(P[4], dy);
  const double leakz_l = jbm::div_normal(P[2], dz), leakz_u = jbm::div_normal(P[5], dz);
  const double th1 = leakx_l + leakx_u;
  const double th2 = th1 + leaky_l;
  const double th3 = th2 + leaky_u;
  const double th4 = th3 + leakz_l;
  const double leak_tot = th4 + leakz_u;
  const double cdf_ddmc = c.ff * c.aa + leak_tot + rmin;
  double xi_dt, xi_b;  // the second draw is xi of an event or z of the census resampling
  rng.draw2(K, xi_dt, xi_b);
  const double dt_ddmc = -jbm::log_unit(xi_dt) / (vv * cdf_ddmc);
  const double dt_end = t_end - p.t;
  const bool is_ddmc_event = dt_ddmc < dt_end;
  const double dt_push = jbm::min2(dt_ddmc, dt_end);
  p.t += dt_push;
  face = -1;
  if (!is_ddmc_event) {
    p.z = c.zl + xi_b * dz;
    return 3;
  }
  const double xi = cdf_ddmc * xi_b;
  if (xi < c.ff * c.aa) return 1;
  if (xi < c.ff * c.aa + leak_tot) {
    const double xim = xi - c.ff * c.aa;
    face = xim < leakx_l ? 0 : (xim < th1 ? 1 : (xim < th2 ? 2 : (xim < th3 ? 3 : (xim < th4 ? 4 : (xim <= leak_tot ? 5 : -1)))));
    if (face >= 0) {
      const int axis = face >> 1;
      const bool upper = (face & 1) != 0;
      const int d = upper ? 1 : -1;
      p.ip += axis == 0 ? d : 0;
      p.jp += axis == 1 ? d * multi_d : 0;
      p.kp += axis == 2 ? d * three_d : 0;
      const double fx = upper ? c.xu + eps * dx : c.xl - eps * dx;
      const double fy = upper ? c.yu + eps * dy : c.yl - eps * dy;
      const double fz = upper ? c.zu + eps * dz : c.zl - eps * dz;
      p.x = axis == 0 ? fx : c.xl + 0.5 * dx;
      p.y = axis == 1 ? fy : c.yl + 0.5 * dy;
      p.z = axis == 2 ? fz : c.zl + 0.5 * dz;
      return 2;
    }
  }
  return 0;
}

template <int NDIM, int MINB, int S, int ABS>
__global__ void __launch_bounds__(POOL_TPB, MINB)
    k_transport_hpool(StepParams P, long long first, long long n, double t_start, double dt, int refill_min,
                      int ready_low) {
  static_assert(S > 32 && S <= 64, "two slots per lane");
  constexpr bool multi_d = NDIM >= 2, three_d = NDIM == 3;
  __shared__ HPool<S> pools[POOL_WARPS];
  __shared__ unsigned sh_cnt[8];  // crossings, absorbed, dead, escaped, sent, error, overflow

  const DMesh &m = P.mesh;
  if (P.range) {  // direct peer exchange: the arrivals' range lives on the device (P2PState::range)
    first = P.range[0];
    n = P.range[1];
  }
  const double t_end = t_start + dt;
  const double vv = P.cl.c_code;
  const int nx1 = m.nxb[0], nx2 = m.nxb[1];
  const unsigned FULL = 0xffffffffu;
  const unsigned lane = threadIdx.x & 31u;
  const unsigned lt = (1u << lane) - 1u;
  const PKeys &K = P.rk;
  HPool<S> &W = pools[threadIdx.x >> 5];

  W.st[lane] = PS_FREE;
  W.st[lane + 32u] = (lane + 32u < (unsigned)S) ? PS_FREE : PS_NONE;
  if (threadIdx.x < 8) sh_cnt[threadIdx.x] = 0u;
  __syncthreads();

  unsigned cur = 0, end = 0;    // read cursor and end of the current chunk (relative to `first`)
  unsigned c_base = 0, wr = 0;  // current chunk: first slot and write cursor
  unsigned p_wr = 0;            // write cursor of the previous chunk
  bool exhausted = false;
  bool flip = false;
  unsigned c_seg = 0, c_ddmc = 0, c_scat = 0;
  // energies stay per lane and are folded by a shuffle tree at the end: stiff problems absorb ~all of a large
  // emitted energy, and the 1e-12 conservation gate is relative to the small census energy
  double e_abs = 0.0, e_abs_c = 0.0, e_esc = 0.0;  // e_abs with a Kahan compensation term

  for (;;) {
    __syncwarp();
    const int s0 = W.st[lane], s1 = W.st[lane + 32u];
    const unsigned r0 = __ballot_sync(FULL, s0 == PS_READY), r1 = __ballot_sync(FULL, s1 == PS_READY);
    const unsigned q0 = __ballot_sync(FULL, s0 == PS_PEND), q1 = __ballot_sync(FULL, s1 == PS_PEND);
    const unsigned z0 = __ballot_sync(FULL, s0 == PS_PEND2), z1 = __ballot_sync(FULL, s1 == PS_PEND2);
    const unsigned f0 = __ballot_sync(FULL, s0 == PS_FREE), f1 = __ballot_sync(FULL, s1 == PS_FREE);
    const unsigned w0 = __ballot_sync(FULL, s0 == PS_CENSUS || s0 == PS_ABS);
    const unsigned w1 = __ballot_sync(FULL, s1 == PS_CENSUS || s1 == PS_ABS);
    const int n_ready = __popc(r0) + __popc(r1), n_pend = __popc(q0) + __popc(q1), n_pend2 = __popc(z0) + __popc(z1);
    const int n_ret = __popc(w0) + __popc(w1), n_nonlive = n_ret + __popc(f0) + __popc(f1);
    bool can_claim = !exhausted;  // (two write cursors: see k_transport_pool)
    if (cur >= end && can_claim) {
      const bool old0 = s0 != PS_FREE && W.idx[lane] < c_base;
      const bool old1 = s1 != PS_FREE && s1 != PS_NONE && W.idx[lane + 32u] < c_base;
      can_claim = !__any_sync(FULL, old0 || old1);
    }
    const bool input_left = cur < end || can_claim;
    const bool idle = n_ready + n_pend + n_pend2 == 0;

    // =====================================================================================
    // retire + refill
    // =====================================================================================
    if ((n_nonlive >= refill_min || idle) && (input_left || n_ret > 0)) {
      int cnt;
      const int s = pool_pick(W.list, input_left ? (w0 | f0) : w0, input_left ? (w1 | f1) : w1, false, lane, lt, cnt);
      {
        const int old = s >= 0 ? (int)W.st[s] : PS_FREE;
        const bool cen = old == PS_CENSUS;
        const bool in_cur = cen && W.idx[s] >= c_base;
        const unsigned m_cur = __ballot_sync(FULL, in_cur), m_prev = __ballot_sync(FULL, cen && !in_cur);
        if (cen) {
          const long long idx = first + (long long)(in_cur ? wr + __popc(m_cur & lt) : p_wr + __popc(m_prev & lt));
          int ip, jp, kp;
          hp_unpack(W.cell[s], ip, jp, kp);
          P.s.x[idx] = W.x[s]; P.s.y[idx] = W.y[s]; P.s.z[idx] = W.z[s]; P.s.t[idx] = W.t[s];
          P.s.vx[idx] = W.vx[s]; P.s.vy[idx] = W.vy[s]; P.s.vz[idx] = W.vz[s];
          P.s.w[idx] = W.w[s]; P.s.e[idx] = W.e[s];
          P.s.loc[idx] = make_int2(W.blk[s].x, (kp * nx2 + jp) * nx1 + ip);
          P.s.ida[idx] = W.ida[s];
          P.s.idb[idx] = W.idb[s];
        } else if (old == PS_ABS) {
          // absorbed (transport_ddmc.cpp:214-219): deposit; the slot became a hole when the photon was loaded
          int ip, jp, kp;
          hp_unpack(W.cell[s], ip, jp, kp);
          const double ww = W.w[s];
          atomicAdd(&P.f.edelta[(long long)W.blk[s].x * m.ncell + ((long long)kp * nx2 + jp) * nx1 + ip], ww);
          atomicAdd(&sh_cnt[1], 1u);
          atomicAdd(&sh_cnt[2], 1u);
          const double y = ww - e_abs_c, tt = e_abs + y;
          e_abs_c = (tt - e_abs) - y;
          e_abs = tt;
        }
        wr += (unsigned)__popc(m_cur);
        p_wr += (unsigned)__popc(m_prev);
      }
      bool in_chunk = false;
      unsigned my32 = 0;
      if (input_left) {
        if (cur >= end) {
          unsigned long long c0 = 0;
          unsigned piece = IMC_CHUNK;  // guided claim, see pool_claim
          if (lane == 0) c0 = pool_claim(P.work, n, (long long)gridDim.x * POOL_WARPS, false, piece);
          c0 = __shfl_sync(FULL, c0, 0);
          piece = __shfl_sync(FULL, piece, 0);
          if ((long long)c0 >= n) {
            exhausted = true;
          } else {
            p_wr = wr;
            cur = (unsigned)c0;
            c_base = wr = cur;
            end = (long long)c0 + piece > n ? (unsigned)n : cur + piece;
          }
        }
        if (cur < end) {
          my32 = cur + lane;
          in_chunk = s >= 0 && my32 < end;
          cur = min(end, cur + (unsigned)cnt);
        }
      }
      int newst = PS_FREE;
      if (in_chunk) {
        const long long my = first + (long long)my32;
        const int2 lc = ld_once(&P.s.loc[my]);
        const double t0 = ld_once(&P.s.t[my]);
        const double l_x = ld_once(&P.s.x[my]), l_y = ld_once(&P.s.y[my]), l_z = ld_once(&P.s.z[my]);
        const double l_vx = ld_once(&P.s.vx[my]), l_vy = ld_once(&P.s.vy[my]), l_vz = ld_once(&P.s.vz[my]);
        const double l_e = ld_once(&P.s.e[my]), l_w = ld_once(&P.s.w[my]);
        const uint2 ia = ld_once(&P.s.ida[my]), ib = ld_once(&P.s.idb[my]);
        if (lc.x >= 0) {
          P.s.loc[my] = make_int2(-1, 0);  // the photon lives in the pool from here on
          W.idx[s] = my32;
          W.t[s] = t0;
          W.x[s] = l_x; W.y[s] = l_y; W.z[s] = l_z;
          W.vx[s] = l_vx; W.vy[s] = l_vy; W.vz[s] = l_vz;
          W.w[s] = l_w;
          W.e[s] = l_e;
          W.bpref[s] = photon_freq(P.cl, l_e).bpref;
          W.ida[s] = ia;
          W.idb[s] = ib;
          uint2 cy = make_uint2(0u, 0u);
          if (ib.y & 1u) {  // an odd draw counter carries the second half of block n >> 1 (Rng2)
            Rng2 g;
            g.id0 = ia.x; g.id1 = ia.y; g.id2 = ib.x;
            uint32_t w0_, w1_;
            g.block(ib.y >> 1, K, w0_, w1_, cy.x, cy.y);
          }
          W.carry[s] = cy;
          const int gid = m.local_gid[lc.x];
          W.blk[s] = make_int2(lc.x, gid);
          // Xtoijk after comms (transport_ddmc.cpp:98)
          const DBlock &b = m.blocks[gid];
          int ip = (int)floor(jbm::div_normal(l_x - b.xmin[0], b.dx[0]));
          int jp = multi_d ? (int)floor(jbm::div_normal(l_y - b.xmin[1], b.dx[1])) : 0;
          int kp = three_d ? (int)floor(jbm::div_normal(l_z - b.xmin[2], b.dx[2])) : 0;
          newst = PS_READY;
          if (!(t0 < t_end)) {  // already at census (transport_ddmc.cpp:100 skips it): passes through with its stored cell
            ip = lc.y % nx1;
            jp = (lc.y / nx1) % nx2;
            kp = lc.y / (nx1 * nx2);
            newst = PS_CENSUS;
          }
          W.cell[s] = hp_pack(ip, jp, kp);
        }
      }
      if (s >= 0) W.st[s] = (unsigned char)newst;
      continue;
    }
    if (idle) break;  // no live photon, nothing to retire, no input left

    // =====================================================================================
    // deferred work: a direction from the draw pair (mu, phi) — IMC scatter (scattering.hpp:21-29), DDMC leak
    // and albedo rejection (cosine law about the face normal, transport_utils.hpp:26-39), DDMC census
    // resampling (x, y drawn first; transport_utils.hpp:264-276) — and photons that left their meshblock
    // =====================================================================================
    if (n_pend >= 32 || n_pend2 >= 32 || (n_pend + n_pend2 > 0 && n_ready < ready_low)) {
      int cnt;
      const bool census_round = n_pend2 >= n_pend;  // the fuller of the two kinds
      const int s = pool_pick(W.list, census_round ? z0 : q0, census_round ? z1 : q1, false, lane, lt, cnt);
      if (s >= 0) {
        const int md = W.mode[s];
        int mode = md & 7;
        const int axis = (md >> 3) & 3;
        const bool upper = (md >> 5) & 1;
        Photon p;
        Rng2 rng;
        {
          const uint2 ia = W.ida[s], ib = W.idb[s], cy = W.carry[s];
          rng.id0 = ia.x; rng.id1 = ia.y; rng.id2 = ib.x; rng.n = ib.y; rng.clo = cy.x; rng.chi = cy.y;
        }
        const int2 bk = W.blk[s];
        int lb = bk.x, gid = bk.y;
        hp_unpack(W.cell[s], p.ip, p.jp, p.kp);
        p.x = W.x[s]; p.y = W.y[s]; p.z = W.z[s]; p.t = W.t[s];
        p.vx = W.vx[s]; p.vy = W.vy[s]; p.vz = W.vz[s];
        // draw pairs: (x, y) of a census resampling first, then (mu, phi) — ONE copy of the generator's code in
        // this phase (the kernel's hot instructions have to fit the 32 KB instruction cache, see the header)
        double xi_mu = 0.0, xi_phi = 0.0;
#pragma unroll 1
        for (int pair = mode == HM_CENSUS ? 0 : (mode != HM_EXIT ? 1 : 2); pair < 2; ++pair) {
          double a, b;
          rng.draw2(K, a, b);
          if (pair == 0) {
            // x, y inside the cell (z was set with the second draw of the step's first pair)
            const BlkGeo *geo = m.geo + gid;
            const double2 gx = __ldg(&geo->xd[0]), gy = __ldg(&geo->xd[1]);
            const double xcen = gx.x + ((p.ip + m.ng[0]) + 0.5) * gx.y;
            const double ycen = gy.x + ((p.jp + m.ng[1]) + 0.5) * gy.y;
            const double xl = xcen - 0.5 * gx.y, xu = xcen + 0.5 * gx.y;
            const double yl = ycen - 0.5 * gy.y, yu = ycen + 0.5 * gy.y;
            p.x = xl + a * (xu - xl);
            p.y = yl + b * (yu - yl);
          } else {
            xi_mu = a;
            xi_phi = b;
          }
        }
        if (mode != HM_EXIT) {
          // the three samplers share phi, sqrt(1 - mu^2) and the sincos
          const double mu = mode == HM_SCAT ? 2.0 * xi_mu - 1.0 : (mode == HM_CENSUS ? 1.0 - 2.0 * xi_mu : sqrt(xi_mu));
          const double nu = sqrt(1.0 - mu * mu);
          const double phi = 2.0 * JB_PI * xi_phi;
          double sn, cs;
          jbm::sincos(phi, &sn, &cs);
          const double V = (mode == HM_LEAK || mode == HM_REJECT) ? (upper ? vv : -vv) : vv;
          const double a1 = V * mu, a2 = V * nu * cs, a3 = V * nu * sn;
          if (mode == HM_SCAT) {          // scatter_iso2: (v sin cos, v sin sin, v mu)
            p.vx = a2; p.vy = a3; p.vz = a1;
            c_scat++;
          } else if (mode == HM_CENSUS) {  // ddmc_step2 census branch: vz = v mu, vx = v nu cos, vy = v nu sin
            p.vz = a1; p.vx = a2; p.vy = a3;
            // Xtoijk (transport_ddmc.cpp:198): the position was redrawn
            const DBlock &b = m.blocks[gid];
            p.ip = (int)floor(jbm::div_normal(p.x - b.xmin[0], b.dx[0]));
            p.jp = multi_d ? (int)floor(jbm::div_normal(p.y - b.xmin[1], b.dx[1])) : 0;
            p.kp = three_d ? (int)floor(jbm::div_normal(p.z - b.xmin[2], b.dx[2])) : 0;
          } else {                         // face_iso_dir2 + unpermute_dir
            unpermute_dir(axis, a1, a2, a3, p.vx, p.vy, p.vz);
          }
        }
        int newst = PS_READY;
        if (!on_block(m, p.ip, p.jp, p.kp)) {
          // ---- left the meshblock (transport_ddmc.cpp:203-211): swarm BCs + hand-over ----
          // zero velocity tags a photon a DDMC step carried out (is_ddmc_step && !is_rejected) for block-face resampling
          const double vmask = !((mode == HM_LEAK || mode == HM_CENSUS) && multi_d) ? 1.0 : 0.0;
          p.vx *= vmask;
          p.vy *= vmask;
          p.vz *= vmask;
          bool gone = false;
          const int bx = p.ip < 0 ? -1 : (p.ip >= nx1 ? 1 : 0);
          const int by = p.jp < 0 ? -1 : (p.jp >= nx2 ? 1 : 0);
          const int bz = p.kp < 0 ? -1 : (p.kp >= m.nxb[2] ? 1 : 0);
          if (!apply_swarm_bc<NDIM>(m, p)) {
            atomicAdd(&sh_cnt[3], 1u);
            e_esc += W.w[s];
            gone = true;
          } else {
            // same-level neighbour from the 27-entry table (as k_transport_pool); the general lattice search only
            // for level changes and non-periodic faces. Xtoijk in the new block is evaluated either way.
            int ngid = __ldg(&m.nbr[lb * 27 + (bz + 1) * 9 + (by + 1) * 3 + (bx + 1)]);
            if (ngid < 0) {
              const double pos[3] = {p.x, p.y, p.z};
              ngid = find_block(m, pos);
            }
            if (ngid < 0) {
              atomicAdd(&sh_cnt[5], 1u);
              gone = true;
            } else {
              atomicAdd(&sh_cnt[0], 1u);
              const DBlock &nb = m.blocks[ngid];
              if (nb.rank != m.my_rank) {
                const unsigned long long slot = atomicAdd(&P.out.count[nb.rank], 1ull);
                if ((long long)slot < P.out.cap) {
                  ExRecord r;
                  r.x = p.x; r.y = p.y; r.z = p.z; r.t = p.t; r.w = W.w[s]; r.e = W.e[s];
                  r.vx = p.vx; r.vy = p.vy; r.vz = p.vz;
                  r.rng[0] = rng.id0; r.rng[1] = rng.id1; r.rng[2] = rng.id2; r.rng[3] = rng.n;
                  r.dest_gid = ngid;
                  r.pad = 0;
                  P.out.rec[(size_t)nb.rank * P.out.cap + slot] = r;
                  atomicAdd(&sh_cnt[4], 1u);
                } else {
                  atomicAdd(&sh_cnt[6], 1u);
                }
                gone = true;
              } else {
                gid = ngid;
                lb = nb.local;
                BlkReg NB;
                load_block(m, gid, NB);
                xtoijk<NDIM>(NB, p, p.ip, p.jp, p.kp);
                Rng g;  // the block-face sampler (rare) runs on the per-draw generator, same stream
                g.init(rng.id0, rng.id1, rng.id2, rng.n, P.key0, P.key1_particle);
                bface_sample<NDIM>(P, NB, p, g);
                if (g.n != rng.n) rng.load(rng.id0, rng.id1, rng.id2, g.n, K);
                W.blk[s] = make_int2(lb, gid);
              }
            }
          }
          if (gone) {
            atomicAdd(&sh_cnt[2], 1u);
            newst = PS_FREE;
          }
        }
        if (newst == PS_READY && !(p.t < t_end)) newst = PS_CENSUS;
        W.x[s] = p.x; W.y[s] = p.y; W.z[s] = p.z;
        W.vx[s] = p.vx; W.vy[s] = p.vy; W.vz[s] = p.vz;
        W.cell[s] = hp_pack(p.ip, p.jp, p.kp);
        W.idb[s].y = rng.n;
        W.carry[s] = make_uint2(rng.clo, rng.chi);
        W.st[s] = (unsigned char)newst;
      }
      continue;
    }

    // =====================================================================================
    // segment: one iteration of the while (t < t_end) loop, transport_ddmc.cpp:100-227, up to its outcome
    // =====================================================================================
    flip = !flip;
    int cnt;
    const int s = pool_pick(W.list, r0, r1, flip, lane, lt, cnt);
    if (s >= 0) {
      c_seg++;
      Photon p;
      Rng2 rng;
      const int2 bk = W.blk[s];
      const int lb = bk.x, gid = bk.y;
      hp_unpack(W.cell[s], p.ip, p.jp, p.kp);
      const long long ci = (long long)lb * m.ncell + ((long long)p.kp * nx2 + p.jp) * nx1 + p.ip;
      const CellProps cp = load_props(P.f.props + ci);
      const BlkGeo *geo = m.geo + gid;
      const double2 gx = __ldg(&geo->xd[0]), gy = __ldg(&geo->xd[1]), gz = __ldg(&geo->xd[2]);
      const double dx_cell = __ldg(&geo->dx_push);
      {
        const uint2 ia = W.ida[s], ib = W.idb[s], cy = W.carry[s];
        rng.id0 = ia.x; rng.id1 = ia.y; rng.id2 = ib.x; rng.n = ib.y; rng.clo = cy.x; rng.chi = cy.y;
      }
      p.x = W.x[s]; p.y = W.y[s]; p.z = W.z[s]; p.t = W.t[s];
      p.vx = W.vx[s]; p.vy = W.vy[s]; p.vz = W.vz[s];
      CellEnv c;
      {
        const double xcen = gx.x + ((p.ip + m.ng[0]) + 0.5) * gx.y;
        const double ycen = gy.x + ((p.jp + m.ng[1]) + 0.5) * gy.y;
        const double zcen = gz.x + ((p.kp + m.ng[2]) + 0.5) * gz.y;
        c.xl = xcen - 0.5 * gx.y; c.xu = xcen + 0.5 * gx.y;
        c.yl = ycen - 0.5 * gy.y; c.yu = ycen + 0.5 * gy.y;
        c.zl = zcen - 0.5 * gz.y; c.zu = zcen + 0.5 * gz.y;
      }
      c.ff = cp.ff;
      c.ss = cp.ss;
      if (ABS == ABS_GRAY_KIRCHHOFF) {
        // absorption_coeff_hoisted() with the gray branch fixed (as k_transport_pool)
        const double hnu = P.cl.h * (W.e[s] * P.cl.freq_unit), bpref = W.bpref[s];
        const double r = P.cl.rho_unit * cp.rho, T = P.cl.temp_unit * cp.temp;
        const double x = hnu / (P.cl.kb * T);
        const double Bnu = bpref / jbm::expm1(x);
        const double jnu = r * P.cl.kappa * Bnu;
        c.aa = robust_ratio_pos(jnu, Bnu) * P.cl.length_unit;
      } else {
        PhotonFreq q;
        q.f = W.e[s] * P.cl.freq_unit; q.hnu = P.cl.h * q.f; q.bpref = W.bpref[s];
        c.aa = absorption_coeff_hoisted(P.cl, cp.rho, cp.temp, q);
      }
      int newst = PS_READY, md = 0;
      const bool is_ddmc_step = dx_cell * (c.ss + c.aa) > P.tau_ddmc;  // transport_ddmc.cpp:135
      if (is_ddmc_step) {
        c_ddmc++;
        const double *fp0 = P.f.fp[0] + (size_t)lb * P.f.nface[0];
        double Pf[6];
        Pf[0] = fp0[((size_t)p.kp * nx2 + p.jp) * (nx1 + 1) + p.ip];
        Pf[3] = fp0[((size_t)p.kp * nx2 + p.jp) * (nx1 + 1) + p.ip + 1];
        if (multi_d) {
          const double *fp1 = P.f.fp[1] + (size_t)lb * P.f.nface[1];
          Pf[1] = fp1[((size_t)p.kp * (nx2 + 1) + p.jp) * nx1 + p.ip];
          Pf[4] = fp1[((size_t)p.kp * (nx2 + 1) + p.jp + 1) * nx1 + p.ip];
        } else {
          Pf[1] = 0.0; Pf[4] = 0.0;
        }
        if (three_d) {
          const double *fp2 = P.f.fp[2] + (size_t)lb * P.f.nface[2];
          Pf[2] = fp2[((size_t)p.kp * nx2 + p.jp) * nx1 + p.ip];
          Pf[5] = fp2[((size_t)(p.kp + 1) * nx2 + p.jp) * nx1 + p.ip];
        } else {
          Pf[2] = 0.0; Pf[5] = 0.0;
        }
        const int rej = hp_albedo<NDIM>(rng, K, c, vv, p);
        if (rej >= 0) {
          // pushed back out of the cell: Xtoijk (transport_ddmc.cpp:198), direction in the deferred phase
          const DBlock &b = m.blocks[gid];
          p.ip = (int)floor(jbm::div_normal(p.x - b.xmin[0], b.dx[0]));
          p.jp = multi_d ? (int)floor(jbm::div_normal(p.y - b.xmin[1], b.dx[1])) : 0;
          p.kp = three_d ? (int)floor(jbm::div_normal(p.z - b.xmin[2], b.dx[2])) : 0;
          newst = PS_PEND;
          md = HM_REJECT | (rej >> 1) << 3 | (rej & 1) << 5;
        } else {
          int face;
          const int outcome = hp_ddmc_core<NDIM>(rng, K, t_end, c, vv, Pf, p, face);
          if (outcome == 1) {
            newst = PS_ABS;
          } else if (outcome == 2) {
            newst = PS_PEND;
            md = HM_LEAK | (face >> 1) << 3 | (face & 1) << 5;
          } else if (outcome == 3) {
            newst = PS_PEND2;
            md = HM_CENSUS;
          }
        }
      } else {
        double xi_abs, xi_sc;
        rng.draw2(K, xi_abs, xi_sc);
        int di, dj, dk;
        unsigned bad_div = 0;
        bool is_absorbed = false, is_scattered = false;
        imc_step_lockstep<NDIM>(xi_abs, xi_sc, t_end, c, vv, dx_cell, p, is_absorbed, is_scattered, di, dj, dk, bad_div);
        p.ip += di; p.jp += dj; p.kp += dk;  // the face nudges move the index exactly (see k_transport_imc)
        if (!on_block(m, p.ip, p.jp, p.kp)) {
          newst = PS_PEND;
          md = HM_EXIT;
        } else if (is_absorbed) {
          newst = PS_ABS;
        } else if (is_scattered) {
          newst = PS_PEND;
          md = HM_SCAT;
        }
      }
      if (newst == PS_READY && !(p.t < t_end)) newst = PS_CENSUS;
      W.x[s] = p.x; W.y[s] = p.y; W.z[s] = p.z; W.t[s] = p.t;
      W.cell[s] = hp_pack(p.ip, p.jp, p.kp);
      W.idb[s].y = rng.n;
      W.carry[s] = make_uint2(rng.clo, rng.chi);
      W.mode[s] = (unsigned char)md;
      W.st[s] = (unsigned char)newst;
    }
  }
  DStats *St = P.stats;
  warp_add_u64(&St->n_segments, c_seg);
  warp_add_u64(&St->n_scatters, c_scat);
  warp_add_u64(&St->n_ddmc, c_ddmc);
  __syncthreads();
  if (threadIdx.x == 0) {
    if (sh_cnt[0]) atomicAdd(&St->n_crossings, (unsigned long long)sh_cnt[0]);
    if (sh_cnt[1]) atomicAdd(&St->n_absorbed, (unsigned long long)sh_cnt[1]);
    if (sh_cnt[2]) atomicAdd(&St->n_dead, (unsigned long long)sh_cnt[2]);
    if (sh_cnt[3]) atomicAdd(&St->n_escaped, (unsigned long long)sh_cnt[3]);
    if (sh_cnt[4]) atomicAdd(&St->n_sent, (unsigned long long)sh_cnt[4]);
    if (sh_cnt[5]) atomicAdd(&St->n_error, (unsigned long long)sh_cnt[5]);
    if (sh_cnt[6]) atomicAdd(&St->n_overflow, (unsigned long long)sh_cnt[6]);
  }
  warp_add_energy(P, EB_ABSORBED, e_abs);
  if (__any_sync(FULL, e_esc != 0.0)) warp_add_energy(P, EB_ESCAPED, e_esc);
}

struct HPoolTuning {
  int ctas;        // resident CTAs per SM the kernel is compiled for (registers: 65536 / (128 * ctas))
  int refill_min, ready_low;
};
static HPoolTuning g_hpool = {0, 24, 16};  // measured: profiles/r02_hpool_first_ab_cells128.md

template <int NDIM, int MINB, int S>
static void launch_transport_hpool_b(const StepParams &P, long long first, long long n, double t_start, double dt,
                                     cudaStream_t st) {
  static int n_sm = 0;
  if (n_sm == 0) {
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&n_sm, cudaDevAttrMultiProcessorCount, dev);
    if (n_sm <= 0) n_sm = 148;
  }
  long long blocks = (long long)n_sm * MINB;
  const long long need = (n + (long long)POOL_WARPS * S - 1) / ((long long)POOL_WARPS * S);
  if (blocks > need) blocks = need;
  if (blocks < 1) blocks = 1;
  static bool carved = false;
  if (!carved) {
    carved = true;
    // the shared-memory carve-out the pools need and no more: the rest of the SM's 256 KB is L1 for the per-cell
    // records and face probabilities (as in k_transport_pool)
    const long long per_cta = (long long)sizeof(HPool<S>) * POOL_WARPS + 32 + 1024;  // + the driver's 1 KB per CTA
    int pct = (int)((per_cta * MINB * 100 + 233471) / 233472);
    if (pct > 100) pct = 100;
    cudaFuncSetAttribute(k_transport_hpool<NDIM, MINB, S, ABS_GENERIC>, cudaFuncAttributePreferredSharedMemoryCarveout, pct);
    cudaFuncSetAttribute(k_transport_hpool<NDIM, MINB, S, ABS_GRAY_KIRCHHOFF>, cudaFuncAttributePreferredSharedMemoryCarveout, pct);
  }
  if (P.cl.opacity_model == JB_OPAC_GRAY && P.cl.gray_mode != JB_GRAY_DIRECT)
    k_transport_hpool<NDIM, MINB, S, ABS_GRAY_KIRCHHOFF><<<(unsigned)blocks, POOL_TPB, 0, st>>>(
        P, first, n, t_start, dt, g_hpool.refill_min, g_hpool.ready_low);
  else
    k_transport_hpool<NDIM, MINB, S, ABS_GENERIC><<<(unsigned)blocks, POOL_TPB, 0, st>>>(
        P, first, n, t_start, dt, g_hpool.refill_min, g_hpool.ready_low);
}
// false: the range or the mesh does not fit the kernel's packed indices, or an event log is wanted -> the caller
// takes k_transport_hybrid
template <int NDIM>
static bool launch_transport_hpool(const StepParams &P, long long first, long long n, double t_start, double dt,
                                   cudaStream_t st) {
  if (n >= 0xffffffffLL || P.mesh.nxb[0] > 1022 || P.mesh.nxb[1] > 1022 || P.mesh.nxb[2] > 1022 || P.log.cap > 0) return false;
  if (g_hpool.ctas == 0) {
    const char *v = getenv("JB_HPOOL_CTAS");
    g_hpool.ctas = v ? atoi(v) : 5;
    if (const char *r = getenv("JB_HPOOL_REFILL_MIN")) g_hpool.refill_min = max(1, min(atoi(r), 32));
    if (const char *r = getenv("JB_HPOOL_READY_LOW")) g_hpool.ready_low = max(1, min(atoi(r), 32));
  }
  switch (g_hpool.ctas) {
    case 6: launch_transport_hpool_b<NDIM, 6, 64>(P, first, n, t_start, dt, st); break;
    default: launch_transport_hpool_b<NDIM, 5, 64>(P, first, n, t_start, dt, st); break;
  }
  return true;
}

#endif  // JAYBENNE_B200_JB_TRANSPORT_HPOOL_CUH_
