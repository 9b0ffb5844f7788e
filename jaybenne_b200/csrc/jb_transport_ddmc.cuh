// jb_transport_ddmc.cuh — the hybrid IMC / DDMC history kernel as persistent warps with lane refill.
// Included by jb_kernels.cu after jb_transport_imc.cuh (shares IMC_TPB, IMC_CHUNK, load_props).
//
// Reference behaviour: TransportPhotons_DDMC, src/jaybenne/transport_ddmc.cpp:28-237, with
// ptcl_transport_step / ptcl_ddmc_step / ptcl_ddmc_albedo (transport_utils.hpp:111-397), scatter
// (scattering.hpp:21-29), SampleDDMCBlockFace for on-rank block changes (sample_ddmc_bface.cpp),
// the swarm boundary conditions (boundaries.hpp:24-84) and the hand-over (jaybenne.cpp:26-61).
//
// Why: the thread-per-photon kernel (k_transport<.,true,.>) ran 9.8 of 32 lanes wide at 168
// registers with instruction-fetch stalls (profiles/r01_v4_ncu_transport_ddmc_cells64.md): history
// lengths differ a lot in DDMC (one census jump or many leaks), and the step functions were six
// inlined copies each. Here a warp claims chunks of slots and refills idle lanes every iteration,
// one iteration = one segment of the while (t < t_end) loop for every lane that holds a photon;
// the per-face copies are folded (jb_device.cuh: ddmc_albedo / ddmc_step); the meshblock geometry
// is re-read per segment instead of living in registers; per-cell data comes from the packed
// CellProps record. Arithmetic and draw order are those of k_transport, so results are bit-identical.
#ifndef JAYBENNE_B200_JB_TRANSPORT_DDMC_CUH_
#define JAYBENNE_B200_JB_TRANSPORT_DDMC_CUH_

// ---- the step functions on the paired generator (Rng2, jb_transport_imc.cuh) -------------------
// Every DDMC outcome consumes its draws in pairs — (dt, xi) + (mu, phi) for a leak, (dt, z) + (x, y) +
// (mu, phi) for census — so one Philox block per pair, evaluated by all lanes of the branch together,
// replaces the per-draw cache test of Rng (half the lanes of every call site re-evaluating a block:
// profiles/r01_v4_ncu_transport_ddmc_cells64.md, Philox lines at 17 of 32 lanes). Same streams, same
// draw order, same arithmetic as ddmc_albedo / ddmc_step / face_iso_dir in jb_device.cuh.
__device__ __forceinline__ double rng2_draw1(Rng2 &r, const PKeys &K) {
  double v;
  if (r.n & 1u) {
    v = Rng2::to_double(r.clo, r.chi);  // the carried half of block n >> 1
  } else {
    uint32_t w0, w1, w2, w3;
    r.block(r.n >> 1, K, w0, w1, w2, w3);
    v = Rng2::to_double(w0, w1);
    r.clo = w2;
    r.chi = w3;
  }
  r.n += 1u;
  return v;
}
__device__ __forceinline__ void face_iso_dir2(const double vv, const double xi_mu, const double xi_phi, double &v1,
                                              double &v2, double &v3) {
  const double mu = sqrt(xi_mu);
  const double nu = sqrt(1.0 - mu * mu);
  const double phi = 2.0 * JB_PI * xi_phi;
  double sn, cs;
  jbm::sincos(phi, &sn, &cs);
  v1 = vv * mu;
  v2 = vv * nu * cs;
  v3 = vv * nu * sn;
}
template <int NDIM>
__device__ __forceinline__ void ddmc_albedo2(Rng2 &rng, const PKeys &K, const CellEnv &c,
                                             const double vv, Photon &p, bool &is_rejected) {
  constexpr bool multi_d = NDIM >= 2, three_d = NDIM == 3;
  const double dx = c.xu - c.xl, dy = c.yu - c.yl, dz = c.zu - c.zl;
  const double tol = 2.5 * EPS_IMC;
  const int face = fuzzy_eq(p.x, c.xl, dx, tol) ? 0
                 : fuzzy_eq(p.x, c.xu, dx, tol) ? 1
                 : (fuzzy_eq(p.y, c.yl, dy, tol) && multi_d) ? 2
                 : (fuzzy_eq(p.y, c.yu, dy, tol) && multi_d) ? 3
                 : (fuzzy_eq(p.z, c.zl, dz, tol) && three_d) ? 4
                 : (fuzzy_eq(p.z, c.zu, dz, tol) && three_d) ? 5 : -1;
  if (face >= 0) {
    const int axis = face >> 1;
    const bool upper = (face & 1) != 0;
    const double dcell = axis == 0 ? dx : (axis == 1 ? dy : dz);
    const double vn = axis == 0 ? p.vx : (axis == 1 ? p.vy : p.vz);
    const double sign_in = upper ? -1.0 : 1.0;
    const double Pf = jbm::div_normal(2.0 / 3.0, (c.aa + c.ss) * dcell + 2.0 * LAM_EXT);  // denominator >= 1.42
    const double P = 2.0 * Pf * (1.0 + jbm::div_normal(sign_in * 1.5 * vn, vv));
    if (rng2_draw1(rng, K) > P) {
      double xi_mu, xi_phi, v1, v2, v3;
      rng.draw2(K, xi_mu, xi_phi);
      face_iso_dir2(upper ? vv : -vv, xi_mu, xi_phi, v1, v2, v3);
      unpermute_dir(axis, v1, v2, v3, p.vx, p.vy, p.vz);
      const double lo = axis == 0 ? c.xl : (axis == 1 ? c.yl : c.zl);
      const double hi = axis == 0 ? c.xu : (axis == 1 ? c.yu : c.zu);
      const double out = upper ? hi + EPS_IMC * dcell : lo - EPS_IMC * dcell;
      p.x = axis == 0 ? out : p.x;
      p.y = axis == 1 ? out : p.y;
      p.z = axis == 2 ? out : p.z;
      is_rejected = true;
    }
  }
  if (!is_rejected) {
    p.x = 0.5 * (c.xl + c.xu);
    p.y = 0.5 * (c.yl + c.yu);
    p.z = 0.5 * (c.zl + c.zu);
  }
}
// `resampled` reports the census branch: the position was redrawn inside the cell and the caller has to
// re-derive the cell index the way the reference does (Xtoijk); in every other outcome ip / jp / kp are
// already what Xtoijk returns (cell centre, or DDMC-offset beyond a face with the index moved).
template <int NDIM>
__device__ __forceinline__ void ddmc_step2(Rng2 &rng, const PKeys &K, const double t_end,
                                           const CellEnv &c, const double vv, const double P[6], Photon &p,
                                           bool &is_absorbed, bool &resampled) {
  constexpr int multi_d = NDIM >= 2 ? 1 : 0, three_d = NDIM == 3 ? 1 : 0;
  const double rmin = DBL_MIN;
  const double eps = EPS_DDMC;
  const double dx = c.xu - c.xl, dy = c.yu - c.yl, dz = c.zu - c.zl;
  // face probabilities are in [0, 1], cell widths normal: range-test-free quotients (jb_math.h)
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
  if (is_ddmc_event) {
    const double xi = cdf_ddmc * xi_b;
    if (xi < c.ff * c.aa) {
      is_absorbed = true;
    } else if (xi < c.ff * c.aa + leak_tot) {
      const double xim = xi - c.ff * c.aa;
      const int face = xim < leakx_l ? 0 : (xim < th1 ? 1 : (xim < th2 ? 2 : (xim < th3 ? 3 : (xim < th4 ? 4 : (xim <= leak_tot ? 5 : -1)))));
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
        double xi_mu, xi_phi, v1, v2, v3;
        rng.draw2(K, xi_mu, xi_phi);
        face_iso_dir2(upper ? vv : -vv, xi_mu, xi_phi, v1, v2, v3);
        unpermute_dir(axis, v1, v2, v3, p.vx, p.vy, p.vz);
      }
    }
  } else {
    resampled = true;
    double xi_x, xi_y, xi_mu, xi_phi;
    rng.draw2(K, xi_x, xi_y);
    rng.draw2(K, xi_mu, xi_phi);
    p.z = c.zl + xi_b * dz;
    p.x = c.xl + xi_x * dx;
    p.y = c.yl + xi_y * dy;
    const double mu = 1.0 - 2.0 * xi_mu;
    const double nu = sqrt(1.0 - mu * mu);
    const double phi = 2.0 * JB_PI * xi_phi;
    double sn, cs;
    jbm::sincos(phi, &sn, &cs);
    p.vz = vv * mu;
    p.vx = vv * nu * cs;
    p.vy = vv * nu * sn;
  }
}

template <int NDIM, bool LOG, int MINB>
__global__ void __launch_bounds__(IMC_TPB, MINB)
    k_transport_hybrid(StepParams P, long long first, long long n, double t_start, double dt, int refill_min) {
  const DMesh &m = P.mesh;
  if (P.range) {
    first = P.range[0];
    n = P.range[1];
  }
  constexpr bool multi_d = NDIM >= 2, three_d = NDIM == 3;
  const double t_end = t_start + dt;
  const double vv = P.cl.c_code;
  const int nx1 = m.nxb[0], nx2 = m.nxb[1];
  const unsigned FULL = 0xffffffffu;
  const unsigned lane = threadIdx.x & 31u;
  const unsigned lt_mask = (1u << lane) - 1u;
  long long cur = 0, end = 0;
  bool exhausted = false;

  bool have = false;
  long long idx = 0;
  Photon p;
  PhotonFreq q;
  Rng2 rng;
  int gid = 0;
  p.x = p.y = p.z = p.t = p.vx = p.vy = p.vz = 0.0;
  p.ip = p.jp = p.kp = 0;
  q.f = q.hnu = q.bpref = 0.0;
  rng.id0 = rng.id1 = rng.id2 = rng.n = rng.clo = rng.chi = 0u;
  const PKeys &K = P.rk;

  unsigned c_seg = 0, c_ddmc = 0, c_scat = 0;
  // energies stay per lane and are folded by a shuffle tree at the end: stiff problems absorb ~all of
  // a large emitted energy, and the 1e-12 conservation gate is relative to the small census energy
  double e_abs = 0.0, e_abs_c = 0.0, e_esc = 0.0;  // e_abs with a Kahan compensation term
  __shared__ unsigned sh_cnt[8];            // crossings, absorbed, dead, escaped, sent, error, overflow
  if (threadIdx.x < 8) sh_cnt[threadIdx.x] = 0u;
  __syncthreads();

  for (;;) {
    // ---- refill idle lanes from the warp's chunk ---------------------------------------------
    for (;;) {
      if (cur >= end) {
        if (exhausted) break;
        unsigned long long c0 = 0;
        if (lane == 0) c0 = atomicAdd(P.work, (unsigned long long)IMC_CHUNK);
        c0 = __shfl_sync(FULL, c0, 0);
        if ((long long)c0 >= n) {
          exhausted = true;
          break;
        }
        cur = first + (long long)c0;
        end = cur + IMC_CHUNK;
        if (end > first + n) end = first + n;
      }
      const unsigned need = __ballot_sync(FULL, !have);
      const int cnt = __popc(need);
      if (cnt < refill_min && cnt < 32) break;
      if (cnt == 0) break;
      const long long my = cur + __popc(need & lt_mask);
      cur += cnt;
      if (!have && my < end) {
        // all of the slot's loads ahead of the first test, as in k_transport_imc: one DRAM round trip per refill
        const int2 lc = ld_once(&P.s.loc[my]);
        const double t0 = ld_once(&P.s.t[my]);
        const double l_x = ld_once(&P.s.x[my]), l_y = ld_once(&P.s.y[my]), l_z = ld_once(&P.s.z[my]);
        const double l_vx = ld_once(&P.s.vx[my]), l_vy = ld_once(&P.s.vy[my]), l_vz = ld_once(&P.s.vz[my]);
        const double l_e = ld_once(&P.s.e[my]);
        const uint2 ia = ld_once(&P.s.ida[my]), ib = ld_once(&P.s.idb[my]);
        {
          if (lc.x >= 0 && t0 < t_end) {  // census photons are skipped (transport_ddmc.cpp:100)
            idx = my;
            p.t = t0;
            p.x = l_x; p.y = l_y; p.z = l_z;
            p.vx = l_vx; p.vy = l_vy; p.vz = l_vz;
            q = photon_freq(P.cl, l_e);
            rng.load(ia.x, ia.y, ib.x, ib.y, K);
            gid = m.local_gid[lc.x];
            BlkReg B;
            load_block(m, gid, B);
            xtoijk<NDIM>(B, p, p.ip, p.jp, p.kp);  // transport_ddmc.cpp:98
            have = true;
          }
        }
      }
    }
    if (__ballot_sync(FULL, have) == 0u) break;

    if (have) {
      // ---- one segment, transport_ddmc.cpp:100-227 -------------------------------------------
      const uint32_t draw0 = rng.n;
      c_seg++;
      BlkReg B;
      load_block(m, gid, B);
      const long long ci = (long long)B.lb * m.ncell + ((long long)p.kp * nx2 + p.jp) * nx1 + p.ip;
      const CellProps cp = load_props(P.f.props + ci);
      CellEnv c;
      {
        const double xcen = xc(B, m.ng[0], 0, p.ip), ycen = xc(B, m.ng[1], 1, p.jp), zcen = xc(B, m.ng[2], 2, p.kp);
        c.xl = xcen - 0.5 * B.dx[0]; c.xu = xcen + 0.5 * B.dx[0];
        c.yl = ycen - 0.5 * B.dx[1]; c.yu = ycen + 0.5 * B.dx[1];
        c.zl = zcen - 0.5 * B.dx[2]; c.zu = zcen + 0.5 * B.dx[2];
      }
      c.ff = cp.ff;
      c.ss = cp.ss;
      c.aa = absorption_coeff_hoisted(P.cl, cp.rho, cp.temp, q);
      bool is_absorbed = false, is_scattered = false, is_rejected = false;
      const bool is_ddmc_step = B.dx_push * (c.ss + c.aa) > P.tau_ddmc;  // transport_ddmc.cpp:135
      if (is_ddmc_step) {
        c_ddmc++;
        const double *fp0 = P.f.fp[0] + (size_t)B.lb * P.f.nface[0];
        double Pf[6];
        Pf[0] = fp0[((size_t)p.kp * nx2 + p.jp) * (nx1 + 1) + p.ip];
        Pf[3] = fp0[((size_t)p.kp * nx2 + p.jp) * (nx1 + 1) + p.ip + 1];
        if (multi_d) {
          const double *fp1 = P.f.fp[1] + (size_t)B.lb * P.f.nface[1];
          Pf[1] = fp1[((size_t)p.kp * (nx2 + 1) + p.jp) * nx1 + p.ip];
          Pf[4] = fp1[((size_t)p.kp * (nx2 + 1) + p.jp + 1) * nx1 + p.ip];
        } else {
          Pf[1] = 0.0; Pf[4] = 0.0;
        }
        if (three_d) {
          const double *fp2 = P.f.fp[2] + (size_t)B.lb * P.f.nface[2];
          Pf[2] = fp2[((size_t)p.kp * nx2 + p.jp) * nx1 + p.ip];
          Pf[5] = fp2[((size_t)(p.kp + 1) * nx2 + p.jp) * nx1 + p.ip];
        } else {
          Pf[2] = 0.0; Pf[5] = 0.0;
        }
        bool resampled = false;
        ddmc_albedo2<NDIM>(rng, K, c, vv, p, is_rejected);
        if (!is_rejected) ddmc_step2<NDIM>(rng, K, t_end, c, vv, Pf, p, is_absorbed, resampled);
        // Xtoijk (transport_ddmc.cpp:198) only where the position was drawn or pushed back out of the cell
        if (is_rejected || resampled) xtoijk<NDIM>(B, p, p.ip, p.jp, p.kp);
      } else {
        double xi_abs, xi_sc;
        rng.draw2(K, xi_abs, xi_sc);
        int di, dj, dk;
        unsigned bad_div = 0;
        imc_step_lockstep<NDIM>(xi_abs, xi_sc, t_end, c, vv, B.dx_push, p, is_absorbed, is_scattered, di, dj, dk, bad_div);
        p.ip += di; p.jp += dj; p.kp += dk;  // the face nudges move the index exactly (see k_transport_imc)
      }
      int kind = (is_ddmc_step ? JB_EV_DDMC : 0) | (is_rejected ? JB_EV_ALBEDO_REJECT : 0);
      bool gone = false;
      if (!on_block(m, p.ip, p.jp, p.kp)) {
        // zero velocity tags a DDMC photon for block-face resampling (transport_ddmc.cpp:203-211)
        const double vmask = !(is_ddmc_step && multi_d && !is_rejected) ? 1.0 : 0.0;
        p.vx *= vmask;
        p.vy *= vmask;
        p.vz *= vmask;
        if (LOG) {
          const uint4 idn = make_uint4(rng.id0, rng.id1, rng.id2, 0u);
          log_event(P, idn, draw0, kind | JB_EV_LEFT_BLOCK | (!(p.t < t_end) ? JB_EV_CENSUS : 0), B.gid, p);
        }
        // ---- MeshSend / MeshReceive, on device ----
        if (!apply_swarm_bc<NDIM>(m, p)) {
          atomicAdd(&sh_cnt[3], 1u);
          e_esc += P.s.w[idx];
          gone = true;
        } else {
          const double pos[3] = {p.x, p.y, p.z};
          const int ngid = find_block(m, pos);
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
                r.x = p.x; r.y = p.y; r.z = p.z; r.t = p.t; r.w = P.s.w[idx]; r.e = P.s.e[idx];
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
              BlkReg NB;
              load_block(m, gid, NB);
              xtoijk<NDIM>(NB, p, p.ip, p.jp, p.kp);
              Rng g;  // the block-face sampler (rare) runs on the per-draw generator, same stream
              g.init(rng.id0, rng.id1, rng.id2, rng.n, P.key0, P.key1_particle);
              bface_sample<NDIM>(P, NB, p, g);
              if (g.n != rng.n) rng.load(rng.id0, rng.id1, rng.id2, g.n, K);
            }
          }
        }
      } else if (is_absorbed) {
        const double ww = P.s.w[idx];
        atomicAdd(&P.f.edelta[(long long)B.lb * m.ncell + ((long long)p.kp * nx2 + p.jp) * nx1 + p.ip], ww);
        atomicAdd(&sh_cnt[1], 1u);
        {
          const double y = ww - e_abs_c, t = e_abs + y;
          e_abs_c = (t - e_abs) - y;
          e_abs = t;
        }
        if (LOG) log_event(P, make_uint4(rng.id0, rng.id1, rng.id2, 0u), draw0, kind | JB_EV_ABSORB, B.gid, p);
        gone = true;
      } else {
        if (is_scattered) {
          double xi_mu, xi_phi;
          rng.draw2(K, xi_mu, xi_phi);
          scatter_iso2(xi_mu, xi_phi, vv, p.vx, p.vy, p.vz);
          c_scat++;
          kind |= JB_EV_SCATTER;
        }
        if (LOG)
          log_event(P, make_uint4(rng.id0, rng.id1, rng.id2, 0u), draw0, kind | (!(p.t < t_end) ? JB_EV_CENSUS : 0),
                    B.gid, p);
      }
      // ---- finished for this call? write back and free the lane -------------------------------
      if (gone) {
        P.s.loc[idx].x = -1;
        atomicAdd(&sh_cnt[2], 1u);
        have = false;
      } else if (!(p.t < t_end)) {
        P.s.x[idx] = p.x; P.s.y[idx] = p.y; P.s.z[idx] = p.z; P.s.t[idx] = p.t;
        P.s.vx[idx] = p.vx; P.s.vy[idx] = p.vy; P.s.vz[idx] = p.vz;
        P.s.loc[idx] = make_int2(m.blocks[gid].local, (p.kp * nx2 + p.jp) * nx1 + p.ip);
        P.s.idb[idx] = make_uint2(rng.id2, rng.n);
        have = false;
      }
    }
  }
  DStats *S = P.stats;
  warp_add_u64(&S->n_segments, c_seg);
  warp_add_u64(&S->n_scatters, c_scat);
  warp_add_u64(&S->n_ddmc, c_ddmc);
  __syncthreads();
  if (threadIdx.x == 0) {
    if (sh_cnt[0]) atomicAdd(&S->n_crossings, (unsigned long long)sh_cnt[0]);
    if (sh_cnt[1]) atomicAdd(&S->n_absorbed, (unsigned long long)sh_cnt[1]);
    if (sh_cnt[2]) atomicAdd(&S->n_dead, (unsigned long long)sh_cnt[2]);
    if (sh_cnt[3]) atomicAdd(&S->n_escaped, (unsigned long long)sh_cnt[3]);
    if (sh_cnt[4]) atomicAdd(&S->n_sent, (unsigned long long)sh_cnt[4]);
    if (sh_cnt[5]) atomicAdd(&S->n_error, (unsigned long long)sh_cnt[5]);
    if (sh_cnt[6]) atomicAdd(&S->n_overflow, (unsigned long long)sh_cnt[6]);
  }
  warp_add_energy(P, EB_ABSORBED, e_abs);
  if (__any_sync(FULL, e_esc != 0.0)) warp_add_energy(P, EB_ESCAPED, e_esc);
}

template <int NDIM, int MINB>
static void launch_transport_hybrid_b(const StepParams &P, long long first, long long n, double t_start, double dt,
                                      cudaStream_t st) {
  static int n_sm = 0;
  if (n_sm == 0) {
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&n_sm, cudaDevAttrMultiProcessorCount, dev);
    if (n_sm <= 0) n_sm = 148;
  }
  long long blocks = (long long)n_sm * MINB;  // persistent grid: MINB CTAs of 4 warps on each SM
  const long long need = (n + IMC_TPB - 1) / IMC_TPB;
  if (blocks > need) blocks = need;
  if (blocks < 1) blocks = 1;
  const int rf = g_tuning.refill_min;
  if (P.log.cap > 0)
    k_transport_hybrid<NDIM, true, 4><<<(unsigned)(blocks > n_sm * 4LL ? n_sm * 4LL : blocks), IMC_TPB, 0, st>>>(P, first, n, t_start, dt, rf);
  else
    k_transport_hybrid<NDIM, false, MINB><<<(unsigned)blocks, IMC_TPB, 0, st>>>(P, first, n, t_start, dt, rf);
}
template <int NDIM>
static void launch_transport_hybrid(const StepParams &P, long long first, long long n, double t_start, double dt,
                                    cudaStream_t st) {
  static int ctas = 0;
  if (ctas == 0) {
    const char *v = getenv("JB_DDMC_CTAS_PER_SM");
    ctas = v ? atoi(v) : 5;
  }
  switch (ctas) {
    case 5: launch_transport_hybrid_b<NDIM, 5>(P, first, n, t_start, dt, st); break;
    case 6: launch_transport_hybrid_b<NDIM, 6>(P, first, n, t_start, dt, st); break;
    default: launch_transport_hybrid_b<NDIM, 4>(P, first, n, t_start, dt, st); break;
  }
}

#endif  // JAYBENNE_B200_JB_TRANSPORT_DDMC_CUH_
