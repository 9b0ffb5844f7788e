// jb_transport_imc.cuh — the IMC history kernel as persistent warps with lane refill.
// Included by jb_kernels.cu after the shared helpers (BlkReg, xtoijk, apply_swarm_bc, ...).
//
// Reference behaviour: TransportPhotons, src/jaybenne/transport.cpp:28-181, with
// ptcl_transport_step / scatter (transport_utils.hpp:111-160, scattering.hpp:21-29), the swarm
// boundary conditions (boundaries.hpp:24-84) and the block hand-over (jaybenne.cpp:26-61).
//
// Why it looks the way it does (profiles/r01_v1_*.md, r01_v2_*.md):
//  * one thread per photon ran 14.6 / 32 lanes wide (history lengths differ) -> a warp owns a
//    contiguous range of swarm slots and refills idle lanes from it (ballot + prefix rank);
//  * v2 still ran 18.8 / 32 wide because of divergence INSIDE a segment, removed here without
//    changing a single rounding:
//      - Philox in lock step: every segment consumes exactly two draws, so every lane evaluates
//        exactly one Philox block per segment (block (n+1)>>1); a lane whose draw counter is odd
//        carries the unused half of its previous block (Rng2);
//      - the vx>0 / vx<0 face-distance branches are one select + one divide per direction;
//      - a photon that leaves its meshblock through a face shared with a same-level block takes
//        the neighbour from a 27-entry table and shifts its cell index by the block size (the
//        general find_block + Xtoijk path remains for level changes and non-periodic faces);
//      - a photon is loaded with the cell index stored in the swarm instead of three divides;
//  * the kernel is latency bound (time ~ 1 / resident warps), so per-lane state is kept small:
//    the meshblock geometry is re-read per segment from a 64-byte BlkGeo line (L1-resident, a
//    warp's photons share a block) instead of living in 28 registers, weight and energy are
//    re-read on the rare paths that need them.
// VALIDATE re-derives every cell index / block with Xtoijk / find_block and counts
// disagreements into n_error (tests/test_parity_gpu.py runs it).
#ifndef JAYBENNE_B200_JB_TRANSPORT_IMC_CUH_
#define JAYBENNE_B200_JB_TRANSPORT_IMC_CUH_

static constexpr int IMC_TPB = 128;
static constexpr int IMC_CHUNK = 1024;  // slots per claim: 8 KB of each swarm array

__device__ __forceinline__ CellProps load_props(const CellProps *p) {
  // 2 x 16-byte read-only loads of one 32-byte aligned record
  const double2 a = __ldg(reinterpret_cast<const double2 *>(p));
  const double2 b = __ldg(reinterpret_cast<const double2 *>(p) + 1);
  CellProps c;
  c.rho = a.x; c.temp = a.y; c.ff = b.x; c.ss = b.y;
  return c;
}

// The photon's Philox stream (same streams as Rng in jb_device.cuh) consumed two draws at a
// time. Invariant: when n is odd, (clo, chi) are words 2, 3 of block n >> 1, i.e. draw #n.
struct Rng2 {
  uint32_t id0, id1, id2, n, clo, chi;

  __device__ __forceinline__ void block(uint32_t blk, const PKeys &K, uint32_t &w0, uint32_t &w1,
                                        uint32_t &w2, uint32_t &w3) const {
    uint32_t c0 = blk, c1 = id0, c2 = id1, c3 = id2;
#pragma unroll
    for (int r = 0; r < 10; ++r) {
      const uint32_t hi0 = __umulhi(0xD2511F53u, c0), lo0 = 0xD2511F53u * c0;
      const uint32_t hi1 = __umulhi(0xCD9E8D57u, c2), lo1 = 0xCD9E8D57u * c2;
      c0 = hi1 ^ c1 ^ K.k[2 * r];  // round keys straight from the kernel-parameter bank (PKeys)
      c1 = lo1;
      c2 = hi0 ^ c3 ^ K.k[2 * r + 1];
      c3 = lo0;
    }
    w0 = c0; w1 = c1; w2 = c2; w3 = c3;
  }
  __device__ __forceinline__ void load(uint32_t a, uint32_t b, uint32_t c, uint32_t draws, const PKeys &K) {
    id0 = a; id1 = b; id2 = c; n = draws; clo = chi = 0u;
    if (draws & 1u) {
      uint32_t w0, w1;
      block(draws >> 1, K, w0, w1, clo, chi);
    }
  }
  static __device__ __forceinline__ double to_double(uint32_t lo, uint32_t hi) {
    const uint64_t u = ((uint64_t)hi << 32) | lo;
    return (double)(u >> 11) * 1.1102230246251565e-16;
  }
  // draws #n and #n+1
  __device__ __forceinline__ void draw2(const PKeys &K, double &a, double &b) {
    uint32_t w0, w1, w2, w3;
    block((n + 1u) >> 1, K, w0, w1, w2, w3);
    const bool odd = (n & 1u) != 0;
    a = to_double(odd ? clo : w0, odd ? chi : w1);
    b = to_double(odd ? w0 : w2, odd ? w1 : w3);
    clo = w2;  // only meaningful while n is odd
    chi = w3;
    n += 2u;
  }
};

// ptcl_transport_step (transport_utils.hpp:111-160) with the two draws handed in and the
// face-distance branches written as selects. Operation for operation the arithmetic of
// imc_step_idx (jb_device.cuh): a direction with v == 0 computes an unused quotient.
// div_normal with the VALIDATE cross-check: the IEEE `/` is evaluated beside it and a differing bit pattern
// counted (`used` is false where the reference never looks at the quotient, e.g. a direction with v == 0)
template <bool CHECK>
__device__ __forceinline__ double divn(const double a, const double b, unsigned &bad, const bool used = true) {
  const double q = jbm::div_normal(a, b);
  if (CHECK) {
    const double r = a / b;
    if (used && __double_as_longlong(q) != __double_as_longlong(r)) bad++;
  }
  return q;
}

template <int NDIM, bool CHECK = false>
__device__ __forceinline__ void imc_step_lockstep(const double xi_abs, const double xi_sc, const double t_end,
                                                  const CellEnv &c, const double vv, const double dx_cell, Photon &p,
                                                  bool &is_absorbed, bool &is_scattered, int &di, int &dj, int &dk,
                                                  unsigned &bad) {
  constexpr bool multi_d = NDIM >= 2, three_d = NDIM == 3;
  const double rmin = DBL_MIN;
  // denominators are >= rmin = DBL_MIN (a normal number) and far below overflow: no range test needed
  const double lam_abs = divn<CHECK>(1.0, c.ff * c.aa + rmin, bad);
  const double lam_sc = divn<CHECK>(1.0, c.ss + (1.0 - c.ff) * c.aa + rmin, bad);
  const double dx_abs = -lam_abs * jbm::log_unit(xi_abs);
  const double dx_sc = -lam_sc * jbm::log_unit(xi_sc);
  const double dx_end = vv * (t_end - p.t);
  double dx_push = jbm::min2(dx_cell, dx_end);
  {
    const double face = p.vx > 0.0 ? c.xu : c.xl;
    const double d = divn<CHECK>(vv * (face - p.x), p.vx, bad, p.vx != 0.0);  // |v| is 0 (result unused) or >= ~1e-7
    const double dm = jbm::min2(dx_push, d);
    dx_push = (p.vx != 0.0) ? dm : dx_push;
  }
  if (multi_d) {
    const double face = p.vy > 0.0 ? c.yu : c.yl;
    const double d = divn<CHECK>(vv * (face - p.y), p.vy, bad, p.vy != 0.0);
    const double dm = jbm::min2(dx_push, d);
    dx_push = (p.vy != 0.0) ? dm : dx_push;
  }
  if (three_d) {
    const double face = p.vz > 0.0 ? c.zu : c.zl;
    const double d = divn<CHECK>(vv * (face - p.z), p.vz, bad, p.vz != 0.0);
    const double dm = jbm::min2(dx_push, d);
    dx_push = (p.vz != 0.0) ? dm : dx_push;
  }
  is_absorbed = ((dx_abs < dx_push) && (dx_abs < dx_sc));
  is_scattered = (!is_absorbed && (dx_sc < dx_push));
  const double dt_push = divn<CHECK>(is_absorbed ? dx_abs : (is_scattered ? dx_sc : dx_push), vv, bad);  // <= dx_cell
  p.t += dt_push;
  p.x += p.vx * dt_push;
  p.y += (multi_d ? 1.0 : 0.0) * p.vy * dt_push;
  p.z += (three_d ? 1.0 : 0.0) * p.vz * dt_push;
  const double fdx = EPS_IMC * (c.xu - c.xl);
  const double fdy = EPS_IMC * (c.yu - c.yl);
  const double fdz = EPS_IMC * (c.zu - c.zl);
  const bool nxl = fabs(p.x - c.xl) < fdx;
  p.x = nxl ? c.xl - fdx : p.x;
  const bool nxu = fabs(p.x - c.xu) < fdx;
  p.x = nxu ? c.xu + fdx : p.x;
  const bool nyl = multi_d && fabs(p.y - c.yl) < fdy;
  p.y = nyl ? c.yl - fdy : p.y;
  const bool nyu = multi_d && fabs(p.y - c.yu) < fdy;
  p.y = nyu ? c.yu + fdy : p.y;
  const bool nzl = three_d && fabs(p.z - c.zl) < fdz;
  p.z = nzl ? c.zl - fdz : p.z;
  const bool nzu = three_d && fabs(p.z - c.zu) < fdz;
  p.z = nzu ? c.zu + fdz : p.z;
  di = nxu ? 1 : (nxl ? -1 : 0);
  dj = nyu ? 1 : (nyl ? -1 : 0);
  dk = nzu ? 1 : (nzl ? -1 : 0);
}

// isotropic scatter (scattering.hpp:21-29) from two draws
__device__ __forceinline__ void scatter_iso2(const double xi_mu, const double xi_phi, const double vv, double &vx,
                                             double &vy, double &vz) {
  const double mu = 2.0 * xi_mu - 1.0;
  const double phi = 2.0 * JB_PI * xi_phi;
  const double stheta = sqrt(1.0 - mu * mu);
  double sn, cs;
  jbm::sincos(phi, &sn, &cs);
  vx = vv * stheta * cs;
  vy = vv * stheta * sn;
  vz = vv * mu;
}

// A load the compiler may not sink into the branch that uses it (asm volatile keeps program order among
// themselves): the refill issues all of a slot's loads back to back and waits once.
__device__ __forceinline__ double ld_once(const double *p) {
  double v;
  asm volatile("ld.global.f64 %0, [%1];" : "=d"(v) : "l"(p));
  return v;
}
__device__ __forceinline__ int2 ld_once(const int2 *p) {
  int2 v;
  asm volatile("ld.global.v2.s32 {%0, %1}, [%2];" : "=r"(v.x), "=r"(v.y) : "l"(p));
  return v;
}
__device__ __forceinline__ uint2 ld_once(const uint2 *p) {
  uint2 v;
  asm volatile("ld.global.v2.u32 {%0, %1}, [%2];" : "=r"(v.x), "=r"(v.y) : "l"(p));
  return v;
}

// ABS selects the absorption closure at compile time: 0 = gray through the reference's
// j_nu / B_nu round trip, 1 = gray as rho * kappa (JB_GRAY_DIRECT), 2 = whatever the
// Closure says (free-free); 0 and 1 need neither the photon frequency nor the EPBremss code.
enum { ABS_GRAY_KIRCHHOFF = 0, ABS_GRAY_DIRECT = 1, ABS_GENERIC = 2 };

template <int NDIM, int MINB, bool VALIDATE, int ABS>
__global__ void __launch_bounds__(IMC_TPB, MINB)
    k_transport_imc(StepParams P, long long first, long long n, double t_start, double dt, int refill_min) {
  const DMesh &m = P.mesh;
  if (P.range) {
    first = P.range[0];
    n = P.range[1];
  }
  const double t_end = t_start + dt;
  const double vv = P.cl.c_code;
  const int nx1 = m.nxb[0], nx2 = m.nxb[1], nx3 = m.nxb[2];
  const bool pow2 = (nx1 & (nx1 - 1)) == 0 && (nx2 & (nx2 - 1)) == 0;
  const int sh1 = pow2 ? __ffs(nx1) - 1 : -1, sh2 = pow2 ? __ffs(nx2) - 1 : 0;
  const unsigned FULL = 0xffffffffu;
  const unsigned lane = threadIdx.x & 31u;
  const unsigned lt_mask = (1u << lane) - 1u;
  const long long gw = (blockIdx.x * (long long)blockDim.x + threadIdx.x) >> 5;
  // a warp works through chunks of IMC_CHUNK contiguous slots claimed from one device counter,
  // so no warp is left with a long static range while the rest of the GPU has drained
  // the chunk cursor is warp-uniform and only touched by the refill: it lives in shared memory, and the
  // slot is kept as a 32-bit offset from `first` (a rank's pool holds < 2^32 photons), which takes five
  // registers off the per-lane state of a kernel that spills at 64
  __shared__ unsigned sh_cur[IMC_TPB / 32], sh_end[IMC_TPB / 32];
  bool exhausted = false;

  bool have = false;     // this lane holds a photon in flight
  bool pend_scat = false, need_carry = false;  // see the shared extra Philox block below
  unsigned idx32 = 0;    // its slot, relative to `first`
  Photon p;
  PhotonFreq q;
  Rng2 rng;
  int gid = 0, lb = 0;   // its meshblock
  p.x = p.y = p.z = p.t = p.vx = p.vy = p.vz = 0.0;
  p.ip = p.jp = p.kp = 0;
  q.f = q.hnu = q.bpref = 0.0;
  rng.id0 = rng.id1 = rng.id2 = rng.n = rng.clo = rng.chi = 0u;

  // per-lane counters only for what happens every segment; the rare events (absorption, block
  // change, removal) count into shared memory so they do not hold registers across the loop
  unsigned c_seg = 0, c_scat = 0;
  __shared__ unsigned sh_cnt[3];                // crossings, absorbed, dead
  double e_abs = 0.0;                           // absorbed energy of this lane's photons (a shared double add is a CAS loop)
  if (threadIdx.x < 3) sh_cnt[threadIdx.x] = 0u;
  if (threadIdx.x < IMC_TPB / 32) sh_cur[threadIdx.x] = sh_end[threadIdx.x] = 0u;
  __syncthreads();
  const unsigned wic = threadIdx.x >> 5;

  for (;;) {
    // ---- refill idle lanes from the warp's range ----------------------------------------
    for (;;) {
      unsigned cur = sh_cur[wic], end = sh_end[wic];
      if (cur >= end) {
        if (exhausted) break;
        unsigned long long c0 = 0;
        if (lane == 0) c0 = atomicAdd(P.work, (unsigned long long)IMC_CHUNK);
        c0 = __shfl_sync(FULL, c0, 0);
        if ((long long)c0 >= n) {
          exhausted = true;
          break;
        }
        cur = (unsigned)c0;
        end = (long long)c0 + IMC_CHUNK > n ? (unsigned)n : cur + IMC_CHUNK;
      }
      const unsigned need = __ballot_sync(FULL, !have);
      const int cnt = __popc(need);
      if ((cnt < refill_min && cnt < 32) || cnt == 0) {
        __syncwarp();
        if (lane == 0) { sh_cur[wic] = cur; sh_end[wic] = end; }
        __syncwarp();
        break;
      }
      const unsigned my32 = cur + __popc(need & lt_mask);
      const long long my = first + (long long)my32;
      __syncwarp();
      if (lane == 0) { sh_cur[wic] = cur + cnt; sh_end[wic] = end; }
      __syncwarp();
      const bool in_chunk = my32 < end;
      if (!have && in_chunk) {
        // every array of the slot is requested before the first test: one DRAM round trip per refill instead
        // of the three a loc -> time -> state chain of dependent branches costs (a dead or already-censused
        // slot wastes 84 B of a bandwidth the kernel uses 7 % of)
        const int2 lc = ld_once(&P.s.loc[my]);
        const double t0 = ld_once(&P.s.t[my]);
        const double l_x = ld_once(&P.s.x[my]), l_y = ld_once(&P.s.y[my]), l_z = ld_once(&P.s.z[my]);
        const double l_vx = ld_once(&P.s.vx[my]), l_vy = ld_once(&P.s.vy[my]), l_vz = ld_once(&P.s.vz[my]);
        const double l_e = ld_once(&P.s.e[my]);
        const uint2 ia = ld_once(&P.s.ida[my]), ib = ld_once(&P.s.idb[my]);
        {
          // census photons of an earlier iteration are skipped (transport.cpp:98)
          if (lc.x >= 0 && t0 < t_end) {
            idx32 = my32;
            p.t = t0;
            p.x = l_x; p.y = l_y; p.z = l_z;
            p.vx = l_vx; p.vy = l_vy; p.vz = l_vz;
            q = photon_freq(P.cl, l_e);
            rng.id0 = ia.x; rng.id1 = ia.y; rng.id2 = ib.x; rng.n = ib.y;
            need_carry = (ib.y & 1u) != 0;  // the half block an odd draw counter carries: evaluated below
            lb = lc.x;
            gid = m.local_gid[lb];
            // the stored flat cell is what Xtoijk (transport.cpp:96) returns for a photon at rest
            if (sh1 >= 0) {  // block sizes that are powers of two (the usual case): shifts
              p.ip = lc.y & (nx1 - 1);
              p.jp = (lc.y >> sh1) & (nx2 - 1);
              p.kp = lc.y >> (sh1 + sh2);
            } else {
              p.ip = lc.y % nx1;
              const int rest = lc.y / nx1;
              p.jp = rest % nx2;
              p.kp = rest / nx2;
            }
            if (VALIDATE) {
              BlkReg B;
              load_block(m, gid, B);
              int ti, tj, tk;
              xtoijk<NDIM>(B, p, ti, tj, tk);
              if (ti != p.ip || tj != p.jp || tk != p.kp) atomicAdd(&P.stats->n_validate, 1ull);
            }
            have = true;
          }
        }
      }
    }
    if (__ballot_sync(FULL, have) == 0u) break;  // range exhausted and nothing in flight

    // ---- the one extra Philox block of the iteration --------------------------------------
    // Two rare needs for a block outside the lock-step one: a photon that scattered in its last segment
    // (draws #n, #n+1 for the new direction) and a photon just loaded with an odd draw counter (the
    // second half of block n >> 1, which it carries). Evaluated separately they were two divergent
    // Philox passes per iteration at ~6 and ~2 lanes; here they share one. The scatter itself is the
    // reference's, applied before the photon's next segment instead of right after the last one.
    if (have && (pend_scat || need_carry)) {
      uint32_t w0, w1, w2, w3;
      rng.block(pend_scat ? (rng.n + 1u) >> 1 : rng.n >> 1, P.rk, w0, w1, w2, w3);
      if (pend_scat) {
        const bool odd = (rng.n & 1u) != 0;
        const double xi_mu = Rng2::to_double(odd ? rng.clo : w0, odd ? rng.chi : w1);
        const double xi_phi = Rng2::to_double(odd ? w0 : w2, odd ? w1 : w3);
        rng.n += 2u;
        scatter_iso2(xi_mu, xi_phi, vv, p.vx, p.vy, p.vz);
        c_scat++;
        pend_scat = false;
      }
      rng.clo = w2;
      rng.chi = w3;
      need_carry = false;
    }

    if (have) {
      // ---- one segment of the while (t < t_end) loop, transport.cpp:112-171 ------------
      c_seg++;
      const long long idx = first + (long long)idx32;
      // loads first: their latency hides behind the Philox block
      const long long ci = (long long)lb * m.ncell + ((long long)p.kp * nx2 + p.jp) * nx1 + p.ip;
      const CellProps cp = load_props(P.f.props + ci);
      const BlkGeo *geo = m.geo + gid;
      const double2 gx = __ldg(&geo->xd[0]), gy = __ldg(&geo->xd[1]), gz = __ldg(&geo->xd[2]);
      const double dx_cell = __ldg(&geo->dx_push);
      double xi_abs, xi_sc;
      rng.draw2(P.rk, xi_abs, xi_sc);
      CellEnv c;
      {
        // coords.Xc<d>(i) -/+ dx/2 (transport.cpp:114-119), Xc as UniformCartesian computes it
        const double xcen = gx.x + ((p.ip + m.ng[0]) + 0.5) * gx.y;
        const double ycen = gy.x + ((p.jp + m.ng[1]) + 0.5) * gy.y;
        const double zcen = gz.x + ((p.kp + m.ng[2]) + 0.5) * gz.y;
        c.xl = xcen - 0.5 * gx.y; c.xu = xcen + 0.5 * gx.y;
        c.yl = ycen - 0.5 * gy.y; c.yu = ycen + 0.5 * gy.y;
        c.zl = zcen - 0.5 * gz.y; c.zu = zcen + 0.5 * gz.y;
      }
      c.ff = cp.ff;
      c.ss = cp.ss;
      if (ABS == ABS_GRAY_DIRECT) {
        c.aa = ((P.cl.rho_unit * cp.rho) * P.cl.kappa) * P.cl.length_unit;
      } else if (ABS == ABS_GRAY_KIRCHHOFF) {
        // absorption_coeff_hoisted() with the gray branch fixed
        const double r = P.cl.rho_unit * cp.rho, T = P.cl.temp_unit * cp.temp;
        const double x = q.hnu / (P.cl.kb * T);
        const double Bnu = q.bpref / jbm::expm1(x);
        const double jnu = r * P.cl.kappa * Bnu;
        c.aa = robust_ratio_pos(jnu, Bnu) * P.cl.length_unit;
      } else {
        c.aa = absorption_coeff_hoisted(P.cl, cp.rho, cp.temp, q);
      }
      bool is_absorbed = false, is_scattered = false;
      int di, dj, dk;
      unsigned bad_div = 0;
      imc_step_lockstep<NDIM, VALIDATE>(xi_abs, xi_sc, t_end, c, vv, dx_cell, p, is_absorbed, is_scattered, di, dj, dk,
                                        bad_div);
      if (VALIDATE && bad_div) atomicAdd(&P.stats->n_validate, (unsigned long long)bad_div);
      if (VALIDATE) {
        BlkReg B;
        load_block(m, gid, B);
        int ti, tj, tk;
        xtoijk<NDIM>(B, p, ti, tj, tk);
        if (ti != p.ip + di || tj != p.jp + dj || tk != p.kp + dk) atomicAdd(&P.stats->n_validate, 1ull);
      }
      p.ip += di; p.jp += dj; p.kp += dk;
      bool gone = false;
      if (!on_block(m, p.ip, p.jp, p.kp)) {
        // ---- left the meshblock: swarm BCs + hand-over (MeshSend / MeshReceive on device) ----
        const int bx = p.ip < 0 ? -1 : (p.ip >= nx1 ? 1 : 0);
        const int by = p.jp < 0 ? -1 : (p.jp >= nx2 ? 1 : 0);
        const int bz = p.kp < 0 ? -1 : (p.kp >= nx3 ? 1 : 0);
        if (!apply_swarm_bc<NDIM>(m, p)) {
          atomicAdd(&P.stats->n_escaped, 1ull);
          atomicAdd(&P.ebins[EB_ESCAPED * EBINS + (unsigned)(gw & (EBINS - 1))], P.s.w[idx]);
          gone = true;
        } else {
          int ngid = __ldg(&m.nbr[lb * 27 + (bz + 1) * 9 + (by + 1) * 3 + (bx + 1)]);
          const bool fast = ngid >= 0;
          if (fast) {
            // same-level neighbour: the photon sits eps_imc_offset cell widths inside it
            p.ip -= bx * nx1; p.jp -= by * nx2; p.kp -= bz * nx3;
          } else {
            const double pos[3] = {p.x, p.y, p.z};
            ngid = find_block(m, pos);
          }
          if (VALIDATE && fast) {
            const double pos[3] = {p.x, p.y, p.z};
            const int vg = find_block(m, pos);
            bool bad = vg != ngid;
            if (!bad) {
              BlkReg B;
              load_block(m, vg, B);
              int ti, tj, tk;
              xtoijk<NDIM>(B, p, ti, tj, tk);
              bad = ti != p.ip || tj != p.jp || tk != p.kp;
            }
            if (bad) atomicAdd(&P.stats->n_validate, 1ull);
          }
          if (ngid < 0) {
            atomicAdd(&P.stats->n_error, 1ull);
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
                atomicAdd(&P.stats->n_sent, 1ull);
              } else {
                atomicAdd(&P.stats->n_overflow, 1ull);
              }
              gone = true;
            } else {
              gid = ngid;
              lb = nb.local;
              if (!fast) {
                BlkReg B;
                load_block(m, gid, B);
                xtoijk<NDIM>(B, p, p.ip, p.jp, p.kp);
              }
            }
          }
        }
      } else if (is_absorbed) {
        const double ww = P.s.w[idx];
        atomicAdd(&P.f.edelta[(long long)lb * m.ncell + ((long long)p.kp * nx2 + p.jp) * nx1 + p.ip], ww);
        atomicAdd(&sh_cnt[1], 1u);
        e_abs += ww;
        gone = true;
      } else if (is_scattered) {
        if (p.t < t_end) {
          pend_scat = true;  // new direction from the shared extra block at the top of the next iteration
        } else {             // scattered exactly into census (rounding): finish it here, the lane is freed below
          double xi_mu, xi_phi;
          rng.draw2(P.rk, xi_mu, xi_phi);
          scatter_iso2(xi_mu, xi_phi, vv, p.vx, p.vy, p.vz);
          c_scat++;
        }
      }
      // ---- photon finished for this call? write it back and free the lane -----------------
      if (gone) {
        P.s.loc[idx].x = -1;
        atomicAdd(&sh_cnt[2], 1u);
        have = false;
      } else if (!(p.t < t_end)) {
        P.s.x[idx] = p.x; P.s.y[idx] = p.y; P.s.z[idx] = p.z; P.s.t[idx] = p.t;
        P.s.vx[idx] = p.vx; P.s.vy[idx] = p.vy; P.s.vz[idx] = p.vz;
        P.s.loc[idx] = make_int2(lb, (p.kp * nx2 + p.jp) * nx1 + p.ip);
        P.s.idb[idx] = make_uint2(rng.id2, rng.n);
        have = false;
      }
    }
  }
  DStats *S = P.stats;
  warp_add_u64(&S->n_segments, c_seg);
  warp_add_u64(&S->n_scatters, c_scat);
  __syncthreads();
  if (threadIdx.x == 0) {
    if (sh_cnt[0]) atomicAdd(&S->n_crossings, (unsigned long long)sh_cnt[0]);
    if (sh_cnt[1]) atomicAdd(&S->n_absorbed, (unsigned long long)sh_cnt[1]);
    if (sh_cnt[2]) atomicAdd(&S->n_dead, (unsigned long long)sh_cnt[2]);
  }
  warp_add_energy(P, EB_ABSORBED, e_abs);
}

static TransportTuning g_tuning = {2, 8, 4, 0};
void set_transport_tuning(const TransportTuning &t) { g_tuning = t; }
// the persistent kernels (variants 1, 2 and the hybrid kernel) can take their slot range from device memory
bool transport_takes_device_range(const StepParams &P) { return g_tuning.variant >= 1 && P.log.cap <= 0; }

template <int NDIM, int MINB, bool VALIDATE>
static void launch_transport_imc_abs(const StepParams &P, long long first, long long n, double t_start, double dt,
                                     cudaStream_t st) {
  // persistent grid: MINB CTAs on each of the SMs (148 on B200); fewer if the range is short
  static int n_sm = 0;
  if (n_sm == 0) {
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&n_sm, cudaDevAttrMultiProcessorCount, dev);
    if (n_sm <= 0) n_sm = 148;
  }
  long long blocks = (long long)n_sm * MINB;
  const long long need = (n + IMC_TPB - 1) / IMC_TPB;
  if (blocks > need) blocks = need;
  if (blocks < 1) blocks = 1;
  const int rf = g_tuning.refill_min;
  if (P.cl.opacity_model != JB_OPAC_GRAY)
    k_transport_imc<NDIM, MINB, VALIDATE, ABS_GENERIC><<<(unsigned)blocks, IMC_TPB, 0, st>>>(P, first, n, t_start, dt, rf);
  else if (P.cl.gray_mode == JB_GRAY_DIRECT)
    k_transport_imc<NDIM, MINB, VALIDATE, ABS_GRAY_DIRECT><<<(unsigned)blocks, IMC_TPB, 0, st>>>(P, first, n, t_start, dt, rf);
  else
    k_transport_imc<NDIM, MINB, VALIDATE, ABS_GRAY_KIRCHHOFF><<<(unsigned)blocks, IMC_TPB, 0, st>>>(P, first, n, t_start, dt, rf);
}
template <int NDIM>
static void launch_transport_imc(const StepParams &P, long long first, long long n, double t_start, double dt,
                                 cudaStream_t st) {
  if (g_tuning.validate) {  // the cross-checked build exists for one occupancy only
    launch_transport_imc_abs<NDIM, 4, true>(P, first, n, t_start, dt, st);
    return;
  }
  switch (g_tuning.ctas_per_sm) {
    case 4: launch_transport_imc_abs<NDIM, 4, false>(P, first, n, t_start, dt, st); break;
    case 6: launch_transport_imc_abs<NDIM, 6, false>(P, first, n, t_start, dt, st); break;
    case 7: launch_transport_imc_abs<NDIM, 7, false>(P, first, n, t_start, dt, st); break;
    default: launch_transport_imc_abs<NDIM, 8, false>(P, first, n, t_start, dt, st); break;
  }
}

#endif  // JAYBENNE_B200_JB_TRANSPORT_IMC_CUH_
