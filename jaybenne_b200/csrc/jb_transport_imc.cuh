// jb_transport_imc.cuh — the IMC history kernel as persistent warps with lane refill.
// Included by jb_kernels.cu after the shared helpers (BlkReg, xtoijk, apply_swarm_bc, ...).
//
// Reference behaviour: TransportPhotons, src/jaybenne/transport.cpp:28-181, with
// ptcl_transport_step / scatter (transport_utils.hpp:111-160, scattering.hpp:21-29), the swarm
// boundary conditions (boundaries.hpp:24-84) and the block hand-over (jaybenne.cpp:26-61).
//
// Why it looks the way it does (profiles/r01_v1_ncu_transport_cells64.md): with one thread
// per photon, 32 histories of very different length share a warp and the warp runs at
// 14.6 / 32 active lanes. Here a warp owns a contiguous range of swarm slots and keeps every
// lane busy: a lane whose photon reached census, was absorbed or left the rank stores its
// result and takes the next slot of the range (ballot + prefix rank), so the segment body
// always runs close to 32 wide. The FP64 work per segment is cut without changing a single
// rounding: (1) rho / temperature / fleck / sigma_s come as one 32-byte CellProps record
// evaluated once per cell per cycle; (2) h*nu and 2 h nu^3 / c^2 are hoisted per photon;
// (3) the cell index after a step follows from which face nudge fired instead of three more
// floor((x - xmin) / dx) — ptcl_transport_step leaves a photon either eps_imc_offset cell
// widths beyond exactly one face per direction or at least that far inside, and rounding is
// seven orders of magnitude below that margin (VALIDATE re-derives the index with Xtoijk and
// counts disagreements into n_error; tests/test_parity_gpu.py runs it).
#ifndef JAYBENNE_B200_JB_TRANSPORT_IMC_CUH_
#define JAYBENNE_B200_JB_TRANSPORT_IMC_CUH_

static constexpr int IMC_TPB = 128;

__device__ __forceinline__ CellProps load_props(const CellProps *p) {
  // 2 x 16-byte read-only loads of one 32-byte aligned record
  const double2 a = __ldg(reinterpret_cast<const double2 *>(p));
  const double2 b = __ldg(reinterpret_cast<const double2 *>(p) + 1);
  CellProps c;
  c.rho = a.x; c.temp = a.y; c.ff = b.x; c.ss = b.y;
  return c;
}

template <int NDIM, int MINB, bool VALIDATE>
__global__ void __launch_bounds__(IMC_TPB, MINB)
    k_transport_imc(StepParams P, long long first, long long n, double t_start, double dt, int refill_min) {
  const DMesh &m = P.mesh;
  const double t_end = t_start + dt;
  const double vv = P.cl.c_code;
  const int nx1 = m.nxb[0], nx2 = m.nxb[1];
  const unsigned FULL = 0xffffffffu;
  const unsigned lane = threadIdx.x & 31u;
  const unsigned lt_mask = (1u << lane) - 1u;
  const long long gw = (blockIdx.x * (long long)blockDim.x + threadIdx.x) >> 5;
  const long long nw = ((long long)gridDim.x * blockDim.x) >> 5;
  // this warp's slots: a contiguous range, a multiple of 32 long so ranges start on 256 B lines
  const long long per = (((n + nw - 1) / nw) + 31) & ~31LL;
  long long cur = first + gw * per;
  long long end = cur + per;
  if (end > first + n) end = first + n;

  bool have = false;     // this lane holds a photon in flight
  long long idx = 0;     // its slot
  Photon p;
  double ww = 0.0;
  PhotonFreq q;
  Rng rng;
  BlkReg B;
  p.x = p.y = p.z = p.t = p.vx = p.vy = p.vz = 0.0;
  p.ip = p.jp = p.kp = 0;
  q.f = q.hnu = q.bpref = 0.0;
  rng.init(0, 0, 0, 0, P.key0, P.key1_particle);
  B.lb = 0; B.gid = 0; B.dx_push = 0.0;
#pragma unroll
  for (int d = 0; d < 3; ++d) B.xmin[d] = B.xmax[d] = B.dx[d] = B.xmin0[d] = 0.0;

  unsigned c_seg = 0, c_scat = 0, c_cross = 0, c_abs = 0, c_dead = 0;
  double e_abs = 0.0;

  for (;;) {
    // ---- refill idle lanes from the warp's range ----------------------------------------
    while (cur < end) {
      const unsigned need = __ballot_sync(FULL, !have);
      const int cnt = __popc(need);
      if (cnt < refill_min && cnt < 32) break;
      if (cnt == 0) break;
      const long long my = cur + __popc(need & lt_mask);
      cur += cnt;
      if (!have && my < end) {
        const int lb = P.s.loc[my].x;
        if (lb >= 0) {
          const double t0 = P.s.t[my];
          if (t0 < t_end) {  // census photons of an earlier iteration are skipped (transport.cpp:98)
            idx = my;
            p.t = t0;
            p.x = P.s.x[my]; p.y = P.s.y[my]; p.z = P.s.z[my];
            p.vx = P.s.vx[my]; p.vy = P.s.vy[my]; p.vz = P.s.vz[my];
            ww = P.s.w[my];
            q = photon_freq(P.cl, P.s.e[my]);
            const uint2 ia = P.s.ida[my], ib = P.s.idb[my];
            rng.init(ia.x, ia.y, ib.x, ib.y, P.key0, P.key1_particle);
            load_block(m, m.local_gid[lb], B);
            xtoijk<NDIM>(B, p, p.ip, p.jp, p.kp);
            have = true;
          }
        }
      }
    }
    if (__ballot_sync(FULL, have) == 0u) break;  // range exhausted and nothing in flight

    if (have) {
      // ---- one segment of the while (t < t_end) loop, transport.cpp:112-171 ------------
      c_seg++;
      CellEnv c;
      {
        const double xcen = xc(B, m.ng[0], 0, p.ip), ycen = xc(B, m.ng[1], 1, p.jp), zcen = xc(B, m.ng[2], 2, p.kp);
        c.xl = xcen - 0.5 * B.dx[0]; c.xu = xcen + 0.5 * B.dx[0];
        c.yl = ycen - 0.5 * B.dx[1]; c.yu = ycen + 0.5 * B.dx[1];
        c.zl = zcen - 0.5 * B.dx[2]; c.zu = zcen + 0.5 * B.dx[2];
      }
      const long long ci = (long long)B.lb * m.ncell + ((long long)p.kp * nx2 + p.jp) * nx1 + p.ip;
      const CellProps cp = load_props(P.f.props + ci);
      c.ff = cp.ff;
      c.ss = cp.ss;
      c.aa = absorption_coeff_hoisted(P.cl, cp.rho, cp.temp, q);
      bool is_absorbed = false, is_scattered = false;
      int di, dj, dk;
      imc_step_idx<NDIM>(rng, t_end, c, vv, B.dx_push, p, is_absorbed, is_scattered, di, dj, dk);
      if (VALIDATE) {
        int ti, tj, tk;
        xtoijk<NDIM>(B, p, ti, tj, tk);
        if (ti != p.ip + di || tj != p.jp + dj || tk != p.kp + dk) atomicAdd(&P.stats->n_error, 1ull);
      }
      p.ip += di; p.jp += dj; p.kp += dk;
      bool gone = false;
      if (!on_block(m, p.ip, p.jp, p.kp)) {
        // ---- left the meshblock: swarm BCs + hand-over (MeshSend / MeshReceive on device) ----
        if (!apply_swarm_bc<NDIM>(m, p)) {
          atomicAdd(&P.stats->n_escaped, 1ull);
          atomicAdd(&P.ebins[EB_ESCAPED * EBINS + (unsigned)(gw & (EBINS - 1))], ww);
          gone = true;
        } else {
          const double pos[3] = {p.x, p.y, p.z};
          const int gid = find_block(m, pos);
          if (gid < 0) {
            atomicAdd(&P.stats->n_error, 1ull);
            gone = true;
          } else {
            c_cross++;
            const DBlock &nb = m.blocks[gid];
            if (nb.rank != m.my_rank) {
              const unsigned long long slot = atomicAdd(&P.out.count[nb.rank], 1ull);
              if ((long long)slot < P.out.cap) {
                ExRecord r;
                r.x = p.x; r.y = p.y; r.z = p.z; r.t = p.t; r.w = ww; r.e = P.s.e[idx];
                r.vx = p.vx; r.vy = p.vy; r.vz = p.vz;
                r.rng[0] = rng.id0; r.rng[1] = rng.id1; r.rng[2] = rng.id2; r.rng[3] = rng.n;
                r.dest_gid = gid;
                r.pad = 0;
                P.out.rec[(size_t)nb.rank * P.out.cap + slot] = r;
                atomicAdd(&P.stats->n_sent, 1ull);
              } else {
                atomicAdd(&P.stats->n_overflow, 1ull);
              }
              gone = true;
            } else {
              load_block(m, gid, B);
              xtoijk<NDIM>(B, p, p.ip, p.jp, p.kp);
            }
          }
        }
      } else if (is_absorbed) {
        atomicAdd(&P.f.edelta[(long long)B.lb * m.ncell + ((long long)p.kp * nx2 + p.jp) * nx1 + p.ip], ww);
        c_abs++;
        e_abs += ww;
        gone = true;
      } else if (is_scattered) {
        scatter_iso(rng, vv, p.vx, p.vy, p.vz);
        c_scat++;
      }
      // ---- photon finished for this call? write it back and free the lane -----------------
      if (gone) {
        P.s.loc[idx].x = -1;
        c_dead++;
        have = false;
      } else if (!(p.t < t_end)) {
        P.s.x[idx] = p.x; P.s.y[idx] = p.y; P.s.z[idx] = p.z; P.s.t[idx] = p.t;
        P.s.vx[idx] = p.vx; P.s.vy[idx] = p.vy; P.s.vz[idx] = p.vz;
        P.s.loc[idx] = make_int2(B.lb, (p.kp * nx2 + p.jp) * nx1 + p.ip);
        P.s.idb[idx] = make_uint2(rng.id2, rng.n);
        have = false;
      }
    }
  }
  DStats *S = P.stats;
  warp_add_u64(&S->n_segments, c_seg);
  warp_add_u64(&S->n_scatters, c_scat);
  warp_add_u64(&S->n_crossings, c_cross);
  warp_add_u64(&S->n_absorbed, c_abs);
  warp_add_u64(&S->n_dead, c_dead);
  warp_add_energy(P, EB_ABSORBED, e_abs);
}

static TransportTuning g_tuning = {1, 4, 4, 0};
void set_transport_tuning(const TransportTuning &t) { g_tuning = t; }

template <int NDIM, int MINB>
static void launch_transport_imc_occ(const StepParams &P, long long first, long long n, double t_start, double dt,
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
  if (g_tuning.validate)
    k_transport_imc<NDIM, MINB, true><<<(unsigned)blocks, IMC_TPB, 0, st>>>(P, first, n, t_start, dt, g_tuning.refill_min);
  else
    k_transport_imc<NDIM, MINB, false><<<(unsigned)blocks, IMC_TPB, 0, st>>>(P, first, n, t_start, dt, g_tuning.refill_min);
}
template <int NDIM>
static void launch_transport_imc(const StepParams &P, long long first, long long n, double t_start, double dt,
                                 cudaStream_t st) {
  switch (g_tuning.ctas_per_sm) {
    case 3: launch_transport_imc_occ<NDIM, 3>(P, first, n, t_start, dt, st); break;
    case 5: launch_transport_imc_occ<NDIM, 5>(P, first, n, t_start, dt, st); break;
    default: launch_transport_imc_occ<NDIM, 4>(P, first, n, t_start, dt, st); break;
  }
}

#endif  // JAYBENNE_B200_JB_TRANSPORT_IMC_CUH_
