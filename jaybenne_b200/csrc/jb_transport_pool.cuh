// jb_transport_pool.cuh — the IMC history kernel, v10: every warp keeps a private POOL of photons in
// shared memory and works on it in phases that each run (nearly) full width.
// Included by jb_kernels.cu after jb_transport_imc.cuh (shares Rng2, imc_step_lockstep, scatter_iso2,
// load_props, ld_once and the ABS_* closure selectors).
//
// Reference behaviour: TransportPhotons, src/jaybenne/transport.cpp:28-181, with ptcl_transport_step /
// scatter (transport_utils.hpp:111-160, scattering.hpp:21-29), the swarm boundary conditions
// (boundaries.hpp:24-84) and the block hand-over (jaybenne.cpp:26-61). Same per-photon Philox streams,
// same draw order per photon and the same arithmetic as k_transport_imc, so results are bit-identical;
// only the order in which a warp visits its photons differs.
//
// Why (profiles/r01_v9_ncu_transport_imc_cells64.md): v9 kept one photon per lane in registers and ran
// 21.8 of 32 lanes wide — 41 % of the issued warp-instructions sat in side paths that a few lanes take
// each iteration (the scatter's extra Philox block + sqrt + sincos at ~6 lanes, the refill at ~4, the
// census write-back at ~3, the absorption deposit at ~2), while 228 KB of shared memory went unused.
// Here the photon state lives in shared memory (S slots per warp, SoA, 120 B per slot) and a slot is in
// one of five states; the warp looks at the state counts and runs ONE phase per iteration:
//   * retire + refill  (>= refill_min slots hold no live photon): absorbed photons deposit their energy,
//     census photons are written back to the swarm arrays, and the freed slots take the next photons of
//     the warp's chunk — coalesced loads, up to 32 at a time;
//   * scatter  (32 photons wait for a new direction, or the ready photons run short): Philox block +
//     scatter for up to 32 photons at once;
//   * segment  (otherwise): up to 32 READY photons advance one iteration of the while (t < t_end) loop.
// Only the block hand-over and the domain boundary conditions (1.5 % of the segments) stay inside the
// segment phase. Registers hold a photon only for the duration of a phase.
//
// v13 — registers before warps: 5 CTAs of 4 warps per SM (96 registers, no spills, ~95 KB of L1 beside 160 KB of
// pools) beat 7 CTAs (72 registers, 140 B spilled, ~30 KB of L1) by 3.5 %: 7.6 instead of 10.6 cycles between a
// warp's instructions (profiles/r02_v13_ncu_transport_pool_cells64.md, profiles/r02_v10_tuning_cells256.md).
//
// v12 — the pool stays ORDERED. A warp owns the 1024-slot chunks it claims, so it writes its census photons
// back to the FRONT of their chunk, in the order they finish (the k-th photon to retire goes to chunk + k,
// which the warp's read cursor has already passed), and marks every slot it loads as a hole. After the
// kernel each chunk holds its survivors packed at the front and holes behind them: nothing ever moves to
// another chunk, so the (meshblock, cell) order the sourcing kernel laid down only blurs by the few cells a
// photon travels per cycle, and the per-cell records of a warp's window stay L2-resident cycle after cycle
// (the tail-into-holes squeeze scattered the last blocks' photons over the whole pool:
// profiles/r02_v10_bench_cells256.json, continuing run 34 -> 54 ps per segment over 8 cycles against
// 34 -> 37 ps with a full sort every cycle). Weight and energy now travel through the pool with the photon
// (their old slot may already hold another photon when they are needed).
#ifndef JAYBENNE_B200_JB_TRANSPORT_POOL_CUH_
#define JAYBENNE_B200_JB_TRANSPORT_POOL_CUH_

static constexpr int POOL_TPB = 128;
static constexpr int POOL_WARPS = POOL_TPB / 32;

// slot states
enum { PS_FREE = 0, PS_READY = 1, PS_PEND = 2, PS_CENSUS = 3, PS_ABS = 4, PS_NONE = 7 };

template <int S, int NQ>
struct __align__(16) WarpPool {
  double x[S], y[S], z[S], t[S], vx[S], vy[S], vz[S];
  double w[S];       // weight (energy the photon carries)
  double q[NQ][S];   // q[0] = photon frequency (swarm field `energy`), q[1] = B_nu prefactor 2 h nu^3 / c^2
  uint2 ida[S];      // (birth index, birth gid)
  uint2 idb[S];      // (source call, draws consumed)
  uint2 carry[S];    // words 2, 3 of Philox block n >> 1 while n is odd (Rng2)
  int2 blk[S];       // (local block, gid)
  unsigned idx[S];   // swarm slot the photon was loaded from, relative to `first` (tells its chunk)
  int cell[S];       // ip | jp << 10 | kp << 20 (interior indices)
  unsigned char st[S > 64 ? 96 : 64];
  unsigned char list[S > 64 ? 96 : 64];
};

// The lanes take the first 32 slots of the set (m0: slots 0-31, m1: slots 32-63), word m1 first when
// `flip` (alternating the order keeps every slot from waiting more than one round). Returns the lane's
// slot or -1; cnt = number of lanes served.
// Three mask words (pools of up to 96 slots): the word order rotates with `rot` (0: 0 1 2, 1: 1 2 0, 2: 2 0 1).
__device__ __forceinline__ int pool_pick3(unsigned char *list, const unsigned m0, const unsigned m1, const unsigned m2,
                                          const unsigned rot, const unsigned lane, const unsigned lt, int &cnt) {
  const int c0 = __popc(m0), c1 = __popc(m1), c2 = __popc(m2);
  const int o0 = rot == 0 ? 0 : (rot == 1 ? c1 + c2 : c2);
  const int o1 = rot == 0 ? c0 : (rot == 1 ? 0 : c2 + c0);
  const int o2 = rot == 0 ? c0 + c1 : (rot == 1 ? c1 : 0);
  if ((m0 >> lane) & 1u) list[__popc(m0 & lt) + o0] = (unsigned char)lane;
  if ((m1 >> lane) & 1u) list[__popc(m1 & lt) + o1] = (unsigned char)(lane + 32u);
  if ((m2 >> lane) & 1u) list[__popc(m2 & lt) + o2] = (unsigned char)(lane + 64u);
  __syncwarp();
  cnt = min(c0 + c1 + c2, 32);
  const int s = (int)lane < cnt ? (int)list[lane] : -1;
  __syncwarp();
  return s;
}
__device__ __forceinline__ int pool_pick(unsigned char *list, const unsigned m0, const unsigned m1, const bool flip,
                                         const unsigned lane, const unsigned lt, int &cnt) {
  const int c0 = __popc(m0), c1 = __popc(m1);
  if ((m0 >> lane) & 1u) list[__popc(m0 & lt) + (flip ? c1 : 0)] = (unsigned char)lane;
  if ((m1 >> lane) & 1u) list[__popc(m1 & lt) + (flip ? 0 : c0)] = (unsigned char)(lane + 32u);
  __syncwarp();
  cnt = min(c0 + c1, 32);
  const int s = (int)lane < cnt ? (int)list[lane] : -1;
  __syncwarp();
  return s;
}

// Guided claim of the next piece of the slot range: IMC_CHUNK slots while plenty remain, smaller pieces towards the
// end (a piece = what is left / twice the warps, at least 64, a multiple of 32), so that the last warps finish
// together — a short range (the arrivals of a later exchange iteration, a rank's 1/8 of the pool) otherwise ran on a
// few warps with 1024 photons each (profiles/r02_v12_bench_cells256_gpus8_*.json: iterations 3+ of a cycle). The
// look at the counter is only a hint; the atomic add is the claim. Out of line: once per ~1000 photons, and its
// 64-bit division must not cost the transport loop registers. `fixed` (JB_POOL_FIXED_CHUNK) keeps IMC_CHUNK for A/B.
__device__ __noinline__ unsigned long long pool_claim(unsigned long long *work, long long n, long long nwarps, bool fixed,
                                                      unsigned &piece) {
  piece = IMC_CHUNK;
  if (!fixed) {
    const long long seen = (long long)*(volatile unsigned long long *)work;
    const long long left = n - seen;
    if (left < 2ll * nwarps * IMC_CHUNK) {
      const long long share = left / (2ll * nwarps);
      piece = share < 64 ? 64u : ((unsigned)share & ~31u);
    }
  }
  return atomicAdd(work, (unsigned long long)piece);
}

template <int NDIM, int MINB, bool VALIDATE, int ABS, int S, int WARPS = POOL_WARPS>
__global__ void __launch_bounds__(WARPS * 32, MINB)
    k_transport_pool(StepParams P, long long first, long long n, double t_start, double dt, int refill_min,
                     int ready_low, int l2_prefetch) {
  constexpr int NQ = ABS == ABS_GRAY_DIRECT ? 1 : 2;
  static_assert(S > 32 && S <= 96, "two or three slots per lane");
  constexpr bool W3 = S > 64;  // three mask words
  constexpr unsigned NW = W3 ? 3u : 2u;
  __shared__ WarpPool<S, NQ> pools[WARPS];
  __shared__ unsigned sh_cnt[4];  // crossings, removed in the segment phase (escaped / sent / lost)

  const DMesh &m = P.mesh;
  if (P.range) {  // direct peer exchange: the arrivals' range lives on the device (P2PState::range)
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
  const unsigned lt = (1u << lane) - 1u;
  const long long gw = (blockIdx.x * (long long)blockDim.x + threadIdx.x) >> 5;
  WarpPool<S, NQ> &W = pools[threadIdx.x >> 5];

  W.st[lane] = PS_FREE;
  W.st[lane + 32u] = (lane + 32u < (unsigned)S) ? PS_FREE : PS_NONE;
  if (W3) W.st[lane + 64u] = (lane + 64u < (unsigned)S) ? PS_FREE : PS_NONE;
  if (threadIdx.x < 4) sh_cnt[threadIdx.x] = 0u;
  __syncthreads();

  // the warp's chunks of swarm slots (warp-uniform): the one being read, and the one before it, whose last
  // photons may still be in flight. Slot offsets are relative to `first`.
  unsigned cur = 0, end = 0;    // read cursor and end of the current chunk
  unsigned c_base = 0, wr = 0;  // current chunk: first slot and write cursor (wr <= cur: only slots already read)
  unsigned p_wr = 0;            // write cursor of the previous chunk (all of its slots have been read)
  bool exhausted = false;
  unsigned rot = 0;  // which mask word the segment phase serves first
  unsigned c_seg = 0, c_scat = 0, c_abs = 0, bad = 0;
  double e_abs = 0.0;

  for (;;) {
    __syncwarp();
    const int s0 = W.st[lane], s1 = W.st[lane + 32u], s2 = W3 ? (int)W.st[lane + 64u] : (int)PS_NONE;
    const unsigned r0 = __ballot_sync(FULL, s0 == PS_READY), r1 = __ballot_sync(FULL, s1 == PS_READY);
    const unsigned q0 = __ballot_sync(FULL, s0 == PS_PEND), q1 = __ballot_sync(FULL, s1 == PS_PEND);
    const unsigned f0 = __ballot_sync(FULL, s0 == PS_FREE), f1 = __ballot_sync(FULL, s1 == PS_FREE);
    const unsigned w0 = __ballot_sync(FULL, s0 == PS_CENSUS || s0 == PS_ABS);
    const unsigned w1 = __ballot_sync(FULL, s1 == PS_CENSUS || s1 == PS_ABS);
    const unsigned r2 = W3 ? __ballot_sync(FULL, s2 == PS_READY) : 0u, q2 = W3 ? __ballot_sync(FULL, s2 == PS_PEND) : 0u;
    const unsigned f2 = W3 ? __ballot_sync(FULL, s2 == PS_FREE) : 0u;
    const unsigned w2 = W3 ? __ballot_sync(FULL, s2 == PS_CENSUS || s2 == PS_ABS) : 0u;
    const int n_ready = __popc(r0) + __popc(r1) + __popc(r2), n_pend = __popc(q0) + __popc(q1) + __popc(q2);
    const int n_ret = __popc(w0) + __popc(w1) + __popc(w2), n_nonlive = n_ret + __popc(f0) + __popc(f1) + __popc(f2);
    // two write cursors are kept, so the next chunk is claimed only once the chunk BEFORE the current one has
    // no photon left in the pool (a chunk is >= 16 refills long: by then they have all retired, except in
    // media where single histories run to hundreds of segments)
    bool can_claim = !exhausted;
    if (cur >= end && can_claim) {
      const bool old0 = s0 != PS_FREE && W.idx[lane] < c_base;
      const bool old1 = s1 != PS_FREE && s1 != PS_NONE && W.idx[lane + 32u] < c_base;
      const bool old2 = W3 && s2 != PS_FREE && s2 != PS_NONE && W.idx[W3 ? lane + 64u : lane] < c_base;
      can_claim = !__any_sync(FULL, old0 || old1 || old2);
    }
    const bool input_left = cur < end || can_claim;
    const bool idle = n_ready + n_pend == 0;

    // =====================================================================================
    // retire + refill
    // =====================================================================================
    if ((n_nonlive >= refill_min || idle) && (input_left || n_ret > 0)) {
      int cnt;
      const int s = pool_pick3(W.list, input_left ? (w0 | f0) : w0, input_left ? (w1 | f1) : w1,
                               input_left ? (w2 | f2) : w2, 0u, lane, lt, cnt);
      {
        const int old = s >= 0 ? (int)W.st[s] : PS_FREE;
        // the photon reached t_end (the reference's loop leaves it in place): back to the swarm arrays, packed
        // at the front of the chunk it came from — position = the chunk's write cursor + rank among the lanes
        const bool cen = old == PS_CENSUS;
        const bool in_cur = cen && W.idx[s] >= c_base;
        const unsigned m_cur = __ballot_sync(FULL, in_cur), m_prev = __ballot_sync(FULL, cen && !in_cur);
        if (cen) {
          const long long idx = first + (long long)(in_cur ? wr + __popc(m_cur & lt) : p_wr + __popc(m_prev & lt));
          const int cl = W.cell[s];
          P.s.x[idx] = W.x[s]; P.s.y[idx] = W.y[s]; P.s.z[idx] = W.z[s]; P.s.t[idx] = W.t[s];
          P.s.vx[idx] = W.vx[s]; P.s.vy[idx] = W.vy[s]; P.s.vz[idx] = W.vz[s];
          P.s.w[idx] = W.w[s]; P.s.e[idx] = W.q[0][s];
          P.s.loc[idx] = make_int2(W.blk[s].x, (((cl >> 20) & 1023) * nx2 + ((cl >> 10) & 1023)) * nx1 + (cl & 1023));
          P.s.ida[idx] = W.ida[s];
          P.s.idb[idx] = W.idb[s];
        } else if (old == PS_ABS) {
          // absorbed (transport.cpp:157-163): deposit; the particle's slot became a hole when it was loaded
          const int cl = W.cell[s];
          const double ww = W.w[s];
          atomicAdd(&P.f.edelta[(long long)W.blk[s].x * m.ncell +
                                ((long long)((cl >> 20) & 1023) * nx2 + ((cl >> 10) & 1023)) * nx1 + (cl & 1023)], ww);
          e_abs += ww;
          c_abs++;
        }
        wr += (unsigned)__popc(m_cur);
        p_wr += (unsigned)__popc(m_prev);
      }
      // the next photons of the warp's chunk
      bool in_chunk = false;
      unsigned my32 = 0;
      if (input_left) {
        if (cur >= end) {
          // guided claim (pool_claim): IMC_CHUNK slots while plenty remain, smaller pieces towards the end of the range
          unsigned long long c0 = 0;
          unsigned piece = IMC_CHUNK;
          if (lane == 0) c0 = pool_claim(P.work, n, (long long)gridDim.x * WARPS, (l2_prefetch & 2) != 0, piece);
          c0 = __shfl_sync(FULL, c0, 0);
          piece = __shfl_sync(FULL, piece, 0);
          if ((long long)c0 >= n) {
            exhausted = true;
          } else {
            p_wr = wr;  // the chunk just read becomes the previous one
            cur = (unsigned)c0;
            c_base = wr = cur;
            end = (long long)c0 + piece > n ? (unsigned)n : cur + piece;
          }
        }
        if (cur < end) {
          my32 = cur + lane;
          in_chunk = s >= 0 && my32 < end;
          cur = min(end, cur + (unsigned)cnt);
          if ((l2_prefetch & 1) && lane < 11u && cur < end) {
            // the window the NEXT refill will read, pulled into L2 by the bulk-copy (TMA) unit while this
            // one's photons run: one cp.async.bulk.prefetch.L2 per swarm array (16-byte aligned, 256 B)
            const double *arr = lane == 0 ? P.s.x : lane == 1 ? P.s.y : lane == 2 ? P.s.z : lane == 3 ? P.s.t
                              : lane == 4 ? P.s.vx : lane == 5 ? P.s.vy : lane == 6 ? P.s.vz : lane == 7 ? P.s.e
                              : lane == 8 ? (const double *)P.s.loc : lane == 9 ? (const double *)P.s.ida
                                                                                 : (const double *)P.s.idb;
            const unsigned long long a0 = (unsigned long long)(arr + first + (long long)cur) & ~15ull;
            const unsigned bytes = min(256u, ((end - cur) * 8u + 15u) & ~15u);
            asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(a0), "r"(bytes) : "memory");
          }
        }
      }
      int newst = PS_FREE;
      if (in_chunk) {
        const long long my = first + (long long)my32;
        // every array of the slot is requested before the first test: one DRAM round trip per refill
        const int2 lc = ld_once(&P.s.loc[my]);
        const double t0 = ld_once(&P.s.t[my]);
        const double l_x = ld_once(&P.s.x[my]), l_y = ld_once(&P.s.y[my]), l_z = ld_once(&P.s.z[my]);
        const double l_vx = ld_once(&P.s.vx[my]), l_vy = ld_once(&P.s.vy[my]), l_vz = ld_once(&P.s.vz[my]);
        const double l_e = ld_once(&P.s.e[my]), l_w = ld_once(&P.s.w[my]);
        const uint2 ia = ld_once(&P.s.ida[my]), ib = ld_once(&P.s.idb[my]);
        if (lc.x >= 0) {
          // the photon lives in the pool from here on and its slot is a hole; one that is already at census
          // (transport.cpp:98 skips it) only passes through to be packed with the others
          P.s.loc[my] = make_int2(-1, 0);
          W.idx[s] = my32;
          W.t[s] = t0;
          W.x[s] = l_x; W.y[s] = l_y; W.z[s] = l_z;
          W.vx[s] = l_vx; W.vy[s] = l_vy; W.vz[s] = l_vz;
          W.w[s] = l_w;
          W.q[0][s] = l_e;
          if (NQ > 1) W.q[NQ - 1][s] = photon_freq(P.cl, l_e).bpref;
          W.ida[s] = ia;
          W.idb[s] = ib;
          // an odd draw counter carries the second half of block n >> 1 (Rng2)
          uint2 cy = make_uint2(0u, 0u);
          if (ib.y & 1u) {
            Rng2 g;
            g.id0 = ia.x; g.id1 = ia.y; g.id2 = ib.x;
            uint32_t w0_, w1_;
            g.block(ib.y >> 1, P.rk, w0_, w1_, cy.x, cy.y);
          }
          W.carry[s] = cy;
          const int gid = m.local_gid[lc.x];
          W.blk[s] = make_int2(lc.x, gid);
          // the stored flat cell is what Xtoijk (transport.cpp:96) returns for a photon at rest
          int ip, jp, kp;
          if (sh1 >= 0) {
            ip = lc.y & (nx1 - 1);
            jp = (lc.y >> sh1) & (nx2 - 1);
            kp = lc.y >> (sh1 + sh2);
          } else {
            ip = lc.y % nx1;
            const int rest = lc.y / nx1;
            jp = rest % nx2;
            kp = rest / nx2;
          }
          W.cell[s] = ip | (jp << 10) | (kp << 20);
          if (VALIDATE) {
            BlkReg B;
            load_block(m, gid, B);
            Photon pv;
            pv.x = l_x; pv.y = l_y; pv.z = l_z;
            int ti, tj, tk;
            xtoijk<NDIM>(B, pv, ti, tj, tk);
            if (ti != ip || tj != jp || tk != kp) atomicAdd(&P.stats->n_validate, 1ull);
          }
          newst = t0 < t_end ? PS_READY : PS_CENSUS;
        }
      }
      if (s >= 0) W.st[s] = (unsigned char)newst;
      continue;
    }
    if (idle) break;  // no live photon, nothing to retire, no input left

    // =====================================================================================
    // scatter (scattering.hpp:21-29): draws #n, #n+1 of the photon's stream give the new direction
    // =====================================================================================
    if (n_pend >= 32 || (n_pend > 0 && n_ready < ready_low)) {
      int cnt;
      const int s = pool_pick3(W.list, q0, q1, q2, 0u, lane, lt, cnt);
      if (s >= 0) {
        Rng2 rng;
        const uint2 ia = W.ida[s], ib = W.idb[s], cy = W.carry[s];
        rng.id0 = ia.x; rng.id1 = ia.y; rng.id2 = ib.x; rng.n = ib.y; rng.clo = cy.x; rng.chi = cy.y;
        double xi_mu, xi_phi, vx, vy, vz;
        rng.draw2(P.rk, xi_mu, xi_phi);
        scatter_iso2(xi_mu, xi_phi, vv, vx, vy, vz);
        W.vx[s] = vx; W.vy[s] = vy; W.vz[s] = vz;
        W.idb[s].y = rng.n;
        W.carry[s] = make_uint2(rng.clo, rng.chi);
        W.st[s] = PS_READY;
        c_scat++;
      }
      continue;
    }

    // =====================================================================================
    // segment: one iteration of the while (t < t_end) loop, transport.cpp:112-171
    // =====================================================================================
    rot = rot + 1u == NW ? 0u : rot + 1u;
    int cnt;
    const int s = pool_pick3(W.list, r0, r1, r2, rot, lane, lt, cnt);
    if (s >= 0) {
      c_seg++;
      Photon p;
      Rng2 rng;
      const int2 bk = W.blk[s];
      const int cl = W.cell[s];
      int lb = bk.x, gid = bk.y;
      p.ip = cl & 1023; p.jp = (cl >> 10) & 1023; p.kp = (cl >> 20) & 1023;
      // loads first: their latency hides behind the Philox block
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
        // h nu as photon_freq() forms it (nu * freq_unit, then * h): two multiplies instead of 8 bytes per slot
        const double hnu = P.cl.h * (W.q[0][s] * P.cl.freq_unit), bpref = W.q[NQ - 1][s];
        const double r = P.cl.rho_unit * cp.rho, T = P.cl.temp_unit * cp.temp;
        const double x = hnu / (P.cl.kb * T);
        const double Bnu = bpref / jbm::expm1(x);
        const double jnu = r * P.cl.kappa * Bnu;
        c.aa = robust_ratio_pos(jnu, Bnu) * P.cl.length_unit;
      } else {
        PhotonFreq q;
        q.f = W.q[0][s] * P.cl.freq_unit; q.hnu = P.cl.h * q.f; q.bpref = W.q[NQ - 1][s];
        c.aa = absorption_coeff_hoisted(P.cl, cp.rho, cp.temp, q);
      }
      bool is_absorbed = false, is_scattered = false;
      int di, dj, dk;
      imc_step_lockstep<NDIM, VALIDATE>(xi_abs, xi_sc, t_end, c, vv, dx_cell, p, is_absorbed, is_scattered, di, dj, dk, bad);
      if (VALIDATE) {
        BlkReg B;
        load_block(m, gid, B);
        int ti, tj, tk;
        xtoijk<NDIM>(B, p, ti, tj, tk);
        if (ti != p.ip + di || tj != p.jp + dj || tk != p.kp + dk) atomicAdd(&P.stats->n_validate, 1ull);
      }
      p.ip += di; p.jp += dj; p.kp += dk;
      int newst = PS_READY;
      if (!on_block(m, p.ip, p.jp, p.kp)) {
        // ---- left the meshblock: swarm BCs + hand-over (MeshSend / MeshReceive on device) ----
        const int bx = p.ip < 0 ? -1 : (p.ip >= nx1 ? 1 : 0);
        const int by = p.jp < 0 ? -1 : (p.jp >= nx2 ? 1 : 0);
        const int bz = p.kp < 0 ? -1 : (p.kp >= nx3 ? 1 : 0);
        bool gone = false;
        if (!apply_swarm_bc<NDIM>(m, p)) {
          atomicAdd(&P.stats->n_escaped, 1ull);
          atomicAdd(&P.ebins[EB_ESCAPED * EBINS + (unsigned)(gw & (EBINS - 1))], W.w[s]);
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
            bool wrong = vg != ngid;
            if (!wrong) {
              BlkReg B;
              load_block(m, vg, B);
              int ti, tj, tk;
              xtoijk<NDIM>(B, p, ti, tj, tk);
              wrong = ti != p.ip || tj != p.jp || tk != p.kp;
            }
            if (wrong) atomicAdd(&P.stats->n_validate, 1ull);
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
                r.x = p.x; r.y = p.y; r.z = p.z; r.t = p.t; r.w = W.w[s]; r.e = W.q[0][s];
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
              W.blk[s] = make_int2(lb, gid);
            }
          }
        }
        if (gone) {  // (its slot became a hole when it was loaded)
          atomicAdd(&sh_cnt[1], 1u);
          newst = PS_FREE;
        } else {
          W.vx[s] = p.vx; W.vy[s] = p.vy; W.vz[s] = p.vz;  // a reflecting boundary turned the photon round
        }
      } else if (is_absorbed) {
        newst = PS_ABS;  // deposit and removal happen in the retire phase
      } else if (is_scattered) {
        if (p.t < t_end) {
          newst = PS_PEND;  // new direction in the scatter phase
        } else {            // scattered exactly into census (rounding): finish it here
          double xi_mu, xi_phi;
          rng.draw2(P.rk, xi_mu, xi_phi);
          scatter_iso2(xi_mu, xi_phi, vv, p.vx, p.vy, p.vz);
          W.vx[s] = p.vx; W.vy[s] = p.vy; W.vz[s] = p.vz;
          c_scat++;
        }
      }
      if (newst == PS_READY && !(p.t < t_end)) newst = PS_CENSUS;
      W.x[s] = p.x; W.y[s] = p.y; W.z[s] = p.z; W.t[s] = p.t;
      W.cell[s] = p.ip | (p.jp << 10) | (p.kp << 20);
      W.idb[s].y = rng.n;
      W.carry[s] = make_uint2(rng.clo, rng.chi);
      W.st[s] = (unsigned char)newst;
    }
  }
  DStats *St = P.stats;
  warp_add_u64(&St->n_segments, c_seg);
  warp_add_u64(&St->n_scatters, c_scat);
  warp_add_u64(&St->n_absorbed, c_abs);
  warp_add_u64(&St->n_dead, c_abs);
  if (VALIDATE) warp_add_u64(&St->n_validate, bad);
  __syncthreads();
  if (threadIdx.x == 0) {
    if (sh_cnt[0]) atomicAdd(&St->n_crossings, (unsigned long long)sh_cnt[0]);
    if (sh_cnt[1]) atomicAdd(&St->n_dead, (unsigned long long)sh_cnt[1]);
  }
  warp_add_energy(P, EB_ABSORBED, e_abs);
}

struct PoolTuning {
  int slots;        // see set_pool_tuning
  int refill_min;   // retire + refill when this many slots of a warp hold no live photon
  int ready_low;    // scatter early when fewer READY photons than this remain
  int l2_prefetch;  // 1: cp.async.bulk.prefetch.L2 of the next refill window (JB_POOL_PREFETCH; measured)
};
static PoolTuning g_pool = {64, 24, 32, 0};
void set_pool_prefetch(int on) { g_pool.l2_prefetch = (g_pool.l2_prefetch & ~1) | (on ? 1 : 0); }
void set_pool_fixed_chunk(int on) { g_pool.l2_prefetch = (g_pool.l2_prefetch & ~2) | (on ? 2 : 0); }
void set_pool_tuning(int slots, int refill_min, int ready_low) {
  // 64 (shipped): 64 slots, 5 CTAs of 4 warps per SM = 96 registers, no spills. A/B configurations
  // (profiles/r02_v10_tuning_cells256.md): 764 / 664 = 64 slots at 7 / 6 CTAs (72 / 80 registers, spills),
  // 588 = 88 slots (three per lane) at 5 CTAs, 56 = 56 slots at 8 CTAs (64 registers)
  g_pool.slots = (slots == 64 || slots == 664 || slots == 764 || slots == 588) ? slots : 56;
  g_pool.refill_min = max(1, min(refill_min, 32));
  g_pool.ready_low = max(1, min(ready_low, 32));
}

template <int NDIM, int MINB, bool VALIDATE, int S, int WARPS = POOL_WARPS>
static void launch_transport_pool_abs(const StepParams &P, long long first, long long n, double t_start, double dt,
                                      cudaStream_t st) {
  static int n_sm = 0;
  if (n_sm == 0) {
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&n_sm, cudaDevAttrMultiProcessorCount, dev);
    if (n_sm <= 0) n_sm = 148;
  }
  // persistent grid: MINB CTAs on each SM; a short range gets one warp per S photons
  long long blocks = (long long)n_sm * MINB;
  const long long need = (n + (long long)WARPS * S - 1) / ((long long)WARPS * S);
  if (blocks > need) blocks = need;
  if (blocks < 1) blocks = 1;
  const int rf = g_pool.refill_min, rl = g_pool.ready_low, pf = g_pool.l2_prefetch;
  // MINB CTAs x 4 pools: ask for the shared-memory carve-out they need (and no more: the rest is L1)
  static bool carved = false;
  if (!carved) {
    carved = true;
    const long long per_cta = (long long)sizeof(WarpPool<S, 2>) * WARPS + 16 + 1024;  // + the driver's 1 KB per CTA
    int pct = (int)((per_cta * MINB * 100 + 233471) / 233472);
    if (pct > 100) pct = 100;
    cudaFuncSetAttribute(k_transport_pool<NDIM, MINB, VALIDATE, ABS_GENERIC, S, WARPS>,
                         cudaFuncAttributePreferredSharedMemoryCarveout, pct);
    cudaFuncSetAttribute(k_transport_pool<NDIM, MINB, VALIDATE, ABS_GRAY_DIRECT, S, WARPS>,
                         cudaFuncAttributePreferredSharedMemoryCarveout, pct);
    cudaFuncSetAttribute(k_transport_pool<NDIM, MINB, VALIDATE, ABS_GRAY_KIRCHHOFF, S, WARPS>,
                         cudaFuncAttributePreferredSharedMemoryCarveout, pct);
  }
  if (P.cl.opacity_model != JB_OPAC_GRAY)
    k_transport_pool<NDIM, MINB, VALIDATE, ABS_GENERIC, S, WARPS><<<(unsigned)blocks, WARPS * 32, 0, st>>>(P, first, n, t_start, dt, rf, rl, pf);
  else if (P.cl.gray_mode == JB_GRAY_DIRECT)
    k_transport_pool<NDIM, MINB, VALIDATE, ABS_GRAY_DIRECT, S, WARPS><<<(unsigned)blocks, WARPS * 32, 0, st>>>(P, first, n, t_start, dt, rf, rl, pf);
  else
    k_transport_pool<NDIM, MINB, VALIDATE, ABS_GRAY_KIRCHHOFF, S, WARPS><<<(unsigned)blocks, WARPS * 32, 0, st>>>(P, first, n, t_start, dt, rf, rl, pf);
}
// false: the range or the mesh does not fit the kernel's packed indices -> the caller takes k_transport_imc
template <int NDIM>
static bool launch_transport_pool(const StepParams &P, long long first, long long n, double t_start, double dt,
                                  cudaStream_t st) {
  if (n >= 0xffffffffLL || P.mesh.nxb[0] > 1024 || P.mesh.nxb[1] > 1024 || P.mesh.nxb[2] > 1024) return false;
  if (g_tuning.validate)
    launch_transport_pool_abs<NDIM, 4, true, 56>(P, first, n, t_start, dt, st);
  else if (g_pool.slots == 64)
    launch_transport_pool_abs<NDIM, 5, false, 64>(P, first, n, t_start, dt, st);
  else if (g_pool.slots == 764 && NDIM == 3)
    launch_transport_pool_abs<3, 7, false, 64>(P, first, n, t_start, dt, st);
  else if (g_pool.slots == 664 && NDIM == 3)
    launch_transport_pool_abs<3, 6, false, 64>(P, first, n, t_start, dt, st);
  else if (g_pool.slots == 588 && NDIM == 3)
    launch_transport_pool_abs<3, 5, false, 88>(P, first, n, t_start, dt, st);
  else
    launch_transport_pool_abs<NDIM, 8, false, 56>(P, first, n, t_start, dt, st);
  return true;
}

#endif  // JAYBENNE_B200_JB_TRANSPORT_POOL_CUH_
