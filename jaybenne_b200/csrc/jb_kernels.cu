// jb_kernels.cu — hand-written CUDA kernels (sm_100a) of the jaybenne IMC/DDMC hot path.
//
//   k_fleck / k_face_prob     UpdateDerivedTransportFields      src/jaybenne/jaybenne.cpp:285-492
//   k_source_count / _emit    SourcePhotons pass 1 / pass 2     src/jaybenne/sourcing.cpp:25-205
//   k_transport               TransportPhotons(_DDMC) + the swarm boundary conditions and the
//                             on-device meshblock hand-over     src/jaybenne/transport.cpp:28-181,
//                             transport_ddmc.cpp:28-237, boundaries.hpp:24-84, jaybenne.cpp:26-61
//   k_bface                   SampleDDMCBlockFace               src/jaybenne/sample_ddmc_bface.cpp:81-427
//   k_tally                   EvaluateRadiationEnergy           src/jaybenne/jaybenne.cpp:514-564
//   k_update_fluid            UpdateFluid                       src/jaybenne/jaybenne.cpp:583-615
//   k_census                  CheckCompletion + census sums     src/jaybenne/transport.cpp:187-216
//
// Compile with -fmad=false: see jb_device.cuh.
#include "jb_kernels.cuh"

namespace jbd {

// ---------------------------------------------------------------------------------------
// small helpers
// ---------------------------------------------------------------------------------------
__device__ __forceinline__ unsigned long long warp_sum_u64(unsigned long long v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ double warp_sum_f64(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
  return v;
}
// all 32 lanes must call
__device__ __forceinline__ void warp_add_u64(unsigned long long *dst, unsigned long long v) {
  v = warp_sum_u64(v);
  if ((threadIdx.x & 31) == 0 && v) atomicAdd(dst, v);
}
__device__ __forceinline__ void warp_add_f64(double *dst, double v) {
  v = warp_sum_f64(v);
  if ((threadIdx.x & 31) == 0 && v != 0.0) atomicAdd(dst, v);
}
// energy sums: warp tree, then one add into the bin of this warp (see EBINS in jb_kernels.cuh)
__device__ __forceinline__ void warp_add_energy(const StepParams &P, int kind, double v) {
  v = warp_sum_f64(v);
  if ((threadIdx.x & 31) == 0 && v != 0.0) {
    const unsigned gw = (unsigned)((blockIdx.x * (unsigned long long)blockDim.x + threadIdx.x) >> 5);
    atomicAdd(&P.ebins[kind * EBINS + (gw & (EBINS - 1))], v);
  }
}
// pairwise fold of one kind's bins: *dst = sum(bins)
__global__ void k_fold_bins(const double *bins, double *dst) {
  __shared__ double sh[EBINS];
  sh[threadIdx.x] = bins[threadIdx.x];
  __syncthreads();
  for (int o = EBINS / 2; o > 0; o >>= 1) {
    if ((int)threadIdx.x < o) sh[threadIdx.x] += sh[threadIdx.x + o];
    __syncthreads();
  }
  if (threadIdx.x == 0) *dst = sh[0];
}
void launch_fold_bins(const StepParams &P, int kind, double *dst, cudaStream_t st) {
  k_fold_bins<<<1, EBINS, 0, st>>>(P.ebins + kind * EBINS, dst);
}

// UniformCartesian::Xc<dir>(idx), idx given as interior index
__device__ __forceinline__ double cell_center(const DBlock &b, const int ng, const int d, const int i) {
  return b.xmin0[d] + ((i + ng) + 0.5) * b.dx[d];
}

__device__ __forceinline__ int lattice_at(const DMesh &m, int lx, int ly, int lz) {
  return m.lattice[((size_t)lz * m.nlat[1] + ly) * m.nlat[0] + lx];
}

// gid of the leaf block containing pos (already inside the domain), -1 on failure.
__device__ int find_block(const DMesh &m, const double pos[3]) {
  int l[3] = {0, 0, 0};
  for (int d = 0; d < m.ndim; ++d) {
    const double w = (m.xmax[d] - m.xmin[d]) / m.nlat[d];
    int v = (int)floor((pos[d] - m.xmin[d]) / w);
    l[d] = min(max(v, 0), m.nlat[d] - 1);
  }
  for (int it = 0; it < 8; ++it) {
    const int gid = lattice_at(m, l[0], l[1], l[2]);
    const DBlock &b = m.blocks[gid];
    bool ok = true;
    for (int d = 0; d < m.ndim; ++d) {
      const double dxd = (b.xmax[d] - b.xmin[d]) / m.nxb[d];
      const int ii = (int)floor((pos[d] - b.xmin[d]) / dxd);
      const int s = 1 << (m.max_level - b.level);
      if (ii < 0) {
        l[d] = b.lx[d] * s - 1;
        ok = false;
      } else if (ii >= m.nxb[d]) {
        l[d] = (b.lx[d] + 1) * s;
        ok = false;
      }
      if (l[d] < 0 || l[d] >= m.nlat[d]) return -1;
    }
    if (ok) return gid;
  }
  return -1;
}

// ---------------------------------------------------------------------------------------
// UpdateDerivedTransportFields: Fleck factor (jaybenne.cpp:304-316)
// ---------------------------------------------------------------------------------------
__global__ void k_fleck(StepParams P, double dt, long long ncells) {
  for (long long c = blockIdx.x * (long long)blockDim.x + threadIdx.x; c < ncells;
       c += (long long)gridDim.x * blockDim.x) {
    const double rho = P.f.rho[c], sie = P.f.sie[c];
    const double temp = eos_temperature(P.cl, sie);
    const double cv = P.cl.cv;
    const double emis = emissivity(P.cl, rho, temp);
    P.f.fleck[c] = 1.0 / (1.0 + (4.0 * emis / (rho * cv * temp)) * dt);
  }
}

// per-cell lookup record of the IMC segment loop (CellProps)
__global__ void k_pack_props(StepParams P, long long ncells) {
  for (long long c = blockIdx.x * (long long)blockDim.x + threadIdx.x; c < ncells;
       c += (long long)gridDim.x * blockDim.x) {
    CellProps cp;
    cp.rho = P.f.rho[c];
    cp.temp = eos_temperature(P.cl, P.f.sie[c]);
    cp.ff = P.f.fleck[c];
    cp.ss = scattering_coeff(P.cl, cp.rho);
    P.f.props[c] = cp;
  }
}

// value of rho (field 0) or sie (field 1) at level-L global cell G with ghost-fill
// semantics: same level copy, coarser neighbour -> covering cell, finer neighbour -> mean
// of the covering fine cells, outflow -> clamp, periodic -> wrap.
__device__ double sample_cell_leaf(const StepParams &P, int field, int L, const long long Gin[3], bool &finer) {
  const DMesh &m = P.mesh;
  long long G[3] = {Gin[0], Gin[1], Gin[2]};
  for (int d = 0; d < 3; ++d) {
    const long long n = ((long long)m.nrb[d] * m.nxb[d]) << (d < m.ndim ? L : 0);
    if (G[d] < 0 || G[d] >= n) {
      const bool periodic = m.mesh_bc[2 * d + (G[d] < 0 ? 0 : 1)] == JB_BC_PERIODIC;
      if (periodic) {
        G[d] = ((G[d] % n) + n) % n;
      } else {
        G[d] = min(max(G[d], 0LL), n - 1);
      }
    }
  }
  int l[3] = {0, 0, 0};
  for (int d = 0; d < m.ndim; ++d) l[d] = (int)((G[d] << (m.max_level - L)) / m.nxb[d]);
  const int gid = lattice_at(m, l[0], l[1], l[2]);
  const DBlock &nb = m.blocks[gid];
  finer = false;
  const double *f;
  size_t blk;
  if (P.f.g_rho) {  // multi-rank: the mirror of all blocks, indexed by gid
    f = field == 0 ? P.f.g_rho : P.f.g_sie;
    blk = (size_t)gid;
  } else {
    if (nb.local < 0) return jbm::mk_nan();
    f = field == 0 ? P.f.rho : P.f.sie;
    blk = (size_t)nb.local;
  }
  if (nb.level <= L) {
    int idx[3] = {0, 0, 0};
    for (int d = 0; d < m.ndim; ++d) idx[d] = (int)((G[d] >> (L - nb.level)) - (long long)nb.lx[d] * m.nxb[d]);
    return f[blk * m.ncell + ((size_t)idx[2] * m.nxb[1] + idx[1]) * m.nxb[0] + idx[0]];
  }
  finer = true;
  return 0.0;
}
__device__ double sample_cell(const StepParams &P, int field, int L, const long long G[3]) {
  bool finer;
  const double v = sample_cell_leaf(P, field, L, G, finer);
  if (!finer) return v;
  // wrap / clamp G first exactly like the leaf lookup did, then average the children
  const DMesh &m = P.mesh;
  long long Gw[3] = {G[0], G[1], G[2]};
  for (int d = 0; d < 3; ++d) {
    const long long n = ((long long)m.nrb[d] * m.nxb[d]) << (d < m.ndim ? L : 0);
    if (Gw[d] < 0 || Gw[d] >= n) {
      const bool periodic = m.mesh_bc[2 * d + (Gw[d] < 0 ? 0 : 1)] == JB_BC_PERIODIC;
      Gw[d] = periodic ? ((Gw[d] % n) + n) % n : min(max(Gw[d], 0LL), n - 1);
    }
  }
  double sum = 0.0;
  int cnt = 0;
  const int nk = m.ndim > 2 ? 2 : 1, nj = m.ndim > 1 ? 2 : 1;
  for (int ck = 0; ck < nk; ++ck)
    for (int cj = 0; cj < nj; ++cj)
      for (int ci = 0; ci < 2; ++ci) {
        const long long Gc[3] = {2 * Gw[0] + ci, m.ndim > 1 ? 2 * Gw[1] + cj : Gw[1],
                                 m.ndim > 2 ? 2 * Gw[2] + ck : Gw[2]};
        bool f2;
        sum += sample_cell_leaf(P, field, L + 1, Gc, f2);
        cnt++;
      }
  return sum / cnt;
}

// level of the leaf across face (dir, side); own level at a non-periodic domain boundary
__device__ int neighbor_level(const DMesh &m, const DBlock &b, int dir, int side) {
  const int s = 1 << (m.max_level - b.level);
  int l[3];
  for (int d = 0; d < 3; ++d) l[d] = d < m.ndim ? b.lx[d] * s : 0;
  l[dir] = side == 0 ? b.lx[dir] * s - 1 : (b.lx[dir] + 1) * s;
  if (l[dir] < 0 || l[dir] >= m.nlat[dir]) {
    if (m.mesh_bc[2 * dir + side] != JB_BC_PERIODIC) return b.level;
    l[dir] = (l[dir] + m.nlat[dir]) % m.nlat[dir];
  }
  return m.blocks[lattice_at(m, l[0], l[1], l[2])].level;
}

// DDMC face leakage probabilities (jaybenne.cpp:319-489), one thread per face
__global__ void k_face_prob(StepParams P, int dir) {
  const DMesh &m = P.mesh;
  const long long nface = P.f.nface[dir];
  const long long total = nface * m.nlocal;
  int n[3] = {m.nxb[0], m.nxb[1], m.nxb[2]};
  const int nu = n[dir];
  n[dir] += 1;
  for (long long q = blockIdx.x * (long long)blockDim.x + threadIdx.x; q < total;
       q += (long long)gridDim.x * blockDim.x) {
    const int lb = (int)(q / nface);
    const long long r = q - (long long)lb * nface;
    const int i = (int)(r % n[0]), j = (int)((r / n[0]) % n[1]), k = (int)(r / ((long long)n[0] * n[1]));
    const DBlock &b = m.blocks[m.local_gid[lb]];
    const int L = b.level;
    const int lev_l = neighbor_level(m, b, dir, 0), lev_u = neighbor_level(m, b, dir, 1);
    const double dxd = b.dx[dir];
    const int f[3] = {i, j, k};
    const double dx_l = f[dir] == 0 ? ldexp(1.0, L - lev_l) * dxd : dxd;
    const double dx_u = f[dir] == nu ? ldexp(1.0, L - lev_u) * dxd : dxd;
    long long Gl[3], Gu[3];
    for (int d = 0; d < 3; ++d) {
      const long long base = d < m.ndim ? (long long)b.lx[d] * m.nxb[d] : 0;
      Gu[d] = base + f[d];
      Gl[d] = base + f[d] - (d == dir ? 1 : 0);
    }
    const double rho_l = sample_cell(P, 0, L, Gl), sie_l = sample_cell(P, 1, L, Gl);
    const double rho_u = sample_cell(P, 0, L, Gu), sie_u = sample_cell(P, 1, L, Gu);
    const double temp_l = eos_temperature(P.cl, sie_l);
    const double temp_u = eos_temperature(P.cl, sie_u);
    const double ss_l = scattering_coeff(P.cl, rho_l);
    const double aa_l = absorption_coeff(P.cl, rho_l, temp_l, 1.0);
    const double ss_u = scattering_coeff(P.cl, rho_u);
    const double aa_u = absorption_coeff(P.cl, rho_u, temp_u, 1.0);
    double tau_l = dx_l * (ss_l + aa_l);
    double tau_u = dx_u * (ss_u + aa_u);
    tau_l = tau_l > P.tau_ddmc ? tau_l : 2.0 * LAM_EXT;
    tau_u = tau_u > P.tau_ddmc ? tau_u : 2.0 * LAM_EXT;
    P.f.fp[dir][q] = 2.0 / (3.0 * (tau_l + tau_u));
  }
}

// ---------------------------------------------------------------------------------------
// SourcePhotons pass 1 (sourcing.cpp:73-119): per-cell source energy and stochastically
// rounded photon count. One thread per cell of the blocks [first_local, first_local+nb).
// ---------------------------------------------------------------------------------------
__global__ void k_source_count(StepParams P, int first_local, long long ncells, int source_type, double dt,
                               double npc, uint32_t call, int *cnt) {
  const DMesh &m = P.mesh;
  const double vv = P.cl.c_code, sb = P.cl.sb_code;
  for (long long q = blockIdx.x * (long long)blockDim.x + threadIdx.x; q < ncells;
       q += (long long)gridDim.x * blockDim.x) {
    const int lb = first_local + (int)(q / m.ncell);
    const int ci = (int)(q % m.ncell);
    const long long c = (long long)lb * m.ncell + ci;
    const DBlock &b = m.blocks[m.local_gid[lb]];
    const double dv = b.dx[0] * b.dx[1] * b.dx[2];
    Rng rng;
    rng.init((uint32_t)ci, (uint32_t)b.gid, call, 0u, P.key0, P.key1_cell);
    const double rho = P.f.rho[c], sie = P.f.sie[c];
    const double temp = eos_temperature(P.cl, sie);
    double erad;
    if (source_type == JB_SOURCE_THERMAL) {
      erad = (4.0 * sb / vv) * jbm::pow4(temp) * dv;
    } else {
      const double emis = emissivity(P.cl, rho, temp);
      erad = P.f.fleck[c] * emis * dv * dt;
    }
    double snpc = floor(npc);
    snpc += ((npc - snpc) > rng.drand()) ? 1.0 : 0.0;
    P.f.src_num[c] = snpc;
    P.f.src_ew[c] = erad / snpc;
    cnt[q] = (int)round(snpc);
  }
}

// ---------------------------------------------------------------------------------------
// exclusive prefix sum of int counts into 64-bit offsets: chunk sums -> scan of chunk
// sums (one CTA) -> per-chunk scan with the chunk offset. Replaces the reference's
// team-scan + device->host->device bounce (sourcing.cpp:110-131).
// ---------------------------------------------------------------------------------------
static constexpr int SCAN_CHUNK = 2048;  // elements per CTA (256 threads x 8)

__global__ void k_scan_chunk_sums(const int *in, long long n, long long *chunk_sum) {
  __shared__ long long sh[8];
  const long long base = (long long)blockIdx.x * SCAN_CHUNK;
  long long s = 0;
  for (int q = threadIdx.x; q < SCAN_CHUNK; q += blockDim.x) {
    const long long i = base + q;
    if (i < n) s += in[i];
  }
  unsigned long long v = warp_sum_u64((unsigned long long)s);
  if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = (long long)v;
  __syncthreads();
  if (threadIdx.x == 0) {
    long long t = 0;
    for (int w = 0; w < (int)(blockDim.x >> 5); ++w) t += sh[w];
    chunk_sum[blockIdx.x] = t;
  }
}
// single CTA: in-place exclusive scan of nchunk values; writes the grand total
__global__ void k_scan_chunk_offsets(long long *chunk_sum, long long nchunk, long long *total) {
  __shared__ long long sh[1024];
  __shared__ long long carry;
  if (threadIdx.x == 0) carry = 0;
  __syncthreads();
  for (long long base = 0; base < nchunk; base += blockDim.x) {
    const long long i = base + threadIdx.x;
    const long long v = i < nchunk ? chunk_sum[i] : 0;
    sh[threadIdx.x] = v;
    __syncthreads();
    for (int o = 1; o < (int)blockDim.x; o <<= 1) {
      const long long a = threadIdx.x >= (unsigned)o ? sh[threadIdx.x - o] : 0;
      __syncthreads();
      sh[threadIdx.x] += a;
      __syncthreads();
    }
    if (i < nchunk) chunk_sum[i] = carry + sh[threadIdx.x] - v;
    __syncthreads();
    if (threadIdx.x == 0) carry += sh[blockDim.x - 1];
    __syncthreads();
  }
  if (threadIdx.x == 0) *total = carry;
}
__global__ void k_scan_apply(const int *in, long long n, const long long *chunk_off, long long *out) {
  // 256 threads, 8 consecutive elements each
  __shared__ long long wsum[8];
  const long long base = (long long)blockIdx.x * SCAN_CHUNK + (long long)threadIdx.x * 8;
  int v[8];
  long long s = 0;
#pragma unroll
  for (int q = 0; q < 8; ++q) {
    v[q] = (base + q < n) ? in[base + q] : 0;
    s += v[q];
  }
  // inclusive scan of s across the CTA
  long long incl = s;
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const long long a = __shfl_up_sync(0xffffffffu, incl, o);
    if (lane >= o) incl += a;
  }
  if (lane == 31) wsum[wid] = incl;
  __syncthreads();
  long long woff = 0;
  for (int w = 0; w < wid; ++w) woff += wsum[w];
  long long run = chunk_off[blockIdx.x] + woff + incl - s;
#pragma unroll
  for (int q = 0; q < 8; ++q) {
    if (base + q < n) out[base + q] = run;
    run += v[q];
  }
}

long long scan_scratch_elems(long long n) { return (n + SCAN_CHUNK - 1) / SCAN_CHUNK + 1; }

void launch_exclusive_scan(const int *cnt, long long *prefix, long long n, long long *scratch, long long *total,
                           cudaStream_t st) {
  const long long nchunk = (n + SCAN_CHUNK - 1) / SCAN_CHUNK;
  k_scan_chunk_sums<<<(unsigned)nchunk, 256, 0, st>>>(cnt, n, scratch);
  k_scan_chunk_offsets<<<1, 1024, 0, st>>>(scratch, nchunk, total);
  k_scan_apply<<<(unsigned)nchunk, 256, 0, st>>>(cnt, n, scratch, prefix);
}

// ---------------------------------------------------------------------------------------
// SourcePhotons pass 2 (sourcing.cpp:141-205): one warp per cell, lanes stride over the
// cell's new photons so the SoA stores coalesce. Photon `np` (its index in the block's
// prefix order) keys its random stream, so the result does not depend on the schedule.
// ---------------------------------------------------------------------------------------
// one new photon: the stream's draws #0..#10 in the reference's order (sourcing.cpp:175-198): x y | z mu |
// phi xi0 | xi1 xi2 | xi3 xi4 | t — one Philox block per pair, straight-line (same values as Rng::drand)
__device__ __forceinline__ void source_one(const StepParams &P, const DBlock &b, const int lb, const int ci,
                                           const double xi, const double yi, const double zi, const double temp,
                                           const double ew, const int np, const long long slot, const int source_type,
                                           const double t_start, const double dt, const uint32_t call) {
  const double vv = P.cl.c_code, sb = P.cl.sb_code;
  const uint32_t id0 = (uint32_t)np, id1 = (uint32_t)b.gid;
  double d0, d1, d2, d3, d4, d5, d6, d7, d8, d9;
  philox_pair(0u, id0, id1, call, P.rk, d0, d1);
  philox_pair(1u, id0, id1, call, P.rk, d2, d3);
  philox_pair(2u, id0, id1, call, P.rk, d4, d5);
  philox_pair(3u, id0, id1, call, P.rk, d6, d7);
  philox_pair(4u, id0, id1, call, P.rk, d8, d9);
  const double x = xi + b.dx[0] * (d0 - 0.5);
  const double y = yi + b.dx[1] * (d1 - 0.5);
  const double z = zi + b.dx[2] * (d2 - 0.5);
  const double theta = jbm::acos(2.0 * d3 - 1.0);
  const double phi = 2.0 * JB_PI * d4;
  double st, ct, sp, cp;
  jbm::sincos(theta, &st, &ct);
  jbm::sincos(phi, &sp, &cp);
  const double energy = planck_energy_draws(P.cl, d5, d6, d7, d8, d9, sb, temp);
  double tp = 0.;
  uint32_t ndraws = 10u;
  if (source_type == JB_SOURCE_EMISSION) {
    double d10, d11;
    philox_pair(5u, id0, id1, call, P.rk, d10, d11);
    tp = t_start + d10 * dt;
    ndraws = 11u;
  }
  P.s.x[slot] = x;
  P.s.y[slot] = y;
  P.s.z[slot] = z;
  P.s.vx[slot] = vv * st * cp;
  P.s.vy[slot] = vv * st * sp;
  P.s.vz[slot] = vv * ct;
  P.s.e[slot] = energy;
  P.s.w[slot] = ew;
  P.s.t[slot] = tp;
  P.s.loc[slot] = make_int2(lb, ci);
  P.s.ida[slot] = make_uint2((uint32_t)np, (uint32_t)b.gid);
  P.s.idb[slot] = make_uint2(call, ndraws);
}

// Sparse launches (an emission pass that gives most cells no photon or one): one THREAD per new photon, which
// finds its cell in the prefix array by bisection; the per-cell energy_delta reset / debit (sourcing.cpp:165-166)
// is a thread-per-cell pass of its own. The warp-per-cell kernel below spent a warp's trip on every cell — 16.7 M
// trips for 2 M photons in the 256^3 benchmark's emission pass (profiles/r02_ncu_source_emit_emission_cells128.md).
__global__ void __launch_bounds__(256) k_source_debit(StepParams P, int first_local, long long ncells, int source_type) {
  const long long base = (long long)first_local * P.mesh.ncell;
  for (long long q = blockIdx.x * (long long)blockDim.x + threadIdx.x; q < ncells; q += (long long)gridDim.x * blockDim.x) {
    const long long c = base + q;
    double dejbn = 0.0;
    if (source_type == JB_SOURCE_EMISSION) {
      // debited serially so the sum rounds exactly as the reference's per-cell loop does
      const int npart = (int)round(P.f.src_num[c]);
      const double ew = P.f.src_ew[c];
      for (int np = 0; np < npart; ++np) dejbn -= ew;
    }
    P.f.edelta[c] = dejbn;
  }
}
__global__ void __launch_bounds__(256, 4) k_source_emit_sparse(StepParams P, int first_local, long long ncells, int source_type,
                                                               double t_start, double dt, uint32_t call,
                                                               const long long *__restrict__ prefix, long long slot0,
                                                               long long total) {
  const DMesh &m = P.mesh;
  unsigned long long my_n = 0;
  double my_e = 0.0;
  for (long long p = blockIdx.x * (long long)blockDim.x + threadIdx.x; p < total; p += (long long)gridDim.x * blockDim.x) {
    // the last cell q with prefix[q] <= p (empty cells share their successor's offset, so this is p's cell)
    long long lo = 0, hi = ncells - 1;
    while (lo < hi) {
      const long long mid = (lo + hi + 1) >> 1;
      if (__ldg(&prefix[mid]) <= p) lo = mid; else hi = mid - 1;
    }
    const long long q = lo;
    const int lbq = (int)(q / m.ncell);
    const int lb = first_local + lbq;
    const int ci = (int)(q - (long long)lbq * m.ncell);
    const long long c = (long long)lb * m.ncell + ci;
    const long long pglobal = __ldg(&prefix[q]);
    const int np = (int)(p - __ldg(&prefix[(long long)lbq * m.ncell]));  // index in the block's prefix order
    const DBlock &b = m.blocks[m.local_gid[lb]];
    const int i = ci % m.nxb[0], j = (ci / m.nxb[0]) % m.nxb[1], k = ci / (m.nxb[0] * m.nxb[1]);
    const double xi = cell_center(b, m.ng[0], 0, i), yi = cell_center(b, m.ng[1], 1, j), zi = cell_center(b, m.ng[2], 2, k);
    const double temp = eos_temperature(P.cl, P.f.sie[c]);
    const double ew = P.f.src_ew[c];
    (void)pglobal;
    source_one(P, b, lb, ci, xi, yi, zi, temp, ew, np, slot0 + p, source_type, t_start, dt, call);
    my_n += 1;
    my_e += ew;
  }
  warp_add_u64(&P.stats->n_sourced, my_n);
  warp_add_energy(P, EB_SOURCED, my_e);
}

__global__ void __launch_bounds__(256, 4) k_source_emit(StepParams P, int first_local, long long ncells, int source_type, double t_start,
                              double dt, uint32_t call, const long long *prefix, long long slot0) {
  const DMesh &m = P.mesh;
  const int lane = threadIdx.x & 31;
  // 32-bit cell arithmetic: a rank holds < 2^31 cells (checked by the launcher), and the 64-bit divide /
  // modulo this loop used to do per cell were ~1/3 of the kernel's instructions (profiles/r02_ncu_source_emit*.md)
  const unsigned warp0 = (unsigned)((blockIdx.x * (long long)blockDim.x + threadIdx.x) >> 5);
  const unsigned nwarps = (unsigned)(((long long)gridDim.x * blockDim.x) >> 5);
  const unsigned ncell_u = (unsigned)m.ncell, ncells_u = (unsigned)ncells;
  const bool pow2 = (m.nxb[0] & (m.nxb[0] - 1)) == 0 && (m.nxb[1] & (m.nxb[1] - 1)) == 0;
  const int sh1 = pow2 ? __ffs(m.nxb[0]) - 1 : -1, sh2 = pow2 ? __ffs(m.nxb[1]) - 1 : 0;
  unsigned long long my_n = 0;
  double my_e = 0.0;
  for (unsigned q = warp0; q < ncells_u; q += nwarps) {
    const int lbq = (int)(q / ncell_u);
    const int lb = first_local + lbq;
    const int ci = (int)(q - (unsigned)lbq * ncell_u);
    const long long c = (long long)lb * m.ncell + ci;
    const int npart = (int)round(P.f.src_num[c]);
    const long long pglobal = prefix[q];
    const long long pblock = prefix[(unsigned)lbq * ncell_u];
    const int pstart = (int)(pglobal - pblock);
    const double ew = P.f.src_ew[c];
    if (lane == 0) {
      // energy_delta is reset here and only here (sourcing.cpp:165-166), then debited
      // serially so the sum rounds exactly as the reference's per-cell loop does
      double dejbn = 0.0;
      if (source_type == JB_SOURCE_EMISSION)
        for (int np = 0; np < npart; ++np) dejbn -= ew;
      P.f.edelta[c] = dejbn;
    }
    if (npart == 0) continue;
    const DBlock &b = m.blocks[m.local_gid[lb]];
    int i, j, k;
    if (sh1 >= 0) {  // power-of-two block: shifts instead of three integer divisions per cell
      i = ci & (m.nxb[0] - 1);
      j = (ci >> sh1) & (m.nxb[1] - 1);
      k = ci >> (sh1 + sh2);
    } else {
      i = ci % m.nxb[0];
      j = (ci / m.nxb[0]) % m.nxb[1];
      k = ci / (m.nxb[0] * m.nxb[1]);
    }
    const double xi = cell_center(b, m.ng[0], 0, i), yi = cell_center(b, m.ng[1], 1, j),
                 zi = cell_center(b, m.ng[2], 2, k);
    const double temp = eos_temperature(P.cl, P.f.sie[c]);
    for (int o = lane; o < npart; o += 32) {
      const int np = pstart + o;
      const long long slot = slot0 + pglobal + o;
      source_one(P, b, lb, ci, xi, yi, zi, temp, ew, np, slot, source_type, t_start, dt, call);
      my_n += 1;
      my_e += ew;
    }
  }
  warp_add_u64(&P.stats->n_sourced, my_n);
  warp_add_energy(P, EB_SOURCED, my_e);
}

// ---------------------------------------------------------------------------------------
// transport
// ---------------------------------------------------------------------------------------
__device__ __forceinline__ uint4 load_id(const Swarm &s, long long idx) {
  const uint2 a = s.ida[idx], b = s.idb[idx];
  return make_uint4(a.x, a.y, b.x, b.y);
}
// per-thread view of the meshblock a photon is in
struct BlkReg {
  double xmin[3], dx[3], xmin0[3], xmax[3];
  double dx_push;
  int gid, lb;
};
__device__ __forceinline__ void load_block(const DMesh &m, int gid, BlkReg &B) {
  const DBlock &b = m.blocks[gid];
#pragma unroll
  for (int d = 0; d < 3; ++d) {
    B.xmin[d] = b.xmin[d];
    B.xmax[d] = b.xmax[d];
    B.dx[d] = b.dx[d];
    B.xmin0[d] = b.xmin0[d];
  }
  B.dx_push = jbm::min2(B.dx[0], jbm::min2(B.dx[1], B.dx[2]));
  B.gid = gid;
  B.lb = b.local;
}
// SwarmDeviceContext::Xtoijk as interior indices; collapsed dimensions give 0
template <int NDIM>
__device__ __forceinline__ void xtoijk(const BlkReg &B, const Photon &p, int &i, int &j, int &k) {
  // cell widths are normal numbers and the offsets finite: the range-test-free quotient (jb_math.h)
  i = (int)floor(jbm::div_normal(p.x - B.xmin[0], B.dx[0]));
  j = (NDIM > 1) ? (int)floor(jbm::div_normal(p.y - B.xmin[1], B.dx[1])) : 0;
  k = (NDIM > 2) ? (int)floor(jbm::div_normal(p.z - B.xmin[2], B.dx[2])) : 0;
}
__device__ __forceinline__ bool on_block(const DMesh &m, int i, int j, int k) {
  return i >= 0 && i < m.nxb[0] && j >= 0 && j < m.nxb[1] && k >= 0 && k < m.nxb[2];
}
__device__ __forceinline__ double xc(const BlkReg &B, const int ng, const int d, const int i) {
  return B.xmin0[d] + ((i + ng) + 0.5) * B.dx[d];
}

// Swarm boundary conditions at the domain faces, applied when a photon leaves its block:
// periodic wrap, PhotonReflectBC (boundaries.hpp:24-84) or outflow removal (false).
template <int NDIM>
__device__ __forceinline__ bool apply_swarm_bc(const DMesh &m, Photon &p) {
  double *pos[3] = {&p.x, &p.y, &p.z};
  double *vel[3] = {&p.vx, &p.vy, &p.vz};
#pragma unroll
  for (int d = 0; d < NDIM; ++d) {
    const double lo = m.xmin[d], hi = m.xmax[d];
    if (*pos[d] < lo) {
      const int bc = m.swarm_bc[2 * d];
      if (bc == JB_BC_PERIODIC) {
        *pos[d] = *pos[d] + (hi - lo);
      } else if (bc == JB_BC_REFLECT) {
        *pos[d] = lo + (lo - *pos[d]);
        *vel[d] = -*vel[d];
      } else {
        return false;
      }
    } else if (*pos[d] >= hi) {
      const int bc = m.swarm_bc[2 * d + 1];
      if (bc == JB_BC_PERIODIC) {
        *pos[d] = *pos[d] - (hi - lo);
      } else if (bc == JB_BC_REFLECT) {
        if (*pos[d] > hi) {
          *pos[d] = hi - (*pos[d] - hi);
          *vel[d] = -*vel[d];
        }
      } else {
        return false;
      }
    }
  }
  return true;
}

__device__ __forceinline__ void log_event(const StepParams &P, const uint4 &id, uint32_t draw, int kind, int gid,
                                          const Photon &p) {
  const unsigned long long slot = atomicAdd(P.log.count, 1ull);
  if ((long long)slot >= P.log.cap) return;
  jb_event e;
  e.id_idx = id.x; e.id_gid = id.y; e.id_call = id.z;
  e.draw = draw;
  e.kind = kind;
  e.block = gid;
  e.i = p.ip; e.j = p.jp; e.k = p.kp;
  e.reserved = 0;
  e.t = p.t; e.x = p.x; e.y = p.y; e.z = p.z;
  e.v0 = p.vx; e.v1 = p.vy; e.v2 = p.vz;
  P.log.ev[slot] = e;
}

// SampleDDMCBlockFace for one photon (sample_ddmc_bface.cpp:81-427): a photon that a DDMC leak carried out
// of a COARSER block arrives with zero velocity, 2 eps_ddmc fine cell widths inside this block and at the
// coarse cell's centre in the transverse directions, i.e. on an edge (2D) or a vertex (3D) between fine
// cells. It gets an inward cosine-law direction and one of the 2 / 4 fine faces that tile the coarse face
// it came through, drawn in proportion to their leakage probabilities, and is spread over that face.
// One evaluation parametrised by the face's axis replaces the reference's per-face copies: `a` is the
// normal axis, (t1, t2) the transverse axes in increasing order — the argument orders of the reference's
// SampleFace2D / SampleFace3D call sites — and draws, operands and operation order are the reference's.
// The reference reads the probabilities through ghost-inclusive face indices; on a 2:1 mesh the pair of
// fine cells under a coarse cell never straddles a block edge, so an index outside the block cannot occur:
// it is counted (DStats::n_bface_ghost, the step fails) instead of being clamped silently.
template <int NDIM>
__device__ void bface_sample(const StepParams &P, const BlkReg &B, Photon &p, Rng &rng) {
  if (NDIM < 2) return;
  const DMesh &m = P.mesh;
  const double vv = P.cl.c_code;
  const double eps = EPS10;
  if (!(p.vx * p.vx + p.vy * p.vy + p.vz * p.vz < eps * vv * vv)) return;
  xtoijk<NDIM>(B, p, p.ip, p.jp, p.kp);
  double pos[3] = {p.x, p.y, p.z};
  int ijk[3] = {p.ip, p.jp, p.kp};
  // which block face (x before y before z, lower before upper, as the reference's if / else-if chain)
  int a = -1;
  bool at_min = false;
#pragma unroll
  for (int d = NDIM - 1; d >= 0; --d) {
    const bool lo = fuzzy_eq(pos[d], B.xmin[d] + 2.0 * EPS_DDMC * B.dx[d], B.dx[d], eps);
    const bool hi = fuzzy_eq(pos[d], B.xmax[d] - 2.0 * EPS_DDMC * B.dx[d], B.dx[d], eps);
    if (lo || hi) {
      a = d;
      at_min = lo;
    }
  }
  if (a < 0) return;
  {
    double v1, v2, v3;
    face_iso_dir(at_min ? vv : -vv, rng, v1, v2, v3);
    unpermute_dir(a, v1, v2, v3, p.vx, p.vy, p.vz);
  }
  const int t1 = a == 0 ? 1 : 0, t2 = a == 2 ? 1 : 2;
  // is the photon on the lower (`up`: it then belongs to the upper candidate) or the upper edge of its cell?
  bool on_edge[3] = {false, false, false};
  int i_l[3] = {ijk[0], ijk[1], ijk[2]};
#pragma unroll
  for (int q = 0; q < 2; ++q) {
    const int t = q == 0 ? t1 : t2;
    if (t >= NDIM) continue;
    const double lo_t = xc(B, m.ng[t], t, ijk[t]) - 0.5 * B.dx[t];
    const bool edge_u = fuzzy_eq(pos[t], lo_t, B.dx[t], eps);
    const bool edge_l = fuzzy_eq(pos[t], lo_t + B.dx[t], B.dx[t], eps);
    on_edge[t] = edge_u || edge_l;
    i_l[t] = edge_u ? ijk[t] - 1 : ijk[t];
  }
  const int nx[3] = {m.nxb[0], m.nxb[1], m.nxb[2]};
  const double *fpa = P.f.fp[a] + (size_t)B.lb * P.f.nface[a];
  int f[3] = {ijk[0], ijk[1], ijk[2]};
  f[a] = at_min ? ijk[a] : ijk[a] + 1;
  auto prob = [&](int o1, int o2) {  // face (i_l[t1] + o1, i_l[t2] + o2) of the block face
    int g[3] = {f[0], f[1], f[2]};
    g[t1] = i_l[t1] + o1;
    if (NDIM > 2) g[t2] = i_l[t2] + o2;
    bool ghost = false;
#pragma unroll
    for (int d = 0; d < 3; ++d) {
      const int hi = nx[d] - 1 + (d == a ? 1 : 0);
      if (g[d] < 0 || g[d] > hi) {
        ghost = true;
        g[d] = min(max(g[d], 0), hi);
      }
    }
    if (ghost) atomicAdd(&P.stats->n_bface_ghost, 1ull);
    return fpa[((size_t)g[2] * (nx[1] + (a == 1 ? 1 : 0)) + g[1]) * (nx[0] + (a == 0 ? 1 : 0)) + g[0]];
  };
  if (NDIM == 2) {
    if (!on_edge[t1]) return;  // SampleFace2D (sample_ddmc_bface.cpp:24-39)
    const double P_l = prob(0, 0), P_u = prob(1, 0);
    const double xi = (P_l + P_u) * rng.drand();
    const int up = xi < P_l ? 0 : 1;
    const double r = B.dx[t1] * rng.drand();
    pos[t1] = up ? pos[t1] + r : pos[t1] - r;
    ijk[t1] = i_l[t1] + up;
  } else {
    if (!(on_edge[t1] && on_edge[t2])) return;  // SampleFace3D (sample_ddmc_bface.cpp:41-78)
    const double P_ll = prob(0, 0), P_lu = prob(1, 0), P_ul = prob(0, 1), P_uu = prob(1, 1);
    const double xi = (P_ll + P_lu + P_ul + P_uu) * rng.drand();
    const int sel = xi < P_ll ? 0 : (xi < P_ll + P_lu ? 1 : (xi < P_ll + P_lu + P_ul ? 2 : 3));
    const int up1 = sel & 1, up2 = sel >> 1;
    const double r1 = B.dx[t1] * rng.drand();
    pos[t1] = up1 ? pos[t1] + r1 : pos[t1] - r1;
    ijk[t1] = i_l[t1] + up1;
    const double r2 = B.dx[t2] * rng.drand();
    pos[t2] = up2 ? pos[t2] + r2 : pos[t2] - r2;
    ijk[t2] = i_l[t2] + up2;
  }
  atomicAdd(&P.stats->n_bface, 1ull);
  p.x = pos[0]; p.y = pos[1]; p.z = pos[2];
  p.ip = ijk[0]; p.jp = ijk[1]; p.kp = ijk[2];
}

// The history kernel. One thread owns one photon from its current state to the first of:
// census (t >= t_end), absorption, escape through an outflow face, or hand-over to a
// block owned by another rank. Crossing into another block of THIS rank applies the swarm
// boundary conditions and continues in place — the reference's break -> Send -> Receive
// -> next transport iteration round trip (transport.cpp:149-155, jaybenne.cpp:113-131)
// collapses to a table lookup because the photon's stream travels with it.
template <int NDIM, bool DDMC, bool LOG>
__global__ void __launch_bounds__(128) k_transport(StepParams P, long long first, long long n, double t_start,
                                                   double dt) {
  const DMesh &m = P.mesh;
  constexpr bool multi_d = NDIM >= 2, three_d = NDIM == 3;
  const double t_end = t_start + dt;
  const double vv = P.cl.c_code;
  const int nx1 = m.nxb[0], nx2 = m.nxb[1];
  unsigned long long c_seg = 0, c_scat = 0, c_ddmc = 0, c_cross = 0, c_abs = 0, c_esc = 0, c_sent = 0, c_err = 0,
                     c_ovf = 0;
  double e_abs = 0.0, e_esc = 0.0;
  for (long long q = blockIdx.x * (long long)blockDim.x + threadIdx.x; q < n;
       q += (long long)gridDim.x * blockDim.x) {
    const long long idx = first + q;
    int lb = P.s.loc[idx].x;
    if (lb < 0) continue;
    Photon p;
    p.t = P.s.t[idx];
    if (!(p.t < t_end)) continue;  // census photon: nothing to do (reference re-derives ijk only)
    p.x = P.s.x[idx]; p.y = P.s.y[idx]; p.z = P.s.z[idx];
    p.vx = P.s.vx[idx]; p.vy = P.s.vy[idx]; p.vz = P.s.vz[idx];
    const uint4 id = load_id(P.s, idx);
    const double ww = P.s.w[idx], ee = P.s.e[idx];
    Rng rng;
    rng.init(id.x, id.y, id.z, id.w, P.key0, P.key1_particle);
    BlkReg B;
    load_block(m, m.local_gid[lb], B);
    xtoijk<NDIM>(B, p, p.ip, p.jp, p.kp);
    bool gone = false;
    while (p.t < t_end) {
      const uint32_t draw0 = rng.n;
      c_seg++;
      CellEnv c;
      {
        const double xcen = xc(B, m.ng[0], 0, p.ip), ycen = xc(B, m.ng[1], 1, p.jp), zcen = xc(B, m.ng[2], 2, p.kp);
        c.xl = xcen - 0.5 * B.dx[0]; c.xu = xcen + 0.5 * B.dx[0];
        c.yl = ycen - 0.5 * B.dx[1]; c.yu = ycen + 0.5 * B.dx[1];
        c.zl = zcen - 0.5 * B.dx[2]; c.zu = zcen + 0.5 * B.dx[2];
      }
      const long long ci = (long long)B.lb * m.ncell + ((long long)p.kp * nx2 + p.jp) * nx1 + p.ip;
      const double rho = __ldg(&P.f.rho[ci]), sie = __ldg(&P.f.sie[ci]);
      const double temp = eos_temperature(P.cl, sie);
      c.ff = __ldg(&P.f.fleck[ci]);
      c.ss = scattering_coeff(P.cl, rho);
      c.aa = absorption_coeff(P.cl, rho, temp, ee);
      bool is_absorbed = false, is_scattered = false, is_rejected = false;
      const bool is_ddmc_step = DDMC && (B.dx_push * (c.ss + c.aa) > P.tau_ddmc);
      if (DDMC && is_ddmc_step) {
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
        ddmc_albedo<NDIM>(rng, c, vv, p, is_rejected);
        if (!is_rejected) ddmc_step<NDIM>(rng, t_end, c, vv, Pf, p, is_absorbed);
      } else {
        imc_step<NDIM>(rng, t_end, c, vv, B.dx_push, p, is_absorbed, is_scattered);
      }
      xtoijk<NDIM>(B, p, p.ip, p.jp, p.kp);
      int kind = (is_ddmc_step ? JB_EV_DDMC : 0) | (is_rejected ? JB_EV_ALBEDO_REJECT : 0);
      if (!on_block(m, p.ip, p.jp, p.kp)) {
        if (DDMC) {
          // zero velocity tags a DDMC photon for block-face resampling (transport_ddmc.cpp:203-211)
          const double vmask = !(is_ddmc_step && multi_d && !is_rejected) ? 1.0 : 0.0;
          p.vx *= vmask;
          p.vy *= vmask;
          p.vz *= vmask;
        }
        if (LOG) {
          uint4 idn = id;
          log_event(P, idn, draw0, kind | JB_EV_LEFT_BLOCK | (!(p.t < t_end) ? JB_EV_CENSUS : 0), B.gid, p);
        }
        // ---- MeshSend / MeshReceive, on device ----
        if (!apply_swarm_bc<NDIM>(m, p)) {
          c_esc++;
          e_esc += ww;
          gone = true;
          break;
        }
        const double pos[3] = {p.x, p.y, p.z};
        const int gid = find_block(m, pos);
        if (gid < 0) {
          c_err++;
          gone = true;
          break;
        }
        c_cross++;
        const DBlock &nb = m.blocks[gid];
        if (nb.rank != m.my_rank) {
          const unsigned long long slot = atomicAdd(&P.out.count[nb.rank], 1ull);
          if ((long long)slot < P.out.cap) {
            ExRecord r;
            r.x = p.x; r.y = p.y; r.z = p.z; r.t = p.t; r.w = ww; r.e = ee;
            r.vx = p.vx; r.vy = p.vy; r.vz = p.vz;
            r.rng[0] = id.x; r.rng[1] = id.y; r.rng[2] = id.z; r.rng[3] = rng.n;
            r.dest_gid = gid;
            r.pad = 0;
            P.out.rec[(size_t)nb.rank * P.out.cap + slot] = r;
            c_sent++;
          } else {
            c_ovf++;
          }
          gone = true;
          break;
        }
        load_block(m, gid, B);
        lb = B.lb;
        xtoijk<NDIM>(B, p, p.ip, p.jp, p.kp);
        if (DDMC) bface_sample<NDIM>(P, B, p, rng);
        continue;
      }
      if (is_absorbed) {
        atomicAdd(&P.f.edelta[(long long)B.lb * m.ncell + ((long long)p.kp * nx2 + p.jp) * nx1 + p.ip], ww);
        c_abs++;
        e_abs += ww;
        if (LOG) log_event(P, id, draw0, kind | JB_EV_ABSORB, B.gid, p);
        gone = true;
        break;
      }
      if (is_scattered) {
        scatter_iso(rng, vv, p.vx, p.vy, p.vz);
        c_scat++;
        kind |= JB_EV_SCATTER;
      }
      if (LOG) log_event(P, id, draw0, kind | (!(p.t < t_end) ? JB_EV_CENSUS : 0), B.gid, p);
    }
    if (gone) {
      P.s.loc[idx].x = -1;
      continue;
    }
    P.s.x[idx] = p.x; P.s.y[idx] = p.y; P.s.z[idx] = p.z; P.s.t[idx] = p.t;
    P.s.vx[idx] = p.vx; P.s.vy[idx] = p.vy; P.s.vz[idx] = p.vz;
    P.s.loc[idx] = make_int2(lb, (p.kp * nx2 + p.jp) * nx1 + p.ip);
    P.s.idb[idx] = make_uint2(id.z, rng.n);
  }
  DStats *S = P.stats;
  warp_add_u64(&S->n_segments, c_seg);
  warp_add_u64(&S->n_scatters, c_scat);
  warp_add_u64(&S->n_crossings, c_cross);
  warp_add_u64(&S->n_absorbed, c_abs);
  warp_add_u64(&S->n_dead, c_abs + c_esc + c_sent + c_err + c_ovf);
  warp_add_energy(P, EB_ABSORBED, e_abs);
  if (DDMC) warp_add_u64(&S->n_ddmc, c_ddmc);
  // rare channels: skip the shuffles unless some lane has something
  if (__any_sync(0xffffffffu, (c_esc | c_sent | c_err | c_ovf) != 0)) {
    warp_add_u64(&S->n_escaped, c_esc);
    warp_add_u64(&S->n_sent, c_sent);
    warp_add_u64(&S->n_error, c_err);
    warp_add_u64(&S->n_overflow, c_ovf);
    warp_add_energy(P, EB_ESCAPED, e_esc);
  }
}

template <int NDIM, bool DDMC>
static void launch_transport_nd(const StepParams &P, long long first, long long n, double t_start, double dt,
                                cudaStream_t st) {
  const int threads = 128;
  long long blocks = (n + threads - 1) / threads;
  if (blocks < 1) blocks = 1;
  if (blocks > 0x7fffffffLL) blocks = 0x7fffffffLL;
  if (P.log.cap > 0)
    k_transport<NDIM, DDMC, true><<<(unsigned)blocks, threads, 0, st>>>(P, first, n, t_start, dt);
  else
    k_transport<NDIM, DDMC, false><<<(unsigned)blocks, threads, 0, st>>>(P, first, n, t_start, dt);
}

#include "jb_transport_imc.cuh"
#include "jb_transport_pool.cuh"
#include "jb_transport_ddmc.cuh"
#include "jb_transport_hpool.cuh"

void launch_transport(const StepParams &P, long long first, long long n, double t_start, double dt,
                      cudaStream_t st) {
  if (n <= 0) return;
  const bool dd = P.use_ddmc != 0;
  if (!dd && P.log.cap <= 0 && g_tuning.variant == 2) {
    bool ok;
    switch (P.mesh.ndim) {
      case 1: ok = launch_transport_pool<1>(P, first, n, t_start, dt, st); break;
      case 2: ok = launch_transport_pool<2>(P, first, n, t_start, dt, st); break;
      default: ok = launch_transport_pool<3>(P, first, n, t_start, dt, st); break;
    }
    if (ok) return;
  }
  // 32-bit slot offsets inside the persistent kernels: longer ranges take the thread-per-photon kernel
  if (!dd && P.log.cap <= 0 && g_tuning.variant >= 1 && n < 0xffffffffLL) {
    switch (P.mesh.ndim) {
      case 1: launch_transport_imc<1>(P, first, n, t_start, dt, st); break;
      case 2: launch_transport_imc<2>(P, first, n, t_start, dt, st); break;
      default: launch_transport_imc<3>(P, first, n, t_start, dt, st); break;
    }
    return;
  }
  if (dd && g_tuning.variant == 2) {  // hybrid IMC / DDMC on warp-private photon pools
    bool ok;
    switch (P.mesh.ndim) {
      case 1: ok = launch_transport_hpool<1>(P, first, n, t_start, dt, st); break;
      case 2: ok = launch_transport_hpool<2>(P, first, n, t_start, dt, st); break;
      default: ok = launch_transport_hpool<3>(P, first, n, t_start, dt, st); break;
    }
    if (ok) return;
  }
  if (dd && g_tuning.variant >= 1) {  // hybrid IMC / DDMC: persistent warps with lane refill
    switch (P.mesh.ndim) {
      case 1: launch_transport_hybrid<1>(P, first, n, t_start, dt, st); break;
      case 2: launch_transport_hybrid<2>(P, first, n, t_start, dt, st); break;
      default: launch_transport_hybrid<3>(P, first, n, t_start, dt, st); break;
    }
    return;
  }
  switch (P.mesh.ndim) {
    case 1: dd ? launch_transport_nd<1, true>(P, first, n, t_start, dt, st) : launch_transport_nd<1, false>(P, first, n, t_start, dt, st); break;
    case 2: dd ? launch_transport_nd<2, true>(P, first, n, t_start, dt, st) : launch_transport_nd<2, false>(P, first, n, t_start, dt, st); break;
    default: dd ? launch_transport_nd<3, true>(P, first, n, t_start, dt, st) : launch_transport_nd<3, false>(P, first, n, t_start, dt, st); break;
  }
}

// ---------------------------------------------------------------------------------------
// SampleDDMCBlockFace over a slot range (photons that arrived from another rank)
// ---------------------------------------------------------------------------------------
template <int NDIM>
__global__ void k_bface(StepParams P, long long first, long long n) {
  const DMesh &m = P.mesh;
  if (P.range) {
    first = P.range[0];
    n = P.range[1];
  }
  for (long long q = blockIdx.x * (long long)blockDim.x + threadIdx.x; q < n;
       q += (long long)gridDim.x * blockDim.x) {
    const long long idx = first + q;
    const int lb = P.s.loc[idx].x;
    if (lb < 0) continue;
    Photon p;
    p.x = P.s.x[idx]; p.y = P.s.y[idx]; p.z = P.s.z[idx]; p.t = P.s.t[idx];
    p.vx = P.s.vx[idx]; p.vy = P.s.vy[idx]; p.vz = P.s.vz[idx];
    const double vv = P.cl.c_code;
    if (!(p.vx * p.vx + p.vy * p.vy + p.vz * p.vz < EPS10 * vv * vv)) continue;
    const uint4 id = load_id(P.s, idx);
    Rng rng;
    rng.init(id.x, id.y, id.z, id.w, P.key0, P.key1_particle);
    BlkReg B;
    load_block(m, m.local_gid[lb], B);
    bface_sample<NDIM>(P, B, p, rng);
    P.s.x[idx] = p.x; P.s.y[idx] = p.y; P.s.z[idx] = p.z;
    P.s.vx[idx] = p.vx; P.s.vy[idx] = p.vy; P.s.vz[idx] = p.vz;
    P.s.loc[idx].y = (p.kp * m.nxb[1] + p.jp) * m.nxb[0] + p.ip;
    P.s.idb[idx] = make_uint2(id.z, rng.n);
  }
}
static inline unsigned grid_for(long long n, int threads) {
  long long b = (n + threads - 1) / threads;
  if (b < 1) b = 1;
  if (b > 148LL * 64) b = 148LL * 64;  // grid-stride beyond that: 148 SMs x 64 CTAs
  return (unsigned)b;
}
void launch_bface_range(const StepParams &P, long long first, long long n, cudaStream_t st) {
  if (n <= 0 || P.mesh.ndim < 2) return;
  if (P.mesh.ndim == 2)
    k_bface<2><<<grid_for(n, 128), 128, 0, st>>>(P, first, n);
  else
    k_bface<3><<<grid_for(n, 128), 128, 0, st>>>(P, first, n);
}
void launch_bface(const StepParams &P, long long n, cudaStream_t st) { launch_bface_range(P, 0, n, st); }

// ---------------------------------------------------------------------------------------
// cell-field kernels
// ---------------------------------------------------------------------------------------
void launch_fleck(const StepParams &P, double dt, cudaStream_t st) {
  const long long ncells = (long long)P.mesh.nlocal * P.mesh.ncell;
  if (ncells <= 0) return;
  k_fleck<<<grid_for(ncells, 256), 256, 0, st>>>(P, dt, ncells);
}
void launch_pack_props(const StepParams &P, cudaStream_t st) {
  const long long ncells = (long long)P.mesh.nlocal * P.mesh.ncell;
  if (ncells > 0) k_pack_props<<<grid_for(ncells, 256), 256, 0, st>>>(P, ncells);
}
void launch_face_prob(const StepParams &P, cudaStream_t st) {
  for (int d = 0; d < P.mesh.ndim; ++d) {
    const long long total = P.f.nface[d] * P.mesh.nlocal;
    if (total > 0) k_face_prob<<<grid_for(total, 128), 128, 0, st>>>(P, d);
  }
}
void launch_source_count(const StepParams &P, int first_local, int nlocal_call, int source_type, double dt,
                         double npc, uint32_t call, int *cnt, cudaStream_t st) {
  const long long ncells = (long long)nlocal_call * P.mesh.ncell;
  if (ncells <= 0) return;
  k_source_count<<<grid_for(ncells, 256), 256, 0, st>>>(P, first_local, ncells, source_type, dt, npc, call, cnt);
}
void launch_source_emit(const StepParams &P, int first_local, int nlocal_call, int source_type, double t_start,
                        double dt, uint32_t call, const long long *prefix, long long slot0, long long total,
                        cudaStream_t st) {
  const long long ncells = (long long)nlocal_call * P.mesh.ncell;
  if (ncells <= 0) return;
  // cumulative j^-4 table of the Planck sampler: once per device, the reference's running sum
  static bool table_ready[64] = {false};
  int dev = 0;
  cudaGetDevice(&dev);
  if (dev >= 0 && dev < 64 && !table_ready[dev]) {
    double cum[100];
    volatile double lhs = 0.0;  // volatile: one rounding per operation, no re-association
    cum[0] = 0.0;
    for (int l = 1; l < 100; ++l) {
      const double j = (double)l;
      lhs = lhs + 1.0 / ((j * j) * (j * j));
      cum[l] = lhs;
    }
    cudaMemcpyToSymbolAsync(d_planck_cum, cum, sizeof(cum), 0, cudaMemcpyHostToDevice, st);
    cudaStreamSynchronize(st);
    table_ready[dev] = true;
  }
  if (total < 4 * ncells) {
    // fewer than 4 photons per cell on average: thread per photon + thread per cell (k_source_emit_sparse)
    k_source_debit<<<grid_for(ncells, 256), 256, 0, st>>>(P, first_local, ncells, source_type);
    if (total > 0)
      k_source_emit_sparse<<<grid_for(total, 256), 256, 0, st>>>(P, first_local, ncells, source_type, t_start, dt, call,
                                                                  prefix, slot0, total);
    return;
  }
  // one warp per cell, 8 warps per CTA
  k_source_emit<<<grid_for(ncells * 32, 256), 256, 0, st>>>(P, first_local, ncells, source_type, t_start, dt, call,
                                                           prefix, slot0);
}

// EvaluateRadiationEnergy (jaybenne.cpp:547-561): tally(cell) += w / dV over census photons.
// Freshly sourced photons sit cell by cell in the pool, so neighbouring lanes mostly hit the
// same cell: runs of equal cells inside a warp are summed with a segmented shuffle scan and
// issue one atomic per run. A warp whose lanes are spread over many cells (after transport)
// sees few equal neighbours in its ballot and takes the plain one-atomic-per-lane path.
__global__ void __launch_bounds__(256) k_tally(StepParams P, long long n, unsigned long long *n_alive) {
  const DMesh &m = P.mesh;
  unsigned long long a = 0;  // census totals ride along: photons in use and the energy they carry
  double e = 0.0;
  const unsigned FULL = 0xffffffffu;
  const unsigned lane = threadIdx.x & 31u;
  // Each warp takes TALLY_U consecutive groups of 32 slots per trip and requests all of their loads (cell and
  // weight) before it looks at any: the kernel is a 16 B / photon stream whose only enemy is load latency
  // (profiles/r02_ncu_tally*.md: 13 of 20 cycles per issue on the long scoreboard with one group in flight).
  constexpr int TALLY_U = 4;
  const long long warp = (blockIdx.x * (long long)blockDim.x + threadIdx.x) >> 5;
  const long long nwarps = ((long long)gridDim.x * blockDim.x) >> 5;
  const long long ngroups = (n + 31) / 32;
  for (long long g0 = warp * TALLY_U; g0 < ngroups; g0 += nwarps * TALLY_U) {
    int2 lc[TALLY_U];
    double ww[TALLY_U];
#pragma unroll
    for (int u = 0; u < TALLY_U; ++u) {
      const long long idx = (g0 + u) * 32 + lane;
      const bool in = idx < n;
      lc[u] = in ? __ldcs(&P.s.loc[idx]) : make_int2(-1, 0);
      ww[u] = in ? __ldcs(&P.s.w[idx]) : 0.0;
    }
#pragma unroll
    for (int u = 0; u < TALLY_U; ++u) {
      if (g0 + u >= ngroups) break;  // warp-uniform
      long long cell = -1;
      double v = 0.0;
      if (lc[u].x >= 0) {
        // cell volume of the photon's block: dx pairs of the 64-byte geometry line (same product, same order,
        // as coords.Volume: dx1 * dx2 * dx3)
        const BlkGeo *geo = m.geo + __ldg(&m.local_gid[lc[u].x]);
        const double dv = __ldg(&geo->xd[0]).y * __ldg(&geo->xd[1]).y * __ldg(&geo->xd[2]).y;
        cell = (long long)lc[u].x * m.ncell + lc[u].y;
        v = ww[u] / dv;
        a++;
        e += ww[u];
      }
      const long long prev = __shfl_up_sync(FULL, cell, 1);
      const bool head = lane == 0 || prev != cell;
      const unsigned heads = __ballot_sync(FULL, head);
      if (__popc(heads) > 16) {  // mostly distinct cells: aggregation would not pay
        if (cell >= 0) atomicAdd(&P.f.tally[cell], v);
        continue;
      }
      // segmented inclusive scan towards higher lanes; the last lane of each run holds the run's sum
      const unsigned seg_start = 31u - __clz(heads & ((2u << lane) - 1u));  // first lane of my run
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const double up = __shfl_up_sync(FULL, v, o);
        if (lane >= seg_start + o) v += up;
      }
      const bool tail = lane == 31u || ((heads >> (lane + 1)) & 1u);
      if (tail && cell >= 0) atomicAdd(&P.f.tally[cell], v);
    }
  }
  warp_add_u64(n_alive, a);
  warp_add_energy(P, EB_CENSUS, e);
}
// EvaluateRadiationEnergy straight after a sourcing pass into an EMPTY pool: every photon of a cell carries that
// cell's weight, so tally(cell) = sum over its src_num photons of src_ew / dV needs no pass over the swarm (the
// same additions the photon loop makes, in photon order). One thread per cell; census totals ride along.
__global__ void __launch_bounds__(256) k_tally_sourced(StepParams P, long long ncells, unsigned long long *n_alive) {
  const DMesh &m = P.mesh;
  unsigned long long a = 0;
  double e = 0.0;
  for (long long c = blockIdx.x * (long long)blockDim.x + threadIdx.x; c < ncells; c += (long long)gridDim.x * blockDim.x) {
    const BlkGeo *geo = m.geo + __ldg(&m.local_gid[(int)(c / m.ncell)]);
    const double dv = __ldg(&geo->xd[0]).y * __ldg(&geo->xd[1]).y * __ldg(&geo->xd[2]).y;
    const int npart = (int)round(P.f.src_num[c]);
    const double ew = P.f.src_ew[c];
    double t = 0.0;
    if (npart > 0) {
      const double v = ew / dv;
      for (int np = 0; np < npart; ++np) {
        t += v;
        e += ew;
      }
      a += (unsigned long long)npart;
    }
    P.f.tally[c] = t;
  }
  warp_add_u64(n_alive, a);
  warp_add_energy(P, EB_CENSUS, e);
}
void launch_tally_sourced(const StepParams &P, unsigned long long *n_alive, double *e_sum, cudaStream_t st) {
  const long long ncells = (long long)P.mesh.nlocal * P.mesh.ncell;
  cudaMemsetAsync(P.ebins + EB_CENSUS * EBINS, 0, sizeof(double) * EBINS, st);
  if (ncells > 0) k_tally_sourced<<<grid_for(ncells, 256), 256, 0, st>>>(P, ncells, n_alive);
  launch_fold_bins(P, EB_CENSUS, e_sum, st);
}
void launch_tally(const StepParams &P, long long n, unsigned long long *n_alive, double *e_sum, cudaStream_t st) {
  const long long ncells = (long long)P.mesh.nlocal * P.mesh.ncell;
  cudaMemsetAsync(P.f.tally, 0, sizeof(double) * ncells, st);
  cudaMemsetAsync(P.ebins + EB_CENSUS * EBINS, 0, sizeof(double) * EBINS, st);
  if (n > 0) k_tally<<<grid_for((n + 3) / 4, 256), 256, 0, st>>>(P, n, n_alive);
  launch_fold_bins(P, EB_CENSUS, e_sum, st);
}

// ---------------------------------------------------------------------------------------
// EvaluateRadiationEnergy without any allocation (jb_evaluate_radiation_energy_noalloc; no counterpart in the
// reference, added for callers that cannot call cudaMalloc: CUDA graph
// capture, tight serving loops, harnesses that forbid it). k_tally's scratch — the per-cell accumulation buffer,
// the 1024 census-energy bins, the folded energy and the photon counter — lives in ONE file-scope device array of
// a fixed, modest size (2 MiB + 8.3 KiB; every GPU that loads the library holds it until exit). In host code the
// array's name is only a stand-in: its device address is looked up once per device with cudaGetSymbolAddress.
// Every call shares the array (it keeps the last call's contents: what must start at zero is cleared here), so
// calls that overlap on the GPU — two streams, two contexts on one device — race.
// ---------------------------------------------------------------------------------------
static constexpr long long NOALLOC_MAX_CELLS = 262144;  // 64^3 cells per rank
static constexpr long long NOALLOC_DOUBLES = NOALLOC_MAX_CELLS + EBINS + 2;  // tally | bins | e_sum, n_alive (as u64)
static __device__ __align__(256) double g_noalloc_scratch[NOALLOC_DOUBLES];

long long tally_noalloc_max_cells() { return NOALLOC_MAX_CELLS; }
// tally_dst: ncells doubles (device or host), n_dst / e_dst: host. Returns the first CUDA error.
cudaError_t launch_tally_noalloc(const StepParams &P0, long long n, double *tally_dst, long long *n_dst, double *e_dst,
                                 cudaStream_t st) {
  static double *scratch_of[64] = {nullptr};  // device address of the array, per device
  int dev = 0;
  cudaError_t err = cudaGetDevice(&dev);
  if (err != cudaSuccess) return err;
  if (dev < 0 || dev >= 64) return cudaErrorInvalidDevice;
  if (!scratch_of[dev]) {
    double *ptr = nullptr;
    err = cudaGetSymbolAddress((void **)&ptr, g_noalloc_scratch);  // with `dev` current
    if (err != cudaSuccess) return err;
    scratch_of[dev] = ptr;
  }
  double *scr = scratch_of[dev];
  const long long ncells = (long long)P0.mesh.nlocal * P0.mesh.ncell;
  if (ncells > NOALLOC_MAX_CELLS) return cudaErrorInvalidValue;
  StepParams P = P0;  // the same kernel, its scratch and outputs redirected into the static array
  P.f.tally = scr;
  // only the census bins are used by k_tally (kind EB_CENSUS): place the bin table so that kind's row is scr + ncells_max
  P.ebins = scr + NOALLOC_MAX_CELLS - (long long)EB_CENSUS * EBINS;
  double *e_sum = scr + NOALLOC_MAX_CELLS + EBINS;
  unsigned long long *n_alive = reinterpret_cast<unsigned long long *>(scr + NOALLOC_MAX_CELLS + EBINS + 1);
  err = cudaMemsetAsync(scr + NOALLOC_MAX_CELLS + EBINS, 0, 2 * sizeof(double), st);
  if (err != cudaSuccess) return err;
  launch_tally(P, n, n_alive, e_sum, st);  // clears the tally cells and the bins it uses, then k_tally + k_fold_bins
  if ((err = cudaGetLastError()) != cudaSuccess) return err;
  if (tally_dst && ncells > 0 &&
      (err = cudaMemcpyAsync(tally_dst, scr, sizeof(double) * ncells, cudaMemcpyDefault, st)) != cudaSuccess)
    return err;
  if (e_dst && (err = cudaMemcpyAsync(e_dst, e_sum, sizeof(double), cudaMemcpyDeviceToHost, st)) != cudaSuccess) return err;
  if (n_dst && (err = cudaMemcpyAsync(n_dst, n_alive, sizeof(long long), cudaMemcpyDeviceToHost, st)) != cudaSuccess) return err;
  return cudaStreamSynchronize(st);
}

// UpdateFluid (jaybenne.cpp:603-612): u += energy_delta / dV
__global__ void k_update_fluid(StepParams P, long long ncells) {
  const DMesh &m = P.mesh;
  for (long long c = blockIdx.x * (long long)blockDim.x + threadIdx.x; c < ncells;
       c += (long long)gridDim.x * blockDim.x) {
    const DBlock &b = m.blocks[m.local_gid[(int)(c / m.ncell)]];
    const double dv = b.dx[0] * b.dx[1] * b.dx[2];
    const double delta = P.f.edelta[c] / dv;
    P.f.u[c] += delta;
  }
}
void launch_update_fluid(const StepParams &P, cudaStream_t st) {
  const long long ncells = (long long)P.mesh.nlocal * P.mesh.ncell;
  if (ncells > 0) k_update_fluid<<<grid_for(ncells, 256), 256, 0, st>>>(P, ncells);
}
// mcblock::UpdateDerived (mcblock.cpp:222-231): sie = u / rho
__global__ void k_update_derived_material(StepParams P, long long ncells) {
  for (long long c = blockIdx.x * (long long)blockDim.x + threadIdx.x; c < ncells;
       c += (long long)gridDim.x * blockDim.x)
    P.f.sie[c] = P.f.u[c] / P.f.rho[c];
}
void launch_update_derived_material(const StepParams &P, cudaStream_t st) {
  const long long ncells = (long long)P.mesh.nlocal * P.mesh.ncell;
  if (ncells > 0) k_update_derived_material<<<grid_for(ncells, 256), 256, 0, st>>>(P, ncells);
}
// mcblock::PostInitialization (mcblock.cpp:252-261): u = rho * sie
__global__ void k_post_initialization(StepParams P, long long ncells) {
  for (long long c = blockIdx.x * (long long)blockDim.x + threadIdx.x; c < ncells;
       c += (long long)gridDim.x * blockDim.x)
    P.f.u[c] = P.f.rho[c] * P.f.sie[c];
}
void launch_post_initialization(const StepParams &P, cudaStream_t st) {
  const long long ncells = (long long)P.mesh.nlocal * P.mesh.ncell;
  if (ncells > 0) k_post_initialization<<<grid_for(ncells, 256), 256, 0, st>>>(P, ncells);
}

// CheckCompletion (transport.cpp:198-209) + census sums: photons in use, photons with
// t < t_end, and the energy they carry.
__global__ void k_census(StepParams P, long long n, double t_end, unsigned long long *n_alive,
                         unsigned long long *n_unfinished) {
  unsigned long long a = 0, u = 0;
  double e = 0.0;
  for (long long idx = blockIdx.x * (long long)blockDim.x + threadIdx.x; idx < n;
       idx += (long long)gridDim.x * blockDim.x) {
    if (P.s.loc[idx].x < 0) continue;
    a++;
    e += P.s.w[idx];
    if (P.s.t[idx] < t_end) u++;
  }
  warp_add_u64(n_alive, a);
  warp_add_u64(n_unfinished, u);
  warp_add_energy(P, EB_CENSUS, e);
}
void launch_census(const StepParams &P, long long n, double t_end, unsigned long long *n_alive,
                   unsigned long long *n_unfinished, double *e_sum, cudaStream_t st) {
  cudaMemsetAsync(P.ebins + EB_CENSUS * EBINS, 0, sizeof(double) * EBINS, st);
  if (n > 0) k_census<<<grid_for(n, 256), 256, 0, st>>>(P, n, t_end, n_alive, n_unfinished);
  launch_fold_bins(P, EB_CENSUS, e_sum, st);
}

// ---------------------------------------------------------------------------------------
// compaction: slots [0, n) -> [0, n_alive). Dead slots below n_alive ("holes") are filled
// with live photons from slots >= n_alive ("tails"); only those photons move. Order is not
// preserved — nothing depends on it (streams are keyed per photon).
// counters[0] = #holes, counters[1] = #tails (must be zero on entry)
// ---------------------------------------------------------------------------------------
__global__ void k_compact_find(StepParams P, long long n, long long n_alive, long long *holes, long long *tails,
                               unsigned long long *counters) {
  // a warp scans 8 x 32 consecutive slots per step and claims list space with ONE atomic per list:
  // every atomic hits the same two addresses, so their number (not the bytes) sets the pace
  constexpr int ROUNDS = 8;
  const unsigned lane = threadIdx.x & 31u, lt = (1u << lane) - 1u;
  const long long warp = (blockIdx.x * (long long)blockDim.x + threadIdx.x) >> 5;
  const long long nwarps = ((long long)gridDim.x * blockDim.x) >> 5;
  for (long long base = warp * (32 * ROUNDS); base < n; base += nwarps * (32 * ROUNDS)) {
    unsigned mh[ROUNDS], mt[ROUNDS];
    int nh = 0, nt = 0;
#pragma unroll
    for (int r = 0; r < ROUNDS; ++r) {
      const long long idx = base + r * 32 + lane;
      const bool valid = idx < n;
      const bool alive = valid && P.s.loc[idx].x >= 0;
      mh[r] = __ballot_sync(0xffffffffu, valid && idx < n_alive && !alive);
      mt[r] = __ballot_sync(0xffffffffu, valid && idx >= n_alive && alive);
      nh += __popc(mh[r]);
      nt += __popc(mt[r]);
    }
    unsigned long long bh = 0, bt = 0;
    if (lane == 0) {
      if (nh) bh = atomicAdd(&counters[0], (unsigned long long)nh);
      if (nt) bt = atomicAdd(&counters[1], (unsigned long long)nt);
    }
    bh = __shfl_sync(0xffffffffu, bh, 0);
    bt = __shfl_sync(0xffffffffu, bt, 0);
#pragma unroll
    for (int r = 0; r < ROUNDS; ++r) {
      const long long idx = base + r * 32 + lane;
      if ((mh[r] >> lane) & 1u) holes[bh + __popc(mh[r] & lt)] = idx;
      if ((mt[r] >> lane) & 1u) tails[bt + __popc(mt[r] & lt)] = idx;
      bh += __popc(mh[r]);
      bt += __popc(mt[r]);
    }
  }
}
__global__ void k_compact_move(StepParams P, const long long *holes, const long long *tails,
                               const unsigned long long *counters) {
  const long long nh = (long long)counters[0];
  for (long long q = blockIdx.x * (long long)blockDim.x + threadIdx.x; q < nh;
       q += (long long)gridDim.x * blockDim.x) {
    const long long d = holes[q], s = tails[q];
    P.s.x[d] = P.s.x[s]; P.s.y[d] = P.s.y[s]; P.s.z[d] = P.s.z[s];
    P.s.t[d] = P.s.t[s]; P.s.w[d] = P.s.w[s]; P.s.e[d] = P.s.e[s];
    P.s.vx[d] = P.s.vx[s]; P.s.vy[d] = P.s.vy[s]; P.s.vz[d] = P.s.vz[s];
    P.s.loc[d] = P.s.loc[s];
    P.s.ida[d] = P.s.ida[s]; P.s.idb[d] = P.s.idb[s];
    P.s.loc[s].x = -1;
  }
}
void launch_compact(const StepParams &P, long long n, long long n_alive, long long *holes, long long *tails,
                    unsigned long long *counters, cudaStream_t st) {
  if (n <= 0 || n_alive >= n) return;
  cudaMemsetAsync(counters, 0, 2 * sizeof(unsigned long long), st);
  k_compact_find<<<grid_for(n, 256), 256, 0, st>>>(P, n, n_alive, holes, tails, counters);
  const long long nmove = n - n_alive;  // upper bound on #holes
  k_compact_move<<<grid_for(nmove, 256), 256, 0, st>>>(P, holes, tails, counters);
}

// ---------------------------------------------------------------------------------------
// Order-preserving squeeze: the live photons keep their relative order (so a pool laid down cell by cell stays
// near-sorted, see jb_transport_pool.cuh v12). Three steps on 1024-slot chunks: (1) one pass over `loc` writes a
// live-bit word per 32 slots and a count per chunk; (2) exclusive scan of the counts; (3) every swarm array is
// streamed through the one scratch array — a lane whose bit is set copies its element to chunk offset + rank —
// and swaps storage with it. The passes read 1 bit per slot instead of `loc`, and the sectors of a chunk's
// hole-only tail are never touched: traffic ~ 16 B per live photon and array.
// ---------------------------------------------------------------------------------------
static constexpr int SQ_CHUNK = 1024;  // 32 mask words; one CTA of 256 threads, 4 rounds
__global__ void __launch_bounds__(256) k_squeeze_flags(const int2 *__restrict__ loc, long long n, unsigned *__restrict__ mask,
                                                       int *__restrict__ cnt) {
  __shared__ int part[8];
  const long long base = blockIdx.x * (long long)SQ_CHUNK;
  const unsigned lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
  int2 lc[4];
#pragma unroll
  for (int r = 0; r < 4; ++r) {
    const long long idx = base + r * 256 + threadIdx.x;
    lc[r] = idx < n ? __ldcs(&loc[idx]) : make_int2(-1, 0);
  }
  int a = 0;
#pragma unroll
  for (int r = 0; r < 4; ++r) {
    const unsigned mk = __ballot_sync(0xffffffffu, lc[r].x >= 0);
    if (lane == 0) mask[blockIdx.x * 32ll + r * 8 + warp] = mk;
    a += __popc(mk);
  }
  if (lane == 0) part[warp] = a;
  __syncthreads();
  if (threadIdx.x == 0) cnt[blockIdx.x] = part[0] + part[1] + part[2] + part[3] + part[4] + part[5] + part[6] + part[7];
}
__global__ void __launch_bounds__(256) k_squeeze_move(const unsigned long long *__restrict__ src,
                                                      unsigned long long *__restrict__ dst,
                                                      const unsigned *__restrict__ mask,
                                                      const long long *__restrict__ prefix) {
  __shared__ unsigned mk[32];
  __shared__ int off[32];
  const long long base = blockIdx.x * (long long)SQ_CHUNK;
  const unsigned lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
  if (threadIdx.x < 32) {
    const unsigned w = mask[blockIdx.x * 32ll + threadIdx.x];
    mk[threadIdx.x] = w;
    // exclusive scan of the 32 word counts (word = round * 8 + warp = slot order)
    int c = __popc(w), inc = c;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int up = __shfl_up_sync(0xffffffffu, inc, o);
      if ((int)lane >= o) inc += up;
    }
    off[threadIdx.x] = inc - c;
  }
  __syncthreads();
  const long long out = prefix[blockIdx.x];
  unsigned long long v[4];
  bool live[4];
#pragma unroll
  for (int r = 0; r < 4; ++r) {
    live[r] = (mk[r * 8 + warp] >> lane) & 1u;
    if (live[r]) v[r] = __ldcs(&src[base + r * 256 + threadIdx.x]);
  }
#pragma unroll
  for (int r = 0; r < 4; ++r)
    if (live[r]) dst[out + off[r * 8 + warp] + __popc(mk[r * 8 + warp] & ((1u << lane) - 1u))] = v[r];
}
// `loc` of the n slots is the still-unsqueezed array; returns nothing the host needs (live count is known)
void launch_squeeze_flags(const int2 *loc, long long n, unsigned *mask, int *cnt, cudaStream_t st) {
  const long long nchunk = (n + SQ_CHUNK - 1) / SQ_CHUNK;
  if (nchunk > 0) k_squeeze_flags<<<(unsigned)nchunk, 256, 0, st>>>(loc, n, mask, cnt);
}
void launch_squeeze_move(const void *src, void *dst, long long n, const unsigned *mask, const long long *prefix,
                         cudaStream_t st) {
  const long long nchunk = (n + SQ_CHUNK - 1) / SQ_CHUNK;
  if (nchunk > 0)
    k_squeeze_move<<<(unsigned)nchunk, 256, 0, st>>>((const unsigned long long *)src, (unsigned long long *)dst, mask, prefix);
}
long long squeeze_chunks(long long n) { return (n + SQ_CHUNK - 1) / SQ_CHUNK; }

// ---------------------------------------------------------------------------------------
// Receive side of the swarm exchange: append records to the pool (Swarm::Receive).
// ---------------------------------------------------------------------------------------
__device__ __forceinline__ void unpack_records(const StepParams &P, const ExRecord *rec, long long nrec, long long slot0,
                                               double t_end) {
  const DMesh &m = P.mesh;
  unsigned long long c_unf = 0, c_err = 0, c_rcv = 0;
  for (long long q = blockIdx.x * (long long)blockDim.x + threadIdx.x; q < nrec;
       q += (long long)gridDim.x * blockDim.x) {
    const ExRecord r = rec[q];
    const long long idx = slot0 + q;
    const int gid = r.dest_gid;
    int lb = -1;
    if (gid >= 0 && gid < m.nblocks) lb = m.blocks[gid].local;
    if (lb < 0) {
      c_err++;
      P.s.loc[idx] = make_int2(-1, 0);
      continue;
    }
    const DBlock &b = m.blocks[gid];
    const int i = (int)floor((r.x - b.xmin[0]) / b.dx[0]);
    const int j = m.ndim > 1 ? (int)floor((r.y - b.xmin[1]) / b.dx[1]) : 0;
    const int k = m.ndim > 2 ? (int)floor((r.z - b.xmin[2]) / b.dx[2]) : 0;
    P.s.x[idx] = r.x; P.s.y[idx] = r.y; P.s.z[idx] = r.z; P.s.t[idx] = r.t;
    P.s.w[idx] = r.w; P.s.e[idx] = r.e;
    P.s.vx[idx] = r.vx; P.s.vy[idx] = r.vy; P.s.vz[idx] = r.vz;
    P.s.loc[idx] = make_int2(lb, (k * m.nxb[1] + j) * m.nxb[0] + i);
    P.s.ida[idx] = make_uint2(r.rng[0], r.rng[1]);
    P.s.idb[idx] = make_uint2(r.rng[2], r.rng[3]);
    c_rcv++;
    if (r.t < t_end) c_unf++;
  }
  warp_add_u64(&P.stats->n_received, c_rcv);
  warp_add_u64(&P.stats->n_unfinished, c_unf);
  warp_add_u64(&P.stats->n_error, c_err);
}
__global__ void k_unpack(StepParams P, const ExRecord *rec, long long nrec, long long slot0, double t_end) {
  unpack_records(P, rec, nrec, slot0, t_end);
}
void launch_unpack(const StepParams &P, const ExRecord *rec, long long nrec, long long slot0, double t_end,
                   cudaStream_t st) {
  if (nrec > 0) k_unpack<<<grid_for(nrec, 128), 128, 0, st>>>(P, rec, nrec, slot0, t_end);
}

// ---------------------------------------------------------------------------------------
// sort photons by (block, cell): counting sort. Photons of one cell end up adjacent, so a
// warp's cell-field loads and tally atomics hit few lines. The permutation is applied
// array by array through one scratch array (out-of-place for 1e9 photons would not fit).
// ---------------------------------------------------------------------------------------
__global__ void k_sort_count(StepParams P, long long n, int *cell_count) {
  const DMesh &m = P.mesh;
  for (long long idx = blockIdx.x * (long long)blockDim.x + threadIdx.x; idx < n;
       idx += (long long)gridDim.x * blockDim.x) {
    const int2 lc = P.s.loc[idx];
    if (lc.x < 0) continue;
    atomicAdd(&cell_count[(long long)lc.x * m.ncell + lc.y], 1);
  }
}
// perm[dest] = source slot
__global__ void k_sort_perm(StepParams P, long long n, long long *cell_offset, unsigned int *perm) {
  const DMesh &m = P.mesh;
  for (long long idx = blockIdx.x * (long long)blockDim.x + threadIdx.x; idx < n;
       idx += (long long)gridDim.x * blockDim.x) {
    const int2 lc = P.s.loc[idx];
    if (lc.x < 0) continue;
    const long long d =
        (long long)atomicAdd((unsigned long long *)&cell_offset[(long long)lc.x * m.ncell + lc.y], 1ull);
    perm[d] = (unsigned int)idx;
  }
}
template <typename T>
__global__ void k_gather(const T *__restrict__ src, T *__restrict__ dst, const unsigned int *__restrict__ perm,
                         long long n) {
  for (long long q = blockIdx.x * (long long)blockDim.x + threadIdx.x; q < n;
       q += (long long)gridDim.x * blockDim.x)
    dst[q] = src[perm[q]];
}
void launch_sort_keys(const StepParams &P, long long n, int *cell_count, cudaStream_t st) {
  if (n > 0) k_sort_count<<<grid_for(n, 256), 256, 0, st>>>(P, n, cell_count);
}
void launch_sort_perm(const StepParams &P, long long n, long long *cell_offset, unsigned int *perm, cudaStream_t st) {
  if (n > 0) k_sort_perm<<<grid_for(n, 256), 256, 0, st>>>(P, n, cell_offset, perm);
}
void launch_gather_b64(const void *src, void *dst, const unsigned int *perm, long long n, cudaStream_t st) {
  if (n > 0)
    k_gather<unsigned long long><<<grid_for(n, 256), 256, 0, st>>>((const unsigned long long *)src,
                                                                  (unsigned long long *)dst, perm, n);
}

}  // namespace jbd

namespace jbd {
// ---------------------------------------------------------------------------------------
// FP64 FMA rate probe (jb_measure_fp64_peak): 8 independent register-resident DFMA chains
// per thread, 256 threads, enough CTAs to fill every SM. 16 flops per inner iteration.
// ---------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_fp64_chain(double *sink, int iters) {
  const double a = 1.0 + 1e-9 * threadIdx.x, b = 1e-12 * (blockIdx.x + 1);
  double r0 = 0.1, r1 = 0.2, r2 = 0.3, r3 = 0.4, r4 = 0.5, r5 = 0.6, r6 = 0.7, r7 = 0.8;
#pragma unroll 4
  for (int i = 0; i < iters; ++i) {
    r0 = fma(r0, a, b); r1 = fma(r1, a, b); r2 = fma(r2, a, b); r3 = fma(r3, a, b);
    r4 = fma(r4, a, b); r5 = fma(r5, a, b); r6 = fma(r6, a, b); r7 = fma(r7, a, b);
  }
  const double s = ((r0 + r1) + (r2 + r3)) + ((r4 + r5) + (r6 + r7));
  if (s == 123.456) sink[0] = s;  // keeps the chains alive without a store in the common case
}
// jb_selftest_division: div_normal(a, b) against a / b, bit for bit. Operands: sign and 52 mantissa bits
// from Philox, exponents uniform in [-300, 300]; every 16th numerator is 0.
__global__ void k_selftest_division(long long n, unsigned long long *mismatches) {
  unsigned long long bad = 0;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    double u0, u1, u2, u3;
    philox_pair(0u, (uint32_t)i, (uint32_t)(i >> 32), 0x5e1f7e57u, 0x9E3779B9u, 0xBB67AE85u, u0, u1);
    philox_pair(1u, (uint32_t)i, (uint32_t)(i >> 32), 0x5e1f7e57u, 0x9E3779B9u, 0xBB67AE85u, u2, u3);
    const int ea = (int)(u2 * 601.0) - 300, eb = (int)(u3 * 601.0) - 300;
    const unsigned long long ma = (unsigned long long)(u0 * 9007199254740992.0), mb = (unsigned long long)(u1 * 9007199254740992.0);
    double a = jbm::u2d(((unsigned long long)(1023 + ea) << 52) | (ma & 0x000fffffffffffffull) | ((ma >> 52) << 63));
    const double b = jbm::u2d(((unsigned long long)(1023 + eb) << 52) | (mb & 0x000fffffffffffffull) | ((mb >> 52) << 63));
    if ((i & 15) == 0) a = 0.0;
    const double q1 = jbm::div_normal(a, b), q2 = a / b;
    if (jbm::d2u(q1) != jbm::d2u(q2)) bad++;
  }
  warp_add_u64(mismatches, bad);
}
void launch_selftest_division(long long n, unsigned long long *mismatches, cudaStream_t st) {
  k_selftest_division<<<148 * 8, 256, 0, st>>>(n, mismatches);
}
void launch_fp64_chain(double *sink, int iters, int blocks, cudaStream_t st) {
  k_fp64_chain<<<blocks, 256, 0, st>>>(sink, iters);
}

// ---------------------------------------------------------------------------------------
// Direct peer exchange over NVLink: replaces the count all-gather + host synchronisation + grouped
// ncclSend/ncclRecv of one transport iteration (jaybenne.cpp:26-61, 113-131 in the reference: swarm
// Send / Receive + CheckCompletion) by stores into the peers' mapped inboxes and flags, so the host
// can queue several iterations without waiting for any of them.
// ---------------------------------------------------------------------------------------
__device__ __forceinline__ unsigned long long ld_acquire_sys(const unsigned long long *p) {
  unsigned long long v;
  asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_release_sys(unsigned long long *p, unsigned long long v) {
  asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}

// one CTA: reserve a range of every peer's inbox for this iteration's records
__global__ void k_p2p_reserve(StepParams P, P2PState *st, const P2PPeers *peers, int R, long long inbox_cap) {
  const int r = threadIdx.x;
  if (r >= R) return;
  unsigned long long cnt = 0, base = 0;
  if (!st->done && r != P.mesh.my_rank) {
    cnt = P.out.count[r];
    if ((long long)cnt > P.out.cap) cnt = (unsigned long long)P.out.cap;  // the overflow itself is counted by the kernels
    if (cnt) {
      const int par = (int)(st->epoch & 1ull);
      base = atomicAdd_system(&peers->mail[r]->inbox_count[par], cnt);
      if ((long long)(base + cnt) > inbox_cap) {  // receiver's inbox is full: send what fits, flag the rest
        atomicOr(&st->error, 1);
        cnt = (long long)base < inbox_cap ? (unsigned long long)inbox_cap - base : 0ull;
      }
    }
  }
  st->cnt[r] = cnt;
  st->base[r] = base;
}
// blockIdx.y = peer: copy my outbox segment for that peer into its inbox, 16 bytes per thread and trip
__global__ void k_p2p_copy(StepParams P, const P2PState *st, const P2PPeers *peers) {
  const int r = blockIdx.y;
  const unsigned long long cnt = st->cnt[r];
  if (cnt == 0) return;
  const int par = (int)(st->epoch & 1ull);
  const uint4 *src = reinterpret_cast<const uint4 *>(P.out.rec + (size_t)r * P.out.cap);
  uint4 *dst = reinterpret_cast<uint4 *>(peers->inbox[par][r] + st->base[r]);
  const unsigned long long nq = cnt * (sizeof(ExRecord) / sizeof(uint4));
  for (unsigned long long q = blockIdx.x * (unsigned long long)blockDim.x + threadIdx.x; q < nq;
       q += (unsigned long long)gridDim.x * blockDim.x)
    dst[q] = src[q];
  __threadfence_system();  // the records are visible to the peer before the flag of k_p2p_signal can be
}
// one CTA: tell every rank how much I sent in total and that my records for it have landed
__global__ void k_p2p_signal(StepParams P, P2PState *st, const P2PPeers *peers, int R) {
  __shared__ unsigned long long total;
  const int r = threadIdx.x;
  if (r == 0) {
    unsigned long long t = 0;
    for (int q = 0; q < R; ++q) t += st->cnt[q];
    total = t;
  }
  __syncthreads();
  if (r >= R) return;
  if (st->done) return;
  const unsigned long long epoch = st->epoch;
  const int par = (int)(epoch & 1ull);
  P.out.count[r] = 0ull;  // my outbox is empty again
  __threadfence_system();
  P2PMail *mail = peers->mail[r];
  *reinterpret_cast<volatile unsigned long long *>(&mail->sent[par][P.mesh.my_rank]) = total;
  __threadfence_system();
  st_release_sys(&mail->arrive[P.mesh.my_rank], epoch + 1ull);
}
// one CTA: wait until every rank's records for me have landed, then advance the loop state
__global__ void k_p2p_wait(StepParams P, P2PState *st, P2PMail *mine, int R, long long inbox_cap) {
  __shared__ int timed_out;
  const int r = threadIdx.x;
  if (r == 0) timed_out = 0;
  __syncthreads();
  if (st->done) return;
  const unsigned long long epoch = st->epoch;
  const int par = (int)(epoch & 1ull);
  if (r < R) {
    const long long t0 = clock64();
    while (ld_acquire_sys(&mine->arrive[r]) < epoch + 1ull) {
      if (clock64() - t0 > 20000000000LL) {  // ~10 s: a peer died; fail the step instead of hanging the GPU
        timed_out = 1;
        break;
      }
      __nanosleep(200);
    }
  }
  __syncthreads();
  if (r != 0) return;
  unsigned long long total = 0;
  for (int q = 0; q < R; ++q) total += *reinterpret_cast<volatile unsigned long long *>(&mine->sent[par][q]);
  unsigned long long nrecv = *reinterpret_cast<volatile unsigned long long *>(&mine->inbox_count[par]);
  mine->inbox_count[par] = 0ull;  // nobody touches this parity again before my next-but-one signal
  if (timed_out) atomicOr(&st->error, 4);
  if ((long long)nrecv > inbox_cap) {
    nrecv = (unsigned long long)inbox_cap;
    atomicOr(&st->error, 1);
  }
  if (st->n + (long long)nrecv > st->capacity) {
    atomicOr(&st->error, 2);
    nrecv = (unsigned long long)(st->capacity - st->n);
  }
  st->range[0] = st->n;
  st->range[1] = (long long)nrecv;
  st->n += (long long)nrecv;
  st->parity = par;
  st->sent_last = total;
  st->iterations += 1;
  st->done = (total == 0ull || timed_out) ? 1 : 0;
  st->epoch = epoch + 1ull;
}
__global__ void k_p2p_unpack(StepParams P, const P2PState *st, const ExRecord *inbox0, const ExRecord *inbox1, double t_end) {
  const long long nrec = st->range[1];
  if (nrec <= 0) return;
  unpack_records(P, st->parity ? inbox1 : inbox0, nrec, st->range[0], t_end);
}

void launch_p2p_push(const StepParams &P, P2PState *state, const P2PPeers *peers, int n_ranks, long long inbox_cap,
                     cudaStream_t st) {
  k_p2p_reserve<<<1, P2P_MAX_RANKS, 0, st>>>(P, state, peers, n_ranks, inbox_cap);
  k_p2p_copy<<<dim3(148, (unsigned)n_ranks), 256, 0, st>>>(P, state, peers);  // CTAs of a peer with nothing to send exit at once
  k_p2p_signal<<<1, P2P_MAX_RANKS, 0, st>>>(P, state, peers, n_ranks);
}
void launch_p2p_wait_unpack(const StepParams &P, P2PState *state, P2PMail *my_mail, ExRecord *inbox0, ExRecord *inbox1,
                            int n_ranks, long long inbox_cap, double t_end, bool ddmc, cudaStream_t st) {
  k_p2p_wait<<<1, P2P_MAX_RANKS, 0, st>>>(P, state, my_mail, n_ranks, inbox_cap);
  k_p2p_unpack<<<148 * 8, 256, 0, st>>>(P, state, inbox0, inbox1, t_end);
  if (ddmc && P.mesh.ndim >= 2) {
    StepParams Q = P;
    Q.range = state->range;  // address of the device-side (first, n)
    if (P.mesh.ndim == 2)
      k_bface<2><<<148, 128, 0, st>>>(Q, 0, 0);
    else
      k_bface<3><<<148, 128, 0, st>>>(Q, 0, 0);
  }
}

}  // namespace jbd
