// jb_kernels.cuh — device data layout shared by the kernels (jb_kernels.cu) and the
// context / C ABI (jb_lib.cu).
//
// HBM layout
//   cell fields   one FP64 array per field, [local block][k][j][i] (interior cells, x1
//                 fastest), local blocks in gid order. Face probabilities: one array per
//                 direction, [local block][...][face].
//   photons       structure of arrays, one slot per photon, slots [0, n) in use:
//                 12 arrays of 8 B: 9 x FP64 (x y z t w e vx vy vz), (int32 local block,
//                 int32 flat cell), and the 16 B stream identity as two 8 B words (birth
//                 index, birth gid | source call, draws used) = 96 B per photon. Every
//                 array has the same element size so one scratch array serves the cell
//                 sort for all of them. loc.x < 0 marks a slot whose photon is gone
//                 (absorbed, escaped or handed to another rank) until the next compaction.
//   block table   one DBlock per GLOBAL block (every rank knows the whole mesh, as every
//                 Parthenon rank does) + the finest-level block lattice -> gid.
#ifndef JAYBENNE_B200_JB_KERNELS_CUH_
#define JAYBENNE_B200_JB_KERNELS_CUH_

#include "jb_device.cuh"

namespace jbd {

struct DBlock {
  double xmin[3], xmax[3], dx[3];
  double xmin0[3];  // coordinate of ghost-inclusive index 0 (UniformCartesian xmin_)
  int level, lx[3], rank, local, gid, pad;
};

// what a segment needs of its meshblock, one 64-byte line: (xmin0[d], dx[d]) pairs so each
// direction arrives in one 16-byte load, and the smallest cell width (the step cap)
struct __align__(16) BlkGeo {
  double2 xd[3];
  double dx_push, pad;
};

struct DMesh {
  int ndim, nxb[3], ng[3], nlat[3], nrb[3];
  int max_level, nblocks, nlocal, my_rank, n_ranks, ncell;
  double xmin[3], xmax[3];
  int swarm_bc[6], mesh_bc[6];
  const DBlock *blocks;   // [nblocks]
  const int *lattice;     // [nlat2][nlat1][nlat0] -> gid
  const int *local_gid;   // [nlocal] -> gid
  const BlkGeo *geo;      // [nblocks]
  // [nlocal][27] gid of the same-level block at offset (bx, by, bz) in {-1,0,1}^3, index
  // (bz+1)*9 + (by+1)*3 + (bx+1), periodic swarm boundaries wrapped; -1 where the neighbour
  // is on another level, or beyond a non-periodic domain face (-> find_block path)
  const int *nbr;
};

// what one IMC segment looks up in its cell, packed so it arrives in one 32-byte sector:
// rho and fleck_factor as stored, temp = EOS temperature of (rho, sie), ss = total scattering
// coefficient — both evaluated with the very operations the per-segment code would use, once
// per cell per cycle instead of once per segment (k_pack_props).
struct __align__(32) CellProps {
  double rho, temp, ff, ss;
};

struct Fields {
  double *rho, *sie, *u, *tally, *fleck, *src_ew, *src_num, *edelta;
  CellProps *props;
  // multi-rank DDMC only: rho and sie of EVERY block, [gid][cell], so the face probabilities at a
  // rank boundary see the neighbour's cells (the reference exchanges ghost zones, jaybenne.cpp:108-110
  // and mcblock_driver.cpp:68); null when one rank owns the whole mesh
  double *g_rho, *g_sie;
  double *fp[3];
  long long nface[3];  // values per block of fp[d]
};

struct Swarm {
  double *x, *y, *z, *t, *w, *e, *vx, *vy, *vz;
  int2 *loc;   // (local block or -1 = slot free, flat interior cell)
  uint2 *ida;  // stream identity: (birth index, birth gid)
  uint2 *idb;  //                  (source call, draws consumed)
};

// record moved between ranks (jb_exchange_*): 96 bytes
struct ExRecord {
  double x, y, z, t, w, e, vx, vy, vz;
  uint32_t rng[4];
  int32_t dest_gid;
  int32_t pad;
};

struct Outbox {
  ExRecord *rec;              // [n_ranks][cap]
  unsigned long long *count;  // [n_ranks]
  long long cap;
};

// ---- direct peer exchange over NVLink (jb_lib.cu: p2p_*) -------------------------------------------
// One mailbox per rank, in that rank's device memory and mapped by every peer (CUDA IPC). A sender
// reserves a range of the receiver's inbox with one system-scope atomicAdd per peer and iteration,
// copies its outbox segment there with plain stores over NVLink, then publishes `sent` and `arrive`.
// Inboxes and their counters are double-buffered by iteration parity: a peer can be one iteration ahead.
static constexpr int P2P_MAX_RANKS = 64;
struct P2PMail {
  unsigned long long inbox_count[2];            // records reserved in inbox[parity] (remote atomics)
  unsigned long long arrive[P2P_MAX_RANKS];     // arrive[src] = last epoch src finished pushing to me
  unsigned long long sent[2][P2P_MAX_RANKS];    // sent[parity][src] = records src sent to ALL ranks that epoch
};
struct P2PPeers {                               // device table of the mapped peer pointers
  ExRecord *inbox[2][P2P_MAX_RANKS];
  P2PMail *mail[P2P_MAX_RANKS];
};
struct P2PState {                               // this rank's loop state, device resident: no host round trips
  long long range[2];                           // (first, n) of the next transport launch = the arrivals
  long long n, capacity;                        // pool size and limit
  unsigned long long epoch;                     // iteration counter across the whole run (never reset)
  unsigned long long sent_last;                 // records ALL ranks sent in the last iteration
  int done, iterations, error, parity;          // error: 1 inbox overflow, 2 pool overflow, 4 peer timed out
  unsigned long long base[P2P_MAX_RANKS];       // reservation results of the current push
  unsigned long long cnt[P2P_MAX_RANKS];        // records to push per peer (outbox counts, clamped)
};

// counters accumulated on the device during one radiation step
struct DStats {
  unsigned long long n_sourced, n_absorbed, n_escaped, n_sent, n_received, n_crossings;
  unsigned long long n_segments, n_scatters, n_ddmc, n_unfinished, n_error, n_dead;
  unsigned long long n_events, n_overflow;
  unsigned long long n_bface;        // photons spread over a fine face by SampleFace2D / SampleFace3D
  unsigned long long n_bface_ghost;  // block-face resampling indexed a face probability outside the block
  unsigned long long n_validate;  // VALIDATE builds: index / division cross-checks that disagreed
  double e_sourced, e_absorbed, e_escaped;
};

struct EventLog {
  jb_event *ev;
  long long cap;
  unsigned long long *count;  // == &DStats::n_events
};

// Energy sums (sourced / absorbed / escaped / census) are accumulated per thread, reduced
// over the warp by shuffles and added to one of EBINS bins chosen by the warp's id, then the
// bins are folded pairwise (k_fold_bins). A sum over 1e9 photons therefore goes through
// ~log2 levels of additions of comparable magnitudes instead of 3e7 atomics onto one double,
// which is what keeps the per-cycle conservation residual below 1e-12 at full size.
static constexpr int EBINS = 1024;
enum { EB_SOURCED = 0, EB_ABSORBED = 1, EB_ESCAPED = 2, EB_CENSUS = 3, EB_KINDS = 4 };

struct StepParams {
  DMesh mesh;
  Fields f;
  Swarm s;
  Closure cl;
  Outbox out;
  EventLog log;
  DStats *stats;
  double *ebins;  // [EB_KINDS][EBINS]
  unsigned long long *work;  // chunk counter of the persistent transport kernel (zero at launch)
  const long long *range;    // non-null: the kernel reads (first, n) here (P2PState::range) instead of its arguments
  uint32_t key0, key1_particle, key1_cell;
  PKeys rk;  // the ten Philox round keys of the particle stream (key0, key1_particle bumped per round)
  int use_ddmc;
  double tau_ddmc;
};

// ---- launchers (defined in jb_kernels.cu) -----------------------------------------------
void launch_fleck(const StepParams &P, double dt, cudaStream_t st);
void launch_face_prob(const StepParams &P, cudaStream_t st);
void launch_source_count(const StepParams &P, int first_local, int nlocal_call, int source_type, double dt,
                         double npc, uint32_t call, int *cnt, cudaStream_t st);
void launch_exclusive_scan(const int *cnt, long long *prefix, long long n, long long *scratch,
                           long long *total, cudaStream_t st);
long long scan_scratch_elems(long long n);
void launch_source_emit(const StepParams &P, int first_local, int nlocal_call, int source_type, double t_start,
                        double dt, uint32_t call, const long long *prefix, long long slot0, long long total,
                        cudaStream_t st);
void launch_transport(const StepParams &P, long long first, long long n, double t_start, double dt,
                      cudaStream_t st);
// knobs of the persistent IMC kernel (jb_lib.cu reads them from the environment once)
struct TransportTuning {
  int variant;      // 0 = thread-per-photon history kernel, 1 = persistent warps with lane refill (v9),
                    // 2 = warp-private photon pools in shared memory, phase by phase (v10, IMC only; default)
  int ctas_per_sm;  // CTAs of 128 threads per SM: 4, 6 or 8 (default)
  int refill_min;   // refill when at least this many lanes of a warp are idle
  int validate;     // 1: cross-check the face-crossing index update against Xtoijk (n_error)
};
void set_transport_tuning(const TransportTuning &t);
bool transport_takes_device_range(const StepParams &P);
void set_pool_tuning(int slots, int refill_min, int ready_low);
void set_pool_prefetch(int on);
void set_pool_fixed_chunk(int on);
void launch_pack_props(const StepParams &P, cudaStream_t st);
// folds ebins[kind] into *dst (dst = &stats->e_* or the census scratch)
void launch_fold_bins(const StepParams &P, int kind, double *dst, cudaStream_t st);
void launch_tally(const StepParams &P, long long n, unsigned long long *n_alive, double *e_sum, cudaStream_t st);
long long tally_noalloc_max_cells();
cudaError_t launch_tally_noalloc(const StepParams &P, long long n, double *tally_dst, long long *n_dst, double *e_dst,
                                 cudaStream_t st);
void launch_tally_sourced(const StepParams &P, unsigned long long *n_alive, double *e_sum, cudaStream_t st);
void launch_update_fluid(const StepParams &P, cudaStream_t st);
void launch_update_derived_material(const StepParams &P, cudaStream_t st);
void launch_post_initialization(const StepParams &P, cudaStream_t st);
void launch_census(const StepParams &P, long long n, double t_end, unsigned long long *n_alive,
                   unsigned long long *n_unfinished, double *e_sum, cudaStream_t st);
// unstable compaction of slots [0, n): returns nothing; n_alive must be known (host)
void launch_compact(const StepParams &P, long long n, long long n_alive, long long *holes, long long *tails,
                    unsigned long long *counters, cudaStream_t st);
void launch_unpack(const StepParams &P, const ExRecord *rec, long long nrec, long long slot0, double t_end,
                   cudaStream_t st);
void launch_bface(const StepParams &P, long long n, cudaStream_t st);
void launch_bface_range(const StepParams &P, long long first, long long n, cudaStream_t st);
// direct peer exchange: one iteration = push (reserve + copy + signal), wait, unpack [+ bface], all on `st`
void launch_p2p_push(const StepParams &P, P2PState *state, const P2PPeers *peers, int n_ranks, long long inbox_cap,
                     cudaStream_t st);
void launch_p2p_wait_unpack(const StepParams &P, P2PState *state, P2PMail *my_mail, ExRecord *inbox0, ExRecord *inbox1,
                            int n_ranks, long long inbox_cap, double t_end, bool ddmc, cudaStream_t st);
// cell sort: key = blk * ncell + cell
void launch_squeeze_flags(const int2 *loc, long long n, unsigned *mask, int *cnt, cudaStream_t st);
void launch_squeeze_move(const void *src, void *dst, long long n, const unsigned *mask, const long long *prefix,
                         cudaStream_t st);
long long squeeze_chunks(long long n);
void launch_sort_keys(const StepParams &P, long long n, int *cell_count, cudaStream_t st);
void launch_sort_perm(const StepParams &P, long long n, long long *cell_offset, unsigned int *perm, cudaStream_t st);
void launch_selftest_division(long long n, unsigned long long *mismatches, cudaStream_t st);
void launch_fp64_chain(double *sink, int iters, int blocks, cudaStream_t st);
void launch_gather_b64(const void *src, void *dst, const unsigned int *perm, long long n, cudaStream_t st);

}  // namespace jbd

#endif  // JAYBENNE_B200_JB_KERNELS_CUH_
