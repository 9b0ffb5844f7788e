/* jaybenne_b200.h — C ABI of the B200-native jaybenne IMC hot path.
 *
 * This is the drop-in boundary: plain C types, pointers and sizes only. Every entry
 * point names the reference interface it replaces (paths relative to lanl/jaybenne
 * @ 04947dfa). The reference reaches this path through Parthenon types
 * (MeshData<Real>*, ParameterInput*, Swarm); a Parthenon-side adapter copies the
 * package parameters, the block table, the host material fields and (optionally) the
 * `photons` swarm across this boundary — see INTEGRATION.md for the stub.
 *
 * Conventions
 *  - every call returns jb_status; nothing throws across the ABI; jb_last_error() gives
 *    the message for the last failing call on that context.
 *  - the library owns all device memory; the caller owns every host buffer it passes.
 *  - a jb_ctx drives ONE GPU (the current CUDA device at jb_create) from one host
 *    thread; multi-GPU = one context (one process) per GPU, see jb_comm_*.
 *  - cell arrays are interior cells only, x1 fastest: idx = (k*nx2 + j)*nx1 + i.
 *  - there is no CPU fallback: without a CUDA device every compute call fails with
 *    JB_ERR_NO_DEVICE.
 */
#ifndef JAYBENNE_B200_H_
#define JAYBENNE_B200_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define JB_ABI_VERSION 1

typedef enum jb_status {
  JB_OK = 0,
  JB_ERR_INVALID = 1,     /* bad argument / PARTHENON_REQUIRE-style input failure      */
  JB_ERR_CUDA = 2,        /* CUDA runtime error                                         */
  JB_ERR_CAPACITY = 3,    /* particle pool / buffer capacity exceeded                   */
  JB_ERR_STATE = 4,       /* call order violated (mesh not set, ...)                    */
  JB_ERR_UNSUPPORTED = 5, /* e.g. source_strategy = energy (reference: sourcing.cpp:38) */
  JB_ERR_ITERATIONS = 6,  /* max_transport_iterations exhausted (jaybenne.cpp:113)      */
  JB_ERR_NO_DEVICE = 7,   /* no CUDA device: the product has no CPU path                */
  JB_ERR_COMM = 8         /* multi-rank exchange failure                                */
} jb_status;

/* mcblock's closure choices: src/mcblock/opacity.hpp:23-30, src/mcblock/mcblock.cpp:95-143 */
enum { JB_OPAC_GRAY = 0, JB_OPAC_EPBREMSS = 1 };
enum { JB_SCAT_GRAY = 0, JB_SCAT_THOMSON = 1 };
/* swarm / mesh boundary kinds: inputs/stepdiff.in:37-51 */
enum { JB_BC_PERIODIC = 0, JB_BC_REFLECT = 1 /* "jaybenne_reflecting" */, JB_BC_OUTFLOW = 2 };
/* jaybenne::SourceType, src/jaybenne/jaybenne.hpp:56 */
enum { JB_SOURCE_THERMAL = 0, JB_SOURCE_EMISSION = 1 };
/* jaybenne::SourceStrategy, src/jaybenne/jaybenne.hpp:55 */
enum { JB_STRATEGY_UNIFORM = 0, JB_STRATEGY_ENERGY = 1 };
/* random streams. PHILOX = per-particle counter-based streams (the product).
 * XORSHIFT_POOL = the reference's Kokkos::Random_XorShift64_Pool borrowed in serial
 * order; implemented by the oracle only (the CUDA library rejects it). */
enum { JB_RNG_PHILOX = 0, JB_RNG_XORSHIFT_POOL = 1,
       /* oracle timing builds only: one XorShift64 pool state per OpenMP thread, which is what
        * the reference's pool gives the Kokkos OpenMP backend (not reproducible) */
       JB_RNG_XORSHIFT_THREADS = 2 };
/* gray absorption evaluation. KIRCHHOFF = the reference's literal j_nu / B_nu round trip
 * (gray_opacity_photons.hpp:50-54); DIRECT = rho*kappa (differs by <= 1 ulp, and where
 * B_nu under/overflows). */
enum { JB_GRAY_KIRCHHOFF = 0, JB_GRAY_DIRECT = 1 };

/* Package parameters: the <jaybenne> input block (src/jaybenne/jaybenne.cpp:163-223)
 * plus the three closures mcblock injects (EOS / Opacity / Scattering,
 * src/jaybenne/jaybenne.cpp:227-233, src/mcblock/mcblock.cpp:78-143).
 * jb_config_default() fills the reference defaults. */
typedef struct jb_config {
  /* <jaybenne> */
  int64_t num_particles;        /* required                                              */
  double dt;                    /* default DBL_MAX                                       */
  double min_swarm_occupancy;   /* default 0, must be in [0,1)                           */
  double numin, numax;          /* stored only (inert, as in the reference)              */
  int32_t unique_rank_seeds;    /* stored only: the pool is seeded with the raw seed     */
  int64_t seed;                 /* default 123                                           */
  int32_t max_transport_iterations; /* default 10000                                     */
  int32_t use_ddmc;             /* default 0                                             */
  double tau_ddmc;              /* default 5.0                                           */
  int32_t source_strategy;      /* JB_STRATEGY_*; ENERGY is rejected at sourcing         */
  int32_t do_emission;          /* default 1                                             */
  int32_t do_feedback;          /* default 1                                             */
  /* Opacity closure: NonCGSUnits<Gray|EPBremss> */
  int32_t opacity_model;        /* JB_OPAC_*                                             */
  double opacity_kappa;         /* Gray kappa (cm^2/g); "none" == Gray(0)                */
  /* Scattering closure: NonCGSUnitsS<GrayS|ThomsonS> */
  int32_t scattering_model;     /* JB_SCAT_*                                             */
  double scattering_sigma;      /* GrayS sigma; "none" == GrayS(0)                       */
  double apm;                   /* average particle mass                                 */
  /* code -> CGS unit scales shared by both closures (mcblock.cpp:84-91) */
  double time_scale, mass_scale, length_scale, temperature_scale;
  /* EOS closure: IdealGas(gm1, Cv) (mcblock.cpp:78-80) */
  double eos_gm1, eos_cv;
  /* implementation knobs (no reference counterpart) */
  int32_t rng_mode;             /* JB_RNG_*                                              */
  int32_t gray_mode;            /* JB_GRAY_*                                             */
  int32_t source_nblocks;       /* `nblocks` in npc = N/num_cells/(nblocks*nbtotal)
                                   (sourcing.cpp:68-69); <=0: this rank's block count    */
  int32_t sort_interval;        /* re-sort particles by cell every n cycles; 0 = never   */
  int32_t reserved[8];
} jb_config;

/* One meshblock (Parthenon MeshBlock + LogicalLocation + RegionSize). */
typedef struct jb_block {
  int32_t gid;        /* global block id, blocks are listed in gid order                */
  int32_t level;      /* refinement level above the root grid                           */
  int64_t lx[3];      /* logical location at `level`                                    */
  double xmin[3];     /* block bounds (inactive dims: the mesh bounds)                  */
  double xmax[3];
  int32_t rank;       /* owning rank                                                     */
  int32_t reserved;
} jb_block;

/* The whole mesh as seen by every rank. */
typedef struct jb_mesh {
  int32_t ndim;
  int32_t nxb[3];         /* cells per block per dim (1 in inactive dims)               */
  int32_t nrb[3];         /* root-grid blocks per dim                                    */
  int32_t max_level;      /* deepest level present                                       */
  int32_t nghost;         /* only used to report ghost-inclusive ijk (Parthenon: 2)     */
  double xmin[3], xmax[3];
  int32_t mesh_bc[6];     /* ix1, ox1, ix2, ox2, ix3, ox3 : JB_BC_PERIODIC / OUTFLOW    */
  int32_t swarm_bc[6];    /* same order: JB_BC_*                                         */
  int32_t nblocks;        /* == Mesh::nbtotal                                            */
  const jb_block *blocks; /* [nblocks]                                                   */
} jb_mesh;

/* Per-cycle bookkeeping returned by jb_radiation_step (no reference counterpart; the
 * energies are what the 1e-12 conservation gate is evaluated on). */
typedef struct jb_cycle_stats {
  int64_t n_start;        /* active particles at cycle start                             */
  int64_t n_sourced;      /* emitted this cycle                                          */
  int64_t n_absorbed;
  int64_t n_escaped;      /* removed by an outflow swarm boundary                        */
  int64_t n_end;          /* active (census) particles at cycle end                      */
  int64_t n_sent;         /* records sent to other ranks                                 */
  int64_t n_received;
  int64_t n_block_crossings; /* meshblock changes (on- or off-rank)                      */
  int64_t n_segments;     /* iterations of the while(t < t_end) loop                     */
  int64_t n_scatters;
  int64_t n_ddmc_segments;
  int32_t n_iterations;   /* transport/exchange iterations                               */
  int32_t n_bface_resampled; /* photons SampleDDMCBlockFace spread over a fine face (SampleFace2D/3D) */
  double e_start;         /* sum of weights at cycle start                               */
  double e_sourced;
  double e_absorbed;
  double e_escaped;
  double e_end;
  double ms_transport;    /* device time in the transport kernels (CUDA events)          */
  double ms_total;        /* device time of the whole step                               */
} jb_cycle_stats;

/* Which cell field jb_get_field / jb_set_field addresses. */
enum {
  JB_FIELD_DENSITY = 0,          /* field.material.density  (mcblock_variables.hpp)      */
  JB_FIELD_SIE = 1,              /* field.material.sie                                   */
  JB_FIELD_INTERNAL_ENERGY = 2,  /* field.material.internal_energy (= update_energy)     */
  JB_FIELD_ENERGY_TALLY = 3,     /* field.jaybenne.energy_tally (jaybenne_variables.hpp) */
  JB_FIELD_FLECK_FACTOR = 4,
  JB_FIELD_SOURCE_EW_PER_CELL = 5,
  JB_FIELD_SOURCE_NUM_PER_CELL = 6,
  JB_FIELD_ENERGY_DELTA = 7,
  JB_FIELD_DDMC_FACE_PROB_X1 = 8, /* (nx1+1)*nx2*nx3 values                              */
  JB_FIELD_DDMC_FACE_PROB_X2 = 9, /* nx1*(nx2+1)*nx3                                     */
  JB_FIELD_DDMC_FACE_PROB_X3 = 10 /* nx1*nx2*(nx3+1)                                     */
};

/* The `photons` swarm as host SoA (src/jaybenne/jaybenne.cpp:236-245). Any pointer may
 * be NULL to skip that variable. */
typedef struct jb_particles {
  double *x, *y, *z;         /* swarm.x / swarm.y / swarm.z                              */
  double *time, *weight, *energy;
  double *v0, *v1, *v2;      /* particle.photons.v                                       */
  int32_t *i, *j, *k;        /* particle.photons.ijk, ghost-inclusive like Parthenon     */
  int32_t *block;            /* gid of the owning block                                  */
  uint32_t *rng;             /* 4 words per particle: birth index, birth gid, source
                                call, draws consumed                                     */
} jb_particles;

/* One record of the optional per-segment event log (debug / parity only). */
typedef struct jb_event {
  uint32_t id_idx, id_gid, id_call; /* particle identity                                 */
  uint32_t draw;                    /* RNG draws consumed BEFORE this segment            */
  int32_t kind;                     /* JB_EV_* bit set                                   */
  int32_t block;                    /* gid of the block after the segment                */
  int32_t i, j, k;                  /* interior cell index after the segment (may be -1
                                       / nx when the particle sits off-block)            */
  int32_t reserved;
  double t, x, y, z, v0, v1, v2;    /* state after the segment                           */
} jb_event;
enum {
  JB_EV_SCATTER = 1, JB_EV_ABSORB = 2, JB_EV_CENSUS = 4, JB_EV_LEFT_BLOCK = 8,
  JB_EV_DDMC = 16, JB_EV_ALBEDO_REJECT = 32, JB_EV_ESCAPED = 64
};

typedef struct jb_ctx jb_ctx;

/* ---- package construction: jaybenne::Initialize, src/jaybenne/jaybenne.cpp:158-266 ---- */
void jb_config_default(jb_config *cfg);
jb_status jb_create(const jb_config *cfg, jb_ctx **out);
void jb_destroy(jb_ctx *ctx);
const char *jb_last_error(const jb_ctx *ctx);
int jb_abi_version(void);
/* device selection for hosts that do not link the CUDA runtime themselves (one rank per
 * GPU: jb_set_device(LOCAL_RANK) before jb_create). Parthenon/Kokkos does this in
 * Kokkos::initialize (src/mcblock/main.cpp:27-35 via ParthenonManager::ParthenonInitEnv). */
jb_status jb_device_count(int32_t *n_devices);
jb_status jb_set_device(int32_t ordinal);
/* speed_of_light / stefan_boltzmann Params in code units (jaybenne.cpp:181-184) */
jb_status jb_get_constants(const jb_ctx *ctx, double *speed_of_light, double *stefan_boltzmann);

/* ---- mesh + fields: what Parthenon owns for the package (Mesh, MeshBlock, fields) ---- */
jb_status jb_set_mesh(jb_ctx *ctx, const jb_mesh *mesh, int32_t my_rank, int32_t n_ranks);
/* particle pool capacity for this rank (Parthenon grows swarm pools on demand; here the
 * pool is sized once; exceeding it returns JB_ERR_CAPACITY) */
jb_status jb_set_capacity(jb_ctx *ctx, int64_t max_particles);
/* host material fields (JAYBENNE_HOST_DENSITY_VARIABLE / ..._SPECIFIC_INTERNAL_ENERGY /
 * ..._UPDATE_ENERGY, src/jaybenne/jaybenne_config.hpp.in:19-30). `src` may be a host or
 * a device pointer. gid must be a block owned by this rank. */
jb_status jb_set_field(jb_ctx *ctx, int32_t gid, int32_t field, const double *src);
jb_status jb_get_field(jb_ctx *ctx, int32_t gid, int32_t field, double *dst_host);
/* The same for `n` blocks in one call (a MeshData partition: every block of the pack at once,
 * src/jaybenne/jaybenne.cpp:285-300 iterates vmesh.GetNBlocks()): n copies queued on the library's
 * stream, one synchronisation. gids[i] pairs with srcs[i] / dsts[i]. */
jb_status jb_set_fields(jb_ctx *ctx, int32_t field, int32_t n, const int32_t *gids, const double *const *srcs);
jb_status jb_get_fields(jb_ctx *ctx, int32_t field, int32_t n, const int32_t *gids, double *const *dsts_host);
/* Multi-rank DDMC: density / sie of a block owned by ANOTHER rank (the ghost-zone exchange of
 * src/mcblock/mcblock_driver.cpp:68 + jaybenne.cpp:108-110 as one whole-block copy). Needed
 * before jb_update_derived_transport_fields only when the host program moves the data itself;
 * with jb_comm_init the library exchanges the blocks over NCCL. No-op on one rank / without DDMC. */
jb_status jb_set_remote_field(jb_ctx *ctx, int32_t gid, int32_t field, const double *src);
/* number of values of `field` per block */
int64_t jb_field_size(const jb_ctx *ctx, int32_t field);

/* ---- the task-list entry points mcblock / RadiationStep call -------------------------- */
/* InitializeRadiation on every local block, src/jaybenne/jaybenne.cpp:570-578 */
jb_status jb_initialize_radiation(jb_ctx *ctx, int32_t is_thermal);
/* RadiationStep(...).Execute(), src/jaybenne/jaybenne.cpp:68-151. stats may be NULL. */
jb_status jb_radiation_step(jb_ctx *ctx, double t_start, double dt, jb_cycle_stats *stats);
/* the individual tasks, src/jaybenne/jaybenne.hpp:59-69 */
jb_status jb_update_derived_transport_fields(jb_ctx *ctx, double dt);      /* jaybenne.cpp:285 */
jb_status jb_source_photons(jb_ctx *ctx, int32_t source_type, double t_start, double dt); /* sourcing.cpp:25 */
jb_status jb_transport_photons(jb_ctx *ctx, double t_start, double dt);    /* transport.cpp:28 / transport_ddmc.cpp:28 by use_ddmc */
jb_status jb_swarm_exchange(jb_ctx *ctx);  /* MeshResetCommunication+MeshSend+MeshReceive, jaybenne.cpp:26-61 */
jb_status jb_sample_ddmc_block_face(jb_ctx *ctx);                          /* sample_ddmc_bface.cpp:81 */
jb_status jb_check_completion(jb_ctx *ctx, double t_end, int64_t *num_unfinished); /* transport.cpp:187 */
jb_status jb_evaluate_radiation_energy(jb_ctx *ctx);                       /* jaybenne.cpp:514 */
/* The same tally for callers that cannot allocate (CUDA graph capture, tight serving loops, harnesses that forbid
 * cudaMalloc). NO counterpart in the reference. The call and everything
 * below it allocates nothing: k_tally's scratch (per-cell accumulation buffer, census-energy bins, counters) is one
 * file-scope device array of JB_NOALLOC_MAX_CELLS doubles + 8.3 KiB (2 MiB, held by every GPU that loads the library
 * until exit). tally_out: tally_len >= local cells doubles, device or host memory, [local block][k][j][i];
 * n_census / e_census (host, may be NULL): photons in the pool and the energy they carry. The context's own
 * energy_tally field is not touched. More than JB_NOALLOC_MAX_CELLS local cells -> JB_ERR_CAPACITY. All calls share
 * the array: calls that overlap on the GPU (two streams or contexts on one device) race. */
#define JB_NOALLOC_MAX_CELLS 262144
jb_status jb_evaluate_radiation_energy_noalloc(jb_ctx *ctx, double *tally_out, int64_t tally_len, int64_t *n_census,
                                               double *e_census);
jb_status jb_update_fluid(jb_ctx *ctx);                                    /* jaybenne.cpp:583 */
/* mcblock's FillDerived after the step: sie = u / rho (mcblock.cpp:208-232) */
jb_status jb_update_derived_material(jb_ctx *ctx);
/* mcblock's PostInitialization: u = rho * sie (mcblock.cpp:237-262) */
jb_status jb_post_initialization(jb_ctx *ctx);

/* ---- swarm access --------------------------------------------------------------------- */
jb_status jb_num_particles(jb_ctx *ctx, int64_t *n_local);
/* copies up to `capacity` particles to host arrays; *n = number written */
jb_status jb_get_particles(jb_ctx *ctx, const jb_particles *dst, int64_t capacity, int64_t *n);
/* replaces the local swarm with `n` host particles (Parthenon adapter: swarm -> ctx) */
jb_status jb_set_particles(jb_ctx *ctx, const jb_particles *src, int64_t n);
/* restart the sourcing-call counter that keys the per-photon Philox streams (birth index, birth gid,
 * source call): a run repeated on the same inputs after this call reproduces the same photons. The
 * reference's pool generator has no counterpart (its streams depend on the thread schedule). */
jb_status jb_reset_streams(jb_ctx *ctx);
/* how the last multi-rank jb_radiation_step moved photons between ranks: 0 = single rank / not yet run,
 * 1 = count all-gather + grouped ncclSend / ncclRecv per iteration, 2 = direct stores into the peers'
 * IPC-mapped inboxes over NVLink with the loop state on the device (needs jb_set_capacity; JB_P2P=0 disables) */
int32_t jb_exchange_mode(const jb_ctx *ctx);
/* Device-timed phases of the last jb_radiation_step, in stream order (CUDA events on the context's stream; no
 * reference counterpart — the reference's timing is Kokkos profiling regions, jaybenne.cpp:87,115,127,145).
 * Entry q: the phase labels[q] (JB_TL_*) took ms[q] milliseconds; *n = number of phases (may exceed cap). A
 * multi-rank step has one JB_TL_TRANSPORT + JB_TL_EXCHANGE pair per transport iteration; the exchange phase
 * includes the time this rank waited for its peers. */
enum {
  JB_TL_BEGIN = 0,      /* (never reported: opens the first phase) */
  JB_TL_FIELDS = 1,     /* census at cycle start, Fleck factor, DDMC face probabilities, packed cell records */
  JB_TL_SOURCE = 2,     /* SourcePhotons (emission) */
  JB_TL_TRANSPORT = 3,  /* one launch of the transport kernel */
  JB_TL_EXCHANGE = 4,   /* swarm exchange of one iteration: counts, send / receive or peer stores, unpack */
  JB_TL_STATS = 5,      /* counters and energy sums folded and read back */
  JB_TL_SQUEEZE = 6,    /* pool squeeze (only when occupancy dropped below the floor) */
  JB_TL_TALLY = 7,      /* EvaluateRadiationEnergy */
  JB_TL_END = 8         /* UpdateFluid (+ cell sort when sort_interval asks for it) */
};
jb_status jb_step_timeline(const jb_ctx *ctx, int32_t *labels, float *ms, int32_t cap, int32_t *n);
/* change jb_config::sort_interval of a live context (0 = never sort the pool by (meshblock, cell)) */
jb_status jb_set_sort_interval(jb_ctx *ctx, int32_t every_n_cycles);
/* cumulative stats since the last jb_radiation_step / jb_stats_begin began */
jb_status jb_get_stats(jb_ctx *ctx, jb_cycle_stats *stats);
/* zero the counters and record n_start / e_start: for hosts that drive the individual
 * tasks themselves (jb_radiation_step does this internally) */
jb_status jb_stats_begin(jb_ctx *ctx);
/* device timer: CUDA events recorded on the library's stream (all kernels of a context
 * run on one private stream); stop waits for the stream and returns the elapsed ms */
jb_status jb_timer_start(jb_ctx *ctx);
jb_status jb_timer_stop(jb_ctx *ctx, double *ms);
/* CUDA kernels launched by this context since jb_create (diagnostics / bench.py) */
int64_t jb_kernel_launches(const jb_ctx *ctx);
/* device self-test: jbm::div_normal (the range-test-free division of the kernels, csrc/jb_math.h) against the
 * compiler's IEEE `/` on n pseudo-random operand pairs spanning the ranges the kernels use it on (exponents within
 * +-300, numerators also 0); *mismatches = pairs whose quotients differ in any bit. tests/test_parity_gpu.py. */
jb_status jb_selftest_division(jb_ctx *ctx, int64_t n, int64_t *mismatches);
/* measurement helper (bench.py): sustained FP64 FMA rate of this GPU in TFLOP/s (2 flops per
 * DFMA), from a register-resident chain kernel timed with CUDA events on the context's
 * stream. It is the denominator of the FP64-pipe fraction reported beside the HBM roofline. */
jb_status jb_measure_fp64_peak(jb_ctx *ctx, double *tflops);

/* ---- event log (parity / debugging) --------------------------------------------------- */
jb_status jb_event_log_enable(jb_ctx *ctx, int64_t capacity); /* 0 disables */
jb_status jb_event_log_read(jb_ctx *ctx, jb_event *dst_host, int64_t capacity, int64_t *n);

/* ---- multi-rank exchange (replaces Parthenon's MPI swarm comms) ----------------------- *
 * Native path: NCCL inside the library. Rank 0 calls jb_comm_unique_id and broadcasts the
 * 128 bytes by any means (torch.distributed, MPI, a file); every rank then calls
 * jb_comm_init. After that jb_swarm_exchange / jb_check_completion / jb_radiation_step
 * are collective over all ranks. */
jb_status jb_comm_unique_id(void *id128);
jb_status jb_comm_init(jb_ctx *ctx, const void *id128, int32_t rank, int32_t n_ranks);
/* Host-staged path (tests, non-NCCL transports): records leaving for `peer` as a packed
 * host buffer of jb_exchange_record_bytes() each; give() appends received records. */
int64_t jb_exchange_record_bytes(void);
jb_status jb_exchange_count(jb_ctx *ctx, int32_t peer, int64_t *n_records);
jb_status jb_exchange_take(jb_ctx *ctx, int32_t peer, void *dst_host, int64_t capacity, int64_t *n_records);
jb_status jb_exchange_give(jb_ctx *ctx, const void *src_host, int64_t n_records);
/* device-resident variant for NCCL driven by the host program (torch.distributed):
 * pointers stay valid until the next jb_transport_photons / jb_exchange_* call */
jb_status jb_exchange_take_device(jb_ctx *ctx, int32_t peer, void **dev_ptr, int64_t *n_records);
jb_status jb_exchange_give_device(jb_ctx *ctx, const void *dev_ptr, int64_t n_records);
jb_status jb_exchange_clear(jb_ctx *ctx);

/* ---- host helpers (no GPU needed): deck parsing and mesh construction ------------------ *
 * ParameterInput-style reader for Parthenon input decks (the inputs/ decks). */
typedef struct jb_deck jb_deck;
jb_status jb_deck_load(const char *path, jb_deck **out);
jb_status jb_deck_parse(const char *text, jb_deck **out);
void jb_deck_free(jb_deck *deck);
/* override or add "block/key=value" (tst/regression_test.py modified_inputs) */
jb_status jb_deck_set(jb_deck *deck, const char *block, const char *key, const char *value);
int32_t jb_deck_has(const jb_deck *deck, const char *block, const char *key);
/* returns NULL if absent; pointer valid until the deck is freed / modified */
const char *jb_deck_get(const jb_deck *deck, const char *block, const char *key);
/* fills cfg from <jaybenne> and <mcblock> exactly like jaybenne::Initialize +
 * mcblock::Initialize; errors mirror their PARTHENON_REQUIRE/FAIL cases */
jb_status jb_deck_config(const jb_deck *deck, jb_config *cfg, char *err, int32_t errlen);
/* builds the block table from <parthenon/mesh>, <parthenon/meshblock>,
 * <parthenon/static_refinementN>, <parthenon/swarm>; blocks are malloc'ed into
 * mesh->blocks (free with jb_mesh_free). Blocks are dealt to ranks in contiguous gid
 * (Z-order) chunks. */
jb_status jb_deck_mesh(const jb_deck *deck, int32_t n_ranks, jb_mesh *mesh, char *err, int32_t errlen);
void jb_mesh_free(jb_mesh *mesh);

#ifdef __cplusplus
}
#endif
#endif /* JAYBENNE_B200_H_ */
