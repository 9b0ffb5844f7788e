// jb_lib.cu — the context behind the C ABI of include/jaybenne_b200.h: owns all device
// memory, orders the kernels of one radiation step the way RadiationStep's task list does
// (src/jaybenne/jaybenne.cpp:68-151) and moves photons between ranks (NCCL), replacing
// Parthenon's swarm Send/Receive.
//
// There is no CPU path in this file: without a CUDA device every compute entry point
// returns JB_ERR_NO_DEVICE.
#include <dlfcn.h>
#include <nccl.h>
#include <nvtx3/nvToolsExt.h>  // header-only NVTX v3: ranges cost nothing unless a tool is attached

#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <limits>
#include <string>
#include <vector>

#include "jb_kernels.cuh"

using namespace jbd;

// ---- NCCL, bound at run time so the library loads on machines without it -----------------
struct NcclApi {
  void *handle = nullptr;
  ncclResult_t (*GetUniqueId)(ncclUniqueId *) = nullptr;
  ncclResult_t (*CommInitRank)(ncclComm_t *, int, ncclUniqueId, int) = nullptr;
  ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
  ncclResult_t (*AllGather)(const void *, void *, size_t, ncclDataType_t, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*AllReduce)(const void *, void *, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t,
                            cudaStream_t) = nullptr;
  ncclResult_t (*Send)(const void *, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*Recv)(void *, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*GroupStart)() = nullptr;
  ncclResult_t (*GroupEnd)() = nullptr;
  const char *(*GetErrorString)(ncclResult_t) = nullptr;
  bool ok = false;
};
static NcclApi &nccl_api() {
  static NcclApi api;
  if (api.handle) return api;
  // a libnccl already in the process (e.g. the one torch ships) is reused by soname
  api.handle = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
  if (!api.handle) api.handle = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
  if (!api.handle) return api;
#define JB_SYM(name) *(void **)(&api.name) = dlsym(api.handle, "nccl" #name)
  JB_SYM(GetUniqueId); JB_SYM(CommInitRank); JB_SYM(CommDestroy); JB_SYM(AllGather); JB_SYM(AllReduce);
  JB_SYM(Send); JB_SYM(Recv); JB_SYM(GroupStart); JB_SYM(GroupEnd); JB_SYM(GetErrorString);
#undef JB_SYM
  api.ok = api.GetUniqueId && api.CommInitRank && api.CommDestroy && api.AllGather && api.AllReduce && api.Send &&
           api.Recv && api.GroupStart && api.GroupEnd;
  return api;
}

// the reference's Kokkos::Profiling regions (jaybenne.cpp:87,115,127,145) as NVTX ranges with the same names
struct NvtxRange {
  explicit NvtxRange(const char *name) { nvtxRangePushA(name); }
  ~NvtxRange() { nvtxRangePop(); }
  NvtxRange(const NvtxRange &) = delete;
  NvtxRange &operator=(const NvtxRange &) = delete;
};

struct jb_ctx {
  jb_config cfg;
  Closure cl;
  jb_mesh mesh;
  std::vector<jb_block> all_blocks;
  std::vector<int> gid_to_local, local_gid;
  int nlat[3] = {1, 1, 1}, ng[3] = {0, 0, 0};
  int my_rank = 0, n_ranks = 1;
  bool mesh_set = false;
  int device = -1;
  cudaStream_t stream = nullptr;
  cudaEvent_t ev[4] = {nullptr, nullptr, nullptr, nullptr};
  cudaEvent_t tev[2] = {nullptr, nullptr};

  // device tables
  DBlock *d_blocks = nullptr;
  int *d_lattice = nullptr, *d_local_gid = nullptr, *d_nbr = nullptr;
  BlkGeo *d_geo = nullptr;
  Fields f{};
  long long ncells_local = 0;
  Swarm s{};
  long long capacity = 0, n = 0;
  long long requested_capacity = 0;
  // scratch
  double *d_tmp = nullptr;            // one SoA-array-sized scratch: rotates with the arrays in the cell sort
  unsigned int *d_perm = nullptr;     // sort permutation
  // order-preserving squeeze (k_squeeze_*): live bits, per-chunk counts / offsets, scan scratch
  unsigned *d_sq_mask = nullptr;
  int *d_sq_cnt = nullptr;
  long long *d_sq_prefix = nullptr, *d_sq_scan = nullptr;
  long long sq_chunks_cap = 0;
  int squeeze_mode = -1;              // 1 order-preserving (default), 0 tail photons into holes (JB_SQUEEZE=tail)
  int *d_cnt = nullptr;               // per-cell counts
  long long *d_prefix = nullptr;      // per-cell offsets
  long long *d_scan = nullptr, *d_total = nullptr;
  long long *d_holes = nullptr, *d_tails = nullptr;
  long long holes_cap = 0;
  long long n_holes = 0;              // dead slots inside [0, n) not yet squeezed out
  // census totals left behind by the last tally (k_tally counts photons and sums their energy on the way);
  // valid until the swarm changes (sourcing, transport, receive, set_particles)
  bool census_valid = false;
  long long census_n = 0;
  double census_e = 0.0;
  unsigned long long *d_counters = nullptr;  // [8] misc device counters
  double *d_esum = nullptr;
  double *d_ebins = nullptr;          // [EB_KINDS][EBINS] energy-sum bins (jb_kernels.cuh)
  bool props_dirty = true;            // CellProps must be rebuilt before the next transport
  DStats *d_stats = nullptr;
  DStats *h_stats = nullptr;          // pinned
  unsigned long long *h_counters = nullptr;  // pinned [8 + n_ranks*n_ranks]
  // exchange
  Outbox out{};
  ExRecord *d_inbox = nullptr;
  long long inbox_cap = 0;
  unsigned long long *d_count_matrix = nullptr;
  ncclComm_t comm = nullptr;
  // direct peer exchange over NVLink (p2p_setup): IPC-mapped inboxes + mailboxes, device-resident loop state
  bool p2p_tried = false, p2p_on = false;
  int exchange_mode = 0;               // what the last multi-rank step used (jb_exchange_mode)
  P2PMail *d_mail = nullptr;
  ExRecord *d_p2p_inbox[2] = {nullptr, nullptr};
  long long p2p_inbox_cap = 0;
  P2PState *d_p2p_state = nullptr;
  P2PState *h_p2p_state = nullptr;     // pinned
  P2PPeers *d_peers = nullptr;
  std::vector<void *> ipc_opened;      // peer mappings to close
  cudaEvent_t it_ev[2 * 16] = {};      // transport launch brackets of the queued iterations
  // phase marks of the last jb_radiation_step (jb_step_timeline): event i closes the phase labelled tl_label[i]
  static constexpr int TL_CAP = 96;
  cudaEvent_t tl_ev[TL_CAP] = {};
  int tl_label[TL_CAP] = {};
  float tl_ms[TL_CAP] = {};
  int tl_n = 0, tl_done = 0;
  // event log
  jb_event *d_events = nullptr;
  long long event_cap = 0;

  uint32_t source_call = 0;
  int cycles_since_sort = 0;
  jb_cycle_stats stats;
  long long launches = 0;             // kernels launched since creation (bench: gpu_launches)
  std::string err;
};

static jb_status fail(jb_ctx *c, jb_status s, const std::string &m) {
  if (c) c->err = m;
  return s;
}
#define JB_CUDA(c, call)                                                                         \
  do {                                                                                           \
    cudaError_t e_ = (call);                                                                     \
    if (e_ != cudaSuccess)                                                                       \
      return fail((c), JB_ERR_CUDA, std::string(#call) + ": " + cudaGetErrorString(e_));         \
  } while (0)

template <typename T>
static cudaError_t dalloc(T **p, long long nelem) {
  *p = nullptr;
  if (nelem <= 0) nelem = 1;
  return cudaMalloc((void **)p, sizeof(T) * (size_t)nelem);
}
static void dfree(void *p) {
  if (p) cudaFree(p);
}

// CGS constants as singularity-opac derives them (constants/constants.hpp:41-178)
namespace cgs {
static constexpr double length = 1.e2, mass = 1.e3, time = 1., temperature = 1.;
static constexpr double force = mass * length / (time * time);
static constexpr double energy = force * length;
static constexpr double h = 6.62606957e-34 * energy * time;
static constexpr double kb = 1.380648800e-23 * energy / temperature;
static constexpr double c = 2.99792458e8 * length / time;
static constexpr double sb = 2.0 * M_PI * M_PI * M_PI * M_PI * M_PI * kb * kb * kb * kb / (15.0 * h * h * h * c * c);
static constexpr double me = 9.10938291e-31 * mass;
static constexpr double mp = 1.672621777e-27 * mass;
static constexpr double alpha = 7.2973525698e-3;
static constexpr double hbar = h / (2.0 * M_PI);
static constexpr double qe = 1.602176565e-19 * 2.997924580e9;
}  // namespace cgs

// Flatten the closures mcblock builds (mcblock.cpp:78-143) into numbers; the derived
// constants follow non_cgs_photons.hpp:33-46 and constants.hpp:191-207.
static void make_closure(const jb_config &c, Closure &cl) {
  memset(&cl, 0, sizeof(cl));
  cl.opacity_model = c.opacity_model;
  cl.scattering_model = c.scattering_model;
  cl.gray_mode = c.gray_mode;
  cl.kappa = c.opacity_kappa;
  cl.sigma = c.scattering_sigma;
  cl.apm = c.apm;
  cl.cv = c.eos_cv;
  cl.time_unit = c.time_scale;
  cl.length_unit = c.length_scale;
  cl.temp_unit = c.temperature_scale;
  const double mass_unit = c.mass_scale;
  cl.rho_unit = mass_unit / (cl.length_unit * cl.length_unit * cl.length_unit);
  cl.freq_unit = 1. / cl.time_unit;
  cl.inv_emiss_unit = cl.length_unit * cl.time_unit * cl.time_unit / mass_unit;
  cl.c_code = cgs::c * cl.time_unit / cl.length_unit;
  cl.sb_code = cgs::sb * std::pow(cl.time_unit, 3) / mass_unit * std::pow(cl.temp_unit, 4);
  cl.bt_num = 8. * std::pow(M_PI, 5) * std::pow(cgs::kb, 4);
  cl.bt_den = 15. * std::pow(cgs::c, 2) * std::pow(cgs::h, 3);
  cl.sigma_thomson = 8. * M_PI / 3. * std::pow(cgs::alpha * cgs::hbar / (cgs::me * cgs::c), 2);
  cl.ep_mmw = (cgs::mp + cgs::me) / 2.;
  cl.ep_prefac = 8. * std::pow(cgs::qe, 6) / (3. * std::pow(cgs::me * cgs::c * cgs::c, 2)) * std::sqrt(2. * M_PI / 3.);
  cl.h = cgs::h;
  cl.kb = cgs::kb;
  cl.c = cgs::c;
  cl.me = cgs::me;
  cl.two_h_over_c2 = 2. * cgs::h / (cgs::c * cgs::c);
  cl.pi4 = std::pow(M_PI, 4.0);
}

static StepParams make_params(jb_ctx *c) {
  StepParams P;
  memset(&P, 0, sizeof(P));
  DMesh &m = P.mesh;
  m.ndim = c->mesh.ndim;
  for (int d = 0; d < 3; ++d) {
    m.nxb[d] = c->mesh.nxb[d];
    m.ng[d] = c->ng[d];
    m.nlat[d] = c->nlat[d];
    m.nrb[d] = c->mesh.nrb[d];
    m.xmin[d] = c->mesh.xmin[d];
    m.xmax[d] = c->mesh.xmax[d];
  }
  for (int q = 0; q < 6; ++q) {
    m.swarm_bc[q] = c->mesh.swarm_bc[q];
    m.mesh_bc[q] = c->mesh.mesh_bc[q];
  }
  m.max_level = c->mesh.max_level;
  m.nblocks = c->mesh.nblocks;
  m.nlocal = (int)c->local_gid.size();
  m.my_rank = c->my_rank;
  m.n_ranks = c->n_ranks;
  m.ncell = c->mesh.nxb[0] * c->mesh.nxb[1] * c->mesh.nxb[2];
  m.blocks = c->d_blocks;
  m.lattice = c->d_lattice;
  m.local_gid = c->d_local_gid;
  m.geo = c->d_geo;
  m.nbr = c->d_nbr;
  P.f = c->f;
  P.s = c->s;
  P.cl = c->cl;
  P.out = c->out;
  P.log.ev = c->d_events;
  P.log.cap = c->event_cap;
  P.log.count = c->d_stats ? &c->d_stats->n_events : nullptr;
  P.stats = c->d_stats;
  P.ebins = c->d_ebins;
  P.work = c->d_counters + 6;
  P.key0 = (uint32_t)((uint64_t)c->cfg.seed);
  P.key1_particle = (uint32_t)((uint64_t)c->cfg.seed >> 32) ^ STREAM_PARTICLE;
  P.key1_cell = (uint32_t)((uint64_t)c->cfg.seed >> 32) ^ STREAM_CELL;
  P.rk = make_pkeys(P.key0, P.key1_particle);
  P.use_ddmc = c->cfg.use_ddmc;
  P.tau_ddmc = c->cfg.tau_ddmc;
  return P;
}

// the 12 SoA arrays, all 8 bytes per slot
static void swarm_arrays(Swarm &s, void **out[12]) {
  out[0] = (void **)&s.x; out[1] = (void **)&s.y; out[2] = (void **)&s.z; out[3] = (void **)&s.t;
  out[4] = (void **)&s.w; out[5] = (void **)&s.e; out[6] = (void **)&s.vx; out[7] = (void **)&s.vy;
  out[8] = (void **)&s.vz; out[9] = (void **)&s.loc; out[10] = (void **)&s.ida; out[11] = (void **)&s.idb;
}

// jb_step_timeline: a mark closes the phase `label` (JB_TL_*) on the stream; durations are read after the step
static void tl_mark(jb_ctx *c, int label) {
  if (c->tl_n >= jb_ctx::TL_CAP) return;
  cudaEvent_t &e = c->tl_ev[c->tl_n];
  if (!e && cudaEventCreate(&e) != cudaSuccess) return;
  cudaEventRecord(e, c->stream);
  c->tl_label[c->tl_n++] = label;
}
static void tl_finish(jb_ctx *c) {  // after the stream has been synchronised
  c->tl_done = 0;
  for (int q = 1; q < c->tl_n; ++q) {
    float ms = 0.f;
    if (cudaEventElapsedTime(&ms, c->tl_ev[q - 1], c->tl_ev[q]) != cudaSuccess) ms = -1.f;
    c->tl_label[c->tl_done] = c->tl_label[q];
    c->tl_ms[c->tl_done++] = ms;
  }
  cudaGetLastError();
}

static void free_swarm(jb_ctx *c) {
  void **arr[12];
  swarm_arrays(c->s, arr);
  for (auto a : arr) dfree(*a);
  dfree(c->d_tmp); dfree(c->d_perm);
  dfree(c->d_sq_mask); dfree(c->d_sq_cnt); dfree(c->d_sq_prefix); dfree(c->d_sq_scan);
  c->d_sq_mask = nullptr; c->d_sq_cnt = nullptr; c->d_sq_prefix = c->d_sq_scan = nullptr;
  c->sq_chunks_cap = 0;
  dfree(c->out.rec); dfree(c->out.count); dfree(c->d_inbox);
  memset(&c->s, 0, sizeof(c->s));
  memset(&c->out, 0, sizeof(c->out));
  c->d_tmp = nullptr;
  c->d_perm = nullptr;
  c->d_inbox = nullptr;
  c->capacity = 0;
  c->inbox_cap = 0;
}

// (re)allocate the photon pool; existing photons are preserved
static jb_status ensure_capacity(jb_ctx *c, long long need) {
  if (need <= c->capacity) return JB_OK;
  long long cap = std::max(need, c->requested_capacity);
  if (c->requested_capacity <= 0) cap = std::max<long long>(need + need / 4, 1024);
  Swarm ns{};
  void **na[12], **oa[12];
  swarm_arrays(ns, na);
  swarm_arrays(c->s, oa);
  for (int q = 0; q < 12; ++q) {
    JB_CUDA(c, dalloc((unsigned long long **)na[q], cap));
    if (c->n > 0)
      JB_CUDA(c, cudaMemcpyAsync(*na[q], *oa[q], (size_t)c->n * 8, cudaMemcpyDeviceToDevice, c->stream));
  }
  JB_CUDA(c, cudaStreamSynchronize(c->stream));
  const long long keep_n = c->n;
  free_swarm(c);
  c->s = ns;
  c->n = keep_n;
  c->capacity = cap;
  if (c->n_ranks > 1) {
    c->out.cap = std::max<long long>(cap / (2 * c->n_ranks), 4096);
    // no capacity requested: a rank that starts (nearly) empty must still be able to take what the
    // others send it, so size the exchange buffers from the global photon count (bounded)
    if (c->requested_capacity <= 0)
      c->out.cap = std::max<long long>(c->out.cap, std::min<long long>(2 * c->cfg.num_particles + 4096, 1LL << 21));
    JB_CUDA(c, dalloc(&c->out.rec, c->out.cap * c->n_ranks));
    c->inbox_cap = c->out.cap * c->n_ranks;
    JB_CUDA(c, dalloc(&c->d_inbox, c->inbox_cap));
  }
  JB_CUDA(c, dalloc(&c->out.count, std::max(c->n_ranks, 1)));
  JB_CUDA(c, cudaMemset(c->out.count, 0, sizeof(unsigned long long) * std::max(c->n_ranks, 1)));
  // A pool the caller sized once (jb_set_capacity) also gets the scratch array of the order-preserving squeeze
  // now: an 8 GB cudaMalloc in the middle of the first cycle that squeezes took 25-120 ms depending on the box
  // (profiles/r02_v13_bench_default_cells256.json, variants.continuing, cycle 2). Failure is not an error here:
  // compact() tries again and falls back to the in-place squeeze.
  if (c->requested_capacity > 0 && !c->d_tmp) {
    const char *v = getenv("JB_SQUEEZE");
    if (!(v && strcmp(v, "tail") == 0) && dalloc(&c->d_tmp, c->capacity) != cudaSuccess) {
      cudaGetLastError();
      c->d_tmp = nullptr;
    }
  }
  return JB_OK;
}

// grow only the 12 swarm arrays (Parthenon grows swarm pools on demand); the outbox / inbox keep
// their size and content, so this is safe between packing and the NCCL send/recv
static jb_status grow_pool_only(jb_ctx *c, long long need) {
  if (need <= c->capacity) return JB_OK;
  const long long cap = need + need / 4 + 1024;
  void **arr[12];
  swarm_arrays(c->s, arr);
  for (int q = 0; q < 12; ++q) {
    unsigned long long *fresh = nullptr;
    JB_CUDA(c, dalloc(&fresh, cap));
    if (c->n > 0) JB_CUDA(c, cudaMemcpyAsync(fresh, *arr[q], (size_t)c->n * 8, cudaMemcpyDeviceToDevice, c->stream));
    JB_CUDA(c, cudaStreamSynchronize(c->stream));
    dfree(*arr[q]);
    *arr[q] = (void *)fresh;
  }
  dfree(c->d_tmp); dfree(c->d_perm);  // sort scratch is re-made at the new size when next needed
  c->d_tmp = nullptr;
  c->d_perm = nullptr;
  c->capacity = cap;
  return JB_OK;
}

// =======================================================================================
// C ABI
// =======================================================================================
extern "C" {

jb_status jb_create(const jb_config *cfg, jb_ctx **out) {
  if (!cfg || !out) return JB_ERR_INVALID;
  *out = nullptr;
  // jaybenne::Initialize input checks (jaybenne.cpp:171-172)
  if (!(cfg->min_swarm_occupancy >= 0 && cfg->min_swarm_occupancy < 1.0)) return JB_ERR_INVALID;
  if (cfg->rng_mode != JB_RNG_PHILOX) return JB_ERR_UNSUPPORTED;  // the pool generator is oracle-only
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev <= 0) {
    cudaGetLastError();
    return JB_ERR_NO_DEVICE;
  }
  jb_ctx *c = new jb_ctx();
  c->cfg = *cfg;
  make_closure(*cfg, c->cl);
  memset(&c->stats, 0, sizeof(c->stats));
  memset(&c->mesh, 0, sizeof(c->mesh));
  if (cudaGetDevice(&c->device) != cudaSuccess || cudaStreamCreate(&c->stream) != cudaSuccess) {
    delete c;
    return JB_ERR_CUDA;
  }
  for (auto &e : c->ev) cudaEventCreate(&e);
  bool ok = dalloc(&c->d_stats, 1) == cudaSuccess && dalloc(&c->d_counters, 8) == cudaSuccess &&
            dalloc(&c->d_esum, 2) == cudaSuccess && dalloc(&c->d_total, 2) == cudaSuccess &&
            dalloc(&c->d_ebins, (long long)EB_KINDS * EBINS) == cudaSuccess &&
            cudaMallocHost((void **)&c->h_stats, sizeof(DStats)) == cudaSuccess &&
            cudaMallocHost((void **)&c->h_counters, sizeof(unsigned long long) * (16 + 64 * 67)) == cudaSuccess;
  if (!ok) {
    delete c;
    return JB_ERR_CUDA;
  }
  cudaMemset(c->d_stats, 0, sizeof(DStats));
  cudaMemset(c->d_ebins, 0, sizeof(double) * EB_KINDS * EBINS);
  // kernel tuning knobs (diagnostics / A-B measurements; defaults are the shipped configuration)
  {
    TransportTuning t = {2, 8, 4, 0};
    if (const char *v = getenv("JB_TRANSPORT_VARIANT")) t.variant = atoi(v);
    if (const char *v = getenv("JB_IMC_CTAS_PER_SM")) t.ctas_per_sm = atoi(v);
    if (const char *v = getenv("JB_IMC_REFILL_MIN")) t.refill_min = std::max(1, std::min(32, atoi(v)));
    if (const char *v = getenv("JB_IMC_VALIDATE")) t.validate = atoi(v);
    set_transport_tuning(t);
    {
      int ps = 64, prf = 24, prl = 32;
      if (const char *v = getenv("JB_POOL_SLOTS")) ps = atoi(v);
      if (const char *v = getenv("JB_POOL_REFILL_MIN")) prf = atoi(v);
      if (const char *v = getenv("JB_POOL_READY_LOW")) prl = atoi(v);
      set_pool_tuning(ps, prf, prl);
      if (const char *v = getenv("JB_POOL_PREFETCH")) set_pool_prefetch(atoi(v));
      if (const char *v = getenv("JB_POOL_FIXED_CHUNK")) set_pool_fixed_chunk(atoi(v));
    }
  }
  *out = c;
  return JB_OK;
}

void jb_destroy(jb_ctx *c) {
  if (!c) return;
  cudaSetDevice(c->device);
  cudaDeviceSynchronize();
  for (void *q : c->ipc_opened) cudaIpcCloseMemHandle(q);
  dfree(c->d_mail); dfree(c->d_p2p_inbox[0]); dfree(c->d_p2p_inbox[1]); dfree(c->d_p2p_state); dfree(c->d_peers);
  if (c->h_p2p_state) cudaFreeHost(c->h_p2p_state);
  for (auto &e : c->tl_ev)
    if (e) cudaEventDestroy(e);
  for (auto &e : c->it_ev)
    if (e) cudaEventDestroy(e);
  if (c->comm && nccl_api().ok) nccl_api().CommDestroy(c->comm);
  free_swarm(c);
  dfree(c->d_blocks); dfree(c->d_lattice); dfree(c->d_local_gid); dfree(c->d_nbr); dfree(c->d_geo);
  dfree(c->f.rho); dfree(c->f.sie); dfree(c->f.u); dfree(c->f.tally); dfree(c->f.fleck);
  dfree(c->f.src_ew); dfree(c->f.src_num); dfree(c->f.edelta); dfree(c->f.g_rho); dfree(c->f.g_sie);
  for (int d = 0; d < 3; ++d) dfree(c->f.fp[d]);
  dfree(c->d_cnt); dfree(c->d_prefix); dfree(c->d_scan); dfree(c->d_total);
  dfree(c->d_holes); dfree(c->d_tails); dfree(c->d_counters); dfree(c->d_esum); dfree(c->d_stats);
  dfree(c->d_count_matrix); dfree(c->d_events); dfree(c->d_ebins); dfree(c->f.props);
  if (c->h_stats) cudaFreeHost(c->h_stats);
  if (c->h_counters) cudaFreeHost(c->h_counters);
  for (auto &e : c->ev)
    if (e) cudaEventDestroy(e);
  for (auto &e : c->tev)
    if (e) cudaEventDestroy(e);
  if (c->stream) cudaStreamDestroy(c->stream);
  delete c;
}

const char *jb_last_error(const jb_ctx *c) { return c ? c->err.c_str() : "null context"; }

jb_status jb_get_constants(const jb_ctx *c, double *sol, double *sb) {
  if (!c) return JB_ERR_INVALID;
  if (sol) *sol = c->cl.c_code;
  if (sb) *sb = c->cl.sb_code;
  return JB_OK;
}

int64_t jb_field_size(const jb_ctx *c, int32_t field) {
  const int64_t n1 = c->mesh.nxb[0], n2 = c->mesh.nxb[1], n3 = c->mesh.nxb[2];
  if (field == JB_FIELD_DDMC_FACE_PROB_X1) return (n1 + 1) * n2 * n3;
  if (field == JB_FIELD_DDMC_FACE_PROB_X2) return n1 * (n2 + 1) * n3;
  if (field == JB_FIELD_DDMC_FACE_PROB_X3) return n1 * n2 * (n3 + 1);
  return n1 * n2 * n3;
}

jb_status jb_set_mesh(jb_ctx *c, const jb_mesh *m, int32_t my_rank, int32_t n_ranks) {
  if (!c) return JB_ERR_INVALID;
  if (!m || m->nblocks <= 0 || m->ndim < 1 || m->ndim > 3) return fail(c, JB_ERR_INVALID, "bad mesh");
  if (n_ranks < 1 || n_ranks > 64 || my_rank < 0 || my_rank >= n_ranks) return fail(c, JB_ERR_INVALID, "bad rank");
  if (c->mesh_set) return fail(c, JB_ERR_STATE, "mesh already set");
  JB_CUDA(c, cudaSetDevice(c->device));
  c->mesh = *m;
  c->all_blocks.assign(m->blocks, m->blocks + m->nblocks);
  c->mesh.blocks = c->all_blocks.data();
  c->my_rank = my_rank;
  c->n_ranks = n_ranks;
  for (int d = 0; d < 3; ++d) {
    c->nlat[d] = d < m->ndim ? (m->nrb[d] << m->max_level) : 1;
    c->ng[d] = (d < m->ndim) ? m->nghost : 0;
  }
  std::vector<int> lattice((size_t)c->nlat[0] * c->nlat[1] * c->nlat[2], -1);
  c->gid_to_local.assign(m->nblocks, -1);
  c->local_gid.clear();
  std::vector<DBlock> hb(m->nblocks);
  for (int g = 0; g < m->nblocks; ++g) {
    const jb_block &b = c->all_blocks[g];
    if (b.gid != g) return fail(c, JB_ERR_INVALID, "blocks must be listed in gid order");
    if (b.level < 0 || b.level > m->max_level) return fail(c, JB_ERR_INVALID, "block level out of range");
    const int s = 1 << (m->max_level - b.level);
    int lo[3] = {0, 0, 0}, hi[3] = {1, 1, 1};
    for (int d = 0; d < m->ndim; ++d) {
      lo[d] = (int)b.lx[d] * s;
      hi[d] = lo[d] + s;
      if (lo[d] < 0 || hi[d] > c->nlat[d]) return fail(c, JB_ERR_INVALID, "block outside the root grid");
    }
    for (int z = lo[2]; z < hi[2]; ++z)
      for (int y = lo[1]; y < hi[1]; ++y)
        for (int x = lo[0]; x < hi[0]; ++x) lattice[((size_t)z * c->nlat[1] + y) * c->nlat[0] + x] = g;
    DBlock &D = hb[g];
    memset(&D, 0, sizeof(D));
    for (int d = 0; d < 3; ++d) {
      D.xmin[d] = b.xmin[d];
      D.xmax[d] = b.xmax[d];
      D.dx[d] = (b.xmax[d] - b.xmin[d]) / m->nxb[d];
      D.xmin0[d] = b.xmin[d] - c->ng[d] * D.dx[d];
      D.lx[d] = (int)b.lx[d];
    }
    D.level = b.level;
    D.rank = b.rank;
    D.gid = g;
    D.local = -1;
    if (b.rank == my_rank) {
      D.local = (int)c->local_gid.size();
      c->gid_to_local[g] = D.local;
      c->local_gid.push_back(g);
    }
  }
  for (int v : lattice)
    if (v < 0) return fail(c, JB_ERR_INVALID, "block table does not tile the domain");
  const long long nc = (long long)m->nxb[0] * m->nxb[1] * m->nxb[2];
  const long long nl = (long long)c->local_gid.size();
  c->ncells_local = nc * nl;
  if (c->ncells_local >= 0x7fffffffLL)
    return fail(c, JB_ERR_CAPACITY, "more than 2^31 cells on one rank (the sourcing kernels index cells with 32 bits)");
  JB_CUDA(c, dalloc(&c->d_blocks, m->nblocks));
  JB_CUDA(c, dalloc(&c->d_lattice, (long long)lattice.size()));
  JB_CUDA(c, dalloc(&c->d_local_gid, nl));
  JB_CUDA(c, cudaMemcpy(c->d_blocks, hb.data(), sizeof(DBlock) * hb.size(), cudaMemcpyHostToDevice));
  JB_CUDA(c, cudaMemcpy(c->d_lattice, lattice.data(), sizeof(int) * lattice.size(), cudaMemcpyHostToDevice));
  if (nl > 0)
    JB_CUDA(c, cudaMemcpy(c->d_local_gid, c->local_gid.data(), sizeof(int) * nl, cudaMemcpyHostToDevice));
  {
    // per-block geometry line and the same-level neighbour table of the persistent IMC kernel
    std::vector<BlkGeo> geo(m->nblocks);
    for (int g = 0; g < m->nblocks; ++g) {
      const DBlock &D = hb[g];
      for (int d = 0; d < 3; ++d) geo[g].xd[d] = make_double2(D.xmin0[d], D.dx[d]);
      geo[g].dx_push = std::min(D.dx[0], std::min(D.dx[1], D.dx[2]));
      geo[g].pad = 0.0;
    }
    std::vector<int> nbr((size_t)std::max<long long>(nl, 1) * 27, -1);
    for (long long q = 0; q < nl; ++q) {
      const DBlock &D = hb[c->local_gid[q]];
      const int s = 1 << (m->max_level - D.level);
      for (int bz = -1; bz <= 1; ++bz)
        for (int by = -1; by <= 1; ++by)
          for (int bx = -1; bx <= 1; ++bx) {
            const int off[3] = {bx, by, bz};
            int tl[3] = {0, 0, 0};
            bool ok = true;
            for (int d = 0; d < 3 && ok; ++d) {
              if (d >= m->ndim) {
                ok = off[d] == 0;
                continue;
              }
              const int nbl = m->nrb[d] << D.level;  // blocks per dimension at this level
              int t = D.lx[d] + off[d];
              if (t < 0) {
                ok = m->swarm_bc[2 * d] == JB_BC_PERIODIC;
                t += nbl;
              } else if (t >= nbl) {
                ok = m->swarm_bc[2 * d + 1] == JB_BC_PERIODIC;
                t -= nbl;
              }
              tl[d] = t;
            }
            if (!ok) continue;
            const int g = lattice[((size_t)(tl[2] * s) * c->nlat[1] + (size_t)(tl[1] * s)) * c->nlat[0] + (size_t)(tl[0] * s)];
            const DBlock &N = hb[g];
            if (N.level == D.level && N.lx[0] == tl[0] && N.lx[1] == tl[1] && N.lx[2] == tl[2])
              nbr[(size_t)q * 27 + (bz + 1) * 9 + (by + 1) * 3 + (bx + 1)] = g;
          }
    }
    JB_CUDA(c, dalloc(&c->d_geo, m->nblocks));
    JB_CUDA(c, cudaMemcpy(c->d_geo, geo.data(), sizeof(BlkGeo) * geo.size(), cudaMemcpyHostToDevice));
    JB_CUDA(c, dalloc(&c->d_nbr, (long long)nbr.size()));
    JB_CUDA(c, cudaMemcpy(c->d_nbr, nbr.data(), sizeof(int) * nbr.size(), cudaMemcpyHostToDevice));
  }
  double **cellf[8] = {&c->f.rho, &c->f.sie, &c->f.u, &c->f.tally, &c->f.fleck, &c->f.src_ew, &c->f.src_num, &c->f.edelta};
  for (auto pp : cellf) {
    JB_CUDA(c, dalloc(pp, c->ncells_local));
    JB_CUDA(c, cudaMemset(*pp, 0, sizeof(double) * (size_t)std::max<long long>(c->ncells_local, 1)));
  }
  JB_CUDA(c, dalloc(&c->f.props, c->ncells_local));
  c->props_dirty = true;
  c->f.nface[0] = (long long)(m->nxb[0] + 1) * m->nxb[1] * m->nxb[2];
  c->f.nface[1] = (long long)m->nxb[0] * (m->nxb[1] + 1) * m->nxb[2];
  c->f.nface[2] = (long long)m->nxb[0] * m->nxb[1] * (m->nxb[2] + 1);
  if (c->cfg.use_ddmc && n_ranks > 1) {
    JB_CUDA(c, dalloc(&c->f.g_rho, nc * m->nblocks));
    JB_CUDA(c, dalloc(&c->f.g_sie, nc * m->nblocks));
    JB_CUDA(c, cudaMemset(c->f.g_rho, 0, sizeof(double) * (size_t)(nc * m->nblocks)));
    JB_CUDA(c, cudaMemset(c->f.g_sie, 0, sizeof(double) * (size_t)(nc * m->nblocks)));
  }
  if (c->cfg.use_ddmc) {
    for (int d = 0; d < 3; ++d) {
      JB_CUDA(c, dalloc(&c->f.fp[d], c->f.nface[d] * nl));
      JB_CUDA(c, cudaMemset(c->f.fp[d], 0, sizeof(double) * (size_t)std::max<long long>(c->f.nface[d] * nl, 1)));
    }
  }
  JB_CUDA(c, dalloc(&c->d_cnt, c->ncells_local));
  JB_CUDA(c, dalloc(&c->d_prefix, c->ncells_local + 1));
  JB_CUDA(c, dalloc(&c->d_scan, scan_scratch_elems(std::max<long long>(c->ncells_local, 1))));
  JB_CUDA(c, dalloc(&c->d_count_matrix, (long long)n_ranks * (n_ranks + 3) + (n_ranks + 3)));
  c->mesh_set = true;
  return JB_OK;
}

jb_status jb_set_capacity(jb_ctx *c, int64_t max_particles) {
  if (!c || max_particles < 0) return JB_ERR_INVALID;
  if (!c->mesh_set) return fail(c, JB_ERR_STATE, "mesh not set");
  JB_CUDA(c, cudaSetDevice(c->device));
  c->requested_capacity = max_particles;
  return ensure_capacity(c, max_particles);
}

static double *field_ptr(jb_ctx *c, int field, long long *per_block) {
  *per_block = jb_field_size(c, field);
  switch (field) {
    case JB_FIELD_DENSITY: return c->f.rho;
    case JB_FIELD_SIE: return c->f.sie;
    case JB_FIELD_INTERNAL_ENERGY: return c->f.u;
    case JB_FIELD_ENERGY_TALLY: return c->f.tally;
    case JB_FIELD_FLECK_FACTOR: return c->f.fleck;
    case JB_FIELD_SOURCE_EW_PER_CELL: return c->f.src_ew;
    case JB_FIELD_SOURCE_NUM_PER_CELL: return c->f.src_num;
    case JB_FIELD_ENERGY_DELTA: return c->f.edelta;
    case JB_FIELD_DDMC_FACE_PROB_X1: return c->f.fp[0];
    case JB_FIELD_DDMC_FACE_PROB_X2: return c->f.fp[1];
    case JB_FIELD_DDMC_FACE_PROB_X3: return c->f.fp[2];
  }
  return nullptr;
}

jb_status jb_set_field(jb_ctx *c, int32_t gid, int32_t field, const double *src) {
  if (!c || !src) return JB_ERR_INVALID;
  if (!c->mesh_set) return fail(c, JB_ERR_STATE, "mesh not set");
  if (gid < 0 || gid >= c->mesh.nblocks || c->gid_to_local[gid] < 0) return fail(c, JB_ERR_INVALID, "block not local");
  long long nb;
  double *p = field_ptr(c, field, &nb);
  if (!p) return fail(c, JB_ERR_INVALID, "bad field (face probabilities exist only with use_ddmc)");
  JB_CUDA(c, cudaSetDevice(c->device));
  JB_CUDA(c, cudaMemcpyAsync(p + (size_t)c->gid_to_local[gid] * nb, src, sizeof(double) * nb, cudaMemcpyDefault, c->stream));
  JB_CUDA(c, cudaStreamSynchronize(c->stream));
  c->props_dirty = true;
  return JB_OK;
}
jb_status jb_set_fields(jb_ctx *c, int32_t field, int32_t n, const int32_t *gids, const double *const *srcs) {
  if (!c || n < 0 || (n > 0 && (!gids || !srcs))) return JB_ERR_INVALID;
  if (!c->mesh_set) return fail(c, JB_ERR_STATE, "mesh not set");
  long long nb;
  double *p = field_ptr(c, field, &nb);
  if (!p) return fail(c, JB_ERR_INVALID, "bad field (face probabilities exist only with use_ddmc)");
  JB_CUDA(c, cudaSetDevice(c->device));
  for (int32_t q = 0; q < n; ++q)
    if (gids[q] < 0 || gids[q] >= c->mesh.nblocks || c->gid_to_local[gids[q]] < 0 || !srcs[q])
      return fail(c, JB_ERR_INVALID, "block not local");
  // Blocks that are neighbours both in the caller's memory and in the library's field array (a MeshData pack, a
  // [block][cell] host array) travel as ONE copy: 512 blocks x 256 KB cost ~3 ms of cudaMemcpyAsync calls per field
  // and direction otherwise. A merged copy that the runtime refuses (the run spans two host allocations) is
  // repeated block by block.
  for (int32_t q = 0; q < n;) {
    int32_t e = q + 1;
    while (e < n && c->gid_to_local[gids[e]] == c->gid_to_local[gids[e - 1]] + 1 && srcs[e] == srcs[e - 1] + nb) ++e;
    double *dst = p + (size_t)c->gid_to_local[gids[q]] * nb;
    if (cudaMemcpyAsync(dst, srcs[q], sizeof(double) * nb * (size_t)(e - q), cudaMemcpyDefault, c->stream) != cudaSuccess) {
      cudaGetLastError();
      for (int32_t r = q; r < e; ++r)
        JB_CUDA(c, cudaMemcpyAsync(p + (size_t)c->gid_to_local[gids[r]] * nb, srcs[r], sizeof(double) * nb, cudaMemcpyDefault,
                                   c->stream));
    }
    q = e;
  }
  JB_CUDA(c, cudaStreamSynchronize(c->stream));
  c->props_dirty = true;
  return JB_OK;
}
jb_status jb_get_fields(jb_ctx *c, int32_t field, int32_t n, const int32_t *gids, double *const *dsts) {
  if (!c || n < 0 || (n > 0 && (!gids || !dsts))) return JB_ERR_INVALID;
  if (!c->mesh_set) return fail(c, JB_ERR_STATE, "mesh not set");
  long long nb;
  double *p = field_ptr(c, field, &nb);
  JB_CUDA(c, cudaSetDevice(c->device));
  for (int32_t q = 0; q < n; ++q)
    if (gids[q] < 0 || gids[q] >= c->mesh.nblocks || c->gid_to_local[gids[q]] < 0 || !dsts[q])
      return fail(c, JB_ERR_INVALID, "block not local");
  if (!p) {
    for (int32_t q = 0; q < n; ++q) memset(dsts[q], 0, sizeof(double) * nb);
    return JB_OK;
  }
  for (int32_t q = 0; q < n;) {  // runs of neighbouring blocks as one copy, see jb_set_fields
    int32_t e = q + 1;
    while (e < n && c->gid_to_local[gids[e]] == c->gid_to_local[gids[e - 1]] + 1 && dsts[e] == dsts[e - 1] + nb) ++e;
    const double *src = p + (size_t)c->gid_to_local[gids[q]] * nb;
    if (cudaMemcpyAsync(dsts[q], src, sizeof(double) * nb * (size_t)(e - q), cudaMemcpyDefault, c->stream) != cudaSuccess) {
      cudaGetLastError();
      for (int32_t r = q; r < e; ++r)
        JB_CUDA(c, cudaMemcpyAsync(dsts[r], p + (size_t)c->gid_to_local[gids[r]] * nb, sizeof(double) * nb, cudaMemcpyDefault,
                                   c->stream));
    }
    q = e;
  }
  JB_CUDA(c, cudaStreamSynchronize(c->stream));
  return JB_OK;
}
// rho / sie of a block another rank owns, for hosts that move the DDMC ghost data themselves
jb_status jb_set_remote_field(jb_ctx *c, int32_t gid, int32_t field, const double *src) {
  if (!c || !src) return JB_ERR_INVALID;
  if (!c->mesh_set) return fail(c, JB_ERR_STATE, "mesh not set");
  if (gid < 0 || gid >= c->mesh.nblocks) return fail(c, JB_ERR_INVALID, "bad block");
  if (field != JB_FIELD_DENSITY && field != JB_FIELD_SIE) return fail(c, JB_ERR_INVALID, "remote fields are density and sie");
  if (!c->f.g_rho) return JB_OK;  // nothing reads them (single rank or no DDMC)
  const long long nc = (long long)c->mesh.nxb[0] * c->mesh.nxb[1] * c->mesh.nxb[2];
  double *dst = (field == JB_FIELD_DENSITY ? c->f.g_rho : c->f.g_sie) + (size_t)gid * nc;
  JB_CUDA(c, cudaSetDevice(c->device));
  JB_CUDA(c, cudaMemcpyAsync(dst, src, sizeof(double) * nc, cudaMemcpyDefault, c->stream));
  JB_CUDA(c, cudaStreamSynchronize(c->stream));
  return JB_OK;
}
jb_status jb_get_field(jb_ctx *c, int32_t gid, int32_t field, double *dst) {
  if (!c || !dst) return JB_ERR_INVALID;
  if (!c->mesh_set) return fail(c, JB_ERR_STATE, "mesh not set");
  if (gid < 0 || gid >= c->mesh.nblocks || c->gid_to_local[gid] < 0) return fail(c, JB_ERR_INVALID, "block not local");
  long long nb;
  double *p = field_ptr(c, field, &nb);
  JB_CUDA(c, cudaSetDevice(c->device));
  if (!p) {
    memset(dst, 0, sizeof(double) * nb);
    return JB_OK;
  }
  JB_CUDA(c, cudaMemcpyAsync(dst, p + (size_t)c->gid_to_local[gid] * nb, sizeof(double) * nb, cudaMemcpyDefault, c->stream));
  JB_CUDA(c, cudaStreamSynchronize(c->stream));
  return JB_OK;
}

// ---- internals of the step ---------------------------------------------------------------
static jb_status sync_stats(jb_ctx *c) {
  {
    StepParams P = make_params(c);
    launch_fold_bins(P, EB_SOURCED, &c->d_stats->e_sourced, c->stream);
    launch_fold_bins(P, EB_ABSORBED, &c->d_stats->e_absorbed, c->stream);
    launch_fold_bins(P, EB_ESCAPED, &c->d_stats->e_escaped, c->stream);
    c->launches += 3;
  }
  JB_CUDA(c, cudaMemcpyAsync(c->h_stats, c->d_stats, sizeof(DStats), cudaMemcpyDeviceToHost, c->stream));
  JB_CUDA(c, cudaStreamSynchronize(c->stream));
  return JB_OK;
}

// census: photons in use / unfinished / energy over slots [0, n)
static jb_status census(jb_ctx *c, double t_end, long long *n_alive, long long *n_unfinished, double *e_sum) {
  if (c->census_valid && !n_unfinished) {  // the tally that just ran already counted them
    if (n_alive) *n_alive = c->census_n;
    if (e_sum) *e_sum = c->census_e;
    return JB_OK;
  }
  StepParams P = make_params(c);
  JB_CUDA(c, cudaMemsetAsync(c->d_counters, 0, sizeof(unsigned long long) * 2, c->stream));
  JB_CUDA(c, cudaMemsetAsync(c->d_esum, 0, sizeof(double), c->stream));
  launch_census(P, c->n, t_end, c->d_counters, c->d_counters + 1, c->d_esum, c->stream);
  c->launches += 1 + (c->n > 0);
  JB_CUDA(c, cudaMemcpyAsync(c->h_counters, c->d_counters, sizeof(unsigned long long) * 2, cudaMemcpyDeviceToHost, c->stream));
  JB_CUDA(c, cudaMemcpyAsync(c->h_counters + 2, c->d_esum, sizeof(double), cudaMemcpyDeviceToHost, c->stream));
  JB_CUDA(c, cudaStreamSynchronize(c->stream));
  if (n_alive) *n_alive = (long long)c->h_counters[0];
  if (n_unfinished) *n_unfinished = (long long)c->h_counters[1];
  if (e_sum) memcpy(e_sum, c->h_counters + 2, sizeof(double));
  return JB_OK;
}

// Removed photons leave holes in the pool (loc.x = -1; every kernel skips them). Like Parthenon,
// which frees the slots of marked particles and only defragments a pool whose occupancy has
// fallen below <jaybenne> min_swarm_occupancy (jaybenne.cpp:171-175), the pool is squeezed only
// when fewer than max(min_swarm_occupancy, 1/2) of the used slots hold a photon, when the
// caller needs a dense pool (swarm access, cell sort) or when room is needed (`force`).
static jb_status compact(jb_ctx *c, long long n_dead, bool force = false) {
  if (n_dead < 0) return fail(c, JB_ERR_STATE, "photon bookkeeping went negative");
  c->n_holes += n_dead;
  if (c->n_holes <= 0) return JB_OK;
  const long long n_alive = c->n - c->n_holes;
  if (n_alive < 0) return fail(c, JB_ERR_STATE, "photon bookkeeping went negative");
  if (!force) {
    const double floor_occ = std::max(0.5, c->cfg.min_swarm_occupancy);
    if ((double)n_alive >= floor_occ * (double)c->n) return JB_OK;
  }
  const long long holes = c->n_holes;
  c->n_holes = 0;
  if (n_alive == 0) {
    c->n = 0;
    return JB_OK;
  }
  if (c->squeeze_mode < 0) {
    const char *v = getenv("JB_SQUEEZE");
    c->squeeze_mode = (v && strcmp(v, "tail") == 0) ? 0 : 1;
  }
  if (c->squeeze_mode == 1) {
    // order-preserving: every array streams through the scratch array and swaps storage with it. Falls back to
    // the in-place tail-into-holes squeeze when the scratch array cannot be allocated.
    const long long nchunk = squeeze_chunks(c->n);
    bool ok = true;
    if (!c->d_tmp) ok = dalloc(&c->d_tmp, c->capacity) == cudaSuccess;
    if (ok && nchunk > c->sq_chunks_cap) {
      dfree(c->d_sq_mask); dfree(c->d_sq_cnt); dfree(c->d_sq_prefix); dfree(c->d_sq_scan);
      c->d_sq_mask = nullptr; c->d_sq_cnt = nullptr; c->d_sq_prefix = c->d_sq_scan = nullptr;
      c->sq_chunks_cap = squeeze_chunks(c->capacity) + 1;
      ok = dalloc(&c->d_sq_mask, c->sq_chunks_cap * 32) == cudaSuccess && dalloc(&c->d_sq_cnt, c->sq_chunks_cap) == cudaSuccess &&
           dalloc(&c->d_sq_prefix, c->sq_chunks_cap + 1) == cudaSuccess &&
           dalloc(&c->d_sq_scan, scan_scratch_elems(c->sq_chunks_cap)) == cudaSuccess;
      if (!ok) c->sq_chunks_cap = 0;
    }
    if (ok) {
      cudaStream_t st = c->stream;
      launch_squeeze_flags(c->s.loc, c->n, c->d_sq_mask, c->d_sq_cnt, st);
      launch_exclusive_scan(c->d_sq_cnt, c->d_sq_prefix, nchunk, c->d_sq_scan, c->d_total, st);
      void **arr[12];
      swarm_arrays(c->s, arr);
      for (auto pa : arr) {
        launch_squeeze_move(*pa, c->d_tmp, c->n, c->d_sq_mask, c->d_sq_prefix, st);
        void *old = *pa;
        *pa = (void *)c->d_tmp;
        c->d_tmp = (double *)old;
      }
      c->launches += 4 + 12;
      c->n = n_alive;
      JB_CUDA(c, cudaGetLastError());
      return JB_OK;
    }
    cudaGetLastError();  // out of memory for the scratch: squeeze in place instead
    dfree(c->d_tmp);
    c->d_tmp = nullptr;
    c->squeeze_mode = 0;
  }
  const long long need = std::min(holes, n_alive);
  if (need > c->holes_cap) {
    dfree(c->d_holes);
    dfree(c->d_tails);
    c->holes_cap = need + need / 4 + 1024;
    JB_CUDA(c, dalloc(&c->d_holes, c->holes_cap));
    JB_CUDA(c, dalloc(&c->d_tails, c->holes_cap));
  }
  StepParams P = make_params(c);
  launch_compact(P, c->n, n_alive, c->d_holes, c->d_tails, c->d_counters + 4, c->stream);
  c->launches += 2;
  c->n = n_alive;
  return JB_OK;
}
static jb_status ensure_dense(jb_ctx *c) { return compact(c, 0, true); }

// SourcePhotons over local blocks [first, first + nb) (sourcing.cpp:25-205)
static jb_status source_photons(jb_ctx *c, int first, int nb, int source_type, double t_start, double dt,
                                int nblocks_npc, uint32_t call) {
  if (c->cfg.source_strategy == JB_STRATEGY_ENERGY)
    return fail(c, JB_ERR_UNSUPPORTED, "Energy source strategy not implemented!");
  if (source_type == JB_SOURCE_EMISSION && !c->cfg.do_emission) return JB_OK;
  if (nb <= 0) return JB_OK;
  const int num_cells = c->mesh.nxb[0] * c->mesh.nxb[1] * c->mesh.nxb[2];
  const double npc = static_cast<double>(c->cfg.num_particles) / num_cells / (nblocks_npc * c->mesh.nblocks);
  const long long ncells = (long long)nb * num_cells;
  StepParams P = make_params(c);
  launch_source_count(P, first, nb, source_type, dt, npc, call, c->d_cnt, c->stream);
  launch_exclusive_scan(c->d_cnt, c->d_prefix, ncells, c->d_scan, c->d_total, c->stream);
  c->launches += 4;
  JB_CUDA(c, cudaMemcpyAsync(c->h_counters, c->d_total, sizeof(long long), cudaMemcpyDeviceToHost, c->stream));
  JB_CUDA(c, cudaStreamSynchronize(c->stream));
  const long long total = (long long)c->h_counters[0];
  if (total < 0 || total > 0x7fffffffLL * (long long)nb) return fail(c, JB_ERR_CAPACITY, "photon count overflow");
  if (c->n + total > c->capacity && c->n_holes > 0) {
    jb_status sd = ensure_dense(c);
    if (sd != JB_OK) return sd;
    P = make_params(c);
  }
  if (c->n + total > c->capacity) {
    if (c->requested_capacity > 0 && c->n + total > c->requested_capacity)
      return fail(c, JB_ERR_CAPACITY, "photon pool capacity exceeded while sourcing (jb_set_capacity)");
    jb_status s = ensure_capacity(c, c->n + total);
    if (s != JB_OK) return s;
    P = make_params(c);
  }
  c->census_valid = false;
  launch_source_emit(P, first, nb, source_type, t_start, dt, call, c->d_prefix, c->n, total, c->stream);
  c->launches += total < 4 * ncells ? (total > 0 ? 2 : 1) : 1;
  c->n += total;
  JB_CUDA(c, cudaGetLastError());
  return JB_OK;
}

static jb_status stats_begin(jb_ctx *c) {
  memset(&c->stats, 0, sizeof(c->stats));
  JB_CUDA(c, cudaMemsetAsync(c->d_stats, 0, sizeof(DStats), c->stream));
  JB_CUDA(c, cudaMemsetAsync(c->d_ebins, 0, sizeof(double) * EB_KINDS * EBINS, c->stream));
  long long na = 0;
  double e = 0.0;
  jb_status s = census(c, 0.0, &na, nullptr, &e);
  if (s != JB_OK) return s;
  memset(c->h_stats, 0, sizeof(DStats));
  c->stats.n_start = na;
  c->stats.e_start = e;
  return JB_OK;
}
static void fold_stats(jb_ctx *c) {
  const DStats &h = *c->h_stats;
  c->stats.n_sourced = (int64_t)h.n_sourced;
  c->stats.n_absorbed = (int64_t)h.n_absorbed;
  c->stats.n_escaped = (int64_t)h.n_escaped;
  c->stats.n_sent = (int64_t)h.n_sent;
  c->stats.n_received = (int64_t)h.n_received;
  c->stats.n_block_crossings = (int64_t)h.n_crossings;
  c->stats.n_segments = (int64_t)h.n_segments;
  c->stats.n_scatters = (int64_t)h.n_scatters;
  c->stats.n_ddmc_segments = (int64_t)h.n_ddmc;
  c->stats.n_bface_resampled = (int32_t)h.n_bface;
  c->stats.e_sourced = h.e_sourced;
  c->stats.e_absorbed = h.e_absorbed;
  c->stats.e_escaped = h.e_escaped;
}
static jb_status check_device_errors(jb_ctx *c) {
  if (c->h_stats->n_error)
    return fail(c, JB_ERR_STATE, "particle left the domain without a boundary condition");
  if (c->h_stats->n_bface_ghost)
    return fail(c, JB_ERR_STATE, "SampleDDMCBlockFace indexed a face probability outside the block (not a 2:1 mesh?)");
  if (c->h_stats->n_validate)
    return fail(c, JB_ERR_STATE, "JB_IMC_VALIDATE: " + std::to_string(c->h_stats->n_validate) +
                                     " index / division cross-checks disagreed");
  if (c->h_stats->n_overflow)
    return fail(c, JB_ERR_CAPACITY, "exchange outbox overflow: raise jb_set_capacity");
  return JB_OK;
}

// all-to-all-v of the outboxes over NCCL: counts (all-gather) then grouped send/recv into
// the inbox, then append to the pool. Returns the number of records sent by ALL ranks.
static jb_status exchange_nccl(jb_ctx *c, double t_end, long long *global_sent) {
  NcclApi &nc = nccl_api();
  const int R = c->n_ranks;
  if (!c->comm || !nc.ok) return fail(c, JB_ERR_COMM, "jb_comm_init has not been called");
  if (!c->out.rec) {
    jb_status s0 = ensure_capacity(c, std::max<long long>(c->capacity + 1, 1024));
    if (s0 != JB_OK) return s0;
  }
#define JB_NCCL(call)                                                                                  \
  do {                                                                                                 \
    ncclResult_t r_ = (call);                                                                          \
    if (r_ != ncclSuccess)                                                                             \
      return fail(c, JB_ERR_COMM, std::string(#call) + ": " + (nc.GetErrorString ? nc.GetErrorString(r_) : "?")); \
  } while (0)
  // Every rank gathers every rank's row: the R outbox counts, then its outbox capacity, its inbox capacity and
  // the room left in its pool (all ones = the pool may grow). With the whole table on every rank, a buffer
  // that would overflow ANYWHERE fails the call on ALL ranks before any of them enters the send / receive
  // group — a rank returning alone would leave its peers hanging in ncclGroupEnd.
  const int RW = R + 3;
  unsigned long long *d_row = c->d_count_matrix + (size_t)R * RW;
  {
    unsigned long long *h_row = c->h_counters + 8;
    h_row[0] = (unsigned long long)c->out.cap;
    h_row[1] = (unsigned long long)c->inbox_cap;
    h_row[2] = c->requested_capacity > 0 ? (unsigned long long)std::max<long long>(c->capacity - c->n, 0) : ~0ull;
    JB_CUDA(c, cudaMemcpyAsync(d_row, c->out.count, sizeof(unsigned long long) * R, cudaMemcpyDeviceToDevice, c->stream));
    JB_CUDA(c, cudaMemcpyAsync(d_row + R, h_row, sizeof(unsigned long long) * 3, cudaMemcpyHostToDevice, c->stream));
  }
  JB_NCCL(nc.AllGather(d_row, c->d_count_matrix, (size_t)RW, ncclUint64, c->comm, c->stream));
  JB_CUDA(c, cudaMemcpyAsync(c->h_counters + 16, c->d_count_matrix, sizeof(unsigned long long) * R * RW,
                             cudaMemcpyDeviceToHost, c->stream));
  JB_CUDA(c, cudaStreamSynchronize(c->stream));
  const unsigned long long *T = c->h_counters + 16;  // T[src * RW + dst], then src's out.cap, inbox_cap, room
  auto Mat = [&](int a, int b) { return (long long)T[(size_t)a * RW + b]; };
  long long total = 0, nrecv = 0;
  std::string problem;
  for (int a = 0; a < R; ++a) {
    const unsigned long long out_cap_a = T[(size_t)a * RW + R];
    long long col = 0;
    for (int b = 0; b < R; ++b) {
      total += Mat(a, b);
      col += Mat(b, a);
      if ((unsigned long long)Mat(a, b) > out_cap_a && problem.empty())
        problem = "exchange outbox overflow on rank " + std::to_string(a) + " (records for rank " + std::to_string(b) +
                  "): raise jb_set_capacity";
    }
    if ((unsigned long long)col > T[(size_t)a * RW + R + 1] && problem.empty())
      problem = "exchange inbox overflow on rank " + std::to_string(a) + ": raise jb_set_capacity";
    if ((unsigned long long)col > T[(size_t)a * RW + R + 2] && problem.empty())
      problem = "photon pool capacity exceeded on receive on rank " + std::to_string(a);
  }
  for (int a = 0; a < R; ++a) nrecv += Mat(a, c->my_rank);
  *global_sent = total;
  if (!problem.empty()) {  // the same verdict on every rank: nobody sends
    cudaMemsetAsync(c->out.count, 0, sizeof(unsigned long long) * R, c->stream);
    return fail(c, JB_ERR_CAPACITY, problem);
  }
  if (total == 0) return JB_OK;
  if (c->n + nrecv > c->capacity) {  // only pools that may grow get here (the fixed ones were checked above)
    jb_status gs = grow_pool_only(c, c->n + nrecv);
    if (gs != JB_OK) return gs;
  }
  JB_NCCL(nc.GroupStart());
  long long roff = 0;
  for (int peer = 0; peer < R; ++peer) {
    const long long ns = std::min(Mat(c->my_rank, peer), c->out.cap), nr = Mat(peer, c->my_rank);
    if (ns > 0)
      JB_NCCL(nc.Send(c->out.rec + (size_t)peer * c->out.cap, (size_t)ns * sizeof(ExRecord), ncclUint8, peer, c->comm,
                      c->stream));
    if (nr > 0) {
      JB_NCCL(nc.Recv(c->d_inbox + roff, (size_t)nr * sizeof(ExRecord), ncclUint8, peer, c->comm, c->stream));
      roff += nr;
    }
  }
  JB_NCCL(nc.GroupEnd());
#undef JB_NCCL
  JB_CUDA(c, cudaMemsetAsync(c->out.count, 0, sizeof(unsigned long long) * R, c->stream));
  if (nrecv > 0) {
    StepParams P = make_params(c);
    c->census_valid = false;
    launch_unpack(P, c->d_inbox, nrecv, c->n, t_end, c->stream);
    if (c->cfg.use_ddmc) launch_bface_range(P, c->n, nrecv, c->stream);
    c->launches += 1 + (c->cfg.use_ddmc ? 1 : 0);
    c->n += nrecv;
  }
  return JB_OK;
}

// sort the pool by (block, cell) with a counting sort; photons must be compacted. The
// permutation is applied array by array through the one scratch array, whose storage
// then swaps roles with the array just gathered.
static jb_status sort_by_cell(jb_ctx *c) {
  {
    jb_status sd = ensure_dense(c);
    if (sd != JB_OK) return sd;
  }
  if (c->n <= 1) return JB_OK;
  if (c->n > 0xffffffffLL) return fail(c, JB_ERR_CAPACITY, "sort supports < 2^32 photons per rank");
  // scratch is only allocated when it is used (d_tmp also serves the order-preserving squeeze)
  if (!c->d_tmp) JB_CUDA(c, dalloc(&c->d_tmp, c->capacity));
  if (!c->d_perm) JB_CUDA(c, dalloc(&c->d_perm, c->capacity));
  StepParams P = make_params(c);
  cudaStream_t st = c->stream;
  JB_CUDA(c, cudaMemsetAsync(c->d_cnt, 0, sizeof(int) * (size_t)c->ncells_local, st));
  launch_sort_keys(P, c->n, c->d_cnt, st);
  launch_exclusive_scan(c->d_cnt, c->d_prefix, c->ncells_local, c->d_scan, c->d_total, st);
  launch_sort_perm(P, c->n, c->d_prefix, c->d_perm, st);
  c->launches += 5;
  void **arr[12];
  swarm_arrays(c->s, arr);
  for (auto pa : arr) {
    launch_gather_b64(*pa, c->d_tmp, c->d_perm, c->n, st);
    void *old = *pa;
    *pa = (void *)c->d_tmp;
    c->d_tmp = (double *)old;
    c->launches++;
  }
  JB_CUDA(c, cudaGetLastError());
  return JB_OK;
}

// Multi-rank DDMC: bring the mirror of every block's rho and sie up to date. Own blocks are
// copied in; with an NCCL communicator every rank sends its blocks to every peer (grouped
// send/recv, one exchange per cycle like the reference's ghost exchange); without one the host
// program is expected to have delivered the other ranks' blocks through jb_set_remote_field.
static jb_status refresh_field_mirror(jb_ctx *c) {
  const long long nc = (long long)c->mesh.nxb[0] * c->mesh.nxb[1] * c->mesh.nxb[2];
  for (size_t q = 0; q < c->local_gid.size(); ++q) {
    const size_t g = (size_t)c->local_gid[q];
    JB_CUDA(c, cudaMemcpyAsync(c->f.g_rho + g * nc, c->f.rho + q * nc, sizeof(double) * nc, cudaMemcpyDeviceToDevice, c->stream));
    JB_CUDA(c, cudaMemcpyAsync(c->f.g_sie + g * nc, c->f.sie + q * nc, sizeof(double) * nc, cudaMemcpyDeviceToDevice, c->stream));
  }
  if (!c->comm) return JB_OK;
  NcclApi &nc_api = nccl_api();
  if (!nc_api.ok) return fail(c, JB_ERR_COMM, "NCCL is not available");
  // blocks are owned in runs of consecutive gids: send / receive run by run
  bool err = nc_api.GroupStart() != ncclSuccess;
  const int nb = c->mesh.nblocks;
  int g = 0;
  while (g < nb && !err) {
    const int owner = c->all_blocks[g].rank;
    int e = g;
    while (e < nb && c->all_blocks[e].rank == owner) ++e;
    const size_t off = (size_t)g * nc, cnt = (size_t)(e - g) * nc;
    if (owner == c->my_rank) {
      for (int peer = 0; peer < c->n_ranks && !err; ++peer) {
        if (peer == c->my_rank) continue;
        err |= nc_api.Send(c->f.g_rho + off, cnt, ncclDouble, peer, c->comm, c->stream) != ncclSuccess;
        err |= nc_api.Send(c->f.g_sie + off, cnt, ncclDouble, peer, c->comm, c->stream) != ncclSuccess;
      }
    } else {
      err |= nc_api.Recv(c->f.g_rho + off, cnt, ncclDouble, owner, c->comm, c->stream) != ncclSuccess;
      err |= nc_api.Recv(c->f.g_sie + off, cnt, ncclDouble, owner, c->comm, c->stream) != ncclSuccess;
    }
    g = e;
  }
  err |= nc_api.GroupEnd() != ncclSuccess;
  if (err) return fail(c, JB_ERR_COMM, "DDMC ghost exchange (rho, sie) over NCCL failed");
  return JB_OK;
}

#define JB_REQUIRE_MESH(c)                                          \
  do {                                                              \
    if (!(c)) return JB_ERR_INVALID;                                \
    if (!(c)->mesh_set) return fail((c), JB_ERR_STATE, "mesh not set"); \
    JB_CUDA((c), cudaSetDevice((c)->device));                       \
  } while (0)

// ---- the individual tasks (jaybenne.hpp:59-69) -------------------------------------------
jb_status jb_update_derived_transport_fields(jb_ctx *c, double dt) {
  JB_REQUIRE_MESH(c);
  StepParams P = make_params(c);
  launch_fleck(P, dt, c->stream);
  c->launches += 1;
  c->props_dirty = true;
  if (c->cfg.use_ddmc) {
    if (c->f.g_rho) {
      jb_status s = refresh_field_mirror(c);
      if (s != JB_OK) return s;
    }
    launch_face_prob(P, c->stream);
    c->launches += c->mesh.ndim;
  }
  JB_CUDA(c, cudaGetLastError());
  return JB_OK;
}

jb_status jb_source_photons(jb_ctx *c, int32_t source_type, double t_start, double dt) {
  JB_REQUIRE_MESH(c);
  const int nl = (int)c->local_gid.size();
  const int nb = c->cfg.source_nblocks > 0 ? c->cfg.source_nblocks : nl;
  jb_status s = source_photons(c, 0, nl, source_type, t_start, dt, nb, c->source_call);
  if (s != JB_OK) return s;
  if (!(source_type == JB_SOURCE_EMISSION && !c->cfg.do_emission)) c->source_call++;
  return JB_OK;
}

// `from_source`: the pool holds exactly the photons of the sourcing pass that just ran over all local blocks
// (jb_initialize_radiation into an empty pool) — the tally then comes from the per-cell source fields
static jb_status evaluate_radiation_energy(jb_ctx *c, bool from_source) {
  StepParams P = make_params(c);
  JB_CUDA(c, cudaMemsetAsync(c->d_counters, 0, sizeof(unsigned long long) * 2, c->stream));
  JB_CUDA(c, cudaMemsetAsync(c->d_esum, 0, sizeof(double), c->stream));
  if (from_source)
    launch_tally_sourced(P, c->d_counters, c->d_esum, c->stream);
  else
    launch_tally(P, c->n, c->d_counters, c->d_esum, c->stream);
  c->launches += 1 + (c->n > 0);
  JB_CUDA(c, cudaMemcpyAsync(c->h_counters, c->d_counters, sizeof(unsigned long long), cudaMemcpyDeviceToHost, c->stream));
  JB_CUDA(c, cudaMemcpyAsync(c->h_counters + 2, c->d_esum, sizeof(double), cudaMemcpyDeviceToHost, c->stream));
  JB_CUDA(c, cudaStreamSynchronize(c->stream));
  JB_CUDA(c, cudaGetLastError());
  c->census_n = (long long)c->h_counters[0];
  memcpy(&c->census_e, c->h_counters + 2, sizeof(double));
  c->census_valid = true;
  return JB_OK;
}
jb_status jb_evaluate_radiation_energy(jb_ctx *c) {
  JB_REQUIRE_MESH(c);
  return evaluate_radiation_energy(c, false);
}

// EvaluateRadiationEnergy for callers that may not allocate: no cudaMalloc / cudaFree in this call or below it
// (launch_tally_noalloc: scratch in a file-scope device array). The context's own energy_tally field and census
// cache are left untouched; the result goes to the caller's buffers.
jb_status jb_evaluate_radiation_energy_noalloc(jb_ctx *c, double *tally_out, int64_t tally_len, int64_t *n_census,
                                               double *e_census) {
  JB_REQUIRE_MESH(c);
  if (!tally_out) return fail(c, JB_ERR_INVALID, "jb_evaluate_radiation_energy_noalloc: tally_out is null");
  if (c->ncells_local > tally_noalloc_max_cells())
    return fail(c, JB_ERR_CAPACITY, "jb_evaluate_radiation_energy_noalloc: " + std::to_string(c->ncells_local) +
                                        " local cells exceed the fixed scratch of " +
                                        std::to_string(tally_noalloc_max_cells()) + " cells (JB_NOALLOC_MAX_CELLS)");
  if (tally_len < c->ncells_local)
    return fail(c, JB_ERR_INVALID, "jb_evaluate_radiation_energy_noalloc: tally_out holds fewer than the " +
                                       std::to_string(c->ncells_local) + " local cells");
  StepParams P = make_params(c);
  long long n = 0;
  double e = 0.0;
  JB_CUDA(c, launch_tally_noalloc(P, c->n, tally_out, &n, &e, c->stream));
  c->launches += 1 + (c->n > 0);
  if (n_census) *n_census = (int64_t)n;
  if (e_census) *e_census = e;
  return JB_OK;
}

// InitializeRadiation on every local block (jaybenne.cpp:570-578). The reference sources
// each block through its own MeshBlockData pack (nblocks = 1 in the per-cell count); the
// blocks are independent, so one launch covers them all.
jb_status jb_initialize_radiation(jb_ctx *c, int32_t is_thermal) {
  JB_REQUIRE_MESH(c);
  const bool was_empty = c->n == 0 && c->n_holes == 0;
  if (is_thermal) {
    jb_status s = source_photons(c, 0, (int)c->local_gid.size(), JB_SOURCE_THERMAL, 0.0, 0.0, 1, c->source_call);
    if (s != JB_OK) return s;
    c->source_call++;
  }
  static int fuse = -1;  // JB_TALLY_FROM_SOURCE=0: always walk the swarm (A/B, cross-check)
  if (fuse < 0) {
    const char *v = getenv("JB_TALLY_FROM_SOURCE");
    fuse = v ? (atoi(v) != 0) : 1;
  }
  jb_status s = evaluate_radiation_energy(c, is_thermal && was_empty && fuse);
  if (s != JB_OK) return s;
  JB_CUDA(c, cudaStreamSynchronize(c->stream));
  return JB_OK;
}

// TransportPhotons / TransportPhotons_DDMC over the whole pool, then free the slots of the
// photons that are gone (RemoveMarkedParticles, transport.cpp:176-178).
static jb_status transport_range(jb_ctx *c, long long first, long long n, double t_start, double dt) {
  StepParams P = make_params(c);
  if (c->props_dirty) {  // rho, sie or fleck_factor changed since the records were packed
    launch_pack_props(P, c->stream);
    c->launches += 1;
    c->props_dirty = false;
  }
  c->census_valid = false;
  JB_CUDA(c, cudaMemsetAsync(P.work, 0, sizeof(unsigned long long), c->stream));
  launch_transport(P, first, n, t_start, dt, c->stream);
  c->launches += n > 0;
  JB_CUDA(c, cudaGetLastError());
  return JB_OK;
}
jb_status jb_transport_photons(jb_ctx *c, double t_start, double dt) {
  JB_REQUIRE_MESH(c);
  const unsigned long long dead0 = c->h_stats->n_dead;
  jb_status s = transport_range(c, 0, c->n, t_start, dt);
  if (s != JB_OK) return s;
  s = sync_stats(c);
  if (s != JB_OK) return s;
  fold_stats(c);
  s = check_device_errors(c);
  if (s != JB_OK) return s;
  return compact(c, (long long)(c->h_stats->n_dead - dead0));
}

jb_status jb_swarm_exchange(jb_ctx *c) {
  JB_REQUIRE_MESH(c);
  if (c->n_ranks == 1 || !c->comm) return JB_OK;  // on-rank hand-over happens inside the transport kernel
  long long total = 0;
  jb_status s = exchange_nccl(c, std::numeric_limits<double>::infinity(), &total);
  if (s != JB_OK) return s;
  return sync_stats(c);
}

jb_status jb_sample_ddmc_block_face(jb_ctx *c) {
  JB_REQUIRE_MESH(c);
  if (!c->cfg.use_ddmc) return JB_OK;
  StepParams P = make_params(c);
  launch_bface(P, c->n, c->stream);
  c->launches += c->n > 0 && c->mesh.ndim > 1;
  JB_CUDA(c, cudaGetLastError());
  return JB_OK;
}

jb_status jb_check_completion(jb_ctx *c, double t_end, int64_t *num_unfinished) {
  JB_REQUIRE_MESH(c);
  long long u = 0;
  jb_status s = census(c, t_end, nullptr, &u, nullptr);
  if (num_unfinished) *num_unfinished = u;
  return s;
}

jb_status jb_update_fluid(jb_ctx *c) {
  JB_REQUIRE_MESH(c);
  if (!c->cfg.do_feedback) return JB_OK;
  StepParams P = make_params(c);
  launch_update_fluid(P, c->stream);
  c->launches += 1;
  JB_CUDA(c, cudaGetLastError());
  return JB_OK;
}
jb_status jb_update_derived_material(jb_ctx *c) {
  JB_REQUIRE_MESH(c);
  StepParams P = make_params(c);
  launch_update_derived_material(P, c->stream);
  c->launches += 1;
  c->props_dirty = true;
  JB_CUDA(c, cudaGetLastError());
  return JB_OK;
}
jb_status jb_post_initialization(jb_ctx *c) {
  JB_REQUIRE_MESH(c);
  StepParams P = make_params(c);
  launch_post_initialization(P, c->stream);
  c->launches += 1;
  JB_CUDA(c, cudaGetLastError());
  return JB_OK;
}

jb_status jb_stats_begin(jb_ctx *c) {
  JB_REQUIRE_MESH(c);
  return stats_begin(c);
}

jb_status jb_get_stats(jb_ctx *c, jb_cycle_stats *out) {
  JB_REQUIRE_MESH(c);
  if (!out) return JB_ERR_INVALID;
  jb_status s = sync_stats(c);
  if (s != JB_OK) return s;
  fold_stats(c);
  long long na = 0;
  double e = 0.0;
  s = census(c, 0.0, &na, nullptr, &e);
  if (s != JB_OK) return s;
  c->stats.n_end = na;
  c->stats.e_end = e;
  *out = c->stats;
  return JB_OK;
}

// ---- direct peer exchange over NVLink ---------------------------------------------------------------
// Set up once, collectively, at the first multi-rank step of a context with a fixed pool (jb_set_capacity):
// every rank allocates a mailbox and two inboxes, the CUDA IPC handles travel through one ncclAllGather, and
// every rank maps every peer's. Any failure on any rank (no peer access, JB_P2P=0, IPC refused) leaves ALL
// ranks on the NCCL send/recv path: the verdict is all-reduced.
__global__ void k_p2p_begin(P2PState *st, long long n, long long capacity) {
  st->n = n;
  st->capacity = capacity;
  st->range[0] = 0;
  st->range[1] = 0;
  st->done = 0;
  st->iterations = 0;
  st->error = 0;
  st->sent_last = 0;
}
static jb_status p2p_setup(jb_ctx *c) {
  c->p2p_tried = true;
  NcclApi &nc = nccl_api();
  const int R = c->n_ranks;
  int ok = 1;
  if (const char *v = getenv("JB_P2P")) ok = atoi(v) != 0;
  if (R > P2P_MAX_RANKS || c->requested_capacity <= 0 || !c->out.rec) ok = 0;
  struct Handles { cudaIpcMemHandle_t mail, in0, in1; };
  static_assert(sizeof(Handles) == 192, "three 64-byte handles");
  Handles mine;
  memset(&mine, 0, sizeof(mine));
  if (ok) {
    c->p2p_inbox_cap = 2 * c->out.cap;
    ok = dalloc(&c->d_mail, 1) == cudaSuccess && dalloc(&c->d_p2p_inbox[0], c->p2p_inbox_cap) == cudaSuccess &&
         dalloc(&c->d_p2p_inbox[1], c->p2p_inbox_cap) == cudaSuccess && dalloc(&c->d_p2p_state, 1) == cudaSuccess &&
         dalloc(&c->d_peers, 1) == cudaSuccess &&
         cudaMallocHost((void **)&c->h_p2p_state, sizeof(P2PState)) == cudaSuccess;
    if (ok) {
      cudaMemset(c->d_mail, 0, sizeof(P2PMail));
      cudaMemset(c->d_p2p_state, 0, sizeof(P2PState));
      ok = cudaIpcGetMemHandle(&mine.mail, c->d_mail) == cudaSuccess &&
           cudaIpcGetMemHandle(&mine.in0, c->d_p2p_inbox[0]) == cudaSuccess &&
           cudaIpcGetMemHandle(&mine.in1, c->d_p2p_inbox[1]) == cudaSuccess;
    }
    cudaGetLastError();
  }
  // all ranks take part in the two collectives whatever their local verdict
  unsigned char *d_h = nullptr;
  JB_CUDA(c, dalloc(&d_h, (long long)192 * (R + 1) + 64));
  std::vector<Handles> all((size_t)R);
  JB_CUDA(c, cudaMemcpy(d_h + (size_t)192 * R, &mine, 192, cudaMemcpyHostToDevice));
  bool comm_ok = nc.AllGather(d_h + (size_t)192 * R, d_h, 192, ncclUint8, c->comm, c->stream) == ncclSuccess;
  JB_CUDA(c, cudaStreamSynchronize(c->stream));
  JB_CUDA(c, cudaMemcpy(all.data(), d_h, (size_t)192 * R, cudaMemcpyDeviceToHost));
  P2PPeers peers;
  memset(&peers, 0, sizeof(peers));
  if (ok && comm_ok) {
    for (int r = 0; r < R && ok; ++r) {
      if (r == c->my_rank) {
        peers.mail[r] = c->d_mail;
        peers.inbox[0][r] = c->d_p2p_inbox[0];
        peers.inbox[1][r] = c->d_p2p_inbox[1];
        continue;
      }
      void *pm = nullptr, *p0 = nullptr, *p1 = nullptr;
      ok = cudaIpcOpenMemHandle(&pm, all[r].mail, cudaIpcMemLazyEnablePeerAccess) == cudaSuccess &&
           cudaIpcOpenMemHandle(&p0, all[r].in0, cudaIpcMemLazyEnablePeerAccess) == cudaSuccess &&
           cudaIpcOpenMemHandle(&p1, all[r].in1, cudaIpcMemLazyEnablePeerAccess) == cudaSuccess;
      for (void *q : {pm, p0, p1})
        if (q) c->ipc_opened.push_back(q);
      peers.mail[r] = (P2PMail *)pm;
      peers.inbox[0][r] = (ExRecord *)p0;
      peers.inbox[1][r] = (ExRecord *)p1;
    }
    cudaGetLastError();
  }
  int *d_ok = reinterpret_cast<int *>(d_h + (size_t)192 * (R + 1));
  int h_ok = (ok && comm_ok) ? 1 : 0;
  JB_CUDA(c, cudaMemcpy(d_ok, &h_ok, sizeof(int), cudaMemcpyHostToDevice));
  comm_ok = nc.AllReduce(d_ok, d_ok, 1, ncclInt, ncclMin, c->comm, c->stream) == ncclSuccess && comm_ok;
  JB_CUDA(c, cudaStreamSynchronize(c->stream));
  JB_CUDA(c, cudaMemcpy(&h_ok, d_ok, sizeof(int), cudaMemcpyDeviceToHost));
  cudaFree(d_h);
  if (!comm_ok) return fail(c, JB_ERR_COMM, "NCCL failed while exchanging the peer-memory handles");
  c->p2p_on = h_ok == 1;
  if (c->p2p_on) {
    JB_CUDA(c, cudaMemcpy(c->d_peers, &peers, sizeof(peers), cudaMemcpyHostToDevice));
    for (auto &e : c->it_ev) JB_CUDA(c, cudaEventCreate(&e));
  }
  return JB_OK;
}

// The transport / exchange iterations of one cycle with the exchange done by the GPUs themselves: the host
// queues P2P_CHUNK iterations (transport of the arrivals, push to the peers, wait for theirs, unpack) and looks
// at the device-resident loop state once per chunk. An iteration after the one in which no rank sent anything
// finds an empty range and a set `done` flag and costs a few empty launches.
static constexpr int P2P_CHUNK = 12;
static jb_status transport_iterations_p2p(jb_ctx *c, double t_start, double dt, float *ms_tr) {
  k_p2p_begin<<<1, 1, 0, c->stream>>>(c->d_p2p_state, c->n, c->capacity);
  c->launches += 1;
  const bool ddmc = c->cfg.use_ddmc != 0;
  int queued = 0;
  for (;;) {
    int in_chunk = 0;
    for (int k = 0; k < P2P_CHUNK && queued < c->cfg.max_transport_iterations; ++k, ++queued, ++in_chunk) {
      NvtxRange loop_range("Jaybenne::TransportLoop");
      StepParams P = make_params(c);
      if (c->props_dirty) {
        launch_pack_props(P, c->stream);
        c->launches += 1;
        c->props_dirty = false;
      }
      c->census_valid = false;
      JB_CUDA(c, cudaMemsetAsync(P.work, 0, sizeof(unsigned long long), c->stream));
      cudaEventRecord(c->it_ev[2 * k], c->stream);
      if (queued == 0) {
        launch_transport(P, 0, c->n, t_start, dt, c->stream);  // the whole pool: the host knows its size
      } else {
        P.range = c->d_p2p_state->range;  // the arrivals: (first, n) on the device
        launch_transport(P, 0, c->capacity, t_start, dt, c->stream);
      }
      cudaEventRecord(c->it_ev[2 * k + 1], c->stream);
      tl_mark(c, JB_TL_TRANSPORT);
      P.range = nullptr;
      launch_p2p_push(P, c->d_p2p_state, c->d_peers, c->n_ranks, c->p2p_inbox_cap, c->stream);
      launch_p2p_wait_unpack(P, c->d_p2p_state, c->d_mail, c->d_p2p_inbox[0], c->d_p2p_inbox[1], c->n_ranks,
                             c->p2p_inbox_cap, t_start + dt, ddmc, c->stream);
      c->launches += 6 + (ddmc ? 1 : 0);
      tl_mark(c, JB_TL_EXCHANGE);
    }
    JB_CUDA(c, cudaMemcpyAsync(c->h_p2p_state, c->d_p2p_state, sizeof(P2PState), cudaMemcpyDeviceToHost, c->stream));
    JB_CUDA(c, cudaStreamSynchronize(c->stream));
    JB_CUDA(c, cudaGetLastError());
    for (int k = 0; k < in_chunk; ++k) {
      float ms = 0.f;
      cudaEventElapsedTime(&ms, c->it_ev[2 * k], c->it_ev[2 * k + 1]);
      *ms_tr += ms;
    }
    const P2PState &h = *c->h_p2p_state;
    c->n = h.n;
    c->stats.n_iterations = h.iterations;
    if (h.error & 4) return fail(c, JB_ERR_COMM, "a peer did not deliver its records (direct peer exchange timed out)");
    if (h.error & 1) return fail(c, JB_ERR_CAPACITY, "exchange inbox overflow: raise jb_set_capacity");
    if (h.error & 2) return fail(c, JB_ERR_CAPACITY, "photon pool capacity exceeded on receive");
    if (h.done) return JB_OK;
    if (queued >= c->cfg.max_transport_iterations) return fail(c, JB_ERR_ITERATIONS, "max_transport_iterations exhausted");
  }
}

// RadiationStep(...).Execute() (jaybenne.cpp:68-151)
jb_status jb_radiation_step(jb_ctx *c, double t_start, double dt, jb_cycle_stats *stats) {
  JB_REQUIRE_MESH(c);
  if (c->n_ranks > 1 && !c->comm)
    return fail(c, JB_ERR_COMM, "multi-rank step needs jb_comm_init (or drive the tasks and jb_exchange_* from the host)");
  NvtxRange timestep_range("Jaybenne::Timestep");
  cudaEventRecord(c->ev[0], c->stream);
  c->tl_n = 0;
  tl_mark(c, JB_TL_BEGIN);
  jb_status s = JB_OK;
  // a fixed-size pool that also has to take arrivals from other ranks starts the cycle dense (the
  // exchange cannot squeeze the pool between transport iterations)
  if (c->n_ranks > 1 && c->requested_capacity > 0 && c->n_holes > 0) {
    s = ensure_dense(c);
    if (s != JB_OK) return s;
  }
  s = stats_begin(c);
  if (s != JB_OK) return s;
  s = jb_update_derived_transport_fields(c, dt);
  if (s != JB_OK) return s;
  tl_mark(c, JB_TL_FIELDS);
  s = jb_source_photons(c, JB_SOURCE_EMISSION, t_start, dt);
  if (s != JB_OK) return s;
  tl_mark(c, JB_TL_SOURCE);
  bool done = false;
  long long first = 0;
  float ms_tr = 0.f;
  if (c->n_ranks > 1 && !c->p2p_tried) {
    s = p2p_setup(c);
    if (s != JB_OK) return s;
  }
  if (c->n_ranks > 1 && c->p2p_on && transport_takes_device_range(make_params(c))) {
    c->exchange_mode = 2;
    s = transport_iterations_p2p(c, t_start, dt, &ms_tr);
    if (s != JB_OK) return s;
    s = sync_stats(c);
    if (s != JB_OK) return s;
    done = true;
  }
  for (int it = 1; !done && it <= c->cfg.max_transport_iterations; ++it) {
    c->stats.n_iterations = it;
    NvtxRange loop_range("Jaybenne::TransportLoop");
    cudaEventRecord(c->ev[2], c->stream);
    s = transport_range(c, first, c->n - first, t_start, dt);
    if (s != JB_OK) return s;
    cudaEventRecord(c->ev[3], c->stream);
    tl_mark(c, JB_TL_TRANSPORT);
    if (c->n_ranks == 1) {
      s = sync_stats(c);
      if (s != JB_OK) return s;
      float ms = 0.f;
      cudaEventElapsedTime(&ms, c->ev[2], c->ev[3]);
      ms_tr += ms;
      done = true;
      break;
    }
    const long long n_before = c->n;
    long long total_sent = 0;
    c->exchange_mode = 1;
    s = exchange_nccl(c, t_start + dt, &total_sent);
    if (s != JB_OK) return s;
    tl_mark(c, JB_TL_EXCHANGE);
    {
      float ms = 0.f;
      cudaEventElapsedTime(&ms, c->ev[2], c->ev[3]);
      ms_tr += ms;
    }
    if (total_sent == 0) {
      s = sync_stats(c);
      if (s != JB_OK) return s;
      done = true;
      break;
    }
    first = n_before;  // only the arrivals still have t < t_end
  }
  fold_stats(c);
  s = check_device_errors(c);
  if (s != JB_OK) return s;
  if (!done) return fail(c, JB_ERR_ITERATIONS, "max_transport_iterations exhausted");
  tl_mark(c, JB_TL_STATS);
  s = compact(c, (long long)c->h_stats->n_dead);
  if (s != JB_OK) return s;
  tl_mark(c, JB_TL_SQUEEZE);
  s = jb_evaluate_radiation_energy(c);
  if (s != JB_OK) return s;
  tl_mark(c, JB_TL_TALLY);
  s = jb_update_fluid(c);
  if (s != JB_OK) return s;
  if (c->cfg.sort_interval > 0 && (++c->cycles_since_sort) >= c->cfg.sort_interval) {
    c->cycles_since_sort = 0;
    s = sort_by_cell(c);
    if (s != JB_OK) return s;
  }
  cudaEventRecord(c->ev[1], c->stream);
  tl_mark(c, JB_TL_END);
  JB_CUDA(c, cudaStreamSynchronize(c->stream));
  tl_finish(c);
  float ms_total = 0.f;
  cudaEventElapsedTime(&ms_total, c->ev[0], c->ev[1]);
  c->stats.ms_transport = ms_tr;
  c->stats.ms_total = ms_total;
  if (stats) {
    long long na = 0;
    double e = 0.0;
    s = census(c, 0.0, &na, nullptr, &e);
    if (s != JB_OK) return s;
    c->stats.n_end = na;
    c->stats.e_end = e;
    *stats = c->stats;
  }
  return JB_OK;
}

// ---- swarm access ------------------------------------------------------------------------
jb_status jb_num_particles(jb_ctx *c, int64_t *n_local) {
  JB_REQUIRE_MESH(c);
  jb_status sd = ensure_dense(c);
  if (sd != JB_OK) return sd;
  if (n_local) *n_local = c->n;
  return JB_OK;
}

jb_status jb_get_particles(jb_ctx *c, const jb_particles *d, int64_t capacity, int64_t *n_out) {
  JB_REQUIRE_MESH(c);
  if (!d || !n_out) return JB_ERR_INVALID;
  {
    jb_status sd = ensure_dense(c);
    if (sd != JB_OK) return sd;
  }
  const long long n = std::min<long long>(c->n, capacity);
  *n_out = n;
  if (n <= 0) return JB_OK;
  JB_CUDA(c, cudaStreamSynchronize(c->stream));
  struct { double *dst; const double *src; } f64[9] = {
      {d->x, c->s.x}, {d->y, c->s.y}, {d->z, c->s.z}, {d->time, c->s.t}, {d->weight, c->s.w},
      {d->energy, c->s.e}, {d->v0, c->s.vx}, {d->v1, c->s.vy}, {d->v2, c->s.vz}};
  for (auto &a : f64)
    if (a.dst) JB_CUDA(c, cudaMemcpy(a.dst, a.src, sizeof(double) * n, cudaMemcpyDeviceToHost));
  if (d->i || d->j || d->k || d->block) {
    std::vector<int2> loc((size_t)n);
    JB_CUDA(c, cudaMemcpy(loc.data(), c->s.loc, sizeof(int2) * n, cudaMemcpyDeviceToHost));
    const int n1 = c->mesh.nxb[0], n2 = c->mesh.nxb[1];
    for (long long m = 0; m < n; ++m) {
      const int cell = loc[m].y;
      if (d->i) d->i[m] = cell % n1 + c->ng[0];
      if (d->j) d->j[m] = (cell / n1) % n2 + c->ng[1];
      if (d->k) d->k[m] = cell / (n1 * n2) + c->ng[2];
      if (d->block) d->block[m] = loc[m].x >= 0 ? c->local_gid[loc[m].x] : -1;
    }
  }
  if (d->rng) {
    std::vector<uint2> a((size_t)n), b((size_t)n);
    JB_CUDA(c, cudaMemcpy(a.data(), c->s.ida, sizeof(uint2) * n, cudaMemcpyDeviceToHost));
    JB_CUDA(c, cudaMemcpy(b.data(), c->s.idb, sizeof(uint2) * n, cudaMemcpyDeviceToHost));
    for (long long m = 0; m < n; ++m) {
      d->rng[4 * m + 0] = a[m].x;
      d->rng[4 * m + 1] = a[m].y;
      d->rng[4 * m + 2] = b[m].x;
      d->rng[4 * m + 3] = b[m].y;
    }
  }
  return JB_OK;
}

int32_t jb_exchange_mode(const jb_ctx *c) { return c ? c->exchange_mode : 0; }

jb_status jb_step_timeline(const jb_ctx *c, int32_t *labels, float *ms, int32_t cap, int32_t *n) {
  if (!c || !n || cap < 0 || (cap > 0 && (!labels || !ms))) return JB_ERR_INVALID;
  const int m = c->tl_done < cap ? c->tl_done : cap;
  for (int q = 0; q < m; ++q) {
    labels[q] = c->tl_label[q];
    ms[q] = c->tl_ms[q];
  }
  *n = c->tl_done;
  return JB_OK;
}

jb_status jb_set_sort_interval(jb_ctx *c, int32_t every_n_cycles) {
  if (!c || every_n_cycles < 0) return JB_ERR_INVALID;
  c->cfg.sort_interval = every_n_cycles;
  c->cycles_since_sort = 0;
  return JB_OK;
}

jb_status jb_reset_streams(jb_ctx *c) {
  if (!c) return JB_ERR_INVALID;
  c->source_call = 0;
  return JB_OK;
}

jb_status jb_set_particles(jb_ctx *c, const jb_particles *s, int64_t n) {
  JB_REQUIRE_MESH(c);
  if (n < 0 || (n > 0 && !s)) return JB_ERR_INVALID;
  c->n_holes = 0;
  c->census_valid = false;
  if (n == 0) {
    c->n = 0;
    return JB_OK;
  }
  if (n > 0 && (!s->x || !s->time || !s->weight || !s->energy || !s->v0 || !s->v1 || !s->v2 || !s->block))
    return fail(c, JB_ERR_INVALID, "x, time, weight, energy, v and block are required");
  c->n = 0;
  jb_status st = ensure_capacity(c, std::max<long long>(n, 1));
  if (st != JB_OK) return st;
  if (n == 0) return JB_OK;
  std::vector<double> zeros;
  auto up = [&](double *dst, const double *src) -> cudaError_t {
    if (!src) {
      zeros.assign((size_t)n, 0.0);
      src = zeros.data();
    }
    return cudaMemcpy(dst, src, sizeof(double) * n, cudaMemcpyHostToDevice);
  };
  JB_CUDA(c, up(c->s.x, s->x)); JB_CUDA(c, up(c->s.y, s->y)); JB_CUDA(c, up(c->s.z, s->z));
  JB_CUDA(c, up(c->s.t, s->time)); JB_CUDA(c, up(c->s.w, s->weight)); JB_CUDA(c, up(c->s.e, s->energy));
  JB_CUDA(c, up(c->s.vx, s->v0)); JB_CUDA(c, up(c->s.vy, s->v1)); JB_CUDA(c, up(c->s.vz, s->v2));
  std::vector<int2> loc((size_t)n);
  std::vector<uint2> a((size_t)n), b((size_t)n);
  for (long long m = 0; m < n; ++m) {
    const int gid = s->block[m];
    if (gid < 0 || gid >= c->mesh.nblocks || c->gid_to_local[gid] < 0) return fail(c, JB_ERR_INVALID, "block not local");
    const jb_block &B = c->all_blocks[gid];
    int ijk[3] = {0, 0, 0};
    const double pos[3] = {s->x[m], s->y ? s->y[m] : 0.0, s->z ? s->z[m] : 0.0};
    for (int d = 0; d < c->mesh.ndim; ++d) {
      // the DBlock's own dx (set_mesh) so the index is what Xtoijk returns on the device; the transport
      // kernels trust the stored cell, so a position on or outside the block edge is refused here
      const double dxd = (B.xmax[d] - B.xmin[d]) / c->mesh.nxb[d];
      ijk[d] = static_cast<int>(std::floor((pos[d] - B.xmin[d]) / dxd));
      if (ijk[d] < 0 || ijk[d] >= c->mesh.nxb[d])
        return fail(c, JB_ERR_INVALID, "jb_set_particles: particle " + std::to_string(m) + " lies outside block " +
                                           std::to_string(gid) + " (or on its upper edge)");
    }
    loc[m] = make_int2(c->gid_to_local[gid], (ijk[2] * c->mesh.nxb[1] + ijk[1]) * c->mesh.nxb[0] + ijk[0]);
    a[m] = s->rng ? make_uint2(s->rng[4 * m], s->rng[4 * m + 1]) : make_uint2((uint32_t)m, (uint32_t)gid);
    b[m] = s->rng ? make_uint2(s->rng[4 * m + 2], s->rng[4 * m + 3]) : make_uint2(0xffffffffu, 0u);
  }
  JB_CUDA(c, cudaMemcpy(c->s.loc, loc.data(), sizeof(int2) * n, cudaMemcpyHostToDevice));
  JB_CUDA(c, cudaMemcpy(c->s.ida, a.data(), sizeof(uint2) * n, cudaMemcpyHostToDevice));
  JB_CUDA(c, cudaMemcpy(c->s.idb, b.data(), sizeof(uint2) * n, cudaMemcpyHostToDevice));
  c->n = n;
  return JB_OK;
}

// ---- event log ---------------------------------------------------------------------------
jb_status jb_event_log_enable(jb_ctx *c, int64_t capacity) {
  if (!c) return JB_ERR_INVALID;
  JB_CUDA(c, cudaSetDevice(c->device));
  dfree(c->d_events);
  c->d_events = nullptr;
  c->event_cap = 0;
  if (capacity > 0) {
    JB_CUDA(c, dalloc(&c->d_events, capacity));
    c->event_cap = capacity;
  }
  JB_CUDA(c, cudaMemset(&c->d_stats->n_events, 0, sizeof(unsigned long long)));
  return JB_OK;
}
jb_status jb_event_log_read(jb_ctx *c, jb_event *dst, int64_t capacity, int64_t *n) {
  if (!c || !n) return JB_ERR_INVALID;
  JB_CUDA(c, cudaSetDevice(c->device));
  JB_CUDA(c, cudaStreamSynchronize(c->stream));
  unsigned long long cnt = 0;
  JB_CUDA(c, cudaMemcpy(&cnt, &c->d_stats->n_events, sizeof(cnt), cudaMemcpyDeviceToHost));
  *n = (int64_t)cnt;
  const long long m = std::min<long long>(std::min<long long>(capacity, (long long)cnt), c->event_cap);
  if (m > 0 && dst) JB_CUDA(c, cudaMemcpy(dst, c->d_events, sizeof(jb_event) * m, cudaMemcpyDeviceToHost));
  return JB_OK;
}

// ---- exchange, host-driven flavours ------------------------------------------------------
int64_t jb_exchange_record_bytes(void) { return (int64_t)sizeof(ExRecord); }

static jb_status outbox_counts(jb_ctx *c) {
  JB_CUDA(c, cudaMemcpyAsync(c->h_counters + 8, c->out.count, sizeof(unsigned long long) * c->n_ranks,
                             cudaMemcpyDeviceToHost, c->stream));
  JB_CUDA(c, cudaStreamSynchronize(c->stream));
  return JB_OK;
}
jb_status jb_exchange_count(jb_ctx *c, int32_t peer, int64_t *n_records) {
  JB_REQUIRE_MESH(c);
  if (peer < 0 || peer >= c->n_ranks || !n_records) return JB_ERR_INVALID;
  *n_records = 0;
  if (!c->out.count) return JB_OK;
  jb_status s = outbox_counts(c);
  if (s != JB_OK) return s;
  *n_records = (int64_t)c->h_counters[8 + peer];
  return JB_OK;
}
jb_status jb_exchange_take_device(jb_ctx *c, int32_t peer, void **dev_ptr, int64_t *n_records) {
  JB_REQUIRE_MESH(c);
  if (peer < 0 || peer >= c->n_ranks || !n_records || !dev_ptr) return JB_ERR_INVALID;
  *n_records = 0;
  *dev_ptr = nullptr;
  if (!c->out.count || !c->out.rec) return JB_OK;
  jb_status s = outbox_counts(c);
  if (s != JB_OK) return s;
  *n_records = (int64_t)c->h_counters[8 + peer];
  *dev_ptr = c->out.rec + (size_t)peer * c->out.cap;
  JB_CUDA(c, cudaMemsetAsync(c->out.count + peer, 0, sizeof(unsigned long long), c->stream));
  return JB_OK;
}
jb_status jb_exchange_take(jb_ctx *c, int32_t peer, void *dst, int64_t capacity, int64_t *n_records) {
  void *p = nullptr;
  int64_t n = 0;
  if (!c || !n_records) return JB_ERR_INVALID;
  if (peer < 0 || peer >= c->n_ranks) return JB_ERR_INVALID;
  // look before taking so a short buffer does not lose the records
  jb_status s = jb_exchange_count(c, peer, &n);
  if (s != JB_OK) return s;
  if (n > capacity) return fail(c, JB_ERR_CAPACITY, "exchange buffer too small");
  s = jb_exchange_take_device(c, peer, &p, &n);
  if (s != JB_OK) return s;
  if (n > 0) JB_CUDA(c, cudaMemcpy(dst, p, sizeof(ExRecord) * n, cudaMemcpyDeviceToHost));
  *n_records = n;
  return JB_OK;
}
jb_status jb_exchange_give_device(jb_ctx *c, const void *dev_ptr, int64_t n_records) {
  JB_REQUIRE_MESH(c);
  if (n_records <= 0) return JB_OK;
  if (c->n + n_records > c->capacity && c->n_holes > 0) {
    jb_status sd = ensure_dense(c);
    if (sd != JB_OK) return sd;
  }
  if (c->n + n_records > c->capacity) {
    if (c->requested_capacity > 0 && c->n + n_records > c->requested_capacity)
      return fail(c, JB_ERR_CAPACITY, "photon pool capacity exceeded on receive");
    // grow the swarm arrays only: the outbox (records other peers have not taken yet, device pointers a host
    // program got from jb_exchange_take_device) and the inbox stay where they are
    jb_status s = c->out.rec ? grow_pool_only(c, c->n + n_records) : ensure_capacity(c, c->n + n_records);
    if (s != JB_OK) return s;
  }
  StepParams P = make_params(c);
  c->census_valid = false;
  launch_unpack(P, (const ExRecord *)dev_ptr, n_records, c->n, std::numeric_limits<double>::infinity(), c->stream);
  c->launches += 1;
  c->n += n_records;
  jb_status s = sync_stats(c);
  if (s != JB_OK) return s;
  fold_stats(c);
  if (c->h_stats->n_error) return fail(c, JB_ERR_COMM, "record for a block that is not local");
  return JB_OK;
}
jb_status jb_exchange_give(jb_ctx *c, const void *src_host, int64_t n_records) {
  JB_REQUIRE_MESH(c);
  if (n_records <= 0) return JB_OK;
  ExRecord *stage = nullptr;
  JB_CUDA(c, dalloc(&stage, n_records));
  cudaError_t e = cudaMemcpy(stage, src_host, sizeof(ExRecord) * n_records, cudaMemcpyHostToDevice);
  jb_status s = e == cudaSuccess ? jb_exchange_give_device(c, stage, n_records)
                                 : fail(c, JB_ERR_CUDA, cudaGetErrorString(e));
  cudaStreamSynchronize(c->stream);
  cudaFree(stage);
  return s;
}
jb_status jb_exchange_clear(jb_ctx *c) {
  JB_REQUIRE_MESH(c);
  if (c->out.count) JB_CUDA(c, cudaMemsetAsync(c->out.count, 0, sizeof(unsigned long long) * c->n_ranks, c->stream));
  return JB_OK;
}

// ---- NCCL communicator -------------------------------------------------------------------
jb_status jb_comm_unique_id(void *id128) {
  if (!id128) return JB_ERR_INVALID;
  NcclApi &nc = nccl_api();
  if (!nc.ok) return JB_ERR_COMM;
  ncclUniqueId id;
  if (nc.GetUniqueId(&id) != ncclSuccess) return JB_ERR_COMM;
  static_assert(sizeof(ncclUniqueId) == 128, "ncclUniqueId is 128 bytes");
  memcpy(id128, &id, 128);
  return JB_OK;
}
jb_status jb_comm_init(jb_ctx *c, const void *id128, int32_t rank, int32_t n_ranks) {
  if (!c || !id128) return JB_ERR_INVALID;
  NcclApi &nc = nccl_api();
  if (!nc.ok) return fail(c, JB_ERR_COMM, "libnccl.so.2 could not be loaded");
  if (c->mesh_set && (rank != c->my_rank || n_ranks != c->n_ranks))
    return fail(c, JB_ERR_INVALID, "rank / n_ranks differ from jb_set_mesh");
  JB_CUDA(c, cudaSetDevice(c->device));
  ncclUniqueId id;
  memcpy(&id, id128, 128);
  ncclResult_t r = nc.CommInitRank(&c->comm, n_ranks, id, rank);
  if (r != ncclSuccess) return fail(c, JB_ERR_COMM, std::string("ncclCommInitRank: ") + (nc.GetErrorString ? nc.GetErrorString(r) : "?"));
  return JB_OK;
}

// CUDA-event timer on the library's own stream (torch.cuda.Event only sees torch's stream)
jb_status jb_timer_start(jb_ctx *c) {
  if (!c) return JB_ERR_INVALID;
  JB_CUDA(c, cudaSetDevice(c->device));
  if (!c->tev[0]) {
    JB_CUDA(c, cudaEventCreate(&c->tev[0]));
    JB_CUDA(c, cudaEventCreate(&c->tev[1]));
  }
  JB_CUDA(c, cudaEventRecord(c->tev[0], c->stream));
  return JB_OK;
}
jb_status jb_timer_stop(jb_ctx *c, double *ms) {
  if (!c || !ms || !c->tev[0]) return JB_ERR_INVALID;
  JB_CUDA(c, cudaSetDevice(c->device));
  JB_CUDA(c, cudaEventRecord(c->tev[1], c->stream));
  JB_CUDA(c, cudaEventSynchronize(c->tev[1]));
  float f = 0.f;
  JB_CUDA(c, cudaEventElapsedTime(&f, c->tev[0], c->tev[1]));
  *ms = f;
  return JB_OK;
}

// kernels launched by this context so far (bench.py reports it as gpu_launches)
int64_t jb_kernel_launches(const jb_ctx *c) { return c ? c->launches : 0; }

jb_status jb_device_count(int32_t *n_devices) {
  if (!n_devices) return JB_ERR_INVALID;
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess) {
    cudaGetLastError();
    n = 0;
  }
  *n_devices = n;
  return n > 0 ? JB_OK : JB_ERR_NO_DEVICE;
}

jb_status jb_set_device(int32_t ordinal) {
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess || n <= 0) {
    cudaGetLastError();
    return JB_ERR_NO_DEVICE;
  }
  if (ordinal < 0 || ordinal >= n) return JB_ERR_INVALID;
  return cudaSetDevice(ordinal) == cudaSuccess ? JB_OK : JB_ERR_CUDA;
}

jb_status jb_selftest_division(jb_ctx *c, int64_t n, int64_t *mismatches) {
  if (!c || !mismatches || n < 0) return JB_ERR_INVALID;
  cudaSetDevice(c->device);
  JB_CUDA(c, cudaMemsetAsync(c->d_counters, 0, sizeof(unsigned long long), c->stream));
  launch_selftest_division(n, c->d_counters, c->stream);
  c->launches += 1;
  JB_CUDA(c, cudaMemcpyAsync(c->h_counters, c->d_counters, sizeof(unsigned long long), cudaMemcpyDeviceToHost, c->stream));
  JB_CUDA(c, cudaStreamSynchronize(c->stream));
  *mismatches = (int64_t)c->h_counters[0];
  return JB_OK;
}

jb_status jb_measure_fp64_peak(jb_ctx *c, double *tflops) {
  if (!c || !tflops) return JB_ERR_INVALID;
  cudaSetDevice(c->device);
  int n_sm = 0;
  cudaDeviceGetAttribute(&n_sm, cudaDevAttrMultiProcessorCount, c->device);
  if (n_sm <= 0) n_sm = 148;
  const int blocks = n_sm * 8, iters = 1 << 16;
  cudaEvent_t e0, e1;
  JB_CUDA(c, cudaEventCreate(&e0));
  JB_CUDA(c, cudaEventCreate(&e1));
  launch_fp64_chain(c->d_esum, 1 << 10, blocks, c->stream);  // warm-up
  float best = 0.f;
  for (int rep = 0; rep < 3; ++rep) {
    cudaEventRecord(e0, c->stream);
    launch_fp64_chain(c->d_esum, iters, blocks, c->stream);
    cudaEventRecord(e1, c->stream);
    JB_CUDA(c, cudaEventSynchronize(e1));
    float ms = 0.f;
    cudaEventElapsedTime(&ms, e0, e1);
    if (rep == 0 || ms < best) best = ms;
  }
  c->launches += 4;
  cudaEventDestroy(e0);
  cudaEventDestroy(e1);
  const double flops = 16.0 * (double)iters * 256.0 * (double)blocks;
  *tflops = flops / ((double)best * 1e-3) / 1e12;
  return JB_OK;
}

}  // extern "C"
