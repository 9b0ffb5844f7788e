// jaybenne_b200.hpp — C++ host side above the C ABI (include/jaybenne_b200.h).
//
// Mirrors the interface the reference exposes to its driver application, so host code written
// against lanl/jaybenne reads the same here (paths relative to the reference @ 04947dfa):
//
//   jaybenne::Initialize(pin, opacity, scattering, eos)   src/jaybenne/jaybenne.hpp:50-52, jaybenne.cpp:158
//   jaybenne::InitializeRadiation(mbd, is_thermal)        jaybenne.hpp:76, jaybenne.cpp:570
//   jaybenne::RadiationStep(pmesh, t_start, dt)           jaybenne.hpp:72, jaybenne.cpp:68
//   the public tasks (TaskStatus f(MeshData*, ...))       jaybenne.hpp:59-69
//   parthenon::ParameterInput GetReal/GetOrAddReal/...    used at jaybenne.cpp:163-223, mcblock.cpp:40-137
//
// Parthenon's types (MeshData<Real>*, StateDescriptor, Swarm) do not exist here: `Package` holds
// what the StateDescriptor + the photons swarm + the package's cell fields are in the
// reference, all owned by the CUDA library behind a jb_ctx. Header-only, C++17, no CUDA
// headers: link with -ljaybenne_b200. Errors: the reference aborts via PARTHENON_REQUIRE/FAIL;
// here the same message is thrown as jaybenne_b200::Error (a host program that wants the
// abort lets it propagate to main()).
#ifndef JAYBENNE_B200_HPP_
#define JAYBENNE_B200_HPP_

#include <cstdint>
#include <cstdlib>
#include <limits>
#include <memory>
#include <stdexcept>
#include <string>
#include <utility>
#include <vector>

#include "jaybenne_b200.h"

namespace jaybenne_b200 {

using Real = double;  // parthenon::Real of a double-precision build

class Error : public std::runtime_error {
 public:
  Error(jb_status st, const std::string &msg) : std::runtime_error(msg), status(st) {}
  jb_status status;
};

// parthenon::TaskStatus (tasks/tasks.hpp); the tasks below return it like the reference's do
enum class TaskStatus { complete, incomplete, iterate };
// jaybenne::SourceType / SourceStrategy, src/jaybenne/jaybenne.hpp:55-56
enum class SourceType { thermal, emission };
enum class SourceStrategy { uniform, energy };

// ---------------------------------------------------------------------------------------
// parthenon::ParameterInput over the library's deck reader (csrc/host/deck.cpp)
// ---------------------------------------------------------------------------------------
class ParameterInput {
 public:
  ParameterInput() { check(jb_deck_parse("", &deck_), "cannot create an empty deck"); }
  static ParameterInput FromFile(const std::string &path) {
    ParameterInput p(nullptr);
    if (jb_deck_load(path.c_str(), &p.deck_) != JB_OK)
      throw Error(JB_ERR_INVALID, "cannot read input deck " + path);
    return p;
  }
  static ParameterInput FromString(const std::string &text) {
    ParameterInput p(nullptr);
    if (jb_deck_parse(text.c_str(), &p.deck_) != JB_OK) throw Error(JB_ERR_INVALID, "cannot parse input deck text");
    return p;
  }
  ParameterInput(ParameterInput &&o) noexcept : deck_(o.deck_) { o.deck_ = nullptr; }
  ParameterInput &operator=(ParameterInput &&o) noexcept {
    if (this != &o) {
      if (deck_) jb_deck_free(deck_);
      deck_ = o.deck_;
      o.deck_ = nullptr;
    }
    return *this;
  }
  ParameterInput(const ParameterInput &) = delete;
  ParameterInput &operator=(const ParameterInput &) = delete;
  ~ParameterInput() {
    if (deck_) jb_deck_free(deck_);
  }

  bool DoesParameterExist(const std::string &block, const std::string &name) const {
    return jb_deck_has(deck_, block.c_str(), name.c_str()) != 0;
  }
  // Parthenon's message for a missing required parameter
  void CheckRequired(const std::string &block, const std::string &name) const {
    if (!DoesParameterExist(block, name))
      throw Error(JB_ERR_INVALID, "Parameter name '" + name + "' not found in block '" + block + "'");
  }
  std::string GetString(const std::string &block, const std::string &name) const {
    CheckRequired(block, name);
    return jb_deck_get(deck_, block.c_str(), name.c_str());
  }
  Real GetReal(const std::string &block, const std::string &name) const { return atof(GetString(block, name).c_str()); }
  int64_t GetInteger(const std::string &block, const std::string &name) const {
    return atoll(GetString(block, name).c_str());
  }
  bool GetBoolean(const std::string &block, const std::string &name) const { return to_bool(GetString(block, name)); }
  std::string GetOrAddString(const std::string &block, const std::string &name, const std::string &dflt) {
    if (!DoesParameterExist(block, name)) SetString(block, name, dflt);
    return GetString(block, name);
  }
  Real GetOrAddReal(const std::string &block, const std::string &name, Real dflt) {
    return DoesParameterExist(block, name) ? GetReal(block, name) : dflt;
  }
  int64_t GetOrAddInteger(const std::string &block, const std::string &name, int64_t dflt) {
    return DoesParameterExist(block, name) ? GetInteger(block, name) : dflt;
  }
  bool GetOrAddBoolean(const std::string &block, const std::string &name, bool dflt) {
    return DoesParameterExist(block, name) ? GetBoolean(block, name) : dflt;
  }
  void SetString(const std::string &block, const std::string &name, const std::string &value) {
    check(jb_deck_set(deck_, block.c_str(), name.c_str(), value.c_str()), "cannot set " + block + "/" + name);
  }
  // command-line override "block/key=value" (parthenon's `mcblock -i deck block/key=value`,
  // what tst/regression_test.py passes as modified_inputs)
  void ModifyFromCmdline(const std::string &arg) {
    const size_t slash = arg.rfind('/', arg.find('=')), eq = arg.find('=');
    if (slash == std::string::npos || eq == std::string::npos || slash > eq)
      throw Error(JB_ERR_INVALID, "cannot parse command-line override '" + arg + "' (want block/key=value)");
    SetString(arg.substr(0, slash), arg.substr(slash + 1, eq - slash - 1), arg.substr(eq + 1));
  }
  const jb_deck *deck() const { return deck_; }

 private:
  explicit ParameterInput(std::nullptr_t) {}
  static bool to_bool(std::string v) {
    for (auto &ch : v) ch = (char)tolower(ch);
    return v == "true" || v == "1" || v == "t" || v == "yes";
  }
  static void check(jb_status st, const std::string &what) {
    if (st != JB_OK) throw Error(st, what);
  }
  jb_deck *deck_ = nullptr;
};

// ---------------------------------------------------------------------------------------
// The block table Parthenon's Mesh would hold: every rank knows every block (gid order =
// Z-order, contiguous chunks per rank), built from <parthenon/mesh|meshblock|static_refinementN|swarm>.
// ---------------------------------------------------------------------------------------
class Mesh {
 public:
  Mesh(const ParameterInput &pin, int n_ranks = 1) {
    char err[512] = {0};
    const jb_status st = jb_deck_mesh(pin.deck(), n_ranks, &m_, err, (int32_t)sizeof(err));
    if (st != JB_OK) throw Error(st, err);
    n_ranks_ = n_ranks;
  }
  Mesh(const Mesh &) = delete;
  Mesh &operator=(const Mesh &) = delete;
  ~Mesh() { jb_mesh_free(&m_); }

  const jb_mesh &c() const { return m_; }
  int nbtotal() const { return m_.nblocks; }  // Mesh::nbtotal
  int ndim() const { return m_.ndim; }
  int n_ranks() const { return n_ranks_; }
  const jb_block &block(int gid) const { return m_.blocks[gid]; }
  int64_t cells_per_block() const { return (int64_t)m_.nxb[0] * m_.nxb[1] * m_.nxb[2]; }
  std::vector<int> local_gids(int rank) const {
    std::vector<int> g;
    for (int b = 0; b < m_.nblocks; ++b)
      if (m_.blocks[b].rank == rank) g.push_back(m_.blocks[b].gid);
    return g;
  }
  // coords.Xc<dir>(i) of interior cell i (0-based) as UniformCartesian computes it from the
  // ghost-inclusive origin: xmin_with_ghosts + (i + ng + 0.5) * dx  (mcblock.cpp:191)
  Real Xc(int gid, int dir, int i) const {
    const jb_block &b = m_.blocks[gid];
    const int n = m_.nxb[dir], ng = dir < m_.ndim ? m_.nghost : 0;
    const Real dx = (b.xmax[dir] - b.xmin[dir]) / n;
    const Real x0 = b.xmin[dir] - ng * dx;
    return x0 + ((i + ng) + 0.5) * dx;
  }

 private:
  jb_mesh m_{};
  int n_ranks_ = 1;
};

// ---------------------------------------------------------------------------------------
// The "jaybenne" package as the driver sees it after jaybenne::Initialize: parameters
// (StateDescriptor Params, jaybenne.cpp:163-233) + the photons swarm + the package's cell
// fields, all resident on ONE GPU behind a jb_ctx. One Package per rank.
// ---------------------------------------------------------------------------------------
class Package {
 public:
  Package(const jb_config &cfg, const Mesh &mesh, int rank = 0, int n_ranks = 1, int64_t capacity = 0)
      : cfg_(cfg), mesh_(&mesh), rank_(rank), n_ranks_(n_ranks) {
    const jb_status st = jb_create(&cfg_, &ctx_);
    if (st != JB_OK) {
      ctx_ = nullptr;
      throw Error(st, st == JB_ERR_NO_DEVICE ? "jb_create: no CUDA device (jaybenne_b200 has no CPU path)"
                                              : "jb_create failed (invalid <jaybenne> parameters)");
    }
    check(jb_set_mesh(ctx_, &mesh.c(), rank, n_ranks));
    if (capacity > 0) check(jb_set_capacity(ctx_, capacity));
    local_gids_ = mesh.local_gids(rank);
  }
  Package(const Package &) = delete;
  Package &operator=(const Package &) = delete;
  ~Package() {
    if (ctx_) jb_destroy(ctx_);
  }

  // pkg->Param<T>("name") for the parameters the driver reads (jaybenne.cpp:176-223)
  const jb_config &params() const { return cfg_; }
  Real speed_of_light() const { return constants().first; }
  Real stefan_boltzmann() const { return constants().second; }
  const Mesh &mesh() const { return *mesh_; }
  int rank() const { return rank_; }
  int n_ranks() const { return n_ranks_; }
  const std::vector<int> &local_gids() const { return local_gids_; }
  jb_ctx *ctx() const { return ctx_; }

  // host material fields (JAYBENNE_HOST_*_VARIABLE) and the package's own fields; `src` may be
  // a host or a device pointer of FieldSize(field) doubles (interior cells, i fastest)
  int64_t FieldSize(int32_t field) const { return jb_field_size(ctx_, field); }
  void SetField(int gid, int32_t field, const Real *src) { check(jb_set_field(ctx_, gid, field, src)); }
  void GetField(int gid, int32_t field, Real *dst) const { check(jb_get_field(ctx_, gid, field, dst)); }
  std::vector<Real> GetField(int gid, int32_t field) const {
    std::vector<Real> v((size_t)FieldSize(field));
    GetField(gid, field, v.data());
    return v;
  }
  // every block of a pack in one call (one synchronisation): gids[i] pairs with ptrs[i]
  void SetFields(int32_t field, const std::vector<int> &gids, const std::vector<const Real *> &srcs) {
    std::vector<int32_t> g(gids.begin(), gids.end());
    check(jb_set_fields(ctx_, field, (int32_t)g.size(), g.data(), srcs.data()));
  }
  void GetFields(int32_t field, const std::vector<int> &gids, const std::vector<Real *> &dsts) const {
    std::vector<int32_t> g(gids.begin(), gids.end());
    check(jb_get_fields(ctx_, field, (int32_t)g.size(), g.data(), dsts.data()));
  }
  void SetRemoteField(int gid, int32_t field, const Real *src) { check(jb_set_remote_field(ctx_, gid, field, src)); }

  int64_t NumParticles() const {
    int64_t n = 0;
    check(jb_num_particles(ctx_, &n));
    return n;
  }
  jb_cycle_stats Stats() const {
    jb_cycle_stats s{};
    check(jb_get_stats(ctx_, &s));
    return s;
  }
  int64_t KernelLaunches() const { return jb_kernel_launches(ctx_); }
  // device-timed phases of the last RadiationStep in stream order: (JB_TL_* label, milliseconds)
  std::vector<std::pair<int, float>> StepTimeline() const {
    int32_t labels[96];
    float ms[96];
    int32_t n = 0;
    check(jb_step_timeline(ctx_, labels, ms, 96, &n));
    std::vector<std::pair<int, float>> out;
    for (int32_t q = 0; q < n && q < 96; ++q) out.emplace_back((int)labels[q], ms[q]);
    return out;
  }
  // 0 single rank, 1 NCCL send / recv per iteration, 2 direct peer stores over NVLink
  int ExchangeMode() const { return (int)jb_exchange_mode(ctx_); }

  // the `photons` swarm of this rank as host SoA (jaybenne.cpp:236-245) for dumps / restart
  struct Photons {
    std::vector<Real> x, y, z, time, weight, energy, v0, v1, v2;
    std::vector<int32_t> i, j, k, block;
    std::vector<uint32_t> rng;  // 4 words per photon: stream identity + draws consumed
    size_t size() const { return x.size(); }
  };
  Photons GetParticles() const {
    const size_t n = (size_t)NumParticles();
    Photons h;
    for (auto *v : {&h.x, &h.y, &h.z, &h.time, &h.weight, &h.energy, &h.v0, &h.v1, &h.v2}) v->resize(n);
    for (auto *v : {&h.i, &h.j, &h.k, &h.block}) v->resize(n);
    h.rng.resize(4 * n);
    jb_particles d{};
    d.x = h.x.data(); d.y = h.y.data(); d.z = h.z.data();
    d.time = h.time.data(); d.weight = h.weight.data(); d.energy = h.energy.data();
    d.v0 = h.v0.data(); d.v1 = h.v1.data(); d.v2 = h.v2.data();
    d.i = h.i.data(); d.j = h.j.data(); d.k = h.k.data(); d.block = h.block.data();
    d.rng = h.rng.data();
    int64_t got = 0;
    check(jb_get_particles(ctx_, &d, (int64_t)n, &got));
    if ((size_t)got != n) throw Error(JB_ERR_STATE, "jb_get_particles returned a different count");
    return h;
  }

  // multi-rank: NCCL inside the library (replaces Parthenon's MPI swarm communication)
  static std::vector<unsigned char> CommUniqueId() {
    std::vector<unsigned char> id(128);
    if (jb_comm_unique_id(id.data()) != JB_OK) throw Error(JB_ERR_COMM, "jb_comm_unique_id failed");
    return id;
  }
  void CommInit(const std::vector<unsigned char> &id) { check(jb_comm_init(ctx_, id.data(), rank_, n_ranks_)); }

  void check(jb_status st) const {
    if (st != JB_OK) throw Error(st, jb_last_error(ctx_));
  }

 private:
  std::pair<Real, Real> constants() const {
    Real c = 0, sb = 0;
    check(jb_get_constants(ctx_, &c, &sb));
    return {c, sb};
  }
  jb_config cfg_;
  const Mesh *mesh_;
  int rank_, n_ranks_;
  jb_ctx *ctx_ = nullptr;
  std::vector<int> local_gids_;
};

// jaybenne::Initialize(pin, opacity, scattering, eos): the closures arrive as the numbers
// mcblock builds them from (<mcblock> block, mcblock.cpp:78-143) because both come from the
// same deck; errors mirror the PARTHENON_REQUIRE / FAIL cases of jaybenne.cpp:171-214.
inline jb_config ConfigFromInput(const ParameterInput &pin) {
  jb_config cfg;
  char err[512] = {0};
  const jb_status st = jb_deck_config(pin.deck(), &cfg, err, (int32_t)sizeof(err));
  if (st != JB_OK) throw Error(st, err);
  return cfg;
}
inline std::shared_ptr<Package> Initialize(const ParameterInput &pin, const Mesh &mesh, int rank = 0, int n_ranks = 1,
                                           int64_t capacity = 0) {
  return std::make_shared<Package>(ConfigFromInput(pin), mesh, rank, n_ranks, capacity);
}

// ---- the public tasks, src/jaybenne/jaybenne.hpp:59-76 (one call covers every local block,
// like one MeshData partition per rank) -----------------------------------------------------
inline void InitializeRadiation(Package &p, bool is_thermal) { p.check(jb_initialize_radiation(p.ctx(), is_thermal)); }
inline TaskStatus UpdateDerivedTransportFields(Package &p, Real dt) {
  p.check(jb_update_derived_transport_fields(p.ctx(), dt));
  return TaskStatus::complete;
}
template <SourceType ST>
inline TaskStatus SourcePhotons(Package &p, Real t_start, Real dt) {
  p.check(jb_source_photons(p.ctx(), ST == SourceType::thermal ? JB_SOURCE_THERMAL : JB_SOURCE_EMISSION, t_start, dt));
  return TaskStatus::complete;
}
// TransportPhotons / TransportPhotons_DDMC: the library picks by use_ddmc as RadiationStep does
inline TaskStatus TransportPhotons(Package &p, Real t_start, Real dt) {
  p.check(jb_transport_photons(p.ctx(), t_start, dt));
  return TaskStatus::complete;
}
// MeshResetCommunication + MeshSend + MeshReceive (jaybenne.cpp:26-61); collective after CommInit
inline TaskStatus MeshSendReceive(Package &p) {
  p.check(jb_swarm_exchange(p.ctx()));
  return TaskStatus::complete;
}
inline TaskStatus SampleDDMCBlockFace(Package &p) {
  p.check(jb_sample_ddmc_block_face(p.ctx()));
  return TaskStatus::complete;
}
// iterate while any rank still holds a photon with time < t_end (transport.cpp:187-216)
inline TaskStatus CheckCompletion(Package &p, Real t_end) {
  int64_t unfinished = 0;
  p.check(jb_check_completion(p.ctx(), t_end, &unfinished));
  return unfinished == 0 ? TaskStatus::complete : TaskStatus::iterate;
}
inline TaskStatus EvaluateRadiationEnergy(Package &p) {
  p.check(jb_evaluate_radiation_energy(p.ctx()));
  return TaskStatus::complete;
}
inline TaskStatus UpdateFluid(Package &p) {
  p.check(jb_update_fluid(p.ctx()));
  return TaskStatus::complete;
}
// EstimateTimestepMesh returns <jaybenne> dt unchanged (jaybenne.cpp:271-275)
inline Real EstimateTimestepMesh(const Package &p) { return p.params().dt; }

// RadiationStep(pmesh, t_start, dt).Execute() in one library call: the whole cycle stays on the
// device (sourcing, the transport / exchange iterations, tallies, fluid update).
inline jb_cycle_stats RadiationStep(Package &p, Real t_start, Real dt) {
  jb_cycle_stats st{};
  p.check(jb_radiation_step(p.ctx(), t_start, dt, &st));
  return st;
}

// The same cycle sequenced task by task from the host, in the order of the reference's task
// list (jaybenne.cpp:104-138): for hosts that interleave their own tasks. Produces the same
// swarm and fields as RadiationStep (tests/test_cpp_host.py checks bit-identity).
inline jb_cycle_stats RadiationStepTaskList(Package &p, Real t_start, Real dt) {
  p.check(jb_stats_begin(p.ctx()));
  UpdateDerivedTransportFields(p, dt);
  SourcePhotons<SourceType::emission>(p, t_start, dt);
  int it = 0;
  for (;;) {
    if (++it > p.params().max_transport_iterations)
      throw Error(JB_ERR_ITERATIONS, "max_transport_iterations exhausted");  // AddSublist {1, max}, jaybenne.cpp:113
    TransportPhotons(p, t_start, dt);
    MeshSendReceive(p);
    if (p.params().use_ddmc) SampleDDMCBlockFace(p);
    if (CheckCompletion(p, t_start + dt) == TaskStatus::complete) break;
  }
  EvaluateRadiationEnergy(p);
  UpdateFluid(p);
  jb_cycle_stats st = p.Stats();
  st.n_iterations = it;
  return st;
}

// mcblock's cell kernels either side of the step (src/mcblock/mcblock.cpp:208-262)
inline void UpdateDerived(Package &p) { p.check(jb_update_derived_material(p.ctx())); }
inline void PostInitialization(Package &p) { p.check(jb_post_initialization(p.ctx())); }

}  // namespace jaybenne_b200
#endif  // JAYBENNE_B200_HPP_
