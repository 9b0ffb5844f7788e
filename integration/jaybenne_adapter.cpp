// jaybenne_adapter.cpp — jaybenne's Parthenon package interface implemented on libjaybenne_b200.so.
//
// Drop-in for src/jaybenne/jaybenne.cpp (+ sourcing.cpp, transport.cpp, transport_ddmc.cpp,
// sample_ddmc_bface.cpp): it defines exactly the symbols src/jaybenne/jaybenne.hpp:50-76 declares —
// Initialize, RadiationStep, InitializeRadiation, EstimateTimestepMesh, the public tasks and
// PhotonReflectBC — so mcblock (src/mcblock/mcblock.cpp:267-297, mcblock_driver.cpp:38-74) or any other
// Parthenon host links against it unchanged. The photons never enter Parthenon: the library owns the swarm
// on the GPU; the adapter moves the package parameters, the block table and the host's material fields
// across the C ABI (include/jaybenne_b200.h) and brings energy_tally / the updated internal energy back.
//
// Build on the reference side: replace the five sources above by this file in src/jaybenne/CMakeLists.txt,
// add -I<jaybenne_b200>/include and link -ljaybenne_b200. Here, where Parthenon is absent, the file is
// type-checked against declaration-only stand-ins (integration/mock/, tests/test_adapter.py:
// g++ -std=c++17 -fsyntax-only -Iintegration/mock -Iinclude).
//
// Host variables (jaybenne_config.hpp.in:19-30): HOST_DENSITY, HOST_SPECIFIC_INTERNAL_ENERGY,
// HOST_UPDATE_ENERGY name the fields read and updated, exactly as in the reference.
#include <cmath>
#include <cstdint>
#include <cstring>
#include <limits>
#include <memory>
#include <string>
#include <vector>

#include <parthenon/driver.hpp>
#include <parthenon/package.hpp>

using namespace parthenon;
using namespace parthenon::driver::prelude;
using namespace parthenon::package::prelude;

#include "jaybenne_config.hpp"

#include <jaybenne_b200.h>

namespace jaybenne {

enum class SourceStrategy { uniform, energy };
enum class SourceType { thermal, emission };
static const char *photons_swarm_name = "photons";  // jaybenne_variables.hpp:23

// ---------------------------------------------------------------------------------------------------
// adapter state: one context per rank (= per GPU), kept as a package Param like the reference's RngPool
// ---------------------------------------------------------------------------------------------------
struct AdapterState {
  jb_ctx *ctx = nullptr;
  jb_config cfg;
  bool mesh_pushed = false;
  int init_calls = 0;                 // InitializeRadiation is called once per local block
  std::vector<int32_t> gids;          // local blocks in block_list order
  ParArray1D<Real> stage[3];          // packed interior cells of density / sie / internal energy, all local blocks
  ParArray1D<Real> stage_tally;
  int64_t ncell = 0;                  // interior cells per block
  int32_t swarm_bc[6] = {0, 0, 0, 0, 0, 0};  // <parthenon/swarm> ix1_bc ... ox3_bc as JB_BC_*
  ~AdapterState() {
    if (ctx) jb_destroy(ctx);
  }
};
using StatePtr = std::shared_ptr<AdapterState>;

static StatePtr GetState(Mesh *pmesh) { return pmesh->packages.Get("jaybenne")->Param<StatePtr>("b200_state"); }

static void Check(AdapterState &s, jb_status st, const char *what) {
  if (st != JB_OK) {
    // the reference aborts through PARTHENON_FAIL with the same message texts (e.g. sourcing.cpp:38-39)
    PARTHENON_FAIL(std::string(what) + ": " + (s.ctx ? jb_last_error(s.ctx) : "no context"));
  }
}

static int32_t SwarmBC(ParameterInput *pin, const char *key) {
  // <parthenon/swarm> ix1_bc ... (inputs/stepdiff.in:44-51): periodic | outflow | jaybenne_reflecting
  const std::string v = pin->GetOrAddString("parthenon/swarm", key, "periodic");
  if (v == "periodic") return JB_BC_PERIODIC;
  if (v == "outflow") return JB_BC_OUTFLOW;
  if (v == "jaybenne_reflecting") return JB_BC_REFLECT;
  PARTHENON_FAIL("swarm boundary condition not supported by jaybenne_b200: " + v);
}

// ---------------------------------------------------------------------------------------------------
// Initialize (jaybenne.cpp:158-266): same <jaybenne> keys, defaults and checks; the three closures are
// identified by type and handed to the library as the numbers they were built from.
// ---------------------------------------------------------------------------------------------------
std::shared_ptr<StateDescriptor> Initialize(ParameterInput *pin, Opacity &opacity, Scattering &scattering, EOS &eos) {
  auto pkg = std::make_shared<StateDescriptor>("jaybenne");
  auto state = std::make_shared<AdapterState>();
  jb_config &cfg = state->cfg;
  jb_config_default(&cfg);

  cfg.num_particles = pin->GetInteger("jaybenne", "num_particles");
  pkg->AddParam<>("num_particles", static_cast<int>(cfg.num_particles));
  cfg.dt = pin->GetOrAddReal("jaybenne", "dt", std::numeric_limits<Real>::max());
  pkg->AddParam<>("dt", cfg.dt);
  cfg.min_swarm_occupancy = pin->GetOrAddReal("jaybenne", "min_swarm_occupancy", 0.);
  PARTHENON_REQUIRE(cfg.min_swarm_occupancy >= 0 && cfg.min_swarm_occupancy < 1.0,
                    "Minimum allowable swarm occupancy must be >= 0 and less than 1");
  pkg->AddParam<>("min_swarm_occupancy", cfg.min_swarm_occupancy);
  cfg.numin = pin->GetOrAddReal("jaybenne", "numin", std::numeric_limits<Real>::min());
  cfg.numax = pin->GetOrAddReal("jaybenne", "numax", std::numeric_limits<Real>::max());
  pkg->AddParam<>("numin", cfg.numin);
  pkg->AddParam<>("numax", cfg.numax);

  const auto units = opacity.GetRuntimePhysicalConstants();
  pkg->AddParam<>("speed_of_light", units.c);
  pkg->AddParam<>("stefan_boltzmann", units.sb);

  cfg.unique_rank_seeds = pin->GetOrAddBoolean("jaybenne", "unique_rank_seeds", true);
  pkg->AddParam<>("unique_rank_seeds", static_cast<bool>(cfg.unique_rank_seeds));
  cfg.seed = pin->GetOrAddInteger("jaybenne", "seed", 123);
  // the reference seeds its pool with the raw seed whatever unique_rank_seeds says (jaybenne.cpp:187-192);
  // the per-photon streams here are keyed by the raw seed on every rank for the same reason
  pkg->AddParam<>("seed", cfg.unique_rank_seeds ? static_cast<int>(cfg.seed) + Globals::my_rank : static_cast<int>(cfg.seed));
  cfg.max_transport_iterations = pin->GetOrAddInteger("jaybenne", "max_transport_iterations", 10000);
  pkg->AddParam<>("max_transport_iterations", static_cast<int>(cfg.max_transport_iterations));
  cfg.use_ddmc = pin->GetOrAddBoolean("jaybenne", "use_ddmc", false);
  pkg->AddParam<>("use_ddmc", static_cast<bool>(cfg.use_ddmc));
  cfg.tau_ddmc = pin->GetOrAddReal("jaybenne", "tau_ddmc", 5.0);
  pkg->AddParam<>("tau_ddmc", cfg.tau_ddmc);

  SourceStrategy source_strategy;
  const std::string strategy = pin->GetOrAddString("jaybenne", "source_strategy", "uniform");
  if (strategy == "uniform") {
    source_strategy = SourceStrategy::uniform;
    cfg.source_strategy = JB_STRATEGY_UNIFORM;
  } else if (strategy == "energy") {
    source_strategy = SourceStrategy::energy;  // accepted here, fatal at sourcing (sourcing.cpp:38-39)
    cfg.source_strategy = JB_STRATEGY_ENERGY;
  } else {
    PARTHENON_FAIL("Only uniform or energy source strategies supported!");
  }
  pkg->AddParam<>("source_strategy", source_strategy);
  cfg.do_emission = pin->GetOrAddBoolean("jaybenne", "do_emission", true);
  pkg->AddParam<>("do_emission", static_cast<bool>(cfg.do_emission));
  cfg.do_feedback = pin->GetOrAddBoolean("jaybenne", "do_feedback", true);
  pkg->AddParam<>("do_feedback", static_cast<bool>(cfg.do_feedback));

  // ---- closures: which alternative of the variant, and the numbers it was built from ----------------
  // The variants do not expose their parameters, so the numbers come from the same input block the host
  // built them from (<mcblock>, mcblock.cpp:78-143; another host names its block in <jaybenne> closure_block)
  // and are then CHECKED against the closure objects actually passed in, at a probe state.
  namespace sp = singularity::photons;
  const std::string hb = pin->GetOrAddString("jaybenne", "closure_block", "mcblock");
  cfg.time_scale = units.time;
  cfg.mass_scale = units.mass;
  cfg.length_scale = units.length;
  cfg.temperature_scale = units.temp;
  PARTHENON_REQUIRE(eos.IsType<singularity::IdealGas>(), "jaybenne_b200 supports the IdealGas EOS");
  cfg.eos_gm1 = eos.GruneisenParamFromDensityTemperature(1.0, 1.0);
  cfg.eos_cv = eos.SpecificHeatFromDensityTemperature(1.0, 1.0);
  const Real rho_p = 1.25, temp_p = 0.75, nu_p = 1.0;  // probe state (code units)
  if (opacity.IsType<sp::NonCGSUnits<sp::Gray>>()) {
    cfg.opacity_model = JB_OPAC_GRAY;
    cfg.opacity_kappa = pin->DoesParameterExist(hb, "opacity_constant_value") ? pin->GetReal(hb, "opacity_constant_value") : 0.0;
    // NonCGSUnits<Gray>::AbsorptionCoefficient = length * (rho_unit * rho) * kappa up to the rounding of B_nu / B_nu
    const Real want = opacity.AbsorptionCoefficient(rho_p, temp_p, nu_p);
    const Real have = units.length * (units.mass / (units.length * units.length * units.length) * rho_p) * cfg.opacity_kappa;
    PARTHENON_REQUIRE(std::abs(want - have) <= 1e-12 * std::max(std::abs(want), std::abs(have)),
                      "gray opacity closure does not match <" + hb + "> opacity_constant_value");
  } else if (opacity.IsType<sp::NonCGSUnits<sp::EPBremss>>()) {
    cfg.opacity_model = JB_OPAC_EPBREMSS;
  } else {
    PARTHENON_FAIL("absorption closure not supported by jaybenne_b200 (Gray or EPBremss in NonCGSUnits)");
  }
  cfg.apm = pin->GetOrAddReal(hb, "apm", 1.);
  if (scattering.IsType<sp::NonCGSUnitsS<sp::GrayS>>()) {
    cfg.scattering_model = JB_SCAT_GRAY;
    cfg.scattering_sigma = pin->DoesParameterExist(hb, "scattering_constant_value") ? pin->GetReal(hb, "scattering_constant_value") : 0.0;
    const Real want = scattering.TotalScatteringCoefficient(rho_p, temp_p, nu_p);
    const Real have = units.length * ((units.mass / (units.length * units.length * units.length) * rho_p) / cfg.apm) * cfg.scattering_sigma;
    PARTHENON_REQUIRE(std::abs(want - have) <= 1e-12 * std::max(std::abs(want), std::abs(have)),
                      "gray scattering closure does not match <" + hb + "> scattering_constant_value / apm");
  } else if (scattering.IsType<sp::NonCGSUnitsS<sp::ThomsonS>>()) {
    cfg.scattering_model = JB_SCAT_THOMSON;
  } else {
    PARTHENON_FAIL("scattering closure not supported by jaybenne_b200 (GrayS or ThomsonS in NonCGSUnitsS)");
  }
  // kept as Params under the reference's names for hosts that read them back
  pkg->AddParam<>("eos_d", eos.GetOnDevice());
  pkg->AddParam<>("opacity_d", opacity.GetOnDevice());
  pkg->AddParam<>("scattering_d", scattering.GetOnDevice());

  // which domain faces reflect / wrap / let photons out: the host registers PhotonReflectBC under the name
  // "jaybenne_reflecting" (mcblock.cpp:271-282) and selects it per face in <parthenon/swarm>
  {
    const char *keys[6] = {"ix1_bc", "ox1_bc", "ix2_bc", "ox2_bc", "ix3_bc", "ox3_bc"};
    for (int f = 0; f < 6; ++f) state->swarm_bc[f] = SwarmBC(pin, keys[f]);
  }

  const jb_status st = jb_create(&cfg, &state->ctx);
  if (st == JB_ERR_NO_DEVICE) PARTHENON_FAIL("jaybenne_b200: no CUDA device (the library has no CPU path)");
  PARTHENON_REQUIRE(st == JB_OK, "jb_create failed");
  double c_code = 0., sb_code = 0.;
  Check(*state, jb_get_constants(state->ctx, &c_code, &sb_code), "jb_get_constants");
  PARTHENON_REQUIRE(std::abs(c_code - units.c) <= 1e-14 * units.c && std::abs(sb_code - units.sb) <= 1e-14 * units.sb,
                    "unit system of the library differs from the closures'");
  pkg->AddParam<>("b200_state", state);

  // The swarm stays registered (outputs / restarts may name it) but holds no particles: the photons live in
  // the library. Fields: what the host reads (energy_tally) and what outputs may name.
  Metadata swarm_metadata({Metadata::Provides, Metadata::None});
  pkg->AddSwarm(photons_swarm_name, swarm_metadata);
  Metadata m({Metadata::Cell, Metadata::Independent});
  pkg->AddField("field.jaybenne.energy_tally", m);
  pkg->AddField("field.jaybenne.fleck_factor", m);
  Metadata m_onecopy({Metadata::Cell, Metadata::OneCopy});
  pkg->AddField("field.jaybenne.energy_delta", m_onecopy);

  pkg->EstimateTimestepMesh = [](MeshData<Real> *md) {
    return md->GetParentPointer()->packages.Get("jaybenne")->Param<Real>("dt");  // jaybenne.cpp:271-275
  };
  return pkg;
}

Real EstimateTimestepMesh(MeshData<Real> *md) {
  return md->GetParentPointer()->packages.Get("jaybenne")->Param<Real>("dt");
}

// ---------------------------------------------------------------------------------------------------
// mesh: one jb_block per GLOBAL block (every rank passes the whole table, like every Parthenon rank holds
// the whole loclist); gids are assigned to ranks in contiguous runs, as Parthenon's load balancer does.
// ---------------------------------------------------------------------------------------------------
static void PushMesh(Mesh *pmesh, AdapterState &s) {
  const auto &loclist = pmesh->GetLocList();
  const RegionSize bs = pmesh->GetBlockSize();
  const CoordinateDirection dirs[3] = {X1DIR, X2DIR, X3DIR};
  jb_mesh m;
  std::memset(&m, 0, sizeof(m));
  m.ndim = pmesh->ndim;
  m.nghost = 2;
  m.nblocks = pmesh->nbtotal;
  int max_level = 0;
  for (int d = 0; d < 3; ++d) {
    m.nxb[d] = bs.nx(dirs[d]);
    m.nrb[d] = d < m.ndim ? pmesh->mesh_size.nx(dirs[d]) / bs.nx(dirs[d]) : 1;
    m.xmin[d] = pmesh->mesh_size.xmin(dirs[d]);
    m.xmax[d] = pmesh->mesh_size.xmax(dirs[d]);
  }
  for (int f = 0; f < 6; ++f) {
    const BoundaryFlag b = pmesh->mesh_bcs[f];
    m.mesh_bc[f] = b == BoundaryFlag::periodic ? JB_BC_PERIODIC : JB_BC_OUTFLOW;
    m.swarm_bc[f] = s.swarm_bc[f];
  }
  std::vector<jb_block> blocks(static_cast<size_t>(m.nblocks));
  // owner of every gid: local blocks are known; with one rank that is all. With several ranks the runs of
  // gids per rank come from the block counts every rank contributes (MPI_Allgather in an MPI build).
  std::vector<int> owner(static_cast<size_t>(m.nblocks), 0);
#ifdef MPI_PARALLEL
  {
    std::vector<int> nb(static_cast<size_t>(Globals::nranks), 0);
    int mine = pmesh->GetNumMeshBlocksThisRank();
    MPI_Allgather(&mine, 1, MPI_INT, nb.data(), 1, MPI_INT, MPI_COMM_WORLD);
    int g = 0;
    for (int r = 0; r < Globals::nranks; ++r)
      for (int q = 0; q < nb[static_cast<size_t>(r)]; ++q) owner[static_cast<size_t>(g++)] = r;
  }
#endif
  for (int g = 0; g < m.nblocks; ++g) {
    const LogicalLocation &loc = loclist[static_cast<size_t>(g)];
    jb_block &b = blocks[static_cast<size_t>(g)];
    std::memset(&b, 0, sizeof(b));
    b.gid = g;
    b.level = loc.level() - pmesh->GetRootLevel();
    max_level = std::max(max_level, b.level);
    const std::int64_t lx[3] = {loc.lx1(), loc.lx2(), loc.lx3()};
    for (int d = 0; d < 3; ++d) {
      b.lx[d] = d < m.ndim ? lx[d] : 0;
      if (d < m.ndim) {
        const double w = (m.xmax[d] - m.xmin[d]) / static_cast<double>(static_cast<std::int64_t>(m.nrb[d]) << b.level);
        b.xmin[d] = m.xmin[d] + static_cast<double>(lx[d]) * w;
        b.xmax[d] = m.xmin[d] + static_cast<double>(lx[d] + 1) * w;
      } else {
        b.xmin[d] = m.xmin[d];
        b.xmax[d] = m.xmax[d];
      }
    }
    b.rank = owner[static_cast<size_t>(g)];
  }
  // the bounds of LOCAL blocks are taken from Parthenon itself (bit-identical coordinates)
  s.gids.clear();
  for (auto &pmb : pmesh->block_list) {
    jb_block &b = blocks[static_cast<size_t>(pmb->gid)];
    for (int d = 0; d < m.ndim; ++d) {
      b.xmin[d] = pmb->block_size.xmin(dirs[d]);
      b.xmax[d] = pmb->block_size.xmax(dirs[d]);
    }
    b.rank = Globals::my_rank;
    s.gids.push_back(pmb->gid);
  }
  m.max_level = max_level;
  m.blocks = blocks.data();
  Check(s, jb_set_mesh(s.ctx, &m, Globals::my_rank, Globals::nranks), "jb_set_mesh");
  s.ncell = static_cast<int64_t>(m.nxb[0]) * m.nxb[1] * m.nxb[2];
  const size_t n = static_cast<size_t>(s.ncell) * s.gids.size();
  for (auto &st : s.stage) st = ParArray1D<Real>("jaybenne_b200::stage", n);
  s.stage_tally = ParArray1D<Real>("jaybenne_b200::stage_tally", n);
  s.mesh_pushed = true;
}

// interior cells of one host variable of every local block -> one packed device array (x1 fastest), or back
static void PackField(Mesh *pmesh, AdapterState &s, const std::string &name, ParArray1D<Real> &dst, bool unpack) {
  int lb = 0;
  for (auto &pmb : pmesh->block_list) {
    auto &v = pmb->meshblock_data.Get()->Get(name).data;
    const IndexShape &cb = pmb->cellbounds;
    const int is = cb.is(IndexDomain::interior), ie = cb.ie(IndexDomain::interior);
    const int js = cb.js(IndexDomain::interior), je = cb.je(IndexDomain::interior);
    const int ks = cb.ks(IndexDomain::interior), ke = cb.ke(IndexDomain::interior);
    const int n1 = ie - is + 1, n2 = je - js + 1;
    const size_t base = static_cast<size_t>(lb) * static_cast<size_t>(s.ncell);
    auto buf = dst;
    parthenon::par_for(
        DEFAULT_LOOP_PATTERN, unpack ? "jaybenne_b200::unpack" : "jaybenne_b200::pack", DevExecSpace(), ks, ke, js, je, is, ie,
        KOKKOS_LAMBDA(const int k, const int j, const int i) {
          const size_t q = base + (static_cast<size_t>(k - ks) * n2 + (j - js)) * n1 + (i - is);
          if (unpack) {
            v(k, j, i) = buf(q);
          } else {
            buf(q) = v(k, j, i);
          }
        });
    ++lb;
  }
}

static void SetOrGet(AdapterState &s, int32_t field, ParArray1D<Real> &buf, bool get) {
  std::vector<double *> ptrs(s.gids.size());
  for (size_t q = 0; q < s.gids.size(); ++q) ptrs[q] = buf.data() + q * static_cast<size_t>(s.ncell);
  // one call for the whole pack: the copies are queued on the library's stream (device -> device)
  if (get) {
    Check(s, jb_get_fields(s.ctx, field, static_cast<int32_t>(s.gids.size()), s.gids.data(), ptrs.data()), "jb_get_fields");
  } else {
    Check(s, jb_set_fields(s.ctx, field, static_cast<int32_t>(s.gids.size()), s.gids.data(), ptrs.data()), "jb_set_fields");
  }
}

static void PushMaterial(Mesh *pmesh, AdapterState &s) {
  PackField(pmesh, s, HOST_DENSITY::name(), s.stage[0], false);
  PackField(pmesh, s, HOST_SPECIFIC_INTERNAL_ENERGY::name(), s.stage[1], false);
  PackField(pmesh, s, HOST_UPDATE_ENERGY::name(), s.stage[2], false);
  DevExecSpace().fence();  // the library's stream is not Kokkos' stream
  SetOrGet(s, JB_FIELD_DENSITY, s.stage[0], false);
  SetOrGet(s, JB_FIELD_SIE, s.stage[1], false);
  SetOrGet(s, JB_FIELD_INTERNAL_ENERGY, s.stage[2], false);
}

static void PullResults(Mesh *pmesh, AdapterState &s) {
  SetOrGet(s, JB_FIELD_ENERGY_TALLY, s.stage_tally, true);
  SetOrGet(s, JB_FIELD_INTERNAL_ENERGY, s.stage[2], true);  // UpdateFluid: u += energy_delta / dV (jaybenne.cpp:583-615)
  PackField(pmesh, s, "field.jaybenne.energy_tally", s.stage_tally, true);
  PackField(pmesh, s, HOST_UPDATE_ENERGY::name(), s.stage[2], true);
}

static void EnsureMesh(Mesh *pmesh, AdapterState &s) {
  if (s.mesh_pushed) return;
  PushMesh(pmesh, s);
}

// ---------------------------------------------------------------------------------------------------
// InitializeRadiation (jaybenne.cpp:570-578): mcblock::ProblemGenerator calls it once per block
// (mcblock.cpp:202); the library sources every local block in one launch, so the call that completes the
// rank's block list does the work.
// ---------------------------------------------------------------------------------------------------
void InitializeRadiation(MeshBlockData<Real> *mbd, const bool is_thermal) {
  Mesh *pmesh = mbd->GetParentPointer();
  auto state = GetState(pmesh);
  AdapterState &s = *state;
  if (++s.init_calls < pmesh->GetNumMeshBlocksThisRank()) return;
  s.init_calls = 0;
  EnsureMesh(pmesh, s);
  PushMaterial(pmesh, s);
  Check(s, jb_initialize_radiation(s.ctx, is_thermal ? 1 : 0), "jb_initialize_radiation");
  PullResults(pmesh, s);
}

// ---------------------------------------------------------------------------------------------------
// RadiationStep (jaybenne.cpp:68-151): the whole task list of one cycle is one library call.
// ---------------------------------------------------------------------------------------------------
TaskCollection RadiationStep(Mesh *pmesh, const Real t_start, const Real dt) {
  TaskCollection tc;
  TaskID none(0);
  auto &region = tc.AddRegion(1);
  auto &tl = region[0];
  tl.AddTask(none, [pmesh, t_start, dt]() {
    auto state = GetState(pmesh);
    AdapterState &s = *state;
    EnsureMesh(pmesh, s);
    PushMaterial(pmesh, s);
    jb_cycle_stats st;
    Check(s, jb_radiation_step(s.ctx, t_start, dt, &st), "jb_radiation_step");
    PullResults(pmesh, s);
    return TaskStatus::complete;
  });
  return tc;
}

// ---------------------------------------------------------------------------------------------------
// the public tasks (jaybenne.hpp:59-69) for hosts that build their own task lists
// ---------------------------------------------------------------------------------------------------
static AdapterState &S(MeshData<Real> *md) { return *GetState(md->GetParentPointer()); }
TaskStatus UpdateDerivedTransportFields(MeshData<Real> *md, const Real dt) {
  PushMaterial(md->GetParentPointer(), S(md));
  Check(S(md), jb_update_derived_transport_fields(S(md).ctx, dt), "jb_update_derived_transport_fields");
  return TaskStatus::complete;
}
template <typename T, SourceType ST>
TaskStatus SourcePhotons(T *md, const Real t_start, const Real dt) {
  AdapterState &s = *GetState(md->GetParentPointer());
  Check(s, jb_source_photons(s.ctx, ST == SourceType::thermal ? JB_SOURCE_THERMAL : JB_SOURCE_EMISSION, t_start, dt),
        "jb_source_photons");
  return TaskStatus::complete;
}
template TaskStatus SourcePhotons<MeshData<Real>, SourceType::emission>(MeshData<Real> *, const Real, const Real);
template TaskStatus SourcePhotons<MeshData<Real>, SourceType::thermal>(MeshData<Real> *, const Real, const Real);
TaskStatus TransportPhotons(MeshData<Real> *md, const Real t_start, const Real dt) {
  Check(S(md), jb_transport_photons(S(md).ctx, t_start, dt), "jb_transport_photons");
  Check(S(md), jb_swarm_exchange(S(md).ctx), "jb_swarm_exchange");  // MeshSend / MeshReceive (jaybenne.cpp:26-61)
  return TaskStatus::complete;
}
TaskStatus TransportPhotons_DDMC(MeshData<Real> *md, const Real t_start, const Real dt) {
  return TransportPhotons(md, t_start, dt);  // the library picks the hybrid kernel from use_ddmc
}
TaskStatus SampleDDMCBlockFace(MeshData<Real> *md) {
  Check(S(md), jb_sample_ddmc_block_face(S(md).ctx), "jb_sample_ddmc_block_face");
  return TaskStatus::complete;
}
TaskStatus CheckCompletion(MeshData<Real> *md, const Real t_end) {
  int64_t unfinished = 0;
  Check(S(md), jb_check_completion(S(md).ctx, t_end, &unfinished), "jb_check_completion");
  return unfinished > 0 ? TaskStatus::iterate : TaskStatus::complete;  // transport.cpp:207-215
}
template <typename T>
TaskStatus EvaluateRadiationEnergy(T *md) {
  AdapterState &s = *GetState(md->GetParentPointer());
  Check(s, jb_evaluate_radiation_energy(s.ctx), "jb_evaluate_radiation_energy");
  return TaskStatus::complete;
}
template TaskStatus EvaluateRadiationEnergy<MeshData<Real>>(MeshData<Real> *);
TaskStatus UpdateFluid(MeshData<Real> *md) {
  Check(S(md), jb_update_fluid(S(md).ctx), "jb_update_fluid");
  PullResults(md->GetParentPointer(), S(md));
  return TaskStatus::complete;
}
TaskStatus DefragParticles(MeshBlock *) { return TaskStatus::complete; }  // unused in the reference too

// ---------------------------------------------------------------------------------------------------
// PhotonReflectBC (boundaries.hpp:24-84): hosts register it under "jaybenne_reflecting"
// (mcblock.cpp:271-282). The reflection itself happens inside the transport kernels (the swarm the
// callback would receive is empty); the registration is what tells the library which faces reflect.
// ---------------------------------------------------------------------------------------------------
template <BoundaryFace BFACE>
void PhotonReflectBC(std::shared_ptr<Swarm> &) {}
template void PhotonReflectBC<BoundaryFace::inner_x1>(std::shared_ptr<Swarm> &);
template void PhotonReflectBC<BoundaryFace::outer_x1>(std::shared_ptr<Swarm> &);
template void PhotonReflectBC<BoundaryFace::inner_x2>(std::shared_ptr<Swarm> &);
template void PhotonReflectBC<BoundaryFace::outer_x2>(std::shared_ptr<Swarm> &);
template void PhotonReflectBC<BoundaryFace::inner_x3>(std::shared_ptr<Swarm> &);
template void PhotonReflectBC<BoundaryFace::outer_x3>(std::shared_ptr<Swarm> &);

}  // namespace jaybenne
