// mcblock_main.cpp — `mcblock_b200`: the reference's driver application re-hosted on the C ABI.
//
//   main / ParthenonManager set-up        src/mcblock/main.cpp:24-75
//   mcblock::Initialize, ProcessPackages  src/mcblock/mcblock.cpp:37-150, 288-297
//   ProblemGenerator, PostInitialization  src/mcblock/mcblock.cpp:155-262
//   McblockDriver::Step, HostUpdateTasks  src/mcblock/mcblock_driver.cpp:23-74
//   EvolutionDriver::Execute              fixed dt = <jaybenne> dt, clipped at tlim, nlim honoured
//
// Usage (same shape as the reference binary, tst/regression_test.py:320-345):
//   mcblock_b200 -i deck.in [block/key=value ...] [--outdir DIR] [--tasklist] [--max-cycles N]
//                [--capacity N] [--parse-only] [--dump-particles] [--quiet]
// One process per GPU: RANK / WORLD_SIZE / LOCAL_RANK from the environment (torchrun or mpirun
// style launchers); the NCCL id travels through a file (JB_ID_FILE, default <outdir>/.jb_nccl_id).
// Dumps are directories of .npy arrays (no HDF5 on this image): the fields <parthenon/output0>
// lists, [local block][k][j][i], plus block bounds as analysis/jhdf.py:36-73 exposes them.
// Pure C++17 host code: everything that computes runs in libjaybenne_b200.so on the GPU.
#include <sys/stat.h>
#include <unistd.h>

#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <fstream>
#include <iostream>
#include <sstream>
#include <thread>

#include "../../../include/jaybenne_b200.hpp"

namespace jb = jaybenne_b200;
using jb::Real;

namespace {

struct Options {
  std::string deck, outdir = ".";
  std::vector<std::string> overrides;
  bool tasklist = false, parse_only = false, quiet = false, dump_particles = false;
  long max_cycles = -1;
  int64_t capacity = 0;
};

int env_int(const char *name, int dflt) {
  const char *v = getenv(name);
  return v && *v ? atoi(v) : dflt;
}

void usage() {
  fprintf(stderr,
          "usage: mcblock_b200 -i deck.in [block/key=value ...] [--outdir DIR] [--tasklist] [--max-cycles N]\n"
          "                    [--capacity N] [--parse-only] [--dump-particles] [--quiet]\n");
}

bool parse_args(int argc, char **argv, Options &o) {
  for (int a = 1; a < argc; ++a) {
    const std::string s = argv[a];
    auto next = [&](std::string &dst) {
      if (a + 1 >= argc) return false;
      dst = argv[++a];
      return true;
    };
    std::string v;
    if (s == "-i") {
      if (!next(o.deck)) return false;
    } else if (s == "--outdir" || s == "-d") {
      if (!next(o.outdir)) return false;
    } else if (s == "--tasklist") {
      o.tasklist = true;
    } else if (s == "--parse-only") {
      o.parse_only = true;
    } else if (s == "--dump-particles") {
      o.dump_particles = true;
    } else if (s == "--quiet") {
      o.quiet = true;
    } else if (s == "--max-cycles") {
      if (!next(v)) return false;
      o.max_cycles = atol(v.c_str());
    } else if (s == "--capacity") {
      if (!next(v)) return false;
      o.capacity = atoll(v.c_str());
    } else if (s == "-h" || s == "--help") {
      return false;
    } else if (s.find('=') != std::string::npos && s.find('/') != std::string::npos) {
      o.overrides.push_back(s);
    } else {
      fprintf(stderr, "mcblock_b200: unknown argument '%s'\n", s.c_str());
      return false;
    }
  }
  return !o.deck.empty();
}

// ---- .npy (format 1.0) ---------------------------------------------------------------------
void write_npy(const std::string &path, const char *descr, size_t itemsize, const std::vector<size_t> &shape,
               const void *data) {
  std::ostringstream h;
  h << "{'descr': '" << descr << "', 'fortran_order': False, 'shape': (";
  size_t count = 1;
  for (size_t d = 0; d < shape.size(); ++d) {
    h << shape[d] << (shape.size() == 1 || d + 1 < shape.size() ? "," : "");
    if (d + 1 < shape.size()) h << " ";
    count *= shape[d];
  }
  h << "), }";
  std::string hdr = h.str();
  const size_t unpadded = 10 + hdr.size() + 1;
  hdr.append((64 - unpadded % 64) % 64, ' ');
  hdr.push_back('\n');
  std::ofstream f(path, std::ios::binary);
  if (!f) throw jb::Error(JB_ERR_INVALID, "cannot write " + path);
  const unsigned char magic[8] = {0x93, 'N', 'U', 'M', 'P', 'Y', 1, 0};
  f.write((const char *)magic, 8);
  const uint16_t hl = (uint16_t)hdr.size();
  f.write((const char *)&hl, 2);
  f.write(hdr.data(), (std::streamsize)hdr.size());
  f.write((const char *)data, (std::streamsize)(count * itemsize));
}

void make_dir(const std::string &p) {
  if (mkdir(p.c_str(), 0777) != 0 && errno != EEXIST) throw jb::Error(JB_ERR_INVALID, "cannot create directory " + p);
}

// "field.material.density" -> (npy name, JB_FIELD_*)
bool output_field(const std::string &var, std::string &name, int32_t &field) {
  static const struct { const char *var, *name; int32_t field; } table[] = {
      {"field.material.density", "density", JB_FIELD_DENSITY},
      {"field.material.sie", "sie", JB_FIELD_SIE},
      {"field.material.internal_energy", "internal_energy", JB_FIELD_INTERNAL_ENERGY},
      {"field.jaybenne.energy_tally", "energy_tally", JB_FIELD_ENERGY_TALLY},
      {"field.jaybenne.fleck_factor", "fleck_factor", JB_FIELD_FLECK_FACTOR},
      {"field.jaybenne.energy_delta", "energy_delta", JB_FIELD_ENERGY_DELTA},
      {"field.jaybenne.source_ew_per_cell", "source_ew_per_cell", JB_FIELD_SOURCE_EW_PER_CELL},
      {"field.jaybenne.source_num_per_cell", "source_num_per_cell", JB_FIELD_SOURCE_NUM_PER_CELL},
  };
  for (const auto &t : table)
    if (var == t.var) {
      name = t.name;
      field = t.field;
      return true;
    }
  return false;
}

std::vector<std::string> split_list(const std::string &s) {
  std::vector<std::string> out;
  std::string cur;
  for (char ch : s) {
    if (ch == ',') {
      if (!cur.empty()) out.push_back(cur);
      cur.clear();
    } else if (ch != ' ' && ch != '\t') {
      cur.push_back(ch);
    }
  }
  if (!cur.empty()) out.push_back(cur);
  return out;
}

}  // namespace

// ---------------------------------------------------------------------------------------
// mcblock: the package that owns the material fields + the driver around jaybenne
// ---------------------------------------------------------------------------------------
class Mcblock {
 public:
  Mcblock(jb::ParameterInput &pin, const Options &opt, int rank, int n_ranks)
      : pin_(pin), opt_(opt), rank_(rank), n_ranks_(n_ranks), mesh_(pin, n_ranks) {
    // McblockDriver's input checks (mcblock_driver.cpp:23-28) and mcblock::Initialize's (mcblock.cpp:40-41)
    for (const char *k : {"ix1_bc", "ox1_bc", "ix2_bc", "ox2_bc", "ix3_bc", "ox3_bc"}) pin.CheckRequired("parthenon/mesh", k);
    if (pin.GetOrAddString("parthenon/time", "integrator", "rk1") != "rk1")
      throw jb::Error(JB_ERR_INVALID, "Only rk1 integrator is supported!");
    problem_id_ = pin.GetOrAddString("parthenon/job", "problem_id", "mcblock");
    rho0_ = pin.GetReal("mcblock", "initial_density");
    tt0_ = pin.GetReal("mcblock", "initial_temperature");
    const std::string ir = pin.GetOrAddString("mcblock", "initial_radiation", "none");
    if (ir != "none" && ir != "thermal") throw jb::Error(JB_ERR_INVALID, "Initial radiation model \"" + ir + "\" not recognized!");
    thermal_ic_ = ir == "thermal";
    tlim_ = pin.GetOrAddReal("parthenon/time", "tlim", std::numeric_limits<Real>::infinity());
    nlim_ = (long)pin.GetOrAddInteger("parthenon/time", "nlim", -1);
    out_dt_ = pin.GetOrAddReal("parthenon/output0", "dt", 0.0);
    for (const auto &v : split_list(pin.GetOrAddString("parthenon/output0", "variables",
                                                       "field.material.density,field.material.sie,"
                                                       "field.material.internal_energy,field.jaybenne.energy_tally"))) {
      std::string name;
      int32_t field;
      if (output_field(v, name, field)) out_fields_.push_back({name, field});
    }
    pkg_ = jb::Initialize(pin, mesh_, rank, n_ranks, opt.capacity);
    dt_ = jb::EstimateTimestepMesh(*pkg_);
  }

  jb::Package &pkg() { return *pkg_; }
  const jb::Mesh &mesh() const { return mesh_; }
  Real time() const { return time_; }
  long ncycle() const { return ncycle_; }

  // mcblock::ProblemGenerator on every local block, then PostInitialization
  void ProblemGenerator() {
    const Real cv = pkg_->params().eos_cv;  // eos.SpecificHeatFromDensityInternalEnergy(rho0, 1.0)
    const jb_mesh &m = mesh_.c();
    const int n1 = m.nxb[0], n2 = m.nxb[1], n3 = m.nxb[2];
    std::vector<Real> rho((size_t)n1 * n2 * n3), sie(rho.size());
    const Real two_pi = 2.0 * 3.14159265358979323846;
    for (int gid : pkg_->local_gids()) {
      for (int k = 0; k < n3; ++k)
        for (int j = 0; j < n2; ++j)
          for (int i = 0; i < n1; ++i) {
            const size_t c = ((size_t)k * n2 + j) * n1 + i;
            rho[c] = rho0_;
            sie[c] = cv * tt0_;
            if (problem_id_ == "stepdiff") {  // mcblock.cpp:186-200
              if (mesh_.Xc(gid, 0, i) >= 0.0) {
                const Real ttlow = 1.0e-5 * tt0_;
                sie[c] = cv * ttlow;
              }
            } else if (problem_id_ == "synthetic") {  // the C5 benchmark medium; not in the reference
              rho[c] = rho0_ * (1.0 + 0.1 * std::sin(two_pi * mesh_.Xc(gid, 0, i)) * std::sin(two_pi * mesh_.Xc(gid, 1, j)) *
                                           std::sin(two_pi * mesh_.Xc(gid, 2, k)));
            }
          }
      pkg_->SetField(gid, JB_FIELD_DENSITY, rho.data());
      pkg_->SetField(gid, JB_FIELD_SIE, sie.data());
    }
    jb::InitializeRadiation(*pkg_, thermal_ic_);
    jb::PostInitialization(*pkg_);
  }

  // McblockDriver::Step: one radiation cycle, then HostUpdateTasks (FillDerived -> EstimateTimestep)
  jb_cycle_stats Step() {
    const jb_cycle_stats st = opt_.tasklist ? jb::RadiationStepTaskList(*pkg_, time_, dt_) : jb::RadiationStep(*pkg_, time_, dt_);
    jb::UpdateDerived(*pkg_);
    return st;
  }

  // EvolutionDriver::Execute
  void Execute() {
    const std::string hst_path = opt_.outdir + "/" + problem_id_ + ".hst.r" + std::to_string(rank_) + ".csv";
    std::ofstream hst(hst_path);
    hst << "cycle,time,dt,n_start,n_sourced,n_absorbed,n_escaped,n_end,n_sent,n_received,n_block_crossings,n_segments,"
           "n_scatters,n_ddmc_segments,n_bface_resampled,n_iterations,e_start,e_sourced,e_absorbed,e_escaped,e_end,ms_transport,ms_total\n";
    hst.precision(17);
    Real next_out = out_dt_ > 0 ? out_dt_ : std::numeric_limits<Real>::infinity();
    if (out_dt_ > 0) Dump(0);
    int nout = 1;
    while (time_ < tlim_ && (nlim_ < 0 || ncycle_ < nlim_)) {
      if (opt_.max_cycles >= 0 && ncycle_ >= opt_.max_cycles) break;
      const jb_cycle_stats s = Step();
      for (const auto &ph : pkg_->StepTimeline())
        if (ph.first >= 0 && ph.first < 9) phase_ms_[ph.first] += ph.second;
      hst << ncycle_ << ',' << time_ << ',' << dt_ << ',' << s.n_start << ',' << s.n_sourced << ',' << s.n_absorbed << ','
          << s.n_escaped << ',' << s.n_end << ',' << s.n_sent << ',' << s.n_received << ',' << s.n_block_crossings << ','
          << s.n_segments << ',' << s.n_scatters << ',' << s.n_ddmc_segments << ',' << s.n_bface_resampled << ',' << s.n_iterations << ',' << s.e_start
          << ',' << s.e_sourced << ',' << s.e_absorbed << ',' << s.e_escaped << ',' << s.e_end << ',' << s.ms_transport
          << ',' << s.ms_total << '\n';
      if (rank_ == 0 && !opt_.quiet)
        printf("cycle=%ld time=%.14e dt=%.14e photons=%lld sourced=%lld absorbed=%lld segments=%lld iterations=%d "
               "device_ms=%.3f\n",
               ncycle_, time_, dt_, (long long)s.n_end, (long long)s.n_sourced, (long long)s.n_absorbed,
               (long long)s.n_segments, s.n_iterations, s.ms_total);
      time_ += dt_;
      ncycle_ += 1;
      dt_ = jb::EstimateTimestepMesh(*pkg_);
      if (time_ < tlim_ && (tlim_ - time_) < dt_) dt_ = tlim_ - time_;
      if (time_ >= next_out) {
        Dump(nout++);
        next_out += out_dt_;
      }
    }
    Dump(-1);
    if (rank_ == 0 && !opt_.quiet && ncycle_ > 0) {
      // where the device time of the run went (jb_step_timeline; the reference's Kokkos regions, jaybenne.cpp:87-145)
      static const char *names[9] = {"", "fields", "source", "transport", "exchange", "stats", "squeeze", "tally", "update"};
      printf("device ms by phase over %ld cycles (rank 0, exchange mode %d):", ncycle_, pkg_->ExchangeMode());
      for (int q = 1; q < 9; ++q) printf(" %s=%.3f", names[q], phase_ms_[q]);
      printf("\n");
    }
  }

  // <problem_id>.out0.<NNNNN|final>/r<rank>.<field>.npy
  void Dump(int number) {
    char tag[32];
    if (number < 0) snprintf(tag, sizeof(tag), "final");
    else snprintf(tag, sizeof(tag), "%05d", number);
    const std::string dir = opt_.outdir + "/" + problem_id_ + ".out0." + tag;
    make_dir(dir);
    const std::string pre = dir + "/r" + std::to_string(rank_) + ".";
    const jb_mesh &m = mesh_.c();
    const auto &gids = pkg_->local_gids();
    const size_t nb = gids.size(), n1 = (size_t)m.nxb[0], n2 = (size_t)m.nxb[1], n3 = (size_t)m.nxb[2];
    std::vector<Real> buf(nb * n1 * n2 * n3);
    for (const auto &f : out_fields_) {
      for (size_t b = 0; b < nb; ++b) pkg_->GetField(gids[b], f.second, buf.data() + b * n1 * n2 * n3);
      write_npy(pre + f.first + ".npy", "<f8", 8, {nb, n3, n2, n1}, buf.data());
    }
    std::vector<Real> lo(nb * 3), hi(nb * 3);
    std::vector<int32_t> lev(nb), gid32(nb);
    for (size_t b = 0; b < nb; ++b) {
      const jb_block &blk = mesh_.block(gids[b]);
      for (int d = 0; d < 3; ++d) {
        lo[b * 3 + d] = blk.xmin[d];
        hi[b * 3 + d] = blk.xmax[d];
      }
      lev[b] = blk.level;
      gid32[b] = blk.gid;
    }
    write_npy(pre + "xmin.npy", "<f8", 8, {nb, 3}, lo.data());
    write_npy(pre + "xmax.npy", "<f8", 8, {nb, 3}, hi.data());
    write_npy(pre + "level.npy", "<i4", 4, {nb}, lev.data());
    write_npy(pre + "gids.npy", "<i4", 4, {nb}, gid32.data());
    if (opt_.dump_particles) {  // the photons swarm (Parthenon writes swarms into the same phdf file)
      const jb::Package::Photons ph = pkg_->GetParticles();
      const size_t n = ph.size();
      const std::pair<const char *, const std::vector<Real> *> reals[] = {
          {"x", &ph.x}, {"y", &ph.y}, {"z", &ph.z}, {"time", &ph.time}, {"weight", &ph.weight},
          {"energy", &ph.energy}, {"v0", &ph.v0}, {"v1", &ph.v1}, {"v2", &ph.v2}};
      for (const auto &r : reals) write_npy(pre + "photons." + r.first + ".npy", "<f8", 8, {n}, r.second->data());
      const std::pair<const char *, const std::vector<int32_t> *> ints[] = {
          {"i", &ph.i}, {"j", &ph.j}, {"k", &ph.k}, {"block", &ph.block}};
      for (const auto &r : ints) write_npy(pre + "photons." + r.first + ".npy", "<i4", 4, {n}, r.second->data());
      write_npy(pre + "photons.rng.npy", "<u4", 4, {n, 4}, ph.rng.data());
    }
    std::ofstream info(pre + "info.json");
    info.precision(17);
    info << "{\"time\": " << time_ << ", \"ncycle\": " << ncycle_ << ", \"dt\": " << dt_ << ", \"rank\": " << rank_
         << ", \"n_ranks\": " << n_ranks_ << ", \"nbtotal\": " << mesh_.nbtotal() << ", \"num_particles\": "
         << pkg_->NumParticles() << ", \"kernel_launches\": " << pkg_->KernelLaunches() << "}\n";
  }

 private:
  jb::ParameterInput &pin_;
  const Options &opt_;
  int rank_, n_ranks_;
  jb::Mesh mesh_;
  std::shared_ptr<jb::Package> pkg_;
  std::string problem_id_;
  Real rho0_ = 0, tt0_ = 0, tlim_ = 0, out_dt_ = 0, time_ = 0, dt_ = 0;
  long nlim_ = -1, ncycle_ = 0;
  double phase_ms_[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};  // device time per JB_TL_* phase, summed over the run
  bool thermal_ic_ = false;
  std::vector<std::pair<std::string, int32_t>> out_fields_;
};

namespace {

// NCCL unique id from rank 0 to every rank through a file (no MPI on this image)
std::vector<unsigned char> rendezvous_id(const std::string &path, int rank) {
  std::vector<unsigned char> id(128);
  if (rank == 0) {
    id = jb::Package::CommUniqueId();
    const std::string tmp = path + ".tmp";
    {
      std::ofstream f(tmp, std::ios::binary);
      f.write((const char *)id.data(), 128);
    }
    if (rename(tmp.c_str(), path.c_str()) != 0) throw jb::Error(JB_ERR_COMM, "cannot publish the NCCL id at " + path);
    return id;
  }
  for (int tries = 0; tries < 1200; ++tries) {  // up to 2 minutes
    std::ifstream f(path, std::ios::binary);
    if (f && f.read((char *)id.data(), 128) && f.gcount() == 128) return id;
    std::this_thread::sleep_for(std::chrono::milliseconds(100));
  }
  throw jb::Error(JB_ERR_COMM, "timed out waiting for the NCCL id at " + path);
}

void print_summary(const jb::ParameterInput &pin, const jb_config &cfg, const jb::Mesh &mesh) {
  const jb_mesh &m = mesh.c();
  printf("problem_id=%s\n", pin.DoesParameterExist("parthenon/job", "problem_id")
                                ? pin.GetString("parthenon/job", "problem_id").c_str() : "mcblock");
  printf("ndim=%d nbtotal=%d max_level=%d block=%dx%dx%d root_blocks=%dx%dx%d\n", m.ndim, m.nblocks, m.max_level, m.nxb[0],
         m.nxb[1], m.nxb[2], m.nrb[0], m.nrb[1], m.nrb[2]);
  printf("num_particles=%lld dt=%.17g seed=%lld use_ddmc=%d tau_ddmc=%.17g do_emission=%d do_feedback=%d "
         "max_transport_iterations=%d\n",
         (long long)cfg.num_particles, cfg.dt, (long long)cfg.seed, cfg.use_ddmc, cfg.tau_ddmc, cfg.do_emission,
         cfg.do_feedback, cfg.max_transport_iterations);
  printf("opacity_model=%d kappa=%.17g scattering_model=%d sigma=%.17g apm=%.17g cv=%.17g gm1=%.17g\n", cfg.opacity_model,
         cfg.opacity_kappa, cfg.scattering_model, cfg.scattering_sigma, cfg.apm, cfg.eos_cv, cfg.eos_gm1);
  printf("swarm_bc=%d,%d,%d,%d,%d,%d mesh_bc=%d,%d,%d,%d,%d,%d\n", m.swarm_bc[0], m.swarm_bc[1], m.swarm_bc[2], m.swarm_bc[3],
         m.swarm_bc[4], m.swarm_bc[5], m.mesh_bc[0], m.mesh_bc[1], m.mesh_bc[2], m.mesh_bc[3], m.mesh_bc[4], m.mesh_bc[5]);
}

}  // namespace

int main(int argc, char **argv) {
  Options opt;
  if (!parse_args(argc, argv, opt)) {
    usage();
    return 2;
  }
  const int rank = env_int("RANK", env_int("OMPI_COMM_WORLD_RANK", 0));
  const int n_ranks = env_int("WORLD_SIZE", env_int("OMPI_COMM_WORLD_SIZE", 1));
  const int local_rank = env_int("LOCAL_RANK", env_int("OMPI_COMM_WORLD_LOCAL_RANK", rank));
  try {
    jb::ParameterInput pin = jb::ParameterInput::FromFile(opt.deck);
    for (const auto &o : opt.overrides) pin.ModifyFromCmdline(o);
    if (opt.parse_only) {  // host logic only: no GPU needed
      jb::Mesh mesh(pin, n_ranks);
      print_summary(pin, jb::ConfigFromInput(pin), mesh);
      return 0;
    }
    make_dir(opt.outdir);
    int32_t ndev = 0;
    if (jb_device_count(&ndev) != JB_OK)
      throw jb::Error(JB_ERR_NO_DEVICE, "no CUDA device: mcblock_b200 has no CPU path");
    if (jb_set_device(local_rank % ndev) != JB_OK) throw jb::Error(JB_ERR_CUDA, "cannot select the CUDA device");
    Mcblock app(pin, opt, rank, n_ranks);
    if (n_ranks > 1) {
      const char *idf = getenv("JB_ID_FILE");
      const char *port = getenv("MASTER_PORT");
      const std::string path = idf && *idf ? std::string(idf) : opt.outdir + "/.jb_nccl_id." + (port ? port : "0");
      app.pkg().CommInit(rendezvous_id(path, rank));
      if (rank == 0) unlink(path.c_str());
    }
    app.ProblemGenerator();
    const auto t0 = std::chrono::steady_clock::now();
    app.Execute();
    const double wall = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    if (rank == 0 && !opt.quiet)
      printf("mcblock_b200 done: cycles=%ld time=%.14e wall=%.3fs kernel_launches=%lld\n", app.ncycle(), app.time(), wall,
             (long long)app.pkg().KernelLaunches());
  } catch (const jb::Error &e) {
    // PARTHENON_FAIL prints the message and aborts; a non-zero exit keeps the harness contract
    fprintf(stderr, "### FATAL ERROR (jb_status %d)\n%s\n", (int)e.status, e.what());
    return 1;
  }
  return 0;
}
