// deck.cpp — host-side mirror of the pieces of Parthenon / mcblock that sit in front of
// the jaybenne package: a ParameterInput-style deck reader, the <jaybenne>/<mcblock>
// parameter handling of jaybenne::Initialize (src/jaybenne/jaybenne.cpp:158-223) and
// mcblock::Initialize (src/mcblock/mcblock.cpp:37-150), and a meshblock-table builder
// for uniform and statically refined meshes (<parthenon/mesh>, <parthenon/meshblock>,
// <parthenon/static_refinementN>, <parthenon/swarm>). Pure C++: usable without a GPU.
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <limits>
#include <map>
#include <set>
#include <sstream>
#include <string>
#include <tuple>
#include <vector>

#include "../../../include/jaybenne_b200.h"

struct jb_deck {
  // block -> key -> value, insertion order is irrelevant for lookups
  std::map<std::string, std::map<std::string, std::string>> blocks;
  std::vector<std::string> block_order;
};

namespace {

std::string trim(const std::string &s) {
  size_t a = s.find_first_not_of(" \t\r\n");
  if (a == std::string::npos) return "";
  size_t b = s.find_last_not_of(" \t\r\n");
  return s.substr(a, b - a + 1);
}

jb_status parse_text(const std::string &text, jb_deck *d, std::string &err) {
  std::istringstream in(text);
  std::string line, block, pending_key, pending_val;
  bool cont = false;
  int lineno = 0;
  while (std::getline(in, line)) {
    ++lineno;
    size_t hash = line.find('#');
    if (hash != std::string::npos) line = line.substr(0, hash);
    line = trim(line);
    if (line.empty()) continue;
    if (cont) {
      bool more = !line.empty() && line.back() == '&';
      if (more) line = trim(line.substr(0, line.size() - 1));
      pending_val += line;
      if (!more) {
        d->blocks[block][pending_key] = pending_val;
        cont = false;
      }
      continue;
    }
    if (line.front() == '<') {
      size_t e = line.find('>');
      if (e == std::string::npos) {
        err = "line " + std::to_string(lineno) + ": unterminated block name";
        return JB_ERR_INVALID;
      }
      block = trim(line.substr(1, e - 1));
      if (!d->blocks.count(block)) d->block_order.push_back(block);
      d->blocks[block];
      continue;
    }
    size_t eq = line.find('=');
    if (eq == std::string::npos || block.empty()) {
      err = "line " + std::to_string(lineno) + ": expected key = value inside a <block>";
      return JB_ERR_INVALID;
    }
    std::string key = trim(line.substr(0, eq));
    std::string val = trim(line.substr(eq + 1));
    if (!val.empty() && val.back() == '&') {
      cont = true;
      pending_key = key;
      pending_val = trim(val.substr(0, val.size() - 1));
      continue;
    }
    d->blocks[block][key] = val;
  }
  return JB_OK;
}

struct Reader {
  const jb_deck *d;
  std::string err;
  bool has(const char *b, const char *k) const {
    auto it = d->blocks.find(b);
    return it != d->blocks.end() && it->second.count(k);
  }
  const std::string &raw(const char *b, const char *k) const { return d->blocks.at(b).at(k); }
  bool need(const char *b, const char *k) {
    if (has(b, k)) return true;
    if (err.empty()) err = std::string("Parameter name '") + k + "' not found in block '" + b + "'";
    return false;
  }
  double real(const char *b, const char *k, double dflt) const { return has(b, k) ? atof(raw(b, k).c_str()) : dflt; }
  long integer(const char *b, const char *k, long dflt) const {
    return has(b, k) ? (long)atoll(raw(b, k).c_str()) : dflt;
  }
  std::string str(const char *b, const char *k, const std::string &dflt) const { return has(b, k) ? raw(b, k) : dflt; }
  bool boolean(const char *b, const char *k, bool dflt) const {
    if (!has(b, k)) return dflt;
    std::string v = raw(b, k);
    std::transform(v.begin(), v.end(), v.begin(), ::tolower);
    return v == "true" || v == "1" || v == "t" || v == "yes";
  }
};

void set_err(char *err, int32_t errlen, const std::string &m) {
  if (err && errlen > 0) {
    snprintf(err, (size_t)errlen, "%s", m.c_str());
  }
}

int bc_from_name(const std::string &name) {
  if (name == "periodic") return JB_BC_PERIODIC;
  if (name == "jaybenne_reflecting" || name == "reflecting") return JB_BC_REFLECT;
  if (name == "outflow") return JB_BC_OUTFLOW;
  return -1;
}

// ---- block tree ------------------------------------------------------------------------
struct Loc {
  int level;
  long lx[3];
  bool operator<(const Loc &o) const {
    return std::tie(level, lx[0], lx[1], lx[2]) < std::tie(o.level, o.lx[0], o.lx[1], o.lx[2]);
  }
};

struct Tree {
  int ndim;
  long nrb[3];
  bool periodic[3];
  std::set<Loc> leaves;

  bool is_leaf(const Loc &l) const { return leaves.count(l) != 0; }
  void split(const Loc &l) {
    leaves.erase(l);
    const int nk = ndim > 2 ? 2 : 1, nj = ndim > 1 ? 2 : 1;
    for (int k = 0; k < nk; ++k)
      for (int j = 0; j < nj; ++j)
        for (int i = 0; i < 2; ++i) {
          Loc c;
          c.level = l.level + 1;
          c.lx[0] = 2 * l.lx[0] + i;
          c.lx[1] = ndim > 1 ? 2 * l.lx[1] + j : 0;
          c.lx[2] = ndim > 2 ? 2 * l.lx[2] + k : 0;
          leaves.insert(c);
        }
  }
  // make sure location t exists as a leaf (or is already refined further)
  void ensure(const Loc &t) {
    for (int l = 0; l < t.level; ++l) {
      Loc a;
      a.level = l;
      for (int d = 0; d < 3; ++d) a.lx[d] = d < ndim ? (t.lx[d] >> (t.level - l)) : 0;
      if (is_leaf(a)) split(a);
    }
  }
  // leaf covering logical position p given at level `level`; false if outside the domain
  bool covering(int level, const long pin[3], Loc &out) const {
    long p[3] = {pin[0], pin[1], pin[2]};
    for (int d = 0; d < ndim; ++d) {
      const long n = nrb[d] << level;
      if (p[d] < 0 || p[d] >= n) {
        if (!periodic[d]) return false;
        p[d] = ((p[d] % n) + n) % n;
      }
    }
    for (int l = level; l >= 0; --l) {
      Loc a;
      a.level = l;
      for (int d = 0; d < 3; ++d) a.lx[d] = d < ndim ? (p[d] >> (level - l)) : 0;
      if (is_leaf(a)) {
        out = a;
        return true;
      }
    }
    return false;  // position is covered by finer leaves
  }
  void balance() {
    bool changed = true;
    while (changed) {
      changed = false;
      std::vector<Loc> snapshot(leaves.begin(), leaves.end());
      for (const Loc &l : snapshot) {
        if (!is_leaf(l) || l.level < 2) continue;
        const int kr = ndim > 2 ? 1 : 0, jr = ndim > 1 ? 1 : 0;
        for (int ok = -kr; ok <= kr; ++ok)
          for (int oj = -jr; oj <= jr; ++oj)
            for (int oi = -1; oi <= 1; ++oi) {
              if (!oi && !oj && !ok) continue;
              long p[3] = {l.lx[0] + oi, l.lx[1] + oj, l.lx[2] + ok};
              Loc cov;
              if (covering(l.level, p, cov) && cov.level < l.level - 1) {
                split(cov);
                changed = true;
              }
            }
      }
    }
  }
};

uint64_t morton(const Loc &l, int max_level, int ndim) {
  uint64_t key = 0;
  uint64_t c[3];
  for (int d = 0; d < 3; ++d) c[d] = d < ndim ? (uint64_t)l.lx[d] << (max_level - l.level) : 0;
  for (int bit = 20; bit >= 0; --bit)
    for (int d = 2; d >= 0; --d) key = (key << 1) | ((c[d] >> bit) & 1u);
  return key;
}

// Parthenon / Athena++ uniform mesh generator: logical fraction -> coordinate
double mesh_generator(long idx, long nrange, double xmin, double xmax) {
  const double x = (static_cast<double>(idx) - 0.5 * static_cast<double>(nrange)) / static_cast<double>(nrange);
  return 0.5 * (xmin + xmax) + (x * xmax - x * xmin);
}

}  // namespace

extern "C" {

jb_status jb_deck_parse(const char *text, jb_deck **out) {
  if (!text || !out) return JB_ERR_INVALID;
  jb_deck *d = new jb_deck();
  std::string err;
  jb_status s = parse_text(text, d, err);
  if (s != JB_OK) {
    delete d;
    *out = nullptr;
    return s;
  }
  *out = d;
  return JB_OK;
}

jb_status jb_deck_load(const char *path, jb_deck **out) {
  if (!path || !out) return JB_ERR_INVALID;
  std::ifstream f(path);
  if (!f) return JB_ERR_INVALID;
  std::stringstream ss;
  ss << f.rdbuf();
  return jb_deck_parse(ss.str().c_str(), out);
}

void jb_deck_free(jb_deck *d) { delete d; }

jb_status jb_deck_set(jb_deck *d, const char *block, const char *key, const char *value) {
  if (!d || !block || !key || !value) return JB_ERR_INVALID;
  if (!d->blocks.count(block)) d->block_order.push_back(block);
  d->blocks[block][key] = value;
  return JB_OK;
}

int32_t jb_deck_has(const jb_deck *d, const char *block, const char *key) {
  Reader r{d, ""};
  return r.has(block, key) ? 1 : 0;
}

const char *jb_deck_get(const jb_deck *d, const char *block, const char *key) {
  Reader r{d, ""};
  return r.has(block, key) ? r.raw(block, key).c_str() : nullptr;
}

void jb_config_default(jb_config *cfg) {
  memset(cfg, 0, sizeof(*cfg));
  cfg->dt = std::numeric_limits<double>::max();
  cfg->numin = std::numeric_limits<double>::min();
  cfg->numax = std::numeric_limits<double>::max();
  cfg->unique_rank_seeds = 1;
  cfg->seed = 123;
  cfg->max_transport_iterations = 10000;
  cfg->tau_ddmc = 5.0;
  cfg->source_strategy = JB_STRATEGY_UNIFORM;
  cfg->do_emission = 1;
  cfg->do_feedback = 1;
  cfg->opacity_model = JB_OPAC_GRAY;
  cfg->scattering_model = JB_SCAT_GRAY;
  cfg->apm = 1.0;
  cfg->time_scale = cfg->mass_scale = cfg->length_scale = cfg->temperature_scale = 1.0;
  cfg->eos_gm1 = 1.66666666667 - 1.;
  cfg->eos_cv = 1. / (1.66666666667 - 1.);
  cfg->rng_mode = JB_RNG_PHILOX;
  cfg->gray_mode = JB_GRAY_KIRCHHOFF;
}

jb_status jb_deck_config(const jb_deck *d, jb_config *cfg, char *err, int32_t errlen) {
  if (!d || !cfg) return JB_ERR_INVALID;
  Reader r{d, ""};
  jb_config_default(cfg);
  // mcblock::Initialize, mcblock.cpp:40-41 (Q9)
  if (r.str("parthenon/time", "integrator", "") != "rk1") {
    set_err(err, errlen, "McBlock driver only supports first order time integration");
    return JB_ERR_INVALID;
  }
  // McblockDriver ctor, mcblock_driver.cpp:23-28
  const char *bck[6] = {"ix1_bc", "ox1_bc", "ix2_bc", "ox2_bc", "ix3_bc", "ox3_bc"};
  for (auto k : bck)
    if (!r.need("parthenon/mesh", k)) break;
  // required mcblock keys (mcblock.cpp:48-56)
  r.need("mcblock", "initial_temperature");
  r.need("mcblock", "initial_density");
  r.need("mcblock", "initial_radiation");
  r.need("mcblock", "opacity_model");
  r.need("jaybenne", "num_particles");
  if (!r.err.empty()) {
    set_err(err, errlen, r.err);
    return JB_ERR_INVALID;
  }
  const std::string ir = r.str("mcblock", "initial_radiation", "");
  if (ir != "none" && ir != "thermal") {
    set_err(err, errlen, "Only none or thermal initial radiation supported!");
    return JB_ERR_INVALID;
  }
  // EOS (mcblock.cpp:78-80)
  const double gamma = r.real("mcblock", "gamma", 1.66666666667);
  const double cv = r.real("mcblock", "cv", 1. / (gamma - 1.));
  cfg->eos_gm1 = gamma - 1.;
  cfg->eos_cv = cv;
  // units (mcblock.cpp:84-91)
  cfg->time_scale = r.real("mcblock", "time_scale", 1.);
  cfg->mass_scale = r.real("mcblock", "mass_scale", 1.);
  cfg->length_scale = r.real("mcblock", "length_scale", 1.);
  cfg->temperature_scale = r.real("mcblock", "temperature_scale", 1.);
  // absorption opacity (mcblock.cpp:95-121)
  const std::string om = r.str("mcblock", "opacity_model", "");
  if (om == "none") {
    cfg->opacity_model = JB_OPAC_GRAY;
    cfg->opacity_kappa = 0.0;
  } else if (om == "constant") {
    if (!r.need("mcblock", "opacity_constant_value")) {
      set_err(err, errlen, r.err);
      return JB_ERR_INVALID;
    }
    cfg->opacity_model = JB_OPAC_GRAY;
    cfg->opacity_kappa = r.real("mcblock", "opacity_constant_value", 0.0);
  } else if (om == "ep_bremss") {
    cfg->opacity_model = JB_OPAC_EPBREMSS;
  } else {
    set_err(err, errlen, "Only none or constant opacity models supported!");
    return JB_ERR_INVALID;
  }
  // scattering (mcblock.cpp:124-143)
  cfg->apm = r.real("mcblock", "apm", 1.);
  const std::string sm = r.str("mcblock", "scattering_model", "none");
  if (sm == "none") {
    cfg->scattering_model = JB_SCAT_GRAY;
    cfg->scattering_sigma = 0.0;
  } else if (sm == "constant") {
    if (!r.need("mcblock", "scattering_constant_value")) {
      set_err(err, errlen, r.err);
      return JB_ERR_INVALID;
    }
    cfg->scattering_model = JB_SCAT_GRAY;
    cfg->scattering_sigma = r.real("mcblock", "scattering_constant_value", 0.0);
  } else {
    set_err(err, errlen, "Only none or constant scattering models supported!");
    return JB_ERR_INVALID;
  }
  // jaybenne::Initialize (jaybenne.cpp:163-223)
  cfg->num_particles = r.integer("jaybenne", "num_particles", 0);
  cfg->dt = r.real("jaybenne", "dt", std::numeric_limits<double>::max());
  cfg->min_swarm_occupancy = r.real("jaybenne", "min_swarm_occupancy", 0.);
  if (!(cfg->min_swarm_occupancy >= 0 && cfg->min_swarm_occupancy < 1.0)) {
    set_err(err, errlen, "Minimum allowable swarm occupancy must be >= 0 and less than 1");
    return JB_ERR_INVALID;
  }
  cfg->numin = r.real("jaybenne", "numin", std::numeric_limits<double>::min());
  cfg->numax = r.real("jaybenne", "numax", std::numeric_limits<double>::max());
  cfg->unique_rank_seeds = r.boolean("jaybenne", "unique_rank_seeds", true);
  cfg->seed = r.integer("jaybenne", "seed", 123);
  cfg->max_transport_iterations = (int32_t)r.integer("jaybenne", "max_transport_iterations", 10000);
  cfg->use_ddmc = r.boolean("jaybenne", "use_ddmc", false);
  cfg->tau_ddmc = r.real("jaybenne", "tau_ddmc", 5.0);
  const std::string strategy = r.str("jaybenne", "source_strategy", "uniform");
  if (strategy == "uniform") {
    cfg->source_strategy = JB_STRATEGY_UNIFORM;
  } else if (strategy == "energy") {
    cfg->source_strategy = JB_STRATEGY_ENERGY;
  } else {
    set_err(err, errlen, "Only uniform or energy source strategies supported!");
    return JB_ERR_INVALID;
  }
  cfg->do_emission = r.boolean("jaybenne", "do_emission", true);
  cfg->do_feedback = r.boolean("jaybenne", "do_feedback", true);
  return JB_OK;
}

jb_status jb_deck_mesh(const jb_deck *d, int32_t n_ranks, jb_mesh *mesh, char *err, int32_t errlen) {
  if (!d || !mesh || n_ranks < 1) return JB_ERR_INVALID;
  Reader r{d, ""};
  memset(mesh, 0, sizeof(*mesh));
  long nx[3], nxb[3];
  const char *nxk[3] = {"nx1", "nx2", "nx3"};
  const char *mink[3] = {"x1min", "x2min", "x3min"};
  const char *maxk[3] = {"x1max", "x2max", "x3max"};
  const char *bck[6] = {"ix1_bc", "ox1_bc", "ix2_bc", "ox2_bc", "ix3_bc", "ox3_bc"};
  if (!r.need("parthenon/mesh", "nx1")) {
    set_err(err, errlen, r.err);
    return JB_ERR_INVALID;
  }
  for (int dd = 0; dd < 3; ++dd) {
    nx[dd] = r.integer("parthenon/mesh", nxk[dd], 1);
    nxb[dd] = r.integer("parthenon/meshblock", nxk[dd], nx[dd]);
    mesh->xmin[dd] = r.real("parthenon/mesh", mink[dd], -0.5);
    mesh->xmax[dd] = r.real("parthenon/mesh", maxk[dd], 0.5);
    if (nx[dd] < 1 || nxb[dd] < 1 || nx[dd] % nxb[dd] != 0) {
      set_err(err, errlen, "mesh must be evenly divisible by the meshblock");
      return JB_ERR_INVALID;
    }
    mesh->nxb[dd] = (int32_t)nxb[dd];
    mesh->nrb[dd] = (int32_t)(nx[dd] / nxb[dd]);
  }
  mesh->ndim = nx[2] > 1 ? 3 : (nx[1] > 1 ? 2 : 1);
  mesh->nghost = (int32_t)r.integer("parthenon/mesh", "nghost", 2);
  for (int f = 0; f < 6; ++f) {
    const std::string mb = r.str("parthenon/mesh", bck[f], "periodic");
    int bc = bc_from_name(mb);
    if (bc < 0 || bc == JB_BC_REFLECT) bc = (mb == "reflecting") ? JB_BC_OUTFLOW : (bc < 0 ? JB_BC_OUTFLOW : bc);
    mesh->mesh_bc[f] = bc;
    const std::string sb = r.str("parthenon/swarm", bck[f], mb);
    const int sbc = bc_from_name(sb);
    if (sbc < 0) {
      set_err(err, errlen, "unknown swarm boundary condition '" + sb + "'");
      return JB_ERR_INVALID;
    }
    mesh->swarm_bc[f] = sbc;
  }
  Tree t;
  t.ndim = mesh->ndim;
  for (int dd = 0; dd < 3; ++dd) {
    t.nrb[dd] = mesh->nrb[dd];
    t.periodic[dd] = mesh->mesh_bc[2 * dd] == JB_BC_PERIODIC;
  }
  for (long k = 0; k < t.nrb[2]; ++k)
    for (long j = 0; j < t.nrb[1]; ++j)
      for (long i = 0; i < t.nrb[0]; ++i) t.leaves.insert(Loc{0, {i, j, k}});
  const std::string refinement = r.str("parthenon/mesh", "refinement", "none");
  if (refinement == "static") {
    for (const std::string &bname : d->block_order) {
      if (bname.rfind("parthenon/static_refinement", 0) != 0) continue;
      const char *bn = bname.c_str();
      const int lev = (int)r.integer(bn, "level", 1);
      if (lev < 1) continue;
      long lo[3] = {0, 0, 0}, hi[3] = {0, 0, 0};
      for (int dd = 0; dd < mesh->ndim; ++dd) {
        const double rmin = r.real(bn, mink[dd], mesh->xmin[dd]);
        const double rmax = r.real(bn, maxk[dd], mesh->xmax[dd]);
        const long lxmax = t.nrb[dd] << lev;
        long a, b;
        // Mesh ctor (Athena++/Parthenon): first block whose upper edge lies beyond rmin,
        // first block whose upper edge reaches rmax, then widen to whole parents
        for (a = 0; a < lxmax; ++a)
          if (mesh_generator(a + 1, lxmax, mesh->xmin[dd], mesh->xmax[dd]) > rmin) break;
        for (b = a; b < lxmax; ++b)
          if (mesh_generator(b + 1, lxmax, mesh->xmin[dd], mesh->xmax[dd]) >= rmax) break;
        if (a % 2 == 1) a--;
        if (b % 2 == 0) b++;
        lo[dd] = a;
        hi[dd] = std::min(b, lxmax - 1);
      }
      for (long k = lo[2]; k <= hi[2]; ++k)
        for (long j = lo[1]; j <= hi[1]; ++j)
          for (long i = lo[0]; i <= hi[0]; ++i) t.ensure(Loc{lev, {i, j, k}});
    }
    t.balance();
  } else if (refinement != "none") {
    set_err(err, errlen, "only refinement = none | static is supported");
    return JB_ERR_UNSUPPORTED;
  }
  int max_level = 0;
  for (const Loc &l : t.leaves) max_level = std::max(max_level, l.level);
  std::vector<Loc> order(t.leaves.begin(), t.leaves.end());
  std::sort(order.begin(), order.end(),
            [&](const Loc &a, const Loc &b) { return morton(a, max_level, mesh->ndim) < morton(b, max_level, mesh->ndim); });
  const int nb = (int)order.size();
  jb_block *blocks = (jb_block *)calloc((size_t)nb, sizeof(jb_block));
  for (int g = 0; g < nb; ++g) {
    jb_block &b = blocks[g];
    b.gid = g;
    b.level = order[g].level;
    for (int dd = 0; dd < 3; ++dd) {
      b.lx[dd] = order[g].lx[dd];
      if (dd < mesh->ndim) {
        const long n = t.nrb[dd] << b.level;
        b.xmin[dd] = mesh_generator(b.lx[dd], n, mesh->xmin[dd], mesh->xmax[dd]);
        b.xmax[dd] = mesh_generator(b.lx[dd] + 1, n, mesh->xmin[dd], mesh->xmax[dd]);
      } else {
        b.xmin[dd] = mesh->xmin[dd];
        b.xmax[dd] = mesh->xmax[dd];
      }
    }
    b.rank = (int32_t)(((int64_t)g * n_ranks) / nb);
  }
  mesh->max_level = max_level;
  mesh->nblocks = nb;
  mesh->blocks = blocks;
  return JB_OK;
}

void jb_mesh_free(jb_mesh *mesh) {
  if (mesh && mesh->blocks) {
    free((void *)mesh->blocks);
    mesh->blocks = nullptr;
    mesh->nblocks = 0;
  }
}

int jb_abi_version(void) { return JB_ABI_VERSION; }

}  // extern "C"
