// DECLARATION-ONLY stand-in for the dozen Parthenon types integration/jaybenne_adapter.cpp touches
// (parthenon-hpc-lab/parthenon @ ed149893, the revision lanl/jaybenne pins as a submodule — absent from
// this image). It exists so that the adapter is real, type-checked code: tests/test_adapter.py compiles
// the adapter against it with `g++ -fsyntax-only`. Nothing here has a body and nothing links; on a machine
// with Parthenon the adapter is compiled against the real <parthenon/package.hpp> instead.
// Names and signatures follow Parthenon's public interface as used by the reference's own sources
// (src/jaybenne/*.cpp, src/mcblock/*.cpp cite them at every call site).
#ifndef JB_MOCK_PARTHENON_PACKAGE_HPP_
#define JB_MOCK_PARTHENON_PACKAGE_HPP_

#include <array>
#include <cstdint>
#include <functional>
#include <memory>
#include <string>
#include <vector>

#define KOKKOS_LAMBDA [=]
#define KOKKOS_INLINE_FUNCTION inline
#define KOKKOS_FORCEINLINE_FUNCTION inline
#define DEFAULT_LOOP_PATTERN parthenon::loop_pattern_mdrange_tag
#define PARTHENON_REQUIRE(cond, msg) \
  do {                               \
    if (!(cond)) parthenon::fail(msg); \
  } while (0)
#define PARTHENON_FAIL(msg) parthenon::fail(msg)

namespace parthenon {
using Real = double;
[[noreturn]] void fail(const char *msg);
[[noreturn]] void fail(const std::string &msg);

struct LoopPatternMDRange {};
constexpr LoopPatternMDRange loop_pattern_mdrange_tag{};
struct DevExecSpace {
  void fence() const;
};
template <typename F>
void par_for(LoopPatternMDRange, const std::string &, DevExecSpace, int, int, int, int, int, int, const F &) {
}  // (a template instantiated with a local lambda type needs a definition to type-check: empty)

// Kokkos::View<Real*, LayoutRight, DevMemSpace>
template <typename T>
struct ParArray1D {
  ParArray1D() = default;
  ParArray1D(const std::string &label, size_t n);
  T *data() const;
  size_t size() const;
  T &operator()(size_t i) const;
};
template <typename T>
struct ParArrayND {
  T &operator()(int k, int j, int i) const;
  T &operator()(int n, int k, int j, int i) const;
};

namespace Globals {
extern int my_rank, nranks;
}

enum CoordinateDirection { NODIR = 0, X1DIR = 1, X2DIR = 2, X3DIR = 3 };
enum class IndexDomain { entire, interior };
enum class BoundaryFace { undef = -1, inner_x1 = 0, outer_x1 = 1, inner_x2 = 2, outer_x2 = 3, inner_x3 = 4, outer_x3 = 5 };
enum class BoundaryFlag { block = -1, undef, outflow, periodic, reflect, user };
enum class TaskStatus { complete, incomplete, iterate };
enum class TaskListStatus { complete, fail };
enum class AmrTag : int { derefine = -1, same = 0, refine = 1 };

struct RegionSize {
  Real xmin(CoordinateDirection d) const;
  Real xmax(CoordinateDirection d) const;
  int nx(CoordinateDirection d) const;
  bool symmetry(CoordinateDirection d) const;
};
struct LogicalLocation {
  int level() const;
  std::int64_t lx1() const;
  std::int64_t lx2() const;
  std::int64_t lx3() const;
};
struct IndexShape {
  int is(IndexDomain d) const;
  int ie(IndexDomain d) const;
  int js(IndexDomain d) const;
  int je(IndexDomain d) const;
  int ks(IndexDomain d) const;
  int ke(IndexDomain d) const;
};

class ParameterInput {
 public:
  int GetInteger(const std::string &block, const std::string &name);
  Real GetReal(const std::string &block, const std::string &name);
  std::string GetString(const std::string &block, const std::string &name);
  int GetOrAddInteger(const std::string &block, const std::string &name, int def);
  Real GetOrAddReal(const std::string &block, const std::string &name, Real def);
  bool GetOrAddBoolean(const std::string &block, const std::string &name, bool def);
  std::string GetOrAddString(const std::string &block, const std::string &name, const std::string &def);
  bool DoesParameterExist(const std::string &block, const std::string &name);
  bool DoesBlockExist(const std::string &block);
};

struct Metadata {
  enum Flag { Cell, Face, Independent, OneCopy, Derived, FillGhost, Provides, None, Real, Integer, Vector };
  Metadata() = default;
  Metadata(std::initializer_list<Flag> flags);
  Metadata(std::initializer_list<Flag> flags, const std::vector<int> &shape);
};

template <typename T>
class MeshData;
template <typename T>
class MeshBlockData;
class Mesh;
class MeshBlock;
class Swarm;

class StateDescriptor {
 public:
  explicit StateDescriptor(const std::string &label);
  template <typename T>
  void AddParam(const std::string &key, T value, bool is_mutable = false);
  template <typename T>
  const T &Param(const std::string &key) const;
  template <typename T>
  T *MutableParam(const std::string &key) const;
  bool AddField(const std::string &name, const Metadata &m);
  bool AddSwarm(const std::string &name, const Metadata &m);
  bool AddSwarmValue(const std::string &value_name, const std::string &swarm_name, const Metadata &m);
  std::function<Real(MeshData<Real> *)> EstimateTimestepMesh;
};
struct Packages_t {
  void Add(const std::shared_ptr<StateDescriptor> &pkg);
  const std::shared_ptr<StateDescriptor> &Get(const std::string &name) const;
};

template <typename T>
struct Variable {
  ParArrayND<T> data;
};
template <typename T>
class MeshBlockData {
 public:
  Variable<T> &Get(const std::string &name) const;
  MeshBlock *GetBlockPointer() const;
  Mesh *GetParentPointer() const;
};
template <typename T>
class MeshData {
 public:
  Mesh *GetParentPointer() const;
  int NumBlocks() const;
  const std::shared_ptr<MeshBlockData<T>> &GetBlockData(int n) const;
};
template <typename T>
struct DataCollection {
  std::shared_ptr<T> &Get(const std::string &label = "base");
  std::shared_ptr<T> &GetOrAdd(const std::string &label, int partition);
};

class MeshBlock {
 public:
  int gid, lid;
  LogicalLocation loc;
  RegionSize block_size;
  IndexShape cellbounds;
  Mesh *pmy_mesh;
  DataCollection<MeshBlockData<Real>> meshblock_data;
};
class Mesh {
 public:
  int ndim, nbtotal;
  RegionSize mesh_size;
  std::array<BoundaryFlag, 6> mesh_bcs;
  std::vector<std::shared_ptr<MeshBlock>> block_list;
  Packages_t packages;
  DataCollection<MeshData<Real>> mesh_data;
  int DefaultNumPartitions();
  int GetNumMeshBlocksThisRank() const;
  const std::vector<LogicalLocation> &GetLocList() const;
  int GetRootLevel() const;
  RegionSize GetBlockSize() const;  // size of every meshblock in cells (<parthenon/meshblock>)
};

class Swarm {};

// tasking
struct TaskID {
  TaskID() = default;
  explicit TaskID(int);
};
class TaskList {
 public:
  template <class F, class... Args>
  TaskID AddTask(TaskID, F &&, Args &&...) {
    return TaskID();  // (see par_for above)
  }
};
class TaskRegion {
 public:
  TaskList &operator[](int i);
};
class TaskCollection {
 public:
  TaskRegion &AddRegion(int num_lists);
  TaskListStatus Execute();
};
}  // namespace parthenon

namespace parthenon::package::prelude {
using parthenon::Real;
}
namespace parthenon::driver::prelude {
using parthenon::Real;
}
#endif  // JB_MOCK_PARTHENON_PACKAGE_HPP_
