// arcane_mock.h — a small working model of the Arcane framework types that ArcaneFEM's modules and the adapter headers
// (include/arcanefem_b200/arcane/*.h) touch.  TEST INFRASTRUCTURE: Arcane (>= 3.14.14) is not installed in this image nor on
// the GPU box, so the adapter and the unchanged module text (tests/cpp/poisson_module_unchanged.cu) are compiled and RUN
// against this model.  Names, argument lists and semantics follow Arcane's public headers as the reference uses them
// (femutils/BSRFormat.h:23-55, femutils/ArcaneFemFunctionsGpu.h:18-49, modules/poisson/FemModule.{h,cc}); everything lives in
// unified memory like Arcane's accelerator-side arrays, so views are usable from host and device code.
#pragma once
#include <chrono>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <iostream>
#include <sstream>
#include <stdexcept>
#include <string>
#include <vector>
#include <cuda_runtime.h>

#ifndef ARCCORE_HOST_DEVICE
#define ARCCORE_HOST_DEVICE __host__ __device__
#endif
#define ARCANE_CHECK_PTR(p) \
  do {                      \
  } while (0)
#define ARCANE_CHECK_POINTER(p)                                  \
  do {                                                           \
    if ((p) == nullptr)                                          \
      throw ::Arcane::FatalErrorException("null pointer: " #p); \
  } while (0)
#define ARCANE_CHECK_AT(i, n)
#define ARCANE_THROW(exception_class, ...) throw exception_class(::Arcane::mock::join(__VA_ARGS__))
#define ARCANE_FATAL(...) throw ::Arcane::FatalErrorException(::Arcane::mock::join(__VA_ARGS__))
#define A_FUNCINFO __func__

namespace Arcane {

using Real = double;
using Int32 = int32_t;
using Int64 = int64_t;
using Integer = int32_t;
using Byte = unsigned char;

namespace mock {
  template <class... A>
  inline std::string join(const A&... a)
  {
    std::ostringstream o;
    ((o << a << ' '), ...);
    return o.str();
  }
  // unified memory when a device is there; plain host memory otherwise (compile + host-only checks on the CPU container)
  inline bool has_device()
  {
    static int n = [] { int c = 0; return cudaGetDeviceCount(&c) == cudaSuccess ? c : 0; }();
    return n > 0;
  }
  template <class T>
  inline T* alloc(size_t n)
  {
    void* p = nullptr;
    if (n == 0)
      n = 1;
    if (has_device()) {
      if (cudaMallocManaged(&p, n * sizeof(T)) != cudaSuccess)
        throw std::bad_alloc();
      memset(p, 0, n * sizeof(T));
    }
    else
      p = calloc(n, sizeof(T));
    return static_cast<T*>(p);
  }
  inline void release(void* p)
  {
    if (!p)
      return;
    if (has_device())
      cudaFree(p);
    else
      free(p);
  }
} // namespace mock

struct FatalErrorException : std::runtime_error { using std::runtime_error::runtime_error; };
struct ArgumentException : std::invalid_argument { using std::invalid_argument::invalid_argument; };
struct NotImplementedException : std::logic_error { using std::logic_error::logic_error; };

class String
{
 public:
  String() = default;
  String(const char* s) : m_s(s) {}
  String(const std::string& s) : m_s(s) {}
  const char* localstr() const { return m_s.c_str(); }
  bool operator==(const String& o) const { return m_s == o.m_s; }
  bool operator!=(const String& o) const { return m_s != o.m_s; }
  bool operator==(const char* o) const { return m_s == o; }
  bool operator!=(const char* o) const { return m_s != o; }
  friend std::ostream& operator<<(std::ostream& o, const String& s) { return o << s.m_s; }

 private:
  std::string m_s;
};

template <class T>
class UniqueArray : public std::vector<T>
{
 public:
  using std::vector<T>::vector;
  Integer size() const { return (Integer)std::vector<T>::size(); }
  void add(const T& v) { this->push_back(v); }
};
template <class T> using ConstArrayView = const UniqueArray<T>&;

class StringList : public UniqueArray<String> {};
class CommandLineArguments
{
 public:
  CommandLineArguments() = default;
  explicit CommandLineArguments(const StringList& l) : m_args(l) {}
  StringList m_args;
};

struct Real2
{
  Real x = 0, y = 0;
  Real2() = default;
  ARCCORE_HOST_DEVICE Real2(Real ax, Real ay) : x(ax), y(ay) {}
};
struct Real3
{
  Real x, y, z;
  ARCCORE_HOST_DEVICE Real operator[](int i) const { return i == 0 ? x : (i == 1 ? y : z); }
  ARCCORE_HOST_DEVICE Real3 operator-(const Real3& o) const { return Real3{ x - o.x, y - o.y, z - o.z }; }
};
namespace math {
  ARCCORE_HOST_DEVICE inline Real3 cross(const Real3& a, const Real3& b) { return Real3{ a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x }; }
  ARCCORE_HOST_DEVICE inline Real dot(const Real3& a, const Real3& b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
} // namespace math

namespace platform {
  inline Real getRealTime() { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); }
} // namespace platform

// --- trace
class TraceMessage
{
 public:
  explicit TraceMessage(bool on) : m_on(on) {}
  ~TraceMessage() { if (m_on) std::cout << "*I-mock  " << m_o.str() << "\n"; }
  template <class T> TraceMessage& operator<<(const T& v) { m_o << v; return *this; }

 private:
  bool m_on;
  std::ostringstream m_o;
};
class ITraceMng { public: bool verbose = false; TraceMessage info() { return TraceMessage(verbose); } };
class TraceAccessor
{
 public:
  explicit TraceAccessor(ITraceMng* tm) : m_tm(tm) {}
  TraceMessage info() const { return m_tm->info(); }
  ITraceMng* traceMng() const { return m_tm; }

 private:
  ITraceMng* m_tm;
};

// --- local ids and items
template <class Tag>
struct LocalIdT
{
  Int32 m_local_id = -1;
  LocalIdT() = default;
  ARCCORE_HOST_DEVICE explicit LocalIdT(Int32 v) : m_local_id(v) {}
  ARCCORE_HOST_DEVICE Int32 localId() const { return m_local_id; }
  ARCCORE_HOST_DEVICE operator Int32() const { return m_local_id; }
};
using CellLocalId = LocalIdT<struct CellTag>;
using NodeLocalId = LocalIdT<struct NodeTag>;
using FaceLocalId = LocalIdT<struct FaceTag>;
using DoFLocalId = LocalIdT<struct DoFTag>;
struct Cell { Int32 lid; Int32 localId() const { return lid; } }; // host-side item of the legacy CPU loops (not on the path)

template <class Id>
struct IdSpan // what `cn_cv.nodes(cell)` / `fn_cv.nodes(face)` iterate
{
  const Int32* p;
  Int32 n;
  struct It
  {
    const Int32* q;
    ARCCORE_HOST_DEVICE Id operator*() const { return Id(*q); }
    ARCCORE_HOST_DEVICE It& operator++() { ++q; return *this; }
    ARCCORE_HOST_DEVICE bool operator!=(const It& o) const { return q != o.q; }
  };
  ARCCORE_HOST_DEVICE It begin() const { return It{ p }; }
  ARCCORE_HOST_DEVICE It end() const { return It{ p + n }; }
  ARCCORE_HOST_DEVICE Int32 size() const { return n; }
  ARCCORE_HOST_DEVICE Id operator[](Int32 i) const { return Id(p[i]); }
};

// Indexed{Cell,Face}NodeConnectivityView: fixed arity in this model (P1 meshes)
template <class ItemId>
struct IndexedItemNodeConnectivityView
{
  const Int32* m_data = nullptr;
  Int32 m_arity = 0;
  ARCCORE_HOST_DEVICE NodeLocalId nodeId(ItemId item, Int32 i) const { return NodeLocalId(m_data[(Int64)item.localId() * m_arity + i]); }
  ARCCORE_HOST_DEVICE Int32 nbNode(ItemId) const { return m_arity; }
  ARCCORE_HOST_DEVICE Int32 nbItem(ItemId) const { return m_arity; }
  ARCCORE_HOST_DEVICE IdSpan<NodeLocalId> nodes(ItemId item) const { return IdSpan<NodeLocalId>{ m_data + (Int64)item.localId() * m_arity, m_arity }; }
  ARCCORE_HOST_DEVICE IdSpan<NodeLocalId> nodeIds(ItemId item) const { return nodes(item); }
};
using IndexedCellNodeConnectivityView = IndexedItemNodeConnectivityView<CellLocalId>;
using IndexedFaceNodeConnectivityView = IndexedItemNodeConnectivityView<FaceLocalId>;

class IItemFamily
{
 public:
  Int32 m_nb_item = 0;
  const Byte* m_is_own = nullptr; // unified memory
  const Byte* m_boundary_outside = nullptr; // faces
  Int32 nbItem() const { return m_nb_item; }
  Int32 maxLocalId() const { return m_nb_item; }
};
struct ItemGenericInfoListView
{
  const Byte* m_is_own = nullptr;
  const Byte* m_outside = nullptr;
  ItemGenericInfoListView() = default;
  explicit ItemGenericInfoListView(IItemFamily* f) : m_is_own(f->m_is_own), m_outside(f->m_boundary_outside) {}
  template <class Id> ARCCORE_HOST_DEVICE bool isOwn(Id id) const { return m_is_own == nullptr || m_is_own[id.localId()] != 0; }
  template <class Id> ARCCORE_HOST_DEVICE bool isSubDomainBoundaryOutside(Id id) const { return m_outside == nullptr || m_outside[id.localId()] != 0; }
};
using NodeInfoListView = ItemGenericInfoListView;
using FaceInfoListView = ItemGenericInfoListView;

// --- groups: a list of local ids in unified memory
struct Int32ConstArrayView
{
  const Int32* m_p = nullptr;
  Int32 m_n = 0;
  const Int32* data() const { return m_p; }
  Int32 size() const { return m_n; }
  Int32 operator[](Int32 i) const { return m_p[i]; }
};
struct ItemVectorView { Int32ConstArrayView m_ids; Int32ConstArrayView localIds() const { return m_ids; } Int32 size() const { return m_ids.size(); } };
class ItemGroup
{
 public:
  ItemGroup() = default;
  ItemGroup(const String& name, const std::vector<Int32>& ids) : m_name(name), m_n((Int32)ids.size())
  {
    m_ids = mock::alloc<Int32>(ids.size());
    for (size_t i = 0; i < ids.size(); ++i)
      m_ids[i] = ids[i];
  }
  Int32 size() const { return m_n; }
  bool null() const { return m_ids == nullptr; }
  const String& name() const { return m_name; }
  ItemVectorView view() const { return ItemVectorView{ Int32ConstArrayView{ m_ids, m_n } }; }

 protected:
  String m_name;
  Int32* m_ids = nullptr; // never freed: groups live as long as the test process
  Int32 m_n = 0;
};
class NodeGroup : public ItemGroup { public: using ItemGroup::ItemGroup; };
class CellGroup : public ItemGroup { public: using ItemGroup::ItemGroup; };
class FaceGroup : public ItemGroup
{
 public:
  FaceGroup() = default;
  FaceGroup(const String& name, const std::vector<Int32>& face_ids, const NodeGroup& nodes) : ItemGroup(name, face_ids), m_nodes(nodes) {}
  NodeGroup nodeGroup() const { return m_nodes; }

 private:
  NodeGroup m_nodes;
};

// --- variables
template <class T>
class MeshVariableArrayT
{
 public:
  MeshVariableArrayT() = default;
  explicit MeshVariableArrayT(Int64 n) : m_n(n), m_p(mock::alloc<T>((size_t)n)) {}
  struct ArrayView { T* m_p; Int64 m_n; T* data() const { return m_p; } Int64 size() const { return m_n; } };
  ArrayView asArray() const { return ArrayView{ m_p, m_n }; }
  template <class Id> T& operator[](Id id) const { return m_p[id.localId()]; }
  void fill(const T& v) { for (Int64 i = 0; i < m_n; ++i) m_p[i] = v; }
  T* m_p_raw() const { return m_p; }

 private:
  Int64 m_n = 0;
  T* m_p = nullptr; // shared by copies (an Arcane variable is a reference to the mesh's storage)
};
using VariableNodeReal3 = MeshVariableArrayT<Real3>;
using VariableNodeReal = MeshVariableArrayT<Real>;

class IMesh;
class UnstructuredMeshConnectivityView
{
 public:
  UnstructuredMeshConnectivityView() = default;
  explicit UnstructuredMeshConnectivityView(IMesh* mesh) { setMesh(mesh); }
  inline void setMesh(IMesh* mesh);
  IndexedCellNodeConnectivityView cellNode() const { return m_cn; }
  IndexedFaceNodeConnectivityView faceNode() const { return m_fn; }

 private:
  IndexedCellNodeConnectivityView m_cn;
  IndexedFaceNodeConnectivityView m_fn;
};

class IMesh
{
 public:
  // host description -> unified-memory arrays
  IMesh(Int32 dim, const std::vector<Int32>& cell_node, const std::vector<Real3>& coords, const std::vector<Int32>& face_node, ITraceMng* tm)
  : m_dim(dim), m_coords((Int64)coords.size()), m_tm(tm)
  {
    const Int32 nv = dim + 1;
    m_cell_node = mock::alloc<Int32>(cell_node.size());
    memcpy(m_cell_node, cell_node.data(), cell_node.size() * sizeof(Int32));
    m_face_node = mock::alloc<Int32>(face_node.size());
    memcpy(m_face_node, face_node.data(), face_node.size() * sizeof(Int32));
    for (size_t i = 0; i < coords.size(); ++i)
      m_coords.asArray().data()[i] = coords[i];
    m_cell_family.m_nb_item = (Int32)(cell_node.size() / nv);
    m_node_family.m_nb_item = (Int32)coords.size();
    m_face_family.m_nb_item = (Int32)(face_node.size() / dim);
    std::vector<Int32> all_c((size_t)m_cell_family.m_nb_item), all_n((size_t)m_node_family.m_nb_item);
    for (size_t i = 0; i < all_c.size(); ++i)
      all_c[i] = (Int32)i;
    for (size_t i = 0; i < all_n.size(); ++i)
      all_n[i] = (Int32)i;
    m_all_cells = CellGroup("AllCells", all_c);
    m_all_nodes = NodeGroup("AllNodes", all_n);
  }
  Int32 dimension() const { return m_dim; }
  Int32 nbNode() const { return m_node_family.m_nb_item; }
  Int32 nbCell() const { return m_cell_family.m_nb_item; }
  Int32 nbFace() const { return m_face_family.m_nb_item; }
  CellGroup allCells() const { return m_all_cells; }
  NodeGroup allNodes() const { return m_all_nodes; }
  IItemFamily* nodeFamily() { return &m_node_family; }
  IItemFamily* cellFamily() { return &m_cell_family; }
  IItemFamily* faceFamily() { return &m_face_family; }
  VariableNodeReal3& nodesCoordinates() { return m_coords; }
  ITraceMng* traceMng() const { return m_tm; }
  const Int32* mockCellNode() const { return m_cell_node; }
  const Int32* mockFaceNode() const { return m_face_node; }

 private:
  Int32 m_dim;
  Int32* m_cell_node = nullptr;
  Int32* m_face_node = nullptr;
  VariableNodeReal3 m_coords;
  IItemFamily m_cell_family, m_node_family, m_face_family;
  CellGroup m_all_cells;
  NodeGroup m_all_nodes;
  ITraceMng* m_tm;
};
inline void UnstructuredMeshConnectivityView::setMesh(IMesh* mesh)
{
  m_cn = IndexedCellNodeConnectivityView{ mesh->mockCellNode(), mesh->dimension() + 1 };
  m_fn = IndexedFaceNodeConnectivityView{ mesh->mockFaceNode(), mesh->dimension() };
}

// --- accelerator
enum class eMemoryRessource { UnifiedMemory };
namespace Accelerator {
  class Runner {};
  class RunQueue
  {
   public:
    RunQueue() = default;
    explicit RunQueue(cudaStream_t s) : m_stream(s) {}
    void* platformStream() const { return m_stream; } // Arcane: the native stream (cudaStream_t) of the queue
    eMemoryRessource memoryRessource() const { return eMemoryRessource::UnifiedMemory; }
    void barrier() const { cudaStreamSynchronize(m_stream); }

   private:
    cudaStream_t m_stream = nullptr;
  };
  struct RunCommand { const RunQueue* queue; };
  inline RunCommand makeCommand(const RunQueue& q) { return RunCommand{ &q }; }
  inline RunCommand makeCommand(const RunQueue* q) { return RunCommand{ q }; }
  struct VariableNodeReal3InView
  {
    const Real3* m_p;
    ARCCORE_HOST_DEVICE Real3 operator[](NodeLocalId n) const { return m_p[n.localId()]; }
  };
  inline VariableNodeReal3InView viewIn(RunCommand&, const VariableNodeReal3& v) { return VariableNodeReal3InView{ v.asArray().data() }; }
  class IAcceleratorMng
  {
   public:
    RunQueue* defaultQueue() { return &m_queue; }
    Runner* defaultRunner() { return &m_runner; }

   private:
    RunQueue m_queue;
    Runner m_runner;
  };
} // namespace Accelerator

// --- sub-domain / module plumbing the module text reaches
class ICaseMng { public: void setTreatWarningAsError(bool) {} void setAllowUnkownRootElelement(bool) {} };
class ITimeLoopMng { public: bool stopped = false; void stopComputeLoop(bool) { stopped = true; } };
class ISubDomain
{
 public:
  ITraceMng* traceMng() { return &m_tm; }
  Accelerator::IAcceleratorMng* acceleratorMng() { return &m_acc; }
  ICaseMng* caseMng() { return &m_case; }
  ITimeLoopMng* timeLoopMng() { return &m_tl; }

 private:
  ITraceMng m_tm;
  Accelerator::IAcceleratorMng m_acc;
  ICaseMng m_case;
  ITimeLoopMng m_tl;
};
class ModuleBuildInfo
{
 public:
  ModuleBuildInfo(ISubDomain* sd, IMesh* mesh) : m_sd(sd), m_mesh(mesh) {}
  ISubDomain* subDomain() const { return m_sd; }
  IMesh* mesh() const { return m_mesh; }

 private:
  ISubDomain* m_sd;
  IMesh* m_mesh;
};
class VersionInfo { public: VersionInfo(int, int, int) {} };

} // namespace Arcane
