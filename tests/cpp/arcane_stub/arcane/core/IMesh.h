// STUB of the part of the Arcane API that include/arcanefem_b200/ArcaneAdapter.h touches — test infrastructure only.
// The Arcane framework (>= 3.14.14, the reference's only hard dependency) is not installed here; these few types have the
// names, members and value semantics the adapter uses (IMesh::dimension/allCells/allNodes, Cell::nodeId/localId/isOwn,
// ENUMERATE_, NumArray<T, MDDim1>, VariableNodeReal3::asArray, RunQueue::memoryRessource/platformStream), so that the
// adapter is at least compiled and run against a small mesh (tests/cpp/arcane_adapter_driver.cc).
#pragma once
#include <cstdint>
#include <vector>

namespace Arcane {
using Int32 = int32_t;
using Int64 = int64_t;
using Byte = unsigned char;
using Real = double;
struct Real3 { Real x, y, z; };
enum class eMemoryRessource { Host, UnifiedMemory, Device };
struct MDDim1 {};

struct StubMeshData
{
  Int32 dim = 2;
  std::vector<Int32> cell_node; // nb_cell * (dim + 1)
  std::vector<Byte> cell_own, node_own;
};

class Node
{
 public:
  Node(const StubMeshData* d, Int32 lid) : m_d(d), m_lid(lid) {}
  Int32 localId() const { return m_lid; }
  bool isOwn() const { return m_d->node_own[m_lid] != 0; }
 private:
  const StubMeshData* m_d;
  Int32 m_lid;
};

class Cell
{
 public:
  Cell(const StubMeshData* d, Int32 lid) : m_d(d), m_lid(lid) {}
  Int32 localId() const { return m_lid; }
  bool isOwn() const { return m_d->cell_own[m_lid] != 0; }
  Int32 nodeId(Int32 a) const { return m_d->cell_node[m_lid * (m_d->dim + 1) + a]; }
 private:
  const StubMeshData* m_d;
  Int32 m_lid;
};

template <class ItemT>
class ItemEnumeratorT
{
 public:
  ItemEnumeratorT(const StubMeshData* d, Int32 n) : m_d(d), m_n(n), m_item(d, 0) {}
  bool hasNext() const { return m_i < m_n; }
  void operator++() { ++m_i; m_item = ItemT(m_d, m_i); }
  ItemT operator*() const { return m_item; }
  const ItemT* operator->() const { return &m_item; }
 private:
  const StubMeshData* m_d;
  Int32 m_n, m_i = 0;
  ItemT m_item;
};

template <class ItemT>
class ItemGroupT
{
 public:
  ItemGroupT(const StubMeshData* d, Int32 n) : m_d(d), m_n(n) {}
  Int32 size() const { return m_n; }
  ItemEnumeratorT<ItemT> enumerator() const { return ItemEnumeratorT<ItemT>(m_d, m_n); }
 private:
  const StubMeshData* m_d;
  Int32 m_n;
};
using CellGroup = ItemGroupT<Cell>;
using NodeGroup = ItemGroupT<Node>;

class IMesh
{
 public:
  explicit IMesh(const StubMeshData* d) : m_d(d) {}
  Int32 dimension() const { return m_d->dim; }
  CellGroup allCells() const { return CellGroup(m_d, (Int32)(m_d->cell_node.size() / (m_d->dim + 1))); }
  NodeGroup allNodes() const { return NodeGroup(m_d, (Int32)m_d->node_own.size()); }
 private:
  const StubMeshData* m_d;
};
} // namespace Arcane

#define ENUMERATE_(type, name, group) for (::Arcane::ItemEnumeratorT<::Arcane::type> name((group).enumerator()); name.hasNext(); ++name)
