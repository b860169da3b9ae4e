// arcanefem_b200/ArcaneAdapter.h — glue between Arcane objects and the C ABI.  NEEDS THE ARCANE FRAMEWORK (>= 3.14.14,
// reference CMakeLists.txt:10), which is not installed here: in this repository it is compiled and run only against a stub
// of the Arcane types it touches (tests/cpp/arcane_stub, tests/test_cpp_shim.py).  It is the code a maintainer adds next to
// femutils/ when switching a module to libb200fem.so.
//
// It fills a b2f_mesh_view from the same Arcane views the reference kernels read (femutils/BSRFormat.h:789-792,
// femutils/ArcaneFemFunctionsGpu.h:66-83): cellNode() connectivity, VariableNodeReal3 coordinates, isOwn flags.
#pragma once
#if __has_include(<arcane/core/IMesh.h>)
#include <arcane/core/IMesh.h>
#include <arcane/core/ItemGroup.h>
#include <arcane/core/UnstructuredMeshConnectivity.h>
#include <arcane/core/VariableTypes.h>
#include <arcane/accelerator/core/RunQueue.h>
#include <arcane/utils/NumArray.h>
#include "BSRFormat.h"
#include "BoundaryConditions.h"

namespace ArcaneFemB200 {

// Owns the flat copies the library reads (Arcane stores connectivity with an indirection and Real3 coordinates already as
// 3 contiguous doubles; the copies live in the queue's memory resource, i.e. device-accessible memory).
struct ArcaneMeshView
{
  Arcane::NumArray<Arcane::Int32, Arcane::MDDim1> cell_node;
  Arcane::NumArray<Arcane::Byte, Arcane::MDDim1> node_is_own, cell_is_own;
  b2f_mesh_view view{};

  ArcaneMeshView(Arcane::IMesh* mesh, Arcane::Accelerator::RunQueue& queue, const Arcane::VariableNodeReal3& node_coord, bool exchange_mode)
  : cell_node(queue.memoryRessource()), node_is_own(queue.memoryRessource()), cell_is_own(queue.memoryRessource())
  {
    const Arcane::Int32 dim = mesh->dimension();
    const Arcane::Int32 nv = dim + 1;
    const Arcane::Int32 nb_cell = mesh->allCells().size(), nb_node = mesh->allNodes().size();
    cell_node.resize(nb_cell * nv);
    node_is_own.resize(nb_node);
    cell_is_own.resize(nb_cell);
    ENUMERATE_ (Cell, icell, mesh->allCells()) {
      Arcane::Cell cell = *icell;
      for (Arcane::Int32 a = 0; a < nv; ++a)
        cell_node[cell.localId() * nv + a] = cell.nodeId(a);
      cell_is_own[cell.localId()] = cell.isOwn() ? 1 : 0;
    }
    ENUMERATE_ (Node, inode, mesh->allNodes())
      node_is_own[inode->localId()] = inode->isOwn() ? 1 : 0;
    view.dim = dim;
    view.nodes_per_cell = nv;
    view.nb_cell = nb_cell;
    view.nb_node = nb_node;
    view.cell_node = cell_node.to1DSpan().data();
    view.coords = reinterpret_cast<const double*>(node_coord.asArray().data()); // Real3 = {x, y, z} doubles
    // reference mode: all cells assembled, own rows kept (BSRFormat.h:815); exchange mode: own cells, all rows, then b2f_ghost_exchange
    view.node_is_own = exchange_mode ? nullptr : node_is_own.to1DSpan().data();
    view.cell_is_own = exchange_mode ? cell_is_own.to1DSpan().data() : nullptr;
  }
};

inline b2f_stream streamOf(Arcane::Accelerator::RunQueue& queue) { return queue.platformStream(); } // cudaStream_t of the RunQueue

} // namespace ArcaneFemB200
#endif
