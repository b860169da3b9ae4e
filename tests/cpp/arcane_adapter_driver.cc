// include/arcanefem_b200/ArcaneAdapter.h compiled and run against the stub of the Arcane API (tests/cpp/arcane_stub): the
// adapter must flatten cellNode(), isOwn() and the Real3 coordinates into the b2f_mesh_view the library reads, in both the
// reference mode (all cells, own rows: node_is_own) and the exchange mode (own cells: cell_is_own).  No device needed.
#include <cstdio>
#include <arcanefem_b200/ArcaneAdapter.h>

int main()
{
  Arcane::StubMeshData d;
  d.dim = 2;
  d.cell_node = { 0, 1, 2, 0, 2, 3, 1, 4, 2 }; // three triangles, the last one a ghost cell
  d.cell_own = { 1, 1, 0 };
  d.node_own = { 1, 1, 1, 1, 0 };
  Arcane::IMesh mesh(&d);
  Arcane::Accelerator::RunQueue queue;
  Arcane::VariableNodeReal3 coords({ { 0, 0, 0 }, { 1, 0, 0 }, { 1, 1, 0 }, { 0, 1, 0 }, { 2, 0.5, 0 } });
  int bad = 0;
  for (int exchange = 0; exchange < 2; ++exchange) {
    ArcaneFemB200::ArcaneMeshView v(&mesh, queue, coords, exchange != 0);
    const b2f_mesh_view& m = v.view;
    if (m.dim != 2 || m.nodes_per_cell != 3 || m.nb_cell != 3 || m.nb_node != 5)
      ++bad;
    for (int q = 0; q < 9; ++q)
      if (m.cell_node[q] != d.cell_node[q])
        ++bad;
    if (m.coords[3 * 4 + 0] != 2.0 || m.coords[3 * 4 + 1] != 0.5 || m.coords[3 * 2 + 1] != 1.0)
      ++bad;
    if (exchange) {
      if (m.node_is_own != nullptr || m.cell_is_own == nullptr || m.cell_is_own[2] != 0 || m.cell_is_own[0] != 1)
        ++bad;
    }
    else if (m.cell_is_own != nullptr || m.node_is_own == nullptr || m.node_is_own[4] != 0 || m.node_is_own[3] != 1)
      ++bad;
  }
  if (ArcaneFemB200::streamOf(queue) != nullptr)
    ++bad;
  // the shim's classes are reachable through the adapter header, as a module would use them
  ArcaneFemB200::BSRFormat<1> bsr(ArcaneFemB200::streamOf(queue));
  (void)bsr;
  std::printf("arcane_adapter_driver: bad=%d\n", bad);
  return bad == 0 ? 0 : 2;
}
