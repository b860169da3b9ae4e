// Model of femutils/FemDoFsOnNodes.h (the reference's own class, not on the hot path): dof = node * nb_dof_per_node + i
// (femutils/FemDoFsOnNodes.cc:79-109).  TEST INFRASTRUCTURE, see arcane_mock.h.
#pragma once
#include "arcane_mock.h"
namespace Arcane::FemUtils {
class FemDoFsOnNodes
{
 public:
  explicit FemDoFsOnNodes(ITraceMng* tm) : m_tm(tm) {}
  void initialize(IMesh* mesh, Int32 nb_dof_per_node)
  {
    m_nb_dof_per_node = nb_dof_per_node;
    m_dof_family.m_nb_item = mesh->nbNode() * nb_dof_per_node;
  }
  IItemFamily* dofFamily() const { return const_cast<IItemFamily*>(&m_dof_family); }
  struct NodeDoFView
  {
    Int32 n;
    ARCCORE_HOST_DEVICE DoFLocalId dofId(NodeLocalId node, Int32 i) const { return DoFLocalId(node.localId() * n + i); }
  };
  NodeDoFView nodeDoFConnectivityView() const { return NodeDoFView{ m_nb_dof_per_node }; }

 private:
  ITraceMng* m_tm;
  Int32 m_nb_dof_per_node = 1;
  IItemFamily m_dof_family;
};
} // namespace Arcane::FemUtils
