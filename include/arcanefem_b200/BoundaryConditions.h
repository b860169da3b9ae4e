// arcanefem_b200/BoundaryConditions.h — mirror of FemUtils::Gpu::BoundaryConditions{,2D,3D}
// (femutils/ArcaneFemFunctionsGpu.h:434-768) over the C ABI.  The reference passes (DoFLinearSystem&, IMesh*, RunQueue*,
// group, ...) and reads the BC values from IArcaneFemBC option objects; here a group is a device list of local node ids
// (Dirichlet) or of boundary-face node tuples (Neumann) and the values are plain arguments.
#pragma once
#include "BSRFormat.h"

namespace ArcaneFemB200 { namespace Gpu {

// node group as device array of Int32 local ids (what `group.nodeGroup()` enumerates)
struct NodeGroup { const Int32* ids = nullptr; Int64 n = 0; };
// face group: node tuples (2 per edge in 2D, 3 per triangle in 3D); outside[f] = isSubDomainBoundaryOutside(), NULL = all true
struct FaceGroup { const Int32* face_node = nullptr; Int64 n = 0; const uint8_t* outside = nullptr; };

namespace BoundaryConditions {
  // applyDirichletViaPenalty / applyDirichletToNodeGroupViaPenalty, ArcaneFemFunctionsGpu.h:434-456,501-512
  inline void applyDirichletViaPenalty(DoFLinearSystem& ls, const NodeGroup& g, Real value, Real penalty, b2f_stream stream = nullptr)
  {
    check(b2f_dirichlet_flag_penalty(ls.handle(), g.ids, g.n, 0, value, penalty, stream));
  }
  // WeakPenalty of modules/testlab/FemModule.cc:687-717 (a_ii += P; the reference has it on the CPU only)
  inline void applyDirichletViaWeakPenalty(DoFLinearSystem& ls, const NodeGroup& g, Real value, Real penalty, b2f_stream stream = nullptr)
  {
    check(b2f_dirichlet_flag_weak_penalty(ls.handle(), g.ids, g.n, 0, value, penalty, stream));
  }
  // vectorial variant :513-545: one call per component whose value string is not "NULL"
  inline void applyDirichletViaPenaltyVectorial(DoFLinearSystem& ls, const NodeGroup& g, int component, Real value, Real penalty, b2f_stream stream = nullptr)
  {
    check(b2f_dirichlet_flag_penalty(ls.handle(), g.ids, g.n, component, value, penalty, stream));
  }
  // applyDirichletToNodeGroupViaRowElimination<ELIMINATE_ROW>, :461-482,546-586
  inline void applyDirichletViaRowElimination(DoFLinearSystem& ls, const NodeGroup& g, Real value, b2f_stream stream = nullptr)
  {
    check(b2f_dirichlet_flag_elimination(ls.handle(), g.ids, g.n, 0, value, B2F_ELIMINATE_ROW, stream));
  }
  inline void applyDirichletViaRowEliminationVectorial(DoFLinearSystem& ls, const NodeGroup& g, int component, Real value, b2f_stream stream = nullptr)
  {
    check(b2f_dirichlet_flag_elimination(ls.handle(), g.ids, g.n, component, value, B2F_ELIMINATE_ROW, stream));
  }
  // row + column elimination with the semantics of femutils/AlephDoFLinearSystem.cc:501-583 (CPU-only in the reference)
  inline void applyDirichletViaRowColumnElimination(DoFLinearSystem& ls, const NodeGroup& g, int component, Real value, b2f_stream stream = nullptr)
  {
    check(b2f_dirichlet_flag_elimination(ls.handle(), g.ids, g.n, component, value, B2F_ELIMINATE_ROW_COLUMN, stream));
  }
  // applyPointDirichletViaPenalty, :560-586: a point group is a node group
  inline void applyPointDirichletViaPenalty(DoFLinearSystem& ls, const NodeGroup& g, Real value, Real penalty, b2f_stream stream = nullptr)
  {
    applyDirichletViaPenalty(ls, g, value, penalty, stream);
  }
} // namespace BoundaryConditions

// BoundaryConditions2D / 3D share the implementation: the dimension is a property of the mesh behind the handle.
namespace BoundaryConditionsND {
  // applyConstantSourceToRhs, :402-429,605-608,695-698
  inline void applyConstantSourceToRhs(Real qdot, DoFLinearSystem& ls, b2f_stream stream = nullptr) { check(b2f_rhs_constant_source(ls.handle(), qdot, stream)); }
  // applyNeumannToRhs with a scalar value, :621-640,711-730
  inline void applyNeumannToRhs(DoFLinearSystem& ls, const FaceGroup& g, Real value, b2f_stream stream = nullptr)
  {
    b2f_neumann bc{};
    bc.nb_face = g.n; bc.face_node = g.face_node; bc.face_outside = g.outside; bc.has_value = 1; bc.value = value;
    check(b2f_rhs_neumann(ls.handle(), &bc, stream));
  }
  // applyNeumannToRhs with a flux vector (valueX, valueY[, valueZ]) . outward normal, :641-677,731-768
  inline void applyNeumannToRhs(DoFLinearSystem& ls, const FaceGroup& g, Real vx, Real vy, Real vz, b2f_stream stream = nullptr)
  {
    b2f_neumann bc{};
    bc.nb_face = g.n; bc.face_node = g.face_node; bc.face_outside = g.outside; bc.has_value = 0;
    bc.value_x = vx; bc.value_y = vy; bc.value_z = vz;
    check(b2f_rhs_neumann(ls.handle(), &bc, stream));
  }
  // Transient heat module (modules/heat/FemModule.cc; CPU-only in the reference).  Source (T_old / dt) * |cell| / Nv, :480-495:
  inline void applyNodalSourceToRhs(DoFLinearSystem& ls, const Real* node_field, Real scale, const uint8_t* node_skip = nullptr, b2f_stream stream = nullptr)
  {
    check(b2f_rhs_nodal_source(ls.handle(), node_field, scale, node_skip, stream));
  }
  // Convection boundary: matrix term h * face mass (_assembleBilinearOperatorEDGE2, :628-672,813-840) and rhs h * Text * |face| / Nvf (:517-537)
  inline void applyConvectionToMatrix(DoFLinearSystem& ls, const FaceGroup& g, Real h, b2f_stream stream = nullptr)
  {
    check(b2f_add_face_mass(ls.handle(), g.n, g.face_node, h, stream));
  }
  inline void applyConvectionToRhs(DoFLinearSystem& ls, const FaceGroup& g, Real h, Real Text, b2f_stream stream = nullptr)
  {
    applyNeumannToRhs(ls, g, h * Text, stream);
  }
} // namespace BoundaryConditionsND
namespace BoundaryConditions2D = BoundaryConditionsND;
namespace BoundaryConditions3D = BoundaryConditionsND;

}} // namespace ArcaneFemB200::Gpu
