// arcanefem_b200/arcane/ArcaneFemFunctionsGpu.h — drop-in for femutils/ArcaneFemFunctionsGpu.h with the reference's argument
// lists (file:line below), on top of the C ABI:
//   Gpu::MeshOperation::computeAreaTria3 / computeVolumeTetra4(cell_lid, cn_cv, in_node_coord)          :75-83,110-122
//   Gpu::FeOperation2D::computeGradient{X,Y}Tria3, Gpu::FeOperation3D::computeGradient{X,Y,Z}Tetra4      :218-252,280-392
//   Gpu::BoundaryConditions::applyDirichletViaPenalty / ...Vectorial / ...ViaRowEliminationVectorial / applyPointDirichlet...   :501-586
//   Gpu::BoundaryConditions2D / 3D::applyConstantSourceToRhs, applyNeumannToRhs                          :605-677,695-768
// The element helpers gather the cell's vertices through the caller's views like the reference's do and evaluate the
// library's formulas (arcanefem_b200/FemFunctionsGpu.cuh).  The boundary functions turn Arcane groups into device id lists
// (one small copy per call) and call b2f_rhs_* / b2f_dirichlet_flag_*; the state they write lives in the handle attached to
// the DoFLinearSystem.
#pragma once
#include <cstdlib>
#include "BSRFormat.h"
#include "FemDoFsOnNodes.h" // femutils/FemDoFsOnNodes.h (the reference's; tests: tests/cpp/arcane_stub)
#include "IArcaneFemBC.h" // femutils/IArcaneFemBC.h

namespace Arcane::FemUtils::Gpu {

namespace b200detail {
  template <int NV>
  ARCCORE_HOST_DEVICE inline void gather(CellLocalId cell_lid, const IndexedCellNodeConnectivityView& cn_cv, const Accelerator::VariableNodeReal3InView& in_node_coord,
                                         Real (&X)[NV][3])
  {
    for (Int32 a = 0; a < NV; ++a) {
      const Real3 v = in_node_coord[cn_cv.nodeId(cell_lid, a)];
      X[a][0] = v.x;
      X[a][1] = v.y;
      X[a][2] = v.z;
    }
  }
  // group ids -> device list (scratch owned by the call; the stream orders the free after the kernels that read it)
  struct DeviceIds
  {
    Int32* p = nullptr;
    Int64 n = 0;
    cudaStream_t s;
    DeviceIds(const Int32* host_or_managed, Int64 count, cudaStream_t stream) : n(count), s(stream)
    {
      if (n == 0)
        return;
      if (cudaMallocAsync(reinterpret_cast<void**>(&p), sizeof(Int32) * (size_t)n, s) != cudaSuccess ||
          cudaMemcpyAsync(p, host_or_managed, sizeof(Int32) * (size_t)n, cudaMemcpyDefault, s) != cudaSuccess)
        throw std::runtime_error("BoundaryConditions: cannot stage the group ids on the device");
    }
    DeviceIds(const DeviceIds&) = delete;
    ~DeviceIds() { if (p) cudaFreeAsync(p, s); }
  };
  __global__ void k_gather_faces(const Int32* face_ids, Int64 n, IndexedFaceNodeConnectivityView fn_cv, FaceInfoListView infos, Int32 nvf, Int32* face_node, unsigned char* outside)
  {
    const Int64 i = (Int64)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n)
      return;
    const FaceLocalId f(face_ids[i]);
    for (Int32 a = 0; a < nvf; ++a)
      face_node[i * nvf + a] = fn_cv.nodeId(f, a).localId();
    outside[i] = infos.isSubDomainBoundaryOutside(f) ? 1 : 0;
  }
  inline cudaStream_t stream_of(Accelerator::RunQueue* queue) { return (cudaStream_t)queue->platformStream(); }

  inline void flag_penalty(DoFLinearSystem& ls, Accelerator::RunQueue* queue, NodeGroup& node_group, Int32 dof_index, Real value, Real penalty)
  {
    auto ids = node_group.view().localIds();
    DeviceIds d(ids.data(), ids.size(), stream_of(queue));
    ArcaneFemB200::check(b2f_dirichlet_flag_penalty(ls.handle(), d.p, d.n, dof_index, value, penalty, stream_of(queue)));
  }
  inline void flag_row_elimination(DoFLinearSystem& ls, Accelerator::RunQueue* queue, NodeGroup& node_group, Int32 dof_index, Real value, Int32 mode)
  {
    auto ids = node_group.view().localIds();
    DeviceIds d(ids.data(), ids.size(), stream_of(queue));
    ArcaneFemB200::check(b2f_dirichlet_flag_elimination(ls.handle(), d.p, d.n, dof_index, value, mode, stream_of(queue)));
  }
  inline void neumann(BC::INeumannBoundaryCondition* bs, VariableDoFReal& rhs, IMesh* mesh, Accelerator::RunQueue* queue)
  {
    ARCANE_CHECK_POINTER(bs);
    ARCANE_CHECK_POINTER(queue);
    ARCANE_CHECK_POINTER(mesh);
    cudaStream_t s = stream_of(queue);
    FaceGroup group = bs->getSurface();
    auto ids = group.view().localIds();
    const Int64 n = ids.size();
    const Int32 nvf = mesh->dimension();
    DeviceIds d(ids.data(), n, s);
    Int32* face_node = nullptr;
    unsigned char* outside = nullptr;
    if (n > 0) {
      if (cudaMallocAsync(reinterpret_cast<void**>(&face_node), sizeof(Int32) * (size_t)(n * nvf), s) != cudaSuccess ||
          cudaMallocAsync(reinterpret_cast<void**>(&outside), (size_t)n, s) != cudaSuccess)
        throw std::runtime_error("applyNeumannToRhs: cannot allocate the face list");
      UnstructuredMeshConnectivityView connectivity_view;
      connectivity_view.setMesh(mesh);
      k_gather_faces<<<(unsigned)((n + 255) / 256), 256, 0, s>>>(d.p, n, connectivity_view.faceNode(), FaceInfoListView(mesh->faceFamily()), nvf, face_node, outside);
    }
    b2f_neumann bc{};
    bc.nb_face = n;
    bc.face_node = face_node;
    bc.face_outside = outside;
    if (bs->hasValue()) { // :627-648,717-738
      bc.has_value = 1;
      bc.value = bs->getValue();
    }
    else { // :649-676,739-767: flux vector . outward normal
      bc.has_value = 0;
      bc.value_x = bs->hasValueX() ? bs->getValueX() : 0.0;
      bc.value_y = bs->hasValueY() ? bs->getValueY() : 0.0;
      bc.value_z = (nvf == 3 && bs->hasValueZ()) ? bs->getValueZ() : 0.0;
    }
    const int st = b2f_rhs_neumann(rhs.handle(), &bc, s);
    if (face_node)
      cudaFreeAsync(face_node, s);
    if (outside)
      cudaFreeAsync(outside, s);
    ArcaneFemB200::check(st);
  }
} // namespace b200detail

namespace MeshOperation {
  ARCCORE_HOST_DEVICE inline Real computeAreaTria3(CellLocalId cell_lid, const IndexedCellNodeConnectivityView& cn_cv, const Accelerator::VariableNodeReal3InView& in_node_coord)
  {
    Real X[3][3];
    b200detail::gather<3>(cell_lid, cn_cv, in_node_coord, X);
    return ArcaneFemB200::Gpu::MeshOperation::computeAreaTria3(X);
  }
  ARCCORE_HOST_DEVICE inline Real computeVolumeTetra4(CellLocalId cell_lid, const IndexedCellNodeConnectivityView& cn_cv, const Accelerator::VariableNodeReal3InView& in_node_coord)
  {
    Real X[4][3];
    b200detail::gather<4>(cell_lid, cn_cv, in_node_coord, X);
    return ArcaneFemB200::Gpu::MeshOperation::computeVolumeTetra4(X);
  }
} // namespace MeshOperation

namespace FeOperation2D {
  ARCCORE_HOST_DEVICE inline Real3 computeGradientXTria3(CellLocalId cell_lid, const IndexedCellNodeConnectivityView& cn_cv, const Accelerator::VariableNodeReal3InView& in_node_coord)
  {
    Real X[3][3];
    b200detail::gather<3>(cell_lid, cn_cv, in_node_coord, X);
    const ArcaneFemB200::Real3 g = ArcaneFemB200::Gpu::FeOperation2D::computeGradientXTria3(X);
    return Real3{ g.x, g.y, g.z };
  }
  ARCCORE_HOST_DEVICE inline Real3 computeGradientYTria3(CellLocalId cell_lid, const IndexedCellNodeConnectivityView& cn_cv, const Accelerator::VariableNodeReal3InView& in_node_coord)
  {
    Real X[3][3];
    b200detail::gather<3>(cell_lid, cn_cv, in_node_coord, X);
    const ArcaneFemB200::Real3 g = ArcaneFemB200::Gpu::FeOperation2D::computeGradientYTria3(X);
    return Real3{ g.x, g.y, g.z };
  }
} // namespace FeOperation2D

namespace FeOperation3D {
  ARCCORE_HOST_DEVICE inline Real4 computeGradientXTetra4(CellLocalId cell_lid, const IndexedCellNodeConnectivityView& cn_cv, const Accelerator::VariableNodeReal3InView& in_node_coord)
  {
    Real X[4][3];
    b200detail::gather<4>(cell_lid, cn_cv, in_node_coord, X);
    return ArcaneFemB200::Gpu::FeOperation3D::computeGradientXTetra4(X);
  }
  ARCCORE_HOST_DEVICE inline Real4 computeGradientYTetra4(CellLocalId cell_lid, const IndexedCellNodeConnectivityView& cn_cv, const Accelerator::VariableNodeReal3InView& in_node_coord)
  {
    Real X[4][3];
    b200detail::gather<4>(cell_lid, cn_cv, in_node_coord, X);
    return ArcaneFemB200::Gpu::FeOperation3D::computeGradientYTetra4(X);
  }
  ARCCORE_HOST_DEVICE inline Real4 computeGradientZTetra4(CellLocalId cell_lid, const IndexedCellNodeConnectivityView& cn_cv, const Accelerator::VariableNodeReal3InView& in_node_coord)
  {
    Real X[4][3];
    b200detail::gather<4>(cell_lid, cn_cv, in_node_coord, X);
    return ArcaneFemB200::Gpu::FeOperation3D::computeGradientZTetra4(X);
  }
} // namespace FeOperation3D

namespace BoundaryConditions {
  // :501-512
  inline void applyDirichletViaPenalty(BC::IDirichletBoundaryCondition* bs, const FemDoFsOnNodes&, DoFLinearSystem& linear_system, IMesh* mesh, Accelerator::RunQueue* queue)
  {
    ARCANE_CHECK_POINTER(bs);
    ARCANE_CHECK_POINTER(queue);
    ARCANE_CHECK_POINTER(mesh);
    FaceGroup face_group = bs->getSurface();
    NodeGroup node_group = face_group.nodeGroup();
    b200detail::flag_penalty(linear_system, queue, node_group, 0, bs->getValue(), bs->getPenalty());
  }
  // :513-528: one call per component whose value string is not "NULL"
  inline void applyDirichletViaPenaltyVectorial(const FemDoFsOnNodes&, DoFLinearSystem& linear_system, IMesh*, Accelerator::RunQueue* queue, FaceGroup face_group, Real penalty,
                                                const UniqueArray<String> u_dirichlet_str)
  {
    NodeGroup node_group = face_group.nodeGroup();
    for (Int32 dof_index = 0; dof_index < u_dirichlet_str.size(); ++dof_index)
      if (u_dirichlet_str[dof_index] != "NULL")
        b200detail::flag_penalty(linear_system, queue, node_group, dof_index, std::stod(u_dirichlet_str[dof_index].localstr()), penalty);
  }
  // :529-545
  inline void applyDirichletViaRowEliminationVectorial(const FemDoFsOnNodes&, DoFLinearSystem& linear_system, IMesh*, Accelerator::RunQueue* queue, FaceGroup face_group,
                                                       const UniqueArray<String> u_dirichlet_str)
  {
    NodeGroup node_group = face_group.nodeGroup();
    for (Int32 dof_index = 0; dof_index < u_dirichlet_str.size(); ++dof_index)
      if (u_dirichlet_str[dof_index] != "NULL")
        b200detail::flag_row_elimination(linear_system, queue, node_group, dof_index, std::stod(u_dirichlet_str[dof_index].localstr()), B2F_ELIMINATE_ROW);
  }
  // :546-559
  inline void applyPointDirichletViaPenalty(BC::IDirichletPointCondition* bs, const FemDoFsOnNodes&, DoFLinearSystem& linear_system, IMesh* mesh, Accelerator::RunQueue* queue)
  {
    ARCANE_CHECK_POINTER(bs);
    ARCANE_CHECK_POINTER(queue);
    ARCANE_CHECK_POINTER(mesh);
    NodeGroup node_group = bs->getNode();
    b200detail::flag_penalty(linear_system, queue, node_group, 0, bs->getValue(), bs->getPenalty());
  }
  // :560-586
  inline void applyPointDirichletViaPenaltyVectorial(const FemDoFsOnNodes&, DoFLinearSystem& linear_system, IMesh*, Accelerator::RunQueue* queue, NodeGroup node_group, Real penalty,
                                                     const UniqueArray<String> u_dirichlet_str)
  {
    for (Int32 dof_index = 0; dof_index < u_dirichlet_str.size(); ++dof_index)
      if (u_dirichlet_str[dof_index] != "NULL")
        b200detail::flag_penalty(linear_system, queue, node_group, dof_index, std::stod(u_dirichlet_str[dof_index].localstr()), penalty);
  }
  inline void applyPointDirichletViaRowEliminationVectorial(const FemDoFsOnNodes&, DoFLinearSystem& linear_system, IMesh*, Accelerator::RunQueue* queue, NodeGroup node_group,
                                                            const UniqueArray<String> u_dirichlet_str)
  {
    for (Int32 dof_index = 0; dof_index < u_dirichlet_str.size(); ++dof_index)
      if (u_dirichlet_str[dof_index] != "NULL")
        b200detail::flag_row_elimination(linear_system, queue, node_group, dof_index, std::stod(u_dirichlet_str[dof_index].localstr()), B2F_ELIMINATE_ROW);
  }
} // namespace BoundaryConditions

// :588-677 and :679-768: the dimension is a property of the mesh behind the handle, so both classes share their bodies
class BoundaryConditions2D
{
 public:
  static inline void applyConstantSourceToRhs(Real qdot, const FemDoFsOnNodes&, VariableNodeReal3, VariableDoFReal& rhs_variable_na, IMesh* mesh, Accelerator::RunQueue* queue)
  {
    ARCANE_CHECK_POINTER(queue);
    ARCANE_CHECK_POINTER(mesh);
    ArcaneFemB200::check(b2f_rhs_constant_source(rhs_variable_na.handle(), qdot, b200detail::stream_of(queue)));
  }
  static inline void applyNeumannToRhs(BC::INeumannBoundaryCondition* bs, const FemDoFsOnNodes&, VariableNodeReal3, VariableDoFReal& rhs_variable_na, IMesh* mesh,
                                       Accelerator::RunQueue* queue)
  {
    b200detail::neumann(bs, rhs_variable_na, mesh, queue);
  }
};
class BoundaryConditions3D : public BoundaryConditions2D {};

} // namespace Arcane::FemUtils::Gpu
