// arcanefem_b200/arcane/BSRFormat.h — drop-in for femutils/BSRFormat.h with the reference's OWN signatures, so that the call
// sites of the modules compile unchanged (modules/poisson/FemModule.h:71,87; FemModule.cc:24-40,76-77,108-109,255-286;
// modules/testlab/FemModule.cc:296-312):
//   BSRFormat<NB_DOF>(ITraceMng*, RunQueue&, const FemDoFsOnNodes&)            femutils/BSRFormat.h:358-363
//   initialize(IMesh*, bool does_linear_system_use_csr, bool use_atomic_free)   :389-409
//   computeSparsity()                                                           :775-781
//   assembleBilinear(Function compute_element_matrix)                           :1105-1111  (any device callable CellLocalId -> FixedMatrix)
//   toLinearSystem(DoFLinearSystem&)                                            :414-430 (CSR view) and :261-280 (per-block values: same device arrays)
//   matrix() / resetMatrixValues() / dumpMatrix(std::string)                    :1116-1127
//   BSRMatrix: values() columns() rowIndex() nbNzPerRow() nbNz() nbCol() nbRow() orderValuePerBlock() setValue getValue   :176-189,312-320
// Needs nvcc (the element callback and two small flattening kernels are device code) and Arcane (real, or the model under
// tests/cpp/arcane_stub).  The reference's callback reads the mesh through its own connectivity view in the cell's own vertex
// order, so it runs on the cell-wise kernel (the reference's own algorithm); passing ArcaneFemB200::Operator::Laplace() /
// Elasticity() / Heat() instead of the lambda selects the library's node-gather kernels - the one-line change INTEGRATION.md shows.
#pragma once
#include <string>
#include <arcane/IMesh.h>
#include <arcane/ItemGroup.h>
#include <arcane/accelerator/core/IAcceleratorMng.h>
#include <arcane/accelerator/VariableViews.h>
#include "../BSRFormat.h"
#include "FemUtils.h"
#include "DoFLinearSystem.h"

namespace Arcane::FemUtils {

class FemDoFsOnNodes; // femutils/FemDoFsOnNodes.h (dof = node * NB_DOF + i, FemDoFsOnNodes.cc:79-109)

namespace b200detail {
  inline void cuda_check(cudaError_t e, const char* what)
  {
    if (e != cudaSuccess)
      throw std::runtime_error(std::string("CUDA error in ") + what + ": " + cudaGetErrorString(e));
  }
  // cellNode() connectivity of Arcane (indexed, with an indirection) -> the flat Int32 array the library reads
  template <class ConnectivityView>
  __global__ void k_flatten_cells(ConnectivityView cn_cv, Int32 nb_cell, Int32 nv, Int32* out)
  {
    const Int32 c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= nb_cell)
      return;
    for (Int32 a = 0; a < nv; ++a)
      out[(Int64)c * nv + a] = cn_cv.nodeId(CellLocalId(c), a).localId();
  }
  template <class InfoView>
  __global__ void k_own_flags(InfoView infos, Int32 nb_node, unsigned char* out)
  {
    const Int32 n = blockIdx.x * blockDim.x + threadIdx.x;
    if (n < nb_node)
      out[n] = infos.isOwn(NodeLocalId(n)) ? 1 : 0;
  }
  struct DeviceArray // owning cudaMalloc buffer
  {
    void* p = nullptr;
    DeviceArray() = default;
    DeviceArray(const DeviceArray&) = delete;
    DeviceArray& operator=(const DeviceArray&) = delete;
    ~DeviceArray() { if (p) cudaFree(p); }
    void resize(size_t bytes)
    {
      if (p)
        cudaFree(p);
      p = nullptr;
      cuda_check(cudaMalloc(&p, bytes ? bytes : 16), "cudaMalloc");
    }
  };

  // The reference's callback `FixedMatrix<N, N> f(CellLocalId)` as an operator of the assembly kernels: OWN_ORDER (the callback
  // gathers through its own views, in the cell's own vertex order) -> cell-wise kernel.
  template <int NV_, int BS_, class F>
  struct CellCallbackOp
  {
    static constexpr int NV = NV_, BS = BS_;
    static constexpr int FAN_WARPS_PER_SM = 8;
    static constexpr bool ROW_SUM_ZERO = false, NEEDS_CELL = true, LINEAR_POST = false, NO_FAN = true, OWN_ORDER = true;
    using Geom = FixedMatrix<NV_ * BS_, NV_ * BS_>;
    F f;
    __device__ __forceinline__ Geom prepare(const double (&)[NV_][3], long long cell) const { return f(CellLocalId((Int32)cell)); }
    __device__ __forceinline__ double cell_value(long long) const { return 0.0; }
    __device__ __forceinline__ void block(const Geom& G, int a, int b, double (&out)[BS_ * BS_]) const
    {
#pragma unroll
      for (int i = 0; i < BS_; ++i)
#pragma unroll
        for (int j = 0; j < BS_; ++j)
          out[i * BS_ + j] = G(BS_ * a + i, BS_ * b + j);
    }
  };
} // namespace b200detail

// NumArray<T, MDDim1>& of the reference's accessors (femutils/BSRFormat.h:317-320): a device array
template <class T>
struct B200NumArray
{
  T* m_p = nullptr;
  Int64 m_n = 0;
  B200Span<T> to1DSpan() const { return B200Span<T>{ m_p, m_n }; }
  B200Span<T> to1DSmallSpan() const { return to1DSpan(); }
  Int64 extent0() const { return m_n; }
  Int64 dim1Size() const { return m_n; }
};

template <int BLOCK_SIZE>
class BSRMatrix // femutils/BSRFormat.h:104-336: views on the arrays the handle owns
{
 public:
  explicit BSRMatrix(b2f_handle* h) : m_h(h) {}
  bool orderValuePerBlock() const { return view().layout == B2F_LAYOUT_PER_BLOCK; } // :315
  Int32 nbNz() const { return (Int32)view().nb_nz; } // :312-314
  Int32 nbCol() const { return (Int32)view().nb_col; }
  Int32 nbRow() const { return (Int32)view().nb_row; }
  B200NumArray<Real> values() const { auto v = view(); return { v.values, v.nb_nz }; } // :317-320
  B200NumArray<const Int32> columns() const { auto v = view(); return { v.columns, v.nb_col }; }
  B200NumArray<const Int32> rowIndex() const { auto v = view(); return { v.row_index, v.nb_row }; } // nb_row entries, :138
  B200NumArray<const Int32> nbNzPerRow() const { auto v = view(); return { v.nb_nz_per_row, v.nb_row }; }
  // :176-189; findValueIndex throws std::runtime_error("BSRMatrix(findValueIndex): Value not found") like :170
  Real getValue(DoFLocalId row, DoFLocalId col) const { Real x = 0; ArcaneFemB200::check(b2f_get_value(*m_h, row.localId(), col.localId(), &x)); return x; }
  void setValue(DoFLocalId row, DoFLocalId col, Real value) { ArcaneFemB200::check(b2f_set_value(*m_h, row.localId(), col.localId(), value)); }
  void dump(const std::string& filename) const { ArcaneFemB200::check(b2f_dump(*m_h, filename.c_str())); } // :285-307
  // :194-256
  CSRFormatView toCsr(b2f_stream stream) const
  {
    b2f_csr_view c{};
    ArcaneFemB200::check(b2f_to_csr(*m_h, &c, stream));
    return CSRFormatView({ c.rows, c.nb_row }, { c.rows_nb_column, c.nb_row }, { c.columns, c.nb_value }, { c.values, c.nb_value });
  }

 private:
  b2f_bsr_view view() const { b2f_bsr_view v{}; ArcaneFemB200::check(b2f_get_bsr(*m_h, &v)); return v; }
  b2f_handle* m_h;
};

template <int NB_DOF>
class BSRFormat : public TraceAccessor // femutils/BSRFormat.h:353-1140
{
 public:
  BSRFormat(ITraceMng* tm, Accelerator::RunQueue& queue, const FemDoFsOnNodes& dofs_on_nodes) // :358-363
  : TraceAccessor(tm), m_queue(queue), m_dofs_on_nodes(dofs_on_nodes), m_bsr_matrix(&m_h) {}
  BSRFormat(const BSRFormat&) = delete;
  BSRFormat& operator=(const BSRFormat&) = delete;
  ~BSRFormat() { if (m_h) b2f_destroy(m_h); }

  // :389-409.  The mesh views are flattened on the device once (two small kernels); coordinates are read in place from
  // mesh->nodesCoordinates() (Real3 = three contiguous doubles) at every assembly.
  void initialize(IMesh* mesh, bool does_linear_system_use_csr, bool use_atomic_free = false)
  {
    ARCANE_CHECK_POINTER(mesh);
    if (mesh->dimension() != 2 && mesh->dimension() != 3)
      ARCANE_THROW(NotImplementedException, "BSRFormat(initialize): Only supports 2D and 3D");
    m_mesh = mesh;
    m_use_csr_in_linear_system = does_linear_system_use_csr;
    m_use_atomic_free = use_atomic_free;
    m_stream = m_queue.platformStream();
    cudaStream_t s = (cudaStream_t)m_stream;
    const Int32 nv = mesh->dimension() + 1, nb_cell = mesh->nbCell(), nb_node = mesh->nbNode();
    UnstructuredMeshConnectivityView connectivity_view(mesh);
    auto cn_cv = connectivity_view.cellNode();
    NodeInfoListView nodes_infos(mesh->nodeFamily());
    m_cell_node.resize(sizeof(Int32) * (size_t)nb_cell * nv);
    m_node_is_own.resize((size_t)nb_node);
    if (nb_cell > 0)
      b200detail::k_flatten_cells<<<(nb_cell + 255) / 256, 256, 0, s>>>(cn_cv, nb_cell, nv, static_cast<Int32*>(m_cell_node.p));
    b200detail::k_own_flags<<<(nb_node + 255) / 256, 256, 0, s>>>(nodes_infos, nb_node, static_cast<unsigned char*>(m_node_is_own.p));
    b200detail::cuda_check(cudaGetLastError(), "BSRFormat(initialize) flattening kernels");
    b2f_mesh_view mv{};
    mv.dim = mesh->dimension();
    mv.nodes_per_cell = nv;
    mv.nb_cell = nb_cell;
    mv.nb_node = nb_node;
    mv.cell_node = static_cast<const Int32*>(m_cell_node.p);
    mv.coords = reinterpret_cast<const double*>(mesh->nodesCoordinates().asArray().data());
    mv.node_is_own = static_cast<const uint8_t*>(m_node_is_own.p); // rows of own nodes only, like :815
    b2f_options opt{};
    opt.block_size = NB_DOF;
    opt.use_csr = does_linear_system_use_csr ? 1 : 0; // order_values_per_block = !does_linear_system_use_csr, :405
    opt.assembly_variant = use_atomic_free ? B2F_ASM_NODE_GATHER : B2F_ASM_DEFAULT;
    if (m_h) {
      b2f_destroy(m_h);
      m_h = nullptr;
    }
    ArcaneFemB200::check(b2f_create(&mv, &opt, &m_h));
  }

  void computeSparsity() // :775-781
  {
    info() << "BSRFormat(computeSparsity): Compute sparsity of BSR matrix on the device (libb200fem)";
    ArcaneFemB200::check(b2f_compute_sparsity(m_h, m_stream));
  }

  // :1105-1111 with the reference's callback: any device callable `FixedMatrix<NB_DOF*NV, NB_DOF*NV>(CellLocalId)`.
  template <class Function>
  void assembleBilinear(Function compute_element_matrix)
  {
    using Matrix = decltype(compute_element_matrix(CellLocalId(0)));
    constexpr int N = Matrix::rows;
    static_assert(Matrix::rows == Matrix::cols && N % NB_DOF == 0, "the element callback returns a square FixedMatrix of NB_DOF * (nodes per cell) rows");
    constexpr int NV = N / NB_DOF;
    static_assert(NV == 3 || NV == 4, "P1 triangles and tetrahedra");
    if (NV != m_mesh->dimension() + 1)
      ARCANE_THROW(ArgumentException, "BSRFormat(assembleBilinear): the element matrix does not fit the mesh dimension");
    ArcaneFemB200::check(b2f::assemble_with(m_h, b200detail::CellCallbackOp<NV, NB_DOF, Function>{ compute_element_matrix }, false, (cudaStream_t)m_stream));
  }
  // The library's own operators (node-gather fan kernel): Operator::Laplace() / Elasticity(E, nu) / Heat(lambda, dt, cell_lambda)
  void assembleBilinear(const ArcaneFemB200::Operator& op) { ArcaneFemB200::check(b2f_assemble(m_h, &op.desc, m_stream)); }

  // :414-430.  CSR consumers (Hypre) get the CSR view; other linear systems (the per-block / Aleph branch, BSRMatrix::toLinearSystem
  // :261-280, which adds every entry through matrixAddValue on the host) are attached to the same device-resident values instead.
  void toLinearSystem(DoFLinearSystem& linear_system)
  {
    info() << "BSRFormat(toLinearSystem): Translate matrix to linear system";
    linear_system.attachHandle(&m_h, &m_stream);
    if (m_use_csr_in_linear_system) {
      if (!linear_system.hasSetCSRValues())
        ARCANE_THROW(ArgumentException, "BSRFormat(toLinearSystem): Linear system was set to use CSR but is incompatible");
      linear_system.setCSRValues(m_bsr_matrix.toCsr(m_stream));
    }
  }
  BSRMatrix<NB_DOF>& matrix() { return m_bsr_matrix; } // :1116
  void resetMatrixValues() { ArcaneFemB200::check(b2f_reset_values(m_h, m_stream)); } // :1120-1123
  void dumpMatrix(const std::string& filename) { m_bsr_matrix.dump(filename); } // :1127
  b2f_handle handle() const { return m_h; }

 private:
  b2f_handle m_h = nullptr;
  b2f_stream m_stream = nullptr;
  IMesh* m_mesh = nullptr;
  Accelerator::RunQueue& m_queue;
  const FemDoFsOnNodes& m_dofs_on_nodes;
  bool m_use_csr_in_linear_system = false, m_use_atomic_free = false;
  b200detail::DeviceArray m_cell_node, m_node_is_own;
  BSRMatrix<NB_DOF> m_bsr_matrix;
};

} // namespace Arcane::FemUtils
