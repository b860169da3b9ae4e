// arcanefem_b200/BSRFormat.h — header-only C++ mirror of the reference's femutils assembly API over the C ABI (b2f.h).
//
// Same class names, method names, argument meaning and error behaviour as
//   femutils/BSRFormat.h            BSRMatrix<BS> (:104-336), BSRFormat<NB_DOF> (:353-1140)
//   femutils/DoFLinearSystem.h      CSRFormatView (:42-76), DoFLinearSystem (:126-286; the part the assembly path touches)
//   femutils/ArcaneFemFunctionsGpu.h  FemUtils::Gpu::BoundaryConditions* (:434-768)
// so a module's FemModule code reads unchanged.  Differences forced by the C ABI:
//   * Arcane handles (ITraceMng*, RunQueue&, IMesh*, node/face groups) become a b2f_mesh_view, a stream and device id
//     lists; the Arcane-typed signatures live in arcanefem_b200/arcane/*.h (they need Arcane or the model of tests/cpp/arcane_stub);
//   * the element lambda of assembleBilinear becomes an Operator descriptor (Laplace / Elasticity / Heat);
//   * NumArray accessors return non-owning device spans.
// Errors: the reference throws ArgumentException / NotImplementedException / std::runtime_error; here every non-zero
// C-ABI status is rethrown as std::invalid_argument (bad arguments), std::logic_error (not implemented / call order) or
// std::runtime_error (everything else) carrying b2f_last_error_string().
#pragma once
#include <cstdint>
#include <stdexcept>
#include <string>
#include "../b2f.h"
#ifdef __CUDACC__
#include "../b2f/user_operator.cuh" // element callbacks written by the caller (needs nvcc: the callback is device code)
#endif

namespace ArcaneFemB200 {

using Real = double;
using Int32 = int32_t;
using Int64 = int64_t;

inline void check(int status)
{
  if (status == B2F_OK)
    return;
  const std::string msg = std::string(b2f_status_string(status)) + ": " + b2f_last_error_string();
  if (status == B2F_ERR_INVALID_ARGUMENT)
    throw std::invalid_argument(msg);
  if (status == B2F_ERR_NOT_IMPLEMENTED || status == B2F_ERR_STATE)
    throw std::logic_error(msg);
  throw std::runtime_error(msg);
}

// Non-owning view of a device array (stands in for NumArray<T, MDDim1>& / Span<T>).
template <class T>
struct DeviceSpan
{
  T* ptr = nullptr;
  Int64 n = 0;
  T* data() const { return ptr; }
  Int64 size() const { return n; }
  Int64 extent0() const { return n; }
};

// femutils/DoFLinearSystem.h:42-76
class CSRFormatView
{
 public:
  CSRFormatView() = default;
  CSRFormatView(DeviceSpan<const Int32> rows, DeviceSpan<const Int32> rows_nb_column, DeviceSpan<const Int32> columns, DeviceSpan<Real> values)
  : m_rows(rows), m_rows_nb_column(rows_nb_column), m_columns(columns), m_values(values) {}
  DeviceSpan<const Int32> rows() const { return m_rows; }
  DeviceSpan<const Int32> rowsNbColumn() const { return m_rows_nb_column; }
  DeviceSpan<const Int32> columns() const { return m_columns; }
  DeviceSpan<Real> values() const { return m_values; }
  Int64 nbRow() const { return m_rows.n; }
  Int64 nbColumn() const { return m_columns.n; }
  Int64 nbValue() const { return m_values.n; }

 private:
  DeviceSpan<const Int32> m_rows, m_rows_nb_column, m_columns;
  DeviceSpan<Real> m_values;
};

// Replaces the element lambda `FixedMatrix<N,N> compute_element_matrix(CellLocalId)` (BSRFormat.h:786-787).
struct Operator
{
  b2f_operator desc{};
  // modules/poisson/FemModule.h:139-147 (Tria3), :177-186 (Tetra4)
  static Operator Laplace() { Operator o; o.desc.kind = B2F_OP_LAPLACE; return o; }
  // modules/elasticity/FemModule.h:112-152 with lambda, mu2 of FemModule.cc:133-134
  static Operator Elasticity(Real E, Real nu)
  {
    Operator o;
    o.desc.kind = B2F_OP_ELASTICITY;
    o.desc.mu2 = (E / (2 * (1 + nu))) * 2;
    o.desc.lambda = E * nu / ((1 + nu) * (1 - 2 * nu));
    return o;
  }
  // modules/heat/FemModule.cc:678-770
  static Operator Heat(Real lambda, Real dt, const Real* cell_lambda_device = nullptr)
  {
    Operator o;
    o.desc.kind = B2F_OP_HEAT;
    o.desc.lambda = lambda;
    o.desc.inv_dt = 1.0 / dt;
    o.desc.cell_lambda = cell_lambda_device;
    return o;
  }
};

class DoFLinearSystem;
template <int NB_DOF> class BSRFormat;

// femutils/BSRFormat.h:104-336 — views on the arrays owned by the handle of its BSRFormat.
template <int BLOCK_SIZE>
class BSRMatrix
{
 public:
  explicit BSRMatrix(b2f_handle* h) : m_h(h) {}
  bool orderValuePerBlock() const { return view().layout == B2F_LAYOUT_PER_BLOCK; }
  Int32 nbNz() const { return (Int32)view().nb_nz; }
  Int32 nbCol() const { return (Int32)view().nb_col; }
  Int32 nbRow() const { return (Int32)view().nb_row; }
  DeviceSpan<Real> values() const { auto v = view(); return { v.values, v.nb_nz }; }
  DeviceSpan<const Int32> columns() const { auto v = view(); return { v.columns, v.nb_col }; }
  DeviceSpan<const Int32> rowIndex() const { auto v = view(); return { v.row_index, v.nb_row }; } // nb_row entries, BSRFormat.h:138
  DeviceSpan<const Int32> nbNzPerRow() const { auto v = view(); return { v.nb_nz_per_row, v.nb_row }; }
  // BSRFormat.h:176-189 (DoFLocalId row, DoFLocalId col); throws std::runtime_error("... Value not found") like :170
  Real getValue(Int64 row, Int64 col) const { Real x = 0; check(b2f_get_value(*m_h, row, col, &x)); return x; }
  void setValue(Int64 row, Int64 col, Real value) { check(b2f_set_value(*m_h, row, col, value)); }
  // BSRFormat.h:194-256
  CSRFormatView toCsr(b2f_stream stream = nullptr) const
  {
    b2f_csr_view c{};
    check(b2f_to_csr(*m_h, &c, stream));
    return CSRFormatView({ c.rows, c.nb_row }, { c.rows_nb_column, c.nb_row }, { c.columns, c.nb_value }, { c.values, c.nb_value });
  }
  void dump(const std::string& filename) const { check(b2f_dump(*m_h, filename.c_str())); } // BSRFormat.h:285-307

 private:
  b2f_bsr_view view() const { b2f_bsr_view v{}; check(b2f_get_bsr(*m_h, &v)); return v; }
  b2f_handle* m_h;
};

// Device-resident part of femutils/DoFLinearSystem.h:126-286 used by the assembly path (no solver behind it).
class DoFLinearSystem
{
 public:
  bool hasSetCSRValues() const { return true; } // DoFLinearSystem.cc:506-511
  void setCSRValues(const CSRFormatView& v) { m_csr = v; } // DoFLinearSystem.cc:451-456: the view is kept, not copied
  const CSRFormatView& getCSRValues() const { return m_csr; }
  DeviceSpan<Real> rhsVariable() const { auto s = sys(); return { s.rhs, s.nb_dof }; }
  DeviceSpan<uint8_t> getForcedInfo() const { auto s = sys(); return { s.forced_info, s.nb_dof }; }
  DeviceSpan<Real> getForcedValue() const { auto s = sys(); return { s.forced_value, s.nb_dof }; }
  DeviceSpan<uint8_t> getEliminationInfo() const { auto s = sys(); return { s.elimination_info, s.nb_dof }; }
  DeviceSpan<Real> getEliminationValue() const { auto s = sys(); return { s.elimination_value, s.nb_dof }; }
  void clearValues(b2f_stream stream = nullptr) { check(b2f_clear_values(handle(), stream)); } // HypreDoFLinearSystem.cc:184-191
  // prelude of HypreDoFLinearSystemImpl::solve(): _applyRowElimination() + _applyForcedValuesToLhs(), :319-382,:387-391
  void applyDirichlet(b2f_stream stream = nullptr) { check(b2f_dirichlet_apply(handle(), stream)); }
  // Device-resident replacement of the host loops + copies in front of HYPRE_IJMatrixSetValues / HYPRE_IJVectorSetValues
  // (HypreDoFLinearSystem.cc:474-519,587-640): own rows only, global ids from `dof_matrix_numbering` (device, may be null).
  b2f_ij_view toHypreIJ(const Int32* dof_matrix_numbering = nullptr, b2f_stream stream = nullptr) const
  {
    b2f_ij_view v{};
    check(b2f_to_hypre_ij(handle(), dof_matrix_numbering, &v, stream));
    return v;
  }
  b2f_handle handle() const
  {
    if (!m_h || !*m_h)
      throw std::logic_error("DoFLinearSystem is not attached: call BSRFormat::toLinearSystem(linear_system) first");
    return *m_h;
  }

 private:
  template <int N> friend class BSRFormat;
  b2f_system_view sys() const { b2f_system_view s{}; check(b2f_get_system(handle(), &s)); return s; }
  b2f_handle* m_h = nullptr;
  CSRFormatView m_csr;
};

// femutils/BSRFormat.h:353-1140
template <int NB_DOF>
class BSRFormat
{
 public:
  explicit BSRFormat(b2f_stream stream = nullptr) : m_stream(stream), m_matrix(&m_h) {}
  BSRFormat(const BSRFormat&) = delete;
  BSRFormat& operator=(const BSRFormat&) = delete;
  ~BSRFormat() { if (m_h) b2f_destroy(m_h); }

  // BSRFormat.h:389-409.  assembly_variant: B2F_ASM_DEFAULT picks the library's choice; use_atomic_free maps to the node gather.
  void initialize(const b2f_mesh_view& mesh, bool does_linear_system_use_csr, bool use_atomic_free = false, int assembly_variant = B2F_ASM_DEFAULT)
  {
    if (m_h) {
      b2f_destroy(m_h);
      m_h = nullptr;
    }
    b2f_options opt{};
    opt.block_size = NB_DOF;
    opt.use_csr = does_linear_system_use_csr ? 1 : 0;
    opt.assembly_variant = (assembly_variant == B2F_ASM_DEFAULT && use_atomic_free) ? B2F_ASM_NODE_GATHER : assembly_variant;
    check(b2f_create(&mesh, &opt, &m_h));
    m_use_csr_in_linear_system = does_linear_system_use_csr;
  }
  void computeSparsity() { check(b2f_compute_sparsity(m_h, m_stream)); } // BSRFormat.h:775-781
  void assembleBilinear(const Operator& op) { check(b2f_assemble(m_h, &op.desc, m_stream)); } // BSRFormat.h:1105-1111
#ifdef __CUDACC__
  // BSRFormat.h:1105-1111 with a callback of the caller, in the reference's shape (the full element matrix of one cell):
  //   __device__ void f(long long cell, const double (&X)[NV][3], double (&Ke)[NB_DOF*NV][NB_DOF*NV])
  // NV = 3 (Triangle3) or 4 (Tetraedron4) is a template argument here; the reference picks the callback from
  // mesh->dimension() at run time (modules/poisson/FemModule.cc:255-286).  See include/b2f/user_operator.cuh.
  template <int NV, class Function>
  void assembleBilinear(const Function& compute_element_matrix, bool accumulate = false)
  {
    check(b2f::assemble_with_element_matrix<NV, NB_DOF>(m_h, compute_element_matrix, accumulate, (cudaStream_t)m_stream));
  }
#endif
  // assembleBilinear + BoundaryConditions{2D,3D}::applyConstantSourceToRhs(qdot, ...) (ArcaneFemFunctionsGpu.h:605,695) fused:
  // the source term is summed along the walk that forms the matrix rows.  Needs toLinearSystem() first (the rhs lives there).
  void assembleBilinearWithSource(const Operator& op, Real qdot) { check(b2f_assemble_with_source(m_h, &op.desc, qdot, m_stream)); }
  // Multi-GPU (one rank per GPU, own cells + ghost-cell layer, b2f.h "Multi-GPU"): the plan once after computeSparsity(), then one
  // call per assembly; the NCCL exchange of the ghost rows runs behind the assembly of the rows that do not travel.
  void setGhostExchange(void* nccl_comm, int32_t rank, int32_t world, const int32_t* node_owner_rank, const int64_t* node_unique_id)
  {
    check(b2f_ghost_plan(m_h, nccl_comm, rank, world, node_owner_rank, node_unique_id, m_stream));
    m_comm = nccl_comm;
  }
  void assembleBilinearAndExchange(const Operator& op) { check(b2f_assemble_exchange(m_h, &op.desc, m_comm, m_stream)); }
  // BSRFormat.h:414-430: hands the CSR view to the linear system (views stay valid until the next computeSparsity)
  // A CSR consumer (m_use_csr_in_linear_system, :416-427) gets the CSR view; any other linear system (the per-block branch,
  // BSRMatrix::toLinearSystem :261-280, which adds entry by entry on the host) is attached to the same device-resident
  // values: the Dirichlet / rhs entries of the C ABI handle both layouts.
  void toLinearSystem(DoFLinearSystem& linear_system)
  {
    linear_system.m_h = &m_h; // attached first: applyDirichlet / rhsVariable / clearValues work for either layout
    if (m_use_csr_in_linear_system || NB_DOF == 1) {
      if (!linear_system.hasSetCSRValues())
        throw std::invalid_argument("BSRFormat(toLinearSystem): Linear system was set to use CSR but is incompatible");
      linear_system.setCSRValues(m_matrix.toCsr(m_stream));
    }
  }
  BSRMatrix<NB_DOF>& matrix() { return m_matrix; } // BSRFormat.h:1116
  void resetMatrixValues() { check(b2f_reset_values(m_h, m_stream)); } // :1120-1123
  void dumpMatrix(const std::string& filename) { m_matrix.dump(filename); } // :1127
  b2f_handle handle() const { return m_h; }
  b2f_stream stream() const { return m_stream; }

 private:
  b2f_handle m_h = nullptr;
  b2f_stream m_stream;
  void* m_comm = nullptr; // ncclComm_t of setGhostExchange
  bool m_use_csr_in_linear_system = true;
  BSRMatrix<NB_DOF> m_matrix;
};

} // namespace ArcaneFemB200
