// arcanefem_b200/arcane/DoFLinearSystem.h — drop-in for femutils/DoFLinearSystem.h + femutils/IDoFLinearSystemFactory.h +
// the Hypre service of femutils/HypreDoFLinearSystem.cc on the assembly path, with the reference's names and argument lists:
//   CSRFormatView                 femutils/DoFLinearSystem.h:42-76
//   DoFLinearSystemImpl           :84-110 (the implementation interface; the entries the BSR path reaches)
//   IDoFLinearSystemFactory       femutils/IDoFLinearSystemFactory.h:34-44 (createInstance(ISubDomain*, IItemFamily*, const String&))
//   DoFLinearSystem               femutils/DoFLinearSystem.h:126-286, femutils/DoFLinearSystem.cc:451-522
//   B200HypreLinearSystem{Impl,Factory}   femutils/HypreDoFLinearSystem.cc:184-242 (state), :319-391 (Dirichlet prelude of solve()),
//                                 :474-640 (own rows, global ids, HYPRE_IJMatrixSetValues / HYPRE_IJVectorSetValues arguments), :767-800 (factory)
// State (rhs, forced / elimination info and values) lives in the library handle the BSRFormat attaches (toLinearSystem), so
// the variables are views on device arrays: fill() runs on the device, the Dirichlet kernels of ArcaneFemFunctionsGpu.h write
// them through the C ABI.  solve() runs the reference's prelude on the device and produces the device-resident argument
// arrays of the HYPRE_IJ calls; with Hypre present (B2F_HAVE_HYPRE, not in this image) it goes on to call them, otherwise it
// stops there and keeps the arrays for the caller (lastHypreArguments()).
#pragma once
#include <arcane/utils/ArcaneGlobal.h>
#include <arcane/IItemFamily.h>
#include <arcane/utils/CommandLineArguments.h>
#include "../BSRFormat.h"

namespace Arcane::FemUtils {

template <class T>
struct B200Span // Span<T> of Arcane on a device array
{
  T* m_p = nullptr;
  Int64 m_n = 0;
  T* data() const { return m_p; }
  Int64 size() const { return m_n; }
};

class CSRFormatView // femutils/DoFLinearSystem.h:42-76
{
 public:
  CSRFormatView() = default;
  CSRFormatView(B200Span<const Int32> rows, B200Span<const Int32> matrix_rows_nb_column, B200Span<const Int32> columns, B200Span<Real> values)
  : m_matrix_rows(rows), m_matrix_rows_nb_column(matrix_rows_nb_column), m_matrix_columns(columns), m_values(values) {}
  B200Span<const Int32> rows() const { return m_matrix_rows; }
  B200Span<const Int32> rowsNbColumn() const { return m_matrix_rows_nb_column; }
  B200Span<const Int32> columns() const { return m_matrix_columns; }
  B200Span<Real> values() { return m_values; }
  Int32 nbRow() { return (Int32)m_matrix_rows.size(); }
  Int32 nbColumn() { return (Int32)m_matrix_columns.size(); }
  Int32 nbValue() { return (Int32)m_values.size(); }

 private:
  B200Span<const Int32> m_matrix_rows, m_matrix_rows_nb_column, m_matrix_columns;
  B200Span<Real> m_values;
};

// A per-DoF variable of the linear system (VariableDoFReal / VariableDoFBool / VariableDoFByte of the reference): a view on
// the device array of the attached handle.  fill() is what the modules call on them (modules/poisson/FemModule.cc:157-160).
template <class T, int KIND> // KIND: 0 rhs, 1 forced info, 2 forced value, 3 elimination info, 4 elimination value, 5 solution
class B200DoFVariable
{
 public:
  void attach(b2f_handle* h, b2f_stream* s) { m_h = h; m_stream = s; }
  bool attached() const { return m_h && *m_h; }
  T* devicePointer() const
  {
    b2f_system_view v{};
    ArcaneFemB200::check(b2f_get_system(handle(), &v));
    if constexpr (KIND == 0) return v.rhs;
    else if constexpr (KIND == 1) return reinterpret_cast<T*>(v.forced_info);
    else if constexpr (KIND == 2) return v.forced_value;
    else if constexpr (KIND == 3) return reinterpret_cast<T*>(v.elimination_info);
    else return v.elimination_value;
  }
  Int64 size() const
  {
    b2f_system_view v{};
    ArcaneFemB200::check(b2f_get_system(handle(), &v));
    return v.nb_dof;
  }
  void fill(T value)
  {
    if constexpr (KIND == 0) {
      ArcaneFemB200::check(b2f_rhs_fill(handle(), (double)value, *m_stream));
    }
    else {
      // byte patterns only (false / 0 / 0.0), like every fill of these variables in the reference
      if (value != T(0))
        throw std::invalid_argument("B200DoFVariable::fill: only 0 / false is supported for the Dirichlet state");
      if (cudaMemsetAsync(devicePointer(), 0, sizeof(T) * (size_t)size(), (cudaStream_t)*m_stream) != cudaSuccess)
        throw std::runtime_error("B200DoFVariable::fill: cudaMemsetAsync failed");
    }
  }
  b2f_handle handle() const
  {
    if (!attached())
      throw std::logic_error("linear-system variable is not attached: call BSRFormat::toLinearSystem(linear_system) first");
    return *m_h;
  }
  b2f_stream stream() const { return *m_stream; }

 private:
  b2f_handle* m_h = nullptr;
  b2f_stream* m_stream = nullptr;
};
using VariableDoFReal = B200DoFVariable<Real, 0>;
using VariableDoFBool = B200DoFVariable<bool, 1>;
using VariableDoFByte = B200DoFVariable<Byte, 3>;
using VariableDoFForcedValue = B200DoFVariable<Real, 2>;
using VariableDoFEliminationValue = B200DoFVariable<Real, 4>;

// femutils/DoFLinearSystem.h:84-110 — the entries the BSR / GPU path reaches (the per-entry CPU entries matrixAddValue,
// eliminateRow ... belong to the legacy formats, SURVEY.md §2 out of scope, and throw).
class DoFLinearSystemImpl
{
 public:
  virtual ~DoFLinearSystemImpl() = default;
  virtual void matrixAddValue(DoFLocalId row, DoFLocalId column, Real value) = 0;
  virtual void matrixSetValue(DoFLocalId row, DoFLocalId column, Real value) = 0;
  virtual void eliminateRow(DoFLocalId row, Real value) = 0;
  virtual void eliminateRowColumn(DoFLocalId row, Real value) = 0;
  virtual void solve() = 0;
  virtual VariableDoFReal& rhsVariable() = 0;
  virtual void setSolverCommandLineArguments(const CommandLineArguments& args) = 0;
  virtual void clearValues() = 0;
  virtual void setCSRValues(const CSRFormatView& csr_view) = 0;
  virtual CSRFormatView& getCSRValues() = 0;
  virtual VariableDoFForcedValue& getForcedValue() = 0;
  virtual VariableDoFBool& getForcedInfo() = 0;
  virtual VariableDoFByte& getEliminationInfo() = 0;
  virtual VariableDoFEliminationValue& getEliminationValue() = 0;
  virtual bool hasSetCSRValues() const = 0;
  virtual void setRunner(Accelerator::Runner* r) = 0;
  virtual Accelerator::Runner* runner() const = 0;
  // attachment of the library handle that owns matrix and state (BSRFormat::toLinearSystem)
  virtual void attachHandle(b2f_handle* h, b2f_stream* s) = 0;
};

class IDoFLinearSystemFactory // femutils/IDoFLinearSystemFactory.h:34-44
{
 public:
  virtual ~IDoFLinearSystemFactory() = default;
  virtual DoFLinearSystemImpl* createInstance(ISubDomain* sd, IItemFamily* dof_family, const String& solver_name) = 0;
};

// The replacement of HypreDoFLinearSystemImpl (femutils/HypreDoFLinearSystem.cc:104-760) on the device-resident path.
class B200HypreLinearSystemImpl : public DoFLinearSystemImpl
{
 public:
  B200HypreLinearSystemImpl(ISubDomain* sd, IItemFamily* dof_family, const String& solver_name) : m_sd(sd), m_dof_family(dof_family), m_name(solver_name) {}
  void matrixAddValue(DoFLocalId, DoFLocalId, Real) override { ARCANE_THROW(NotImplementedException, "B200HypreLinearSystem: matrixAddValue (per-entry CPU assembly) is not on the BSR path"); }
  void matrixSetValue(DoFLocalId row, DoFLocalId column, Real value) override { ArcaneFemB200::check(b2f_set_value(handle(), row.localId(), column.localId(), value)); }
  void eliminateRow(DoFLocalId, Real) override { ARCANE_THROW(NotImplementedException, "B200HypreLinearSystem: use applyDirichletViaRowElimination (device flags)"); }
  void eliminateRowColumn(DoFLocalId, Real) override { ARCANE_THROW(NotImplementedException, "B200HypreLinearSystem: use the device elimination flags"); }
  VariableDoFReal& rhsVariable() override { return m_rhs; }
  void setSolverCommandLineArguments(const CommandLineArguments& args) override { m_args = args; }
  void clearValues() override { ArcaneFemB200::check(b2f_clear_values(handle(), *m_stream)); } // HypreDoFLinearSystem.cc:184-191
  void setCSRValues(const CSRFormatView& v) override { m_csr = v; m_has_csr = true; } // the view is kept, not copied (:199-202)
  CSRFormatView& getCSRValues() override { return m_csr; }
  VariableDoFForcedValue& getForcedValue() override { return m_forced_value; }
  VariableDoFBool& getForcedInfo() override { return m_forced_info; }
  VariableDoFByte& getEliminationInfo() override { return m_elim_info; }
  VariableDoFEliminationValue& getEliminationValue() override { return m_elim_value; }
  bool hasSetCSRValues() const override { return true; } // a CSR consumer (:235)
  void setRunner(Accelerator::Runner* r) override { m_runner = r; }
  Accelerator::Runner* runner() const override { return m_runner; }
  void attachHandle(b2f_handle* h, b2f_stream* s) override
  {
    m_h = h;
    m_stream = s;
    m_rhs.attach(h, s);
    m_forced_info.attach(h, s);
    m_forced_value.attach(h, s);
    m_elim_info.attach(h, s);
    m_elim_value.attach(h, s);
  }
  // HypreDoFLinearSystemImpl::solve(): prelude _applyRowElimination() + _applyForcedValuesToLhs() (:319-391) on the device-resident
  // matrix, then the own rows with global ids as the argument arrays of HYPRE_IJMatrixSetValues / HYPRE_IJVectorSetValues
  // (:474-640).  dof_matrix_numbering: device array first_own_row + own index / owner's number (:264-303), nullptr = one rank.
  void solve() override
  {
    ArcaneFemB200::check(b2f_dirichlet_apply(handle(), *m_stream));
    ArcaneFemB200::check(b2f_to_hypre_ij(handle(), m_dof_matrix_numbering, &m_ij, *m_stream));
    m_solved = true;
#ifdef B2F_HAVE_HYPRE
    hypreSolve(); // HYPRE_IJMatrixSetValues(ij_A, nb_row, ncols, rows, cols, values) ... on the arrays above (arcanefem_b200/arcane/HypreCalls.h)
#endif
  }
  void setDoFMatrixNumbering(const Int32* device_numbering) { m_dof_matrix_numbering = device_numbering; }
  const b2f_ij_view& lastHypreArguments() const
  {
    if (!m_solved)
      throw std::logic_error("B200HypreLinearSystem: solve() has not run");
    return m_ij;
  }
  b2f_handle handle() const
  {
    if (!m_h || !*m_h)
      throw std::logic_error("B200HypreLinearSystem is not attached: call BSRFormat::toLinearSystem(linear_system) first");
    return *m_h;
  }
  const String& solverName() const { return m_name; }

 private:
#ifdef B2F_HAVE_HYPRE
  void hypreSolve();
#endif
  ISubDomain* m_sd;
  IItemFamily* m_dof_family;
  String m_name;
  b2f_handle* m_h = nullptr;
  b2f_stream* m_stream = nullptr;
  Accelerator::Runner* m_runner = nullptr;
  CommandLineArguments m_args;
  CSRFormatView m_csr;
  bool m_has_csr = false, m_solved = false;
  const Int32* m_dof_matrix_numbering = nullptr;
  b2f_ij_view m_ij{};
  VariableDoFReal m_rhs;
  VariableDoFBool m_forced_info;
  VariableDoFForcedValue m_forced_value;
  VariableDoFByte m_elim_info;
  VariableDoFEliminationValue m_elim_value;
};

// femutils/HypreDoFLinearSystem.cc:767-800: the service a case file selects with <linear-system name="HypreLinearSystem">.
// With Arcane present it is registered under that name in place of the reference's service:
//   ARCANE_REGISTER_SERVICE_HYPREDOFLINEARSYSTEMFACTORY(HypreLinearSystem, B200HypreLinearSystemFactory);
class B200HypreLinearSystemFactory : public IDoFLinearSystemFactory
{
 public:
  DoFLinearSystemImpl* createInstance(ISubDomain* sd, IItemFamily* dof_family, const String& solver_name) override
  {
    return new B200HypreLinearSystemImpl(sd, dof_family, solver_name); // the caller (DoFLinearSystem) owns it, like :775-780
  }
};

// femutils/DoFLinearSystem.h:126-286
class DoFLinearSystem
{
 public:
  DoFLinearSystem() = default;
  DoFLinearSystem(const DoFLinearSystem&) = delete;
  DoFLinearSystem& operator=(const DoFLinearSystem&) = delete;
  ~DoFLinearSystem() { delete m_p; }
  // femutils/DoFLinearSystem.cc:405-425
  void initialize(ISubDomain* sd, Accelerator::Runner* runner, IItemFamily* dof_family, const String& solver_name)
  {
    ARCANE_CHECK_POINTER(sd);
    ARCANE_CHECK_POINTER(dof_family);
    if (m_p)
      ARCANE_FATAL("The instance is already initialized");
    if (!m_linear_system_factory)
      ARCANE_FATAL("No linear system factory: call setLinearSystemFactory() first");
    m_p = m_linear_system_factory->createInstance(sd, dof_family, solver_name);
    m_p->setRunner(runner);
    if (m_h)
      m_p->attachHandle(m_h, m_stream);
  }
  bool isInitialized() const { return m_p != nullptr; }
  void reset() // :516-522: destroys the implementation; the matrix itself belongs to the BSRFormat
  {
    delete m_p;
    m_p = nullptr;
  }
  void setLinearSystemFactory(IDoFLinearSystemFactory* factory) { m_linear_system_factory = factory; }
  void setSolverCommandLineArguments(const CommandLineArguments& args) { impl()->setSolverCommandLineArguments(args); }
  void matrixAddValue(DoFLocalId row, DoFLocalId column, Real value) { impl()->matrixAddValue(row, column, value); }
  void matrixSetValue(DoFLocalId row, DoFLocalId column, Real value) { impl()->matrixSetValue(row, column, value); }
  void eliminateRow(DoFLocalId row, Real value) { impl()->eliminateRow(row, value); }
  void eliminateRowColumn(DoFLocalId row, Real value) { impl()->eliminateRowColumn(row, value); }
  void solve() { impl()->solve(); }
  VariableDoFReal& rhsVariable() { return impl()->rhsVariable(); }
  void clearValues() { impl()->clearValues(); }
  void setCSRValues(const CSRFormatView& csr_view) { impl()->setCSRValues(csr_view); } // :451-456
  CSRFormatView& getCSRValues() { return impl()->getCSRValues(); }
  VariableDoFForcedValue& getForcedValue() { return impl()->getForcedValue(); }
  VariableDoFBool& getForcedInfo() { return impl()->getForcedInfo(); }
  VariableDoFByte& getEliminationInfo() { return impl()->getEliminationInfo(); }
  VariableDoFEliminationValue& getEliminationValue() { return impl()->getEliminationValue(); }
  bool hasSetCSRValues() { return impl()->hasSetCSRValues(); } // :506-511
  DoFLinearSystemImpl* implementation() const { return m_p; }
  // the library handle behind matrix and state (set by BSRFormat::toLinearSystem)
  void attachHandle(b2f_handle* h, b2f_stream* s)
  {
    m_h = h;
    m_stream = s;
    if (m_p)
      m_p->attachHandle(h, s);
  }
  b2f_handle handle() const
  {
    if (!m_h || !*m_h)
      throw std::logic_error("DoFLinearSystem is not attached: call BSRFormat::toLinearSystem(linear_system) first");
    return *m_h;
  }
  b2f_stream stream() const { return m_stream ? *m_stream : nullptr; }

 private:
  DoFLinearSystemImpl* impl() const
  {
    if (!m_p)
      ARCANE_FATAL("The instance is not initialized. You need to call initialize() before using this class");
    return m_p;
  }
  DoFLinearSystemImpl* m_p = nullptr;
  IDoFLinearSystemFactory* m_linear_system_factory = nullptr;
  b2f_handle* m_h = nullptr;
  b2f_stream* m_stream = nullptr;
};

} // namespace Arcane::FemUtils
