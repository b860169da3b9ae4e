// Model of the header arcane_generate_axl makes from modules/poisson/Fem.axl: the module base class ArcaneFemObject with its
// options (bsr, bsr-atomic-free, f, linear-system service, boundary-conditions), mesh / sub-domain accessors and the
// variables the module text uses.  TEST INFRASTRUCTURE, see arcane_mock.h.
#pragma once
#include "arcane_mock.h"
#include <arcanefem_b200/arcane/DoFLinearSystem.h>
#include "IArcaneFemBC.h"

template <class T>
struct MockCaseOption
{
  T value{};
  bool present = true;
  T operator()() const { return value; }
  operator T() const { return value; } // the generated option converts to its value (modules/poisson/FemModule.cc:76 relies on it)
  bool isPresent() const { return present; }
};
struct MockLinearSystemOption
{
  Arcane::String name = "HypreLinearSystem";
  Arcane::FemUtils::IDoFLinearSystemFactory* factory = nullptr;
  Arcane::String serviceName() const { return name; }
  Arcane::FemUtils::IDoFLinearSystemFactory* operator()() const { return factory; }
};
struct CaseOptionsFem
{
  MockCaseOption<bool> bsr, bsrAtomicFree;
  MockCaseOption<Arcane::Real> f;
  MockLinearSystemOption linearSystem;
  Arcane::FemUtils::BC::IArcaneFemBC* bc = nullptr;
  Arcane::FemUtils::BC::IArcaneFemBC* boundaryConditions() const { return bc; }
};

class ArcaneFemObject : public Arcane::TraceAccessor
{
 public:
  explicit ArcaneFemObject(const Arcane::ModuleBuildInfo& mbi)
  : Arcane::TraceAccessor(mbi.subDomain()->traceMng()), m_node_coord(mbi.mesh()->nodesCoordinates()), m_sd(mbi.subDomain()), m_mesh(mbi.mesh()) {}
  virtual ~ArcaneFemObject() = default;
  virtual void startInit() = 0;
  virtual void compute() = 0;
  virtual Arcane::VersionInfo versionInfo() const = 0;
  CaseOptionsFem* options() { return &m_options; }
  Arcane::IMesh* mesh() const { return m_mesh; }
  Arcane::ISubDomain* subDomain() const { return m_sd; }
  Arcane::Accelerator::IAcceleratorMng* acceleratorMng() const { return m_sd->acceleratorMng(); }
  Arcane::CellGroup allCells() const { return m_mesh->allCells(); }

  Arcane::VariableNodeReal3 m_node_coord;
  struct { Arcane::Int32 value = 0; Arcane::Int32 operator()() const { return value; } } m_global_iteration;

 private:
  Arcane::ISubDomain* m_sd;
  Arcane::IMesh* m_mesh;
  CaseOptionsFem m_options;
};
