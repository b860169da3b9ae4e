// Model of femutils/IArcaneFemBC.h:19-72 (the boundary-condition option interfaces a case file fills).  TEST INFRASTRUCTURE.
#pragma once
#include "arcane_mock.h"
namespace Arcane::FemUtils::BC {
class IDirichletPointCondition
{
 public:
  virtual ~IDirichletPointCondition() = default;
  virtual Arcane::NodeGroup getNode() = 0;
  virtual Real getValue() = 0;
  virtual Real getPenalty() = 0;
  virtual String getEnforceDirichletMethod() = 0;
};
class IDirichletBoundaryCondition
{
 public:
  virtual ~IDirichletBoundaryCondition() = default;
  virtual Arcane::FaceGroup getSurface() = 0;
  virtual Real getValue() = 0;
  virtual Real getPenalty() = 0;
  virtual String getEnforceDirichletMethod() = 0;
};
class INeumannBoundaryCondition
{
 public:
  virtual ~INeumannBoundaryCondition() = default;
  virtual Arcane::FaceGroup getSurface() = 0;
  virtual Real getValue() = 0;
  virtual Real getValueX() = 0;
  virtual Real getValueY() = 0;
  virtual Real getValueZ() = 0;
  virtual bool hasValue() const = 0;
  virtual bool hasValueX() const = 0;
  virtual bool hasValueY() const = 0;
  virtual bool hasValueZ() const = 0;
};
class IArcaneFemBC
{
 public:
  virtual ~IArcaneFemBC() = default;
  virtual ConstArrayView<BC::IDirichletPointCondition*> dirichletPointConditions() = 0;
  virtual ConstArrayView<BC::IDirichletBoundaryCondition*> dirichletBoundaryConditions() = 0;
  virtual ConstArrayView<BC::INeumannBoundaryCondition*> neumannBoundaryConditions() = 0;
};
} // namespace Arcane::FemUtils::BC
