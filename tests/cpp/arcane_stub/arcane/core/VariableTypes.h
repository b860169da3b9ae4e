#pragma once
#include <vector>
#include "IMesh.h"
namespace Arcane {
// stub of VariableNodeReal3: asArray() views the per-node Real3 values in local-id order
class VariableNodeReal3
{
 public:
  struct ArrayView { const Real3* m_p; const Real3* data() const { return m_p; } };
  explicit VariableNodeReal3(std::vector<Real3> v) : m_v(std::move(v)) {}
  ArrayView asArray() const { return ArrayView{ m_v.data() }; }
 private:
  std::vector<Real3> m_v;
};
} // namespace Arcane
