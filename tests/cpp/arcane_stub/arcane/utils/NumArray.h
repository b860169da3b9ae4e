#pragma once
#include <vector>
#include "../core/IMesh.h"
namespace Arcane {
// stub of NumArray<T, MDDim1>: host std::vector behind the calls the adapter makes
template <class T, class Extents>
class NumArray
{
 public:
  struct Span1 { T* m_p; T* data() const { return m_p; } };
  explicit NumArray(eMemoryRessource) {}
  void resize(Int64 n) { m_v.resize((size_t)n); }
  T& operator[](Int64 i) { return m_v[(size_t)i]; }
  Span1 to1DSpan() { return Span1{ m_v.data() }; }
 private:
  std::vector<T> m_v;
};
} // namespace Arcane
