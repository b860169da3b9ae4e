// arcanefem_b200/arcane/FemUtils.h — drop-in for the part of femutils/FemUtils.h the element callbacks use (:34-40 Real4,
// :53-186 FixedMatrix, :191-226 operator^, :266-313 matrixAddition / matrixMultiplication / matrixTranspose), under the
// reference's own names (global Real4, Arcane::FemUtils::FixedMatrix) so that a module's callbacks
// (modules/poisson/FemModule.h:139-186, modules/elasticity/FemModule.h:112-152) compile unchanged.  Needs Arcane's Real3
// (real Arcane or tests/cpp/arcane_stub); the types themselves are the library's (arcanefem_b200/FemFunctionsGpu.cuh).
#pragma once
#include <arcane/utils/ArcaneGlobal.h>
#include "../FemFunctionsGpu.cuh"

using Real4 = ArcaneFemB200::Real4; // the reference declares Real4 at global scope (femutils/FemUtils.h:34)

namespace Arcane::FemUtils {

using ArcaneFemB200::FixedMatrix;
using ArcaneFemB200::matrixAddition;
using ArcaneFemB200::matrixMultiplication;
using ArcaneFemB200::matrixTranspose;

// femutils/FemUtils.h:191-200 on Arcane's own Real3
ARCCORE_HOST_DEVICE inline FixedMatrix<3, 3> operator^(const Arcane::Real3& lhs, const Arcane::Real3& rhs)
{
  FixedMatrix<3, 3> result;
  for (Arcane::Int32 i = 0; i < 3; ++i)
    for (Arcane::Int32 j = 0; j < 3; ++j)
      result(i, j) = lhs[i] * rhs[j];
  return result;
}

} // namespace Arcane::FemUtils

namespace Arcane {
using FemUtils::operator^; // found by argument-dependent lookup on Arcane::Real3, like the reference's (declared in Arcane::FemUtils, used with `using namespace`)
}
