#pragma once
#include "../../core/IMesh.h"
namespace Arcane { namespace Accelerator {
// stub of RunQueue: the two members the adapter uses
class RunQueue
{
 public:
  eMemoryRessource memoryRessource() const { return eMemoryRessource::UnifiedMemory; }
  void* platformStream() const { return nullptr; }
};
}} // namespace Arcane::Accelerator
