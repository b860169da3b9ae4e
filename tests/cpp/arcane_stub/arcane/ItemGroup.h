// forwards to the model of the Arcane types (tests/cpp/arcane_stub/arcane_mock.h); Arcane itself is not installed here
#pragma once
#include "../arcane_mock.h"
