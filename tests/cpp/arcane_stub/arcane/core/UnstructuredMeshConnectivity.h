#pragma once
#include "IMesh.h" // stub: the adapter reads connectivity through Cell::nodeId
