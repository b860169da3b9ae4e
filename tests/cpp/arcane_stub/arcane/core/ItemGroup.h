#pragma once
#include "IMesh.h" // stub: groups live with the mesh
