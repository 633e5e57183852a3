// Test-infrastructure shim: the reference's model.h needs assimp (absent here).  The host
// oracle build only compiles the InitWorld/InitModel part of scenes/bunny.cu, which uses no
// Model<> member, so an empty stand-in suffices.
#pragma once
#include "bvh.cuh"
#include "utils.cuh"
