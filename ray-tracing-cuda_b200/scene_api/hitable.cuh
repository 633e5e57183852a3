// Forwarding header: the reference declares this interface in ray-tracing-cuda/hitable.cuh; here every class lives in
// rt_scene_api.cuh (see its header comment).
#pragma once
#include "rt_scene_api.cuh"
