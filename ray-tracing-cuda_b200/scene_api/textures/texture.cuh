// Forwarding header for ray-tracing-cuda/textures/texture.cuh (see rt_scene_api.cuh).
#pragma once
#include "../rt_scene_api.cuh"
