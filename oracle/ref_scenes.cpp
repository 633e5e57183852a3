// ref_scenes.cpp — TEST INFRASTRUCTURE, NOT PRODUCT CODE.
//
// Wraps the scene-construction part of the reference's scenes/*.cu (everything above their host
// main()/MPI/texture-upload code) so that the reference's own InitWorld / InitModel run on the
// host.  The *_head.inc files are generated at build time by oracle/Makefile from
// /root/reference/scenes/*.cu into oracle/_ref/gen/ (git-ignored; never committed): the file is
// cut before the first host-only function and its #include lines are dropped, because every
// header it needs is included here at global scope first.
#include <cstdio>
#include <fstream>
#include <iostream>
#include <string>
#include <tuple>
#include <vector>

#include <glm/glm.hpp>
#include <glm/gtc/constants.hpp>
#ifndef GLM_ENABLE_EXPERIMENTAL
#define GLM_ENABLE_EXPERIMENTAL
#endif
#include <glm/gtx/rotate_vector.hpp>

#include "bvh.cuh"
#include "camera.cuh"
#include "dielectric.cuh"
#include "diffuse_light.cuh"
#include "hitable_list.cuh"
#include "lambertian.cuh"
#include "metal.cuh"
#include "parallelepiped.cuh"
#include "parallelogram.cuh"
#include "ray_tracing.cuh"
#include "sky.cuh"
#include "sphere.cuh"
#include "textures/constant_texture.cuh"
#include "textures/image_texture.cuh"
#include "utils.cuh"

namespace scene_spheres {
#include "gen/spheres_head.inc"
}
namespace scene_cornell_box {
#include "gen/cornell_box_head.inc"
}
namespace scene_bunny {
#include "gen/bunny_head.inc"
}
namespace scene_birthday {
#include "gen/birthday_head.inc"
}
