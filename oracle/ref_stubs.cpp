// oracle/ref_stubs.cpp -- TEST INFRASTRUCTURE ONLY.
// Link-time stubs for the two reference subsystems that cannot be built in this image
// (SURVEY.md F3): EmbreeAccel (needs libembree, src/pine/impl/accel/embree.cpp) and the PRL JIT
// (needs libLLVM-18, src/pine/core/jit.cpp). Neither is on the oracle's path: the harness always
// constructs PathIntegrator with pine::BVH (accel.h:10) through the C++ API.
#include <pine/core/accel.h>
#include <pine/core/jit.h>

#include <cstdio>
#include <cstdlib>

namespace pine {
[[noreturn]] static void unavailable(const char* what) {
  std::fprintf(stderr, "[oracle] %s is not available in this build\n", what);
  std::abort();
}
#ifndef PINE_REF_WITH_EMBREE  // libpine_ref_embree.so compiles the reference's own embree.cpp instead (oracle/Makefile)
void EmbreeAccel::build(const Scene*) { unavailable("EmbreeAccel::build"); }
bool EmbreeAccel::hit(Ray) const { unavailable("EmbreeAccel::hit"); }
uint8_t EmbreeAccel::hit8(psl::span<const Ray>) const { unavailable("EmbreeAccel::hit8"); }
bool EmbreeAccel::intersect(Ray&, SurfaceInteraction&) const { unavailable("EmbreeAccel::intersect"); }
#endif
void jit_interpret(Context&, psl::string) { unavailable("jit_interpret (PRL)"); }
}  // namespace pine
