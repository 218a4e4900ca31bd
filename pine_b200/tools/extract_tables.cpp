// pine_b200/tools/extract_tables.cpp -- PRODUCT BUILD STEP (pine_b200/csrc/Makefile `tables`).
// Extracts the Heitz et al. 2019 blue-noise Sobol tables (third-party data vendored by the
// reference under src/contrib/bluesobol/, see CITATION.txt there) by COMPILING the vendored table
// sources into this program and dumping the arrays as u8. Nothing is retyped or copied as source.
//
// Output layout (pine_b200/data/bluesobol_tables.bin), little endian:
//   char magic[8] = "PB2BLUE1"; u32 n_variants = 9; u32 reserved;
//   per variant, in order spp = 1,2,4,8,16,32,64,128,256:
//     u32 spp; u32 reserved; u8 sobol[65536]; u8 scramble[131072]; u8 rank[131072];
// Reference consumer of the same arrays: bluenoise_Nspp(), bluenoise_16spp.cpp:14-34.
#include <cstdint>
#include <cstdio>
#include <cstring>

namespace t1 {
#include "bluesobol/bluenoise_1spp.cpp"
}
namespace t2 {
#include "bluesobol/bluenoise_2spp.cpp"
}
namespace t4 {
#include "bluesobol/bluenoise_4spp.cpp"
}
namespace t8 {
#include "bluesobol/bluenoise_8spp.cpp"
}
namespace t16 {
#include "bluesobol/bluenoise_16spp.cpp"
}
namespace t32 {
#include "bluesobol/bluenoise_32spp.cpp"
}
namespace t64 {
#include "bluesobol/bluenoise_64spp.cpp"
}
namespace t128 {
#include "bluesobol/bluenoise_128spp.cpp"
}
namespace t256 {
#include "bluesobol/bluenoise_256spp.cpp"
}

static int dump(FILE* f, uint32_t spp, const int* sobol, const int* scr, const int* rank) {
  uint32_t hdr[2] = {spp, 0};
  fwrite(hdr, 4, 2, f);
  static uint8_t buf[131072];
  auto put = [&](const int* src, int n) {
    for (int i = 0; i < n; i++) {
      if (src[i] < 0 || src[i] > 255) return 1;
      buf[i] = uint8_t(src[i]);
    }
    fwrite(buf, 1, n, f);
    return 0;
  };
  return put(sobol, 65536) | put(scr, 131072) | put(rank, 131072);
}

int main(int argc, char** argv) {
  if (argc < 2) return 2;
  FILE* f = fopen(argv[1], "wb");
  if (!f) return 3;
  fwrite("PB2BLUE1", 1, 8, f);
  uint32_t hdr[2] = {9, 0};
  fwrite(hdr, 4, 2, f);
  int bad = 0;
#define D(N) bad |= dump(f, N, t##N::sobol_256spp_256d, t##N::scramblingTile, t##N::rankingTile)
  D(1); D(2); D(4); D(8); D(16); D(32); D(64); D(128); D(256);
  fclose(f);
  if (bad) { fprintf(stderr, "table value out of u8 range\n"); return 4; }
  return 0;
}
