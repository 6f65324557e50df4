// fuzz_loaders.cpp — sanitizer harness for the host loaders (not part of the library): every file named on the command
// line goes through the image decoders or, by extension, the model loaders; exceptions are the expected answer to a
// malformed file, a sanitizer report is a bug.  Built and driven by scripts/fuzz_loaders.py:
//   g++ -O1 -g -std=c++17 -fsanitize=address,undefined -fno-sanitize-recover=undefined -I ray_tracer_b200/host \
//       tests/native/fuzz_loaders.cpp ray_tracer_b200/host/rt_resource.cpp -lz -o /tmp/fuzz_loaders
#include <cstdio>
#include <cstring>
#include <stdexcept>
#include "rt_host.hpp"

static bool ends(const char* s, const char* suf) { size_t a = strlen(s), b = strlen(suf); return a >= b && !strcmp(s + a - b, suf); }

int main(int argc, char** argv)
{
    int ok = 0, err = 0;
    for (int i = 1; i < argc; i++) {
        const bool model = ends(argv[i], ".glb") || ends(argv[i], ".gltf") || ends(argv[i], ".obj");
        for (int mode = 0; mode < (model ? 3 : 1); mode++) {
            std::vector<rt_vertex> v; std::vector<uint32_t> ix; std::vector<rt_mesh> m;
            try {
                if (!model) rtb::load_image(argv[i]);
                else if (mode == 0) rtb::load_model(argv[i], v, ix, m);
                else if (mode == 1) rtb::load_scene_glb(argv[i], v, ix, m);
                else rtb::load_glb_images(argv[i]);
                ok++;
            } catch (const std::exception&) { err++; }
        }
    }
    printf("ok %d err %d\n", ok, err);
    return 0;
}
